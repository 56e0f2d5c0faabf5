// Native training-step executor for the bf16 engine.
//
// The reference drives one minibatch from C++ (parallel_train's loop body, ref
// fpcode/neural_network.cpp:484-541: 6 H2D copies, 13 synchronised launches, 4 D2H copies,
// 4 host all-reduces, host SGD).  cme_mlp_step is the same unit of work for this library's
// C/C++ users: given a plan with every device buffer of one rank it issues the whole step
// (staging, forward, backward, fused all-reduce + SGD, re-cast) with no host synchronisation.
//
// Schedule (the same one MLPEngineBF16._step_body issues from Python):
//   side   gradient memset, hidden under staging + the forward GEMM
//   main   staging, forward GEMM + sigmoid, layer 2 + softmax + CE
//   main   backward of hidden units [0, hs)           (plan->overlap and more than one rank)
//   side   barrier, all-reduce + SGD of W1 rows [0, hs), barrier      } concurrently
//   main   backward of hidden units [hs, H)                           }
//   main   barrier, all-reduce + SGD of the rest, barrier, join side, re-cast W2
// Without a side stream the same launches go to the one stream in that order, so the update
// ranges (and with them the owner of every slice of the sharded fp32 master) never depend on
// how the step was issued.
#include "common.cuh"

namespace {

struct Geometry {
    int D, H, C, K, ldx, ldh;
    int64_t n_w1, o_w2, o_b2, n_total;
};

Geometry geometry(const CmeMlpPlan* p) {
    Geometry g;
    g.D = p->D, g.H = p->H, g.C = p->C, g.K = p->D + 1, g.ldx = p->ldx, g.ldh = p->ldh;
    g.n_w1 = (int64_t)p->Hp * p->ldx;
    g.o_w2 = g.n_w1;
    g.o_b2 = g.o_w2 + (((int64_t)g.C * g.H + 3) / 4) * 4;
    g.n_total = g.o_b2 + ((g.C + 3) / 4) * 4;
    return g;
}

// Backward for hidden units [h0, h0 + hb): dZ1 and its share of dW2 in one pass, then dW1 rows.
int backward_block(const CmeMlpPlan* p, const Geometry& g, int64_t rows, int h0, int hb, void* stream) {
    auto* a1 = static_cast<__nv_bfloat16*>(p->a1) + h0;
    auto* dz1 = static_cast<__nv_bfloat16*>(p->dz1) + h0;
    auto* w2t = static_cast<__nv_bfloat16*>(p->w2t) + (int64_t)h0 * 64;
    float* grads = p->grads[p->rank];
    int rc = cme_mlp_bwd_tail_tc(a1, p->dz2b, w2t, dz1, grads + g.o_w2 + h0, g.H, rows, hb, g.ldh, g.C, stream);
    if (rc) return rc;
    return cme_mlp_dw1(dz1, p->xb, grads + (int64_t)h0 * g.ldx, rows, g.K, hb, g.ldx, g.ldh, g.ldx, stream);
}

// Barrier, fused all-reduce + L2 + SGD + re-cast of [r0, r1) of the flat vector, barrier.
int reduce_range(const CmeMlpPlan* p, const Geometry& g, float lr, float reg, int64_t r0, int64_t r1,
                 int channel, void* stream) {
    int rc;
    if (p->nranks > 1)
        if ((rc = cme_peer_barrier(p->pads_dev, p->rank, p->nranks, channel, 0, 1, 1, stream))) return rc;
    const int64_t w2_end = g.o_w2 + (int64_t)g.C * g.H;
    if (p->use_mc && p->nranks > 1)
        rc = cme_allreduce_sgd_mc(p->grads_mc, p->params[p->rank], p->params_mc, p->w1c_mc, p->rank, p->nranks,
                                  g.n_total, g.n_w1, g.ldx, g.D, g.o_w2, w2_end, lr, reg, r0, r1, stream);
    else
        rc = cme_allreduce_sgd(p->grads, p->params, p->w1c, p->rank, p->nranks, g.n_total, g.n_w1, g.ldx,
                               g.D, g.o_w2, w2_end, lr, reg, r0, r1, stream);
    if (rc) return rc;
    if (p->nranks > 1) return cme_peer_barrier(p->pads_dev, p->rank, p->nranks, channel + 1, 0, 1, 1, stream);
    return 0;
}

}  // namespace

extern "C" {

int cme_mlp_layout(const CmeMlpPlan* p, int64_t out[4]) {
    CME_CHECK_ARG(p && out && p->D > 0 && p->H > 0 && p->C > 0, "cme_mlp_layout: bad plan");
    CME_CHECK_ARG(p->ldx >= p->D + 1 && p->ldx % 64 == 0 && p->Hp >= p->H && p->Hp % 64 == 0 && p->ldh >= p->Hp,
                  "cme_mlp_layout: ldx=%d Hp=%d ldh=%d do not fit D=%d H=%d", p->ldx, p->Hp, p->ldh, p->D, p->H);
    const Geometry g = geometry(p);
    out[0] = g.n_w1, out[1] = g.o_w2, out[2] = g.o_b2, out[3] = g.n_total;
    return 0;
}

int cme_mlp_split_point(const CmeMlpPlan* p) {
    if (!p || !p->overlap || p->nranks <= 1) return 0;
    const int hs = (p->H / 2 + 255) / 256 * 256;
    return (hs > 0 && hs < p->H) ? hs : 0;
}

int cme_mlp_step_ex(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels, int64_t rows,
                    float lr, float reg, float scale, float* loss_sum, void* stream, void* side_stream,
                    void* const* events) {
    CME_CHECK_ARG(p && x_u8 && labels, "cme_mlp_step: null argument");
    CME_CHECK_ARG(rows > 0 && rows <= p->max_rows, "cme_mlp_step: rows %lld not in 1..%lld",
                  (long long)rows, (long long)p->max_rows);
    CME_CHECK_ARG(p->nranks >= 1 && p->nranks <= 16 && p->rank >= 0 && p->rank < p->nranks,
                  "cme_mlp_step: rank %d of %d", p->rank, p->nranks);
    CME_CHECK_ARG(p->nranks == 1 || p->pads_dev, "cme_mlp_step: pads_dev is required with more than one rank");
    CME_CHECK_ARG(!(p->use_mc && p->nranks > 1) || (p->grads_mc && p->params_mc && p->w1c_mc),
                  "cme_mlp_step: use_mc without multicast addresses");
    CME_CHECK_ARG(!side_stream || events, "cme_mlp_step: a side stream needs 4 events");
    const Geometry g = geometry(p);
    float* params = p->params[p->rank];
    float* grads = p->grads[p->rank];
    cudaStream_t main_s = cme::as_stream(stream);
    cudaStream_t side_s = side_stream ? cme::as_stream(side_stream) : main_s;
    auto ev = [&](int i) { return reinterpret_cast<cudaEvent_t>(events[i]); };
    int rc;

    if (side_stream) {
        CME_CUDA(cudaEventRecord(ev(0), main_s));
        CME_CUDA(cudaStreamWaitEvent(side_s, ev(0), 0));
    }
    CME_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * (size_t)g.n_total, side_s));
    if (side_stream) CME_CUDA(cudaEventRecord(ev(1), side_s));
    if ((rc = cme_u8_to_bf16_rows(x_u8, p->xb, rows, g.D, g.ldx, 1, stream))) return rc;
    if ((rc = cme_mlp_fwd1(p->xb, p->w1c[p->rank], nullptr, p->a1, rows, g.K, g.H, g.ldx, g.ldh, stream))) return rc;
    if (side_stream) CME_CUDA(cudaStreamWaitEvent(main_s, ev(1), 0));
    if ((rc = cme_mlp_fwd2_tc(p->a1, p->w2c, params + g.o_b2, labels, p->dz2, p->dz2b, loss_sum, p->pred,
                              grads + g.o_b2, scale, rows, g.H, g.ldh, g.C, stream))) return rc;

    const int hs = cme_mlp_split_point(p);
    if (hs == 0) {
        if ((rc = backward_block(p, g, rows, 0, g.H, stream))) return rc;
        if ((rc = reduce_range(p, g, lr, reg, 0, g.n_total, 0, stream))) return rc;
    } else {
        const int64_t cut = (int64_t)hs * g.ldx;
        if ((rc = backward_block(p, g, rows, 0, hs, stream))) return rc;
        if (side_stream) {
            CME_CUDA(cudaEventRecord(ev(2), main_s));
            CME_CUDA(cudaStreamWaitEvent(side_s, ev(2), 0));
        }
        if ((rc = reduce_range(p, g, lr, reg, 0, cut, 0, side_s))) return rc;
        if (side_stream) CME_CUDA(cudaEventRecord(ev(3), side_s));
        if ((rc = backward_block(p, g, rows, hs, g.H - hs, stream))) return rc;
        if ((rc = reduce_range(p, g, lr, reg, cut, g.n_total, 2, stream))) return rc;
        if (side_stream) CME_CUDA(cudaStreamWaitEvent(main_s, ev(3), 0));
    }
    return cme_mlp_recast_w2(params + g.o_w2, p->w2c, g.ldh, p->w2t, g.H, g.C, stream);
}

int cme_mlp_step(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels, int64_t rows, float lr,
                 float reg, float scale, float* loss_sum, void* stream) {
    return cme_mlp_step_ex(p, x_u8, labels, rows, lr, reg, scale, loss_sum, stream, nullptr, nullptr);
}

int cme_mlp_step_launches(const CmeMlpPlan* p) {
    if (!p) return 0;
    const int blocks = cme_mlp_split_point(p) ? 2 : 1;
    // staging, forward GEMM, layer 2, re-cast + per block: fused tail, dW1, update (+ 2 barriers)
    return 4 + blocks * (3 + (p->nranks > 1 ? 2 : 0));
}

}  // extern "C"
