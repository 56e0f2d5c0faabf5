// Native training-step executor for the bf16 engine.
//
// The reference drives one minibatch from C++ (parallel_train's loop body, ref
// fpcode/neural_network.cpp:484-541: 6 H2D copies, 13 synchronised launches, 4 D2H copies,
// 4 host all-reduces, host SGD).  cme_mlp_step is the same unit of work for this library's
// C/C++ users: given a plan with every device buffer of one rank it issues the whole step
// (staging, forward, backward, fused all-reduce + SGD, re-cast) on one stream with no host
// synchronisation.  The Python engine issues the same sequence (plus a side stream for
// overlap); this entry point exists so that a host program does not need Python at all.
#include "common.cuh"

extern "C" int cme_mlp_step(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels,
                            int64_t rows, float lr, float reg, float scale, float* loss_sum,
                            void* stream) {
    CME_CHECK_ARG(p && x_u8 && labels, "cme_mlp_step: null argument");
    CME_CHECK_ARG(rows > 0 && rows <= p->max_rows, "cme_mlp_step: rows %lld not in 1..%lld",
                  (long long)rows, (long long)p->max_rows);
    CME_CHECK_ARG(p->nranks >= 1 && p->nranks <= 16 && p->rank >= 0 && p->rank < p->nranks,
                  "cme_mlp_step: rank %d of %d", p->rank, p->nranks);
    const int D = p->D, H = p->H, C = p->C, K = D + 1;
    const int64_t n_w1 = (int64_t)p->Hp * p->ldx;
    const int64_t o_w2 = n_w1, o_b2 = o_w2 + (((int64_t)C * H + 3) / 4) * 4;
    const int64_t n_total = o_b2 + ((C + 3) / 4) * 4;
    float* params = p->params[p->rank];
    float* grads = p->grads[p->rank];
    float* w2 = params + o_w2;
    float* b2 = params + o_b2;
    float* gw2 = grads + o_w2;
    float* gb2 = grads + o_b2;
    cudaStream_t st = cme::as_stream(stream);
    int rc;
    if ((rc = cme_u8_to_bf16_rows(x_u8, p->xb, rows, D, p->ldx, 1, stream))) return rc;
    CME_CUDA(cudaMemsetAsync(grads, 0, sizeof(float) * (size_t)n_total, st));
    if ((rc = cme_mlp_fwd1(p->xb, p->w1c[p->rank], nullptr, p->a1, rows, K, H, p->ldx, p->ldh, stream))) return rc;
    if ((rc = cme_mlp_fwd2_tc(p->a1, p->w2c, b2, labels, p->dz2, p->dz2b, loss_sum, p->pred, gb2, scale,
                              rows, H, p->ldh, C, stream))) return rc;
    if ((rc = cme_mlp_bwd_tail_tc(p->a1, p->dz2b, p->w2t, p->dz1, gw2, H, rows, H, p->ldh, C, stream))) return rc;
    if ((rc = cme_mlp_dw1(p->dz1, p->xb, grads, rows, K, H, p->ldx, p->ldh, p->ldx, stream))) return rc;
    if (p->nranks > 1) {
        CME_CHECK_ARG(p->pads_dev, "cme_mlp_step: pads_dev is required with more than one rank");
        if ((rc = cme_peer_barrier(p->pads_dev, p->rank, p->nranks, 0, 0, 1, 1, stream))) return rc;
    }
    if ((rc = cme_allreduce_sgd(p->grads, p->params, p->w1c, p->rank, p->nranks, n_total, n_w1, p->ldx, D,
                                o_w2, o_w2 + (int64_t)C * H, lr, reg, 0, n_total, stream))) return rc;
    if (p->nranks > 1)
        if ((rc = cme_peer_barrier(p->pads_dev, p->rank, p->nranks, 1, 0, 1, 1, stream))) return rc;
    return cme_mlp_recast_w2(w2, p->w2c, p->ldh, p->w2t, H, C, stream);
}
