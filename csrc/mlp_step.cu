// Native training-step executor for the bf16 engine.
//
// The reference drives one minibatch from C++ (parallel_train's loop body, ref
// fpcode/neural_network.cpp:484-541: 6 H2D copies, 13 synchronised launches, 4 D2H copies,
// 4 host all-reduces, host SGD).  cme_mlp_step is the same unit of work for this library's
// C/C++ users: given a plan with every device buffer of one rank it issues the whole step
// (staging, forward, backward, streaming all-reduce + SGD, re-cast) with no host
// synchronisation.  The Python engine issues its steps through this function too.
//
// Schedule:
//   main   staging, forward GEMM + sigmoid with the layer-2 logits riding on its staged tiles,
//          per-row softmax + CE tail, backward tail (dZ1 + dW2)
//   main   dW1 GEMM + update in ONE kernel: the GEMM publishes every unit (256 rows of W1) of the
//          gradient as it completes, and four extra warps in each of its CTAs reduce the unit
//          over NVLink as soon as every rank has published it, apply SGD and replicate the weights
//   main   re-cast W2
// Nothing zero-fills a gradient (the first k-chunk of every dW1 tile stores; dW2 is added up from
// per-CTA slots in CTA order; db2 and the loss from per-block partials), so the step is
// bit-reproducible; shapes the one-kernel forward does not take fall back to atomics on a zeroed
// tail.  There are no
// barrier launches: the flags live inside the kernel.  With plan->overlap == 0 the update runs
// as a kernel of its own behind the GEMM; the arithmetic and the owner of every slice of the
// sharded fp32 master are the same either way.
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

int cme_mlp_unit_slice(const CmeMlpPlan* p, int unit, int rank, int64_t out[2]) {
    CME_CHECK_ARG(p && out && p->nranks >= 1 && rank >= 0 && rank < p->nranks, "cme_mlp_unit_slice: rank %d", rank);
    CME_CHECK_ARG(p->unit_rows == 128 || p->unit_rows == 256, "cme_mlp_unit_slice: unit_rows %d", p->unit_rows);
    const Geometry g = geometry(p);
    const int nunits = (p->Hp + p->unit_rows - 1) / p->unit_rows;
    CME_CHECK_ARG(unit >= 0 && unit <= nunits, "cme_mlp_unit_slice: unit %d not in 0..%d", unit, nunits);
    const int64_t ue = (int64_t)p->unit_rows * p->ldx;
    const int64_t r0 = unit == nunits ? g.n_w1 : (int64_t)unit * ue;
    const int64_t r1 = unit == nunits ? g.n_total : ((int64_t)(unit + 1) * ue < g.n_w1 ? (int64_t)(unit + 1) * ue : g.n_w1);
    const int64_t quads = (r1 - r0) / 4;
    out[0] = r0 + quads * rank / p->nranks * 4;
    out[1] = r0 + quads * (rank + 1) / p->nranks * 4;
    return 0;
}

int cme_mlp_step_ex(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels, int64_t rows,
                    float lr, float reg, float scale, float* loss_sum, void* stream, void* side_stream,
                    void* const* events) {
    CME_CHECK_ARG(p && ((x_u8 && labels) || (!x_u8 && p->step_args)),
                  "cme_mlp_step: null inputs (and no plan->step_args to read them from)");
    const CmeStepArgs* sa = x_u8 ? nullptr : p->step_args;
    CME_CHECK_ARG(rows > 0 && rows <= p->max_rows, "cme_mlp_step: rows %lld not in 1..%lld",
                  (long long)rows, (long long)p->max_rows);
    CME_CHECK_ARG(p->nranks >= 1 && p->nranks <= 16 && p->rank >= 0 && p->rank < p->nranks,
                  "cme_mlp_step: rank %d of %d", p->rank, p->nranks);
    CME_CHECK_ARG(p->pads_dev && p->pad_local && p->sync_scratch,
                  "cme_mlp_step: the plan needs pads_dev, pad_local and sync_scratch (one rank too)");
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

    const bool fused_fwd = p->fwd_scratch && cme_mlp_fwd_fused_ok(rows, g.H);
    // With the one-kernel forward (db2 stored by its tail) and the slot-wise dW2 every gradient is
    // STORED in a fixed order: nothing to zero, nothing on the side stream.  Otherwise dW2 / db2
    // are accumulated with atomics and the tail of the gradient vector is zeroed first.
    const bool need_zero = !(fused_fwd && p->dw2_scratch);
    if (need_zero) {
        if (side_stream) {
            CME_CUDA(cudaEventRecord(ev(0), main_s));
            CME_CUDA(cudaStreamWaitEvent(side_s, ev(0), 0));
        }
        CME_CUDA(cudaMemsetAsync(grads + g.o_w2, 0, sizeof(float) * (size_t)(g.n_total - g.o_w2), side_s));
        if (side_stream) CME_CUDA(cudaEventRecord(ev(1), side_s));
    }
    // staging, and in the same launch the bf16 copies of the W2 the previous step's update left
    if ((rc = cme_mlp_stage_recast(x_u8, sa, p->xb, rows, g.D, g.ldx, params + g.o_w2, p->w2c, g.ldh, p->w2t, g.H, g.C,
                                   stream))) return rc;
    if (fused_fwd) {
        // forward pass in one kernel: the layer-2 logits ride on the A1 tiles while they are staged
        if (need_zero && side_stream) CME_CUDA(cudaStreamWaitEvent(main_s, ev(1), 0));
        if ((rc = cme_mlp_fwd_fused(p->xb, p->w1c[p->rank], p->w2c, params + g.o_b2, labels, sa, p->a1, p->dz2,
                                    p->dz2b, loss_sum, p->pred, grads + g.o_b2, p->fwd_scratch, p->max_rows, scale,
                                    rows, g.K, g.H, g.ldx, g.ldh, g.C, stream))) return rc;
    } else {
        if ((rc = cme_mlp_fwd1(p->xb, p->w1c[p->rank], nullptr, p->a1, rows, g.K, g.H, g.ldx, g.ldh, stream))) return rc;
        if (side_stream) CME_CUDA(cudaStreamWaitEvent(main_s, ev(1), 0));
        if ((rc = cme_mlp_fwd2_tc_ind(p->a1, p->w2c, params + g.o_b2, labels, sa, p->dz2, p->dz2b, loss_sum, p->pred,
                                      grads + g.o_b2, scale, rows, g.H, g.ldh, g.C, stream))) return rc;
    }
    // the slot-wise (bit-reproducible) dW2 costs one more small launch: only where the rest of the
    // step is reproducible too, i.e. with the one-kernel forward
    if ((rc = cme_mlp_bwd_tail_det(p->a1, p->dz2b, p->w2t, p->dz1, grads + g.o_w2, g.H, rows, g.H, g.ldh, g.C,
                                   need_zero ? nullptr : p->dw2_scratch, stream))) return rc;
    if (p->overlap) {
        // ONE kernel: the dW1 GEMM with the update warps inside it
        if ((rc = cme_mlp_dw1_update(p, rows, lr, reg, sa != nullptr, stream))) return rc;
    } else {
        if ((rc = cme_mlp_dw1_sync(p->dz1, p->xb, grads, rows, g.K, g.H, g.ldx, g.ldh, g.ldx, p->sync_scratch,
                                   p->pads_dev, p->nranks, p->unit_rows, p->pad_local + cme::kPadStatus, stream))) return rc;
        if ((rc = cme_mlp_update_stream_ind(p, lr, reg, sa != nullptr, stream))) return rc;
    }
    return 0;
}

int cme_mlp_step(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels, int64_t rows, float lr,
                 float reg, float scale, float* loss_sum, void* stream) {
    return cme_mlp_step_ex(p, x_u8, labels, rows, lr, reg, scale, loss_sum, stream, nullptr, nullptr);
}

int cme_mlp_plan_init_sync(CmeMlpPlan* p) {
    CME_CHECK_ARG(p && p->nranks == 1 && p->rank == 0, "cme_mlp_plan_init_sync: one-rank plans only "
                  "(several ranks need peer-mapped pads, see cme213_sp18_b200/parallel/dist.py)");
    CME_CHECK_ARG(p->Hp > 0 && p->Hp % 64 == 0, "cme_mlp_plan_init_sync: Hp %d", p->Hp);
    uint32_t* pad = nullptr;
    uint32_t** table = nullptr;
    void* scratch = nullptr;
    CmeStepArgs* sa = nullptr;
    CME_CUDA(cudaMalloc(&pad, sizeof(uint32_t) * cme::kPadWords));
    CME_CUDA(cudaMemset(pad, 0, sizeof(uint32_t) * cme::kPadWords));
    CME_CUDA(cudaMalloc(&table, sizeof(uint32_t*)));
    CME_CUDA(cudaMemcpy(table, &pad, sizeof(uint32_t*), cudaMemcpyHostToDevice));
    CME_CUDA(cudaMalloc(&scratch, cme_mlp_dw1_scratch_bytes()));
    CME_CUDA(cudaMemset(scratch, 0, cme_mlp_dw1_scratch_bytes()));
    CME_CUDA(cudaMalloc(&sa, sizeof(CmeStepArgs)));
    CME_CUDA(cudaMemset(sa, 0, sizeof(CmeStepArgs)));
    p->pad_local = pad;
    p->pads_dev = table;
    p->sync_scratch = scratch;
    p->step_args = sa;
    p->fwd_scratch = nullptr;
    if (p->Hp % 256 == 0 && p->max_rows >= 256) {
        const size_t n = cme_mlp_fwd_scratch_floats(p->max_rows);
        CME_CUDA(cudaMalloc(&p->fwd_scratch, n * sizeof(float)));
        CME_CUDA(cudaMemset(p->fwd_scratch, 0, n * sizeof(float)));
    }
    {
        const size_t n = cme_mlp_bwd_tail_scratch_floats(p->H);
        CME_CUDA(cudaMalloc(&p->dw2_scratch, n * sizeof(float)));
        CME_CUDA(cudaMemset(p->dw2_scratch, 0, n * sizeof(float)));
    }
    p->unit_rows = p->Hp >= 512 ? 256 : 128;
    p->overlap = 1;
    return 0;
}

int cme_mlp_plan_free_sync(CmeMlpPlan* p) {
    if (!p) return 0;
    if (p->pad_local) cudaFree(p->pad_local);
    if (p->pads_dev) cudaFree((void*)p->pads_dev);
    if (p->sync_scratch) cudaFree(p->sync_scratch);
    if (p->step_args) cudaFree(p->step_args);
    if (p->fwd_scratch) cudaFree(p->fwd_scratch);
    if (p->dw2_scratch) cudaFree(p->dw2_scratch);
    p->fwd_scratch = nullptr, p->dw2_scratch = nullptr;
    p->pad_local = nullptr, p->pads_dev = nullptr, p->sync_scratch = nullptr, p->step_args = nullptr;
    return 0;
}

int cme_mlp_step_launches(const CmeMlpPlan* p) {
    // staging (+ re-cast of W2), forward GEMM, layer 2 (or its per-row tail), backward tail, dW1 (+ update)
    return p ? (p->overlap ? 5 : 6) + (p->dw2_scratch ? 1 : 0) : 0;
}

}  // extern "C"
