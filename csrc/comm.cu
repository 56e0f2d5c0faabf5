// Cross-GPU pieces over NVLink 5 / NVSwitch peer memory (no MPI, no NCCL here).
//
// The reference moves every minibatch through host memory: D2H of four gradient
// buffers, four MPI_Allreduce calls on host memory, SGD on the host, and a full
// H2D re-upload of the weights next step (ref fpcode/neural_network.cpp:484-541).
// Here the gradient all-reduce, the L2 term, the SGD step and the bf16 re-cast
// of the weights are ONE kernel that reads its slice of every rank's gradient
// buffer through peer-mapped pointers (or one multimem.ld_reduce through the
// NVSwitch multicast address) and writes the updated weights into every rank's
// replica.  Flags for ordering between ranks are plain uint32 slots in a
// peer-mapped pad, bumped with release atomics and polled with acquire loads.
#include "common.cuh"
#include "update_stream.cuh"

#include <algorithm>

namespace {

using upd::kMaxRanks;
using upd::ld_peer_f4;
using upd::pack2;
using upd::reg_for;

struct PeerPtrs {
    float* grads[kMaxRanks];
    float* params[kMaxRanks];
    __nv_bfloat16* w1c[kMaxRanks];
};

__global__ void __launch_bounds__(256)
allreduce_sgd_p2p_kernel(PeerPtrs pp, int rank, int nranks, int64_t begin, int64_t end,
                         int64_t n_w1, int ldx, int kreg, int64_t w2_begin, int64_t w2_end,
                         float lr, float reg) {
    const int64_t gstride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = begin + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < end;
         i += gstride) {
        // fixed summation order (rank 0, 1, ...) so every configuration is reproducible
        float4 s = ld_peer_f4(pp.grads[0] + i);
        for (int r = 1; r < nranks; ++r) {
            const float4 g = ld_peer_f4(pp.grads[r] + i);
            s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
        }
        const float4 p = *reinterpret_cast<const float4*>(pp.params[rank] + i);
        float4 q;
        q.x = p.x - lr * (s.x + reg_for(i, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.x);
        q.y = p.y - lr * (s.y + reg_for(i + 1, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.y);
        q.z = p.z - lr * (s.z + reg_for(i + 2, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.z);
        q.w = p.w - lr * (s.w + reg_for(i + 3, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.w);
        // The fp32 master of W1 is sharded: only its owner keeps it current (nobody else reads
        // it between steps).  What every rank needs is the bf16 compute copy of W1 and the small
        // fp32 tail (W2, b2), so only those cross NVLink.
        if (i < n_w1) {
            *reinterpret_cast<float4*>(pp.params[rank] + i) = q;
            const uint2 qb = make_uint2(pack2(q.x, q.y), pack2(q.z, q.w));
            for (int r = 0; r < nranks; ++r) *reinterpret_cast<uint2*>(pp.w1c[r] + i) = qb;
        } else {
            for (int r = 0; r < nranks; ++r) *reinterpret_cast<float4*>(pp.params[r] + i) = q;
        }
    }
}

// NVSwitch multicast variant: the switch performs the sum (multimem.ld_reduce)
// and replicates the stores (multimem.st) to every GPU bound to the address.
__global__ void __launch_bounds__(256)
allreduce_sgd_mc_kernel(const float* __restrict__ mc_grads, float* __restrict__ local_params,
                        float* __restrict__ mc_params, __nv_bfloat16* __restrict__ mc_w1c,
                        int64_t begin, int64_t end, int64_t n_w1, int ldx, int kreg,
                        int64_t w2_begin, int64_t w2_end, float lr, float reg) {
    const int64_t gstride = (int64_t)gridDim.x * blockDim.x * 4;
    for (int64_t i = begin + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < end;
         i += gstride) {
        float4 s;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(s.x), "=f"(s.y), "=f"(s.z), "=f"(s.w)
                     : "l"(mc_grads + i)
                     : "memory");
        const float4 p = *reinterpret_cast<const float4*>(local_params + i);
        float4 q;
        q.x = p.x - lr * (s.x + reg_for(i, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.x);
        q.y = p.y - lr * (s.y + reg_for(i + 1, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.y);
        q.z = p.z - lr * (s.z + reg_for(i + 2, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.z);
        q.w = p.w - lr * (s.w + reg_for(i + 3, n_w1, ldx, kreg, w2_begin, w2_end, reg) * p.w);
        if (i >= n_w1) {
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_params + i),
                         "f"(q.x), "f"(q.y), "f"(q.z), "f"(q.w)
                         : "memory");
        } else {
            *reinterpret_cast<float4*>(local_params + i) = q;   // sharded fp32 master, see above
            const float b0 = __uint_as_float(pack2(q.x, q.y)), b1 = __uint_as_float(pack2(q.z, q.w));
            asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc_w1c + i),
                         "f"(b0), "f"(b1)
                         : "memory");
        }
    }
}

// signal every peer, then wait until every peer has signalled `epoch` times.
// epoch == 0: use (and advance) the device-side call counter kept in slot 63 of
// this rank's own channel row, so the launch has no per-call host argument and
// can be captured in a CUDA graph.
constexpr unsigned long long kBarrierTimeoutNs = 30ull * 1000000000ull;

__global__ void peer_barrier_kernel(uint32_t* const* __restrict__ pads, int rank, int nranks,
                                    int channel, uint32_t epoch, int do_signal, int do_wait) {
    __shared__ uint32_t target;
    if (threadIdx.x == 0) {
        uint32_t e = epoch;
        if (e == 0) {
            uint32_t* counter = pads[rank] + channel * 64 + 63;
            e = *counter + 1;
            *counter = e;
        }
        target = e;
    }
    __syncthreads();
    const int r = threadIdx.x;
    if (r >= nranks) return;
    if (do_signal) {
        __threadfence_system();
        uint32_t* slot = pads[r] + channel * 64 + rank;
        asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(slot) : "memory");
    }
    if (do_wait) {
        const uint32_t* mine = pads[rank] + channel * 64 + r;
        const uint32_t want = target;
        // bounded: a peer that died would otherwise leave this GPU spinning for good.  After
        // kBarrierTimeoutNs the kernel traps, and the host sees a launch failure at its next
        // synchronisation instead of a hang.
        const unsigned long long t0 = cme::global_timer_ns();
        uint32_t v = cme::ld_relaxed_sys(mine);
        while ((int32_t)(v - want) < 0) {
            __nanosleep(100);
            v = cme::ld_relaxed_sys(mine);
            if (cme::global_timer_ns() - t0 > kBarrierTimeoutNs) {
                printf("cme_peer_barrier: rank %d waited %llu s for rank %d on channel %d (epoch %u, saw %u)\n",
                       rank, kBarrierTimeoutNs / 1000000000ull, r, channel, want, v);
                __trap();
            }
        }
        cme::fence_acquire_sys();
    }
}

__global__ void __launch_bounds__(256)
peer_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
    constexpr int U = 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n16; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = cme::ld_stream_u4(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) dst[i + u * stride] = v[u];
    }
    for (; i < n16; i += stride) dst[i] = cme::ld_stream_u4(src + i);
}

// The streaming update as a kernel of its own (plan->overlap == 0, tests): the same code the
// update warps of the dW1 GEMM run (update_stream.cuh), one CTA per worker group.
constexpr int kUpdateThreads = 256;

__global__ void __launch_bounds__(kUpdateThreads)
update_stream_kernel(const __grid_constant__ upd::UpdateArgs a) {
    __shared__ uint32_t s_ok;
    upd::run_update(a, blockIdx.x, gridDim.x, threadIdx.x, kUpdateThreads, 1, &s_ok);
}

// ---- fp64 parity engine: gradient all-reduce + L2 + SGD, handshake included ----------------
// One launch per step, no barrier launches around it.  Channel 6 of the pad carries "rank r's
// gradients of step e are complete", channel 7 "rank r has written its slice of step e to every
// replica"; slot 63 of channel 6 is this rank's step counter, slot 62 the CTA counter, slot 61
// the status word (0, or 1 + the rank that never showed up).  Every rank reduces 1/nranks of the
// vector in fixed rank order and stores the new parameters into every replica, so the replicas
// are bit-identical and a run is reproducible.
// Replaces D2H + MPI_Allreduce x4 + host SGD + H2D (ref fpcode/neural_network.cpp:496-541) at
// the reference's own precision.
struct PeerPtrsD {
    double* grads[kMaxRanks];
    double* params[kMaxRanks];
};
constexpr int kF64Ready = 6 * 64, kF64Done = 7 * 64;
constexpr int kF64Epoch = kF64Ready + 63, kF64Ctr = kF64Ready + 62, kF64Status = kF64Ready + 61;

__device__ __forceinline__ double2 ld_peer_d2(const double* p) {
    double2 r;
    asm volatile("ld.global.relaxed.sys.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p) : "memory");
    return r;
}

__global__ void __launch_bounds__(256)
allreduce_sgd_f64_kernel(PeerPtrsD pp, uint32_t* const* __restrict__ pads, int rank, int nranks, int64_t n,
                         int64_t w1_end, int64_t w2_begin, int64_t w2_end, double lr, double reg,
                         unsigned long long timeout_ns) {
    __shared__ uint32_t s_epoch, s_ok, s_last;
    uint32_t* pad = pads[rank];
    if (threadIdx.x == 0) {
        s_epoch = cme::ld_relaxed_gpu(pad + kF64Epoch) + 1;
        s_ok = 1;
        s_last = 0;
    }
    __syncthreads();
    const uint32_t e = s_epoch;
    if (blockIdx.x == 0 && threadIdx.x < nranks) {
        // the kernels that wrote this rank's gradients ran before this one in the stream
        __threadfence_system();
        cme::red_relaxed_sys_add(pads[threadIdx.x] + kF64Ready + rank, 1u);
    }
    if (threadIdx.x < nranks && !upd::wait_flag_sys(pad + kF64Ready + threadIdx.x, e, timeout_ns)) {
        pad[kF64Status] = 1u + threadIdx.x;
        s_ok = 0;
    }
    __syncthreads();
    if (s_ok) {
        const int64_t pairs = n / 2;
        const int64_t b = pairs * rank / nranks * 2, end = pairs * (rank + 1) / nranks * 2;
        const int64_t gstride = (int64_t)gridDim.x * blockDim.x * 2;
        for (int64_t i = b + ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2; i < end; i += gstride) {
            double2 s = ld_peer_d2(pp.grads[0] + i);
            for (int r = 1; r < nranks; ++r) {
                const double2 g = ld_peer_d2(pp.grads[r] + i);
                s.x += g.x;
                s.y += g.y;
            }
            const double2 p = *reinterpret_cast<const double2*>(pp.params[rank] + i);
            const double r0 = (i < w1_end || (i >= w2_begin && i < w2_end)) ? reg : 0.0;
            const double r1 = (i + 1 < w1_end || (i + 1 >= w2_begin && i + 1 < w2_end)) ? reg : 0.0;
            double2 q;
            q.x = p.x - lr * (s.x + r0 * p.x);
            q.y = p.y - lr * (s.y + r1 * p.y);
            for (int r = 0; r < nranks; ++r) *reinterpret_cast<double2*>(pp.params[r] + i) = q;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(pad + kF64Ctr, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    // last CTA of this rank: every slice of mine is written; tell every rank, and leave only when
    // every rank has said the same (the next kernel may read the parameters and overwrite the
    // gradients)
    if (threadIdx.x < nranks) {
        __threadfence_system();
        cme::red_relaxed_sys_add(pads[threadIdx.x] + kF64Done + rank, 1u);
        if (!upd::wait_flag_sys(pad + kF64Done + threadIdx.x, e, timeout_ns)) pad[kF64Status] = 1u + threadIdx.x;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        pad[kF64Ctr] = 0;
        pad[kF64Epoch] = e;
        __threadfence();
    }
}

int slice_bounds(int64_t r_begin, int64_t r_end, int rank, int nranks, int64_t* b, int64_t* e) {
    // the range [r_begin, r_end) is cut into nranks slices of whole 16-byte vectors
    const int64_t quads = (r_end - r_begin) / 4;
    *b = r_begin + quads * rank / nranks * 4;
    *e = r_begin + quads * (rank + 1) / nranks * 4;
    return 0;
}

}  // namespace

namespace cme {

int make_update_args(const CmeMlpPlan* p, float lr, float reg, int indirect, upd::UpdateArgs* out) {
    CME_CHECK_ARG(p && p->pads_dev && p->nranks >= 1 && p->nranks <= upd::kMaxRanks && p->rank >= 0 &&
                      p->rank < p->nranks, "cme_mlp_update_stream: bad plan");
    CME_CHECK_ARG(p->unit_rows == 128 || p->unit_rows == 256, "cme_mlp_update_stream: unit_rows %d", p->unit_rows);
    CME_CHECK_ARG(!(p->use_mc && p->nranks > 1) || (p->grads_mc && p->params_mc && p->w1c_mc),
                  "cme_mlp_update_stream: use_mc without multicast addresses");
    CME_CHECK_ARG(!indirect || p->step_args, "cme_mlp_update_stream: indirect arguments need plan->step_args");
    int64_t lay[4];
    int rc = cme_mlp_layout(p, lay);
    if (rc) return rc;
    upd::UpdateArgs a{};
    a.enabled = 1;
    a.use_mc = (p->use_mc && p->nranks > 1) ? 1 : 0;
    for (int r = 0; r < p->nranks; ++r) {
        a.grads[r] = p->grads[r];
        a.params_peer[r] = p->params[r];
        a.w1c[r] = (__nv_bfloat16*)p->w1c[r];
        CME_CHECK_ARG(a.grads[r] && a.params_peer[r] && a.w1c[r], "cme_mlp_update_stream: null peer %d", r);
    }
    a.mc_grads = p->grads_mc, a.mc_params = p->params_mc, a.mc_w1c = (__nv_bfloat16*)p->w1c_mc;
    a.pads = p->pads_dev;
    a.rank = p->rank, a.nranks = p->nranks;
    a.n_w1 = lay[0], a.n_total = lay[3];
    a.ldx = p->ldx, a.kreg = p->D;
    CME_CHECK_ARG(lay[3] < (1ll << 31), "cme_mlp_update_stream: %lld parameters; the update indexes with 31 bits", (long long)lay[3]);
    a.ldx_magic = (uint32_t)(((1ull << 32) + (uint64_t)p->ldx - 1) / (uint64_t)p->ldx);
    a.w2_begin = lay[1], a.w2_end = lay[1] + (int64_t)p->C * p->H;
    a.unit_elems = (int64_t)p->unit_rows * p->ldx;
    a.nunits = (p->Hp + p->unit_rows - 1) / p->unit_rows;
    CME_CHECK_ARG(a.nunits <= cme::kPadMaxUnits, "cme_mlp_update_stream: %d units", a.nunits);
    a.lr = lr, a.reg = reg;
    a.sa = indirect ? p->step_args : nullptr;
    a.timeout_ns = (unsigned long long)(p->timeout_ms > 0 ? p->timeout_ms : 20000) * 1000000ull;
    *out = a;
    return 0;
}

}  // namespace cme

extern "C" {

int cme_update_slice(int64_t range_begin, int64_t range_end, int rank, int nranks, int64_t out[2]) {
    CME_CHECK_ARG(out && nranks >= 1 && rank >= 0 && rank < nranks, "cme_update_slice: rank %d of %d", rank, nranks);
    CME_CHECK_ARG(0 <= range_begin && range_begin <= range_end && range_begin % 4 == 0 && range_end % 4 == 0,
                  "cme_update_slice: bad range [%lld, %lld)", (long long)range_begin, (long long)range_end);
    return slice_bounds(range_begin, range_end, rank, nranks, &out[0], &out[1]);
}

int cme_allreduce_sgd(float* const* grads_host, float* const* params_host, void* const* w1c_host,
                      int rank, int nranks, int64_t n_total, int64_t n_w1, int ldx, int kreg,
                      int64_t w2_begin, int64_t w2_end, float lr, float reg, int64_t range_begin,
                      int64_t range_end, void* stream) {
    CME_CHECK_ARG(grads_host && params_host && w1c_host, "cme_allreduce_sgd: null pointer table");
    CME_CHECK_ARG(nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks,
                  "cme_allreduce_sgd: rank %d of %d", rank, nranks);
    CME_CHECK_ARG(n_total % 4 == 0 && n_w1 % 4 == 0 && ldx % 4 == 0 && n_w1 <= n_total,
                  "cme_allreduce_sgd: sizes must be multiples of 4 floats");
    PeerPtrs pp;
    for (int r = 0; r < nranks; ++r) {
        pp.grads[r] = grads_host[r];
        pp.params[r] = params_host[r];
        pp.w1c[r] = (__nv_bfloat16*)w1c_host[r];
        CME_CHECK_ARG(pp.grads[r] && pp.params[r] && pp.w1c[r], "cme_allreduce_sgd: null peer %d", r);
    }
    CME_CHECK_ARG(0 <= range_begin && range_begin <= range_end && range_end <= n_total &&
                      range_begin % 4 == 0 && range_end % 4 == 0,
                  "allreduce_sgd: bad range [%lld, %lld)", (long long)range_begin, (long long)range_end);
    int64_t b, e;
    slice_bounds(range_begin, range_end, rank, nranks, &b, &e);
    if (e <= b) return 0;
    const int grid = cme::stride_grid((e - b) / 4, 256, 4, 2);
    allreduce_sgd_p2p_kernel<<<grid, 256, 0, cme::as_stream(stream)>>>(
        pp, rank, nranks, b, e, n_w1, ldx, kreg, w2_begin, w2_end, lr, reg);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_allreduce_sgd_mc(const float* mc_grads, float* local_params, float* mc_params,
                         void* mc_w1c, int rank, int nranks, int64_t n_total, int64_t n_w1, int ldx,
                         int kreg, int64_t w2_begin, int64_t w2_end, float lr, float reg,
                         int64_t range_begin, int64_t range_end, void* stream) {
    CME_CHECK_ARG(mc_grads && local_params && mc_params && mc_w1c, "cme_allreduce_sgd_mc: null pointer");
    CME_CHECK_ARG(nranks >= 1 && rank >= 0 && rank < nranks, "cme_allreduce_sgd_mc: rank %d of %d",
                  rank, nranks);
    CME_CHECK_ARG(n_total % 4 == 0 && n_w1 % 4 == 0 && ldx % 4 == 0 && n_w1 <= n_total,
                  "cme_allreduce_sgd_mc: sizes must be multiples of 4 floats");
    CME_CHECK_ARG(0 <= range_begin && range_begin <= range_end && range_end <= n_total &&
                      range_begin % 4 == 0 && range_end % 4 == 0,
                  "allreduce_sgd: bad range [%lld, %lld)", (long long)range_begin, (long long)range_end);
    int64_t b, e;
    slice_bounds(range_begin, range_end, rank, nranks, &b, &e);
    if (e <= b) return 0;
    const int grid = cme::stride_grid((e - b) / 4, 256, 4, 2);
    allreduce_sgd_mc_kernel<<<grid, 256, 0, cme::as_stream(stream)>>>(
        mc_grads, local_params, mc_params, (__nv_bfloat16*)mc_w1c, b, e, n_w1, ldx, kreg, w2_begin,
        w2_end, lr, reg);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_allreduce_sgd_f64(double* const* grads_host, double* const* params_host, uint32_t* const* pads_dev,
                          int rank, int nranks, int64_t n, int64_t w1_end, int64_t w2_begin, int64_t w2_end,
                          double lr, double reg, int timeout_ms, void* stream) {
    CME_CHECK_ARG(grads_host && params_host && pads_dev, "cme_allreduce_sgd_f64: null pointer table");
    CME_CHECK_ARG(nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks,
                  "cme_allreduce_sgd_f64: rank %d of %d", rank, nranks);
    CME_CHECK_ARG(n > 0 && n % 2 == 0, "cme_allreduce_sgd_f64: the vector must hold an even number of doubles (n = %lld)",
                  (long long)n);
    CME_CHECK_ARG(0 <= w1_end && w1_end <= w2_begin && w2_begin <= w2_end && w2_end <= n,
                  "cme_allreduce_sgd_f64: bad segment bounds");
    PeerPtrsD pp;
    for (int r = 0; r < nranks; ++r) {
        pp.grads[r] = grads_host[r];
        pp.params[r] = params_host[r];
        CME_CHECK_ARG(pp.grads[r] && pp.params[r], "cme_allreduce_sgd_f64: null peer %d", r);
        CME_CHECK_ARG(((uintptr_t)pp.grads[r] | (uintptr_t)pp.params[r]) % 16 == 0,
                      "cme_allreduce_sgd_f64: buffers of rank %d are not 16-byte aligned", r);
    }
    const int64_t slice_pairs = std::max<int64_t>(1, n / 2 / nranks);
    const int grid = (int)std::min<int64_t>(cme::sm_count(), (slice_pairs + 255) / 256);
    allreduce_sgd_f64_kernel<<<grid, 256, 0, cme::as_stream(stream)>>>(
        pp, pads_dev, rank, nranks, n, w1_end, w2_begin, w2_end, lr, reg,
        (unsigned long long)(timeout_ms > 0 ? timeout_ms : 20000) * 1000000ull);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_update_stream(const CmeMlpPlan* p, float lr, float reg, void* stream) {
    return cme_mlp_update_stream_ind(p, lr, reg, 0, stream);
}

namespace {
__global__ void set_step_args_kernel(CmeStepArgs* dst, CmeStepArgs v) { *dst = v; }
}  // namespace

int cme_mlp_set_step_args(const CmeMlpPlan* p, const uint8_t* x_u8, const int32_t* labels, float lr, float reg,
                          float scale, void* stream) {
    CME_CHECK_ARG(p && p->step_args && x_u8 && labels, "cme_mlp_set_step_args: the plan has no step_args (or null inputs)");
    CmeStepArgs v{x_u8, labels, lr, reg, scale, 0};
    set_step_args_kernel<<<1, 1, 0, cme::as_stream(stream)>>>(p->step_args, v);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_update_stream_ind(const CmeMlpPlan* p, float lr, float reg, int indirect, void* stream) {
    upd::UpdateArgs a;
    int rc = cme::make_update_args(p, lr, reg, indirect, &a);
    if (rc) return rc;
    // a slice is unit_elems / nranks floats; small CTAs (256 threads, no shared memory), every one
    // taking its share of every unit
    const int64_t slice_quads = std::max<int64_t>(1, a.unit_elems / 4 / a.nranks);
    const int grid = (int)std::min<int64_t>(cme::sm_count(), (slice_quads + kUpdateThreads - 1) / kUpdateThreads);
    update_stream_kernel<<<grid, kUpdateThreads, 0, cme::as_stream(stream)>>>(a);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_status(const CmeMlpPlan* p, uint32_t* out) {
    CME_CHECK_ARG(p && out && p->pad_local, "cme_mlp_status: bad argument");
    CME_CUDA(cudaMemcpy(out, p->pad_local + cme::kPadStatus, sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return 0;
}

int cme_peer_barrier(uint32_t* const* pads, int rank, int nranks, int channel, uint32_t epoch,
                     int do_signal, int do_wait, void* stream) {
    CME_CHECK_ARG(pads && nranks >= 1 && nranks <= 32 && rank >= 0 && rank < nranks,
                  "cme_peer_barrier: rank %d of %d", rank, nranks);
    CME_CHECK_ARG(channel >= 0 && channel < 8, "cme_peer_barrier: channel %d not in 0..7", channel);
    peer_barrier_kernel<<<1, 32, 0, cme::as_stream(stream)>>>(pads, rank, nranks, channel, epoch,
                                                             do_signal, do_wait);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_peer_copy(const void* src, void* dst, size_t bytes, void* stream) {
    return cme_peer_copy_bg(src, dst, bytes, 0, stream);
}

/* The same transfer on a copy engine instead of SMs (plain cudaMemcpyAsync, device to device through
 * the peer mapping): for copies that must make progress WHILE persistent GEMM kernels hold every
 * SM's registers (the owner-mode prefetch: a copy kernel only got scheduled between those kernels). */
int cme_peer_memcpy(const void* src, void* dst, size_t bytes, void* stream) {
    CME_CHECK_ARG(src && dst, "cme_peer_memcpy: null pointer");
    if (bytes == 0) return 0;
    CME_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, cme::as_stream(stream)));
    return 0;
}

/* max_blocks > 0 caps the grid: a copy that runs next to compute kernels (the owner-mode prefetch)
 * is bound by NVLink, not by SMs; 64 blocks keep ~1 MB in flight and leave the SMs to the GEMMs. */
int cme_peer_copy_bg(const void* src, void* dst, size_t bytes, int max_blocks, void* stream) {
    CME_CHECK_ARG(src && dst && bytes % 16 == 0, "cme_peer_copy: 16-byte multiples only");
    if (bytes == 0) return 0;
    const size_t n16 = bytes / 16;
    int grid = cme::stride_grid((int64_t)n16, 256, 8, 2);
    if (max_blocks > 0 && grid > max_blocks) grid = max_blocks;
    peer_copy_kernel<<<grid, 256, 0, cme::as_stream(stream)>>>((const uint4*)src, (uint4*)dst, n16);
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
