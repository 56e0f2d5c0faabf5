// Streaming update: gradient all-reduce + L2 + SGD + bf16 re-cast, unit by unit.
//
// The flat gradient vector is cut into units (unit_rows rows of W1 each, plus the W2/b2 tail).
// The dW1 GEMM bumps a unit's flag on every rank's pad when the last k-chunk of the unit's last
// tile has landed (gemm_tc.cu, dw_publish).  The code below runs in "update warps": 128 extra
// threads inside every CTA of that same GEMM (the product path: ONE kernel computes dW1 on the
// tensor cores and reduces it across GPUs while the MMAs of later tiles are still running), or
// the threads of a kernel of its own (update_stream_kernel in comm.cu: plan->overlap == 0).
// Workers wait for "all ranks have bumped unit u", then reduce this rank's 1/nranks slice of the
// unit over NVLink (multimem.ld_reduce through the switch, or peer loads in fixed rank order),
// apply p -= lr * (g + reg * p), keep the fp32 master of the slice, and replicate the bf16
// compute copy (and the fp32 tail) to every rank (multimem.st or peer stores).  There are no
// barrier launches: worker 0 signals that this rank's tail gradients are ready, and the last
// worker to finish tells every rank "my slices are written", waits until every rank has said
// so, and closes the step; whatever follows in the stream may then read the new weights and
// overwrite the gradients.  Every cross-GPU wait is bounded (timeout_ns): a wait that expires
// records a code in the pad's status word, the workers drop the rest of their work, and the
// host reports it (cme_mlp_status).
// Replaces D2H x4 + MPI_Allreduce x4 + host SGD + H2D of the weights
// (ref fpcode/neural_network.cpp:496-541).
#pragma once

#include "common.cuh"

namespace upd {

constexpr int kMaxRanks = 16;

struct UpdateArgs {
    int enabled;                  // 0: the update warps of a GEMM have nothing to do
    int use_mc;                   // multimem instructions on the multicast mappings below
    float* grads[kMaxRanks];
    float* params_peer[kMaxRanks];
    __nv_bfloat16* w1c[kMaxRanks];
    const float* mc_grads;
    float* mc_params;
    __nv_bfloat16* mc_w1c;
    uint32_t* const* pads;        // device array of nranks pad pointers
    int rank, nranks;
    int64_t n_w1, n_total;
    int ldx, kreg;
    uint32_t ldx_magic;           // ceil(2^32 / ldx), see sgd4
    int64_t w2_begin, w2_end;
    int64_t unit_elems;
    int nunits;
    float lr, reg;
    unsigned long long timeout_ns;
    const CmeStepArgs* sa;        // non-null: lr and reg come from the device-side step arguments
};

__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
    float4 r;
    // peer data changes every step: never serve it from a stale L1 line
    asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// reg applies to W1 entries (column < kreg of each ldx-long row) and to the W2
// block [w2_begin, w2_end); biases and pad entries get none
__device__ __forceinline__ float reg_for(int64_t i, int64_t n_w1, int ldx, int kreg,
                                         int64_t w2_begin, int64_t w2_end, float reg) {
    if (i < n_w1) return ((int)(i % ldx) < kreg) ? reg : 0.f;
    return (i >= w2_begin && i < w2_end) ? reg : 0.f;
}

__device__ __forceinline__ bool wait_flag_sys(const uint32_t* p, uint32_t want, unsigned long long timeout_ns) {
    bool ok = (int32_t)(cme::ld_relaxed_sys(p) - want) >= 0;
    if (!ok) {
        const unsigned long long t0 = cme::global_timer_ns();
        while (!ok) {
#pragma unroll 1
            for (int i = 0; i < 16 && !ok; ++i) {
                __nanosleep(64);
                ok = (int32_t)(cme::ld_relaxed_sys(p) - want) >= 0;
            }
            if (!ok && cme::global_timer_ns() - t0 > timeout_ns) return false;
        }
    }
    // No acquire fence here, on purpose.  A sys-scope fence costs 3-6 us on this machine while
    // stores are in flight (ncu: the update's threads spent their time behind thread 0's MEMBAR.SYS),
    // and nothing it would order can go wrong: the producer made its data visible at system scope
    // before it bumped the flag (fence + red.release.sys), the flag is observed by a load that
    // completes before the group's barrier, and every load of the data is issued after that
    // barrier as a sys-scope / multimem access that is served by the owning GPU's L2, never by a
    // cache on this side.
    return true;
}

// p - lr * (s + reg * p) for one aligned vector of four; rows are ldx (a multiple of 4) long, so
// the four share a row and ONE column computation serves them.  The column comes from a
// multiply-high with ldx_magic = ceil(2^32 / ldx) plus a fix-up (the flat vector is < 2^31 long,
// checked on the host): a 64-bit remainder per element was most of this kernel's instructions.
__device__ __forceinline__ float4 sgd4(const float4& p, const float4& s, int64_t i, const UpdateArgs& a,
                                       float lr, float reg) {
    float r0, r1, r2, r3;
    if (i < a.n_w1) {
        const uint32_t u = (uint32_t)i;
        uint32_t q = __umulhi(u, a.ldx_magic);
        int col = (int)(u - q * (uint32_t)a.ldx);
        if (col < 0) col += a.ldx;
        if (col >= a.ldx) col -= a.ldx;
        r0 = col < a.kreg ? reg : 0.f;
        r1 = col + 1 < a.kreg ? reg : 0.f;
        r2 = col + 2 < a.kreg ? reg : 0.f;
        r3 = col + 3 < a.kreg ? reg : 0.f;
    } else {
        r0 = (i >= a.w2_begin && i < a.w2_end) ? reg : 0.f;
        r1 = (i + 1 >= a.w2_begin && i + 1 < a.w2_end) ? reg : 0.f;
        r2 = (i + 2 >= a.w2_begin && i + 2 < a.w2_end) ? reg : 0.f;
        r3 = (i + 3 >= a.w2_begin && i + 3 < a.w2_end) ? reg : 0.f;
    }
    float4 q4;
    q4.x = p.x - lr * (s.x + r0 * p.x);
    q4.y = p.y - lr * (s.y + r1 * p.y);
    q4.z = p.z - lr * (s.z + r2 * p.z);
    q4.w = p.w - lr * (s.w + r3 * p.w);
    return q4;
}

template <bool MC>
__device__ __forceinline__ float4 gather4(const UpdateArgs& a, int64_t i) {
    float4 s;
    if (MC) {
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(s.x), "=f"(s.y), "=f"(s.z), "=f"(s.w)
                     : "l"(a.mc_grads + i)
                     : "memory");
    } else {
        s = ld_peer_f4(a.grads[0] + i);          // fixed summation order: rank 0, 1, ...
        for (int r = 1; r < a.nranks; ++r) {
            const float4 g = ld_peer_f4(a.grads[r] + i);
            s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
        }
    }
    return s;
}

template <bool MC>
__device__ __forceinline__ void scatter4(const UpdateArgs& a, int64_t i, const float4& q) {
    float* mine = a.params_peer[a.rank];
    if (i < a.n_w1) {
        // The fp32 master of W1 is sharded: only the slice's owner keeps it (nobody else reads it
        // between steps).  What every rank needs is the bf16 compute copy of W1 and the small
        // fp32 tail (W2, b2), so only those cross NVLink.
        *reinterpret_cast<float4*>(mine + i) = q;
        const uint32_t lo = pack2(q.x, q.y), hi = pack2(q.z, q.w);
        if (MC) {
            asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(a.mc_w1c + i),
                         "f"(__uint_as_float(lo)), "f"(__uint_as_float(hi))
                         : "memory");
        } else {
            for (int r = 0; r < a.nranks; ++r) *reinterpret_cast<uint2*>(a.w1c[r] + i) = make_uint2(lo, hi);
        }
    } else if (MC) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a.mc_params + i),
                     "f"(q.x), "f"(q.y), "f"(q.z), "f"(q.w)
                     : "memory");
    } else {
        for (int r = 0; r < a.nranks; ++r) *reinterpret_cast<float4*>(a.params_peer[r] + i) = q;
    }
}

// Up to kBatch consecutive units are reduced together: every thread first issues the gathers of
// all its vectors in the batch (one NVLink round trip for the lot), then the SGD, then the stores.
constexpr int kBatch = 4;

template <bool MC>
__device__ __forceinline__ void reduce_batch(const UpdateArgs& a, const int64_t (&b)[kBatch], const int64_t (&e)[kBatch],
                                             int n, int64_t first, int64_t step, float lr, float reg) {
    const float* mine = a.params_peer[a.rank];
    for (int64_t off = first;; off += step) {
        float4 s[kBatch], p[kBatch];
        bool any = false;
#pragma unroll
        for (int j = 0; j < kBatch; ++j)
            if (j < n && b[j] + off < e[j]) {
                s[j] = gather4<MC>(a, b[j] + off);
                any = true;
            }
        if (!any) break;
#pragma unroll
        for (int j = 0; j < kBatch; ++j)
            if (j < n && b[j] + off < e[j]) p[j] = *reinterpret_cast<const float4*>(mine + b[j] + off);
#pragma unroll
        for (int j = 0; j < kBatch; ++j)
            if (j < n && b[j] + off < e[j]) scatter4<MC>(a, b[j] + off, sgd4(p[j], s[j], b[j] + off, a, lr, reg));
    }
}

// The whole update of one step, run by `nworkers` groups of `nthreads` threads (a group = the
// update warps of one CTA, or one CTA of the standalone kernel); tid is the thread's index in its
// group, bar_id a named barrier private to the group, s_ok a shared-memory word of the group.
// Every group takes the same share of every unit's slice, in unit order (tail first), so the
// reduction of a unit starts on all SMs the moment its flag is complete.
__device__ __forceinline__ void run_update(const UpdateArgs& a, int worker, int nworkers, int tid, int nthreads,
                                           int bar_id, uint32_t* s_ok) {
    uint32_t* pad = a.pads[a.rank];
    const uint32_t epoch = *reinterpret_cast<volatile uint32_t*>(pad + cme::kPadEpoch) + 1u;
    const float lr = a.sa ? a.sa->lr : a.lr, reg = a.sa ? a.sa->reg : a.reg;
    bool aborted = false;

    // stream order put this kernel after the backward tail: my W2 / b2 gradients are complete
    // (one rank: nobody on the other side of a link, no fence to pay for; the flag still counts)
    if (worker == 0 && tid < a.nranks) {
        if (a.nranks > 1) __threadfence_system();
        cme::red_relaxed_sys_add(a.pads[tid] + cme::kPadTail, 1u);
    }
    const int64_t first = ((int64_t)worker * nthreads + tid) * 4, step = (int64_t)nworkers * nthreads * 4;
    // position 0 is the tail (ready first), position u + 1 is unit u
    for (int pos0 = 0; pos0 <= a.nunits && !aborted; pos0 += kBatch) {
        const int n = min(kBatch, a.nunits + 1 - pos0);
        int64_t b[kBatch], e[kBatch];
        bool mine_any = false;
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const int pos = pos0 + j;
            const int64_t r0 = pos == 0 ? a.n_w1 : (int64_t)(pos - 1) * a.unit_elems;
            const int64_t r1 = pos == 0 ? a.n_total : min((int64_t)pos * a.unit_elems, a.n_w1);
            // this rank's slice of the unit
            const int64_t quads = (r1 - r0) / 4;
            b[j] = r0 + quads * a.rank / a.nranks * 4;
            e[j] = j < n ? r0 + quads * (a.rank + 1) / a.nranks * 4 : b[j];
            mine_any = mine_any || (b[j] + (int64_t)worker * nthreads * 4 < e[j]);
        }
        if (!mine_any) continue;                                   // nothing of these units for my group
        if (tid == 0) {
            bool ok = true;
            for (int j = 0; j < n && ok; ++j) {
                const int pos = pos0 + j;
                const int flag = pos == 0 ? cme::kPadTail : cme::kPadUnit0 + pos - 1;
                ok = wait_flag_sys(pad + flag, epoch * (uint32_t)a.nranks, a.timeout_ns);
                if (!ok) atomicCAS(pad + cme::kPadStatus, 0u, 1u | ((uint32_t)pos << 8));
            }
            *s_ok = ok;
        }
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(nthreads) : "memory");
        const bool ok = *reinterpret_cast<volatile uint32_t*>(s_ok) != 0;
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(nthreads) : "memory");
        if (!ok) { aborted = true; break; }
        if (a.use_mc) reduce_batch<true>(a, b, e, n, first, step, lr, reg);
        else reduce_batch<false>(a, b, e, n, first, step, lr, reg);
    }

    // every group counts itself out; the last one tells every rank "my slices are written",
    // waits until every rank has said so, and closes the step
    asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(nthreads) : "memory");
    if (tid < 32) {
        uint32_t last = 0;
        if (tid == 0) {
            // one rank: the end of the kernel orders these stores for whatever runs next, and a
            // system-scope fence costs microseconds (most of this kernel's time on a small model)
            if (a.nranks > 1) __threadfence_system();
            last = atomicAdd(pad + cme::kPadDoneCtr, 1u) == (uint32_t)nworkers - 1 ? 1u : 0u;
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last && a.nranks == 1) {
            if (tid == 0) {
                pad[cme::kPadDoneCtr] = 0;
                pad[cme::kPadDone0] += 1u;
                *reinterpret_cast<volatile uint32_t*>(pad + cme::kPadEpoch) = epoch;
            }
        } else if (last) {
            if (tid == 0) pad[cme::kPadDoneCtr] = 0;
            // every group fenced its stores at system scope before it counted itself out, and this
            // thread has seen all the counts: one more fence orders those observations before the bumps
            __threadfence_system();
            if (tid < a.nranks) cme::red_relaxed_sys_add(a.pads[tid] + cme::kPadDone0 + a.rank, 1u);
            bool ok = true;
            const bool skip = aborted || *reinterpret_cast<volatile uint32_t*>(pad + cme::kPadStatus) != 0;
            if (tid < a.nranks && !skip) ok = wait_flag_sys(pad + cme::kPadDone0 + tid, epoch, a.timeout_ns);
            if (!ok) atomicCAS(pad + cme::kPadStatus, 0u, 2u | ((uint32_t)tid << 8));
            __syncwarp();
            if (tid == 0) {
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(pad + cme::kPadEpoch) = epoch;
            }
        }
    }
}

}  // namespace upd

namespace cme {
// host: the update arguments of one step of `plan` (comm.cu)
int make_update_args(const CmeMlpPlan* plan, float lr, float reg, int indirect, upd::UpdateArgs* out);
}  // namespace cme
