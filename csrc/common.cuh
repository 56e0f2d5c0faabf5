// Shared host/device helpers for the cme213b device library (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "cme213b.h"

namespace cme {

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...);

#define CME_CHECK_ARG(cond, ...)              \
    do {                                      \
        if (!(cond)) {                        \
            ::cme::set_error(__VA_ARGS__);    \
            return -1;                        \
        }                                     \
    } while (0)

#define CME_CUDA(call)                                                                  \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            ::cme::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,               \
                             cudaGetErrorString(e__));                                  \
            return (int)e__;                                                            \
        }                                                                               \
    } while (0)

#define CME_LAUNCH_CHECK()                                                              \
    do {                                                                                \
        cudaError_t e__ = cudaGetLastError();                                           \
        if (e__ != cudaSuccess) {                                                       \
            ::cme::set_error("%s:%d launch -> %s", __FILE__, __LINE__,                  \
                             cudaGetErrorString(e__));                                  \
            return (int)e__;                                                            \
        }                                                                               \
    } while (0)

// Cached per-process device facts (queried once for the current device).
int sm_count();
// cuTensorMapEncodeTiled from the driver (no link-time dependency on libcuda); null if absent.
void* tensor_map_encoder();

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Grid for a grid-stride kernel: enough blocks to cover `work_items` but never
// more than `waves` full waves of the machine (148 SMs x blocks_per_sm).
static inline int stride_grid(int64_t work_items, int block, int blocks_per_sm, int waves = 4) {
    int64_t need = ceil_div64(work_items, block);
    int64_t cap = (int64_t)sm_count() * blocks_per_sm * waves;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

// ---- flag pad of the MLP step (one pad of kPadWords uint32 per rank, peer-mapped) ----------
// Words [0, 512) belong to cme_peer_barrier (8 channels x 64).  The streaming update uses:
//   kPadUnit0 + u   cumulative count of ranks whose gradient unit u is complete (u < kPadMaxUnits;
//                   kPadTail for the W2/b2 tail), bumped by every rank on every rank's pad
//   kPadDone0 + r   steps whose update rank r has finished (bumped by r on every pad)
//   kPadEpoch       local: steps this rank has finished     kPadDoneCtr   local scratch counter
//   kPadStatus      local: 0, or the code of the wait that timed out (see cme_mlp_status)
constexpr int kPadWords = 4096;
constexpr int kPadUnit0 = 2048;
constexpr int kPadMaxUnits = 1023;
constexpr int kPadTail = kPadUnit0 + kPadMaxUnits;
constexpr int kPadDone0 = 3072;
constexpr int kPadEpoch = 3104;
constexpr int kPadDoneCtr = 3105;
constexpr int kPadStatus = 3106;

// ---- device helpers ---------------------------------------------------------
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Polling loads are RELAXED: an acquire load is a load plus a fence, and a fence of gpu or sys
// scope is expensive here (it invalidates the SM's L1 and waits for the stores in flight;
// measured 3-6 us per MEMBAR.SYS inside the update kernel).
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_gpu(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_acquire_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_acquire_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// flag bump without a fence of its own: the caller issues ONE fence.sys and then bumps the flag on
// every rank (a .release on each of 8 bumps would be 8 fences in a row, microseconds each)
__device__ __forceinline__ void red_relaxed_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_release_sys_add(uint32_t* p, uint32_t v) {
    asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Streaming 128-bit accesses that do not allocate in L1.
__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream_u4(uint4* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

}  // namespace cme
