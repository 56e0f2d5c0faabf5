// hw2 q1 byte-shift bandwidth kernels (ref hw2code/shift.cu:13-99).
//
// The reference has three kernels that differ only in the width of the memory
// transaction (1, 4, 8 bytes).  Here one templated grid-stride kernel covers
// 1/4/8/16-byte words; the packed variants add per byte with __vadd4 so a byte
// that overflows never carries into its neighbour.  The 128-bit variant uses
// L1-bypassing streaming loads/stores, which is the shape that reaches HBM3e
// bandwidth on B200.
#include "common.cuh"

namespace {

__device__ __forceinline__ uint32_t add4(uint32_t w, uint32_t s4) { return __vadd4(w, s4); }

__global__ void __launch_bounds__(256) shift_u8(const uint8_t* __restrict__ in,
                                                uint8_t* __restrict__ out, size_t n, uint8_t s) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = (uint8_t)(in[i] + s);
}

__global__ void __launch_bounds__(256) shift_u32(const uint32_t* __restrict__ in,
                                                 uint32_t* __restrict__ out, size_t nw,
                                                 uint32_t s4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nw; i += stride)
        out[i] = add4(in[i], s4);
}

__global__ void __launch_bounds__(256) shift_u64(const uint2* __restrict__ in,
                                                 uint2* __restrict__ out, size_t nw, uint32_t s4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nw; i += stride) {
        uint2 v = in[i];
        v.x = add4(v.x, s4);
        v.y = add4(v.y, s4);
        out[i] = v;
    }
}

// 4 independent 128-bit loads in flight per thread before the first store.
__global__ void __launch_bounds__(256) shift_u128(const uint4* __restrict__ in,
                                                  uint4* __restrict__ out, size_t nw,
                                                  uint32_t s4) {
    constexpr int U = 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < nw; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = cme::ld_stream_u4(in + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            v[u].x = add4(v[u].x, s4);
            v[u].y = add4(v[u].y, s4);
            v[u].z = add4(v[u].z, s4);
            v[u].w = add4(v[u].w, s4);
            cme::st_stream_u4(out + i + u * stride, v[u]);
        }
    }
    for (; i < nw; i += stride) {
        uint4 v = cme::ld_stream_u4(in + i);
        v.x = add4(v.x, s4);
        v.y = add4(v.y, s4);
        v.z = add4(v.z, s4);
        v.w = add4(v.w, s4);
        cme::st_stream_u4(out + i, v);
    }
}

}  // namespace

extern "C" int cme_shift_bytes(const uint8_t* in, uint8_t* out, size_t n, uint8_t shift,
                               int variant, void* stream) {
    CME_CHECK_ARG(variant >= 0 && variant <= 3, "cme_shift_bytes: variant %d not in 0..3", variant);
    CME_CHECK_ARG((in && out) || n == 0, "cme_shift_bytes: null buffer");
    if (n == 0) return 0;
    cudaStream_t st = cme::as_stream(stream);
    const uint32_t s4 = 0x01010101u * shift;
    const int word = variant == 0 ? 1 : variant == 1 ? 4 : variant == 2 ? 8 : 16;
    if (word > 1)
        CME_CHECK_ARG(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 16 == 0),
                      "cme_shift_bytes: packed variants need 16-byte aligned buffers");
    const size_t nw = n / word, tail = n - nw * word;
    if (nw) {
        const int grid = cme::stride_grid((int64_t)nw, 256, 8, 2);
        switch (variant) {
            case 0: shift_u8<<<grid, 256, 0, st>>>(in, out, nw, shift); break;
            case 1: shift_u32<<<grid, 256, 0, st>>>((const uint32_t*)in, (uint32_t*)out, nw, s4); break;
            case 2: shift_u64<<<grid, 256, 0, st>>>((const uint2*)in, (uint2*)out, nw, s4); break;
            default: shift_u128<<<grid, 256, 0, st>>>((const uint4*)in, (uint4*)out, nw, s4); break;
        }
        CME_LAUNCH_CHECK();
    }
    if (tail) {
        shift_u8<<<1, 32, 0, st>>>(in + nw * word, out + nw * word, tail, shift);
        CME_LAUNCH_CHECK();
    }
    return 0;
}
