// hw1 on the GPU: parity sums and a stable LSD radix sort.
//
// The reference's hw1 is an OpenMP program (ref hw1code/main_q1.cpp:27-45,
// hw1code/main_q2.cpp:26-159); the native host versions live in
// csrc/host/host_ops.cpp.  These kernels are the device counterparts with the
// same block structure as the reference's parallel radix pass: per-block
// histograms -> global exclusive scan -> stable per-block scatter, 8 bits per
// pass.  Histograms are stored bucket-major ([bucket][block]) so one flat
// exclusive scan yields every block's write offset for every bucket.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kRounds = 16;
constexpr int kTile = kThreads * kRounds;  // keys per block
constexpr int kBuckets = 256;

__global__ void __launch_bounds__(256)
parity_kernel(const uint32_t* __restrict__ v, size_t n, unsigned long long* __restrict__ out2) {
    __shared__ unsigned long long warp_even[8], warp_odd[8];
    unsigned long long even = 0, odd = 0;
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride) {
        const uint32_t x = v[i];
        const uint32_t m = 0u - (x & 1u);
        odd += x & m;
        even += x & ~m;
    }
    even = cme::warp_sum(even);
    odd = cme::warp_sum(odd);
    if ((threadIdx.x & 31) == 0) {
        warp_even[threadIdx.x >> 5] = even;
        warp_odd[threadIdx.x >> 5] = odd;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long e = 0, o = 0;
        for (int w = 0; w < 8; ++w) {
            e += warp_even[w];
            o += warp_odd[w];
        }
        atomicAdd(&out2[0], e);
        atomicAdd(&out2[1], o);
    }
}

__global__ void __launch_bounds__(kThreads)
radix_histogram(const uint32_t* __restrict__ keys, size_t n, int shift, uint32_t nblocks,
                uint32_t* __restrict__ histo /* [bucket][block] */) {
    __shared__ uint32_t h[kBuckets];
    h[threadIdx.x] = 0;
    __syncthreads();
    const size_t base = (size_t)blockIdx.x * kTile;
#pragma unroll 4
    for (int r = 0; r < kRounds; ++r) {
        const size_t i = base + (size_t)r * kThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    histo[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Exclusive scan of `count` uint32 values by one block of 1024 threads.  Every warp owns one
// contiguous segment and walks it 128 values at a time with shuffles only (its running total stays
// in a register), the 32 segment totals are scanned once, and a second sweep adds each segment's
// offset.  The first version synchronised the whole block four times per 4096 values: 64 rounds
// per pass for 4 M keys, about 110 us of the 170 us a pass took.
__global__ void __launch_bounds__(1024) radix_scan(uint32_t* __restrict__ data, size_t count) {
    __shared__ uint32_t warp_tot[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t seg = ((count + 31) / 32 + 127) / 128 * 128;
    const size_t lo = min(count, (size_t)warp * seg), hi = min(count, lo + seg);
    uint32_t carry = 0;
#pragma unroll 4
    for (size_t base = lo; base < hi; base += 128) {
        const size_t i = base + (size_t)lane * 4;
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (i + k < hi) ? data[i + k] : 0;
        const uint32_t tsum = v[0] + v[1] + v[2] + v[3];
        uint32_t inc = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        uint32_t run = carry + inc - tsum;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (i + k < hi) data[i + k] = run;
            run += v[k];
        }
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) warp_tot[warp] = carry;
    __syncthreads();
    if (warp == 0) {
        const uint32_t mine = warp_tot[lane];
        uint32_t w = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        warp_tot[lane] = w - mine;                 // exclusive: what precedes this warp's segment
    }
    __syncthreads();
    const uint32_t off = warp_tot[warp];
    if (off != 0)
        for (size_t i = lo + lane; i < hi; i += 32) data[i] += off;
}

// Stable scatter: keys of a tile are ranked round by round (round, warp, lane
// order = input order).  Inside a warp __match_any_sync groups equal digits.
__global__ void __launch_bounds__(kThreads)
radix_scatter(const uint32_t* __restrict__ keys, uint32_t* __restrict__ out, size_t n, int shift,
              uint32_t nblocks, const uint32_t* __restrict__ offsets /* [bucket][block] */) {
    __shared__ uint32_t cursor[kBuckets];          // next free slot per bucket for this block
    __shared__ uint32_t warp_cnt[kThreads / 32][kBuckets];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    cursor[threadIdx.x] = offsets[(size_t)threadIdx.x * nblocks + blockIdx.x];
    const size_t base = (size_t)blockIdx.x * kTile;
    for (int r = 0; r < kRounds; ++r) {
        for (int w = 0; w < kThreads / 32; ++w) warp_cnt[w][threadIdx.x] = 0;
        __syncthreads();
        const size_t i = base + (size_t)r * kThreads + threadIdx.x;
        const bool valid = i < n;
        const uint32_t key = valid ? keys[i] : 0;
        const uint32_t digit = (key >> shift) & 0xffu;
        // lanes holding the same digit (invalid lanes get a digit no key can have)
        const uint32_t peers = __match_any_sync(0xffffffffu, valid ? digit : 0x100u + lane);
        const uint32_t rank_in_warp = __popc(peers & ((1u << lane) - 1));
        if (valid && rank_in_warp == 0) warp_cnt[warp][digit] = __popc(peers);
        __syncthreads();
        if (valid) {
            uint32_t pos = cursor[digit] + rank_in_warp;
            for (int w = 0; w < warp; ++w) pos += warp_cnt[w][digit];
            out[pos] = key;
        }
        __syncthreads();
        uint32_t add = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) add += warp_cnt[w][threadIdx.x];
        cursor[threadIdx.x] += add;
        __syncthreads();
    }
}

}  // namespace

extern "C" {

int cme_parity_sum(const uint32_t* v, size_t n, unsigned long long* out2, void* stream) {
    CME_CHECK_ARG(out2 && (v || n == 0), "cme_parity_sum: bad argument");
    cudaStream_t st = cme::as_stream(stream);
    CME_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(unsigned long long), st));
    if (n == 0) return 0;
    parity_kernel<<<cme::stride_grid((int64_t)n, 256, 4, 2), 256, 0, st>>>(v, n, out2);
    CME_LAUNCH_CHECK();
    return 0;
}

size_t cme_radix_scratch_bytes(size_t n) {
    const size_t nblocks = (n + kTile - 1) / kTile;
    return (nblocks ? nblocks : 1) * kBuckets * sizeof(uint32_t);
}

int cme_radix_sort(uint32_t* keys, uint32_t* tmp, size_t n, void* scratch, size_t scratch_bytes,
                   void* stream) {
    if (n == 0) return 0;
    CME_CHECK_ARG(keys && tmp && scratch, "cme_radix_sort: null buffer");
    CME_CHECK_ARG(scratch_bytes >= cme_radix_scratch_bytes(n),
                  "cme_radix_sort: scratch needs %zu bytes, got %zu", cme_radix_scratch_bytes(n),
                  scratch_bytes);
    const size_t nblocks = (n + kTile - 1) / kTile;
    // histogram offsets and key positions are 32-bit
    CME_CHECK_ARG(n < (1ull << 32) && nblocks < (1u << 22),
                  "cme_radix_sort: %zu keys; one call sorts fewer than 2^32 keys", n);
    cudaStream_t st = cme::as_stream(stream);
    uint32_t* histo = (uint32_t*)scratch;
    uint32_t* src = keys;
    uint32_t* dst = tmp;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = pass * 8;
        radix_histogram<<<(unsigned)nblocks, kThreads, 0, st>>>(src, n, shift, (uint32_t)nblocks,
                                                                 histo);
        CME_LAUNCH_CHECK();
        radix_scan<<<1, 1024, 0, st>>>(histo, nblocks * kBuckets);
        CME_LAUNCH_CHECK();
        radix_scatter<<<(unsigned)nblocks, kThreads, 0, st>>>(src, dst, n, shift,
                                                               (uint32_t)nblocks, histo);
        CME_LAUNCH_CHECK();
        uint32_t* t = src;
        src = dst;
        dst = t;
    }
    return 0;  // 4 passes: the result is back in `keys`
}

}  // extern "C"
