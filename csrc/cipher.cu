// hw4 Vigenere cipher primitives (ref hw4code/create_cipher.cu, solve_cipher.cu).
//
// The reference builds every step from Thrust sorts and segmented reductions
// (sort + reduce_by_key to count 26 letters).  Counting letters is a 256-bin
// histogram, so here it is a shared-memory privatised histogram kernel; the
// per-key-position analysis does all `period` histograms in one pass and never
// reorders the text (the reference sorted it in place and then decrypted the
// sorted buffer, ref solve_cipher.cu:118-173).  Sanitising is a stable stream
// compaction, the index-of-coincidence search evaluates many lags per launch.
#include "common.cuh"

#include <mutex>

namespace {

constexpr int kCompactThreads = 256;
constexpr int kCompactItems = 8;
constexpr int kCompactTile = kCompactThreads * kCompactItems;  // 2048 bytes per block

__device__ __forceinline__ bool keep_lower(uint8_t c, uint8_t& lowered) {
    const uint8_t l = (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c;
    lowered = l;
    return l >= 'a' && l <= 'z';
}

__global__ void __launch_bounds__(kCompactThreads)
compact_count(const uint8_t* __restrict__ in, size_t n, uint32_t* __restrict__ block_counts) {
    __shared__ uint32_t warp_tot[kCompactThreads / 32];
    const size_t base = (size_t)blockIdx.x * kCompactTile + (size_t)threadIdx.x * kCompactItems;
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
        uint8_t l;
        if (base + k < n && keep_lower(in[base + k], l)) ++c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int w = 0; w < kCompactThreads / 32; ++w) t += warp_tot[w];
        block_counts[blockIdx.x] = t;
    }
}

// One block scans all block counts (a few hundred entries for MB-sized text).
__global__ void __launch_bounds__(1024)
compact_scan(uint32_t* __restrict__ counts, uint32_t nblocks, int64_t* __restrict__ total) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < nblocks ? counts[i] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if ((threadIdx.x & 31) >= o) inc += t;
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_tot[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
                if (threadIdx.x >= o) w += t;
            }
            warp_tot[threadIdx.x] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t warp_excl = (threadIdx.x >> 5) ? warp_tot[(threadIdx.x >> 5) - 1] : 0;
        const uint32_t c0 = carry;
        if (i < nblocks) counts[i] = c0 + warp_excl + inc - v;  // exclusive
        __syncthreads();
        if (threadIdx.x == 0) carry = c0 + warp_tot[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = (int64_t)carry;
}

__global__ void __launch_bounds__(kCompactThreads)
compact_scatter(const uint8_t* __restrict__ in, size_t n, const uint32_t* __restrict__ block_off,
                uint8_t* __restrict__ out) {
    __shared__ uint32_t warp_tot[kCompactThreads / 32];
    const size_t base = (size_t)blockIdx.x * kCompactTile + (size_t)threadIdx.x * kCompactItems;
    uint8_t kept[kCompactItems];
    uint32_t c = 0;
#pragma unroll
    for (int k = 0; k < kCompactItems; ++k) {
        uint8_t l;
        if (base + k < n && keep_lower(in[base + k], l)) kept[c++] = l;
    }
    uint32_t inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t off = block_off[blockIdx.x] + inc - c;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) off += warp_tot[w];
    for (uint32_t k = 0; k < c; ++k) out[off + k] = kept[k];
}

// counts of text[offset + k*stride]; per-warp sub-histograms in shared memory
__global__ void __launch_bounds__(256)
histogram_kernel(const uint8_t* __restrict__ text, size_t n, size_t stride, size_t offset,
                 unsigned long long* __restrict__ counts) {
    __shared__ uint32_t sub[8][256];
    for (int i = threadIdx.x; i < 8 * 256; i += 256) (&sub[0][0])[i] = 0;
    __syncthreads();
    uint32_t* mine = sub[threadIdx.x >> 5];
    const size_t items = offset < n ? (n - offset + stride - 1) / stride : 0;
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < items; k += gstride)
        atomicAdd(&mine[text[offset + k * stride]], 1u);
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sub[w][threadIdx.x];
    if (t) atomicAdd(&counts[threadIdx.x], (unsigned long long)t);
}

// counts[p][byte] for p = i % period, all key positions in one pass
__global__ void __launch_bounds__(256)
strided_histogram_kernel(const uint8_t* __restrict__ text, size_t n, int period,
                         unsigned long long* __restrict__ counts) {
    extern __shared__ uint32_t hist[];  // period * 256
    const int bins = period * 256;
    for (int i = threadIdx.x; i < bins; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride)
        atomicAdd(&hist[(int)(i % (size_t)period) * 256 + text[i]], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < bins; i += blockDim.x)
        if (hist[i]) atomicAdd(&counts[i], (unsigned long long)hist[i]);
}

__global__ void __launch_bounds__(256)
vigenere_kernel(const uint8_t* __restrict__ text, uint8_t* __restrict__ out, size_t n,
                const int32_t* __restrict__ shifts, int period, int wrap26) {
    extern __shared__ int32_t key[];
    for (int i = threadIdx.x; i < period; i += blockDim.x) key[i] = shifts[i];
    __syncthreads();
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride) {
        const int s = key[i % (size_t)period];
        const uint8_t c = text[i];
        if (wrap26 && c >= 'a' && c <= 'z') {
            int v = ((int)c - 'a' + s) % 26;
            if (v < 0) v += 26;
            out[i] = (uint8_t)('a' + v);
        } else {
            out[i] = (uint8_t)(c + s);  // byte arithmetic, ref create_cipher.cu:58-61
        }
    }
}

__global__ void __launch_bounds__(256)
lag_match_kernel(const uint8_t* __restrict__ text, size_t n, int lag0,
                 unsigned long long* __restrict__ matches) {
    __shared__ unsigned long long warp_tot[8];
    const size_t lag = (size_t)lag0 + blockIdx.y;
    unsigned long long c = 0;
    if (lag < n) {
        const size_t m = n - lag;
        const size_t gstride = (size_t)gridDim.x * blockDim.x;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gstride)
            c += (text[i] == text[i + lag]);
    }
    c = cme::warp_sum(c);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; ++w) t += warp_tot[w];
        if (t) atomicAdd(&matches[blockIdx.y], t);
    }
}

}  // namespace

extern "C" {

int cme_sanitize_compact(const uint8_t* in, size_t n, uint8_t* out, int64_t* out_count,
                         void* scratch, size_t scratch_bytes, void* stream) {
    CME_CHECK_ARG(out_count != nullptr, "cme_sanitize_compact: out_count is null");
    cudaStream_t st = cme::as_stream(stream);
    if (n == 0) {
        CME_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int64_t), st));
        return 0;
    }
    CME_CHECK_ARG(in && out && scratch, "cme_sanitize_compact: null buffer");
    const size_t nblocks = (n + kCompactTile - 1) / kCompactTile;
    // block counts, scanned offsets and scatter positions are 32-bit
    CME_CHECK_ARG(n < (1ull << 32), "cme_sanitize_compact: %zu bytes; one call handles fewer than 2^32", n);
    CME_CHECK_ARG(scratch_bytes >= nblocks * sizeof(uint32_t),
                  "cme_sanitize_compact: scratch needs %zu bytes, got %zu",
                  nblocks * sizeof(uint32_t), scratch_bytes);
    uint32_t* counts = (uint32_t*)scratch;
    compact_count<<<(unsigned)nblocks, kCompactThreads, 0, st>>>(in, n, counts);
    CME_LAUNCH_CHECK();
    compact_scan<<<1, 1024, 0, st>>>(counts, (uint32_t)nblocks, out_count);
    CME_LAUNCH_CHECK();
    compact_scatter<<<(unsigned)nblocks, kCompactThreads, 0, st>>>(in, n, counts, out);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_byte_histogram(const uint8_t* text, size_t n, size_t stride, size_t offset,
                       unsigned long long* counts, void* stream) {
    CME_CHECK_ARG(counts && stride > 0 && (text || n == 0), "cme_byte_histogram: bad argument");
    cudaStream_t st = cme::as_stream(stream);
    CME_CUDA(cudaMemsetAsync(counts, 0, 256 * sizeof(unsigned long long), st));
    if (n == 0 || offset >= n) return 0;
    const size_t items = (n - offset + stride - 1) / stride;
    histogram_kernel<<<cme::stride_grid((int64_t)items, 256, 4, 2), 256, 0, st>>>(text, n, stride,
                                                                                 offset, counts);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_strided_histograms(const uint8_t* text, size_t n, int period, unsigned long long* counts,
                           void* stream) {
    CME_CHECK_ARG(counts && period > 0 && period <= 200 && (text || n == 0),
                  "cme_strided_histograms: period must be in 1..200");
    cudaStream_t st = cme::as_stream(stream);
    CME_CUDA(cudaMemsetAsync(counts, 0, (size_t)period * 256 * sizeof(unsigned long long), st));
    if (n == 0) return 0;
    const size_t smem = (size_t)period * 256 * sizeof(uint32_t);
    if (smem > 48 * 1024)
        CME_CUDA(cudaFuncSetAttribute(strided_histogram_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    strided_histogram_kernel<<<cme::stride_grid((int64_t)n, 256, 1, 2), 256, smem, st>>>(
        text, n, period, counts);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_vigenere_shift(const uint8_t* text, uint8_t* out, size_t n, const int32_t* shifts,
                       int period, int wrap26, void* stream) {
    CME_CHECK_ARG(shifts && period > 0 && period <= 8192 && ((text && out) || n == 0),
                  "cme_vigenere_shift: bad argument");
    if (n == 0) return 0;
    vigenere_kernel<<<cme::stride_grid((int64_t)n, 256, 8, 2), 256, period * sizeof(int32_t),
                      cme::as_stream(stream)>>>(text, out, n, shifts, period, wrap26);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_lag_match_counts(const uint8_t* text, size_t n, int lag0, int num_lags,
                         unsigned long long* matches, void* stream) {
    CME_CHECK_ARG(matches && lag0 >= 0 && num_lags > 0 && num_lags <= 65535 && (text || n == 0),
                  "cme_lag_match_counts: bad argument");
    cudaStream_t st = cme::as_stream(stream);
    CME_CUDA(cudaMemsetAsync(matches, 0, (size_t)num_lags * sizeof(unsigned long long), st));
    if (n == 0) return 0;
    int gx = cme::stride_grid((int64_t)n, 256, 2, 1);
    if (gx > 128) gx = 128;  // num_lags rows of blocks already fill the machine
    lag_match_kernel<<<dim3(gx, num_lags), 256, 0, st>>>(text, n, lag0, matches);
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"

// =============================================================================
// Sanitise without any allocation
// =============================================================================
// cme_sanitize_compact() needs a small scratch array (one 32-bit counter per
// 2048 input bytes) and asks the caller to bring it.  Some callers have no way
// to do that: a plug-in that is handed device pointers by its host program and
// is not allowed to call the CUDA allocator itself, or a service that pins its
// whole memory budget at start-up and treats any later cudaMalloc (which also
// stalls every stream of the device while it runs) as a fault.  For them the
// scratch below lives inside the library image: every GPU that loads
// libcme213b.so carries these 128 KiB from load until the process exits.
//
// kNoAllocMaxBytes is the largest input the fixed array can serve.
constexpr size_t kNoAllocMaxBytes = (size_t)64 << 20;                          // 64 MiB of text
constexpr size_t kNoAllocSlots = kNoAllocMaxBytes / kCompactTile;             // 32768 counters
static __device__ __align__(256) uint32_t g_noalloc_counts[kNoAllocSlots];    // 128 KiB per GPU

namespace {

// In host code `g_noalloc_counts` is only a handle for the runtime, never an
// address a kernel may dereference: the device address has to be asked for, with
// the device that will run the kernels current.  It is asked for on EVERY call
// and never cached: once the library's module is on the device this is a host-side
// table lookup, and an address remembered across cudaDeviceReset (or any other
// re-creation of the primary context, which a plug-in cannot prevent its host
// program from doing) would point into memory that no longer belongs to the array.
int noalloc_scratch_for_current_device(uint32_t** out) {
    uint32_t* p = nullptr;
    CME_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_counts));
    CME_CHECK_ARG(p != nullptr, "cme_sanitize_compact_noalloc: the scratch array has no device address");
    *out = p;
    return 0;
}

}  // namespace

extern "C" int cme_sanitize_compact_noalloc(const uint8_t* in, size_t n, uint8_t* out,
                                            int64_t* out_count, void* stream) {
    CME_CHECK_ARG(out_count != nullptr, "cme_sanitize_compact_noalloc: out_count is null");
    CME_CHECK_ARG(n <= kNoAllocMaxBytes,
                  "cme_sanitize_compact_noalloc: %zu bytes exceed the built-in scratch limit of %zu "
                  "bytes; use cme_sanitize_compact with caller scratch",
                  n, kNoAllocMaxBytes);
    cudaStream_t st = cme::as_stream(stream);
    if (n == 0) {
        CME_CUDA(cudaMemsetAsync(out_count, 0, sizeof(int64_t), st));
        return 0;
    }
    CME_CHECK_ARG(in && out, "cme_sanitize_compact_noalloc: null buffer");
    uint32_t* counts = nullptr;
    int rc = noalloc_scratch_for_current_device(&counts);
    if (rc) return rc;
    // Nothing in the shared array has to start at zero: the first kernel writes
    // every counter this call reads (slots 0..nblocks-1) before the scan reads
    // them, so whatever an earlier call left behind is overwritten.
    const size_t nblocks = (n + kCompactTile - 1) / kCompactTile;
    compact_count<<<(unsigned)nblocks, kCompactThreads, 0, st>>>(in, n, counts);
    CME_LAUNCH_CHECK();
    compact_scan<<<1, 1024, 0, st>>>(counts, (uint32_t)nblocks, out_count);
    CME_LAUNCH_CHECK();
    compact_scatter<<<(unsigned)nblocks, kCompactThreads, 0, st>>>(in, n, counts, out);
    CME_LAUNCH_CHECK();
    return 0;
}
