// Native host (CPU / OpenMP) implementations behind include/cme213b_host.h.
//
// These are the CPU-side workloads of the reference (hw1 OpenMP programs) and
// the CPU oracles every GPU kernel in this repo is checked against (hw2 host
// shift / PageRank, hw3 cpuComputation, hw4 letter histogram, ULP compares).
// Design notes: radix passes use per-block private histograms laid out
// [block][bucket] exactly like the reference's golden files, but the scatter
// keeps a per-thread offset table on the stack of the worker instead of a
// heap vector per block; parity sums use an OpenMP reduction.
#include "cme213b_host.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

// Every exported function is declared extern "C" in cme213b_host.h, so the
// definitions below pick up C linkage from those declarations.

int cmeh_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// ---------------------------------------------------------------- parity sums
int cmeh_parity_sum_serial(const uint32_t* v, size_t n, uint32_t out[2], uint64_t out64[2]) {
    if (!v && n) return -1;
    uint64_t s[2] = {0, 0};
    for (size_t i = 0; i < n; ++i) s[v[i] & 1u] += v[i];
    if (out) { out[0] = (uint32_t)s[0]; out[1] = (uint32_t)s[1]; }
    if (out64) { out64[0] = s[0]; out64[1] = s[1]; }
    return 0;
}

int cmeh_parity_sum_omp(const uint32_t* v, size_t n, int num_threads, uint32_t out[2],
                        uint64_t out64[2]) {
    if (!v && n) return -1;
    uint64_t even = 0, odd = 0;
#ifdef _OPENMP
    if (num_threads <= 0) num_threads = omp_get_max_threads();
#pragma omp parallel for reduction(+ : even, odd) num_threads(num_threads) schedule(static)
#endif
    for (long long i = 0; i < (long long)n; ++i) {
        const uint32_t x = v[i];
        // branch-free split: the mask selects which accumulator receives x
        const uint64_t m = (uint64_t)0 - (uint64_t)(x & 1u);
        odd += x & m;
        even += x & ~m;
    }
    if (out) { out[0] = (uint32_t)even; out[1] = (uint32_t)odd; }
    if (out64) { out64[0] = even; out64[1] = odd; }
    return 0;
}

// ------------------------------------------------------------------ radix sort
int cmeh_radix_block_histograms(const uint32_t* keys, size_t n, uint32_t num_blocks,
                                uint32_t num_bits, uint32_t start_bit, size_t block_size,
                                uint32_t* block_histo) {
    if (!keys || !block_histo || num_bits == 0 || num_bits > 16 || block_size == 0) return -1;
    const uint32_t buckets = 1u << num_bits, mask = buckets - 1;
    std::memset(block_histo, 0, sizeof(uint32_t) * (size_t)num_blocks * buckets);
#pragma omp parallel for schedule(static)
    for (long long b = 0; b < (long long)num_blocks; ++b) {
        const size_t lo = std::min(n, (size_t)b * block_size);
        const size_t hi = std::min(n, lo + block_size);
        uint32_t* h = block_histo + (size_t)b * buckets;
        for (size_t i = lo; i < hi; ++i) ++h[(keys[i] >> start_bit) & mask];
    }
    return 0;
}

int cmeh_radix_reduce_histograms(const uint32_t* block_histo, uint32_t num_blocks,
                                 uint32_t num_buckets, uint32_t* global_histo) {
    if (!block_histo || !global_histo) return -1;
    for (uint32_t k = 0; k < num_buckets; ++k) {
        uint32_t s = 0;
        for (uint32_t b = 0; b < num_blocks; ++b) s += block_histo[(size_t)b * num_buckets + k];
        global_histo[k] = s;
    }
    return 0;
}

int cmeh_radix_scan_global(const uint32_t* global_histo, uint32_t num_buckets,
                           uint32_t* global_exscan) {
    if (!global_histo || !global_exscan) return -1;
    uint32_t run = 0;
    for (uint32_t k = 0; k < num_buckets; ++k) {
        global_exscan[k] = run;
        run += global_histo[k];
    }
    return 0;
}

int cmeh_radix_block_exscan(const uint32_t* global_exscan, const uint32_t* block_histo,
                            uint32_t num_blocks, uint32_t num_buckets, uint32_t* block_exscan) {
    if (!global_exscan || !block_histo || !block_exscan) return -1;
    // column-wise running sum: bucket k of block b starts where bucket k of
    // block b-1 ended; block 0 starts at the global exclusive scan
    for (uint32_t k = 0; k < num_buckets; ++k) {
        uint32_t run = global_exscan[k];
        for (uint32_t b = 0; b < num_blocks; ++b) {
            block_exscan[(size_t)b * num_buckets + k] = run;
            run += block_histo[(size_t)b * num_buckets + k];
        }
    }
    return 0;
}

int cmeh_radix_populate(const uint32_t* block_exscan, uint32_t num_blocks, uint32_t num_bits,
                        uint32_t start_bit, size_t block_size, const uint32_t* keys, size_t n,
                        uint32_t* sorted) {
    if (!block_exscan || !keys || !sorted || num_bits == 0 || num_bits > 16) return -1;
    const uint32_t buckets = 1u << num_bits, mask = buckets - 1;
#pragma omp parallel
    {
        std::vector<uint32_t> cursor(buckets);
#pragma omp for schedule(static)
        for (long long b = 0; b < (long long)num_blocks; ++b) {
            std::memcpy(cursor.data(), block_exscan + (size_t)b * buckets,
                        sizeof(uint32_t) * buckets);
            const size_t lo = std::min(n, (size_t)b * block_size);
            const size_t hi = std::min(n, lo + block_size);
            for (size_t i = lo; i < hi; ++i) {
                const uint32_t key = keys[i];
                sorted[cursor[(key >> start_bit) & mask]++] = key;
            }
        }
    }
    return 0;
}

static void serial_pass(const uint32_t* in, uint32_t* out, size_t n, uint32_t start_bit,
                        uint32_t num_bits, std::vector<uint32_t>& cursor) {
    const uint32_t buckets = 1u << num_bits, mask = buckets - 1;
    std::fill(cursor.begin(), cursor.end(), 0u);
    for (size_t i = 0; i < n; ++i) ++cursor[(in[i] >> start_bit) & mask];
    uint32_t run = 0;
    for (uint32_t k = 0; k < buckets; ++k) {
        const uint32_t c = cursor[k];
        cursor[k] = run;
        run += c;
    }
    for (size_t i = 0; i < n; ++i) out[cursor[(in[i] >> start_bit) & mask]++] = in[i];
}

int cmeh_radix_sort_serial(uint32_t* keys, uint32_t* tmp, size_t n, uint32_t num_bits) {
    if (!keys || !tmp || num_bits == 0 || num_bits > 16 || (32 % (2 * num_bits)) != 0) return -1;
    std::vector<uint32_t> cursor(1u << num_bits);
    for (uint32_t start = 0; start < 32; start += 2 * num_bits) {
        serial_pass(keys, tmp, n, start, num_bits, cursor);
        serial_pass(tmp, keys, n, start + num_bits, num_bits, cursor);
    }
    return 0;
}

static int parallel_pass(const uint32_t* in, uint32_t* out, size_t n, uint32_t num_bits,
                         uint32_t start_bit, size_t block_size, std::vector<uint32_t>& bh,
                         std::vector<uint32_t>& bx) {
    const uint32_t buckets = 1u << num_bits;
    const uint32_t num_blocks = (uint32_t)((n + block_size - 1) / block_size);
    bh.resize((size_t)num_blocks * buckets);
    bx.resize((size_t)num_blocks * buckets);
    std::vector<uint32_t> gh(buckets), gx(buckets);
    cmeh_radix_block_histograms(in, n, num_blocks, num_bits, start_bit, block_size, bh.data());
    cmeh_radix_reduce_histograms(bh.data(), num_blocks, buckets, gh.data());
    cmeh_radix_scan_global(gh.data(), buckets, gx.data());
    cmeh_radix_block_exscan(gx.data(), bh.data(), num_blocks, buckets, bx.data());
    return cmeh_radix_populate(bx.data(), num_blocks, num_bits, start_bit, block_size, in, n, out);
}

int cmeh_radix_sort_parallel(uint32_t* keys, uint32_t* tmp, size_t n, uint32_t num_bits,
                             uint32_t num_blocks, int num_threads) {
    if (!keys || !tmp || num_bits == 0 || num_bits > 16 || (32 % (2 * num_bits)) != 0 ||
        num_blocks == 0)
        return -1;
    if (n == 0) return 0;
#ifdef _OPENMP
    const int saved = omp_get_max_threads();
    if (num_threads > 0) omp_set_num_threads(num_threads);
#endif
    // ref hw1code/main_q2.cpp:151-159: blockSize = size / numBlocks (floor), the
    // pass then rounds the block count up so the tail keys get their own block
    size_t block_size = n / num_blocks;
    if (block_size == 0) block_size = 1;
    std::vector<uint32_t> bh, bx;
    for (uint32_t start = 0; start < 32; start += 2 * num_bits) {
        parallel_pass(keys, tmp, n, num_bits, start, block_size, bh, bx);
        parallel_pass(tmp, keys, n, num_bits, start + num_bits, block_size, bh, bx);
    }
#ifdef _OPENMP
    omp_set_num_threads(saved);
#endif
    return 0;
}

int cmeh_std_sort(uint32_t* keys, size_t n) {
    if (!keys && n) return -1;
    std::sort(keys, keys + n);
    return 0;
}

// ------------------------------------------------------------------------ hw2
int cmeh_shift_bytes(const uint8_t* in, uint8_t* out, size_t n, uint8_t shift) {
    if ((!in || !out) && n) return -1;
    for (size_t i = 0; i < n; ++i) out[i] = (uint8_t)(in[i] + shift);
    return 0;
}

static void pagerank_step(const int64_t* row_ptr, const uint32_t* edges, const float* in,
                          float* out, const float* inv_deg, int64_t n) {
    const float base = 0.5f / (float)n;
#pragma omp parallel for schedule(static, 4096)
    for (long long i = 0; i < (long long)n; ++i) {
        float sum = 0.f;
        for (int64_t j = row_ptr[i]; j < row_ptr[i + 1]; ++j) {
            const uint32_t e = edges[j];
            sum += in[e] * inv_deg[e];
        }
        out[i] = base + 0.5f * sum;
    }
}

int cmeh_pagerank(const int64_t* row_ptr, const uint32_t* edges, const float* rank_in,
                  float* rank_out, const float* inv_deg, int iterations, int64_t num_nodes) {
    if (!row_ptr || !edges || !rank_in || !rank_out || !inv_deg || iterations < 0 || num_nodes <= 0)
        return -1;
    std::vector<float> a(rank_in, rank_in + num_nodes), b((size_t)num_nodes);
    float* cur = a.data();
    float* nxt = b.data();
    for (int it = 0; it < iterations; ++it) {
        pagerank_step(row_ptr, edges, cur, nxt, inv_deg, num_nodes);
        std::swap(cur, nxt);
    }
    std::memcpy(rank_out, cur, sizeof(float) * (size_t)num_nodes);
    return 0;
}

int cmeh_pagerank_graph(int64_t num_nodes, int avg_edges, uint64_t seed, int64_t* row_ptr,
                        uint32_t* edges, float* inv_deg) {
    if (num_nodes <= 0 || avg_edges <= 0 || !row_ptr) return -1;
    const int64_t nodes_per_block = num_nodes / (2 * (int64_t)avg_edges - 1) + 1;
    row_ptr[0] = 0;
    for (int64_t i = 0; i < num_nodes; ++i) {
        const int64_t deg = i / nodes_per_block + 1;
        row_ptr[i + 1] = row_ptr[i] + deg;
        if (inv_deg) inv_deg[i] = 1.f / (float)deg;
    }
    if (edges) {
        const int64_t m = row_ptr[num_nodes];
        // counter-based generator (splitmix64 of the edge index) so the edge list
        // is the same whatever the thread count
#pragma omp parallel for schedule(static)
        for (long long j = 0; j < (long long)m; ++j) {
            uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(j + 1);
            z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
            z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
            z ^= z >> 31;
            edges[j] = (uint32_t)(z % (uint64_t)num_nodes);
        }
    }
    return 0;
}

// ------------------------------------------------------------------------ hw3
int cmeh_heat_init(float* grid, int gx, int gy, double dx, double dy) {
    if (!grid || gx <= 0 || gy <= 0) return -1;
#pragma omp parallel for schedule(static)
    for (int j = 0; j < gy; ++j) {
        const double sy = std::sin(j * dy);
        for (int i = 0; i < gx; ++i) grid[(size_t)i + (size_t)gx * j] = (float)(std::sin(i * dx) * sy);
    }
    return 0;
}

template <int ORDER>
static inline float stencil_point(const float* c, int gx, float xcfl, float ycfl) {
    // expression order follows ref hw3code/gpuStencil.cu:15-43 term by term so
    // the fp32 cancellation pattern is the same as the reference's
    if (ORDER == 2) {
        return c[0] + xcfl * (c[-1] + c[1] - 2.f * c[0]) + ycfl * (c[gx] + c[-gx] - 2.f * c[0]);
    } else if (ORDER == 4) {
        return c[0] +
               xcfl * (-c[2] + 16.f * c[1] - 30.f * c[0] + 16.f * c[-1] - c[-2]) +
               ycfl * (-c[2 * gx] + 16.f * c[gx] - 30.f * c[0] + 16.f * c[-gx] - c[-2 * gx]);
    } else {
        return c[0] +
               xcfl * (-9.f * c[4] + 128.f * c[3] - 1008.f * c[2] + 8064.f * c[1] -
                       14350.f * c[0] + 8064.f * c[-1] - 1008.f * c[-2] + 128.f * c[-3] -
                       9.f * c[-4]) +
               ycfl * (-9.f * c[4 * gx] + 128.f * c[3 * gx] - 1008.f * c[2 * gx] +
                       8064.f * c[gx] - 14350.f * c[0] + 8064.f * c[-gx] - 1008.f * c[-2 * gx] +
                       128.f * c[-3 * gx] - 9.f * c[-4 * gx]);
    }
}

template <int ORDER>
static void heat_sweep(float* next, const float* curr, int gx, int gy, float xcfl, float ycfl) {
    const int b = ORDER / 2;
#pragma omp parallel for schedule(static)
    for (int y = b; y < gy - b; ++y) {
        const float* c = curr + (size_t)gx * y;
        float* o = next + (size_t)gx * y;
        for (int x = b; x < gx - b; ++x) o[x] = stencil_point<ORDER>(c + x, gx, xcfl, ycfl);
    }
}

static void border_scale(float* dst, const float* src, int gx, int gy, int b, float s) {
    for (int j = 0; j < gy; ++j) {
        const bool full = j < b || j >= gy - b;
        const size_t row = (size_t)gx * j;
        if (full) {
            for (int i = 0; i < gx; ++i) dst[row + i] = src[row + i] * s;
        } else {
            for (int i = 0; i < b; ++i) dst[row + i] = src[row + i] * s;
            for (int i = gx - b; i < gx; ++i) dst[row + i] = src[row + i] * s;
        }
    }
}

int cmeh_heat_cpu(float* grid, float* scratch, int gx, int gy, int order, int iters, float xcfl,
                  float ycfl, float bc_scale, int semantics, int num_threads) {
    if (!grid || !scratch || (order != 2 && order != 4 && order != 8) || iters < 0) return -1;
    const int b = order / 2;
    if (gx <= 2 * b || gy <= 2 * b) return -1;
#ifdef _OPENMP
    const int saved = omp_get_max_threads();
    if (num_threads > 0) omp_set_num_threads(num_threads);
#endif
    float* curr = grid;
    float* next = scratch;
    // the reference copy-constructs next from curr (ref hw3code/main.cu:168)
    std::memcpy(next, curr, sizeof(float) * (size_t)gx * gy);
    for (int it = 0; it < iters; ++it) {
        if (semantics == 1)
            border_scale(curr, next, gx, gy, b, bc_scale);  // CPU path: main.cu:184
        else
            border_scale(next, curr, gx, gy, b, bc_scale);  // GPU path: gpuStencil.cu:115
        if (order == 2) heat_sweep<2>(next, curr, gx, gy, xcfl, ycfl);
        else if (order == 4) heat_sweep<4>(next, curr, gx, gy, xcfl, ycfl);
        else heat_sweep<8>(next, curr, gx, gy, xcfl, ycfl);
        std::swap(curr, next);
    }
    if (curr != grid) std::memcpy(grid, curr, sizeof(float) * (size_t)gx * gy);
#ifdef _OPENMP
    omp_set_num_threads(saved);
#endif
    return 0;
}

// ------------------------------------------------------------------------ ULPs
static inline int64_t ordered_f32(float f) {
    int32_t i;
    std::memcpy(&i, &f, 4);
    return i < 0 ? (int64_t)INT32_MIN - (int64_t)i : (int64_t)i;
}
static inline int64_t ordered_f64(double f) {
    int64_t i;
    std::memcpy(&i, &f, 8);
    return i < 0 ? INT64_MIN - i : i;
}

int64_t cmeh_max_ulp_f32(const float* a, const float* b, size_t n, size_t* argmax) {
    int64_t worst = 0;
    size_t where = 0;
    for (size_t i = 0; i < n; ++i) {
        int64_t d;
        if (std::isnan(a[i]) || std::isnan(b[i])) d = INT64_MAX;
        else d = std::llabs(ordered_f32(a[i]) - ordered_f32(b[i]));
        if (d > worst) { worst = d; where = i; }
    }
    if (argmax) *argmax = where;
    return worst;
}

int64_t cmeh_max_ulp_f64(const double* a, const double* b, size_t n, size_t* argmax) {
    int64_t worst = 0;
    size_t where = 0;
    for (size_t i = 0; i < n; ++i) {
        int64_t d;
        if (std::isnan(a[i]) || std::isnan(b[i])) {
            d = INT64_MAX;
        } else {
            const int64_t x = ordered_f64(a[i]), y = ordered_f64(b[i]);
            // the difference of two ordered doubles can overflow int64: saturate
            const __int128 w = (__int128)x - (__int128)y;
            const __int128 aw = w < 0 ? -w : w;
            d = aw > (__int128)INT64_MAX ? INT64_MAX : (int64_t)aw;
        }
        if (d > worst) { worst = d; where = i; }
    }
    if (argmax) *argmax = where;
    return worst;
}

// ------------------------------------------------------------------------ hw4
int cmeh_letter_histogram(const uint8_t* text, size_t n, size_t stride, size_t offset,
                          uint64_t counts[256]) {
    if ((!text && n) || !counts || stride == 0) return -1;
    std::memset(counts, 0, sizeof(uint64_t) * 256);
    for (size_t i = offset; i < n; i += stride) ++counts[text[i]];
    return 0;
}


// ---- IDX files ---------------------------------------------------------------
namespace {
bool read_be32(FILE* f, uint32_t* v) {
    unsigned char b[4];
    if (std::fread(b, 1, 4, f) != 4) return false;
    *v = ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | (uint32_t)b[3];
    return true;
}

int64_t idx_header(FILE* f, int32_t* magic, int64_t dims[3]) {
    uint32_t m = 0;
    if (!read_be32(f, &m)) return -2;
    const int nd = (int)(m & 0xFF);               // low byte of the magic = number of dimensions
    if (((m >> 8) & 0xFF) != 0x08 || nd < 1 || nd > 3) return -3;   // 0x08 = unsigned byte payload
    dims[0] = dims[1] = dims[2] = 1;
    for (int i = 0; i < nd; ++i) {
        uint32_t d = 0;
        if (!read_be32(f, &d)) return -2;
        dims[i] = d;
    }
    if (magic) *magic = (int32_t)m;
    return 4 + 4 * (int64_t)nd;
}
}  // namespace

int64_t cmeh_idx_info(const char* path, int32_t* magic, int64_t dims[3]) {
    if (!path || !dims) return -1;
    FILE* f = std::fopen(path, "rb");
    if (!f) return -4;
    const int64_t off = idx_header(f, magic, dims);
    std::fclose(f);
    return off;
}

int64_t cmeh_idx_read(const char* path, int64_t first_item, int64_t num_items, uint8_t* dst) {
    if (!path || !dst || first_item < 0 || num_items < 0) return -1;
    FILE* f = std::fopen(path, "rb");
    if (!f) return -4;
    int64_t dims[3];
    int32_t magic = 0;
    const int64_t off = idx_header(f, &magic, dims);
    if (off < 0) {
        std::fclose(f);
        return off;
    }
    const int64_t item = dims[1] * dims[2];
    if (first_item + num_items > dims[0]) {
        std::fclose(f);
        return -5;
    }
    int64_t got = -2;
    if (std::fseek(f, (long)(off + first_item * item), SEEK_SET) == 0)
        got = (int64_t)std::fread(dst, 1, (size_t)(num_items * item), f);
    std::fclose(f);
    return got == num_items * item ? got : -2;
}
