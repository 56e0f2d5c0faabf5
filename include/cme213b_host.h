/*
 * cme213b host library: native C++/OpenMP implementations of the CPU-side
 * workloads and oracles.  Pure C ABI so it can be loaded with ctypes on a box
 * with no GPU (libcme213b_host.so has no CUDA dependency).
 *
 * Every function returns 0 on success and a negative value on a bad argument.
 * Reference behaviour is cited as file:line relative to the reference tree.
 */
#ifndef CME213B_HOST_H_
#define CME213B_HOST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- hw1 q1: parity sums (ref hw1code/main_q1.cpp:16-45) ------------------
 * out[0] = sum of even values, out[1] = sum of odd values, both modulo 2^32
 * like the reference's `uint` accumulators; out64 keeps the exact sums. */
int cmeh_parity_sum_serial(const uint32_t* v, size_t n, uint32_t out[2], uint64_t out64[2]);
int cmeh_parity_sum_omp(const uint32_t* v, size_t n, int num_threads, uint32_t out[2],
                        uint64_t out64[2]);

/* ---- hw1 q2: LSD radix sort (ref hw1code/main_q2.cpp:26-218) -------------- */
/* Stage functions, same contracts as the reference's five pipeline stages so
 * the golden files of hw1code/test_files can be replayed against them. */
int cmeh_radix_block_histograms(const uint32_t* keys, size_t n, uint32_t num_blocks,
                                uint32_t num_bits, uint32_t start_bit, size_t block_size,
                                uint32_t* block_histo /* num_blocks << num_bits */);
int cmeh_radix_reduce_histograms(const uint32_t* block_histo, uint32_t num_blocks,
                                 uint32_t num_buckets, uint32_t* global_histo);
int cmeh_radix_scan_global(const uint32_t* global_histo, uint32_t num_buckets,
                           uint32_t* global_exscan);
int cmeh_radix_block_exscan(const uint32_t* global_exscan, const uint32_t* block_histo,
                            uint32_t num_blocks, uint32_t num_buckets, uint32_t* block_exscan);
int cmeh_radix_populate(const uint32_t* block_exscan, uint32_t num_blocks, uint32_t num_bits,
                        uint32_t start_bit, size_t block_size, const uint32_t* keys, size_t n,
                        uint32_t* sorted);
/* Whole sorts.  `tmp` is caller scratch of n keys; the result lands in `keys`. */
int cmeh_radix_sort_serial(uint32_t* keys, uint32_t* tmp, size_t n, uint32_t num_bits);
int cmeh_radix_sort_parallel(uint32_t* keys, uint32_t* tmp, size_t n, uint32_t num_bits,
                             uint32_t num_blocks, int num_threads);
int cmeh_std_sort(uint32_t* keys, size_t n);

/* ---- hw2: host shift + PageRank oracle (ref hw2code/main_q1.cu:31-38,
 *      hw2code/main_q2.cu:30-85) -------------------------------------------- */
int cmeh_shift_bytes(const uint8_t* in, uint8_t* out, size_t n, uint8_t shift);
int cmeh_pagerank(const int64_t* row_ptr, const uint32_t* edges, const float* rank_in,
                  float* rank_out, const float* inv_deg, int iterations, int64_t num_nodes);
/* Reference graph generator: node i has i/(N/(2*avg-1)+1)+1 edges; targets come
 * from a 64-bit LCG seeded by `seed` (the reference used unseeded rand()). */
int cmeh_pagerank_graph(int64_t num_nodes, int avg_edges, uint64_t seed, int64_t* row_ptr,
                        uint32_t* edges /* may be NULL: only row_ptr+inv_deg */, float* inv_deg);

/* ---- hw3: CPU heat-diffusion oracle (ref hw3code/main.cu:74-214) ----------
 * grid is gx*gy floats, border = order/2.  semantics: 0 = GPU path of the
 * reference (next.border = curr.border*s), 1 = CPU path of the reference
 * (curr.border = next.border*s before each sweep).  Result is left in `grid`. */
int cmeh_heat_cpu(float* grid, float* scratch, int gx, int gy, int order, int iters, float xcfl,
                  float ycfl, float bc_scale, int semantics, int num_threads);
int cmeh_heat_init(float* grid, int gx, int gy, double dx, double dy);

/* ---- ULP comparison (ref fpcode/utils/test_utils.h:18-44,
 *      hw2code/common/mp1-util.h:47-69) ------------------------------------- */
int64_t cmeh_max_ulp_f32(const float* a, const float* b, size_t n, size_t* argmax);
int64_t cmeh_max_ulp_f64(const double* a, const double* b, size_t n, size_t* argmax);

/* ---- hw4: CPU cipher oracles (ref hw4code/create_cipher.cu:69-96) --------- */
int cmeh_letter_histogram(const uint8_t* text, size_t n, size_t stride, size_t offset,
                          uint64_t counts[256]);

/* ---- fpcode: IDX (MNIST) files (ref fpcode/utils/mnist.cpp:9-66) -----------
 * cmeh_idx_info: magic (2051 images / 2049 labels) and up to 3 big-endian dimensions
 * (count, rows, cols; rows = cols = 1 for labels); returns the payload offset in bytes,
 * negative on failure.  cmeh_idx_read: copies `count` payload bytes starting at item
 * `first_item` (item = rows*cols bytes) into dst; pixels stay raw 0..255 as in the reference. */
int64_t cmeh_idx_info(const char* path, int32_t* magic, int64_t dims[3]);
int64_t cmeh_idx_read(const char* path, int64_t first_item, int64_t num_items, uint8_t* dst);

/* ---- fpcode: the serial fp64 trainer of the two-layer MLP (ref fpcode/neural_network.cpp:91-279) ----
 * Row-major buffers: X [N][D] (one sample per row), labels [N] class ids, W1 [H][D], b1 [H],
 * W2 [C][H], b2 [C]; the weights are updated in place.  epochs x ceil(N / batch) minibatches, the
 * last one may be short; sigmoid hidden layer, softmax + cross-entropy, plain SGD with the L2 term
 * reg * W in the weight gradients.  loss_every > 0: before the update of every loss_every-th
 * iteration the minibatch loss (mean cross-entropy + 0.5 reg ||W||^2) is appended to losses
 * (at most losses_cap values; losses may be NULL).  Returns the number of iterations, -1 on a bad
 * argument, -2 on a label outside [0, C). */
int64_t cmeh_mlp_train(const double* X, const int32_t* labels, int64_t N, int D, int H, int C, double* W1,
                       double* b1, double* W2, double* b2, double lr, double reg, int epochs, int batch,
                       int loss_every, double* losses, int64_t losses_cap);
/* Forward pass: probs [N][C] (may be NULL) and arg-max classes pred [N] (may be NULL)
 * (ref fpcode/neural_network.cpp:91-114, 159-169). */
int cmeh_mlp_forward(const double* X, int64_t N, int D, int H, int C, const double* W1, const double* b1,
                     const double* W2, const double* b2, double* probs, int32_t* pred);

int cmeh_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif /* CME213B_HOST_H_ */
