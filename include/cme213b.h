/*
 * cme213b device library (libcme213b.so): hand-written sm_100a kernels behind a
 * plain C ABI.  Pointers are device pointers unless a parameter says "host".
 * `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *
 * Return value: 0 on success, a positive cudaError_t value on a CUDA failure,
 * a negative value on a rejected argument.  cme_last_error() returns a
 * thread-local message for the last non-zero return.
 *
 * Nothing here synchronises the device unless the comment says so.
 * Reference behaviour is cited as file:line relative to the reference tree.
 */
#ifndef CME213B_H_
#define CME213B_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* cme_last_error(void);
int cme_version(void);
/* sm_count / cc_major / cc_minor of the current device (any may be NULL). */
int cme_device_info(int* sm_count, int* cc_major, int* cc_minor);
/* Get (and, if bytes > 0, set) the L2 fetch granularity of the current device: 32, 64 or 128 bytes.
 * Device-wide hint; the previous value comes back in *old_bytes. */
int cme_l2_fetch_granularity(int bytes, int* old_bytes);
/* Creates the context, loads the module and runs one trivial kernel, then waits for it
 * (the reference's device_add_one warm-up, ref fpcode/gpu_func.cu:11-36). *value_inout += 1. */
int cme_warmup(int* value_inout);

/* ---- hw2 q1: byte shift (ref hw2code/shift.cu:13-99) -----------------------
 * out[i] = in[i] + shift (mod 256) for i < n.  variant: 0 = one byte per thread
 * step, 1 = 32-bit words, 2 = 64-bit words, 3 = 128-bit vectors (the B200
 * path).  Packed variants use per-byte SIMD adds, so bytes never carry into
 * their neighbours (the reference's plain 32-bit add did, ref shift.cu:29).
 * in/out must be 16-byte aligned for variants 1-3; the tail is handled. */
int cme_shift_bytes(const uint8_t* in, uint8_t* out, size_t n, uint8_t shift, int variant,
                    void* stream);

/* ---- hw2 q2: PageRank (ref hw2code/pagerank.cu:9-136) ----------------------
 * One propagation step over rows [row_begin,row_end) of a CSR graph:
 *   rank_out[i] = 0.5/N + 0.5 * sum_j contrib_in[edges[j]]
 * with contrib = rank*inv_deg precomputed per node (bitwise the same products
 * the reference forms per edge).  Also writes contrib_out[i] = rank_out[i] *
 * inv_deg[i] for the next step.  If nranks > 1, peer_contrib is a DEVICE array
 * of nranks pointers to every rank's contrib_out buffer (peer-mapped) and the
 * slice is stored to all of them (the rank all-gather fused into the SpMV). */
int cme_pagerank_prescale(const float* rank, const float* inv_deg, float* contrib, int64_t n,
                          void* stream);
int cme_pagerank_step(const int64_t* row_ptr, const uint32_t* edges, const float* contrib_in,
                      const float* inv_deg, float* rank_out, float* contrib_out,
                      float* const* peer_contrib, int nranks, int64_t row_begin, int64_t row_end,
                      int64_t num_nodes, void* stream);
/* `iterations` steps on one GPU, ping-ponging contrib_a / contrib_b (n floats each): the pre-scale
 * and every iteration in ONE cooperative launch with grid barriers, no host synchronisation
 * (ref hw2code/pagerank.cu:59-136 device_graph_iterate).  rank_out gets the result. */
int cme_pagerank_iterate(const int64_t* row_ptr, const uint32_t* edges, const float* rank0,
                         const float* inv_deg, float* contrib_a, float* contrib_b, float* rank_out,
                         int iterations, int64_t num_nodes, void* stream);
/* One pass of the column-blocked propagation (graphs larger than L2): blk_ptr holds nrows + 1 offsets
 * into edges_b for THIS block of source nodes (local rows), partial carries the row sums between
 * passes; the last pass writes rank_out / the new contributions like cme_pagerank_step. */
int cme_pagerank_block_step(const int32_t* blk_ptr, const uint32_t* edges_b, const float* contrib_in, float* partial,
                            const float* inv_deg, float* rank_out, float* contrib_out, float* const* peer_contrib,
                            int nranks, int64_t row_begin, int64_t nrows, int64_t num_nodes, int first, int last,
                            void* stream);
/* Device-side graph generator matching cmeh_pagerank_graph (same splitmix64). */
int cme_pagerank_gen_edges(uint32_t* edges, int64_t edge_begin, int64_t edge_count,
                           int64_t num_nodes, uint64_t seed, void* stream);

/* ---- hw3: 2-D heat diffusion (ref hw3code/gpuStencil.cu, hw3code/BC.h) -----
 * Grids are gx*gy floats, border b = order/2.  One call = one time step of the
 * whole interior.  variant: 0 = one point per thread (ref gpuStencil),
 * 1 = several rows per thread (ref gpuStencilLoop), 2 = shared-memory tile
 * with halo (what the reference left as a stub, ref gpuStencil.cu:263-308).
 * bc: if non-zero the same launch also writes next.border = curr.border *
 * bc_scale (ref BC.h:7-38) for the border cells adjacent to the rows it owns. */
int cme_heat_step(float* next, const float* curr, int gx, int gy, int order, float xcfl,
                  float ycfl, int variant, int bc, float bc_scale, void* stream);
int cme_heat_bc(float* next, const float* curr, int gx, int gy, int order, float bc_scale,
                void* stream);
/* Slab step for domain decomposition along y.  The local grid holds
 * rows_local interior rows plus b ghost rows above and below.  After computing,
 * the kernel's edge blocks store the first/last b computed rows straight into
 * the neighbours' ghost rows (peer-mapped pointers, may be NULL at the domain
 * ends) -- the halo exchange fused into the sweep.  Physical top/bottom border
 * rows (ranks at the ends) and the left/right border columns get the BC decay. */
int cme_heat_slab_step(float* next, const float* curr, float* up_next_ghost,
                       float* down_next_ghost, int gx, int rows_local, int order, float xcfl,
                       float ycfl, float bc_scale, int is_top, int is_bottom, void* stream);

/* ---- hw4: cipher primitives (ref hw4code/create_cipher.cu, solve_cipher.cu) */
/* Lower-case A-Z, drop everything that is not a letter; *out_count (device
 * int64) receives the number of bytes kept.  Stable. scratch: >= 4*(n/2048+2)
 * +16 bytes of device memory. */
int cme_sanitize_compact(const uint8_t* in, size_t n, uint8_t* out, int64_t* out_count,
                         void* scratch, size_t scratch_bytes, void* stream);
/* The same operation for callers that cannot provide scratch memory: a plug-in
 * that only receives device pointers from its host application and may not call
 * the CUDA allocator, or a long-running service that fixes its memory budget at
 * start-up and forbids later cudaMalloc/cudaFree (each of which also stalls the
 * whole device while it runs).  The counters live in a 128 KiB array inside the
 * library image, held by every GPU that loads the library until the process
 * exits, which limits one call to n <= 64 MiB (67108864 bytes); a larger n is
 * rejected with a negative return and a message naming the limit.  The call
 * itself performs no cudaMalloc, cudaFree or other allocating call.
 *   - Every call asks the runtime for the array's address on the current device
 *     (that device must be the one the stream belongs to) and keeps nothing
 *     between calls, so it stays correct across cudaDeviceReset or a re-created
 *     context.  The first call on a device also loads the library's code onto
 *     it: make that first call before any stream capture is begun there.
 *   - All calls on a device share the one array.  It keeps whatever the last call
 *     left in it; this is harmless because a call overwrites every counter it
 *     later reads.  Two calls that can run at the same time on one GPU (different
 *     streams, or different threads) race on it: serialise them.
 * Results are identical to cme_sanitize_compact.  This entry point has no
 * counterpart in the reference. */
int cme_sanitize_compact_noalloc(const uint8_t* in, size_t n, uint8_t* out, int64_t* out_count,
                                 void* stream);
/* counts[256] (device uint64, zeroed by the call) of text[offset + k*stride]. */
int cme_byte_histogram(const uint8_t* text, size_t n, size_t stride, size_t offset,
                       unsigned long long* counts, void* stream);
/* All key positions at once: counts[period][256]. */
int cme_strided_histograms(const uint8_t* text, size_t n, int period,
                           unsigned long long* counts, void* stream);
/* out[i] = text[i] + shifts[i % period] as unsigned char, no mod-26 wrap
 * (ref create_cipher.cu:58-61); wrap26 != 0 keeps letters inside a-z. */
int cme_vigenere_shift(const uint8_t* text, uint8_t* out, size_t n, const int32_t* shifts,
                       int period, int wrap26, void* stream);
/* matches[k] = #{i : text[i] == text[i + lag0 + k]}, k < num_lags
 * (ref solve_cipher.cu:83-88, many lags per launch). */
int cme_lag_match_counts(const uint8_t* text, size_t n, int lag0, int num_lags,
                         unsigned long long* matches, void* stream);

/* ---- hw1 on the GPU: parity sums and LSD radix sort ------------------------ */
int cme_parity_sum(const uint32_t* v, size_t n, unsigned long long* out2, void* stream);
/* 4 passes of 8 bits; tmp = n keys, scratch >= cme_radix_scratch_bytes(n). */
size_t cme_radix_scratch_bytes(size_t n);
int cme_radix_sort(uint32_t* keys, uint32_t* tmp, size_t n, void* scratch, size_t scratch_bytes,
                   void* stream);

/* ---- fp64 parity path (ref fpcode/gpu_func.cu) -----------------------------
 * Column-major C = alpha*op(A)*op(B) + beta*C, the reference's myGEMM contract
 * (ref gpu_func.cu:259-297) with both transposes allowed.  M,N,K are the
 * dimensions of op(A) (MxK), op(B) (KxN), C (MxN); lda/ldb/ldc leading dims. */
int cme_dgemm(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
              const double* B, int ldb, double beta, double* C, int ldc, void* stream);
/* The round-1 CUDA-core kernel behind the same contract, kept for A/B timing (cli/gemm_bench). */
int cme_dgemm_simt(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
              const double* B, int ldb, double beta, double* C, int ldc, void* stream);
/* Fused fp64 MLP pieces, column-major like the reference's device_cache.
 * z is rows x cols (one sample per column). */
int cme_d_bias_sigmoid(double* z, const double* bias, int rows, int cols, void* stream);
/* In place: z <- softmax(z + bias) per column (max-subtracted); then
 * diff = (softmax - onehot(label)) * scale; loss_sum += -log p[label] (may be NULL). */
int cme_d_bias_softmax_ce(double* z, const double* bias, const int32_t* labels, double* diff,
                          double scale, double* loss_sum, int rows, int cols, void* stream);
/* dz = da * a * (1-a) elementwise. */
int cme_d_dsigmoid(const double* da, const double* a, double* dz, size_t n, void* stream);
/* out[r] = sum_c m[r + c*rows]  (ref gpu_func.cu:359-382). */
int cme_d_row_sum(const double* m, double* out, int rows, int cols, void* stream);
/* p -= lr * (g + reg * p) for n values (reg = 0 for biases). */
int cme_d_sgd(double* p, const double* g, size_t n, double lr, double reg, void* stream);
int cme_d_axpby(double* c, const double* a, const double* b, double alpha, double beta, size_t n,
                void* stream);

/* Per-step arguments kept in device memory, so that one recorded CUDA graph serves every
 * minibatch position and every learning rate: the kernels of a step issued with x_u8 == NULL
 * read them from plan->step_args; cme_mlp_set_step_args (a one-thread kernel, stream-ordered)
 * writes them before the step. */
typedef struct CmeStepArgs {
    const uint8_t* x_u8;
    const int32_t* labels;
    float lr, reg, scale;
    int reserved;
} CmeStepArgs;

/* ---- bf16 tensor-core path --------------------------------------------------
 * Row-major activations: X[B][ldx] bf16, A1[B][ldh] bf16, compute weights
 * W1c[Hp][ldx] bf16 (K-major for tcgen05), W2[C][H] fp32.  ldx, ldh are
 * multiples of 64 elements, Hp = H rounded up to 64.  The bias b1 is folded
 * into the GEMMs: X carries a column of ones at index K-1 and W1c carries b1
 * there, so K below is "pixels + 1" (785 for MNIST). */

/* Minibatch staging: uint8 pixels (rows x cols, dense) -> bf16 rows with leading
 * dimension ld; if add_ones, column `cols` is set to 1.0; remaining pad columns
 * are zeroed.  src may be a peer-mapped pointer into another GPU's memory: each
 * rank pulls its slice of the minibatch from the owner over NVLink, which
 * replaces the host MPI_Scatter (ref fpcode/neural_network.cpp:460-481). */
int cme_u8_to_bf16_rows(const uint8_t* src, void* dst_bf16, int64_t rows, int cols, int ld,
                        int add_ones, void* stream);

/* tcgen05 GEMM #1:  A1 = sigmoid(X * W1c^T), bf16 out through TMA stores, fp32
 * accumulate in TMEM.  X: rows x K, W1c: Hp x K (pad rows zero), A1: rows x Hp.
 * b1 must be NULL (the bias is inside W1c).  Replaces ref gpu_func.cu K1+K8+K6
 * (sGEMM_kernel :92-147, gpuMatVecSum :385-408, gpuSigmoid :332-356). */
int cme_mlp_fwd1(const void* X, const void* W1c, const float* b1, void* A1, int64_t rows, int K,
                 int H, int ldx, int ldh, void* stream);
/* Plain bf16 GEMM on the same kernel: C[M,N] = A[M,K] * B[N,K]^T (N % 64 == 0).
 * Used by the GEMM benchmark against cuBLAS (ref fpcode/utils/tests.cpp:126-280). */
int cme_gemm_bf16(const void* A, const void* B, void* C, int64_t M, int N, int K, int lda, int ldb,
                  int ldc, void* stream);
/* Layer 2 + softmax + cross-entropy gradient in one pass over A1:
 * dz2[r][c] = (softmax(A1[r]*W2^T + b2)[c] - [c==label[r]]) * scale as fp32 rows
 * of 16; loss_sum (may be NULL) += -log p[label]; pred (may be NULL) = argmax;
 * gb2 (may be NULL) += column sums of dz2.  Replaces K1(fwd2)+K8+K5+K9+K7(db2). */
int cme_mlp_fwd2_softmax(const void* A1, const float* W2, const float* b2, const int32_t* labels,
                         float* dz2, float* loss_sum, int32_t* pred, float* gb2, float scale,
                         int64_t rows, int H, int ldh, int C, void* stream);
/* Backward elementwise + dW2 in one pass over A1:
 * dZ1[r][h] = (sum_c dz2[r][c]*W2[c][h]) * a*(1-a) (bf16 out),
 * gW2[c][h] += sum_r dz2[r][c]*A1[r][h].  Replaces K2+K10+K3(dW2).  gW2 must be
 * zeroed by the caller. */
int cme_mlp_bwd_elem(const void* A1, const float* dz2, const float* W2, void* dZ1, float* gW2,
                     int64_t rows, int H, int ldh, int C, void* stream);
/* The same backward pass with gW2 == NULL only produces dZ1 (half the FMA work);
 * pair it with cme_mlp_dw2_tc. */
/* Layer 2 on the tensor cores: logits = A1 * W2c^T accumulate in TMEM (N = 16,
 * K = H, the kernel streams A1 once at HBM speed), and the epilogue does bias,
 * softmax, cross-entropy, arg-max, dz2 (fp32 rows of 16), dz2b (the same values
 * as bf16 rows of 64 for cme_mlp_dw2_tc; may be NULL), the b2 gradient and the
 * loss in registers.  W2c: bf16 [16][ldh], rows >= C zero. */
int cme_mlp_fwd2_tc(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                    float* dz2, void* dz2b, float* loss_sum, int32_t* pred, float* gb2, float scale,
                    int64_t rows, int H, int ldh, int C, void* stream);
/* gW2[c][h] += sum_r dz2b[r][c] * A1[r][h] on the tensor cores (stream-K over the
 * batch, result tile stored transposed).  gW2 must be zeroed first. */
int cme_mlp_dw2_tc(const void* A1, const void* dz2b, float* gW2, int64_t rows, int H, int ldh, int C,
                   void* stream);
/* dZ1 = (dz2b * W2T^T) .* A1 .* (1 - A1) on the tensor cores: the K = 16 product is
 * one MMA per 128 x 256 tile, A1 chunks are prefetched by TMA into a staging ring,
 * combined with the accumulator in place and written back by TMA stores, so the
 * pass runs at HBM speed.  W2T: bf16 [H][64] with W2T[h][c] = W2[c][h], columns
 * >= 16 unused.  dz2b: bf16 [rows][64] from cme_mlp_fwd2_tc. */
int cme_mlp_bwd_dz1_tc(const void* A1, const void* dz2b, const void* W2T, void* dZ1, int64_t rows,
                       int H, int ldh, void* stream);
/* cme_mlp_bwd_dz1_tc and cme_mlp_dw2_tc in ONE pass over A1: every 128 x 128 tile of A1 is
 * staged once, used as the MN-major operand of the dW2 product (accumulated in TMEM across
 * the row tiles of a column block) and then overwritten in place with dZ1.  gW2 rows have
 * pitch ldg2 (>= H) so a column block of a wider layer can be processed; gW2 must be zeroed
 * first.  Replaces ref K2 + K10 + K3(dW2) (fpcode/neural_network.cpp:357-378). */
int cme_mlp_bwd_tail_tc(const void* A1, const void* dz2b, const void* W2T, void* dZ1, float* gW2,
                        int ldg2, int64_t rows, int H, int ldh, int C, void* stream);
/* The same pass with a bit-reproducible dW2: every CTA stores its share of a column block into
 * a slot of `part` (cme_mlp_bwd_tail_scratch_floats(H) floats) and a small second kernel adds
 * the slots in CTA order and STORES gW2 (no zero-fill needed).  part == NULL: as above. */
size_t cme_mlp_bwd_tail_scratch_floats(int H);
int cme_mlp_bwd_tail_det(const void* A1, const void* dz2b, const void* W2T, void* dZ1, float* gW2,
                         int ldg2, int64_t rows, int H, int ldh, int C, float* part, void* stream);
/* The whole forward pass in one tensor-core kernel plus a per-row tail: A1 = sigmoid(X W1c^T) as
 * cme_mlp_fwd1, and the layer-2 logits from the same staged tiles (a 256 x 16 x 64 "rider" MMA on
 * every 128 x 64 chunk of A1 while it sits in shared memory for its TMA store), so A1 is not
 * read again for layer 2.  The tail adds the partial logits in a fixed order and writes dz2, dz2b,
 * pred, gb2 (STORED: no zero-fill needed) and loss_sum (+= once) exactly like cme_mlp_fwd2_tc, but
 * bit-reproducibly.  Needs H rounded up to 64 to be a multiple of 256 and rows >= 256
 * (cme_mlp_fwd_fused_ok); scratch: cme_mlp_fwd_scratch_floats(max_rows) floats, zeroed once.
 * Replaces ref fpcode/neural_network.cpp:289-323 (two GEMMs, two bias passes, sigmoid, softmax). */
int cme_mlp_fwd_fused_ok(int64_t rows, int H);
size_t cme_mlp_fwd_scratch_floats(int64_t max_rows);
int cme_mlp_fwd_fused(const void* X, const void* W1c, const void* W2c, const float* b2, const int32_t* labels,
                      const CmeStepArgs* step_args, void* A1, float* dz2, void* dz2b, float* loss_sum,
                      int32_t* pred, float* gb2, float* scratch, int64_t max_rows, float scale, int64_t rows,
                      int K, int H, int ldx, int ldh, int C, void* stream);
/* W2 (fp32 [C][H]) -> w2c (bf16 [16][ldh]) and, if not NULL, w2t (bf16 [H][64]). */
int cme_mlp_recast_w2(const float* W2, void* w2c, int ldh, void* w2t, int H, int C, void* stream);
/* fp32 [rows][ld_src] -> bf16 [rows][ld_dst], first `cols` columns (W2 -> W2c). */
int cme_f32_to_bf16_2d(const float* src, int ld_src, void* dst_bf16, int ld_dst, int rows, int cols,
                       void* stream);
/* tcgen05 GEMM #2: gW1[h][k] = sum_r dZ1[r][h] * X[r][k]  (H x K fp32 with ld = ldg; every
 * column below ldg of the rows below H is WRITTEN, the pad columns with zeros), both operands
 * MN-major, the batch cut into k-chunks dealt to all SMs.  The chunks of a tile are combined in
 * chunk order (chunk 0 stores, chunk c adds after chunk c-1 has landed), so the result is
 * bit-reproducible and gW1 needs no zero-fill.  Column K-1 (the ones column) receives the b1
 * gradient.  Replaces K3 (sGEMM_BT_kernel :205-256) for dW1 and K7 for db1.
 * cme_mlp_dw1 keeps its chunk counters in a per-device buffer allocated on first use: do not
 * run two of these calls concurrently on one device.  cme_mlp_dw1_sync takes the counters from
 * the caller (scratch: cme_mlp_dw1_scratch_bytes() bytes, zeroed once) and, when pads != NULL,
 * bumps flag kPadUnit0 + u (word 2048 + u) on each of the nranks pads as soon as rows
 * [u*unit_rows, (u+1)*unit_rows) of gW1 are complete (unit_rows = 128, or 256 for H >= 512):
 * the publication the streaming update (cme_mlp_update_stream) waits for.  status (may be NULL):
 * word that receives code 3 if a chunk waits ~2 s for its turn (the grid was not fully resident,
 * e.g. another large kernel shares the GPU); with NULL the kernel traps instead. */
int cme_mlp_dw1(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                int ldh, int ldg, void* stream);
/* Tiling of a dW1 product of `rows` samples without launching anything (works without a GPU; 148 SMs
 * assumed then): out = {CTA-pair path 0/1, row tiles, column tiles, column-tile width, k-split, CTAs}. */
int cme_mlp_dw1_plan_query(int64_t rows, int K, int H, int out[6]);
size_t cme_mlp_dw1_scratch_bytes(void);
int cme_mlp_dw1_sync(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                     int ldh, int ldg, void* scratch, uint32_t* const* pads, int nranks, int unit_rows,
                     uint32_t* status, void* stream);

/* Flat fp32 parameter / gradient vector of the bf16 trainer:
 *   [ W1aug (Hp x ldx, b1 in column K-1) | W2 (C x H) | b2 (C) ]  each segment
 * padded to a multiple of 4 floats.
 * Fused all-reduce + L2 + SGD + bf16 re-cast.  Rank r owns the r-th 1/nranks
 * slice: it sums that slice of every rank's gradient buffer straight from peer
 * memory (fixed rank order), applies  p -= lr * (sum_g + reg * p)  (reg only on
 * W1 columns < kreg and on [w2_begin, w2_end)).  For the W1 part the new fp32
 * value is stored only in the owner's master copy (the fp32 master of W1 is
 * sharded by slice; gather the slices before reading it on the host) and the
 * bf16 value is stored into every rank's W1c; for the tail (W2, b2) the fp32
 * value is stored into every rank's replica.  Replaces the four MPI_Allreduce calls and the host
 * SGD (ref fpcode/neural_network.cpp:501-541).  grads/params/w1c are HOST
 * arrays of nranks peer-mapped device pointers.  The caller orders ranks with
 * cme_peer_barrier before (gradients complete) and after (weights visible).
 * Only [range_begin, range_end) of the flat vector is processed (cut into nranks
 * slices), so the trainer can reduce one bucket while later gradients are still
 * being computed; pass 0, n_total for everything. */
int cme_allreduce_sgd(float* const* grads, float* const* params, void* const* w1c, int rank,
                      int nranks, int64_t n_total, int64_t n_w1, int ldx, int kreg,
                      int64_t w2_begin, int64_t w2_end, float lr, float reg, int64_t range_begin,
                      int64_t range_end, void* stream);
/* The slice [out[0], out[1]) of the update range [range_begin, range_end) that `rank` owns (sums,
 * updates, and keeps the fp32 master of): whole 16-byte vectors, contiguous, in rank order. */
int cme_update_slice(int64_t range_begin, int64_t range_end, int rank, int nranks, int64_t out[2]);
/* Same update through the NVSwitch multicast mapping: multimem.ld_reduce sums in
 * the switch, multimem.st replicates the result to all GPUs. */
int cme_allreduce_sgd_mc(const float* mc_grads, float* local_params, float* mc_params,
                         void* mc_w1c, int rank, int nranks, int64_t n_total, int64_t n_w1, int ldx,
                         int kreg, int64_t w2_begin, int64_t w2_end, float lr, float reg,
                         int64_t range_begin, int64_t range_end, void* stream);

/* fp64 parity engine: gradient all-reduce + L2 + SGD in ONE launch that carries its own peer
 * handshake (pad channels 6 and 7; no barrier launches).  grads/params: one pointer per rank
 * (peer-mapped), n doubles each (n even, buffers 16-byte aligned).  reg applies to [0, w1_end)
 * and [w2_begin, w2_end).  Every rank reduces 1/nranks of the vector in rank order and writes
 * every replica.  Waits are bounded by timeout_ms (0 = 20 s); an expired wait leaves 1 + the
 * missing rank in word 6*64+61 of this rank's pad. */
int cme_allreduce_sgd_f64(double* const* grads_host, double* const* params_host, uint32_t* const* pads_dev,
                          int rank, int nranks, int64_t n, int64_t w1_end, int64_t w2_begin, int64_t w2_end,
                          double lr, double reg, int timeout_ms, void* stream);

/* ---- native step executor ------------------------------------------------------
 * Every device buffer of one rank's bf16 engine.  D inputs, H hidden units, C classes;
 * ldx = (D+1) rounded up to 64, Hp = ldh = H rounded up to 64.  params / grads are the
 * flat fp32 vectors described above, one pointer per rank (peer-mapped for nranks > 1);
 * w1c: bf16 [Hp][ldx] per rank.  Local scratch: xb bf16 [max_rows][ldx], a1 and dz1 bf16
 * [max_rows][ldh], dz2 fp32 [max_rows][16], dz2b bf16 [max_rows][64] (zero-initialised),
 * pred int32 [max_rows], w2c bf16 [16][ldh] and w2t bf16 [Hp][64] (zero-initialised).
 * pads_dev: device array of per-rank flag pads (see cme_peer_barrier), NULL for one rank. */
typedef struct CmeMlpPlan {
    int D, H, C, Hp, ldx, ldh;
    int rank, nranks;
    int64_t max_rows;
    float* params[16];
    float* grads[16];
    void* w1c[16];
    void *xb, *a1, *dz1, *dz2b, *w2c, *w2t;
    float* dz2;
    int32_t* pred;
    uint32_t* const* pads_dev;
    /* overlap != 0: the streaming update runs INSIDE the dW1 GEMM (four extra warps per CTA);
     * 0: as a kernel of its own after it, on the same stream.  use_mc: reduce through the NVSwitch multicast mappings below
     * instead of peer loads. */
    int overlap, use_mc;
    const float* grads_mc;
    float* params_mc;
    void* w1c_mc;
    /* Streaming update (cme_mlp_update_stream).  unit_rows: rows of W1 per publication unit, 256
     * when Hp >= 512 else 128; it fixes which rank owns which slice of the sharded fp32 master, so
     * it must not change during a run.  pad_local = this rank's flag pad (pads_dev[rank], 4096
     * zero-initialised uint32; required for ONE rank too).  sync_scratch: cme_mlp_dw1_scratch_bytes()
     * bytes, zeroed once.  timeout_ms: bound on every cross-GPU wait (0 = 20 s). */
    int unit_rows;
    int timeout_ms;
    uint32_t* pad_local;
    void* sync_scratch;
    /* Optional: device memory for the indirect step arguments (see CmeStepArgs). */
    CmeStepArgs* step_args;
    /* Optional: cme_mlp_fwd_scratch_floats(max_rows) floats, zeroed once.  When present (and the
     * shape allows it, cme_mlp_fwd_fused_ok) the step runs the forward pass as cme_mlp_fwd_fused. */
    float* fwd_scratch;
    /* Optional: cme_mlp_bwd_tail_scratch_floats(H) floats.  When present dW2 is summed in a fixed
     * order (cme_mlp_bwd_tail_det); with fwd_scratch too the whole step is bit-reproducible and
     * nothing in it needs a zero-fill. */
    float* dw2_scratch;
} CmeMlpPlan;
/* One whole training step on `rows` local samples (uint8 pixels x_u8 [rows][D], labels
 * int32 [rows], both device pointers; x_u8 may be peer-mapped): staging, forward, backward,
 * streaming all-reduce + L2 + SGD, bf16 re-cast, with no host synchronisation.
 * scale = 1 / (global samples in the minibatch); loss_sum (may be NULL) += cross-entropy.
 * The C++ counterpart of one iteration of parallel_train's loop
 * (ref fpcode/neural_network.cpp:484-541).  cme_mlp_step issues everything on `stream`;
 * cme_mlp_step_ex also takes a side stream and 4 cudaEvent_t (disable-timing is fine), used only
 * by models that still zero the tail of the gradient vector (hidden width not a multiple of 256:
 * the memset then hides under the forward GEMM); both are capturable into a CUDA graph from
 * `stream`.  With several ranks every rank must issue the same steps. */
int cme_mlp_step(const CmeMlpPlan* plan, const uint8_t* x_u8, const int32_t* labels, int64_t rows,
                 float lr, float reg, float scale, float* loss_sum, void* stream);
/* x_u8 == NULL selects the indirect form: inputs, labels, lr, reg and scale are read on the
 * device from plan->step_args (the lr / reg / scale arguments of the call are ignored). */
int cme_mlp_set_step_args(const CmeMlpPlan* plan, const uint8_t* x_u8, const int32_t* labels, float lr,
                          float reg, float scale, void* stream);
int cme_mlp_step_ex(const CmeMlpPlan* plan, const uint8_t* x_u8, const int32_t* labels,
                    int64_t rows, float lr, float reg, float scale, float* loss_sum, void* stream,
                    void* side_stream, void* const* events);
/* dW1 GEMM and the update of the step in ONE kernel (the product path): the GEMM's CTAs carry
 * four extra "update warps" that reduce each unit of the gradient across the GPUs (multimem /
 * peer loads), apply the SGD step and replicate the new weights as soon as every rank has
 * published the unit, while the tensor cores work on later units.  Reads plan->dz1 / xb, writes
 * plan->grads, params, w1c.  indirect != 0: lr / reg from plan->step_args. */
int cme_mlp_dw1_update(const CmeMlpPlan* plan, int64_t rows, float lr, float reg, int indirect, void* stream);
/* The update of one step on its own: waits (on the device) for every unit of every rank's
 * gradient to be published, reduces this rank's slices over NVLink, applies
 * p -= lr * (g + reg * p), keeps the fp32 master of its slices and stores the bf16 compute copy
 * of W1 (and the fp32 W2 / b2) into every rank's replica; returns to the stream only when
 * every rank has written its slices.  Replaces the four MPI_Allreduce calls, the host SGD and
 * the weight re-upload (ref fpcode/neural_network.cpp:501-541, 486-489). */
int cme_mlp_update_stream(const CmeMlpPlan* plan, float lr, float reg, void* stream);
/* The building blocks of a step in their indirect form (step_args != NULL / indirect != 0:
 * inputs, labels, lr, reg, scale are read from CmeStepArgs in device memory, see above). */
int cme_mlp_update_stream_ind(const CmeMlpPlan* plan, float lr, float reg, int indirect, void* stream);
int cme_u8_to_bf16_rows_ind(const uint8_t* src, const CmeStepArgs* step_args, void* dst_bf16, int64_t rows,
                            int cols, int ld, int add_ones, void* stream);
/* The staging launch of a training step: uint8 rows -> bf16 rows with the ones column, and in
 * extra blocks of the same launch the bf16 copies of W2 (cme_mlp_recast_w2). */
int cme_mlp_stage_recast(const uint8_t* src, const CmeStepArgs* step_args, void* dst_bf16, int64_t rows, int cols,
                         int ld, const float* W2, void* w2c, int ldh, void* w2t, int H, int C, void* stream);
int cme_mlp_fwd2_tc_ind(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                        const CmeStepArgs* step_args, float* dz2, void* dz2b, float* loss_sum, int32_t* pred,
                        float* gb2, float scale, int64_t rows, int H, int ldh, int C, void* stream);
/* *out = 0, or the code of a wait that gave up (low byte 1: a unit of some rank never arrived
 * within plan->timeout_ms, 2: a rank never finished its update, 3: a k-chunk of the dW1 GEMM
 * never got its turn; bits 8.. = which).  After a
 * non-zero status the replicas are not consistent: tear the job down.  Synchronous copy. */
int cme_mlp_status(const CmeMlpPlan* plan, uint32_t* out);
/* out = {n_w1, o_w2, o_b2, n_total}: float offsets of the flat parameter / gradient vector the
 * step reads (W1aug at 0, W2 at o_w2, b2 at o_b2, n_total floats in all); rejects a plan whose
 * leading dimensions do not fit its sizes. */
int cme_mlp_layout(const CmeMlpPlan* plan, int64_t out[4]);
/* Float range [out[0], out[1]) of the flat vector whose fp32 master `rank` keeps inside update
 * unit `unit` (unit == number of W1 units: the W2 / b2 tail). */
int cme_mlp_unit_slice(const CmeMlpPlan* plan, int unit, int rank, int64_t out[2]);
/* One-rank convenience for C/C++ callers: allocates (cudaMalloc) and zeroes the flag pad, the
 * pad pointer table, the split-K counters and the step arguments, and sets unit_rows and
 * overlap.  Several ranks need peer-mapped pads (cme213_sp18_b200/parallel/dist.py builds them
 * from symmetric memory). */
int cme_mlp_plan_init_sync(CmeMlpPlan* plan);
int cme_mlp_plan_free_sync(CmeMlpPlan* plan);
/* Kernels one step launches (memsets not counted). */
int cme_mlp_step_launches(const CmeMlpPlan* plan);

/* Native streaming trainer (csrc/mlp_runtime.cu): the C++ counterpart of parallel_train's
 * whole minibatch loop (ref fpcode/neural_network.cpp:443-567) for data that lives in pinned
 * host memory.  create() allocates two device staging slots (max_rows x D uint8 + max_rows
 * int32 each), two streams and the events; it copies *plan (the buffers it names must outlive
 * the runtime).  run() issues `nsteps` steps without waiting for the device: step i trains on
 * rows [row_begin[i], row_begin[i] + rows[i]) of host_x [.][D] / host_labels (this rank's
 * slice, pinned), with gradient scale 1 / global_rows[i]; the H2D copy of step i+1 overlaps
 * step i; each step is replayed from a CUDA graph (use_graphs) recorded on first use; the
 * summed cross-entropy of step i is stored by the device into loss_host[i] (pinned and
 * device-mapped, i.e. cudaHostAlloc / cudaHostRegister memory; may be NULL).  Work is ordered
 * after everything already queued on `stream`, and `stream` waits for the last step; the host
 * arrays must stay valid until then.  With several ranks every rank runs the same steps.
 * A non-zero return means a step could not be queued: the call has then waited for everything
 * it did queue (the host arrays are free again), but with several ranks the peers are left
 * waiting for this rank's step and give up after plan->timeout_ms: tear all ranks down. */
typedef struct CmeMlpRuntime CmeMlpRuntime;
int cme_mlp_runtime_create(const CmeMlpPlan* plan, int use_graphs, CmeMlpRuntime** out);
int cme_mlp_runtime_run(CmeMlpRuntime* rt, const uint8_t* host_x, const int32_t* host_labels,
                        const int64_t* row_begin, const int64_t* rows, const int64_t* global_rows,
                        int nsteps, float lr, float reg, float* loss_host, void* stream);
/* out = {h2d bytes, d2h bytes, kernels launched, steps, graphs recorded} since create(). */
int cme_mlp_runtime_stats(const CmeMlpRuntime* rt, int64_t out[5]);
int cme_mlp_runtime_destroy(CmeMlpRuntime* rt);

/* Cross-GPU flag barrier over a peer-mapped pad of uint32 slots.  pads: DEVICE
 * array of nranks pointers (one pad per rank, >= 8*64 slots, zero-initialised).
 * Slot layout pads[dst][channel*64 + src].  do_signal: bump my slot on every
 * peer (release); do_wait: spin until every peer's slot on my pad has reached
 * `epoch` (acquire).  `epoch` counts barriers on that channel, starting at 1;
 * epoch == 0 means "use and advance the device-side counter in slot 63 of my own
 * channel row", which keeps the launch free of per-call host arguments (CUDA graphs).
 * A channel must be used either always with signal+wait together or split
 * consistently, so that every rank signals exactly once per epoch.
 * One rank per GPU only: ranks spinning on each other must not share a GPU. */
int cme_peer_barrier(uint32_t* const* pads, int rank, int nranks, int channel, uint32_t epoch,
                     int do_signal, int do_wait, void* stream);
/* P2P copy bandwidth probe: dst (local) <- src (peer), 128-bit loads. */
int cme_peer_copy(const void* src, void* dst, size_t bytes, void* stream);
/* The same copy with the grid capped at max_blocks (0 = no cap), for copies that run beside compute. */
int cme_peer_copy_bg(const void* src, void* dst, size_t bytes, int max_blocks, void* stream);
/* The same transfer on a copy engine (cudaMemcpyAsync through the peer mapping): makes progress
 * while persistent kernels occupy every SM. */
int cme_peer_memcpy(const void* src, void* dst, size_t bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CME213B_H_ */
