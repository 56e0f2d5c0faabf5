// tcgen05 / TMEM / TMA GEMM kernels for the bf16 MLP path (sm_100a).
//
// Two persistent, warp-specialised kernels replace the reference's three 32x32
// shared-memory fp64 GEMM kernels (ref fpcode/gpu_func.cu:92-256):
//
//   gemm_kmajor   C[M,N] = epi(A[M,K] * B[N,K]^T), both operands K-major.
//                 Forward layer 1: A = X (minibatch rows), B = W1c, epilogue =
//                 sigmoid, bf16 result written with TMA stores.  The bias rides
//                 in the GEMM: X carries a column of ones and W1c carries b1 in
//                 the matching column, so there is no bias pass (ref K8) and no
//                 separate sigmoid pass (ref K6).
//   gemm_mnmajor  G[M,N] += A[Kd,M]^T * B[Kd,N], both operands MN-major, the
//                 reduction dimension is the (huge) batch.  dW1 = dZ1^T * X.
//                 The batch is cut into a few k-chunks and the (k-chunk, m, n)
//                 items are dealt round-robin to 148 persistent CTAs, n fastest,
//                 so CTAs running together share operand rows through L2;
//                 partial tiles are added with vector fp32 reductions.  The ones column of X makes column K of the result
//                 the bias gradient (ref K7 row sums) for free.
//
// Roles per CTA (192 threads): warp 0 = TMA producer (one lane), warp 1 = MMA
// issuer (one lane; also owns the TMEM allocation), warps 2-5 = epilogue (each
// reads its 32-lane TMEM sub-partition).  Pipelines: smem full/empty ring
// (TMA <-> MMA) and a 2-deep TMEM accumulator ring (MMA <-> epilogue).
#include "common.cuh"
#include "tc_common.cuh"
#include "update_stream.cuh"
#include "tc_host.cuh"

#include <map>
#include <mutex>
#include <tuple>

namespace {

using namespace tc;

constexpr int kThreads = 192;
constexpr int kUpdWarps = 4;                          // update warps of the dW1 GEMMs (warps 6..9)
constexpr int kThreadsDw = kThreads + 32 * kUpdWarps; // launched with kThreads when there is no fused update
constexpr int kStages = 4;
constexpr int BM = 128;
constexpr int BK = 64;                       // bf16 elements = one 128-byte swizzle row
constexpr int kABytes = BM * BK * 2;         // 16 KB
constexpr int kBBytesMax = 256 * BK * 2;     // 32 KB
constexpr int kStageBytes = 64 * 128 * 2;    // 128 rows x 64 cols bf16 epilogue staging
constexpr int kSmemKMajor = kStages * (kABytes + kBBytesMax) + 2 * kStageBytes + 1024 + 512;
constexpr int kSmemSoftmax = 10 * (kABytes + 16 * BK * 2) + 1024 + 512 + 128 * 16 * 4;
constexpr int kSmemMNMajor = kStages * (kABytes + kBBytesMax) + 1024 + 512;

constexpr int kMaxStages = 10;
struct Barriers {
    uint64_t full[kMaxStages];
    uint64_t empty[kMaxStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint64_t x_full, x_empty;     // split-K exchange between the two CTAs of a cluster (softmax kernel)
    uint32_t tmem_base;
    uint32_t upd_ok;              // update warps of the dW1 GEMM (update_stream.cuh)
};

__device__ __forceinline__ uint8_t* align1024(uint8_t* p) {
    return (uint8_t*)(((uintptr_t)p + 1023) & ~(uintptr_t)1023);
}

__device__ __forceinline__ void setup(Barriers* bars, uint32_t tmem_cols, int warp, int lane,
                                      int stages = kStages) {
    if (warp == 0 && lane == 0) {
        for (int s = 0; s < stages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], 4);   // one arrive per epilogue warp
        }
        mbar_init(&bars->x_full, 4);
        mbar_init(&bars->x_empty, 4);
        fence_barrier_init();
    }
    if (warp == 1) {   // whole warp: .sync.aligned
        tmem_alloc(&bars->tmem_base, tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
}

__device__ __forceinline__ void teardown(Barriers* bars, uint32_t tmem_cols, int warp) {
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(bars->tmem_base, tmem_cols);
    }
}

// Arguments of the layer-2 epilogue (EPI == 2): the 16 accumulator columns of a
// row are its 10 logits; the epilogue finishes the forward pass in registers.
struct SoftmaxEpi {
    const float* b2;
    const int32_t* labels;
    float* dz2;              // [rows][16] fp32, may be null
    __nv_bfloat16* dz2b;     // [rows][64] bf16 copy for the tensor-core dW2 product, may be null
    float* loss_sum;         // may be null
    int32_t* pred;           // may be null
    float* gb2;              // may be null
    float scale;
    long long rows;
    int classes;
    const CmeStepArgs* sa;   // non-null: labels and scale come from the device-side step arguments
};

// =============================================================================
// K-major GEMM with fused epilogue and TMA store
// =============================================================================
template <int EPI /* 0 = identity, 1 = sigmoid, 2 = bias + softmax + cross-entropy gradient */>
__global__ void __launch_bounds__(kThreads, 1)
gemm_kmajor_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmC, int m_tiles, int n_tiles, int K, int BN,
                   uint32_t tmem_cols, SoftmaxEpi sm, int ksplit) {
    // The softmax variant (EPI == 2) has a 16-row B tile and no output staging: it spends its
    // shared memory on a deep ring instead (10 x 18 KB in flight per SM), because it is a pure
    // stream over A and per-SM bandwidth is set by the bytes in flight.
    constexpr int S = (EPI == 2) ? kMaxStages : kStages;
    constexpr int kBSlot = (EPI == 2) ? 16 * BK * 2 : kBBytesMax;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + S * kABytes;
    uint8_t* sC = sB + S * kBSlot;
    Barriers* bars = (Barriers*)(sC + (EPI == 2 ? 0 : 2 * kStageBytes));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        if (EPI != 2) prefetch_tmap(&tmC);
    }
    setup(bars, tmem_cols, warp, lane, S);
    const uint32_t tmem_base = bars->tmem_base;

    const int num_tiles = m_tiles * n_tiles;
    const int nk = (K + BK - 1) / BK;
    const uint32_t b_bytes = (uint32_t)BN * BK * 2;
    // ksplit == 2 (softmax kernel, few row tiles): the two CTAs of a cluster take the two halves
    // of K for the same row tile; CTA 1 hands its partial logits to CTA 0 through distributed
    // shared memory and CTA 0 finishes the row.
    const uint32_t cta = (EPI == 2 && ksplit == 2) ? cluster_ctarank() : 0u;
    const int tile_first = (ksplit == 2) ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int tile_step = (ksplit == 2) ? (int)(gridDim.x >> 1) : (int)gridDim.x;
    const int kb_lo = (ksplit == 2) ? (int)cta * (nk / 2) : 0;
    const int kb_hi = (ksplit == 2 && cta == 0) ? nk / 2 : nk;
    // 128 x 16 fp32 exchange tile (EPI == 2), 16-byte aligned for the vector stores
    float* xbuf = reinterpret_cast<float*>(((uintptr_t)(bars + 1) + 15) & ~(uintptr_t)15);
    if (ksplit == 2) cluster_sync_all();                  // peer barriers are initialised

    if (warp == 0) {
        if (lane == 0) {
            // ---------------- TMA producer ----------------
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
                const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
                for (int kb = kb_lo; kb < kb_hi; ++kb) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    mbar_expect_tx(&bars->full[stage], kABytes + b_bytes);
                    tma_load_2d(sA + stage * kABytes, &tmA, &bars->full[stage], kb * BK, m0);
                    tma_load_2d(sB + stage * kBSlot, &tmB, &bars->full[stage], kb * BK, n0);
                    if (++stage == S) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();   // reconverge before the CTA-wide barrier in teardown()
    } else if (warp == 1) {
        if (lane == 0) {
            // ---------------- MMA issuer ----------------
            const uint32_t idesc = idesc_bf16(BM, BN, 0, 0);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
                for (int kb = kb_lo; kb < kb_hi; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const int kvalid = min(BK, K - kb * BK);
                    const int nmma = (kvalid + 15) >> 4;
                    const uint64_t da = smem_desc_sw128(smem_u32(sA + stage * kABytes), 16, 1024);
                    const uint64_t db = smem_desc_sw128(smem_u32(sB + stage * kBSlot), 16, 1024);
                    for (int j = 0; j < nmma; ++j) {
                        // +32 bytes per 16-element K step inside the 128-byte swizzle row
                        umma_bf16(d_tmem, da + (uint64_t)(2 * j), db + (uint64_t)(2 * j), idesc,
                                  (kb > kb_lo) || (j > 0));
                    }
                    umma_commit(&bars->empty[stage]);
                    if (++stage == S) { stage = 0; phase ^= 1; }
                }
                umma_commit(&bars->tmem_full[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue: TMEM -> regs -> smem -> TMA store ----------------
        const int q = warp & 3;                    // TMEM sub-partition of this warp
        const int row = q * 32 + lane;             // row inside the 128-row tile
        const bool leader = (warp == 2 && lane == 0);
        int acc = 0;
        uint32_t acc_phase = 0, xphase = 0;
        int cbuf = 0;
        for (int tile = tile_first; tile < num_tiles; tile += tile_step) {
            const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
            if constexpr (EPI == 2) {
                uint32_t v[16];
                tmem_ld_32x16(t_row, v);
                tmem_ld_wait();
                // logits are in registers: give the accumulator back before the math
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                if (ksplit == 2) {
                    if (cta == 1) {
                        // hand my half of the K sum to CTA 0 (its buffer is free again once it
                        // has arrived on my x_empty), then go on to the next tile
                        mbar_wait_cluster(&bars->x_empty, xphase ^ 1);
                        const uint32_t dst = map_to_cta(smem_u32(xbuf + row * 16), 0);
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4)
                            st_cluster_f4(dst + k4 * 16, __uint_as_float(v[4 * k4]), __uint_as_float(v[4 * k4 + 1]),
                                          __uint_as_float(v[4 * k4 + 2]), __uint_as_float(v[4 * k4 + 3]));
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&bars->x_full), 0));
                        xphase ^= 1;
                        continue;
                    }
                    mbar_wait_cluster(&bars->x_full, xphase);
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {
                        const float4 o = *reinterpret_cast<const float4*>(xbuf + row * 16 + 4 * k4);
                        v[4 * k4] = __float_as_uint(__uint_as_float(v[4 * k4]) + o.x);
                        v[4 * k4 + 1] = __float_as_uint(__uint_as_float(v[4 * k4 + 1]) + o.y);
                        v[4 * k4 + 2] = __float_as_uint(__uint_as_float(v[4 * k4 + 2]) + o.z);
                        v[4 * k4 + 3] = __float_as_uint(__uint_as_float(v[4 * k4 + 3]) + o.w);
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(map_to_cta(smem_u32(&bars->x_empty), 1));
                    xphase ^= 1;
                }
                const long long grow = (long long)m0 + row;
                const bool valid = grow < sm.rows;
                const int C = sm.classes;
                const int32_t* labels = sm.sa ? sm.sa->labels : sm.labels;
                const float scale = sm.sa ? sm.sa->scale : sm.scale;
                float z[16];
                float mx = -3.0e38f;
                int arg = 0;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    z[c] = (c < C) ? __uint_as_float(v[c]) + __ldg(sm.b2 + c) : -3.0e38f;
                    if (z[c] > mx) { mx = z[c]; arg = c; }
                }
                float ssum = 0.f;
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    z[c] = (c < C) ? __expf(z[c] - mx) : 0.f;
                    ssum += z[c];
                }
                const float inv = 1.f / ssum;
                const int lab = (valid && labels) ? labels[grow] : -1;
                float my_loss = 0.f;
                float out[16];
#pragma unroll
                for (int c = 0; c < 16; ++c) {
                    const float p = z[c] * inv;
                    out[c] = (valid && c < C) ? (p - (c == lab ? 1.f : 0.f)) * scale : 0.f;
                    if (c == lab) my_loss = -__logf(fmaxf(p, 1e-30f));
                }
                if (valid) {
                    if (sm.dz2) {
                        float4* d = reinterpret_cast<float4*>(sm.dz2 + grow * 16);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            d[k] = make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
                    }
                    if (sm.dz2b) {
                        uint4* d = reinterpret_cast<uint4*>(sm.dz2b + grow * 64);
                        d[0] = make_uint4(pack_bf16(out[0], out[1]), pack_bf16(out[2], out[3]),
                                          pack_bf16(out[4], out[5]), pack_bf16(out[6], out[7]));
                        d[1] = make_uint4(pack_bf16(out[8], out[9]), pack_bf16(out[10], out[11]),
                                          pack_bf16(out[12], out[13]), pack_bf16(out[14], out[15]));
                    }
                    if (sm.pred) sm.pred[grow] = arg;
                }
                if (sm.gb2) {
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        const float t = cme::warp_sum(out[c]);
                        if (lane == 0 && c < C && t != 0.f) atomicAdd(sm.gb2 + c, t);
                    }
                }
                if (sm.loss_sum) {
                    const float t = cme::warp_sum(my_loss);
                    if (lane == 0 && t != 0.f) atomicAdd(sm.loss_sum, t);
                }
                continue;
            }
            for (int c = 0; c < BN / 64; ++c) {
                // the staging buffer we are about to fill was handed to a TMA store
                // two chunks ago: wait until that store has finished reading it
                if (leader) tma_store_wait_read<1>();
                named_bar_sync(1, 128);
                uint8_t* stg = sC + cbuf * kStageBytes;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t v[32];
                    tmem_ld_32x32(t_row + (uint32_t)(c * 64 + half * 32), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {   // 4 x 16-byte chunks (8 bf16 each)
                        uint32_t p[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float x0 = __uint_as_float(v[g * 8 + 2 * e]);
                            float x1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
                            if (EPI == 1) { x0 = sigmoid_fast(x0); x1 = sigmoid_fast(x1); }
                            p[e] = pack_bf16(x0, x1);
                        }
                        const int chunk = half * 4 + g;                 // 16-byte chunk in the row
                        const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
                        *reinterpret_cast<uint4*>(stg + off) = make_uint4(p[0], p[1], p[2], p[3]);
                    }
                }
                fence_proxy_async_smem();
                named_bar_sync(1, 128);
                if (leader) {
                    tma_store_2d(&tmC, stg, n0 + c * 64, m0);
                    tma_store_commit();
                }
                cbuf ^= 1;
            }
            // accumulator fully read: hand the TMEM stage back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        if (EPI != 2 && leader) tma_store_wait_all<0>();
    }
    if (ksplit == 2) cluster_sync_all();     // nobody leaves while the peer may still touch its smem
    teardown(bars, tmem_cols, warp);
}

// =============================================================================
// K-major GEMM on CTA pairs (cta_group::2): 256 x 256 tiles over two SMs
// =============================================================================
// Both 1-CTA GEMMs are limited by operand delivery from L2 to the SMs (ncu: ~15.4 TB/s of
// L2->SM traffic in both, tensor pipe 55-65 % busy).  A pair of CTAs on one tile halves the B
// bytes each SM has to receive: CTA r loads its own 128 rows of A and rows r*128.. of the
// 256-row B tile, the leader issues tcgen05.mma.cta_group::2 with M = 256, and the hardware
// reads the two halves of B from both CTAs' shared memory.  Each CTA keeps its 128 x 256
// accumulator in its own TMEM and runs its own epilogue.
constexpr int kStages2 = 6;                  // MN-major pair kernel: no output staging
constexpr int kStagesK2 = 6;                 // K-major pair kernel: six stages + two output staging buffers
constexpr int kThreadsK2 = 320;              // producer warp + MMA warp + 8 epilogue warps
constexpr int kB2Bytes = 128 * BK * 2;       // this CTA's half of the B tile: 16 KB
constexpr int kSmemKMajor2 = kStagesK2 * (kABytes + kB2Bytes) + 2 * kStageBytes + 1024 + 256;
constexpr int kSmemMNMajor2 = kStages2 * (kABytes + kB2Bytes) + 1024 + 256;

struct Barriers2 {
    uint64_t full[kStages2];      // used in the leader: TMA bytes of both CTAs
    uint64_t empty[kStages2];     // one multicast commit per consumption, in both CTAs
    uint64_t tmem_full[2];        // multicast commit, in both CTAs
    uint64_t tmem_empty[2];       // used in the leader: 4 epilogue warps x 2 CTAs
    uint32_t tmem_base;
    uint32_t upd_ok;              // update warps of the dW1 GEMM (update_stream.cuh)
};

// Epilogue: TMEM -> sigmoid -> bf16 -> swizzled smem -> TMA store of 128 x 256 outputs per CTA
// and tile.  Eight epilogue warps work as two groups, each owning two of the four 64-column
// chunks with its own staging buffer, named barrier and TMA-store leader; the sigmoid is
// evaluated two at a time (tanh.approx.bf16x2).  The main loop, not the epilogue, sets the
// pace (profiles/fwd1_probe.md): with every epilogue instruction removed the 65536 x 8192 x 785
// product still takes 0.61-0.63 ms (0.70-0.71 with the epilogue; cuBLAS 0.63), and that floor
// does not move with the ring depth (5 or 6 stages), the operand layout, the accumulator
// hand-over between tiles, or even when every load hits the same lines: 1.43 PFLOP/s, the rate
// this board sustains on bf16 MMAs under its power limit (cuBLAS: 1.38 PFLOP/s).
template <int EPI /* 0 = identity, 1 = sigmoid */>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsK2, 1)
gemm_kmajor2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, int m_tiles, int n_tiles, int K) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + kStagesK2 * kABytes;
    uint8_t* sC = sB + kStagesK2 * kB2Bytes;
    Barriers2* bars = (Barriers2*)(sC + 2 * kStageBytes);
    constexpr int BN = 256;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();          // 0 = leader
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmC);
        for (int s = 0; s < kStagesK2; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], 16);   // 8 epilogue warps x 2 CTAs
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_2cta(&bars->tmem_base, 512);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    const int num_tiles = m_tiles * n_tiles;         // tiles of 256 x 256
    const int nk = (K + BK - 1) / BK;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = pair; tile < num_tiles; tile += npairs) {
                const int m0 = (tile / n_tiles) * 256 + (int)cta * 128;
                const int n0 = (tile % n_tiles) * BN + (int)cta * 128;
                for (int kb = 0; kb < nk; ++kb) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    if (cta == 0) mbar_expect_tx(&bars->full[stage], 2 * (kABytes + kB2Bytes));
                    tma_load_2d_2cta(sA + stage * kABytes, &tmA, &bars->full[stage], kb * BK, m0);
                    tma_load_2d_2cta(sB + stage * kB2Bytes, &tmB, &bars->full[stage], kb * BK, n0);
                    if (++stage == kStagesK2) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0 && cta == 0) {
            const uint32_t idesc = idesc_bf16(256, BN, 0, 0);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = pair; tile < num_tiles; tile += npairs) {
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
                for (int kb = 0; kb < nk; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const int kvalid = min(BK, K - kb * BK);
                    const int nmma = (kvalid + 15) >> 4;
                    const uint64_t da = smem_desc_sw128(smem_u32(sA + stage * kABytes), 16, 1024);
                    const uint64_t db = smem_desc_sw128(smem_u32(sB + stage * kB2Bytes), 16, 1024);
                    for (int j = 0; j < nmma; ++j)
                        umma_bf16_2cta(d_tmem, da + (uint64_t)(2 * j), db + (uint64_t)(2 * j), idesc,
                                       (kb | j) != 0);
                    umma_commit_2cta(&bars->empty[stage]);
                    if (++stage == kStagesK2) { stage = 0; phase ^= 1; }
                }
                umma_commit_2cta(&bars->tmem_full[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3;                          // TMEM sub-partition of this warp
        const int grp = (warp - 2) >> 2;                 // 0: chunks 0,1   1: chunks 2,3
        const int row = q * 32 + lane;
        const bool leader = (((warp - 2) & 3) == 0 && lane == 0);
        uint8_t* stg = sC + grp * kStageBytes;           // one staging buffer per group
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = pair; tile < num_tiles; tile += npairs) {
            const int m0 = (tile / n_tiles) * 256 + (int)cta * 128;
            const int n0 = (tile % n_tiles) * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll
            for (int cc = 0; cc < 2; ++cc) {
                const int c = grp * 2 + cc;
                uint32_t v0[32], v1[32];
                tmem_ld_32x32(t_row + (uint32_t)(c * 64), v0);
                tmem_ld_32x32(t_row + (uint32_t)(c * 64 + 32), v1);
                tmem_ld_wait();
                // 64 outputs of this thread's row, packed, before the staging buffer is needed
                uint32_t pk[32];
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                    const float a0 = __uint_as_float(v0[2 * e]), a1 = __uint_as_float(v0[2 * e + 1]);
                    const float b0 = __uint_as_float(v1[2 * e]), b1 = __uint_as_float(v1[2 * e + 1]);
                    pk[e] = (EPI == 1) ? sigmoid2_bf16(a0, a1) : pack_bf16(a0, a1);
                    pk[16 + e] = (EPI == 1) ? sigmoid2_bf16(b0, b1) : pack_bf16(b0, b1);
                }
                if (leader) tma_store_wait_read<0>();    // the previous chunk has left the buffer
                named_bar_sync(1 + grp, 128);
#pragma unroll
                for (int chunk = 0; chunk < 8; ++chunk) {
                    const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
                    *reinterpret_cast<uint4*>(stg + off) =
                        make_uint4(pk[4 * chunk], pk[4 * chunk + 1], pk[4 * chunk + 2], pk[4 * chunk + 3]);
                }
                fence_proxy_async_smem();
                named_bar_sync(1 + grp, 128);
                if (leader) {
                    tma_store_2d(&tmC, stg, n0 + c * 64, m0);
                    tma_store_commit();
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&bars->tmem_empty[acc]);   // the leader's MMA lane waits for both CTAs
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        if (leader) tma_store_wait_all<0>();
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, 512);
    }
}

// =============================================================================
// Ordered split-K: how the k-chunks of one output tile are combined
// =============================================================================
// The partial tiles of a split-K product are combined in chunk order, so the fp32 result does
// not depend on which CTA finished first: chunk 0 STORES its partial tile (no zero-fill pass
// over the gradient buffer), chunk c > 0 waits until chunk c - 1 of the same (tile, CTA) part
// has landed (a per-part counter, acquire/release at gpu scope) and adds on top with vector
// reductions.  Items are dealt (super-block, k-chunk, m, n) round-robin to persistent CTAs that
// are all resident, so the part a CTA waits for was started a whole chunk phase earlier and is
// always in progress on a running CTA.  When the last chunk of a part is in, the part counts
// towards its row block ("unit": every row of W1 the block covers, all column tiles); the CTA
// that completes a unit bumps that unit's flag on every rank's pad, which is what the update
// warps (update_stream.cuh) wait for before they reduce the unit across GPUs.
struct DwSync {
    uint32_t* part_ctr;          // [tiles * ctas_per_tile], zero between launches
    uint32_t* unit_ctr;          // [units], zero between launches; null = no unit signalling
    uint32_t* const* pads;       // device array of nranks pad pointers; null = no signalling
    uint32_t* status;            // this rank's status word (cme_mlp_status); null = trap on a stuck wait
    int nranks;
    int parts_per_tile;          // (tile, CTA) parts per row block of tiles: column tiles x CTAs per tile
    int tiles_per_unit;          // row blocks of tiles per unit (a unit is unit_rows rows of the result)
    int m_tiles;
    int sblocks;                 // row blocks are walked in this many groups (see dw_item)
};

struct DwItem {
    int kc, m, n;
};
// item -> (k-chunk, row block m, column tile n).  Row blocks are cut into `sblocks` groups of
// msb; inside a group the order is k-chunk, then m, then n (fastest), so CTAs that run together
// read the same k rows of B and neighbouring rows of A through L2, and a group's row blocks are
// complete before the next group starts.
__device__ __forceinline__ DwItem dw_item(int item, int m_tiles, int n_tiles, int ksplit, int sblocks) {
    const int msb = (m_tiles + sblocks - 1) / sblocks;
    const int per_sb = msb * n_tiles * ksplit;
    int sb = item / per_sb, r = item - sb * per_sb, mcount = msb;
    if (sb >= sblocks - 1) {                 // the last group may hold fewer row blocks
        sb = sblocks - 1;
        r = item - sb * per_sb;
        mcount = m_tiles - sb * msb;
    }
    DwItem it;
    it.kc = r / (mcount * n_tiles);
    const int rr = r - it.kc * (mcount * n_tiles);
    it.m = sb * msb + rr / n_tiles;
    it.n = rr % n_tiles;
    return it;
}

// lane 0 polls, the warp follows.  A wait that outlives ~2 s of SM clocks means a CTA that owns an
// earlier chunk is not running (the grid is not fully resident): record it in the status word
// (cme_mlp_status code 3) and go on, or trap when there is no status word, rather than hang.
__device__ __forceinline__ void dw_wait_turn(const DwSync& sy, const uint32_t* ctr, int kc, int lane) {
    if (kc > 0) {
        if (lane == 0) {
            const long long t0 = clock64();
            while (cme::ld_relaxed_gpu(ctr) != (uint32_t)kc) {
                if (clock64() - t0 > (4ll << 30)) {
                    if (!sy.status) __trap();
                    atomicCAS(sy.status, 0u, 3u | ((uint32_t)kc << 8) | ((uint32_t)(ctr - sy.part_ctr) << 12));
                    break;
                }
            }
            // no acquire fence: what follows are reductions performed at L2, which apply in
            // arrival order behind the stores the previous chunk released before it set the counter
        }
        __syncwarp();
    }
}

// called by ONE epilogue thread after a CTA-level barrier over the epilogue warps
__device__ __forceinline__ void dw_publish(const DwSync& sy, uint32_t* ctr, int kc, int ksplit, int m) {
    if (kc + 1 < ksplit) {
        cme::st_release_gpu(ctr, (uint32_t)(kc + 1));      // cumulative: the barrier ordered the CTA's writes
        return;
    }
    __threadfence();
    *ctr = 0;                                              // ready for the next launch
    if (!sy.unit_ctr) return;
    const int unit = m / sy.tiles_per_unit;
    const int parts = min(sy.tiles_per_unit, sy.m_tiles - unit * sy.tiles_per_unit) * sy.parts_per_tile;
    const uint32_t old = atomicAdd(sy.unit_ctr + unit, 1u);
    if ((int)old + 1 == parts) {
        sy.unit_ctr[unit] = 0;
        if (sy.pads) {
            // one fence, then the bumps (see red_relaxed_sys_add); a single rank's readers are on this GPU
            if (sy.nranks > 1) __threadfence_system();
            else __threadfence();
            for (int r = 0; r < sy.nranks; ++r) cme::red_relaxed_sys_add(sy.pads[r] + cme::kPadUnit0 + unit, 1u);
        }
    }
}

__device__ __forceinline__ void dw_store4(float* p, bool first, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    if (first)
        asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__uint_as_float(a)),
                     "f"(__uint_as_float(b)), "f"(__uint_as_float(c)), "f"(__uint_as_float(d))
                     : "memory");
    else
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(__uint_as_float(a)),
                     "f"(__uint_as_float(b)), "f"(__uint_as_float(c)), "f"(__uint_as_float(d))
                     : "memory");
}

// =============================================================================
// MN-major split-K GEMM with fp32 reduction epilogue
// =============================================================================
__global__ void __launch_bounds__(kThreadsDw, 1)
gemm_mnmajor_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                    int kb_total, int BN, uint32_t tmem_cols, int out_transposed, int ksplit, DwSync sy,
                    const __grid_constant__ upd::UpdateArgs ua) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + kStages * kABytes;
    Barriers* bars = (Barriers*)(sB + kStages * kBBytesMax);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
    }
    setup(bars, tmem_cols, warp, lane);
    const uint32_t tmem_base = bars->tmem_base;

    // Work items are (k-chunk, m-tile, n-tile) with n fastest, then m, then the k-chunk (inside a
    // group of row blocks, see dw_item), handed out round-robin.  CTAs that run at the same time
    // therefore sit on the same k-chunk: the n-tiles of one m-tile read the same A rows and every
    // CTA reads the same B rows, so each operand byte comes out of HBM once and is served to the
    // other CTAs by L2.
    const int tiles_mn = m_tiles * n_tiles;
    const int items = tiles_mn * ksplit;
    const int nb_boxes = (BN + 63) / 64;
    const uint32_t b_bytes = (uint32_t)nb_boxes * 64 * BK * 2;
    const int sblocks = sy.part_ctr ? sy.sblocks : 1;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const DwItem it = dw_item(item, m_tiles, n_tiles, ksplit, sblocks);
                const int kc = it.kc;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                const int m0 = it.m * BM, n0 = it.n * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    mbar_expect_tx(&bars->full[stage], kABytes + b_bytes);
                    uint8_t* a = sA + stage * kABytes;
                    uint8_t* b = sB + stage * kBBytesMax;
                    // boxes are {64 MN elements} x {64 k rows}: 8 KB each
                    tma_load_2d(a, &tmA, &bars->full[stage], m0, kb * BK);
                    tma_load_2d(a + 8192, &tmA, &bars->full[stage], m0 + 64, kb * BK);
                    for (int x = 0; x < nb_boxes; ++x)
                        tma_load_2d(b + x * 8192, &tmB, &bars->full[stage], n0 + x * 64, kb * BK);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = idesc_bf16(BM, BN, 1, 1);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int kc = dw_item(item, m_tiles, n_tiles, ksplit, sblocks).kc;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(sA + stage * kABytes);
                    const uint32_t b0 = smem_u32(sB + stage * kBBytesMax);
#pragma unroll
                    for (int j = 0; j < BK / 16; ++j) {
                        // MN-major: 64-element MN groups are 8 KB apart (LBO), 8-row k
                        // groups 1 KB apart (SBO); one MMA consumes two k groups
                        const uint64_t da = smem_desc_sw128(a0 + j * 2048, 8192, 1024);
                        const uint64_t db = smem_desc_sw128(b0 + j * 2048, 8192, 1024);
                        umma_bf16(d_tmem, da, db, idesc, (kb > kb0) || (j > 0));
                    }
                    umma_commit(&bars->empty[stage]);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(&bars->tmem_full[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
        __syncwarp();
    } else if (warp < 6) {
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const DwItem it = dw_item(item, m_tiles, n_tiles, ksplit, sblocks);
            const int m0 = it.m * BM, n0 = it.n * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const int h = m0 + q * 32 + lane;
            float* grow = G + (size_t)h * ldg;
            if (sy.part_ctr) {
                // ordered mode: G = sum of the k-chunks in chunk order, every column below ldg
                // written (columns past N receive the zeros TMA filled in)
                uint32_t* ctr = sy.part_ctr + (it.m * n_tiles + it.n);
                dw_wait_turn(sy, ctr, it.kc, lane);
                for (int c = 0; c < BN / 16; ++c) {
                    if (n0 + c * 16 >= ldg) break;
                    uint32_t v[16];
                    tmem_ld_32x16(t_row + (uint32_t)(c * 16), v);
                    tmem_ld_wait();
                    if (h < M) {
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const int d = n0 + c * 16 + g * 4;
                            if (d < ldg) dw_store4(grow + d, it.kc == 0, v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                named_bar_sync(1, 128);
                if (warp == 2 && lane == 0) dw_publish(sy, ctr, it.kc, ksplit, it.m);
                continue;
            }
            for (int c = 0; c < BN / 16; ++c) {
                uint32_t v[16];
                tmem_ld_32x16(t_row + (uint32_t)(c * 16), v);
                tmem_ld_wait();
                if (h < M && out_transposed) {
                    // G[d][h]: consecutive lanes hit consecutive addresses
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const int d = n0 + c * 16 + e;
                        if (d < N) atomicAdd(G + (size_t)d * ldg + h, __uint_as_float(v[e]));
                    }
                } else if (h < M) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int d = n0 + c * 16 + g * 4;
                        if (d + 3 < N) {
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(grow + d),
                                         "f"(__uint_as_float(v[g * 4])), "f"(__uint_as_float(v[g * 4 + 1])),
                                         "f"(__uint_as_float(v[g * 4 + 2])), "f"(__uint_as_float(v[g * 4 + 3]))
                                         : "memory");
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e)
                                if (d + e < N) atomicAdd(grow + d + e, __uint_as_float(v[g * 4 + e]));
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    } else if (ua.enabled) {
        // update warps: reduce every unit of the gradient across the GPUs as soon as all ranks
        // have published it, while the warps above are still computing later units
        upd::run_update(ua, blockIdx.x, gridDim.x, threadIdx.x - kThreads, 32 * kUpdWarps, 2, &bars->upd_ok);
    }
    teardown(bars, tmem_cols, warp);
}

// =============================================================================
// MN-major split-K GEMM on CTA pairs: 256 x 256 tiles, fp32 reduction epilogue
// =============================================================================
// Same operand layout as gemm_mnmajor_kernel; the pair covers 256 rows of M (CTA r owns rows
// r*128..) and 256 columns of N, of which each CTA stages 128.  Items are (k-chunk, m, n) dealt
// round-robin to the pairs, n fastest.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsDw, 1)
gemm_mnmajor2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                     int kb_total, int ksplit, int BN, DwSync sy,
                     const __grid_constant__ upd::UpdateArgs ua) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + kStages2 * kABytes;
    Barriers2* bars = (Barriers2*)(sB + kStages2 * kB2Bytes);
    // BN (a multiple of 16, <= 256) is the width of the pair's tile; each CTA stages BN/2 of
    // the columns (one or two 64-wide boxes) and keeps its accumulators 256 TMEM columns apart
    const int half_n = BN / 2;
    const int nb_boxes = (half_n + 63) / 64;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        for (int s = 0; s < kStages2; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], 8);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc_2cta(&bars->tmem_base, 512);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;

    const int tiles_mn = m_tiles * n_tiles;
    const int items = tiles_mn * ksplit;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int item = pair; item < items; item += npairs) {
                const DwItem it = dw_item(item, m_tiles, n_tiles, ksplit, sy.sblocks);
                const int kc = it.kc;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                const int m0 = it.m * 256 + (int)cta * 128;
                const int n0 = it.n * BN + (int)cta * half_n;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->empty[stage], phase ^ 1);
                    if (cta == 0)
                        mbar_expect_tx(&bars->full[stage], 2 * (kABytes + (uint32_t)nb_boxes * 8192u));
                    uint8_t* a = sA + stage * kABytes;
                    uint8_t* b = sB + stage * kB2Bytes;
                    tma_load_2d_2cta(a, &tmA, &bars->full[stage], m0, kb * BK);
                    tma_load_2d_2cta(a + 8192, &tmA, &bars->full[stage], m0 + 64, kb * BK);
                    for (int x = 0; x < nb_boxes; ++x)
                        tma_load_2d_2cta(b + x * 8192, &tmB, &bars->full[stage], n0 + x * 64, kb * BK);
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0 && cta == 0) {
            const uint32_t idesc = idesc_bf16(256, BN, 1, 1);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int item = pair; item < items; item += npairs) {
                const int kc = dw_item(item, m_tiles, n_tiles, ksplit, sy.sblocks).kc;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(sA + stage * kABytes);
                    const uint32_t b0 = smem_u32(sB + stage * kB2Bytes);
#pragma unroll
                    for (int j = 0; j < BK / 16; ++j) {
                        const uint64_t da = smem_desc_sw128(a0 + j * 2048, 8192, 1024);
                        const uint64_t db = smem_desc_sw128(b0 + j * 2048, 8192, 1024);
                        umma_bf16_2cta(d_tmem, da, db, idesc, (kb > kb0) || (j > 0));
                    }
                    umma_commit_2cta(&bars->empty[stage]);
                    if (++stage == kStages2) { stage = 0; phase ^= 1; }
                }
                umma_commit_2cta(&bars->tmem_full[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
        __syncwarp();
    } else if (warp < 6) {
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = pair; item < items; item += npairs) {
            const DwItem it = dw_item(item, m_tiles, n_tiles, ksplit, sy.sblocks);
            const int m0 = it.m * 256 + (int)cta * 128;
            const int n0 = it.n * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const int h = m0 + q * 32 + lane;
            float* grow = G + (size_t)h * ldg;
            // this CTA's half (128 rows) of the tile is one "part": chunk 0 stores, later chunks
            // wait for their turn and add (see "Ordered split-K" above)
            uint32_t* ctr = sy.part_ctr + ((it.m * n_tiles + it.n) * 2 + (int)cta);
            dw_wait_turn(sy, ctr, it.kc, lane);
            for (int c = 0; c < BN / 16; ++c) {
                if (n0 + c * 16 >= ldg) break;        // warp-uniform: nothing left in this row block
                uint32_t v[16];
                tmem_ld_32x16(t_row + (uint32_t)(c * 16), v);
                tmem_ld_wait();
                if (h < M) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int d = n0 + c * 16 + g * 4;
                        if (d < ldg) dw_store4(grow + d, it.kc == 0, v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(&bars->tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
            named_bar_sync(1, 128);
            if (warp == 2 && lane == 0) dw_publish(sy, ctr, it.kc, ksplit, it.m);
        }
    } else if (ua.enabled) {
        // update warps (see gemm_mnmajor_kernel): every CTA of the GEMM takes a share of every unit
        upd::run_update(ua, blockIdx.x, gridDim.x, threadIdx.x - kThreads, 32 * kUpdWarps, 2, &bars->upd_ok);
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, 512);
    }
}

// =============================================================================
// Backward through the sigmoid: dZ1 = (dz2 * W2) .* a .* (1 - a)
// =============================================================================
// The product has K = 16 (the classes), so one MMA fills a 128 x 256 TMEM tile and
// the kernel is a stream over A1: the producer lane prefetches A1 chunks
// (128 rows x 64 columns, TMA, 128-byte swizzle) into a ring of staging buffers,
// the epilogue warps combine them with the accumulator in place and hand the
// buffers to TMA stores.  Replaces the CUDA-core pass (ref K2 + K10,
// fpcode/neural_network.cpp:365-378) and keeps HBM as the only limit.
constexpr int kDzStages = 2;                 // operand stages (tiny operands)
constexpr int kDzBufs = 6;                   // A1 / dZ1 staging ring
constexpr int kDzPending = 2;                // TMA stores allowed in flight
constexpr int kSmemDz = kDzStages * (kABytes + kBBytesMax) + kDzBufs * kStageBytes + 1024 + 512;

struct DzBarriers {
    uint64_t full[kDzStages];
    uint64_t empty[kDzStages];
    uint64_t tmem_full[2];
    uint64_t tmem_empty[2];
    uint64_t ld_full[kDzBufs];
    uint64_t ld_empty[kDzBufs];
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
bwd_dz1_kernel(const __grid_constant__ CUtensorMap tmD /* dz2b  [rows][16 of 64] */,
               const __grid_constant__ CUtensorMap tmW /* W2T   [H][16 of 64]    */,
               const __grid_constant__ CUtensorMap tmA1 /* A1   [rows][H] load   */,
               const __grid_constant__ CUtensorMap tmZ /* dZ1   [rows][H] store  */,
               int m_tiles, int n_tiles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + kDzStages * kABytes;
    uint8_t* sC = sB + kDzStages * kBBytesMax;
    DzBarriers* bars = (DzBarriers*)(sC + kDzBufs * kStageBytes);
    constexpr int BN = 256;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmD);
        prefetch_tmap(&tmW);
        prefetch_tmap(&tmA1);
        prefetch_tmap(&tmZ);
        for (int s = 0; s < kDzStages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], 4);
        }
        for (int b = 0; b < kDzBufs; ++b) {
            mbar_init(&bars->ld_full[b], 1);
            mbar_init(&bars->ld_empty[b], 1);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(&bars->tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    const int num_tiles = m_tiles * n_tiles;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0, buf = 0;
            uint32_t phase = 0, bphase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
                mbar_wait(&bars->empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->full[stage], kABytes + kBBytesMax);
                tma_load_2d(sA + stage * kABytes, &tmD, &bars->full[stage], 0, m0);
                tma_load_2d(sB + stage * kBBytesMax, &tmW, &bars->full[stage], 0, n0);
                if (++stage == kDzStages) { stage = 0; phase ^= 1; }
                for (int c = 0; c < BN / 64; ++c) {
                    mbar_wait(&bars->ld_empty[buf], bphase ^ 1);
                    mbar_expect_tx(&bars->ld_full[buf], kStageBytes);
                    tma_load_2d(sC + buf * kStageBytes, &tmA1, &bars->ld_full[buf], n0 + c * 64, m0);
                    if (++buf == kDzBufs) { buf = 0; bphase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = idesc_bf16(BM, BN, 0, 0);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                mbar_wait(&bars->full[stage], phase);
                tc_fence_after();
                const uint64_t da = smem_desc_sw128(smem_u32(sA + stage * kABytes), 16, 1024);
                const uint64_t db = smem_desc_sw128(smem_u32(sB + stage * kBBytesMax), 16, 1024);
                umma_bf16(tmem_base + (uint32_t)(acc * BN), da, db, idesc, 0);   // K = 16: one MMA
                umma_commit(&bars->empty[stage]);
                umma_commit(&bars->tmem_full[acc]);
                if (++stage == kDzStages) { stage = 0; phase ^= 1; }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
        }
        __syncwarp();
    } else {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const bool leader = (warp == 2 && lane == 0);
        int acc = 0, buf = 0;
        uint32_t acc_phase = 0, bphase = 0;
        long long chunk_no = 0;            // chunks stored so far (leader bookkeeping)
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
            for (int c = 0; c < BN / 64; ++c) {
                mbar_wait(&bars->ld_full[buf], bphase);
                uint8_t* stg = sC + buf * kStageBytes;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t v[32];
                    tmem_ld_32x32(t_row + (uint32_t)(c * 64 + half * 32), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int chunk = half * 4 + g;
                        const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
                        uint4 a4 = *reinterpret_cast<const uint4*>(stg + off);
                        uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
                        uint32_t p[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float a0 = __uint_as_float(aw[e] << 16);
                            const float a1 = __uint_as_float(aw[e] & 0xffff0000u);
                            const float g0 = __uint_as_float(v[g * 8 + 2 * e]);
                            const float g1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
                            p[e] = pack_bf16(g0 * a0 * (1.f - a0), g1 * a1 * (1.f - a1));
                        }
                        *reinterpret_cast<uint4*>(stg + off) = make_uint4(p[0], p[1], p[2], p[3]);
                    }
                }
                fence_proxy_async_smem();
                named_bar_sync(1, 128);
                if (leader) {
                    tma_store_2d(&tmZ, stg, n0 + c * 64, m0);
                    tma_store_commit();
                    // stores older than the last kDzPending have finished reading their
                    // buffer: give that buffer back to the producer
                    tma_store_wait_read<kDzPending>();
                    if (chunk_no >= kDzPending)
                        mbar_arrive(&bars->ld_empty[(int)((chunk_no - kDzPending) % kDzBufs)]);
                    ++chunk_no;
                }
                if (++buf == kDzBufs) { buf = 0; bphase ^= 1; }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
        if (leader) tma_store_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// =============================================================================
// Fused backward tail: dZ1 = (dz2 * W2) .* a .* (1 - a)  AND  gW2 += dz2^T * A1
// =============================================================================
// Both need one pass over A1, so they share it.  A tile is 128 batch rows x 128 hidden
// columns = two A1 chunks (128 x 64, TMA, 128-byte swizzle) in adjacent staging buffers.
//   MMA1  g[128 b][128 h] = dz2b[128 b][16] * W2T[128 h][16]^T      (K = 16, one instruction)
//   MMA2  D2[128 h][16 c] += A1_tile^T[128 h][128 b] * dz2b[128 b][16 c]   (8 instructions)
// MMA2 reads the A1 chunks as an MN-major operand straight out of the staging ring and
// the dz2b tile (already in shared memory for MMA1) as its B operand; the epilogue then
// overwrites the chunks in place with dZ1 and TMA-stores them.  Each CTA walks a
// contiguous range of tiles in column-major order, so D2 stays in TMEM across all the
// row tiles of one column block and is flushed (transposed, fp32 reductions) only when
// the column block changes.
constexpr int kFtBufs = 8;                    // A1 / dZ1 staging ring, used in aligned pairs
constexpr int kFtPending = 2;
constexpr int kFtOpBytes = 128 * 128;         // one 128-row x 64-col bf16 box (16 live columns)
constexpr int kSmemFt = 2 * kFtOpBytes + kFtOpBytes + kFtBufs * kStageBytes + 1024 + 512;

struct FtBarriers {
    uint64_t op_full[2], op_empty[2];         // dz2b tile ring
    uint64_t w_full, w_empty;                 // W2T tile of the current column block
    uint64_t g_full[2], g_empty[2];           // TMEM g accumulators
    uint64_t ld_full[kFtBufs], ld_empty[kFtBufs];
    uint64_t a1_done[kFtBufs / 2];            // MMA2 finished reading a chunk pair
    uint64_t d2_full, d2_empty;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
bwd_tail_kernel(const __grid_constant__ CUtensorMap tmD /* dz2b [rows][16 of 64] */,
                const __grid_constant__ CUtensorMap tmW /* W2T  [H][16 of 64]    */,
                const __grid_constant__ CUtensorMap tmA1 /* A1  [rows][H] load   */,
                const __grid_constant__ CUtensorMap tmZ /* dZ1  [rows][H] store  */,
                float* __restrict__ gW2, int ldg2, int H, int C, int m_tiles, int n_tiles,
                float* __restrict__ part /* null: add into gW2; else slot (CTA + column block) x 16 x 128 */) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* sD = smem;                               // 2 x 16 KB: dz2b tiles
    uint8_t* sW = sD + 2 * kFtOpBytes;                // 16 KB: W2T tile
    uint8_t* sC = sW + kFtOpBytes;                    // staging ring
    FtBarriers* bars = (FtBarriers*)(sC + kFtBufs * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmD);
        prefetch_tmap(&tmW);
        prefetch_tmap(&tmA1);
        prefetch_tmap(&tmZ);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&bars->op_full[s], 1);
            mbar_init(&bars->op_empty[s], 1);
            mbar_init(&bars->g_full[s], 1);
            mbar_init(&bars->g_empty[s], 4);
        }
        mbar_init(&bars->w_full, 1);
        mbar_init(&bars->w_empty, 1);
        mbar_init(&bars->d2_full, 1);
        mbar_init(&bars->d2_empty, 4);
        for (int b = 0; b < kFtBufs; ++b) {
            mbar_init(&bars->ld_full[b], 1);
            mbar_init(&bars->ld_empty[b], 1);
        }
        for (int b = 0; b < kFtBufs / 2; ++b) mbar_init(&bars->a1_done[b], 1);
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(&bars->tmem_base, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    constexpr uint32_t kD2Col = 256;                  // g stages at columns 0 and 128, D2 at 256

    // contiguous range of tiles, column block (n) slowest: tile = n * m_tiles + m
    const long long total = (long long)m_tiles * n_tiles;
    const long long t_begin = total * blockIdx.x / gridDim.x;
    const long long t_end = total * (blockIdx.x + 1) / gridDim.x;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0, pair = 0, last_n = -1;
            uint32_t phase = 0, pphase = 0, wphase = 0;
            for (long long t = t_begin; t < t_end; ++t) {
                const int n = (int)(t / m_tiles), m = (int)(t % m_tiles);
                const int m0 = m * 128, n0 = n * 128;
                if (n != last_n) {
                    mbar_wait(&bars->w_empty, wphase ^ 1);
                    mbar_expect_tx(&bars->w_full, kFtOpBytes);
                    tma_load_2d(sW, &tmW, &bars->w_full, 0, n0);
                    wphase ^= 1;
                    last_n = n;
                }
                mbar_wait(&bars->op_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->op_full[stage], kFtOpBytes);
                tma_load_2d(sD + stage * kFtOpBytes, &tmD, &bars->op_full[stage], 0, m0);
                if (++stage == 2) { stage = 0; phase ^= 1; }
                for (int c = 0; c < 2; ++c) {
                    const int buf = pair * 2 + c;
                    mbar_wait(&bars->ld_empty[buf], pphase ^ 1);
                    mbar_expect_tx(&bars->ld_full[buf], kStageBytes);
                    tma_load_2d(sC + buf * kStageBytes, &tmA1, &bars->ld_full[buf], n0 + c * 64, m0);
                }
                if (++pair == kFtBufs / 2) { pair = 0; pphase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc_g = idesc_bf16(128, 128, 0, 0);     // dz2b (K-major) x W2T (K-major)
            const uint32_t idesc_d2 = idesc_bf16(128, 16, 1, 1);     // A1^T (MN-major) x dz2b (MN-major)
            int stage = 0, acc = 0, pair = 0, last_n = -1;
            uint32_t phase = 0, acc_phase = 0, pphase = 0, wphase = 0, d2phase = 0;
            bool d2_open = false;                                    // D2 holds partial sums of last_n
            for (long long t = t_begin; t < t_end; ++t) {
                const int n = (int)(t / m_tiles);
                if (n != last_n) {
                    if (d2_open) {                                   // finish the previous column block
                        umma_commit(&bars->d2_full);
                        d2_open = false;
                    }
                    mbar_wait(&bars->w_full, wphase);
                    wphase ^= 1;
                    mbar_wait(&bars->d2_empty, d2phase ^ 1);         // previous flush has drained D2
                    d2phase ^= 1;
                    last_n = n;
                }
                mbar_wait(&bars->op_full[stage], phase);
                mbar_wait(&bars->g_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_addr = smem_u32(sD + stage * kFtOpBytes);
                umma_bf16(tmem_base + (uint32_t)(acc * 128), smem_desc_sw128(d_addr, 16, 1024),
                          smem_desc_sw128(smem_u32(sW), 16, 1024), idesc_g, 0);
                umma_commit(&bars->g_full[acc]);
                // the W2T tile may be replaced once the last MMA1 of this column block is done
                const bool last_of_n = (t + 1 == t_end) || ((int)((t + 1) / m_tiles) != n);
                if (last_of_n) umma_commit(&bars->w_empty);
                // MMA2 over the 128 batch rows of this tile
                mbar_wait(&bars->ld_full[pair * 2], pphase);
                mbar_wait(&bars->ld_full[pair * 2 + 1], pphase);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(sC + pair * 2 * kStageBytes);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint64_t da = smem_desc_sw128(a_addr + j * 2048, kStageBytes, 1024);
                    const uint64_t db = smem_desc_sw128(d_addr + j * 2048, 8192, 1024);
                    umma_bf16(tmem_base + kD2Col, da, db, idesc_d2, (d2_open || j > 0) ? 1u : 0u);
                }
                d2_open = true;
                umma_commit(&bars->a1_done[pair]);
                umma_commit(&bars->op_empty[stage]);
                if (++stage == 2) { stage = 0; phase ^= 1; }
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
                if (++pair == kFtBufs / 2) { pair = 0; pphase ^= 1; }
            }
            if (d2_open) umma_commit(&bars->d2_full);
        }
        __syncwarp();
    } else {
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const bool leader = (warp == 2 && lane == 0);
        int acc = 0, pair = 0;
        uint32_t acc_phase = 0, pphase = 0, d2phase = 0;
        long long chunk_no = 0;
        for (long long t = t_begin; t < t_end; ++t) {
            const int n = (int)(t / m_tiles), m = (int)(t % m_tiles);
            const int m0 = m * 128, n0 = n * 128;
            mbar_wait(&bars->g_full[acc], acc_phase);
            mbar_wait(&bars->a1_done[pair], pphase);      // MMA2 no longer reads the pair
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 128);
            for (int c = 0; c < 2; ++c) {
                const int buf = pair * 2 + c;
                mbar_wait(&bars->ld_full[buf], pphase);
                uint8_t* stg = sC + buf * kStageBytes;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t v[32];
                    tmem_ld_32x32(t_row + (uint32_t)(c * 64 + half * 32), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int chunk = half * 4 + g;
                        const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
                        uint4 a4 = *reinterpret_cast<const uint4*>(stg + off);
                        uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
                        uint32_t p[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float a0 = __uint_as_float(aw[e] << 16);
                            const float a1 = __uint_as_float(aw[e] & 0xffff0000u);
                            const float g0 = __uint_as_float(v[g * 8 + 2 * e]);
                            const float g1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
                            p[e] = pack_bf16(g0 * a0 * (1.f - a0), g1 * a1 * (1.f - a1));
                        }
                        *reinterpret_cast<uint4*>(stg + off) = make_uint4(p[0], p[1], p[2], p[3]);
                    }
                }
                fence_proxy_async_smem();
                named_bar_sync(1, 128);
                if (leader) {
                    tma_store_2d(&tmZ, stg, n0 + c * 64, m0);
                    tma_store_commit();
                    tma_store_wait_read<kFtPending>();
                    if (chunk_no >= kFtPending)
                        mbar_arrive(&bars->ld_empty[(int)((chunk_no - kFtPending) % kFtBufs)]);
                    ++chunk_no;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->g_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
            if (++pair == kFtBufs / 2) { pair = 0; pphase ^= 1; }
            // last tile of this column block: flush D2 = gW2^T block, transposed
            const bool last_of_n = (t + 1 == t_end) || ((int)((t + 1) / m_tiles) != n);
            if (last_of_n) {
                mbar_wait(&bars->d2_full, d2phase);
                d2phase ^= 1;
                tc_fence_after();
                uint32_t v[16];
                tmem_ld_32x16(tmem_base + ((uint32_t)(q * 32) << 16) + kD2Col, v);
                tmem_ld_wait();
                const int h = n0 + row;
                if (part) {
                    // deterministic mode: this CTA's share of the column block goes to a slot of its
                    // own (CTA index + column block is unique, see dw2_reduce_kernel), stored
                    float* dst = part + ((size_t)blockIdx.x + (size_t)n) * 2048 + row;
#pragma unroll
                    for (int c = 0; c < 16; ++c) dst[c * 128] = __uint_as_float(v[c]);
                } else if (h < H) {
#pragma unroll
                    for (int c = 0; c < 16; ++c)
                        if (c < C) atomicAdd(gW2 + (size_t)c * ldg2 + h, __uint_as_float(v[c]));
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->d2_empty);
            }
        }
        if (leader) tma_store_wait_all<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// gW2[c][h] = sum over the CTAs of bwd_tail_kernel that worked on h's column block, in CTA order.
// CTA b covered tiles [T*b/G, T*(b+1)/G) of the column-major tile order (T = m_tiles * n_tiles),
// so the CTAs of column block n are consecutive and (b, n) -> slot b + n is one-to-one.
__global__ void __launch_bounds__(128)
dw2_reduce_kernel(const float* __restrict__ part, float* __restrict__ gW2, int ldg2, int H, int C,
                  int m_tiles, int n_tiles, int G) {
    const int n = blockIdx.x, c = blockIdx.y, r = threadIdx.x;
    const int h = n * 128 + r;
    const long long T = (long long)m_tiles * n_tiles, lo = (long long)n * m_tiles, hi = lo + m_tiles;
    int b0 = (int)(lo * G / T);
    while (b0 < G && T * (b0 + 1) / G <= lo) ++b0;
    int b1 = b0;
    while (b1 < G && T * b1 / G < hi) ++b1;                      // every CTA has tiles (G <= T)
    const float* src = part + (size_t)n * 2048 + (size_t)c * 128 + r;
    float acc = 0.f;
    int b = b0;
    for (; b + 8 <= b1; b += 8) {                                // loads first, then the adds in CTA order
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = src[(size_t)(b + k) * 2048];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc += v[k];
    }
    for (; b < b1; ++b) acc += src[(size_t)b * 2048];
    if (h < H && c < C) gW2[(size_t)c * ldg2 + h] = acc;
}

// =============================================================================
// host side: tensor maps and launches
// =============================================================================
using tc_host::make_map;
using tc_host::opt_in_smem;

// Number of k-chunks for the (k-chunk, m, n) schedule: the smallest split (at most 16, at least
// 8 k-blocks per chunk) whose item count fills `sms` CTAs with the least idle tail.
int choose_ksplit(int tiles_mn, int kb_total, int sms) {
    // Small problems (every item of every chunk runs at once): the chunks of a tile then finish
    // together and their epilogues go one after the other (ordered split-K), so the time is about
    // kb/s k-blocks (~0.27 us each) plus s epilogues (~2.5 us each): pick the s that minimises it.
    if ((long long)tiles_mn * 16 <= sms || (long long)tiles_mn * std::max(1, kb_total / 8) <= sms) {
        int bs = 1;
        double bt = 1e30;
        for (int s = 1; s <= 16 && (long long)tiles_mn * s <= sms; ++s) {
            if (s > 1 && kb_total / s < 8) break;
            const double t = 0.27 * ((kb_total + s - 1) / s) + 2.5 * s;
            if (t < bt) { bt = t; bs = s; }
        }
        return bs;
    }
    int best = 1;
    double best_eff = 0.0;
    for (int s = 1; s <= 16; ++s) {
        if (s > 1 && kb_total / s < 8) break;
        const long long items = (long long)tiles_mn * s;
        const long long waves = (items + sms - 1) / sms;
        const double eff = (double)items / (double)(waves * sms);
        if (eff > best_eff + 0.02) { best_eff = eff; best = s; }
    }
    return best;
}

// k-split of the CTA-pair dW1 kernel.  Measured on B200 (8192-wide model, dW1 alone / whole step):
//   8192 rows   s = 1: 92 / 311 us    s = 4 (what filling every wave suggests): 98 / 341 us
//   16384 rows  s = 1: 172 / 534      s = 4: 181 / 552
//   65536 rows  s = 1: 656            s = 4: 651
//   1024-wide model, 8192 rows (16 tiles for 74 pairs)   s = 1: 51 / 114    s = 4: 69 / 132
// Two things the wave count does not see: the board is power-bound, so a ragged last wave costs
// little (idle SMs lend their power to the busy ones), and the epilogue of a 256 x 256 fp32 tile
// takes ~10 us, which a chunk shorter than ~40 k-blocks no longer hides, while the chunks of one
// tile that run at the same time queue up behind each other (ordered split-K).  So: no split once
// the tiles cover half the machine; below that the split that minimises
// (kb / s) * 0.33 us + s * 13 us among those that keep every item resident.
int choose_ksplit_pairs(int tiles, int kb_total, int npairs) {
    if (2 * tiles >= npairs) return 1;
    int bs = 1;
    double bt = 1e30;
    for (int s = 1; s <= 16 && (long long)tiles * s <= npairs; ++s) {
        if (s > 1 && kb_total / s < 16) break;
        const double t = 0.33 * ((kb_total + s - 1) / s) + 13.0 * s;
        if (t < bt) { bt = t; bs = s; }
    }
    return bs;
}

// Small dW1 products (a handful of tiles): one SM pulls its operands out of L2 at ~100 GB/s, so a
// k-block costs about c = 0.10 + 0.0018 * BN microseconds and every chained epilogue of the ordered
// split-K about e = 0.9 + 0.009 * BN (fitted on the 784-100-10 model, 3200 rows: 34 us with one
// 208-wide chunk per tile, 15 us with 64-wide tiles and 2-3 chunks).  Narrow tiles put more SMs
// to work on the same bytes; pick the column-tile width and the k-split that minimise
// c * ceil(kb / s) + e * s with every item resident at once.
void choose_small_tiling(int m_tiles, int K, int kb_total, int sms, int* bn_cap, int* ksplit) {
    double best = 1e30;
    *bn_cap = 256, *ksplit = 1;
    for (int cap : {64, 128, 256}) {
        const int nt = (K + cap - 1) / cap;
        const int bn = ((K + nt - 1) / nt + 15) / 16 * 16;
        const double c = 0.10 + 0.0018 * bn, e = 0.9 + 0.009 * bn;
        for (int s = 1; s <= 16 && (long long)m_tiles * nt * s <= sms; ++s) {
            if (s > 1 && kb_total / s < 4) break;
            const double t = c * ((kb_total + s - 1) / s) + e * s;
            if (t < best - 1e-9) { best = t; *bn_cap = cap; *ksplit = s; }
        }
    }
}

uint32_t pow2_cols(int need) {
    uint32_t c = 32;
    while ((int)c < need) c <<= 1;
    return c;
}

int launch_kmajor(int epi, const void* A, const void* B, void* Cout, int64_t M, int N, int K,
                  int lda, int ldb, int ldc, cudaStream_t st) {
    CME_CHECK_ARG(A && B && Cout && M > 0 && N > 0 && K > 0, "gemm_kmajor: bad argument");
    CME_CHECK_ARG(N % 64 == 0, "gemm_kmajor: N (%d) must be a multiple of 64", N);
    CME_CHECK_ARG(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 && lda >= K && ldb >= K && ldc >= N,
                  "gemm_kmajor: leading dimensions must be multiples of 8 and cover the rows");
    CME_CHECK_ARG(M < (1ll << 31), "gemm_kmajor: M too large");
    const int BN = N >= 256 ? 256 : N;
    const int m_tiles = (int)((M + BM - 1) / BM), n_tiles = (N + BN - 1) / BN;
    CUtensorMap tmA, tmB, tmC;
    int rc;
    if ((rc = make_map(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, BK, BM))) return rc;
    if ((rc = make_map(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, (uint32_t)BN))) return rc;
    if ((rc = make_map(&tmC, Cout, (uint64_t)N, (uint64_t)M, (uint64_t)ldc, 64, BM))) return rc;
    // CTA-pair path: full 256-wide column tiles and enough row tiles to keep every pair busy
    static const bool pairs_off = getenv("CME_NO_PAIRS") != nullptr;
    if (!pairs_off && N % 256 == 0 && M >= 256 && epi <= 1) {
        const int m2 = (int)((M + 255) / 256), n2 = N / 256;
        CUtensorMap tmB2;
        if ((rc = make_map(&tmB2, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, BK, 128))) return rc;
        const long long t2 = (long long)m2 * n2;
        int pairs = (int)std::min<long long>(t2, cme::sm_count() / 2);
        if (epi == 1) {
            if ((rc = opt_in_smem(gemm_kmajor2_kernel<1>, kSmemKMajor2))) return rc;
            gemm_kmajor2_kernel<1><<<2 * pairs, kThreadsK2, kSmemKMajor2, st>>>(tmA, tmB2, tmC, m2, n2, K);
        } else {
            if ((rc = opt_in_smem(gemm_kmajor2_kernel<0>, kSmemKMajor2))) return rc;
            gemm_kmajor2_kernel<0><<<2 * pairs, kThreadsK2, kSmemKMajor2, st>>>(tmA, tmB2, tmC, m2, n2, K);
        }
        CME_LAUNCH_CHECK();
        return 0;
    }
    const uint32_t cols = pow2_cols(2 * BN);
    const long long tiles = (long long)m_tiles * n_tiles;
    const int grid = (int)std::min<long long>(tiles, cme::sm_count());
    SoftmaxEpi none{};
    if (epi == 1) {
        if ((rc = opt_in_smem(gemm_kmajor_kernel<1>, kSmemKMajor))) return rc;
        gemm_kmajor_kernel<1><<<grid, kThreads, kSmemKMajor, st>>>(tmA, tmB, tmC, m_tiles, n_tiles, K,
                                                                  BN, cols, none, 1);
    } else {
        if ((rc = opt_in_smem(gemm_kmajor_kernel<0>, kSmemKMajor))) return rc;
        gemm_kmajor_kernel<0><<<grid, kThreads, kSmemKMajor, st>>>(tmA, tmB, tmC, m_tiles, n_tiles, K,
                                                                  BN, cols, none, 1);
    }
    CME_LAUNCH_CHECK();
    return 0;
}

}  // namespace

extern "C" {

int cme_mlp_fwd1(const void* X, const void* W1c, const float* b1, void* A1, int64_t rows, int K,
                 int H, int ldx, int ldh, void* stream) {
    // b1 is folded into W1c (column K-1 of the augmented operands); the pointer
    // is accepted for API symmetry and must be NULL
    CME_CHECK_ARG(b1 == nullptr, "cme_mlp_fwd1: the bias lives in W1c's ones column; pass NULL");
    const int Hp = (H + 63) / 64 * 64;
    CME_CHECK_ARG(ldh >= Hp, "cme_mlp_fwd1: ldh (%d) must cover H rounded up to 64 (%d)", ldh, Hp);
    return launch_kmajor(1, X, W1c, A1, rows, Hp, K, ldx, ldx, ldh, cme::as_stream(stream));
}

int cme_gemm_bf16(const void* A, const void* B, void* Cout, int64_t M, int N, int K, int lda,
                  int ldb, int ldc, void* stream) {
    return launch_kmajor(0, A, B, Cout, M, N, K, lda, ldb, ldc, cme::as_stream(stream));
}

// Scratch of the ordered split-K (see DwSync): kDwPartWords part counters followed by
// kDwUnitWords unit counters, all zero between launches.
constexpr int kDwPartWords = 8192;
constexpr int kDwUnitWords = 1024;

static int dw1_check_args(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                          int ldh, int ldg) {
    CME_CHECK_ARG(dZ1 && X && gW1 && rows > 0 && K > 0 && H > 0, "cme_mlp_dw1: bad argument");
    CME_CHECK_ARG(ldx % 8 == 0 && ldh % 8 == 0 && ldg % 4 == 0 && ldx >= K && ldh >= H && ldg >= K,
                  "cme_mlp_dw1: leading dimensions (ldx=%d ldh=%d ldg=%d)", ldx, ldh, ldg);
    CME_CHECK_ARG(((uintptr_t)gW1 & 15) == 0, "cme_mlp_dw1: gW1 must be 16-byte aligned");
    CME_CHECK_ARG(rows < (1ll << 31), "cme_mlp_dw1: too many rows");
    return 0;
}

static int dw1_launch(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                      int ldh, int ldg, uint32_t* scratch, uint32_t* const* pads, int nranks,
                      int unit_rows, uint32_t* status, const upd::UpdateArgs* fused, cudaStream_t st,
                      bool* fused_used = nullptr) {
    CME_CHECK_ARG(unit_rows == 128 || unit_rows == 256, "cme_mlp_dw1: unit_rows is 128 or 256 (got %d)", unit_rows);
    int rc0 = dw1_check_args(dZ1, X, gW1, rows, K, H, ldx, ldh, ldg);
    if (rc0) return rc0;
    CME_CHECK_ARG(scratch, "cme_mlp_dw1: no scratch");
    upd::UpdateArgs ua{};
    if (fused) ua = *fused;
    // The update warps are 128 threads per CTA of THIS grid.  A small GEMM (few CTAs) would leave
    // each of them many vectors per unit (more than four here), one NVLink round trip after the other (measured:
    // 60 us for the 784-100-10 model on 8 CTAs): such launches leave the update to the kernel of
    // its own, which sizes its grid for the slice (fused_used tells the caller).
    auto enough_workers = [&](long long ctas) {
        const long long slice_quads = ua.unit_elems / 4 / std::max(1, ua.nranks);
        return slice_quads <= 4 * ctas * 32 * kUpdWarps;
    };
    // output tile: 128 rows of H  x  BN columns of K, BN a multiple of 16, <= 256
    static const int dev_bn = getenv("CME_DW1_BN") ? atoi(getenv("CME_DW1_BN")) : 0;      // tuning hooks
    static const int dev_ks = getenv("CME_DW1_KS") ? atoi(getenv("CME_DW1_KS")) : 0;
    const int m_tiles = (H + BM - 1) / BM;
    const int kb_total = (int)((rows + BK - 1) / BK);
    int bn_cap = 256, ks_small = 0;
    if ((long long)m_tiles * ((K + 255) / 256) * 16 <= cme::sm_count())
        choose_small_tiling(m_tiles, K, kb_total, cme::sm_count(), &bn_cap, &ks_small);
    if (dev_bn >= 16 && dev_bn <= 256) bn_cap = dev_bn, ks_small = 0;
    const int n_tiles = (K + bn_cap - 1) / bn_cap;
    const int BN = ((K + n_tiles - 1) / n_tiles + 15) / 16 * 16;
    CUtensorMap tmA, tmB;
    int rc;
    if ((rc = make_map(&tmA, dZ1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, BK))) return rc;
    if ((rc = make_map(&tmB, X, (uint64_t)K, (uint64_t)rows, (uint64_t)ldx, 64, BK))) return rc;
    DwSync sy{};
    sy.part_ctr = scratch;
    sy.unit_ctr = scratch + kDwPartWords;
    sy.pads = pads;
    sy.status = status;
    sy.nranks = nranks;
    // Row blocks are walked in `sblocks` groups; a group's units are complete (and their update
    // starts) before the next group begins, at the price of one more pass over B (the minibatch)
    // per group.  Inside a group the k-chunks of a tile are a whole round of items apart, so
    // nobody waits for its turn, PROVIDED the group holds at least as many tiles as there are
    // CTAs; a smaller group puts the chunks of one tile on neighbouring CTAs at the same time and
    // their epilogues queue up behind each other.  That is harmless when a chunk is long (>= 128
    // k-blocks: 4 groups then, measured 2.02 vs 2.06 ms per step at 65536 rows) and ruinous when
    // it is short (8192 rows per GPU: 32 k-blocks per chunk, dW1 0.33 instead of 0.12 ms with one
    // group per row block), so short chunks use a single group: all units then complete during
    // the last chunk phase, the last quarter of the kernel, in row-block order.
    // CTA-pair path for big problems: 256 x 256 tiles (columns past K are masked in the epilogue)
    static const bool pairs_off = getenv("CME_NO_PAIRS") != nullptr;
    if (!pairs_off && H >= 512 && K >= 256 && kb_total >= 16 && unit_rows == 256) {
        // 256-wide column tiles.  Narrower equal-width tiles (785 columns -> 4 x 208, 5.6 % padding
        // instead of 23 %) were measured SLOWER (0.71 vs 0.62 ms): with cta_group::2 the MMA time
        // does not shrink with N = 208 and each CTA's 104 columns leave a partial swizzle group.
        // A 32-wide last tile (3 x 256 + 32, per-item instruction descriptor) was also built and
        // measured: correct, and exactly as fast as four full tiles (2.0977 vs 2.0974 ms per
        // step).  MMAs on zero padding cost pipe time but no power, and power is what bounds
        // this kernel, so the simpler uniform tiling stays.
        const int m2 = (H + 255) / 256, n2 = (K + 255) / 256;
        const int BN2 = 256;
        const int npairs = cme::sm_count() / 2;
        const int ks = dev_ks > 0 ? std::min(dev_ks, std::max(1, kb_total / 8)) : choose_ksplit_pairs(m2 * n2, kb_total, npairs);
        CME_CHECK_ARG(m2 * n2 * 2 <= kDwPartWords && m2 <= kDwUnitWords && m2 <= cme::kPadMaxUnits,
                      "cme_mlp_dw1: %d x %d tiles exceed the counter scratch", m2, n2);
        sy.parts_per_tile = n2 * 2;
        sy.tiles_per_unit = 1;
        sy.m_tiles = m2;
        sy.sblocks = (ks > 1 && kb_total / ks >= 128) ? std::min(m2, 4) : 1;
        if ((rc = opt_in_smem(gemm_mnmajor2_kernel, kSmemMNMajor2))) return rc;
        const long long items2 = (long long)m2 * n2 * ks;
        const int pairs = (int)std::min<long long>(items2, npairs);
        if (ua.enabled && !enough_workers(2 * pairs)) ua.enabled = 0;
        if (fused_used) *fused_used = ua.enabled != 0;
        const int threads = ua.enabled ? kThreadsDw : kThreads;
        gemm_mnmajor2_kernel<<<2 * pairs, threads, kSmemMNMajor2, st>>>(
            tmA, tmB, gW1, ldg, H, K, m2, n2, kb_total, ks, BN2, sy, ua);
        CME_LAUNCH_CHECK();
        return 0;
    }
    CME_CHECK_ARG(m_tiles * n_tiles <= kDwPartWords && m_tiles <= kDwUnitWords && m_tiles <= cme::kPadMaxUnits,
                  "cme_mlp_dw1: %d x %d tiles exceed the counter scratch", m_tiles, n_tiles);
    sy.parts_per_tile = n_tiles;
    sy.tiles_per_unit = unit_rows / BM;
    sy.m_tiles = m_tiles;
    sy.sblocks = 1;                      // filled in below, once ksplit is known
    if ((rc = opt_in_smem(gemm_mnmajor_kernel, kSmemMNMajor))) return rc;
    const int ksplit = dev_ks > 0 ? std::min(dev_ks, std::max(1, kb_total))
                       : ks_small > 0 ? ks_small
                                      : choose_ksplit(m_tiles * n_tiles, kb_total, cme::sm_count());
    sy.sblocks = (kb_total / ksplit >= 128) ? std::min(m_tiles, 4) : 1;
    const long long items = (long long)m_tiles * n_tiles * ksplit;
    const int grid = (int)std::min<long long>(items, cme::sm_count());
    if (ua.enabled && !enough_workers(grid)) ua.enabled = 0;
    if (fused_used) *fused_used = ua.enabled != 0;
    const int threads = ua.enabled ? kThreadsDw : kThreads;
    gemm_mnmajor_kernel<<<grid, threads, kSmemMNMajor, st>>>(
        tmA, tmB, gW1, ldg, H, K, m_tiles, n_tiles, kb_total, BN, 512u, 0, ksplit, sy, ua);
    CME_LAUNCH_CHECK();
    return 0;
}

// counter scratch for callers that bring none: one per device, allocated and zeroed on first use
static uint32_t* dw1_default_scratch() {
    static std::mutex mu;
    static std::map<int, uint32_t*> per_dev;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lk(mu);
    auto it = per_dev.find(dev);
    if (it != per_dev.end()) return it->second;
    uint32_t* p = nullptr;
    const size_t bytes = sizeof(uint32_t) * (kDwPartWords + kDwUnitWords);
    if (cudaMalloc(&p, bytes) != cudaSuccess || cudaMemset(p, 0, bytes) != cudaSuccess) {
        cme::set_error("cme_mlp_dw1: cannot allocate the split-K counters");
        return nullptr;
    }
    per_dev[dev] = p;
    return p;
}

/* How cme_mlp_dw1 / cme_mlp_dw1_update would tile a dW1 product of `rows` samples (no launch, no
 * GPU needed): out = {pair path (0/1), row tiles, column tiles, column-tile width, k-split, CTAs}.
 * The same arithmetic as dw1_launch; exists so the tiling rules can be tested and inspected. */
int cme_mlp_dw1_plan_query(int64_t rows, int K, int H, int out[6]) {
    CME_CHECK_ARG(out && rows > 0 && K > 0 && H > 0, "cme_mlp_dw1_plan_query: bad argument");
    const int unit_rows = ((H + 63) / 64 * 64) >= 512 ? 256 : 128;
    const int m_tiles = (H + BM - 1) / BM;
    const int kb_total = (int)((rows + BK - 1) / BK);
    if (H >= 512 && K >= 256 && kb_total >= 16 && unit_rows == 256) {
        const int m2 = (H + 255) / 256, n2 = (K + 255) / 256, npairs = cme::sm_count() / 2;
        const int ks = choose_ksplit_pairs(m2 * n2, kb_total, npairs);
        const long long items = (long long)m2 * n2 * ks;
        out[0] = 1, out[1] = m2, out[2] = n2, out[3] = 256, out[4] = ks;
        out[5] = 2 * (int)std::min<long long>(items, npairs);
        return 0;
    }
    int bn_cap = 256, ks_small = 0;
    if ((long long)m_tiles * ((K + 255) / 256) * 16 <= cme::sm_count())
        choose_small_tiling(m_tiles, K, kb_total, cme::sm_count(), &bn_cap, &ks_small);
    const int n_tiles = (K + bn_cap - 1) / bn_cap;
    const int BN = ((K + n_tiles - 1) / n_tiles + 15) / 16 * 16;
    const int ks = ks_small > 0 ? ks_small : choose_ksplit(m_tiles * n_tiles, kb_total, cme::sm_count());
    out[0] = 0, out[1] = m_tiles, out[2] = n_tiles, out[3] = BN, out[4] = ks;
    out[5] = (int)std::min<long long>((long long)m_tiles * n_tiles * ks, cme::sm_count());
    return 0;
}

size_t cme_mlp_dw1_scratch_bytes(void) { return sizeof(uint32_t) * (kDwPartWords + kDwUnitWords); }

int cme_mlp_dw1(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                int ldh, int ldg, void* stream) {
    int rc0 = dw1_check_args(dZ1, X, gW1, rows, K, H, ldx, ldh, ldg);
    if (rc0) return rc0;
    uint32_t* scratch = dw1_default_scratch();
    if (!scratch) return -2;
    return dw1_launch(dZ1, X, gW1, rows, K, H, ldx, ldh, ldg, scratch, nullptr, 0, H >= 512 ? 256 : 128,
                      nullptr, nullptr, cme::as_stream(stream));
}

int cme_mlp_dw1_sync(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,
                     int ldh, int ldg, void* scratch, uint32_t* const* pads, int nranks, int unit_rows,
                     uint32_t* status, void* stream) {
    CME_CHECK_ARG(scratch && (!pads || (nranks >= 1 && nranks <= 16)), "cme_mlp_dw1_sync: bad sync arguments");
    return dw1_launch(dZ1, X, gW1, rows, K, H, ldx, ldh, ldg, (uint32_t*)scratch, pads, nranks, unit_rows,
                      status, nullptr, cme::as_stream(stream));
}

int cme_mlp_dw1_update(const CmeMlpPlan* p, int64_t rows, float lr, float reg, int indirect, void* stream) {
    CME_CHECK_ARG(p && p->sync_scratch && p->pad_local, "cme_mlp_dw1_update: the plan has no sync buffers");
    upd::UpdateArgs ua;
    int rc = cme::make_update_args(p, lr, reg, indirect, &ua);
    if (rc) return rc;
    bool fused_used = false;
    rc = dw1_launch(p->dz1, p->xb, p->grads[p->rank], rows, p->D + 1, p->H, p->ldx, p->ldh, p->ldx,
                    (uint32_t*)p->sync_scratch, p->pads_dev, p->nranks, p->unit_rows,
                    p->pad_local + cme::kPadStatus, &ua, cme::as_stream(stream), &fused_used);
    if (rc || fused_used) return rc;
    return cme_mlp_update_stream_ind(p, lr, reg, indirect, stream);      // small GEMM: update on its own grid
}

int cme_mlp_fwd2_tc(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                    float* dz2, void* dz2b, float* loss_sum, int32_t* pred, float* gb2, float scale,
                    int64_t rows, int H, int ldh, int C, void* stream) {
    return cme_mlp_fwd2_tc_ind(A1, W2c, b2, labels, nullptr, dz2, dz2b, loss_sum, pred, gb2, scale, rows, H,
                               ldh, C, stream);
}

int cme_mlp_fwd2_tc_ind(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                        const CmeStepArgs* step_args, float* dz2, void* dz2b, float* loss_sum, int32_t* pred,
                        float* gb2, float scale, int64_t rows, int H, int ldh, int C, void* stream) {
    CME_CHECK_ARG(A1 && W2c && b2 && rows > 0 && H > 0, "cme_mlp_fwd2_tc: bad argument");
    CME_CHECK_ARG(C >= 1 && C <= 16, "cme_mlp_fwd2_tc: 1..16 classes (got %d)", C);
    CME_CHECK_ARG(ldh % 8 == 0 && ldh >= H, "cme_mlp_fwd2_tc: ldh (%d)", ldh);
    CME_CHECK_ARG((!dz2 && !dz2b) || labels || step_args, "cme_mlp_fwd2_tc: gradients need labels");
    CME_CHECK_ARG(rows < (1ll << 31), "cme_mlp_fwd2_tc: too many rows");
    const int m_tiles = (int)((rows + BM - 1) / BM);
    CUtensorMap tmA, tmB;
    int rc;
    // the K extent is the true H: pad columns of A1 hold sigmoid(0) and must read as zero
    if ((rc = make_map(&tmA, A1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, BK, BM))) return rc;
    if ((rc = make_map(&tmB, W2c, (uint64_t)H, 16, (uint64_t)ldh, BK, 16))) return rc;
    if ((rc = opt_in_smem(gemm_kmajor_kernel<2>, kSmemSoftmax))) return rc;
    SoftmaxEpi sm{b2, labels, dz2, (__nv_bfloat16*)dz2b, loss_sum, pred, gb2, scale, (long long)rows, C, step_args};
    const int nk = (H + BK - 1) / BK;
    if (2 * m_tiles <= cme::sm_count() && nk >= 8) {
        // few row tiles: two CTAs per tile, each streaming half of K (cluster of 2)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * m_tiles);
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = kSmemSoftmax;
        cfg.stream = cme::as_stream(stream);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        CME_CUDA(cudaLaunchKernelEx(&cfg, gemm_kmajor_kernel<2>, tmA, tmB, tmA, m_tiles, 1, H, 16, 32u, sm, 2));
        return 0;
    }
    const int grid = std::min(m_tiles, cme::sm_count());
    gemm_kmajor_kernel<2><<<grid, kThreads, kSmemSoftmax, cme::as_stream(stream)>>>(
        tmA, tmB, tmA, m_tiles, 1, H, 16, 32u, sm, 1);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_bwd_dz1_tc(const void* A1, const void* dz2b, const void* W2T, void* dZ1, int64_t rows,
                       int H, int ldh, void* stream) {
    CME_CHECK_ARG(A1 && dz2b && W2T && dZ1 && rows > 0 && H > 0, "cme_mlp_bwd_dz1_tc: bad argument");
    CME_CHECK_ARG(ldh % 8 == 0 && ldh >= H, "cme_mlp_bwd_dz1_tc: ldh (%d)", ldh);
    CME_CHECK_ARG(rows < (1ll << 31), "cme_mlp_bwd_dz1_tc: too many rows");
    const int m_tiles = (int)((rows + BM - 1) / BM), n_tiles = (H + 255) / 256;
    CUtensorMap tmD, tmW, tmA1, tmZ;
    int rc;
    // both small operands are rows of 64 bf16 with 16 live values: bound the maps at 16
    if ((rc = make_map(&tmD, dz2b, 16, (uint64_t)rows, 64, 64, BM))) return rc;
    if ((rc = make_map(&tmW, W2T, 16, (uint64_t)H, 64, 64, 256))) return rc;
    if ((rc = make_map(&tmA1, A1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, BM))) return rc;
    if ((rc = make_map(&tmZ, dZ1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, BM))) return rc;
    if ((rc = opt_in_smem(bwd_dz1_kernel, kSmemDz))) return rc;
    const long long tiles = (long long)m_tiles * n_tiles;
    const int grid = (int)std::min<long long>(tiles, cme::sm_count());
    bwd_dz1_kernel<<<grid, kThreads, kSmemDz, cme::as_stream(stream)>>>(tmD, tmW, tmA1, tmZ, m_tiles,
                                                                      n_tiles);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_bwd_tail_tc(const void* A1, const void* dz2b, const void* W2T, void* dZ1, float* gW2,
                        int ldg2, int64_t rows, int H, int ldh, int C, void* stream) {
    return cme_mlp_bwd_tail_det(A1, dz2b, W2T, dZ1, gW2, ldg2, rows, H, ldh, C, nullptr, stream);
}

size_t cme_mlp_bwd_tail_scratch_floats(int H) {
    // one 16 x 128 slot per (CTA, column block) pair: at most SMs + column blocks of them
    return ((size_t)cme::sm_count() + (size_t)(H + 127) / 128 + 1) * 2048;
}

int cme_mlp_bwd_tail_det(const void* A1, const void* dz2b, const void* W2T, void* dZ1, float* gW2,
                         int ldg2, int64_t rows, int H, int ldh, int C, float* part, void* stream) {
    CME_CHECK_ARG(A1 && dz2b && W2T && dZ1 && gW2 && rows > 0 && H > 0, "cme_mlp_bwd_tail_tc: bad argument");
    CME_CHECK_ARG(ldh % 8 == 0 && ldh >= H && C >= 1 && C <= 16, "cme_mlp_bwd_tail_tc: ldh (%d), C (%d)", ldh, C);
    CME_CHECK_ARG(rows < (1ll << 31), "cme_mlp_bwd_tail_tc: too many rows");
    const int m_tiles = (int)((rows + 127) / 128), n_tiles = (H + 127) / 128;
    CUtensorMap tmD, tmW, tmA1, tmZ;
    int rc;
    if ((rc = make_map(&tmD, dz2b, 16, (uint64_t)rows, 64, 64, 128))) return rc;
    if ((rc = make_map(&tmW, W2T, 16, (uint64_t)H, 64, 64, 128))) return rc;
    if ((rc = make_map(&tmA1, A1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, 128))) return rc;
    if ((rc = make_map(&tmZ, dZ1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, 128))) return rc;
    if ((rc = opt_in_smem(bwd_tail_kernel, kSmemFt))) return rc;
    const long long tiles = (long long)m_tiles * n_tiles;
    const int grid = (int)std::min<long long>(tiles, cme::sm_count());
    CME_CHECK_ARG(ldg2 >= H, "cme_mlp_bwd_tail_tc: ldg2 (%d) < H (%d)", ldg2, H);
    bwd_tail_kernel<<<grid, kThreads, kSmemFt, cme::as_stream(stream)>>>(tmD, tmW, tmA1, tmZ, gW2, ldg2,
                                                                        H, C, m_tiles, n_tiles, part);
    CME_LAUNCH_CHECK();
    if (part) {
        dw2_reduce_kernel<<<dim3(n_tiles, 16), 128, 0, cme::as_stream(stream)>>>(part, gW2, ldg2, H, C, m_tiles,
                                                                                 n_tiles, grid);
        CME_LAUNCH_CHECK();
    }
    return 0;
}

int cme_mlp_dw2_tc(const void* A1, const void* dz2b, float* gW2, int64_t rows, int H, int ldh, int C,
                   void* stream) {
    CME_CHECK_ARG(A1 && dz2b && gW2 && rows > 0 && H > 0, "cme_mlp_dw2_tc: bad argument");
    CME_CHECK_ARG(C >= 1 && C <= 16, "cme_mlp_dw2_tc: 1..16 classes (got %d)", C);
    CME_CHECK_ARG(ldh % 8 == 0 && ldh >= H, "cme_mlp_dw2_tc: ldh (%d)", ldh);
    CME_CHECK_ARG(rows < (1ll << 31), "cme_mlp_dw2_tc: too many rows");
    const int m_tiles = (H + BM - 1) / BM;
    const int kb_total = (int)((rows + BK - 1) / BK);
    CUtensorMap tmA, tmB;
    int rc;
    if ((rc = make_map(&tmA, A1, (uint64_t)H, (uint64_t)rows, (uint64_t)ldh, 64, BK))) return rc;
    // dz2b rows are 64 bf16 wide of which 16 carry data: bound the map at 16 so the rest is
    // zero-filled on chip and never read from memory
    if ((rc = make_map(&tmB, dz2b, 16, (uint64_t)rows, 64, 64, BK))) return rc;
    if ((rc = opt_in_smem(gemm_mnmajor_kernel, kSmemMNMajor))) return rc;
    const int ksplit = choose_ksplit(m_tiles, kb_total, cme::sm_count());
    const long long items = (long long)m_tiles * ksplit;
    const int grid = (int)std::min<long long>(items, cme::sm_count());
    // result tile is [h][c]; store it transposed into gW2[c][h]
    gemm_mnmajor_kernel<<<grid, kThreads, kSmemMNMajor, cme::as_stream(stream)>>>(
        tmA, tmB, gW2, H, H, C, m_tiles, 1, kb_total, 16, 512u, 1, ksplit, DwSync{}, upd::UpdateArgs{});
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
