// Forward pass in one tensor-core kernel plus a tiny tail (sm_100a).
//
//   fwd_rider_kernel   A1 = sigmoid(X * W1c^T) on CTA pairs (tcgen05.mma.cta_group::2, 256 x 256
//                      tiles, TMA-fed ring, TMEM accumulators) AND the partial logits
//                      z2 += A1_tile * W2c_tile^T from the SAME staged tile: when an epilogue group
//                      has written a 128 x 64 chunk of A1 into shared memory for its TMA store, that
//                      chunk is already in the K-major 128-byte-swizzled layout an MMA operand
//                      wants, so the MMA lane issues a 256 x 16 x 64 "rider" product on it.  The
//                      rider accumulates in TMEM columns of the accumulator stage that the
//                      epilogue has just drained, so it needs no TMEM of its own.  A1 is therefore
//                      never re-read for layer 2: the reference runs a second GEMM over it
//                      (ref fpcode/neural_network.cpp:300-323), round 1 of this repo streamed it once
//                      more through gemm_kmajor_kernel<2> (1.07 GB, 0.20 ms at batch 65536).
//   softmax_tail_kernel  sums the partial logits in a fixed order, adds b2, and finishes the
//                      forward pass per row: softmax, cross-entropy, arg-max, dz2 (fp32 and
//                      bf16), and the loss and db2 through per-block partials that the last block
//                      adds up in block order, so all of it is bit-reproducible
//                      (ref K5 + K9 + K7(db2), fpcode/gpu_func.cu:300-329, 411-439, 359-382).
//
// Work items are (row block of 256, slice of the column tiles): one CTA pair walks the tiles
// of its slice for one row block, so the partial logits of a row stay in the epilogue's
// registers across tiles and are written once per item, [slice][row][16] fp32.
#include "common.cuh"
#include "tc_common.cuh"
#include "tc_host.cuh"

#include <algorithm>
#include <cstdlib>

namespace {

using namespace tc;

constexpr int kThreadsF = 352;               // producer warp + MMA warp + 8 epilogue warps + rider warp
constexpr int kStagesF = 4;                  // 4 x 32 KB ring + 4 staging chunks + W2c slots fit 227 KB
constexpr int BKf = 64;
constexpr int kAB = 128 * BKf * 2;           // 16 KB: this CTA's 128 rows of A, or its half of B
constexpr int kStg = 128 * 64 * 2;           // 16 KB staging chunk (128 rows x 64 bf16)
constexpr int kW2Chunk = 8 * 64 * 2;         // 1 KB: this CTA's 8 of the 16 W2c rows, 64 columns
constexpr int kW2Slot = 4 * kW2Chunk;        // one 256-column tile
constexpr int kXbuf = 128 * 16 * 4;          // group 1 -> group 0 hand-over of partial logits
constexpr int kSmemF = kStagesF * 2 * kAB + 4 * kStg + 2 * kW2Slot + kXbuf + 1024 + 512;

struct BarriersF {
    uint64_t full[kStagesF];      // leader: TMA bytes of both CTAs
    uint64_t empty[kStagesF];     // both CTAs (multicast commit)
    uint64_t tmem_full[2];        // both CTAs (multicast commit)
    uint64_t tmem_empty[2];       // leader: 8 epilogue warps x 2 CTAs
    uint64_t w_full[2];           // leader: W2c tile bytes of both CTAs
    uint64_t w_empty[2];          // both CTAs (multicast commit after the tile's riders)
    uint64_t rider_ready[4];      // [group * 2 + chunk], leader: one arrive per CTA when the chunk is staged.
                                  // One barrier per (group, chunk): with one per group a CTA that runs ahead
                                  // could arrive twice (chunks 0 and 1) and complete a phase on its own
    uint64_t rider_done[4];       // [group * 2 + chunk], both CTAs (multicast commit): the rider has read the chunk
    uint32_t tmem_base;
};

__device__ __forceinline__ uint8_t* align1024f(uint8_t* p) {
    return (uint8_t*)(((uintptr_t)p + 1023) & ~(uintptr_t)1023);
}
__device__ __forceinline__ void mbar_arrive_leader_release(uint64_t* bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(smem_u32(bar) & kPeerBitMask)
                 : "memory");
}

struct RiderArgs {
    float* z2part;          // [slices][part_stride] fp32, rows of 16
    long long part_stride;  // floats between slices
    long long rows;
    int slices;             // column tiles of a row block are cut into this many items
    int tiles_per_slice;
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsF, 1)
fwd_rider_kernel(const __grid_constant__ CUtensorMap tmA /* X   [rows][K]  box 64 x 128 */,
                 const __grid_constant__ CUtensorMap tmB /* W1c [Hp][K]    box 64 x 128 */,
                 const __grid_constant__ CUtensorMap tmC /* A1  [rows][Hp] box 64 x 128 (store) */,
                 const __grid_constant__ CUtensorMap tmW /* W2c [16][Hp]   box 64 x 8 */,
                 int m_blocks, int K, RiderArgs ra) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024f(smem_raw);
    uint8_t* sA = smem;
    uint8_t* sB = sA + kStagesF * kAB;
    uint8_t* sC = sB + kStagesF * kAB;
    uint8_t* sW = sC + 4 * kStg;
    float* xbuf = reinterpret_cast<float*>(sW + 2 * kW2Slot);
    BarriersF* bars = (BarriersF*)(reinterpret_cast<uint8_t*>(xbuf) + kXbuf);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();          // 0 = leader
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tmA);
        prefetch_tmap(&tmB);
        prefetch_tmap(&tmC);
        prefetch_tmap(&tmW);
        for (int s = 0; s < kStagesF; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&bars->tmem_full[a], 1);
            mbar_init(&bars->tmem_empty[a], 16);     // 8 epilogue warps x 2 CTAs
            mbar_init(&bars->w_full[a], 1);
            mbar_init(&bars->w_empty[a], 1);
        }
        for (int a = 0; a < 4; ++a) {
            mbar_init(&bars->rider_ready[a], 2);     // the group's leader thread of each CTA
            mbar_init(&bars->rider_done[a], 1);
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

    const int items = m_blocks * ra.slices;
    const int nk = (K + BKf - 1) / BKf;
    const int tps = ra.tiles_per_slice;

    if (warp == 0) {
        if (lane == 0) {
            // ---------------- TMA producer (both CTAs) ----------------
            int stage = 0;
            uint32_t phase = 0;
            int tl = 0;                                   // tiles this pair has started
            for (int item = pair; item < items; item += npairs) {
                const int mb = item / ra.slices, sl = item % ra.slices;
                const int m0 = mb * 256 + (int)cta * 128;
                for (int t = 0; t < tps; ++t, ++tl) {
                    const int n_tile = sl * tps + t;
                    const int n0 = n_tile * 256 + (int)cta * 128;
                    for (int kb = 0; kb < nk; ++kb) {
                        mbar_wait_bounded(&bars->empty[stage], phase ^ 1);
                        if (cta == 0) mbar_expect_tx(&bars->full[stage], 4 * kAB);
                        tma_load_2d_2cta(sA + stage * kAB, &tmA, &bars->full[stage], kb * BKf, m0);
                        tma_load_2d_2cta(sB + stage * kAB, &tmB, &bars->full[stage], kb * BKf, n0);
                        if (++stage == kStagesF) { stage = 0; phase ^= 1; }
                    }
                    // this CTA's 8 rows of the W2c tile, after the tile's operands: the riders need
                    // it only once the tile's accumulator is complete, and the slot it goes into
                    // was released by the riders of two tiles ago
                    const int ws = tl & 1;
                    mbar_wait_bounded(&bars->w_empty[ws], (((uint32_t)tl >> 1) & 1) ^ 1);
                    if (cta == 0) mbar_expect_tx(&bars->w_full[ws], 2 * kW2Slot);
                    for (int c = 0; c < 4; ++c)
                        tma_load_2d_2cta(sW + ws * kW2Slot + c * kW2Chunk, &tmW, &bars->w_full[ws],
                                         n_tile * 256 + c * 64, (int)cta * 8);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0 && cta == 0) {
            // ---------------- MMA issuer (leader CTA): the main products ----------------
            const uint32_t idesc_main = idesc_bf16(256, 256, 0, 0);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int item = pair; item < items; item += npairs) {
                for (int t = 0; t < tps; ++t) {
                    mbar_wait_bounded(&bars->tmem_empty[acc], acc_phase ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                    for (int kb = 0; kb < nk; ++kb) {
                        mbar_wait_bounded(&bars->full[stage], phase);
                        tc_fence_after();
                        const int kvalid = min(BKf, K - kb * BKf);
                        const int nmma = (kvalid + 15) >> 4;
                        const uint64_t da = smem_desc_sw128(smem_u32(sA + stage * kAB), 16, 1024);
                        const uint64_t db = smem_desc_sw128(smem_u32(sB + stage * kAB), 16, 1024);
                        for (int j = 0; j < nmma; ++j)
                            umma_bf16_2cta(d_tmem, da + (uint64_t)(2 * j), db + (uint64_t)(2 * j), idesc_main,
                                           (kb | j) != 0);
                        umma_commit_2cta(&bars->empty[stage]);
                        if (++stage == kStagesF) { stage = 0; phase ^= 1; }
                    }
                    umma_commit_2cta(&bars->tmem_full[acc]);
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                }
            }
        }
        __syncwarp();
    } else if (warp == 10) {
        if (lane == 0 && cta == 0) {
            // ---------------- rider issuer (leader CTA) ----------------
            // A lane of its own, so the main issuer never polls: it sleeps on the "chunk staged"
            // barriers and issues the 256 x 16 x 64 product of each staged chunk with the W2c tile.
            // Riders of one group run in chunk order (chunk 1 accumulates on chunk 0); against the
            // main products no order is needed, they write other TMEM columns.
            const uint32_t idesc_rider = idesc_bf16(256, 16, 0, 0);
            uint32_t tl = 0;
            for (int item = pair; item < items; item += npairs) {
                for (int t = 0; t < tps; ++t, ++tl) {
                    const uint32_t ws = tl & 1, acc = tl & 1;
                    mbar_wait_bounded(&bars->w_full[ws], (tl >> 1) & 1);
                    uint32_t pending = 0xF;                       // bit g * 2 + cc
                    const long long t0 = clock64();
                    while (pending) {
                        if (clock64() - t0 > (4ll << 30)) __trap();
#pragma unroll
                        for (int g = 0; g < 2; ++g) {
                            const int cc = (pending >> (g * 2)) & 1 ? 0 : 1;       // chunk 0 before chunk 1
                            const int idx = g * 2 + cc;
                            if (!((pending >> idx) & 1)) continue;
                            if (!mbar_try_wait(&bars->rider_ready[idx], tl & 1)) continue;
                            tc_fence_after();
                            const uint32_t d = tmem_base + acc * 256u + (uint32_t)g * 128u;
                            const uint32_t a0 = smem_u32(sC + idx * kStg);
                            const uint32_t b0 = smem_u32(sW + ws * kW2Slot + idx * kW2Chunk);
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                umma_bf16_2cta(d, smem_desc_sw128(a0, 16, 1024) + (uint64_t)(2 * j),
                                               smem_desc_sw128(b0, 16, 1024) + (uint64_t)(2 * j), idesc_rider,
                                               (cc | j) != 0);
                            umma_commit_2cta(&bars->rider_done[idx]);
                            pending &= ~(1u << idx);
                        }
                    }
                    umma_commit_2cta(&bars->w_empty[ws]);         // the W2c slot is free again
                }
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue: 8 warps, two groups of four ----------------
        // Group g owns the 64-column chunks 2g and 2g + 1 of every tile, one staging buffer, one
        // named barrier and one TMA-store leader; its four warps cover the four TMEM sub-partitions.
        const int q = warp & 3;
        const int grp = (warp - 2) >> 2;
        const int row = q * 32 + lane;
        const bool leader = (((warp - 2) & 3) == 0 && lane == 0);
        int acc = 0;
        uint32_t acc_phase = 0;
        uint32_t rid = 0;                                  // chunks this group has staged
        for (int item = pair; item < items; item += npairs) {
            const int mb = item / ra.slices, sl = item % ra.slices;
            const int m0 = mb * 256 + (int)cta * 128;
            float zacc[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) zacc[c] = 0.f;
            for (int t = 0; t < tps; ++t) {
                const int n0 = (sl * tps + t) * 256;
                mbar_wait_bounded(&bars->tmem_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
#pragma unroll
                for (int cc = 0; cc < 2; ++cc) {
                    const int c = grp * 2 + cc;
                    uint32_t v0[32], v1[32];
                    tmem_ld_32x32(t_row + (uint32_t)(c * 64), v0);
                    tmem_ld_32x32(t_row + (uint32_t)(c * 64 + 32), v1);
                    tmem_ld_wait();
                    uint32_t pk[32];
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        pk[e] = sigmoid2_bf16(__uint_as_float(v0[2 * e]), __uint_as_float(v0[2 * e + 1]));
                        pk[16 + e] = sigmoid2_bf16(__uint_as_float(v1[2 * e]), __uint_as_float(v1[2 * e + 1]));
                    }
                    // Two staging buffers per group (one per chunk of the tile), so staging chunk 1
                    // does not wait for the rider of chunk 0, which sits in the tensor pipe behind
                    // every main MMA already issued.  Buffer cc is free once the TMA store of the
                    // previous tile's chunk cc has read it (at most one younger store pending) AND
                    // that chunk's rider has completed (one barrier per (group, chunk): it completes
                    // once per tile, so its parity names the tile).
                    uint8_t* stg = sC + (grp * 2 + cc) * kStg;
                    const uint32_t tl = rid >> 1;                      // tiles this pair has finished
                    if (leader) tma_store_wait_read<1>();
                    if (tl > 0) mbar_wait_bounded(&bars->rider_done[grp * 2 + cc], (tl - 1) & 1);
                    named_bar_sync(1 + grp, 128);
#pragma unroll
                    for (int chunk = 0; chunk < 8; ++chunk) {
                        const uint32_t off = (uint32_t)row * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
                        *reinterpret_cast<uint4*>(stg + off) =
                            make_uint4(pk[4 * chunk], pk[4 * chunk + 1], pk[4 * chunk + 2], pk[4 * chunk + 3]);
                    }
                    fence_proxy_async_smem();
                    tc_fence_before();         // my reads of these TMEM columns precede the rider that reuses them
                    named_bar_sync(1 + grp, 128);
                    if (leader) {
                        tma_store_2d(&tmC, stg, n0 + c * 64, m0);
                        tma_store_commit();
                        mbar_arrive_leader_release(&bars->rider_ready[grp * 2 + cc]);
                    }
                    ++rid;
                }
                // both riders of this group for this tile: 16 partial logits per row
                mbar_wait_bounded(&bars->rider_done[grp * 2 + 1], ((rid >> 1) - 1) & 1);
                tc_fence_after();
                uint32_t z[16];
                tmem_ld_32x16(t_row + (uint32_t)(grp * 128), z);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 16; ++c) zacc[c] += __uint_as_float(z[c]);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_leader(&bars->tmem_empty[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
            // group 1 hands its partial logits to group 0, which writes the item's partial rows
            if (grp == 1) {
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4)
                    *reinterpret_cast<float4*>(xbuf + row * 16 + 4 * k4) =
                        make_float4(zacc[4 * k4], zacc[4 * k4 + 1], zacc[4 * k4 + 2], zacc[4 * k4 + 3]);
            }
            named_bar_sync(3, 256);
            if (grp == 0) {
                const long long grow = (long long)m0 + row;
                if (grow < ra.rows) {
                    float* dst = ra.z2part + (long long)sl * ra.part_stride + grow * 16;
#pragma unroll
                    for (int k4 = 0; k4 < 4; ++k4) {
                        const float4 o = *reinterpret_cast<const float4*>(xbuf + row * 16 + 4 * k4);
                        *reinterpret_cast<float4*>(dst + 4 * k4) =
                            make_float4(zacc[4 * k4] + o.x, zacc[4 * k4 + 1] + o.y, zacc[4 * k4 + 2] + o.z,
                                        zacc[4 * k4 + 3] + o.w);
                    }
                }
            }
            named_bar_sync(3, 256);
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
// Tail of the forward pass: one thread per row
// =============================================================================
constexpr int kTailThreads = 256;

struct TailArgs {
    const float* z2part;
    long long part_stride;
    int slices;
    const float* b2;
    const int32_t* labels;
    const CmeStepArgs* sa;     // non-null: labels and scale from the device-side step arguments
    float* dz2;                // [rows][16]
    __nv_bfloat16* dz2b;       // [rows][64], 16 live
    int32_t* pred;
    float* loss_sum;           // += (one add per launch), may be null
    float* gb2;                // stored, may be null
    float* block_part;         // [blocks][20]: 16 column sums of dz2 + the loss of each block
    uint32_t* counter;         // zero between launches
    float scale;
    long long rows;
    int classes;
};

__global__ void __launch_bounds__(kTailThreads)
softmax_tail_kernel(const TailArgs a) {
    __shared__ float warp_part[kTailThreads / 32][17];
    __shared__ uint32_t is_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long row = (long long)blockIdx.x * kTailThreads + threadIdx.x;
    const bool valid = row < a.rows;
    const int C = a.classes;
    const int32_t* labels = a.sa ? a.sa->labels : a.labels;
    const float scale = a.sa ? a.sa->scale : a.scale;
    float z[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) z[c] = 0.f;
    if (valid) {
        // fixed order: slice 0, 1, ...  Four slices' loads are issued before their adds: with one
        // slice at a time the 16 slices of an 8192-row step were 16 dependent trips to L2 (16 us)
        int s = 0;
        for (; s + 4 <= a.slices; s += 4) {
            float4 v[4][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float4* p = reinterpret_cast<const float4*>(a.z2part + (long long)(s + u) * a.part_stride + row * 16);
#pragma unroll
                for (int k = 0; k < 4; ++k) v[u][k] = p[k];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    z[4 * k] += v[u][k].x; z[4 * k + 1] += v[u][k].y; z[4 * k + 2] += v[u][k].z; z[4 * k + 3] += v[u][k].w;
                }
        }
        for (; s < a.slices; ++s) {
            const float4* p = reinterpret_cast<const float4*>(a.z2part + (long long)s * a.part_stride + row * 16);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float4 v = p[k];
                z[4 * k] += v.x; z[4 * k + 1] += v.y; z[4 * k + 2] += v.z; z[4 * k + 3] += v.w;
            }
        }
    }
    float mx = -3.0e38f;
    int arg = 0;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        z[c] = (c < C) ? z[c] + __ldg(a.b2 + c) : -3.0e38f;
        if (z[c] > mx) { mx = z[c]; arg = c; }
    }
    float ssum = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        z[c] = (c < C) ? __expf(z[c] - mx) : 0.f;
        ssum += z[c];
    }
    const float inv = 1.f / ssum;
    const int lab = (valid && labels) ? labels[row] : -1;
    float my_loss = 0.f;
    float out[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        const float p = z[c] * inv;
        out[c] = (valid && c < C) ? (p - (c == lab ? 1.f : 0.f)) * scale : 0.f;
        if (valid && c == lab) my_loss = -__logf(fmaxf(p, 1e-30f));
    }
    if (valid) {
        if (a.dz2) {
            float4* d = reinterpret_cast<float4*>(a.dz2 + row * 16);
#pragma unroll
            for (int k = 0; k < 4; ++k) d[k] = make_float4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
        }
        if (a.dz2b) {
            uint4* d = reinterpret_cast<uint4*>(a.dz2b + row * 64);
            d[0] = make_uint4(pack_bf16(out[0], out[1]), pack_bf16(out[2], out[3]), pack_bf16(out[4], out[5]),
                              pack_bf16(out[6], out[7]));
            d[1] = make_uint4(pack_bf16(out[8], out[9]), pack_bf16(out[10], out[11]), pack_bf16(out[12], out[13]),
                              pack_bf16(out[14], out[15]));
        }
        if (a.pred) a.pred[row] = arg;
    }
    if (!a.block_part) return;
    // loss and db2: warp sums (fixed shuffle tree), warps in order, blocks in order
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        const float t = cme::warp_sum(out[c]);
        if (lane == 0) warp_part[warp][c] = t;
    }
    {
        const float t = cme::warp_sum(my_loss);
        if (lane == 0) warp_part[warp][16] = t;
    }
    __syncthreads();
    if (threadIdx.x < 17) {
        float t = 0.f;
        for (int w = 0; w < kTailThreads / 32; ++w) t += warp_part[w][threadIdx.x];
        a.block_part[(size_t)blockIdx.x * 20 + threadIdx.x] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        is_last = atomicAdd(a.counter, 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    if (threadIdx.x < 17) {
        float t = 0.f;
        for (unsigned b = 0; b < gridDim.x; ++b) t += __ldcg(a.block_part + (size_t)b * 20 + threadIdx.x);
        if (threadIdx.x < 16) {
            if (a.gb2 && threadIdx.x < C) a.gb2[threadIdx.x] = t;
        } else if (a.loss_sum) {
            atomicAdd(a.loss_sum, t);
        }
    }
    if (threadIdx.x == 0) *a.counter = 0;
}

// slices of the column tiles: the smallest power-of-two divisor of n2 whose item count fills
// the CTA pairs to within 5 % of the best any divisor reaches
int choose_slices(int m_blocks, int n2, int npairs) {
    double best = 0.0;
    for (int s = 1; s <= n2; s *= 2) {
        if (n2 % s) break;
        const long long items = (long long)m_blocks * s;
        best = std::max(best, (double)items / (double)((items + npairs - 1) / npairs * npairs));
    }
    for (int s = 1; s <= n2; s *= 2) {
        if (n2 % s) break;
        const long long items = (long long)m_blocks * s;
        if ((double)items / (double)((items + npairs - 1) / npairs * npairs) >= best - 0.05) return s;
    }
    return 1;
}

}  // namespace

extern "C" {

int cme_mlp_fwd_fused_ok(int64_t rows, int H) {
    const int Hp = (H + 63) / 64 * 64;
    return (Hp % 256 == 0 && rows >= 256) ? 1 : 0;
}

size_t cme_mlp_fwd_scratch_floats(int64_t max_rows) {
    // partial logits: up to 32 slices of 16 floats per row (fewer slices for many rows), then the
    // tail's per-block partials and its counter
    const int64_t part = std::max<int64_t>(4 * max_rows, 32 * std::min<int64_t>(max_rows, 16384)) * 16;
    return (size_t)(part + ((max_rows + kTailThreads - 1) / kTailThreads) * 20 + 64);
}

int cme_mlp_fwd_fused(const void* X, const void* W1c, const void* W2c, const float* b2, const int32_t* labels,
                      const CmeStepArgs* step_args, void* A1, float* dz2, void* dz2b, float* loss_sum,
                      int32_t* pred, float* gb2, float* scratch, int64_t max_rows, float scale, int64_t rows,
                      int K, int H, int ldx, int ldh, int C, void* stream) {
    CME_CHECK_ARG(X && W1c && W2c && b2 && A1 && scratch && rows > 0 && rows <= max_rows && K > 0 && H > 0,
                  "cme_mlp_fwd_fused: bad argument");
    CME_CHECK_ARG(C >= 1 && C <= 16, "cme_mlp_fwd_fused: 1..16 classes (got %d)", C);
    CME_CHECK_ARG(labels || step_args || (!dz2 && !dz2b), "cme_mlp_fwd_fused: gradients need labels");
    const int Hp = (H + 63) / 64 * 64;
    CME_CHECK_ARG(cme_mlp_fwd_fused_ok(rows, H), "cme_mlp_fwd_fused: needs H rounded up to 64 to be a multiple "
                  "of 256 and at least 256 rows (H=%d rows=%lld); use cme_mlp_fwd1 + cme_mlp_fwd2_tc", H, (long long)rows);
    CME_CHECK_ARG(ldx % 8 == 0 && ldh % 8 == 0 && ldx >= K && ldh >= Hp && rows < (1ll << 31),
                  "cme_mlp_fwd_fused: leading dimensions (ldx=%d ldh=%d)", ldx, ldh);
    cudaStream_t st = cme::as_stream(stream);
    const int m_blocks = (int)((rows + 255) / 256), n2 = Hp / 256;
    const int npairs = cme::sm_count() / 2;
    int slices = choose_slices(m_blocks, n2, npairs);
    static const int dev_slices = getenv("CME_FWD_SLICES") ? atoi(getenv("CME_FWD_SLICES")) : 0;      // tuning hook
    if (dev_slices > 0 && n2 % dev_slices == 0) slices = dev_slices;
    const int64_t part_floats = std::max<int64_t>(4 * max_rows, 32 * std::min<int64_t>(max_rows, 16384)) * 16;
    while (slices > 1 && (int64_t)slices * max_rows * 16 > part_floats) slices /= 2;
    CUtensorMap tmA, tmB, tmC, tmW;
    int rc;
    if ((rc = tc_host::make_map(&tmA, X, (uint64_t)K, (uint64_t)rows, (uint64_t)ldx, 64, 128))) return rc;
    if ((rc = tc_host::make_map(&tmB, W1c, (uint64_t)K, (uint64_t)Hp, (uint64_t)ldx, 64, 128))) return rc;
    if ((rc = tc_host::make_map(&tmC, A1, (uint64_t)Hp, (uint64_t)rows, (uint64_t)ldh, 64, 128))) return rc;
    // the K extent of the rider is the true H: pad columns of A1 hold sigmoid(0) and must meet zeros
    if ((rc = tc_host::make_map(&tmW, W2c, (uint64_t)H, 16, (uint64_t)ldh, 64, 8))) return rc;
    if ((rc = tc_host::opt_in_smem(fwd_rider_kernel, kSmemF))) return rc;
    RiderArgs ra{scratch, (long long)max_rows * 16, (long long)rows, slices, n2 / slices};
    const long long items = (long long)m_blocks * slices;
    const int pairs = (int)std::min<long long>(items, npairs);
    fwd_rider_kernel<<<2 * pairs, kThreadsF, kSmemF, st>>>(tmA, tmB, tmC, tmW, m_blocks, K, ra);
    CME_LAUNCH_CHECK();
    const int blocks = (int)((rows + kTailThreads - 1) / kTailThreads);
    float* block_part = scratch + part_floats;
    uint32_t* counter = reinterpret_cast<uint32_t*>(block_part + (size_t)((max_rows + kTailThreads - 1) / kTailThreads) * 20);
    TailArgs ta{scratch, (long long)max_rows * 16, slices, b2, labels, step_args, dz2, (__nv_bfloat16*)dz2b, pred,
                loss_sum, gb2, (gb2 || loss_sum) ? block_part : nullptr, counter, scale, (long long)rows, C};
    softmax_tail_kernel<<<blocks, kTailThreads, 0, st>>>(ta);
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
