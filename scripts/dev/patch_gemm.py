p='csrc/gemm_tc.cu'
s=open(p).read()

def rep(old, new, count=1):
    global s
    assert old in s, old[:80]
    s = s.replace(old, new, count)

rep('#include "common.cuh"\n#include "tc_common.cuh"\n', '#include "common.cuh"\n#include "tc_common.cuh"\n#include "update_stream.cuh"\n')
rep("""constexpr int kThreads = 192;
""","""constexpr int kThreads = 192;
constexpr int kUpdWarps = 4;                          // update warps of the dW1 GEMMs (warps 6..9)
constexpr int kThreadsDw = kThreads + 32 * kUpdWarps; // launched with kThreads when there is no fused update
""")
rep("""    uint64_t x_full, x_empty;     // split-K exchange between the two CTAs of a cluster (softmax kernel)
    uint32_t tmem_base;
};""","""    uint64_t x_full, x_empty;     // split-K exchange between the two CTAs of a cluster (softmax kernel)
    uint32_t tmem_base;
    uint32_t upd_ok;              // update warps of the dW1 GEMM (update_stream.cuh)
};""")
rep("""    uint64_t tmem_empty[2];       // used in the leader: 4 epilogue warps x 2 CTAs
    uint32_t tmem_base;
};""","""    uint64_t tmem_empty[2];       // used in the leader: 4 epilogue warps x 2 CTAs
    uint32_t tmem_base;
    uint32_t upd_ok;              // update warps of the dW1 GEMM (update_stream.cuh)
};""")

# ---- SoftmaxEpi indirect args ----
rep("""    float scale;
    long long rows;
    int classes;
};""","""    float scale;
    long long rows;
    int classes;
    const CmeStepArgs* sa;   // non-null: labels and scale come from the device-side step arguments
};""")
rep("""                const long long grow = (long long)m0 + row;
                const bool valid = grow < sm.rows;
                const int C = sm.classes;""","""                const long long grow = (long long)m0 + row;
                const bool valid = grow < sm.rows;
                const int C = sm.classes;
                const int32_t* labels = sm.sa ? sm.sa->labels : sm.labels;
                const float scale = sm.sa ? sm.sa->scale : sm.scale;""")
rep("const int lab = (valid && sm.labels) ? sm.labels[grow] : -1;","const int lab = (valid && labels) ? labels[grow] : -1;")
rep("out[c] = (valid && c < C) ? (p - (c == lab ? 1.f : 0.f)) * sm.scale : 0.f;","out[c] = (valid && c < C) ? (p - (c == lab ? 1.f : 0.f)) * scale : 0.f;")

# ---- ordered split-K helpers ----
marker = "// =============================================================================\n// MN-major split-K GEMM with fp32 reduction epilogue\n"
helpers = r'''// =============================================================================
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
            while (cme::ld_acquire_gpu(ctr) != (uint32_t)kc) {
                if (clock64() - t0 > (4ll << 30)) {
                    if (!sy.status) __trap();
                    atomicCAS(sy.status, 0u, 3u | ((uint32_t)kc << 8) | ((uint32_t)(ctr - sy.part_ctr) << 12));
                    break;
                }
            }
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
            __threadfence_system();
            for (int r = 0; r < sy.nranks; ++r) cme::red_release_sys_add(sy.pads[r] + cme::kPadUnit0 + unit, 1u);
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

'''
rep(marker, helpers + marker)

# ---- single-CTA kernel ----
rep("""__global__ void __launch_bounds__(kThreads, 1)
gemm_mnmajor_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                    int kb_total, int BN, uint32_t tmem_cols, int out_transposed, int ksplit) {""","""__global__ void __launch_bounds__(kThreadsDw, 1)
gemm_mnmajor_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                    int kb_total, int BN, uint32_t tmem_cols, int out_transposed, int ksplit, DwSync sy,
                    const __grid_constant__ upd::UpdateArgs ua) {""")
rep("""    // Work items are (k-chunk, m-tile, n-tile) with n fastest, then m, then the k-chunk, handed
    // out round-robin.  CTAs that run at the same time therefore sit on the same k-chunk: the
    // n-tiles of one m-tile read the same A rows and every CTA reads the same B rows, so each
    // operand byte comes out of HBM once and is served to the other CTAs by L2.
    const int tiles_mn = m_tiles * n_tiles;
    const int items = tiles_mn * ksplit;
    const int nb_boxes = (BN + 63) / 64;
    const uint32_t b_bytes = (uint32_t)nb_boxes * 64 * BK * 2;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int kc = item / tiles_mn, tile = item % tiles_mn;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
                for (int kb = kb0; kb < kb1; ++kb) {""","""    // Work items are (k-chunk, m-tile, n-tile) with n fastest, then m, then the k-chunk (inside a
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
                for (int kb = kb0; kb < kb1; ++kb) {""")
rep("""            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int kc = item / tiles_mn;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(sA + stage * kABytes);
                    const uint32_t b0 = smem_u32(sB + stage * kBBytesMax);""","""            for (int item = blockIdx.x; item < items; item += gridDim.x) {
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
                    const uint32_t b0 = smem_u32(sB + stage * kBBytesMax);""")
rep("""        __syncwarp();
    } else {
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int tile = item % tiles_mn;
            const int m0 = (tile / n_tiles) * BM, n0 = (tile % n_tiles) * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const int h = m0 + q * 32 + lane;
            float* grow = G + (size_t)h * ldg;
            for (int c = 0; c < BN / 16; ++c) {
                uint32_t v[16];
                tmem_ld_32x16(t_row + (uint32_t)(c * 16), v);
                tmem_ld_wait();
                if (h < M && out_transposed) {""","""        __syncwarp();
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
                if (h < M && out_transposed) {""")
rep("""            if (acc == 0) acc_phase ^= 1;
        }
    }
    teardown(bars, tmem_cols, warp);
}

// =============================================================================
// MN-major split-K GEMM on CTA pairs: 256 x 256 tiles, fp32 reduction epilogue""","""            if (acc == 0) acc_phase ^= 1;
        }
    } else if (ua.enabled) {
        // update warps: reduce every unit of the gradient across the GPUs as soon as all ranks
        // have published it, while the warps above are still computing later units
        upd::run_update(ua, blockIdx.x, gridDim.x, threadIdx.x - kThreads, 32 * kUpdWarps, 2, &bars->upd_ok);
    }
    teardown(bars, tmem_cols, warp);
}

// =============================================================================
// MN-major split-K GEMM on CTA pairs: 256 x 256 tiles, fp32 reduction epilogue""")

# ---- pair kernel ----
rep("""__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm_mnmajor2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                     int kb_total, int ksplit, int BN) {""","""__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreadsDw, 1)
gemm_mnmajor2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     float* __restrict__ G, int ldg, int M, int N, int m_tiles, int n_tiles,
                     int kb_total, int ksplit, int BN, DwSync sy,
                     const __grid_constant__ upd::UpdateArgs ua) {""")
rep("""            for (int item = pair; item < items; item += npairs) {
                const int kc = item / tiles_mn, tile = item % tiles_mn;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                const int m0 = (tile / n_tiles) * 256 + (int)cta * 128;
                const int n0 = (tile % n_tiles) * BN + (int)cta * half_n;""","""            for (int item = pair; item < items; item += npairs) {
                const DwItem it = dw_item(item, m_tiles, n_tiles, ksplit, sy.sblocks);
                const int kc = it.kc;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                const int m0 = it.m * 256 + (int)cta * 128;
                const int n0 = it.n * BN + (int)cta * half_n;""")
rep("""            for (int item = pair; item < items; item += npairs) {
                const int kc = item / tiles_mn;
                const int kb0 = (int)((long long)kb_total * kc / ksplit);
                const int kb1 = (int)((long long)kb_total * (kc + 1) / ksplit);
                mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&bars->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a0 = smem_u32(sA + stage * kABytes);
                    const uint32_t b0 = smem_u32(sB + stage * kB2Bytes);""","""            for (int item = pair; item < items; item += npairs) {
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
                    const uint32_t b0 = smem_u32(sB + stage * kB2Bytes);""")
rep("""        __syncwarp();
    } else {
        const int q = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = pair; item < items; item += npairs) {
            const int tile = item % tiles_mn;
            const int m0 = (tile / n_tiles) * 256 + (int)cta * 128;
            const int n0 = (tile % n_tiles) * BN;
            mbar_wait(&bars->tmem_full[acc], acc_phase);
            tc_fence_after();
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
            const int h = m0 + q * 32 + lane;
            float* grow = G + (size_t)h * ldg;
            for (int c = 0; c < BN / 16; ++c) {
                if (n0 + c * 16 >= N) break;          // warp-uniform: nothing left in this row block
                uint32_t v[16];
                tmem_ld_32x16(t_row + (uint32_t)(c * 16), v);
                tmem_ld_wait();
                if (h < M) {
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
            if (lane == 0) mbar_arrive_leader(&bars->tmem_empty[acc]);
            acc ^= 1;
            if (acc == 0) acc_phase ^= 1;
        }
    }""","""        __syncwarp();
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
    }""")

# ---- host ----
old_start = s.index("int cme_mlp_dw1(const void* dZ1, const void* X, float* gW1, int64_t rows, int K, int H, int ldx,")
old_end = s.index("int cme_mlp_fwd2_tc(const void* A1, const void* W2c")
host = r'''// Scratch of the ordered split-K (see DwSync): kDwPartWords part counters followed by
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
                      int unit_rows, uint32_t* status, const upd::UpdateArgs* fused, cudaStream_t st) {
    CME_CHECK_ARG(unit_rows == 128 || unit_rows == 256, "cme_mlp_dw1: unit_rows is 128 or 256 (got %d)", unit_rows);
    int rc0 = dw1_check_args(dZ1, X, gW1, rows, K, H, ldx, ldh, ldg);
    if (rc0) return rc0;
    CME_CHECK_ARG(scratch, "cme_mlp_dw1: no scratch");
    upd::UpdateArgs ua{};
    if (fused) ua = *fused;
    const int threads = ua.enabled ? kThreadsDw : kThreads;
    // output tile: 128 rows of H  x  BN columns of K, BN a multiple of 16, <= 256
    const int n_tiles = (K + 255) / 256;
    const int BN = ((K + n_tiles - 1) / n_tiles + 15) / 16 * 16;
    const int m_tiles = (H + BM - 1) / BM;
    const int kb_total = (int)((rows + BK - 1) / BK);
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
    // B (the minibatch) is re-read once per group of row blocks.  When it stays in L2 anyway, every
    // row block is its own group, so row blocks complete one after the other while the kernel runs
    // and the update warps overlap almost all of it; a minibatch that does not fit is walked
    // in at most 4 groups (4 passes over X from HBM, hidden under a compute-bound kernel).
    const bool x_in_l2 = (double)rows * ldx * 2.0 <= 32.0e6;
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
        const int ks = choose_ksplit(m2 * n2, kb_total, npairs);
        CME_CHECK_ARG(m2 * n2 * 2 <= kDwPartWords && m2 <= kDwUnitWords && m2 <= cme::kPadMaxUnits,
                      "cme_mlp_dw1: %d x %d tiles exceed the counter scratch", m2, n2);
        sy.parts_per_tile = n2 * 2;
        sy.tiles_per_unit = 1;
        sy.m_tiles = m2;
        sy.sblocks = x_in_l2 ? m2 : std::min(m2, 4);
        if ((rc = opt_in_smem(gemm_mnmajor2_kernel, kSmemMNMajor2))) return rc;
        const long long items2 = (long long)m2 * n2 * ks;
        const int pairs = (int)std::min<long long>(items2, npairs);
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
    sy.sblocks = x_in_l2 ? m_tiles : std::min(m_tiles, 4);
    if ((rc = opt_in_smem(gemm_mnmajor_kernel, kSmemMNMajor))) return rc;
    const int ksplit = choose_ksplit(m_tiles * n_tiles, kb_total, cme::sm_count());
    const long long items = (long long)m_tiles * n_tiles * ksplit;
    const int grid = (int)std::min<long long>(items, cme::sm_count());
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
    return dw1_launch(p->dz1, p->xb, p->grads[p->rank], rows, p->D + 1, p->H, p->ldx, p->ldh, p->ldx,
                      (uint32_t*)p->sync_scratch, p->pads_dev, p->nranks, p->unit_rows,
                      p->pad_local + cme::kPadStatus, &ua, cme::as_stream(stream));
}

'''
s = s[:old_start] + host + s[old_end:]

rep("""    gemm_mnmajor_kernel<<<grid, kThreads, kSmemMNMajor, cme::as_stream(stream)>>>(
        tmA, tmB, gW2, H, H, C, m_tiles, 1, kb_total, 16, 512u, 1, ksplit);""","""    gemm_mnmajor_kernel<<<grid, kThreads, kSmemMNMajor, cme::as_stream(stream)>>>(
        tmA, tmB, gW2, H, H, C, m_tiles, 1, kb_total, 16, 512u, 1, ksplit, DwSync{}, upd::UpdateArgs{});""")

# fwd2 indirect
rep("""int cme_mlp_fwd2_tc(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                    float* dz2, void* dz2b, float* loss_sum, int32_t* pred, float* gb2, float scale,
                    int64_t rows, int H, int ldh, int C, void* stream) {
    CME_CHECK_ARG(A1 && W2c && b2 && rows > 0 && H > 0, "cme_mlp_fwd2_tc: bad argument");""","""int cme_mlp_fwd2_tc(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                    float* dz2, void* dz2b, float* loss_sum, int32_t* pred, float* gb2, float scale,
                    int64_t rows, int H, int ldh, int C, void* stream) {
    return cme_mlp_fwd2_tc_ind(A1, W2c, b2, labels, nullptr, dz2, dz2b, loss_sum, pred, gb2, scale, rows, H,
                               ldh, C, stream);
}

int cme_mlp_fwd2_tc_ind(const void* A1, const void* W2c, const float* b2, const int32_t* labels,
                        const CmeStepArgs* step_args, float* dz2, void* dz2b, float* loss_sum, int32_t* pred,
                        float* gb2, float scale, int64_t rows, int H, int ldh, int C, void* stream) {
    CME_CHECK_ARG(A1 && W2c && b2 && rows > 0 && H > 0, "cme_mlp_fwd2_tc: bad argument");""")
rep("""    CME_CHECK_ARG((!dz2 && !dz2b) || labels, "cme_mlp_fwd2_tc: gradients need labels");""","""    CME_CHECK_ARG((!dz2 && !dz2b) || labels || step_args, "cme_mlp_fwd2_tc: gradients need labels");""")
rep("""    SoftmaxEpi sm{b2, labels, dz2, (__nv_bfloat16*)dz2b, loss_sum, pred, gb2, scale, (long long)rows, C};""","""    SoftmaxEpi sm{b2, labels, dz2, (__nv_bfloat16*)dz2b, loss_sum, pred, gb2, scale, (long long)rows, C, step_args};""")
open(p,'w').write(s)
print("patched")
