// hw3 2-D heat diffusion sweeps (ref hw3code/gpuStencil.cu, hw3code/BC.h).
//
// Three sweep shapes, all sharing one arithmetic routine (combine<ORDER>) whose
// term order follows the reference's Stencil<order> (ref gpuStencil.cu:15-43),
// because the order-8 weights cancel heavily in fp32:
//   variant 0  one point per thread straight from global memory (ref K16)
//   variant 1  several rows per thread (ref K17, gpuStencilLoop)
//   variant 2  shared-memory tiles with halo plus register marching along y --
//              the tiled kernel the reference left as an empty stub (ref
//              gpuStencil.cu:263-308).  Each input value is read from global
//              memory once per tile.
// The border decay (ref BC.h:7-38) can ride in the same launch as extra blocks.
// The slab kernel adds the multi-GPU piece: edge rows are stored directly into
// the neighbours' ghost rows over NVLink from inside the sweep.
#include "common.cuh"
#include "tc_common.cuh"

#include <algorithm>
#include <cstdlib>

namespace {

template <int ORDER>
__device__ __forceinline__ float combine(const float* xv, const float* yv, float xcfl,
                                         float ycfl) {
    // xv[k], yv[k] hold the values at offset k-R along x / y, R = ORDER/2
    constexpr int R = ORDER / 2;
    const float c = xv[R];
    if (ORDER == 2) {
        return c + xcfl * (xv[R - 1] + xv[R + 1] - 2.f * c) +
               ycfl * (yv[R + 1] + yv[R - 1] - 2.f * c);
    } else if (ORDER == 4) {
        return c +
               xcfl * (-xv[R + 2] + 16.f * xv[R + 1] - 30.f * c + 16.f * xv[R - 1] - xv[R - 2]) +
               ycfl * (-yv[R + 2] + 16.f * yv[R + 1] - 30.f * c + 16.f * yv[R - 1] - yv[R - 2]);
    } else {
        return c +
               xcfl * (-9.f * xv[R + 4] + 128.f * xv[R + 3] - 1008.f * xv[R + 2] +
                       8064.f * xv[R + 1] - 14350.f * c + 8064.f * xv[R - 1] -
                       1008.f * xv[R - 2] + 128.f * xv[R - 3] - 9.f * xv[R - 4]) +
               ycfl * (-9.f * yv[R + 4] + 128.f * yv[R + 3] - 1008.f * yv[R + 2] +
                       8064.f * yv[R + 1] - 14350.f * c + 8064.f * yv[R - 1] -
                       1008.f * yv[R - 2] + 128.f * yv[R - 3] - 9.f * yv[R - 4]);
    }
}

template <int ORDER>
__device__ __forceinline__ float point_from_global(const float* __restrict__ c, int gx,
                                                   float xcfl, float ycfl) {
    constexpr int R = ORDER / 2;
    float xv[ORDER + 1], yv[ORDER + 1];
#pragma unroll
    for (int k = -R; k <= R; ++k) {
        xv[k + R] = c[k];
        yv[k + R] = c[(ptrdiff_t)k * gx];
    }
    return combine<ORDER>(xv, yv, xcfl, ycfl);
}

// Border decay for the cells of rows [row_lo,row_hi) (grid rows, inclusive of
// the physical border rows when the range touches them): work item t maps to
// one border cell.  Used both by the standalone BC kernel and as extra blocks.
__device__ __forceinline__ void bc_item(float* __restrict__ next, const float* __restrict__ curr,
                                        int gx, int gy, int b, float s, int64_t t) {
    const int ny = gy - 2 * b;
    const int64_t n_x = (int64_t)gx * b;  // one full-width band
    const int64_t n_y = (int64_t)ny * b;  // one side band
    int64_t idx;
    if (t < n_x) {
        idx = t;
    } else if (t < 2 * n_x) {
        idx = t - n_x + (int64_t)gx * (gy - b);
    } else if (t < 2 * n_x + n_y) {
        const int64_t u = t - 2 * n_x;
        idx = (u % b) + (int64_t)gx * (b + u / b);
    } else if (t < 2 * (n_x + n_y)) {
        const int64_t u = t - 2 * n_x - n_y;
        idx = (u % b) + (gx - b) + (int64_t)gx * (b + u / b);
    } else {
        return;
    }
    next[idx] = curr[idx] * s;
}

__global__ void __launch_bounds__(256) bc_kernel(float* __restrict__ next,
                                                 const float* __restrict__ curr, int gx, int gy,
                                                 int b, float s) {
    const int64_t total = 2 * ((int64_t)gx * b + (int64_t)(gy - 2 * b) * b);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride)
        bc_item(next, curr, gx, gy, b, s, t);
}

// ---- variant 0: global memory, one point per thread -------------------------
template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_global(float* __restrict__ next, const float* __restrict__ curr, int gx, int gy, float xcfl,
             float ycfl, int tiles_x, int tiles, int bc, float s) {
    constexpr int R = ORDER / 2;
    const int nx = gx - 2 * R, ny = gy - 2 * R;
    if ((int)blockIdx.x >= tiles) {  // appended border blocks
        if (!bc) return;
        const int64_t total = 2 * ((int64_t)gx * R + (int64_t)ny * R);
        const int64_t stride = (int64_t)(gridDim.x - tiles) * blockDim.x;
        for (int64_t t = (int64_t)(blockIdx.x - tiles) * blockDim.x + threadIdx.x; t < total;
             t += stride)
            bc_item(next, curr, gx, gy, R, s, t);
        return;
    }
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const int x = tx * 32 + (threadIdx.x & 31);
    const int y = ty * 8 + (threadIdx.x >> 5);
    if (x < nx && y < ny) {
        const size_t i = (size_t)gx * (y + R) + (x + R);
        next[i] = point_from_global<ORDER>(curr + i, gx, xcfl, ycfl);
    }
}

// ---- variant 1: several rows per thread -------------------------------------
template <int ORDER, int ROWS>
__global__ void __launch_bounds__(256)
sweep_loop(float* __restrict__ next, const float* __restrict__ curr, int gx, int gy, float xcfl,
           float ycfl, int tiles_x, int tiles, int bc, float s) {
    constexpr int R = ORDER / 2;
    const int nx = gx - 2 * R, ny = gy - 2 * R;
    if ((int)blockIdx.x >= tiles) {
        if (!bc) return;
        const int64_t total = 2 * ((int64_t)gx * R + (int64_t)ny * R);
        const int64_t stride = (int64_t)(gridDim.x - tiles) * blockDim.x;
        for (int64_t t = (int64_t)(blockIdx.x - tiles) * blockDim.x + threadIdx.x; t < total;
             t += stride)
            bc_item(next, curr, gx, gy, R, s, t);
        return;
    }
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    const int x = tx * 32 + (threadIdx.x & 31);
    const int y0 = ty * (8 * ROWS) + (threadIdx.x >> 5);
    if (x >= nx) return;
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
        const int y = y0 + r * 8;
        if (y < ny) {
            const size_t i = (size_t)gx * (y + R) + (x + R);
            next[i] = point_from_global<ORDER>(curr + i, gx, xcfl, ycfl);
        }
    }
}

// ---- variant 2: shared-memory tile, four columns per thread, register marching -----------
// A block of 256 threads owns a tile of kTX x kTY interior points.  The tile and its halo are
// staged in shared memory once (128-bit loads, one barrier per tile).  A thread owns four
// adjacent columns and kRY consecutive rows: it marches down with the 2R+1 vertical neighbours
// of its four columns in registers (one 128-bit shared load per new row) and takes the x
// neighbours of the centre row with three more (order 8).  That is 5 shared-memory wavefronts
// per 32 points where one column per thread with scalar loads needed 18: the first version of
// this kernel was bound by shared-memory load issue (3.3 TB/s of grid traffic, 51 % of the
// HBM copy rate).  If push_* pointers are set, the rows listed as edge rows are also stored to
// a neighbour GPU.
constexpr int kTX = 128, kTY = 64, kRY = 8;   // 32 x 8 threads, 4 columns x 8 rows per thread
// (48-row tiles with three blocks per SM: 38.9 instead of 40.4 ms at 32768^2, but 3.13 instead of 2.42 ms at
// 4096^2, where 2752 tiles on 444 blocks leave a ragged last round; measured, not kept)

template <int R>
constexpr int tile_pitch() { return (kTX + 2 * R + 3) / 4 * 4; }   // rows stay 16-byte aligned

// four floats starting at column c (any alignment) of a 16-byte aligned shared row
template <int R>
__device__ __forceinline__ float4 lds4_at_r(const float* __restrict__ row, int c4) {
    // c4 is a multiple of 4; the four values start at row[c4 + R]
    if constexpr (R % 4 == 0) {
        return *reinterpret_cast<const float4*>(row + c4 + R);
    } else if constexpr (R % 2 == 0) {
        const float2 lo = *reinterpret_cast<const float2*>(row + c4 + R);
        const float2 hi = *reinterpret_cast<const float2*>(row + c4 + R + 2);
        return make_float4(lo.x, lo.y, hi.x, hi.y);
    } else {
        const float4 a = *reinterpret_cast<const float4*>(row + c4);
        const float4 b4 = *reinterpret_cast<const float4*>(row + c4 + 4);
        const float t[8] = {a.x, a.y, a.z, a.w, b4.x, b4.y, b4.z, b4.w};
        return make_float4(t[R % 4], t[R % 4 + 1], t[R % 4 + 2], t[R % 4 + 3]);
    }
}

__device__ __forceinline__ float f4_get(const float4& v, int j) {
    return j == 0 ? v.x : j == 1 ? v.y : j == 2 ? v.z : v.w;
}

// four results of one row to `dst` (points x .. x+3, the first `nvalid` exist)
__device__ __forceinline__ void store_quad(float* __restrict__ dst, const float (&v)[4], int nvalid,
                                           bool vec_ok) {
    if (vec_ok && nvalid == 4) {
        *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (j < nvalid) dst[j] = v[j];
    }
}

// global -> shared staging of one tile with its halo (the non-TMA path)
template <int ORDER>
__device__ __forceinline__ void tile_fill(const float* __restrict__ curr, int gx, int gy_local, int x0,
                                          int y0, float* smem) {
    constexpr int R = ORDER / 2;
    constexpr int W = tile_pitch<R>(), HT = kTY + 2 * R;
    const int tid = threadIdx.x;
    // interior point (x, y) lives at grid (x + R, y + R): the tile with its halo starts at
    // grid column x0 and grid row y0
    const bool rows_aligned = (gx % 4 == 0) && ((reinterpret_cast<uintptr_t>(curr) & 15) == 0);
    if (rows_aligned) {
        constexpr int W4 = W / 4;
        for (int i = tid; i < W4 * HT; i += 256) {
            const int r = i / W4, c = (i - r * W4) * 4;
            const int gr = y0 + r, gc = x0 + c;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gr < gy_local) {
                const float* src = curr + (size_t)gr * gx + gc;
                if (gc + 3 < gx) {
                    v = __ldg(reinterpret_cast<const float4*>(src));
                } else {
                    if (gc < gx) v.x = src[0];
                    if (gc + 1 < gx) v.y = src[1];
                    if (gc + 2 < gx) v.z = src[2];
                }
            }
            *reinterpret_cast<float4*>(smem + r * W + c) = v;
        }
    } else {
        for (int i = tid; i < W * HT; i += 256) {
            const int r = i / W, c = i - r * W;
            const int gr = y0 + r, gc = x0 + c;
            smem[i] = (gr < gy_local && gc < gx) ? curr[(size_t)gr * gx + gc] : 0.f;
        }
    }
}

// the marching phase: every thread computes its 4 x kRY points from the staged tile
template <int ORDER>
__device__ __forceinline__ void tile_march(float* __restrict__ next, int gx, int x0, int y0, int nx,
                                           int ny_local, float xcfl, float ycfl,
                                           float* __restrict__ push_up, int push_up_rows,
                                           float* __restrict__ push_down, int push_down_from,
                                           const float* smem) {
    constexpr int R = ORDER / 2;
    constexpr int W = tile_pitch<R>();
    constexpr int XW = (4 + 2 * R + 3) / 4;          // float4s covering the x window of a quad
    const int tid = threadIdx.x;
    const int c4 = (tid & 31) * 4, tyg = tid >> 5;
    const int x = x0 + c4;
    if (x >= nx) return;
    const int yb = tyg * kRY;                        // first row of this thread inside the tile
    // 16-byte stores need the output row start and R to be multiples of 4 floats
    const bool st_vec = (R % 4 == 0) && (gx % 4 == 0) && ((reinterpret_cast<uintptr_t>(next) & 15) == 0);
    float4 win[2 * R + 1];                           // vertical neighbours of the four columns
#pragma unroll
    for (int k = 0; k < 2 * R; ++k) win[k + 1] = lds4_at_r<R>(smem + (yb + k) * W, c4);
    const float* srow = smem + (yb + R) * W + c4;    // centre row of the first output
    const float* vrow = smem + (yb + 2 * R) * W;     // row entering the vertical window
    float* outp = next + (size_t)gx * (y0 + yb + R) + (x + R);

    // one output row of four points from the registers + the centre row's x window
    auto quad = [&](float (&out)[4]) {
        float xw[4 * XW];
#pragma unroll
        for (int q = 0; q < XW; ++q) {
            const float4 t = *reinterpret_cast<const float4*>(srow + 4 * q);
            xw[4 * q] = t.x, xw[4 * q + 1] = t.y, xw[4 * q + 2] = t.z, xw[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float xv[2 * R + 1], yv[2 * R + 1];
#pragma unroll
            for (int k = 0; k <= 2 * R; ++k) {
                xv[k] = xw[j + k];
                yv[k] = f4_get(win[k], j);
            }
            out[j] = combine<ORDER>(xv, yv, xcfl, ycfl);
        }
    };

    // Tiles that lie fully inside the grid and push nothing take a loop without any bounds
    // test or address arithmetic beyond three pointer increments per row.
    const bool pushes = (push_up && y0 < push_up_rows) || (push_down && y0 + kTY > push_down_from);
    if (st_vec && !pushes && x0 + kTX <= nx && y0 + kTY <= ny_local) {
#pragma unroll
        for (int i = 0; i < kRY; ++i) {
#pragma unroll
            for (int k = 0; k < 2 * R; ++k) win[k] = win[k + 1];
            win[2 * R] = lds4_at_r<R>(vrow, c4);
            float out[4];
            quad(out);
            *reinterpret_cast<float4*>(outp) = make_float4(out[0], out[1], out[2], out[3]);
            srow += W, vrow += W, outp += gx;
        }
        return;
    }
    const int nvalid = min(4, nx - x);
    const bool up_vec = st_vec && ((reinterpret_cast<uintptr_t>(push_up) & 15) == 0);
    const bool dn_vec = st_vec && ((reinterpret_cast<uintptr_t>(push_down) & 15) == 0);
#pragma unroll
    for (int i = 0; i < kRY; ++i) {
#pragma unroll
        for (int k = 0; k < 2 * R; ++k) win[k] = win[k + 1];
        win[2 * R] = lds4_at_r<R>(vrow, c4);
        const int y = y0 + yb + i;
        if (y < ny_local) {
            float out[4];
            quad(out);
            store_quad(outp, out, nvalid, st_vec);
            if (push_up && y < push_up_rows)
                store_quad(push_up + (size_t)gx * y + (x + R), out, nvalid, up_vec);
            if (push_down && y >= push_down_from)
                store_quad(push_down + (size_t)gx * (y - push_down_from) + (x + R), out, nvalid, dn_vec);
        }
        srow += W, vrow += W, outp += gx;
    }
}

template <int ORDER>
__device__ __forceinline__ void tile_compute(float* __restrict__ next,
                                             const float* __restrict__ curr, int gx, int gy_local,
                                             int x0, int y0, int nx, int ny_local, float xcfl,
                                             float ycfl, float* __restrict__ push_up,
                                             int push_up_rows, float* __restrict__ push_down,
                                             int push_down_from, float* smem) {
    tile_fill<ORDER>(curr, gx, gy_local, x0, y0, smem);
    __syncthreads();
    tile_march<ORDER>(next, gx, x0, y0, nx, ny_local, xcfl, ycfl, push_up, push_up_rows, push_down,
                      push_down_from, smem);
}

template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_shared(float* __restrict__ next, const float* __restrict__ curr, int gx, int gy, float xcfl,
             float ycfl, int tiles_x, int tiles, int bc, float s) {
    constexpr int R = ORDER / 2;
    __shared__ __align__(16) float smem[tile_pitch<R>() * (kTY + 2 * R)];
    const int nx = gx - 2 * R, ny = gy - 2 * R;
    if ((int)blockIdx.x >= tiles) {
        if (!bc) return;
        const int64_t total = 2 * ((int64_t)gx * R + (int64_t)ny * R);
        const int64_t stride = (int64_t)(gridDim.x - tiles) * blockDim.x;
        for (int64_t t = (int64_t)(blockIdx.x - tiles) * blockDim.x + threadIdx.x; t < total;
             t += stride)
            bc_item(next, curr, gx, gy, R, s, t);
        return;
    }
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    tile_compute<ORDER>(next, curr, gx, gy, tx * kTX, ty * kTY, nx, ny, xcfl, ycfl, nullptr, 0,
                        nullptr, 0, smem);
}

// Border decay of one slab (multi-GPU): side columns of every owned row; the full-width bands
// only where this slab touches the physical top / bottom border.  Work item t of `total`.
__device__ __forceinline__ int64_t slab_bc_total(int gx, int rows, int R, int is_top, int is_bottom) {
    return (int64_t)rows * R * 2 + (is_top ? (int64_t)gx * R : 0) + (is_bottom ? (int64_t)gx * R : 0);
}
__device__ __forceinline__ void slab_bc_item(float* __restrict__ next, const float* __restrict__ curr,
                                             int gx, int rows, int R, int is_top, float s, int64_t t) {
    const int64_t n_side = (int64_t)rows * R * 2;
    const int64_t n_band = (int64_t)gx * R;
    int64_t idx;
    if (t < n_side) {
        const int64_t row = t / (2 * R);
        const int k = (int)(t % (2 * R));
        const int colg = k < R ? k : gx - 2 * R + k;
        idx = (int64_t)gx * (row + R) + colg;
    } else if (is_top && t < n_side + n_band) {
        idx = t - n_side;
    } else {
        const int64_t u = t - n_side - (is_top ? n_band : 0);
        idx = u + (int64_t)gx * (rows + R);
    }
    next[idx] = curr[idx] * s;
}

struct SlabArgs {
    float* up_ghost;      // upper neighbour's bottom ghost rows (or null)
    float* down_ghost;    // lower neighbour's top ghost rows (or null)
    int slab, is_top, is_bottom;
};

// ---- variant 2 on TMA: persistent blocks, tiles prefetched by the copy engine -------------
// ncu on sweep_shared: 60 % of the warp stalls are the global loads of the staging loop, with
// DRAM at 46 % and the issue slots at 46 %: four resident blocks per SM do not hide it.  Here
// one cp.async.bulk.tensor per tile brings the (kTX + 2R) x (kTY + 2R) box, halo and
// out-of-range zeros included, into one of two shared buffers while the block computes the
// other; nothing in the block waits on a global load.  Needs 16-byte aligned rows (gx % 4 == 0).
template <int R>
constexpr int tma_smem_bytes() { return 2 * tile_pitch<R>() * (kTY + 2 * R) * 4 + 256; }

template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_tma(const __grid_constant__ CUtensorMap tm, float* __restrict__ next,
          const float* __restrict__ curr, int gx, int gy, float xcfl, float ycfl, int tiles_x,
          int tiles, int bc, float s, SlabArgs sa, int strips) {
    constexpr int R = ORDER / 2;
    constexpr int W = tile_pitch<R>(), HT = kTY + 2 * R;
    constexpr uint32_t kTileBytes = W * HT * 4;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full = reinterpret_cast<uint64_t*>(base + 2 * kTileBytes);
    const int nx = gx - 2 * R, ny = gy - 2 * R;
    const int tid = threadIdx.x;
    if (tid == 0) {
        tc::prefetch_tmap(&tm);
        tc::mbar_init(&full[0], 1);
        tc::mbar_init(&full[1], 1);
        tc::fence_barrier_init();
    }
    __syncthreads();
    // Tile order.  Default: tiles numbered along the rows, dealt round-robin, so blocks that run
    // together read neighbouring 512-byte pieces of the same grid rows.  strips != 0
    // (CME_STENCIL_STRIPS=1): a block walks down a column strip, which keeps the halo rows it shares
    // with the tile below in L2 -- measured SLOWER (32768^2, order 8, 20 sweeps: 53.3 vs 40.4 ms):
    // 296 blocks each streaming its own 544-byte-wide column scatter the DRAM accesses over pages.
    const int tiles_y = tiles / tiles_x;
    const int per = strips ? (tiles + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const int t_begin = strips ? (int)blockIdx.x * per : (int)blockIdx.x;
    const int t_end = strips ? min(tiles, t_begin + per) : tiles;
    const int t_step = strips ? 1 : (int)gridDim.x;
    auto tile_xy = [&](int t, int& x0, int& y0) {
        if (strips) { x0 = (t / tiles_y) * kTX; y0 = (t % tiles_y) * kTY; }
        else { x0 = (t % tiles_x) * kTX; y0 = (t / tiles_x) * kTY; }
    };
    int t = t_begin;
    uint32_t b = 0, phases = 0;                      // current buffer, one parity bit per buffer
    if (t < t_end && tid == 0) {
        int x0, y0;
        tile_xy(t, x0, y0);
        tc::mbar_expect_tx(&full[0], kTileBytes);
        tc::tma_load_2d(base, &tm, &full[0], x0, y0);
    }
    for (; t < t_end; t += t_step) {
        const int tn = t + t_step;
        if (tn < t_end && tid == 0) {
            // the other buffer was last read before the barrier that ended the previous iteration
            int x0, y0;
            tile_xy(tn, x0, y0);
            tc::mbar_expect_tx(&full[b ^ 1], kTileBytes);
            tc::tma_load_2d(base + (b ^ 1) * kTileBytes, &tm, &full[b ^ 1], x0, y0);
        }
        tc::mbar_wait(&full[b], (phases >> b) & 1u);
        phases ^= 1u << b;
        // slab mode: my first R rows also go to the upper neighbour's bottom ghost rows, my last R
        // rows to the lower neighbour's top ghost rows
        int x0, y0;
        tile_xy(t, x0, y0);
        tile_march<ORDER>(next, gx, x0, y0, nx, ny, xcfl, ycfl, sa.up_ghost, R,
                          sa.down_ghost, ny - R, reinterpret_cast<const float*>(base + b * kTileBytes));
        __syncthreads();
        b ^= 1;
    }
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (sa.slab) {
        const int64_t total = slab_bc_total(gx, ny, R, sa.is_top, sa.is_bottom);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + tid; i < total; i += stride)
            slab_bc_item(next, curr, gx, ny, R, sa.is_top, s, i);
    } else if (bc) {
        const int64_t total = 2 * ((int64_t)gx * R + (int64_t)ny * R);
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + tid; i < total; i += stride)
            bc_item(next, curr, gx, gy, R, s, i);
    }
}

// ---- multi-GPU slab: sweep + BC + halo push in one launch ---------------------
template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_slab(float* __restrict__ next, const float* __restrict__ curr,
           float* __restrict__ up_ghost, float* __restrict__ down_ghost, int gx, int rows,
           float xcfl, float ycfl, float s, int is_top, int is_bottom, int tiles_x, int tiles) {
    constexpr int R = ORDER / 2;
    __shared__ __align__(16) float smem[tile_pitch<R>() * (kTY + 2 * R)];
    const int nx = gx - 2 * R;
    const int gy = rows + 2 * R;
    if ((int)blockIdx.x >= tiles) {
        const int64_t total = slab_bc_total(gx, rows, R, is_top, is_bottom);
        const int64_t stride = (int64_t)(gridDim.x - tiles) * blockDim.x;
        for (int64_t t = (int64_t)(blockIdx.x - tiles) * blockDim.x + threadIdx.x; t < total; t += stride)
            slab_bc_item(next, curr, gx, rows, R, is_top, s, t);
        return;
    }
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    // my first R rows go to the upper neighbour's bottom ghost rows, my last R
    // rows to the lower neighbour's top ghost rows
    tile_compute<ORDER>(next, curr, gx, gy, tx * kTX, ty * kTY, nx, rows, xcfl, ycfl, up_ghost, R,
                        down_ghost, rows - R, smem);
}

// fp32 2-D tensor map over a grid (gx x gy, rows contiguous) with a box of one tile + halo
int make_tile_map(CUtensorMap* out, const float* grid, int gx, int gy, int box_w, int box_h) {
    typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                      CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                      CUtensorMapFloatOOBfill);
    EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(cme::tensor_map_encoder());
    if (!fn) return -2;
    cuuint64_t dims[2] = {(cuuint64_t)gx, (cuuint64_t)gy};
    cuuint64_t strides[1] = {(cuuint64_t)gx * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h};
    cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(grid), dims, strides, box,
                          estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -3;
}

// the TMA kernel needs more dynamic shared memory than the default limit (set per device)
template <int ORDER>
int opt_in_tma() {
    static bool done[64] = {};
    int dev = 0;
    CME_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !done[dev]) {
        CME_CUDA(cudaFuncSetAttribute(sweep_tma<ORDER>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      tma_smem_bytes<ORDER / 2>()));
        if (dev >= 0 && dev < 64) done[dev] = true;
    }
    return 0;
}

template <int ORDER>
int launch_step(float* next, const float* curr, int gx, int gy, float xcfl, float ycfl,
                int variant, int bc, float s, cudaStream_t st) {
    constexpr int R = ORDER / 2;
    const int nx = gx - 2 * R, ny = gy - 2 * R;
    const int bc_blocks = bc ? 2 * cme::sm_count() : 0;
    if (variant == 0) {
        const int tiles_x = (nx + 31) / 32, tiles_y = (ny + 7) / 8;
        const int64_t tiles = (int64_t)tiles_x * tiles_y;
        CME_CHECK_ARG(tiles + bc_blocks < INT32_MAX, "cme_heat_step: grid too large");
        sweep_global<ORDER><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
            next, curr, gx, gy, xcfl, ycfl, tiles_x, (int)tiles, bc, s);
    } else if (variant == 1) {
        constexpr int ROWS = 4;
        const int tiles_x = (nx + 31) / 32, tiles_y = (ny + 8 * ROWS - 1) / (8 * ROWS);
        const int64_t tiles = (int64_t)tiles_x * tiles_y;
        CME_CHECK_ARG(tiles + bc_blocks < INT32_MAX, "cme_heat_step: grid too large");
        sweep_loop<ORDER, ROWS><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
            next, curr, gx, gy, xcfl, ycfl, tiles_x, (int)tiles, bc, s);
    } else {
        const int tiles_x = (nx + kTX - 1) / kTX, tiles_y = (ny + kTY - 1) / kTY;
        const int64_t tiles = (int64_t)tiles_x * tiles_y;
        CME_CHECK_ARG(tiles + bc_blocks < INT32_MAX, "cme_heat_step: grid too large");
        static const bool tma_off = getenv("CME_STENCIL_NO_TMA") != nullptr;
        static const int strips = getenv("CME_STENCIL_STRIPS") ? 1 : 0;
        CUtensorMap tm;
        if (!tma_off && gx % 4 == 0 && ((uintptr_t)curr & 15) == 0 &&
            make_tile_map(&tm, curr, gx, gy, tile_pitch<R>(), kTY + 2 * R) == 0) {
            constexpr int kBytes = tma_smem_bytes<R>();
            if (int rc = opt_in_tma<ORDER>()) return rc;
            const int grid = (int)std::min<int64_t>(tiles, 2 * (int64_t)cme::sm_count());
            sweep_tma<ORDER><<<grid, 256, kBytes, st>>>(tm, next, curr, gx, gy, xcfl, ycfl, tiles_x, (int)tiles,
                                                         bc, s, SlabArgs{nullptr, nullptr, 0, 0, 0}, strips);
        } else {
            sweep_shared<ORDER><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
                next, curr, gx, gy, xcfl, ycfl, tiles_x, (int)tiles, bc, s);
        }
    }
    CME_LAUNCH_CHECK();
    return 0;
}

template <int ORDER>
int launch_slab(float* next, const float* curr, float* up, float* down, int gx, int rows,
                float xcfl, float ycfl, float s, int is_top, int is_bottom, cudaStream_t st) {
    constexpr int R = ORDER / 2;
    const int nx = gx - 2 * R;
    const int tiles_x = (nx + kTX - 1) / kTX, tiles_y = (rows + kTY - 1) / kTY;
    const int64_t tiles = (int64_t)tiles_x * tiles_y;
    const int bc_blocks = 2 * cme::sm_count();
    CME_CHECK_ARG(tiles + bc_blocks < INT32_MAX, "cme_heat_slab_step: grid too large");
    static const bool tma_off = getenv("CME_STENCIL_NO_TMA") != nullptr;
    static const int strips = getenv("CME_STENCIL_STRIPS") ? 1 : 0;
    CUtensorMap tm;
    if (!tma_off && gx % 4 == 0 && ((uintptr_t)curr & 15) == 0 &&
        make_tile_map(&tm, curr, gx, rows + 2 * R, tile_pitch<R>(), kTY + 2 * R) == 0) {
        if (int rc = opt_in_tma<ORDER>()) return rc;
        const int grid = (int)std::min<int64_t>(tiles, 2 * (int64_t)cme::sm_count());
        sweep_tma<ORDER><<<grid, 256, tma_smem_bytes<R>(), st>>>(tm, next, curr, gx, rows + 2 * R, xcfl, ycfl,
                                                                 tiles_x, (int)tiles, 1, s,
                                                                 SlabArgs{up, down, 1, is_top, is_bottom}, strips);
    } else {
        sweep_slab<ORDER><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
            next, curr, up, down, gx, rows, xcfl, ycfl, s, is_top, is_bottom, tiles_x, (int)tiles);
    }
    CME_LAUNCH_CHECK();
    return 0;
}

}  // namespace

extern "C" {

int cme_heat_step(float* next, const float* curr, int gx, int gy, int order, float xcfl,
                  float ycfl, int variant, int bc, float bc_scale, void* stream) {
    CME_CHECK_ARG(next && curr && next != curr, "cme_heat_step: bad grid pointers");
    CME_CHECK_ARG(order == 2 || order == 4 || order == 8, "cme_heat_step: order %d", order);
    CME_CHECK_ARG(variant >= 0 && variant <= 2, "cme_heat_step: variant %d", variant);
    CME_CHECK_ARG(gx > order && gy > order, "cme_heat_step: grid smaller than the border");
    cudaStream_t st = cme::as_stream(stream);
    if (order == 2) return launch_step<2>(next, curr, gx, gy, xcfl, ycfl, variant, bc, bc_scale, st);
    if (order == 4) return launch_step<4>(next, curr, gx, gy, xcfl, ycfl, variant, bc, bc_scale, st);
    return launch_step<8>(next, curr, gx, gy, xcfl, ycfl, variant, bc, bc_scale, st);
}

int cme_heat_bc(float* next, const float* curr, int gx, int gy, int order, float bc_scale,
                void* stream) {
    CME_CHECK_ARG(next && curr, "cme_heat_bc: null grid");
    CME_CHECK_ARG(order == 2 || order == 4 || order == 8, "cme_heat_bc: order %d", order);
    const int b = order / 2;
    CME_CHECK_ARG(gx > 2 * b && gy > 2 * b, "cme_heat_bc: grid smaller than the border");
    const int64_t total = 2 * ((int64_t)gx * b + (int64_t)(gy - 2 * b) * b);
    bc_kernel<<<cme::stride_grid(total, 256, 4, 2), 256, 0, cme::as_stream(stream)>>>(
        next, curr, gx, gy, b, bc_scale);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_heat_slab_step(float* next, const float* curr, float* up_next_ghost,
                       float* down_next_ghost, int gx, int rows_local, int order, float xcfl,
                       float ycfl, float bc_scale, int is_top, int is_bottom, void* stream) {
    CME_CHECK_ARG(next && curr && next != curr, "cme_heat_slab_step: bad grid pointers");
    CME_CHECK_ARG(order == 2 || order == 4 || order == 8, "cme_heat_slab_step: order %d", order);
    CME_CHECK_ARG(rows_local >= order / 2 && gx > order, "cme_heat_slab_step: slab too small");
    cudaStream_t st = cme::as_stream(stream);
    if (order == 2)
        return launch_slab<2>(next, curr, up_next_ghost, down_next_ghost, gx, rows_local, xcfl,
                              ycfl, bc_scale, is_top, is_bottom, st);
    if (order == 4)
        return launch_slab<4>(next, curr, up_next_ghost, down_next_ghost, gx, rows_local, xcfl,
                              ycfl, bc_scale, is_top, is_bottom, st);
    return launch_slab<8>(next, curr, up_next_ghost, down_next_ghost, gx, rows_local, xcfl, ycfl,
                          bc_scale, is_top, is_bottom, st);
}

}  // extern "C"
