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

// ---- variant 2: shared-memory tile + register marching ---------------------------
// A block of 256 threads owns a tile of kTX x kTY interior points.  The tile and its
// halo are staged in shared memory once (one barrier per tile); each thread then owns
// one column of kRY consecutive rows and marches down it with the 2R+1 vertical
// neighbours in registers, taking the x neighbours from the shared row.  Per output:
// one new shared load for the window plus 2R+1 for the row, no global re-reads.  If
// push_* pointers are set, the rows listed as edge rows are also stored to a
// neighbour GPU.
constexpr int kTX = 64, kTY = 64, kRY = 16;   // 64 x 4 threads, 16 rows per thread

template <int ORDER>
__device__ __forceinline__ void tile_compute(float* __restrict__ next,
                                             const float* __restrict__ curr, int gx, int gy_local,
                                             int x0, int y0, int nx, int ny_local, float xcfl,
                                             float ycfl, float* __restrict__ push_up,
                                             int push_up_rows, float* __restrict__ push_down,
                                             int push_down_from, float* smem) {
    constexpr int R = ORDER / 2;
    constexpr int W = kTX + 2 * R, HT = kTY + 2 * R;
    const int tid = threadIdx.x;
    // interior point (x, y) lives at grid (x + R, y + R): the tile with its halo starts at
    // grid column x0 and grid row y0
    for (int i = tid; i < W * HT; i += 256) {
        const int r = i / W, c = i - r * W;
        const int gr = y0 + r, gc = x0 + c;
        smem[i] = (gr < gy_local && gc < gx) ? curr[(size_t)gr * gx + gc] : 0.f;
    }
    __syncthreads();
    const int tx = tid & (kTX - 1), tyg = tid / kTX;
    const int x = x0 + tx;
    if (x >= nx) return;
    const int yb = tyg * kRY;                       // first row of this thread inside the tile
    const float* colp = smem + (size_t)yb * W + tx + R;
    float win[2 * R + 1];
#pragma unroll
    for (int k = 0; k < 2 * R; ++k) win[k + 1] = colp[k * W];
#pragma unroll
    for (int i = 0; i < kRY; ++i) {
#pragma unroll
        for (int k = 0; k < 2 * R; ++k) win[k] = win[k + 1];
        win[2 * R] = colp[(i + 2 * R) * W];
        const int y = y0 + yb + i;
        if (y < ny_local) {
            const float* rowp = smem + (size_t)(yb + i + R) * W + tx;
            float xv[2 * R + 1];
#pragma unroll
            for (int k = 0; k <= 2 * R; ++k) xv[k] = rowp[k];
            const float v = combine<ORDER>(xv, win, xcfl, ycfl);
            next[(size_t)gx * (y + R) + (x + R)] = v;
            if (push_up && y < push_up_rows) push_up[(size_t)gx * y + (x + R)] = v;
            if (push_down && y >= push_down_from)
                push_down[(size_t)gx * (y - push_down_from) + (x + R)] = v;
        }
    }
}

template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_shared(float* __restrict__ next, const float* __restrict__ curr, int gx, int gy, float xcfl,
             float ycfl, int tiles_x, int tiles, int bc, float s) {
    constexpr int R = ORDER / 2;
    __shared__ float smem[(kTX + 2 * R) * (kTY + 2 * R)];
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

// ---- multi-GPU slab: sweep + BC + halo push in one launch ---------------------
template <int ORDER>
__global__ void __launch_bounds__(256)
sweep_slab(float* __restrict__ next, const float* __restrict__ curr,
           float* __restrict__ up_ghost, float* __restrict__ down_ghost, int gx, int rows,
           float xcfl, float ycfl, float s, int is_top, int is_bottom, int tiles_x, int tiles) {
    constexpr int R = ORDER / 2;
    __shared__ float smem[(kTX + 2 * R) * (kTY + 2 * R)];
    const int nx = gx - 2 * R;
    const int gy = rows + 2 * R;
    if ((int)blockIdx.x >= tiles) {
        // border decay: side columns of every owned row; the full-width bands
        // only where this slab touches the physical top / bottom border
        const int64_t n_side = (int64_t)rows * R * 2;
        const int64_t n_band = (int64_t)gx * R;
        const int64_t total = n_side + (is_top ? n_band : 0) + (is_bottom ? n_band : 0);
        const int64_t stride = (int64_t)(gridDim.x - tiles) * blockDim.x;
        for (int64_t t = (int64_t)(blockIdx.x - tiles) * blockDim.x + threadIdx.x; t < total;
             t += stride) {
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
                idx = u + (int64_t)gx * (gy - R);
            }
            next[idx] = curr[idx] * s;
        }
        return;
    }
    const int tx = blockIdx.x % tiles_x, ty = blockIdx.x / tiles_x;
    // my first R rows go to the upper neighbour's bottom ghost rows, my last R
    // rows to the lower neighbour's top ghost rows
    tile_compute<ORDER>(next, curr, gx, gy, tx * kTX, ty * kTY, nx, rows, xcfl, ycfl, up_ghost, R,
                        down_ghost, rows - R, smem);
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
        sweep_shared<ORDER><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
            next, curr, gx, gy, xcfl, ycfl, tiles_x, (int)tiles, bc, s);
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
    sweep_slab<ORDER><<<(unsigned)(tiles + bc_blocks), 256, 0, st>>>(
        next, curr, up, down, gx, rows, xcfl, ycfl, s, is_top, is_bottom, tiles_x, (int)tiles);
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
