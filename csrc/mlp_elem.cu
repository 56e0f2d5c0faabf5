// bf16 MLP path: everything around the two tensor-core GEMMs.
//
//   u8_to_bf16_rows   minibatch staging: uint8 pixels -> bf16 rows, with a ones
//                     column appended (the bias rides in the GEMMs) and the pad
//                     columns zeroed.  The source may be another GPU's memory:
//                     each rank pulls its own slice of the minibatch straight
//                     from the owner over NVLink, which replaces the host-side
//                     MPI_Scatter (ref fpcode/neural_network.cpp:460-481).
//   fwd2_softmax      z2 = A1*W2^T + b2, softmax, cross-entropy, (p - y)*scale,
//                     argmax and the b2 gradient in ONE pass over A1 (replaces
//                     ref K1(fwd2) + K8 + K5 + D2D copy + K9 + K7(db2),
//                     ref neural_network.cpp:309-323, 348, 361).
//   bwd_elem          dZ1 = (dz2*W2) .* a .* (1-a) and gW2 += dz2^T * A1 in ONE
//                     pass over A1 (replaces ref K2 + K10 + K3(dW2),
//                     ref neural_network.cpp:357-378).
// Softmax subtracts the row max (the reference does not, ref gpu_func.cu:307-314).
#include "common.cuh"

namespace {

constexpr int kMaxClasses = 16;

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

// one thread per hidden unit: column h of W2 -> bf16 row entries of w2c and row h of w2t
__device__ __forceinline__ void recast_w2_column(const float* __restrict__ W2, __nv_bfloat16* __restrict__ w2c, int ldh,
                                                 __nv_bfloat16* __restrict__ w2t, int H, int C, int h) {
    if (h >= H) return;
    float v[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) v[c] = (c < C) ? W2[(size_t)c * H + h] : 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c)
        if (c < C) w2c[(size_t)c * ldh + h] = __float2bfloat16_rn(v[c]);
    if (w2t) {
        uint4* d = reinterpret_cast<uint4*>(w2t + (size_t)h * 64);
        d[0] = make_uint4(pack2(v[0], v[1]), pack2(v[2], v[3]), pack2(v[4], v[5]), pack2(v[6], v[7]));
        d[1] = make_uint4(pack2(v[8], v[9]), pack2(v[10], v[11]), pack2(v[12], v[13]), pack2(v[14], v[15]));
    }
}

// The bf16 copies of W2 ride in the staging launch of the NEXT step (blocks past `first_block`):
// the update kernel that wrote W2 has finished by then, and a launch of its own cost ~4 us, a
// tenth of a step of the 784-100-10 model.
struct RecastArgs {
    const float* W2;
    __nv_bfloat16* w2c;
    __nv_bfloat16* w2t;
    int ldh, H, C;
    int first_block;      // 0 = no recast in this launch
};

// ---- staging -------------------------------------------------------------------
// one thread per 16-element chunk of a destination row
__global__ void __launch_bounds__(256)
u8_to_bf16_kernel(const uint8_t* __restrict__ src_arg, const CmeStepArgs* __restrict__ sa,
                  __nv_bfloat16* __restrict__ dst, int64_t rows, int cols, int ld, int add_ones, RecastArgs rc) {
    if (rc.first_block && (int)blockIdx.x >= rc.first_block) {
        recast_w2_column(rc.W2, rc.w2c, rc.ldh, rc.w2t, rc.H, rc.C,
                         ((int)blockIdx.x - rc.first_block) * blockDim.x + threadIdx.x);
        return;
    }
    const uint8_t* __restrict__ src = sa ? sa->x_u8 : src_arg;      // indirect step arguments
    const int chunks = ld / 16;
    const int64_t total = rows * chunks;
    const int64_t gstride = (int64_t)(rc.first_block ? rc.first_block : gridDim.x) * blockDim.x;
    const bool vec_ok = (cols % 16 == 0) && (((uintptr_t)src & 15) == 0);
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += gstride) {
        const int64_t r = t / chunks;
        const int c0 = (int)(t % chunks) * 16;
        uint32_t out[8];
        if (vec_ok && c0 + 16 <= cols) {
            const uint4 v = cme::ld_stream_u4(reinterpret_cast<const uint4*>(src + r * cols + c0));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                out[2 * k] = pack2((float)(w[k] & 0xff), (float)((w[k] >> 8) & 0xff));
                out[2 * k + 1] = pack2((float)((w[k] >> 16) & 0xff), (float)(w[k] >> 24));
            }
        } else {
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const int c = c0 + k;
                f[k] = c < cols ? (float)src[r * cols + c] : ((add_ones && c == cols) ? 1.f : 0.f);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) out[k] = pack2(f[2 * k], f[2 * k + 1]);
        }
        uint4* d = reinterpret_cast<uint4*>(dst + r * ld + c0);
        d[0] = make_uint4(out[0], out[1], out[2], out[3]);
        d[1] = make_uint4(out[4], out[5], out[6], out[7]);
    }
}

// ---- layer 2 + softmax + CE gradient ----------------------------------------------
// Block = 8 warps x 4 rows.  W2 is staged through shared memory in chunks of
// HC columns; each lane covers columns {4*lane..+3} and {128+4*lane..+3} of every
// 256-column piece so both the A1 loads (8 bytes, coalesced) and the W2 reads
// (LDS.128, stride 16 bytes) are conflict-free.
constexpr int kF2Rows = 4;       // rows per warp
constexpr int kF2HC = 1024;      // W2 columns staged per pass

template <int C>
__global__ void __launch_bounds__(256)
fwd2_softmax_kernel(const __nv_bfloat16* __restrict__ A1, const float* __restrict__ W2,
                    const float* __restrict__ b2, const int32_t* __restrict__ labels,
                    float* __restrict__ dz2, float* __restrict__ loss_sum,
                    int32_t* __restrict__ pred, float* __restrict__ gb2, float scale, int64_t rows,
                    int H, int ldh) {
    __shared__ __align__(16) float w2s[C][kF2HC];
    __shared__ float gb2_s[kMaxClasses];
    __shared__ float loss_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < kMaxClasses) gb2_s[threadIdx.x] = 0.f;
    if (threadIdx.x == 0) loss_s = 0.f;
    const int64_t row0 = ((int64_t)blockIdx.x * 8 + warp) * kF2Rows;

    float acc[kF2Rows][C];
#pragma unroll
    for (int r = 0; r < kF2Rows; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) acc[r][c] = 0.f;

    for (int h0 = 0; h0 < H; h0 += kF2HC) {
        __syncthreads();
        for (int i = threadIdx.x; i < C * kF2HC; i += 256) {
            const int c = i / kF2HC, k = i % kF2HC;
            w2s[c][k] = (h0 + k < H) ? W2[(size_t)c * H + h0 + k] : 0.f;
        }
        __syncthreads();
        const int hc = min(kF2HC, H - h0);
        for (int p0 = 0; p0 < hc; p0 += 256) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int k = p0 + half * 128 + lane * 4;     // column inside the chunk
                if (k >= hc) continue;                         // pad columns carry garbage: skip
                float w[C][4];
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const float4 t = *reinterpret_cast<const float4*>(&w2s[c][k]);
                    w[c][0] = t.x; w[c][1] = t.y; w[c][2] = t.z; w[c][3] = t.w;
                }
#pragma unroll
                for (int r = 0; r < kF2Rows; ++r) {
                    const int64_t row = row0 + r;
                    if (row >= rows) continue;
                    const uint2 av = *reinterpret_cast<const uint2*>(A1 + row * ldh + h0 + k);
                    float a[4] = {bf16_lo(av.x), bf16_hi(av.x), bf16_lo(av.y), bf16_hi(av.y)};
                    // columns past H inside the last 4-group hold pad values: mask them
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (h0 + k + e >= H) a[e] = 0.f;
#pragma unroll
                    for (int c = 0; c < C; ++c)
#pragma unroll
                        for (int e = 0; e < 4; ++e) acc[r][c] = fmaf(a[e], w[c][e], acc[r][c]);
                }
            }
        }
    }
    // reduce the per-lane partial dot products across the warp
#pragma unroll
    for (int r = 0; r < kF2Rows; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) acc[r][c] = cme::warp_sum(acc[r][c]);

    float my_loss = 0.f;
#pragma unroll
    for (int r = 0; r < kF2Rows; ++r) {
        const int64_t row = row0 + r;
        if (lane == r && row < rows) {
            float z[C];
            float mx = -3.0e38f;
            int arg = 0;
#pragma unroll
            for (int c = 0; c < C; ++c) {
                z[c] = acc[r][c] + b2[c];
                if (z[c] > mx) { mx = z[c]; arg = c; }
            }
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < C; ++c) { z[c] = __expf(z[c] - mx); s += z[c]; }
            const float inv = 1.f / s;
            const int lab = labels ? labels[row] : -1;
            float out[kMaxClasses];
#pragma unroll
            for (int c = 0; c < kMaxClasses; ++c) out[c] = 0.f;
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const float p = z[c] * inv;
                out[c] = (p - (c == lab ? 1.f : 0.f)) * scale;
                if (c == lab) my_loss = -__logf(fmaxf(p, 1e-30f));
                atomicAdd(&gb2_s[c], out[c]);
            }
            if (dz2) {
                float4* d = reinterpret_cast<float4*>(dz2 + row * kMaxClasses);
#pragma unroll
                for (int v = 0; v < 4; ++v)
                    d[v] = make_float4(out[4 * v], out[4 * v + 1], out[4 * v + 2], out[4 * v + 3]);
            }
            if (pred) pred[row] = arg;
        }
    }
    if (my_loss != 0.f) atomicAdd(&loss_s, my_loss);
    __syncthreads();
    if (gb2 && threadIdx.x < C) atomicAdd(&gb2[threadIdx.x], gb2_s[threadIdx.x]);
    if (loss_sum && threadIdx.x == 0 && loss_s != 0.f) atomicAdd(loss_sum, loss_s);
}

// ---- backward elementwise + dW2 ------------------------------------------------------
// Thread owns 4 hidden columns and walks a chunk of rows; W2 for its columns and
// the gW2 partial sums live in registers, dz2 rows are broadcast loads.
template <int C, bool WITH_GW2>
__global__ void __launch_bounds__(256)
bwd_elem_kernel(const __nv_bfloat16* __restrict__ A1, const float* __restrict__ dz2,
                const float* __restrict__ W2, __nv_bfloat16* __restrict__ dZ1,
                float* __restrict__ gW2, int64_t rows, int64_t rows_per_block, int H, int ldh) {
    const int h = (blockIdx.x * 256 + threadIdx.x) * 4;
    if (h >= H) return;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_block;
    const int64_t r_end = min(rows, r_begin + rows_per_block);
    float w[C][4], g2[C][4];
#pragma unroll
    for (int c = 0; c < C; ++c)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            w[c][e] = (h + e < H) ? W2[(size_t)c * H + h + e] : 0.f;
            g2[c][e] = 0.f;
        }
    for (int64_t r = r_begin; r < r_end; ++r) {
        const uint2 av = *reinterpret_cast<const uint2*>(A1 + r * ldh + h);
        const float a[4] = {bf16_lo(av.x), bf16_hi(av.x), bf16_lo(av.y), bf16_hi(av.y)};
        float d[kMaxClasses];
        const float4* dp = reinterpret_cast<const float4*>(dz2 + r * kMaxClasses);
#pragma unroll
        for (int v = 0; v < (C + 3) / 4; ++v) {
            const float4 t = __ldg(dp + v);
            d[4 * v] = t.x; d[4 * v + 1] = t.y; d[4 * v + 2] = t.z; d[4 * v + 3] = t.w;
        }
        float dz[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float g = 0.f;
#pragma unroll
            for (int c = 0; c < C; ++c) {
                g = fmaf(d[c], w[c][e], g);
                if (WITH_GW2) g2[c][e] = fmaf(d[c], a[e], g2[c][e]);
            }
            dz[e] = g * a[e] * (1.f - a[e]);
        }
        *reinterpret_cast<uint2*>(dZ1 + r * ldh + h) = make_uint2(pack2(dz[0], dz[1]), pack2(dz[2], dz[3]));
    }
    if (WITH_GW2) {
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (h + e < H) atomicAdd(&gW2[(size_t)c * H + h + e], g2[c][e]);
    }
}

// fp32 [rows][ld_src] -> bf16 [rows][ld_dst] (first `cols` columns; the rest untouched)
__global__ void __launch_bounds__(256)
f32_to_bf16_2d_kernel(const float* __restrict__ src, int ld_src, __nv_bfloat16* __restrict__ dst,
                      int ld_dst, int rows, int cols) {
    const int64_t total = (int64_t)rows * cols;
    const int64_t gstride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gstride) {
        const int r = (int)(i / cols), c = (int)(i % cols);
        dst[(size_t)r * ld_dst + c] = __float2bfloat16_rn(src[(size_t)r * ld_src + c]);
    }
}

// W2 (fp32 [C][H]) -> both bf16 layouts the tensor-core tail wants, one thread per column h:
//   w2c[c][h] (K-major rows for the forward product) and w2t[h][c] (16 of 64 columns live,
//   K-major rows for the dz2*W2 product)
__global__ void __launch_bounds__(256)
recast_w2_kernel(const float* __restrict__ W2, __nv_bfloat16* __restrict__ w2c, int ldh,
                 __nv_bfloat16* __restrict__ w2t, int H, int C) {
    recast_w2_column(W2, w2c, ldh, w2t, H, C, blockIdx.x * blockDim.x + threadIdx.x);
}

}  // namespace

extern "C" {

int cme_u8_to_bf16_rows(const uint8_t* src, void* dst_bf16, int64_t rows, int cols, int ld,
                        int add_ones, void* stream) {
    return cme_u8_to_bf16_rows_ind(src, nullptr, dst_bf16, rows, cols, ld, add_ones, stream);
}

int cme_u8_to_bf16_rows_ind(const uint8_t* src, const CmeStepArgs* step_args, void* dst_bf16, int64_t rows,
                            int cols, int ld, int add_ones, void* stream) {
    CME_CHECK_ARG((src || step_args) && dst_bf16 && rows >= 0 && cols > 0, "cme_u8_to_bf16_rows: bad argument");
    CME_CHECK_ARG(ld % 16 == 0 && ld >= cols + (add_ones ? 1 : 0),
                  "cme_u8_to_bf16_rows: ld (%d) must be a multiple of 16 and cover %d columns", ld,
                  cols + (add_ones ? 1 : 0));
    CME_CHECK_ARG(((uintptr_t)dst_bf16 & 15) == 0, "cme_u8_to_bf16_rows: dst must be 16-byte aligned");
    if (rows == 0) return 0;
    const int64_t total = rows * (ld / 16);
    u8_to_bf16_kernel<<<cme::stride_grid(total, 256, 8, 4), 256, 0, cme::as_stream(stream)>>>(
        src, step_args, (__nv_bfloat16*)dst_bf16, rows, cols, ld, add_ones, RecastArgs{});
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_stage_recast(const uint8_t* src, const CmeStepArgs* step_args, void* dst_bf16, int64_t rows, int cols,
                         int ld, const float* W2, void* w2c, int ldh, void* w2t, int H, int C, void* stream) {
    CME_CHECK_ARG((src || step_args) && dst_bf16 && rows > 0 && cols > 0, "cme_mlp_stage_recast: bad argument");
    CME_CHECK_ARG(ld % 16 == 0 && ld >= cols + 1, "cme_mlp_stage_recast: ld (%d) must be a multiple of 16 and cover %d columns",
                  ld, cols + 1);
    CME_CHECK_ARG(((uintptr_t)dst_bf16 & 15) == 0, "cme_mlp_stage_recast: dst must be 16-byte aligned");
    CME_CHECK_ARG(W2 && w2c && H > 0 && C >= 1 && C <= 16 && ldh >= H, "cme_mlp_stage_recast: bad W2 argument");
    const int stage_blocks = cme::stride_grid(rows * (ld / 16), 256, 8, 4);
    RecastArgs rc{W2, (__nv_bfloat16*)w2c, (__nv_bfloat16*)w2t, ldh, H, C, stage_blocks};
    u8_to_bf16_kernel<<<stage_blocks + (H + 255) / 256, 256, 0, cme::as_stream(stream)>>>(
        src, step_args, (__nv_bfloat16*)dst_bf16, rows, cols, ld, 1, rc);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_fwd2_softmax(const void* A1, const float* W2, const float* b2, const int32_t* labels,
                         float* dz2, float* loss_sum, int32_t* pred, float* gb2, float scale,
                         int64_t rows, int H, int ldh, int C, void* stream) {
    CME_CHECK_ARG(A1 && W2 && b2 && rows >= 0 && H > 0, "cme_mlp_fwd2_softmax: bad argument");
    CME_CHECK_ARG(C == 10, "cme_mlp_fwd2_softmax: built for 10 classes (got %d)", C);
    CME_CHECK_ARG(ldh % 4 == 0 && ldh >= H, "cme_mlp_fwd2_softmax: ldh (%d)", ldh);
    CME_CHECK_ARG(!dz2 || labels, "cme_mlp_fwd2_softmax: dz2 needs labels");
    if (rows == 0) return 0;
    const int64_t blocks = cme::ceil_div64(rows, 8 * kF2Rows);
    CME_CHECK_ARG(blocks < INT32_MAX, "cme_mlp_fwd2_softmax: too many rows");
    fwd2_softmax_kernel<10><<<(unsigned)blocks, 256, 0, cme::as_stream(stream)>>>(
        (const __nv_bfloat16*)A1, W2, b2, labels, dz2, loss_sum, pred, gb2, scale, rows, H, ldh);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_bwd_elem(const void* A1, const float* dz2, const float* W2, void* dZ1, float* gW2,
                     int64_t rows, int H, int ldh, int C, void* stream) {
    CME_CHECK_ARG(A1 && dz2 && W2 && dZ1 && rows >= 0 && H > 0, "cme_mlp_bwd_elem: bad argument");
    CME_CHECK_ARG(C == 10, "cme_mlp_bwd_elem: built for 10 classes (got %d)", C);
    CME_CHECK_ARG(ldh % 4 == 0 && ldh >= (H + 3) / 4 * 4, "cme_mlp_bwd_elem: ldh (%d)", ldh);
    if (rows == 0) return 0;
    const int col_blocks = (H + 1023) / 1024;
    int64_t row_blocks = (4ll * cme::sm_count() + col_blocks - 1) / col_blocks;
    int64_t rpb = cme::ceil_div64(rows, row_blocks);
    if (rpb < 64) rpb = 64;
    row_blocks = cme::ceil_div64(rows, rpb);
    CME_CHECK_ARG(row_blocks < 65536, "cme_mlp_bwd_elem: too many row blocks");
    if (gW2)
        bwd_elem_kernel<10, true><<<dim3(col_blocks, (unsigned)row_blocks), 256, 0, cme::as_stream(stream)>>>(
            (const __nv_bfloat16*)A1, dz2, W2, (__nv_bfloat16*)dZ1, gW2, rows, rpb, H, ldh);
    else
        bwd_elem_kernel<10, false><<<dim3(col_blocks, (unsigned)row_blocks), 256, 0, cme::as_stream(stream)>>>(
            (const __nv_bfloat16*)A1, dz2, W2, (__nv_bfloat16*)dZ1, nullptr, rows, rpb, H, ldh);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_mlp_recast_w2(const float* W2, void* w2c, int ldh, void* w2t, int H, int C,
                                 void* stream) {
    CME_CHECK_ARG(W2 && w2c && H > 0 && C >= 1 && C <= 16 && ldh >= H, "cme_mlp_recast_w2: bad argument");
    recast_w2_kernel<<<(H + 255) / 256, 256, 0, cme::as_stream(stream)>>>(
        W2, (__nv_bfloat16*)w2c, ldh, (__nv_bfloat16*)w2t, H, C);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_f32_to_bf16_2d(const float* src, int ld_src, void* dst_bf16, int ld_dst, int rows, int cols,
                       void* stream) {
    CME_CHECK_ARG(src && dst_bf16 && rows >= 0 && cols >= 0 && ld_src >= cols && ld_dst >= cols,
                  "cme_f32_to_bf16_2d: bad argument");
    if (rows == 0 || cols == 0) return 0;
    f32_to_bf16_2d_kernel<<<cme::stride_grid((int64_t)rows * cols, 256, 8), 256, 0,
                            cme::as_stream(stream)>>>(src, ld_src, (__nv_bfloat16*)dst_bf16, ld_dst,
                                                      rows, cols);
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
