// fp64 parity path (ref fpcode/gpu_func.cu).
//
// The reference trains in fp64 end to end and grades the parallel run against
// the serial one at 1e-7.  tcgen05 has no fp64 kind, so the fp64 tensor path of
// this chip is the warp-level DMMA (mma.sync.m8n8k4.f64); the GEMM below is built
// on it and covers all four transpose cases of the reference's myGEMM contract
// (ref gpu_func.cu:259-297 dispatches to three 32x32 one-output-per-thread
// kernels), with guarded loads on every edge (the reference read out of bounds
// in M/N on edge tiles, ref gpu_func.cu:108-124):
//   dgemm_mma_kernel     128x128 or 64x64 block tiles, 8 warps, operands through shared
//                        memory with the next k-block prefetched into registers
//   dgemm_skinny_kernel  min(M, N) <= 16 (layer 2, dW2, the -g 4 (800,10,1000) shape):
//                        64x16 tiles, the K dimension split over the 8 warps of a CTA and
//                        over a cluster of up to 8 CTAs, partial tiles summed through
//                        (distributed) shared memory in a fixed order, so the result is
//                        reproducible and needs no workspace and no atomics
//   dgemm_kernel         the round-1 CUDA-core kernel, kept as cme_dgemm_simt for A/B timing
// Fused elementwise kernels: bias+sigmoid, bias+softmax+cross-entropy gradient, sigmoid
// derivative, row sums, SGD.
#include "common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16;  // block tile; 256 threads, 4x4 outputs each

template <bool TA, bool TB>
__global__ void __launch_bounds__(256)
dgemm_kernel(int M, int N, int K, double alpha, const double* __restrict__ A, int lda,
             const double* __restrict__ B, int ldb, double beta, double* __restrict__ C, int ldc) {
    __shared__ double As[BK][BM + 1];
    __shared__ double Bs[BK][BN + 1];
    const int tid = threadIdx.x;
    const int tm = tid % 16, tn = tid / 16;   // thread owns rows tm+16*i, cols tn+16*j
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;

    for (int k0 = 0; k0 < K; k0 += BK) {
        // A tile: BM x BK values of op(A); pick the load order that is contiguous
        // in memory for the given transpose
#pragma unroll
        for (int t = 0; t < (BM * BK) / 256; ++t) {
            const int e = tid + t * 256;
            int m, k;
            if (!TA) { m = e % BM; k = e / BM; } else { k = e % BK; m = e / BK; }
            const int gm = m0 + m, gk = k0 + k;
            double v = 0.0;
            if (gm < M && gk < K) v = TA ? A[(size_t)gk + (size_t)lda * gm] : A[(size_t)gm + (size_t)lda * gk];
            As[k][m] = v;
        }
#pragma unroll
        for (int t = 0; t < (BN * BK) / 256; ++t) {
            const int e = tid + t * 256;
            int n, k;
            if (!TB) { k = e % BK; n = e / BK; } else { n = e % BN; k = e / BN; }
            const int gn = n0 + n, gk = k0 + k;
            double v = 0.0;
            if (gn < N && gk < K) v = TB ? B[(size_t)gn + (size_t)ldb * gk] : B[(size_t)gk + (size_t)ldb * gn];
            Bs[k][n] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[k][tm + 16 * i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[k][tn + 16 * j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int gn = n0 + tn + 16 * j;
        if (gn >= N) continue;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int gm = m0 + tm + 16 * i;
            if (gm >= M) continue;
            double* c = C + (size_t)gm + (size_t)ldc * gn;
            // beta == 0 must not read C (it may hold NaNs from a fresh allocation)
            *c = (beta == 0.0) ? alpha * acc[i][j] : alpha * acc[i][j] + beta * (*c);
        }
    }
}

// ---- DMMA ------------------------------------------------------------------------------------
// D(8x8) += A(8x4) * B(4x8).  Lane l: g = l / 4, t = l % 4 holds A(g, t), B(t, g), D(g, 2t), D(g, 2t+1).
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

constexpr int MK = 16;   // k-block of the shared-memory kernel

__device__ __forceinline__ double ld_dsmem_f64(const double* local, uint32_t cta) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(local);
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(cta));
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(ra) : "memory");
    return v;
}
__device__ __forceinline__ void cluster_barrier() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}


// 8 warps as 4 (M) x 2 (N); a warp owns (BM/4) x (BN/2) of the block tile as m8n8 fragments.
// Shared layouts keep the dimension that is contiguous in global memory contiguous in shared
// memory (conflict-free stores), with row strides chosen so that a fragment read (8 values of g,
// 4 values of t) touches 32 distinct 8-byte words in two wavefronts:
//   [k][m] rows of BM + 8 doubles (stride = 8 mod 16),  [m][k] rows of MK + 4 doubles (stride = 4 mod 16)
template <int BM_, int BN_, bool TA, bool TB>
__global__ void __launch_bounds__(256, (BM_ == 64 ? 2 : 1))
dgemm_mma_kernel(int M, int N, int K, double alpha, const double* __restrict__ A, int lda,
                 const double* __restrict__ B, int ldb, double beta, double* __restrict__ C, int ldc,
                 int k_per_cta) {
    constexpr int SA = TA ? MK + 4 : BM_ + 8;
    constexpr int SB = TB ? BN_ + 8 : MK + 4;
    constexpr int AE = BM_ * MK / 256, BE = BN_ * MK / 256;
    constexpr int WM = BM_ / 4, WN = BN_ / 2, MT = WM / 8, NT = WN / 8;
    constexpr int kAsz = TA ? BM_ * SA : MK * SA, kBsz = TB ? MK * SB : BN_ * SB;
    // split-K over a cluster (gridDim.z CTAs, 64 x 64 tiles only): the operand buffers are reused
    // for the partial tile, [n][m] with rows of BM_ + 2 doubles
    constexpr int RS = BM_ + 2;
    constexpr int kRed = BM_ == 64 ? BN_ * RS : 0;
    __shared__ double smem_all[(kAsz + kBsz > kRed) ? kAsz + kBsz : kRed];
    double* As = smem_all;
    double* Bs = smem_all + kAsz;
    const int k_begin = (int)blockIdx.z * k_per_cta, k_end = min(K, k_begin + k_per_cta);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int m0 = blockIdx.x * BM_, n0 = blockIdx.y * BN_;
    const int wm0 = (warp & 3) * WM, wn0 = (warp >> 2) * WN;
    double acc[MT][NT][2];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    double ra[AE], rb[BE];

    auto gload = [&](int k0) {
#pragma unroll
        for (int x = 0; x < AE; ++x) {
            const int e = tid + x * 256;
            const int m = TA ? e / MK : e % BM_, k = TA ? e % MK : e / BM_;
            const int gm = m0 + m, gk = k0 + k;
            ra[x] = (gm < M && gk < k_end) ? (TA ? A[(size_t)gk + (size_t)lda * gm] : A[(size_t)gm + (size_t)lda * gk]) : 0.0;
        }
#pragma unroll
        for (int x = 0; x < BE; ++x) {
            const int e = tid + x * 256;
            const int n = TB ? e % BN_ : e / MK, k = TB ? e / BN_ : e % MK;
            const int gn = n0 + n, gk = k0 + k;
            rb[x] = (gn < N && gk < k_end) ? (TB ? B[(size_t)gn + (size_t)ldb * gk] : B[(size_t)gk + (size_t)ldb * gn]) : 0.0;
        }
    };

    gload(k_begin);
    for (int k0 = k_begin; k0 < k_end; k0 += MK) {
#pragma unroll
        for (int x = 0; x < AE; ++x) {
            const int e = tid + x * 256;
            const int m = TA ? e / MK : e % BM_, k = TA ? e % MK : e / BM_;
            As[TA ? m * SA + k : k * SA + m] = ra[x];
        }
#pragma unroll
        for (int x = 0; x < BE; ++x) {
            const int e = tid + x * 256;
            const int n = TB ? e % BN_ : e / MK, k = TB ? e / BN_ : e % MK;
            Bs[TB ? k * SB + n : n * SB + k] = rb[x];
        }
        __syncthreads();
        if (k0 + MK < k_end) gload(k0 + MK);       // in flight while the tensor pipe works on this block
#pragma unroll
        for (int kk = 0; kk < MK; kk += 4) {
            double a[MT], b[NT];
            const int k = kk + t;
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                const int m = wm0 + i * 8 + g;
                a[i] = As[TA ? m * SA + k : k * SA + m];
            }
#pragma unroll
            for (int j = 0; j < NT; ++j) {
                const int n = wn0 + j * 8 + g;
                b[j] = Bs[TB ? k * SB + n : n * SB + k];
            }
#pragma unroll
            for (int i = 0; i < MT; ++i)
#pragma unroll
                for (int j = 0; j < NT; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        __syncthreads();
    }
    if (BM_ == 64 && gridDim.z > 1) {
        // every CTA of the cluster leaves its partial tile in shared memory; CTA r then adds rows
        // [r * BM_/S, (r+1) * BM_/S) of all of them in CTA order and writes that part of C
        const int S = gridDim.z, r = blockIdx.z;
#pragma unroll
        for (int j = 0; j < NT; ++j)
#pragma unroll
            for (int x = 0; x < 2; ++x)
#pragma unroll
                for (int i = 0; i < MT; ++i)
                    smem_all[(wn0 + j * 8 + 2 * t + x) * RS + wm0 + i * 8 + g] = acc[i][j][x];
        cluster_barrier();
        const int rows_per = BM_ / S;                 // S is 2, 4 or 8
        for (int e = tid; e < rows_per * BN_; e += 256) {
            const int m = r * rows_per + e % rows_per, n = e / rows_per;
            const double* cell = &smem_all[n * RS + m];
            double v = ld_dsmem_f64(cell, 0);
            for (int c = 1; c < S; ++c) v += ld_dsmem_f64(cell, (uint32_t)c);
            const int gm = m0 + m, gn = n0 + n;
            if (gm < M && gn < N) {
                double* cp = C + (size_t)gm + (size_t)ldc * gn;
                *cp = (beta == 0.0) ? alpha * v : alpha * v + beta * (*cp);
            }
        }
        cluster_barrier();                            // nobody leaves while its tile is being read
        return;
    }
#pragma unroll
    for (int j = 0; j < NT; ++j)
#pragma unroll
        for (int x = 0; x < 2; ++x) {
            const int gn = n0 + wn0 + j * 8 + 2 * t + x;
            if (gn >= N) continue;
#pragma unroll
            for (int i = 0; i < MT; ++i) {
                const int gm = m0 + wm0 + i * 8 + g;
                if (gm >= M) continue;
                double* c = C + (size_t)gm + (size_t)ldc * gn;
                // beta == 0 must not read C (it may hold NaNs from a fresh allocation)
                *c = (beta == 0.0) ? alpha * acc[i][j][x] : alpha * acc[i][j][x] + beta * (*c);
            }
        }
}

// ---- skinny products ---------------------------------------------------------------------
// C(i, j) = alpha * sum_k A(i, k) B(k, j) + beta * C(i, j) for i < I, j < J <= 16, all three
// operands addressed through element strides (the host maps the transposes, and the M <= 16
// case as the transposed product, onto them).  One cluster per 64 rows of i; CTA c of the
// cluster and warp w of the CTA take the k-blocks of four q = (x * S + c) * 8 + w.  Fragments
// come straight from global memory: every operand element is used once per cluster, and a
// fragment load is eight 32-byte sectors, all of them used, whichever dimension is contiguous.
constexpr int SK_I = 64, SK_J = 16, SK_LD = SK_I + 4;      // partial tile [j][i], rows of 68 doubles

__global__ void __launch_bounds__(256)
dgemm_skinny_kernel(int I, int J, int K, double alpha, const double* __restrict__ A, int64_t ai, int64_t ak,
                    const double* __restrict__ B, int64_t bk, int64_t bj, double beta, double* __restrict__ C,
                    int64_t ci, int64_t cj) {
    __shared__ double red[4][SK_J * SK_LD];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int S = gridDim.x;                     // cluster = the CTAs of one row of the grid
    const int cta = blockIdx.x;
    const int i0 = blockIdx.y * SK_I;
    double acc[8][2][2];
#pragma unroll
    for (int m = 0; m < 8; ++m)
#pragma unroll
        for (int n = 0; n < 2; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;

    const int nq = (K + 3) >> 2;
    const bool j0_in = g < J, j1_in = 8 + g < J;
#pragma unroll 2
    for (int q = cta * 8 + warp; q < nq; q += S * 8) {
        const int k = q * 4 + t;
        const bool kin = k < K;
        double a[8], b[2];
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            const int i = i0 + m * 8 + g;
            a[m] = (kin && i < I) ? A[(int64_t)i * ai + (int64_t)k * ak] : 0.0;
        }
        b[0] = (kin && j0_in) ? B[(int64_t)k * bk + (int64_t)g * bj] : 0.0;
        b[1] = (kin && j1_in) ? B[(int64_t)k * bk + (int64_t)(8 + g) * bj] : 0.0;
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            dmma884(acc[m][0][0], acc[m][0][1], a[m], b[0]);
            dmma884(acc[m][1][0], acc[m][1][1], a[m], b[1]);
        }
    }

    // the 8 warps' partial tiles, added in a fixed order: (0+4, 1+5, 2+6, 3+7), (+2, +3), (+1)
#pragma unroll
    for (int half = 4; half >= 1; half >>= 1) {
        if (warp >= half && warp < 2 * half) {
            double* dst = red[warp - half];
#pragma unroll
            for (int m = 0; m < 8; ++m)
#pragma unroll
                for (int n = 0; n < 2; ++n)
#pragma unroll
                    for (int x = 0; x < 2; ++x) dst[(n * 8 + 2 * t + x) * SK_LD + m * 8 + g] = acc[m][n][x];
        }
        __syncthreads();
        if (warp < half) {
            const double* src = red[warp];
#pragma unroll
            for (int m = 0; m < 8; ++m)
#pragma unroll
                for (int n = 0; n < 2; ++n)
#pragma unroll
                    for (int x = 0; x < 2; ++x) acc[m][n][x] += src[(n * 8 + 2 * t + x) * SK_LD + m * 8 + g];
        }
        __syncthreads();
    }
    if (warp == 0) {
#pragma unroll
        for (int m = 0; m < 8; ++m)
#pragma unroll
            for (int n = 0; n < 2; ++n)
#pragma unroll
                for (int x = 0; x < 2; ++x) red[0][(n * 8 + 2 * t + x) * SK_LD + m * 8 + g] = acc[m][n][x];
    }
    __syncthreads();
    if (S > 1) cluster_barrier();                // every CTA's tile is in its red[0]
    if (cta == 0) {
        for (int e = tid; e < SK_I * SK_J; e += 256) {
            const int i = e % SK_I, j = e / SK_I;
            const double* cell = &red[0][j * SK_LD + i];
            double v = *cell;
            for (int r = 1; r < S; ++r) v += ld_dsmem_f64(cell, (uint32_t)r);   // fixed order
            if (i0 + i < I && j < J) {
                double* c = C + (int64_t)(i0 + i) * ci + (int64_t)j * cj;
                *c = (beta == 0.0) ? alpha * v : alpha * v + beta * (*c);
            }
        }
    }
    if (S > 1) cluster_barrier();                // nobody leaves while CTA 0 still reads its tile
}

__device__ __forceinline__ double sigmoid_d(double x) { return 1.0 / (1.0 + exp(-x)); }

__global__ void __launch_bounds__(256)
bias_sigmoid_kernel(double* __restrict__ z, const double* __restrict__ bias, int rows, size_t n) {
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride)
        z[i] = sigmoid_d(z[i] + bias[i % rows]);
}

// one warp per column (sample): rows = number of classes (10 in the MLP)
__global__ void __launch_bounds__(256)
bias_softmax_ce_kernel(double* __restrict__ z, const double* __restrict__ bias,
                       const int32_t* __restrict__ labels, double* __restrict__ diff, double scale,
                       double* __restrict__ loss_sum, int rows, int cols) {
    const int lane = threadIdx.x & 31;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    double loss_local = 0.0;
    for (int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; c < cols; c += warps) {
        double* col = z + (size_t)c * rows;
        double mx = -1e300;
        for (int r = lane; r < rows; r += 32) mx = fmax(mx, col[r] + bias[r]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        double s = 0.0;
        for (int r = lane; r < rows; r += 32) s += exp(col[r] + bias[r] - mx);
        s = cme::warp_sum(s);
        const int lab = labels ? labels[c] : -1;
        for (int r = lane; r < rows; r += 32) {
            const double p = exp(col[r] + bias[r] - mx) / s;
            col[r] = p;
            if (diff) diff[(size_t)c * rows + r] = (p - (r == lab ? 1.0 : 0.0)) * scale;
            if (r == lab) loss_local -= log(p);
        }
    }
    if (loss_sum) {
        loss_local = cme::warp_sum(loss_local);
        if (lane == 0 && loss_local != 0.0) atomicAdd(loss_sum, loss_local);
    }
}

__global__ void __launch_bounds__(256)
dsigmoid_kernel(const double* __restrict__ da, const double* __restrict__ a,
                double* __restrict__ dz, size_t n) {
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride) {
        const double y = a[i];
        dz[i] = da[i] * y * (1.0 - y);
    }
}

// out[r] = sum over columns; one block per 32 rows, 8 warps stride the columns
// and the partial sums meet in shared memory (the reference used one thread per
// row with a serial loop, ref gpu_func.cu:359-369)
__global__ void __launch_bounds__(256)
row_sum_kernel(const double* __restrict__ m, double* __restrict__ out, int rows, int cols) {
    __shared__ double part[8][33];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int r = blockIdx.x * 32 + lane;
    double s = 0.0;
    if (r < rows)
        for (int c = w; c < cols; c += 8) s += m[(size_t)r + (size_t)rows * c];
    part[w][lane] = s;
    __syncthreads();
    if (w == 0 && r < rows) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += part[k][lane];
        out[r] = t;
    }
}

__global__ void __launch_bounds__(256)
sgd_kernel(double* __restrict__ p, const double* __restrict__ g, size_t n, double lr, double reg) {
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride)
        p[i] = p[i] - lr * (g[i] + reg * p[i]);
}

__global__ void __launch_bounds__(256)
axpby_kernel(double* __restrict__ c, const double* __restrict__ a, const double* __restrict__ b,
             double alpha, double beta, size_t n) {
    const size_t gstride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gstride)
        c[i] = alpha * a[i] + beta * b[i];
}

}  // namespace

namespace {

int dgemm_check(int transA, int transB, int M, int N, int K, const double* A, int lda, const double* B, int ldb,
                double* C, int ldc) {
    CME_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "cme_dgemm: negative dimension");
    if (M == 0 || N == 0) return 1;
    CME_CHECK_ARG(A && B && C, "cme_dgemm: null matrix");
    CME_CHECK_ARG(lda >= (transA ? K : M) && ldb >= (transB ? N : K) && ldc >= M,
                  "cme_dgemm: leading dimension too small (lda=%d ldb=%d ldc=%d)", lda, ldb, ldc);
    return 0;
}

template <int BM_, int BN_, bool TA, bool TB>
int launch_mma_t(dim3 grid, int S, int k_per, int M, int N, int K, double alpha, const double* A, int lda,
                 const double* B, int ldb, double beta, double* C, int ldc, cudaStream_t st) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(256, 1, 1);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = S;
    cfg.attrs = attr;
    cfg.numAttrs = S > 1 ? 1 : 0;
    CME_CUDA(cudaLaunchKernelEx(&cfg, dgemm_mma_kernel<BM_, BN_, TA, TB>, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc,
                                k_per));
    return 0;
}

// S > 1 (64 x 64 tiles only): K is split over a cluster of S CTAs, see the kernel's epilogue
template <int BM_, int BN_>
int launch_mma(bool ta, bool tb, int S, int M, int N, int K, double alpha, const double* A, int lda, const double* B,
               int ldb, double beta, double* C, int ldc, cudaStream_t st) {
    dim3 grid((M + BM_ - 1) / BM_, (N + BN_ - 1) / BN_, S);
    CME_CHECK_ARG(grid.y < 65536, "cme_dgemm: N too large for one launch");
    const int k_per = S > 1 ? ((K + S - 1) / S + MK - 1) / MK * MK : (K > 0 ? K : 1);
    if (!ta && !tb) return launch_mma_t<BM_, BN_, false, false>(grid, S, k_per, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
    if (ta && !tb) return launch_mma_t<BM_, BN_, true, false>(grid, S, k_per, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
    if (!ta && tb) return launch_mma_t<BM_, BN_, false, true>(grid, S, k_per, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
    return launch_mma_t<BM_, BN_, true, true>(grid, S, k_per, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
}

int launch_skinny(bool ta, bool tb, int M, int N, int K, double alpha, const double* A, int lda, const double* B,
                  int ldb, double beta, double* C, int ldc, cudaStream_t st) {
    // element strides of op(A)(m, k) and op(B)(k, n)
    const int64_t am = ta ? lda : 1, ak = ta ? 1 : lda;
    const int64_t bk = tb ? ldb : 1, bn = tb ? 1 : ldb;
    int I, J;
    const double *P, *Q;
    int64_t pi, pk, qk, qj, ci, cj;
    if (N <= SK_J) {            // tall C: rows of C are the long dimension
        I = M, J = N, P = A, pi = am, pk = ak, Q = B, qk = bk, qj = bn, ci = 1, cj = ldc;
    } else {                    // wide C: the transposed product, C^T = op(B)^T op(A)^T
        I = N, J = M, P = B, pi = bn, pk = bk, Q = A, qk = ak, qj = am, ci = ldc, cj = 1;
    }
    // split K over a cluster while every warp keeps at least four k-blocks of four
    int S = 1;
    while (S < 8 && K >= 128 * 2 * S) S *= 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(S, (I + SK_I - 1) / SK_I, 1);
    cfg.blockDim = dim3(256, 1, 1);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CME_CHECK_ARG(cfg.gridDim.y < 65536, "cme_dgemm: long dimension too large for one launch");
    CME_CUDA(cudaLaunchKernelEx(&cfg, dgemm_skinny_kernel, I, J, K, alpha, P, pi, pk, Q, qk, qj, beta, C, ci, cj));
    return 0;
}

}  // namespace

extern "C" {

int cme_dgemm(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
              const double* B, int ldb, double beta, double* C, int ldc, void* stream) {
    const int chk = dgemm_check(transA, transB, M, N, K, A, lda, B, ldb, C, ldc);
    if (chk) return chk < 0 ? chk : 0;
    cudaStream_t st = cme::as_stream(stream);
    const bool ta = transA != 0, tb = transB != 0;
    if (M <= SK_J || N <= SK_J) return launch_skinny(ta, tb, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
    // 128x128 tiles once they fill the machine; below that 64x64 tiles, two CTAs per SM
    const int64_t big_tiles = (int64_t)((M + 127) / 128) * ((N + 127) / 128);
    if (big_tiles >= cme::sm_count())
        return launch_mma<128, 128>(ta, tb, 1, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
    // 64 x 64 tiles, two CTAs per SM.  Too few of them leave SMs idle or unevenly loaded (ncu on
    // (1000,800,784): DMMA pipe 67 % busy on the SMs that got two tiles, 33 % on those with one),
    // so K is split over a cluster until the CTAs cover the machine twice, at least 64 deep each.
    const int64_t tiles = (int64_t)((M + 63) / 64) * ((N + 63) / 64);
    int S = 1;
    while (S < 8 && tiles * S < 2 * cme::sm_count() && K / (2 * S) >= 64) S *= 2;
    return launch_mma<64, 64>(ta, tb, S, M, N, K, alpha, A, lda, B, ldb, beta, C, ldc, st);
}

/* The round-1 CUDA-core kernel (64x64 tiles, 4x4 outputs per thread), for A/B timing only. */
int cme_dgemm_simt(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
                   const double* B, int ldb, double beta, double* C, int ldc, void* stream) {
    const int chk = dgemm_check(transA, transB, M, N, K, A, lda, B, ldb, C, ldc);
    if (chk) return chk < 0 ? chk : 0;
    dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN);
    CME_CHECK_ARG(grid.y < 65536, "cme_dgemm: N too large for one launch");
    cudaStream_t st = cme::as_stream(stream);
    if (!transA && !transB) dgemm_kernel<false, false><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
    else if (transA && !transB) dgemm_kernel<true, false><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
    else if (!transA && transB) dgemm_kernel<false, true><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
    else dgemm_kernel<true, true><<<grid, 256, 0, st>>>(M, N, K, alpha, A, lda, B, ldb, beta, C, ldc);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_bias_sigmoid(double* z, const double* bias, int rows, int cols, void* stream) {
    CME_CHECK_ARG(z && bias && rows > 0 && cols >= 0, "cme_d_bias_sigmoid: bad argument");
    const size_t n = (size_t)rows * cols;
    if (n == 0) return 0;
    bias_sigmoid_kernel<<<cme::stride_grid((int64_t)n, 256, 8), 256, 0, cme::as_stream(stream)>>>(
        z, bias, rows, n);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_bias_softmax_ce(double* z, const double* bias, const int32_t* labels, double* diff,
                          double scale, double* loss_sum, int rows, int cols, void* stream) {
    CME_CHECK_ARG(z && bias && rows > 0 && cols >= 0, "cme_d_bias_softmax_ce: bad argument");
    CME_CHECK_ARG(!diff || labels, "cme_d_bias_softmax_ce: diff needs labels");
    if (cols == 0) return 0;
    bias_softmax_ce_kernel<<<cme::stride_grid((int64_t)cols * 32, 256, 8), 256, 0,
                             cme::as_stream(stream)>>>(z, bias, labels, diff, scale, loss_sum,
                                                       rows, cols);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_dsigmoid(const double* da, const double* a, double* dz, size_t n, void* stream) {
    CME_CHECK_ARG((da && a && dz) || n == 0, "cme_d_dsigmoid: null buffer");
    if (n == 0) return 0;
    dsigmoid_kernel<<<cme::stride_grid((int64_t)n, 256, 8), 256, 0, cme::as_stream(stream)>>>(
        da, a, dz, n);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_row_sum(const double* m, double* out, int rows, int cols, void* stream) {
    CME_CHECK_ARG(m && out && rows > 0 && cols >= 0, "cme_d_row_sum: bad argument");
    row_sum_kernel<<<(rows + 31) / 32, 256, 0, cme::as_stream(stream)>>>(m, out, rows, cols);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_sgd(double* p, const double* g, size_t n, double lr, double reg, void* stream) {
    CME_CHECK_ARG((p && g) || n == 0, "cme_d_sgd: null buffer");
    if (n == 0) return 0;
    sgd_kernel<<<cme::stride_grid((int64_t)n, 256, 8), 256, 0, cme::as_stream(stream)>>>(p, g, n,
                                                                                        lr, reg);
    CME_LAUNCH_CHECK();
    return 0;
}

int cme_d_axpby(double* c, const double* a, const double* b, double alpha, double beta, size_t n,
                void* stream) {
    CME_CHECK_ARG((c && a && b) || n == 0, "cme_d_axpby: null buffer");
    if (n == 0) return 0;
    axpby_kernel<<<cme::stride_grid((int64_t)n, 256, 8), 256, 0, cme::as_stream(stream)>>>(
        c, a, b, alpha, beta, n);
    CME_LAUNCH_CHECK();
    return 0;
}

}  // extern "C"
