// fp64 parity path (ref fpcode/gpu_func.cu).
//
// The reference trains in fp64 end to end and grades the parallel run against
// the serial one at 1e-7, and tcgen05 has no fp64 kind, so grade mode keeps an
// fp64 SIMT path: one register-tiled GEMM that covers all four transpose cases
// of the reference's myGEMM contract (ref gpu_func.cu:259-297 dispatches to
// three 32x32 one-output-per-thread kernels), with guarded loads on every edge
// (the reference read out of bounds in M/N on edge tiles, ref gpu_func.cu:108-124),
// and fused elementwise kernels: bias+sigmoid, bias+softmax+cross-entropy
// gradient, sigmoid derivative, row sums, SGD.
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

extern "C" {

int cme_dgemm(int transA, int transB, int M, int N, int K, double alpha, const double* A, int lda,
              const double* B, int ldb, double beta, double* C, int ldc, void* stream) {
    CME_CHECK_ARG(M >= 0 && N >= 0 && K >= 0, "cme_dgemm: negative dimension");
    if (M == 0 || N == 0) return 0;
    CME_CHECK_ARG(A && B && C, "cme_dgemm: null matrix");
    CME_CHECK_ARG(lda >= (transA ? K : M) && ldb >= (transB ? N : K) && ldc >= M,
                  "cme_dgemm: leading dimension too small (lda=%d ldb=%d ldc=%d)", lda, ldb, ldc);
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
