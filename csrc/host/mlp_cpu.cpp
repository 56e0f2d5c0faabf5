// Native serial trainer of the two-layer MLP on the CPU, fp64.
//
// The reference's serial path is C++ on Armadillo (ref fpcode/neural_network.cpp:91-279:
// feedforward, backprop, loss, predict, train).  The oracle of this repo is the same math on torch
// CPU tensors (cme213_sp18_b200/models/mlp.py); this file is its counterpart for hosts without
// Python: plain C++17 + OpenMP, no BLAS dependency, row-major buffers (one sample per row, which is
// the same bytes as the reference's column-major D x N matrix).
//
// The three GEMM shapes of a step are written out separately because their contiguous dimension
// differs; each keeps a small block of the result in registers and is parallel over rows of it:
//   nt:  C[i][j]  = sum_k A[i][k] * B[j][k]      (forward: X W1^T, A1 W2^T)
//   nn:  C[i][j]  = sum_k A[i][k] * B[k][j]      (dA1 = diff W2)
//   tn:  C[i][j]  = sum_k A[k][i] * B[k][j]      (dW = dZ^T A)
#include "cme213b_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

// The build has no -march flag (the library must load on whatever CPU the GPU box has), so the three
// kernels are compiled once per instruction set and picked at load time.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define CME_MULTIVERSION __attribute__((target_clones("avx512f", "avx2,fma", "default")))
#else
#define CME_MULTIVERSION
#endif

namespace {

// C[M][N] = A[M][K] * B[N][K]^T, all row-major.  4 x 4 register block, k innermost (both operands
// are read along their contiguous dimension).
CME_MULTIVERSION void gemm_nt(int64_t M, int64_t N, int64_t K, const double* A, int64_t lda, const double* B, int64_t ldb,
             double* C, int64_t ldc) {
#pragma omp parallel for schedule(static)
    for (int64_t i0 = 0; i0 < M; i0 += 4) {
        const int64_t mi = std::min<int64_t>(4, M - i0);
        for (int64_t j0 = 0; j0 < N; j0 += 4) {
            const int64_t nj = std::min<int64_t>(4, N - j0);
            double acc[4][4] = {};
            if (mi == 4 && nj == 4) {
                const double *a0 = A + (i0 + 0) * lda, *a1 = A + (i0 + 1) * lda, *a2 = A + (i0 + 2) * lda,
                             *a3 = A + (i0 + 3) * lda;
                const double *b0 = B + (j0 + 0) * ldb, *b1 = B + (j0 + 1) * ldb, *b2 = B + (j0 + 2) * ldb,
                             *b3 = B + (j0 + 3) * ldb;
                double s00 = 0, s01 = 0, s02 = 0, s03 = 0, s10 = 0, s11 = 0, s12 = 0, s13 = 0;
                double s20 = 0, s21 = 0, s22 = 0, s23 = 0, s30 = 0, s31 = 0, s32 = 0, s33 = 0;
                // sixteen dot products at once, vectorised along k (each sum becomes a vector of partial sums)
#pragma omp simd reduction(+ : s00, s01, s02, s03, s10, s11, s12, s13, s20, s21, s22, s23, s30, s31, s32, s33)
                for (int64_t k = 0; k < K; ++k) {
                    const double x0 = a0[k], x1 = a1[k], x2 = a2[k], x3 = a3[k];
                    const double y0 = b0[k], y1 = b1[k], y2 = b2[k], y3 = b3[k];
                    s00 += x0 * y0; s01 += x0 * y1; s02 += x0 * y2; s03 += x0 * y3;
                    s10 += x1 * y0; s11 += x1 * y1; s12 += x1 * y2; s13 += x1 * y3;
                    s20 += x2 * y0; s21 += x2 * y1; s22 += x2 * y2; s23 += x2 * y3;
                    s30 += x3 * y0; s31 += x3 * y1; s32 += x3 * y2; s33 += x3 * y3;
                }
                acc[0][0] = s00; acc[0][1] = s01; acc[0][2] = s02; acc[0][3] = s03;
                acc[1][0] = s10; acc[1][1] = s11; acc[1][2] = s12; acc[1][3] = s13;
                acc[2][0] = s20; acc[2][1] = s21; acc[2][2] = s22; acc[2][3] = s23;
                acc[3][0] = s30; acc[3][1] = s31; acc[3][2] = s32; acc[3][3] = s33;
            } else {
                for (int64_t i = 0; i < mi; ++i)
                    for (int64_t j = 0; j < nj; ++j) {
                        const double* a = A + (i0 + i) * lda;
                        const double* b = B + (j0 + j) * ldb;
                        double s = 0.0;
                        for (int64_t k = 0; k < K; ++k) s += a[k] * b[k];
                        acc[i][j] = s;
                    }
            }
            for (int64_t i = 0; i < mi; ++i)
                for (int64_t j = 0; j < nj; ++j) C[(i0 + i) * ldc + j0 + j] = acc[i][j];
        }
    }
}

// C[M][N] = A[M][K] * B[K][N], all row-major: for every row of C, axpy over the rows of B
// (j contiguous, vectorisable); rows of C are independent.
CME_MULTIVERSION void gemm_nn(int64_t M, int64_t N, int64_t K, const double* A, int64_t lda, const double* B, int64_t ldb,
             double* C, int64_t ldc) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < M; ++i) {
        double* c = C + i * ldc;
        for (int64_t j = 0; j < N; ++j) c[j] = 0.0;
        const double* a = A + i * lda;
        for (int64_t k = 0; k < K; ++k) {
            const double x = a[k];
            const double* b = B + k * ldb;
#pragma omp simd
            for (int64_t j = 0; j < N; ++j) c[j] += x * b[j];
        }
    }
}

// C[M][N] = A[K][M]^T * B[K][N] + beta_c * Cadd[M][N] (Cadd may be null), row-major.  Threads take
// column blocks of C: a block of B (K rows x ~N/threads columns) then stays in the core's L2 while
// the rows of C are walked four at a time (a 4-wide column strip of A, j contiguous).  Splitting
// over the rows of C instead made every thread stream all of B once per strip: 125 MB of traffic
// per step of the 784-100-10 model, which was most of its time.
CME_MULTIVERSION void gemm_tn(int64_t M, int64_t N, int64_t K, const double* A, int64_t lda, const double* B,
                              int64_t ldb, double* C, int64_t ldc, double beta_c, const double* Cadd,
                              int64_t ldadd) {
    int nthreads = 1;
#ifdef _OPENMP
    nthreads = omp_get_max_threads();
#endif
    int64_t NB = ((N + nthreads - 1) / nthreads + 7) / 8 * 8;
    if (NB < 8) NB = 8;
#pragma omp parallel for schedule(static)
    for (int64_t jb = 0; jb < N; jb += NB) {
        const int64_t je = std::min<int64_t>(N, jb + NB);
        for (int64_t i0 = 0; i0 < M; i0 += 4) {
            const int64_t mi = std::min<int64_t>(4, M - i0);
            for (int64_t i = 0; i < mi; ++i) {
                double* c = C + (i0 + i) * ldc;
                if (Cadd) {
                    const double* w = Cadd + (i0 + i) * ldadd;
                    for (int64_t j = jb; j < je; ++j) c[j] = beta_c * w[j];
                } else {
                    for (int64_t j = jb; j < je; ++j) c[j] = 0.0;
                }
            }
            for (int64_t k = 0; k < K; ++k) {
                const double* b = B + k * ldb;
                const double* a = A + k * lda + i0;
                if (mi == 4) {
                    const double x0 = a[0], x1 = a[1], x2 = a[2], x3 = a[3];
                    double *c0 = C + (i0 + 0) * ldc, *c1 = C + (i0 + 1) * ldc, *c2 = C + (i0 + 2) * ldc,
                           *c3 = C + (i0 + 3) * ldc;
#pragma omp simd
                    for (int64_t j = jb; j < je; ++j) {
                        const double y = b[j];
                        c0[j] += x0 * y;
                        c1[j] += x1 * y;
                        c2[j] += x2 * y;
                        c3[j] += x3 * y;
                    }
                } else {
                    for (int64_t i = 0; i < mi; ++i) {
                        const double x = a[i];
                        double* c = C + (i0 + i) * ldc;
#pragma omp simd
                        for (int64_t j = jb; j < je; ++j) c[j] += x * b[j];
                    }
                }
            }
        }
    }
}

struct Net {
    int D, H, C;
    double *W1, *b1, *W2, *b2;
};

struct Work {
    std::vector<double> a1, z2, diff, da1, dW1, dW2, db1, db2;
    void size_for(int64_t B, const Net& n) {
        a1.resize((size_t)B * n.H);
        z2.resize((size_t)B * n.C);
        diff.resize((size_t)B * n.C);
        da1.resize((size_t)B * n.H);
        dW1.resize((size_t)n.H * n.D);
        dW2.resize((size_t)n.C * n.H);
        db1.resize(n.H);
        db2.resize(n.C);
    }
};

// a1 = sigmoid(X W1^T + b1), yc = softmax(a1 W2^T + b2) left in w.z2 (ref neural_network.cpp:91-114;
// the row maximum is subtracted before exponentiating, like models/mlp.py::softmax)
void forward(const Net& n, const double* X, int64_t B, Work& w) {
    // (layer 1 through a materialised W1^T, so that the inner loop runs along the hidden units, was
    // tried and is slower here: 0.54 vs 0.33 s per epoch of 54 000 samples)
    gemm_nt(B, n.H, n.D, X, n.D, n.W1, n.D, w.a1.data(), n.H);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < B; ++i) {
        double* a = w.a1.data() + i * n.H;
        for (int h = 0; h < n.H; ++h) a[h] = 1.0 / (1.0 + std::exp(-(a[h] + n.b1[h])));
    }
    gemm_nt(B, n.C, n.H, w.a1.data(), n.H, n.W2, n.H, w.z2.data(), n.C);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < B; ++i) {
        double* z = w.z2.data() + i * n.C;
        double mx = -1e300;
        for (int c = 0; c < n.C; ++c) {
            z[c] += n.b2[c];
            mx = std::max(mx, z[c]);
        }
        double s = 0.0;
        for (int c = 0; c < n.C; ++c) {
            z[c] = std::exp(z[c] - mx);
            s += z[c];
        }
        for (int c = 0; c < n.C; ++c) z[c] /= s;
    }
}

// mean cross-entropy of the batch + 0.5 * reg * (||W1||^2 + ||W2||^2) (ref neural_network.cpp:144-154)
double batch_loss(const Net& n, const int32_t* labels, int64_t B, const Work& w, double reg) {
    double ce = 0.0;
    for (int64_t i = 0; i < B; ++i) ce -= std::log(w.z2[(size_t)i * n.C + labels[i]]);
    double l2 = 0.0;
    const int64_t n1 = (int64_t)n.H * n.D, n2 = (int64_t)n.C * n.H;
#pragma omp parallel for reduction(+ : l2) schedule(static)
    for (int64_t i = 0; i < n1; ++i) l2 += n.W1[i] * n.W1[i];
    for (int64_t i = 0; i < n2; ++i) l2 += n.W2[i] * n.W2[i];
    return ce / (double)B + 0.5 * reg * l2;
}

// gradients of one minibatch (ref neural_network.cpp:123-139), L2 term included
void backward(const Net& n, const double* X, const int32_t* labels, int64_t B, double reg, Work& w) {
    const double inv = 1.0 / (double)B;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < B; ++i)
        for (int c = 0; c < n.C; ++c)
            w.diff[(size_t)i * n.C + c] = (w.z2[(size_t)i * n.C + c] - (c == labels[i] ? 1.0 : 0.0)) * inv;
    gemm_tn(n.C, n.H, B, w.diff.data(), n.C, w.a1.data(), n.H, w.dW2.data(), n.H, reg, n.W2, n.H);
    for (int c = 0; c < n.C; ++c) {
        double s = 0.0;
        for (int64_t i = 0; i < B; ++i) s += w.diff[(size_t)i * n.C + c];
        w.db2[c] = s;
    }
    gemm_nn(B, n.H, n.C, w.diff.data(), n.C, n.W2, n.H, w.da1.data(), n.H);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < B; ++i) {
        double* d = w.da1.data() + i * n.H;
        const double* a = w.a1.data() + i * n.H;
        for (int h = 0; h < n.H; ++h) d[h] *= a[h] * (1.0 - a[h]);
    }
    gemm_tn(n.H, n.D, B, w.da1.data(), n.H, X, n.D, w.dW1.data(), n.D, reg, n.W1, n.D);
#pragma omp parallel for schedule(static)
    for (int h = 0; h < n.H; ++h) {
        double s = 0.0;
        for (int64_t i = 0; i < B; ++i) s += w.da1[(size_t)i * n.H + h];
        w.db1[h] = s;
    }
}

bool bad_net(int64_t N, int D, int H, int C, const void* a, const void* b, const void* c, const void* d) {
    return N < 0 || D <= 0 || H <= 0 || C <= 0 || !a || !b || !c || !d;
}

}  // namespace

int64_t cmeh_mlp_train(const double* X, const int32_t* labels, int64_t N, int D, int H, int C, double* W1,
                       double* b1, double* W2, double* b2, double lr, double reg, int epochs, int batch,
                       int loss_every, double* losses, int64_t losses_cap) {
    if (bad_net(N, D, H, C, W1, b1, W2, b2) || (N > 0 && (!X || !labels)) || epochs < 0 || batch <= 0) return -1;
    for (int64_t i = 0; i < N; ++i)
        if (labels[i] < 0 || labels[i] >= C) return -2;
    Net n{D, H, C, W1, b1, W2, b2};
    Work w;
    w.size_for(std::min<int64_t>(batch, std::max<int64_t>(N, 1)), n);
    int64_t it = 0, nloss = 0;
    const int64_t nb = (N + batch - 1) / batch;
    for (int e = 0; e < epochs; ++e) {
        for (int64_t bi = 0; bi < nb; ++bi) {
            const int64_t lo = bi * batch, B = std::min<int64_t>(batch, N - lo);   // the last batch may be short
            const double* Xb = X + lo * D;
            forward(n, Xb, B, w);
            if (loss_every > 0 && it % loss_every == 0 && losses && nloss < losses_cap)
                losses[nloss++] = batch_loss(n, labels + lo, B, w, reg);
            backward(n, Xb, labels + lo, B, reg, w);
            const int64_t n1 = (int64_t)H * D, n2 = (int64_t)C * H;
#pragma omp parallel for schedule(static)
            for (int64_t i = 0; i < n1; ++i) W1[i] -= lr * w.dW1[i];
            for (int64_t i = 0; i < n2; ++i) W2[i] -= lr * w.dW2[i];
            for (int h = 0; h < H; ++h) b1[h] -= lr * w.db1[h];
            for (int c = 0; c < C; ++c) b2[c] -= lr * w.db2[c];
            ++it;
        }
    }
    return it;
}

int cmeh_mlp_forward(const double* X, int64_t N, int D, int H, int C, const double* W1, const double* b1,
                     const double* W2, const double* b2, double* probs, int32_t* pred) {
    if (bad_net(N, D, H, C, W1, b1, W2, b2) || (N > 0 && !X)) return -1;
    Net n{D, H, C, const_cast<double*>(W1), const_cast<double*>(b1), const_cast<double*>(W2), const_cast<double*>(b2)};
    Work w;
    const int64_t chunk = 4096;
    w.size_for(std::min<int64_t>(chunk, std::max<int64_t>(N, 1)), n);
    for (int64_t lo = 0; lo < N; lo += chunk) {
        const int64_t B = std::min<int64_t>(chunk, N - lo);
        forward(n, X + lo * D, B, w);
        if (probs) std::memcpy(probs + lo * C, w.z2.data(), sizeof(double) * (size_t)B * C);
        if (pred)
            for (int64_t i = 0; i < B; ++i) {
                const double* z = w.z2.data() + i * C;
                int best = 0;
                for (int c = 1; c < C; ++c)
                    if (z[c] > z[best]) best = c;
                pred[lo + i] = best;
            }
    }
    return 0;
}
