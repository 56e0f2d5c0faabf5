// The serial fp64 trainer from C++ with no GPU and no Python: the host library's C ABI only
// (include/cme213b_host.h, libcme213b_host.so).  Counterpart of the reference's `./main -s`
// sequential run (ref fpcode/main.cpp:212-231, fpcode/neural_network.cpp:219-279) on synthetic
// data: every class lights up its own block of pixels, plus noise.
//
//   train_mlp_cpu_example [hidden=64] [samples=6000] [epochs=20] [lr=1.0]
#include "cme213b_host.h"

#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

int main(int argc, char** argv) {
    const int D = 784, C = 10;
    const int H = argc > 1 ? std::atoi(argv[1]) : 64;
    const long N = argc > 2 ? std::atol(argv[2]) : 6000;
    const int epochs = argc > 3 ? std::atoi(argv[3]) : 20;
    const double lr = argc > 4 ? std::atof(argv[4]) : 1.0;
    if (H <= 0 || N <= 0 || epochs <= 0) {
        std::fprintf(stderr, "usage: %s [hidden] [samples] [epochs] [lr]\n", argv[0]);
        return 2;
    }
    std::mt19937_64 rng(7);
    std::uniform_real_distribution<double> noise(0.0, 0.3);
    std::vector<double> X((size_t)N * D);
    std::vector<int32_t> y(N);
    for (long i = 0; i < N; ++i) {
        y[i] = (int32_t)(rng() % C);
        for (int d = 0; d < D; ++d) X[(size_t)i * D + d] = noise(rng) + ((d * C) / D == y[i] ? 0.7 : 0.0);
    }
    // the reference's initialisation: W = 0.01 * randn, b = 0 (ref fpcode/inc/neural_network.h:20-30)
    std::normal_distribution<double> g(0.0, 0.01);
    std::vector<double> W1((size_t)H * D), b1(H, 0.0), W2((size_t)C * H), b2(C, 0.0);
    for (auto& w : W1) w = g(rng);
    for (auto& w : W2) w = g(rng);
    const long n_train = N - N / 10;                      // last 10 % held out, like the reference
    const int batch = 800, every = 4;
    std::vector<double> losses(1024);
    const long long it = cmeh_mlp_train(X.data(), y.data(), n_train, D, H, C, W1.data(), b1.data(), W2.data(),
                                        b2.data(), lr, 1e-4, epochs, batch, every, losses.data(), (long long)losses.size());
    if (it < 0) {
        std::fprintf(stderr, "cmeh_mlp_train failed (%lld)\n", it);
        return 1;
    }
    const long nl = (it + every - 1) / every;
    std::printf("threads %d, %lld iterations, loss %.4f -> %.4f\n", cmeh_num_threads(), it, losses[0],
                losses[nl > 0 ? nl - 1 : 0]);
    std::vector<int32_t> pred(N - n_train);
    if (cmeh_mlp_forward(X.data() + (size_t)n_train * D, N - n_train, D, H, C, W1.data(), b1.data(), W2.data(),
                         b2.data(), nullptr, pred.data()) != 0)
        return 1;
    long hit = 0;
    for (long i = 0; i < N - n_train; ++i) hit += pred[i] == y[n_train + i];
    const double acc = (double)hit / (double)(N - n_train);
    std::printf("accuracy %.3f on %ld held-out samples\n", acc, N - n_train);
    std::printf(acc > 0.9 && losses[nl - 1] < losses[0] ? "TRAIN_CPU_OK\n" : "TRAIN_CPU_FAILED\n");
    return acc > 0.9 ? 0 : 1;
}
