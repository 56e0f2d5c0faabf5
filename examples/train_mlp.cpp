// Training the two-layer MLP from plain C++ through the C ABI of libcme213b.so: no Python, no
// torch.  The program plays the role of the reference's `main` + `parallel_train` for one rank
// (ref fpcode/main.cpp:34-276, fpcode/neural_network.cpp:401-575): it builds a synthetic
// MNIST-shaped dataset in pinned host memory, allocates one rank's device buffers, fills a
// CmeMlpPlan, and lets the native runtime stream every minibatch to the GPU.
//
//   train_mlp_example [hidden=256] [batch=1024] [epochs=4] [samples=16384] [lr=0.3]
//
// Prints the per-sample loss of the first and last step, the accuracy of the last minibatch and
// the throughput; exits 0 when the loss went down.
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "cme213b.h"

#define CK(call)                                                                        \
    do {                                                                                \
        int rc__ = (int)(call);                                                         \
        if (rc__ != 0) {                                                                \
            std::fprintf(stderr, "%s:%d %s -> %d (%s / %s)\n", __FILE__, __LINE__, #call, rc__, \
                         cme_last_error(), rc__ > 0 ? cudaGetErrorString((cudaError_t)rc__) : "");  \
            return 1;                                                                   \
        }                                                                               \
    } while (0)

namespace {

struct Rng {
    uint64_t s;
    uint32_t next() {
        s = s * 6364136223846793005ull + 1442695040888963407ull;
        return (uint32_t)(s >> 33);
    }
    float uniform() { return (next() & 0xFFFFFF) / 16777216.0f; }
    float normal() {  // Box-Muller
        const float u = uniform() + 1e-7f, v = uniform();
        return std::sqrt(-2.0f * std::log(u)) * std::cos(6.2831853f * v);
    }
};

int round_up(int v, int m) { return (v + m - 1) / m * m; }

template <typename T>
int dev_zeros(T** p, size_t count) {
    cudaError_t e = cudaMalloc((void**)p, count * sizeof(T));
    if (e != cudaSuccess) return (int)e;
    return (int)cudaMemset(*p, 0, count * sizeof(T));
}

}  // namespace

int main(int argc, char** argv) {
    const int hidden = argc > 1 ? std::atoi(argv[1]) : 256;
    const int batch = argc > 2 ? std::atoi(argv[2]) : 1024;
    const int epochs = argc > 3 ? std::atoi(argv[3]) : 4;
    const int samples = argc > 4 ? std::atoi(argv[4]) : 16384;
    const float lr = argc > 5 ? (float)std::atof(argv[5]) : 0.3f;
    const float reg = 1e-4f;
    const int D = 784, C = 10;
    if (hidden < 16 || batch < 1 || epochs < 1 || samples < batch) {
        std::fprintf(stderr, "usage: %s [hidden>=16] [batch] [epochs] [samples>=batch] [lr]\n", argv[0]);
        return 2;
    }

    int warm = 0;
    CK(cme_warmup(&warm));

    // ---- data: ten random prototype images, samples are noisy copies (pixels in {0, 1}) ----
    uint8_t* hx = nullptr;
    int32_t* hy = nullptr;
    float* hloss = nullptr;
    CK(cudaHostAlloc((void**)&hx, (size_t)samples * D, cudaHostAllocDefault));
    CK(cudaHostAlloc((void**)&hy, (size_t)samples * sizeof(int32_t), cudaHostAllocDefault));
    Rng rng{42};
    std::vector<uint8_t> proto((size_t)C * D);
    for (auto& p : proto) p = rng.uniform() < 0.25f;
    for (int i = 0; i < samples; ++i) {
        const int c = (int)(rng.next() % C);
        hy[i] = c;
        for (int j = 0; j < D; ++j) {
            uint8_t v = proto[(size_t)c * D + j];
            if (rng.uniform() < 0.1f) v ^= 1;
            hx[(size_t)i * D + j] = v;
        }
    }

    // ---- one rank's buffers ----
    CmeMlpPlan plan{};
    plan.D = D, plan.H = hidden, plan.C = C;
    plan.ldx = round_up(D + 1, 64);
    plan.Hp = plan.ldh = round_up(hidden, 64);
    plan.rank = 0, plan.nranks = 1, plan.max_rows = batch;
    const int64_t n_w1 = (int64_t)plan.Hp * plan.ldx;
    const int64_t o_w2 = n_w1, o_b2 = o_w2 + round_up(C * hidden, 4);
    const int64_t n_total = o_b2 + round_up(C, 4);
    float *params = nullptr, *grads = nullptr;
    uint16_t *w1c = nullptr, *xb = nullptr, *a1 = nullptr, *dz1 = nullptr, *dz2b = nullptr, *w2c = nullptr,
             *w2t = nullptr;
    CK(dev_zeros(&params, n_total));
    CK(dev_zeros(&grads, n_total));
    CK(dev_zeros(&w1c, n_w1));
    CK(dev_zeros(&xb, (size_t)batch * plan.ldx));
    CK(dev_zeros(&a1, (size_t)batch * plan.ldh));
    CK(dev_zeros(&dz1, (size_t)batch * plan.ldh));
    CK(dev_zeros(&dz2b, (size_t)batch * 64));
    CK(dev_zeros(&w2c, (size_t)16 * plan.ldh));
    CK(dev_zeros(&w2t, (size_t)plan.Hp * 64));
    CK(dev_zeros(&plan.dz2, (size_t)batch * 16));
    CK(dev_zeros(&plan.pred, (size_t)batch));
    plan.params[0] = params, plan.grads[0] = grads, plan.w1c[0] = w1c;
    plan.xb = xb, plan.a1 = a1, plan.dz1 = dz1, plan.dz2b = dz2b, plan.w2c = w2c, plan.w2t = w2t;
    CK(cme_mlp_plan_init_sync(&plan));   // flag pad, split-K counters, device-side step arguments

    // weights: 0.01 * N(0,1), biases 0 (ref fpcode/inc/neural_network.h:20-30); b1 lives in
    // column D of the augmented W1, padding stays zero
    std::vector<float> hp((size_t)n_total, 0.0f);
    for (int h = 0; h < hidden; ++h)
        for (int j = 0; j < D; ++j) hp[(size_t)h * plan.ldx + j] = 0.01f * rng.normal();
    for (int i = 0; i < C * hidden; ++i) hp[(size_t)o_w2 + i] = 0.01f * rng.normal();
    CK(cudaMemcpy(params, hp.data(), sizeof(float) * (size_t)n_total, cudaMemcpyHostToDevice));
    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    CK(cme_f32_to_bf16_2d(params, plan.ldx, w1c, plan.ldx, plan.Hp, plan.ldx, stream));
    CK(cme_mlp_recast_w2(params + o_w2, w2c, plan.ldh, w2t, hidden, C, stream));

    // ---- the run: every minibatch of every epoch, queued in one call ----
    std::vector<int64_t> begin, rows, global;
    for (int e = 0; e < epochs; ++e)
        for (int64_t s = 0; s < samples; s += batch) {
            const int64_t n = std::min<int64_t>(batch, samples - s);   // the last batch may be short
            begin.push_back(s), rows.push_back(n), global.push_back(n);
        }
    const int nsteps = (int)rows.size();
    CK(cudaHostAlloc((void**)&hloss, sizeof(float) * (size_t)nsteps, cudaHostAllocDefault));
    CmeMlpRuntime* rt = nullptr;
    CK(cme_mlp_runtime_create(&plan, 1, &rt));
    const auto t0 = std::chrono::steady_clock::now();
    CK(cme_mlp_runtime_run(rt, hx, hy, begin.data(), rows.data(), global.data(), nsteps, lr, reg, hloss,
                           stream));
    CK(cudaStreamSynchronize(stream));
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // accuracy of the last minibatch from the arg-max the layer-2 epilogue leaves behind
    std::vector<int32_t> pred((size_t)rows.back());
    CK(cudaMemcpy(pred.data(), plan.pred, sizeof(int32_t) * pred.size(), cudaMemcpyDeviceToHost));
    int hit = 0;
    for (size_t i = 0; i < pred.size(); ++i) hit += pred[i] == hy[begin.back() + (int64_t)i];
    int64_t st[5];
    CK(cme_mlp_runtime_stats(rt, st));
    const double first = hloss[0] / rows.front(), last = hloss[nsteps - 1] / rows.back();
    std::printf("steps %d  loss/sample first %.4f last %.4f  last-batch accuracy %.3f\n", nsteps, first, last,
                (double)hit / pred.size());
    std::printf("%.1f k samples/s end to end (first run records the graphs)  h2d %lld B  kernels %lld  graphs %lld\n",
                (double)samples * epochs / sec / 1e3, (long long)st[0], (long long)st[2], (long long)st[4]);
    CK(cme_mlp_runtime_destroy(rt));
    uint32_t status = 0;
    CK(cme_mlp_status(&plan, &status));
    CK(cme_mlp_plan_free_sync(&plan));
    if (status != 0) std::fprintf(stderr, "update kernel reported status 0x%x\n", status);
    const bool ok = std::isfinite(last) && last < first;
    std::printf(ok ? "TRAIN_OK\n" : "TRAIN_FAILED\n");
    return ok ? 0 : 1;
}
