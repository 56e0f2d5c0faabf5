// Native streaming trainer runtime for the bf16 MLP engine.
//
// The reference's parallel_train is a C++ loop that, per minibatch, scatters the inputs on the
// host, copies them to the device with blocking cudaMemcpy, runs the kernels one by one with a
// device-wide sync after each, and copies the gradients back (ref
// fpcode/neural_network.cpp:443-567).  This is the same loop for this library, built the
// other way round: the host never waits inside the loop.
//
//   copy stream     H2D of step i+1's uint8 pixels + int32 labels from pinned host memory into
//                   one of two device staging slots
//   compute stream  one CUDA graph per distinct row count holding the whole step
//                   (cme_mlp_step_ex in its indirect form: staging, forward, backward, the
//                   streaming all-reduce + SGD inside the dW1 GEMM;
//                   the staging slot, labels, lr, reg and scale are read from CmeStepArgs in
//                   device memory, written by a one-thread kernel before each replay, so a
//                   learning-rate schedule or a long dataset never adds graphs)
//                   and a one-thread kernel that stores the step's summed cross-entropy
//                   straight into the caller's pinned (device-mapped) loss array.  A
//                   cudaMemcpyAsync per step for those 4 bytes was measured at +17 us of host
//                   time and +21 us of device time per step (copy-engine round trip between
//                   two graphs); the store from inside the graph costs neither.
//
// Two events per slot (copied / consumed) order the two streams; the caller's stream is joined
// at both ends, so from outside cme_mlp_runtime_run behaves like one asynchronous operation on
// `stream`.  Per step the host issues 7 driver calls (~41 us measured), below the 49 us the
// H2D copy of a 784-100-10 minibatch takes, so that config runs at the copy rate.
//
// The graphs are recorded on the runtime's own compute stream (capture on the legacy default
// stream is not allowed).  With more than one rank every rank must run the same sequence of
// steps: the barrier kernels inside the step count epochs on the device.
#include "common.cuh"

#include <cstdlib>

#include <map>
#include <new>
#include <tuple>

struct LossCursor {
    float* base;               // pinned host array mapped into the device address space, or null
    unsigned long long next;   // index of the next step
};

struct CmeMlpRuntime {
    CmeMlpPlan plan;
    int device = 0;
    int use_graphs = 1;
    cudaStream_t copy = nullptr, copy2 = nullptr, compute = nullptr, side = nullptr;
    cudaEvent_t step_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t copied[2] = {nullptr, nullptr}, copied2[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr};
    cudaEvent_t join_in = nullptr, join_out = nullptr;
    uint8_t* x_stage[2] = {nullptr, nullptr};
    int32_t* y_stage[2] = {nullptr, nullptr};
    float* loss_dev = nullptr;
    struct LossCursor* cursor = nullptr;  // device: where the next step's loss goes
    CmeStepArgs* step_args = nullptr;     // owned only if the plan brought none
    using Key = int64_t;                  // rows
    std::map<Key, cudaGraphExec_t> graphs;
    int64_t h2d_bytes = 0, d2h_bytes = 0, launches = 0, steps = 0;
};

namespace {

__global__ void set_cursor_kernel(LossCursor* c, float* base) {
    c->base = base;
    c->next = 0;
}

// Last node of every step: hand the loss to the host without a copy-engine round trip.
// It also clears the accumulator for the next step (a memset node per step plus a system-scope
// fence behind the host store cost ~10 us per step on the 784-100-10 model; the store is visible
// to the host once the stream is synchronised, which is when the caller may read it).
__global__ void publish_loss_kernel(LossCursor* c, float* loss) {
    float* base = c->base;
    const unsigned long long i = c->next;
    if (base) base[i] = *loss;
    *loss = 0.f;
    c->next = i + 1;
}

int kernels_per_step(const CmeMlpPlan& p) { return cme_mlp_step_launches(&p) + 2; }   // + set args, publish loss

// the step proper, with its per-step arguments read from rt->plan.step_args on the device
// development probe (CME_RT_PROBE bit mask): 1 = skip the H2D copies, 2 = skip the loss hand-over
int probe_mask() {
    static const int m = getenv("CME_RT_PROBE") ? atoi(getenv("CME_RT_PROBE")) : 0;
    return m;
}

int issue_step(CmeMlpRuntime* rt, int64_t rows) {
    if (probe_mask() & 2) {
        return cme_mlp_step_ex(&rt->plan, nullptr, nullptr, rows, 0.f, 0.f, 0.f, nullptr, rt->compute, rt->side,
                               reinterpret_cast<void* const*>(rt->step_ev));
    }
    const int rc = cme_mlp_step_ex(&rt->plan, nullptr, nullptr, rows, 0.f, 0.f, 0.f, rt->loss_dev, rt->compute,
                                   rt->side, reinterpret_cast<void* const*>(rt->step_ev));
    if (rc) return rc;
    publish_loss_kernel<<<1, 1, 0, rt->compute>>>(rt->cursor, rt->loss_dev);
    CME_LAUNCH_CHECK();
    return 0;
}

int graph_for(CmeMlpRuntime* rt, int64_t rows, cudaGraphExec_t* out) {
    const CmeMlpRuntime::Key key = rows;
    auto it = rt->graphs.find(key);
    if (it != rt->graphs.end()) {
        *out = it->second;
        return 0;
    }
    CME_CUDA(cudaStreamBeginCapture(rt->compute, cudaStreamCaptureModeThreadLocal));
    const int rc = issue_step(rt, rows);
    cudaGraph_t g = nullptr;
    const cudaError_t e = cudaStreamEndCapture(rt->compute, &g);
    if (rc != 0) {
        if (g) cudaGraphDestroy(g);
        return rc;
    }
    CME_CUDA(e);
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, g, 0);
    cudaGraphDestroy(g);
    CME_CUDA(ei);
    rt->graphs.emplace(key, exec);
    *out = exec;
    return 0;
}

int prefetch(CmeMlpRuntime* rt, int slot, const uint8_t* host_x, const int32_t* host_labels,
             int64_t row_begin, int64_t rows) {
    const size_t D = (size_t)rt->plan.D;
    CME_CUDA(cudaStreamWaitEvent(rt->copy, rt->consumed[slot], 0));
    if (probe_mask() & 1) {
        CME_CUDA(cudaEventRecord(rt->copied[slot], rt->copy));
        CME_CUDA(cudaEventRecord(rt->copied2[slot], rt->copy2));
        return 0;
    }
    CME_CUDA(cudaMemcpyAsync(rt->x_stage[slot], host_x + (size_t)row_begin * D, (size_t)rows * D,
                             cudaMemcpyHostToDevice, rt->copy));
    CME_CUDA(cudaEventRecord(rt->copied[slot], rt->copy));
    // The labels travel on a copy stream of their own: behind the pixels on one stream, the two
    // transfers and their event made a 75 us chain per step of the 784-100-10 model (2.5 MB of
    // pixels are 49 us of PCIe time), longer than the 54 us of compute it was meant to hide under.
    CME_CUDA(cudaStreamWaitEvent(rt->copy2, rt->consumed[slot], 0));
    CME_CUDA(cudaMemcpyAsync(rt->y_stage[slot], host_labels + row_begin, (size_t)rows * sizeof(int32_t),
                             cudaMemcpyHostToDevice, rt->copy2));
    CME_CUDA(cudaEventRecord(rt->copied2[slot], rt->copy2));
    rt->h2d_bytes += (int64_t)rows * (int64_t)(D + sizeof(int32_t));
    return 0;
}

}  // namespace

extern "C" {

int cme_mlp_runtime_destroy(CmeMlpRuntime* rt) {
    if (!rt) return 0;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(rt->device);
    if (rt->compute) cudaStreamSynchronize(rt->compute);
    if (rt->copy) cudaStreamSynchronize(rt->copy);
    if (rt->copy2) cudaStreamSynchronize(rt->copy2);
    if (rt->side) cudaStreamSynchronize(rt->side);
    for (auto& e : rt->step_ev)
        if (e) cudaEventDestroy(e);
    for (auto& kv : rt->graphs) cudaGraphExecDestroy(kv.second);
    for (int s = 0; s < 2; ++s) {
        if (rt->copied[s]) cudaEventDestroy(rt->copied[s]);
        if (rt->copied2[s]) cudaEventDestroy(rt->copied2[s]);
        if (rt->consumed[s]) cudaEventDestroy(rt->consumed[s]);
        if (rt->x_stage[s]) cudaFree(rt->x_stage[s]);
        if (rt->y_stage[s]) cudaFree(rt->y_stage[s]);
    }
    if (rt->join_in) cudaEventDestroy(rt->join_in);
    if (rt->join_out) cudaEventDestroy(rt->join_out);
    if (rt->loss_dev) cudaFree(rt->loss_dev);
    if (rt->step_args) cudaFree(rt->step_args);
    if (rt->cursor) cudaFree(rt->cursor);
    if (rt->copy) cudaStreamDestroy(rt->copy);
    if (rt->copy2) cudaStreamDestroy(rt->copy2);
    if (rt->compute) cudaStreamDestroy(rt->compute);
    if (rt->side) cudaStreamDestroy(rt->side);
    cudaSetDevice(prev);
    delete rt;
    return 0;
}

int cme_mlp_runtime_create(const CmeMlpPlan* plan, int use_graphs, CmeMlpRuntime** out) {
    CME_CHECK_ARG(plan && out, "cme_mlp_runtime_create: null argument");
    CME_CHECK_ARG(plan->max_rows > 0 && plan->D > 0, "cme_mlp_runtime_create: empty plan");
    CmeMlpRuntime* rt = new (std::nothrow) CmeMlpRuntime();
    CME_CHECK_ARG(rt, "cme_mlp_runtime_create: out of host memory");
    rt->plan = *plan;
    rt->use_graphs = use_graphs;
    int rc = 0;
    // everything below returns through the macros; tidy up on the way out
    auto build = [&]() -> int {
        CME_CUDA(cudaGetDevice(&rt->device));
        CME_CUDA(cudaStreamCreateWithFlags(&rt->copy, cudaStreamNonBlocking));
        CME_CUDA(cudaStreamCreateWithFlags(&rt->copy2, cudaStreamNonBlocking));
        CME_CUDA(cudaStreamCreateWithFlags(&rt->compute, cudaStreamNonBlocking));
        CME_CUDA(cudaStreamCreateWithFlags(&rt->side, cudaStreamNonBlocking));
        for (auto& e : rt->step_ev) CME_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (int s = 0; s < 2; ++s) {
            CME_CUDA(cudaEventCreateWithFlags(&rt->copied[s], cudaEventDisableTiming));
            CME_CUDA(cudaEventCreateWithFlags(&rt->copied2[s], cudaEventDisableTiming));
            CME_CUDA(cudaEventCreateWithFlags(&rt->consumed[s], cudaEventDisableTiming));
            CME_CUDA(cudaMalloc(&rt->x_stage[s], (size_t)plan->max_rows * plan->D + 16));
            CME_CUDA(cudaMalloc(&rt->y_stage[s], (size_t)plan->max_rows * sizeof(int32_t)));
        }
        CME_CUDA(cudaEventCreateWithFlags(&rt->join_in, cudaEventDisableTiming));
        CME_CUDA(cudaEventCreateWithFlags(&rt->join_out, cudaEventDisableTiming));
        CME_CUDA(cudaMalloc(&rt->loss_dev, sizeof(float)));
        CME_CUDA(cudaMemset(rt->loss_dev, 0, sizeof(float)));
        if (!rt->plan.step_args) {
            CME_CUDA(cudaMalloc(&rt->step_args, sizeof(CmeStepArgs)));
            rt->plan.step_args = rt->step_args;
        }
        CME_CUDA(cudaMalloc(&rt->cursor, sizeof(LossCursor)));
        return 0;
    };
    rc = build();
    if (rc != 0) {
        cme_mlp_runtime_destroy(rt);
        return rc;
    }
    *out = rt;
    return 0;
}

int cme_mlp_runtime_run(CmeMlpRuntime* rt, const uint8_t* host_x, const int32_t* host_labels,
                        const int64_t* row_begin, const int64_t* rows, const int64_t* global_rows,
                        int nsteps, float lr, float reg, float* loss_host, void* stream) {
    CME_CHECK_ARG(rt && host_x && host_labels && row_begin && rows && global_rows,
                  "cme_mlp_runtime_run: null argument");
    CME_CHECK_ARG(nsteps >= 0, "cme_mlp_runtime_run: nsteps %d", nsteps);
    for (int i = 0; i < nsteps; ++i)
        CME_CHECK_ARG(rows[i] > 0 && rows[i] <= rt->plan.max_rows && row_begin[i] >= 0 &&
                          global_rows[i] >= rows[i],
                      "cme_mlp_runtime_run: step %d has rows %lld (max %lld), begin %lld, global %lld",
                      i, (long long)rows[i], (long long)rt->plan.max_rows, (long long)row_begin[i],
                      (long long)global_rows[i]);
    if (nsteps == 0) return 0;
    cudaStream_t user = cme::as_stream(stream);
    int rc;
    // join: nothing of ours starts before the caller's stream reaches this point
    CME_CUDA(cudaEventRecord(rt->join_in, user));
    CME_CUDA(cudaStreamWaitEvent(rt->compute, rt->join_in, 0));
    CME_CUDA(cudaStreamWaitEvent(rt->copy, rt->join_in, 0));
    CME_CUDA(cudaStreamWaitEvent(rt->copy2, rt->join_in, 0));
    CME_CUDA(cudaMemsetAsync(rt->loss_dev, 0, sizeof(float), rt->compute));      // once per run; every step leaves it at zero
    float* loss_mapped = nullptr;
    if (loss_host) {
        // pinned memory is mapped under unified addressing; ask rather than assume
        CME_CUDA(cudaHostGetDevicePointer((void**)&loss_mapped, loss_host, 0));
    }
    set_cursor_kernel<<<1, 1, 0, rt->compute>>>(rt->cursor, loss_mapped);
    CME_LAUNCH_CHECK();
    rt->launches += 1;
    for (int s = 0; s < 2; ++s) CME_CUDA(cudaEventRecord(rt->consumed[s], rt->compute));
    // The loop proper.  Whatever goes wrong inside it, the steps already queued keep reading
    // host_x / host_labels: the caller's stream (and the copy stream) must still be made to wait
    // for them before the error is returned, or a caller that frees its arrays races with copies
    // in flight.
    auto queue_steps = [&]() -> int {
        if ((rc = prefetch(rt, 0, host_x, host_labels, row_begin[0], rows[0]))) return rc;
        for (int i = 0; i < nsteps; ++i) {
            const int slot = i & 1;
            if (i + 1 < nsteps)
                if ((rc = prefetch(rt, slot ^ 1, host_x, host_labels, row_begin[i + 1], rows[i + 1]))) return rc;
            CME_CUDA(cudaStreamWaitEvent(rt->compute, rt->copied[slot], 0));
            CME_CUDA(cudaStreamWaitEvent(rt->compute, rt->copied2[slot], 0));
            const float scale = 1.0f / (float)global_rows[i];
            if ((rc = cme_mlp_set_step_args(&rt->plan, rt->x_stage[slot], rt->y_stage[slot], lr, reg, scale,
                                            rt->compute))) return rc;
            if (rt->use_graphs) {
                cudaGraphExec_t exec = nullptr;
                if ((rc = graph_for(rt, rows[i], &exec))) return rc;
                CME_CUDA(cudaGraphLaunch(exec, rt->compute));
            } else {
                if ((rc = issue_step(rt, rows[i]))) return rc;
            }
            CME_CUDA(cudaEventRecord(rt->consumed[slot], rt->compute));
            if (loss_host) rt->d2h_bytes += sizeof(float);
            rt->launches += kernels_per_step(rt->plan);
            rt->steps += 1;
        }
        return 0;
    };
    const int run_rc = queue_steps();
    if (run_rc != 0) {
        // best effort, keeping the first error message: drain what was queued
        cudaStreamSynchronize(rt->copy);
        cudaStreamSynchronize(rt->copy2);
        cudaStreamSynchronize(rt->compute);
        cudaStreamSynchronize(rt->side);
        return run_rc;
    }
    CME_CUDA(cudaEventRecord(rt->join_out, rt->compute));
    CME_CUDA(cudaStreamWaitEvent(user, rt->join_out, 0));
    return 0;
}

int cme_mlp_runtime_stats(const CmeMlpRuntime* rt, int64_t out[5]) {
    CME_CHECK_ARG(rt && out, "cme_mlp_runtime_stats: null argument");
    out[0] = rt->h2d_bytes;
    out[1] = rt->d2h_bytes;
    out[2] = rt->launches;
    out[3] = rt->steps;
    out[4] = (int64_t)rt->graphs.size();
    return 0;
}

}  // extern "C"
