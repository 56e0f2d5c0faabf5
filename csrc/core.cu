// Library plumbing: error string, device facts.
#include "common.cuh"

#include <mutex>

namespace cme {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int sm_count() {
    // per device: a process may drive several GPUs in tests
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

void* tensor_map_encoder() {
    static void* fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = p;
    });
    return fn;
}

}  // namespace cme

namespace {
__global__ void add_one_kernel(int* v) { *v += 1; }
}  // namespace

extern "C" {

int cme_warmup(int* value_inout) {
    CME_CHECK_ARG(value_inout, "cme_warmup: null argument");
    int* d = nullptr;
    CME_CUDA(cudaMalloc(&d, sizeof(int)));
    cudaError_t e = cudaMemcpy(d, value_inout, sizeof(int), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        add_one_kernel<<<1, 1>>>(d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(value_inout, d, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(d);
    CME_CUDA(e);
    return 0;
}

const char* cme_last_error(void) { return cme::g_err; }

int cme_version(void) { return 100; }

int cme_device_info(int* sms, int* cc_major, int* cc_minor) {
    int dev = 0;
    CME_CUDA(cudaGetDevice(&dev));
    if (sms) CME_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
    if (cc_major) CME_CUDA(cudaDeviceGetAttribute(cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_minor) CME_CUDA(cudaDeviceGetAttribute(cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
    return 0;
}

/* L2 fetch granularity of the current device (cudaLimitMaxL2FetchGranularity; 32, 64 or 128 bytes;
 * the driver's default is 64).  Random 4-byte gathers (PageRank at 64 M nodes: ncu shows 3 DRAM
 * sectors per edge at the default) pay for whole fetches; 32 halves that.  A hint, device-wide:
 * returns the previous value in *old_bytes so the caller can restore it. */
int cme_l2_fetch_granularity(int bytes, int* old_bytes) {
    size_t cur = 0;
    CME_CUDA(cudaDeviceGetLimit(&cur, cudaLimitMaxL2FetchGranularity));
    if (old_bytes) *old_bytes = (int)cur;
    if (bytes > 0) {
        CME_CHECK_ARG(bytes == 32 || bytes == 64 || bytes == 128, "cme_l2_fetch_granularity: 32, 64 or 128 (got %d)", bytes);
        CME_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes));
    }
    return 0;
}

}  // extern "C"
