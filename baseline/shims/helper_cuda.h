// Stand-in for the CUDA samples header the reference includes (fpcode/inc/gpu_func.h:5):
// only checkCudaErrors is used.
#ifndef CME213B_HELPER_CUDA_SHIM_
#define CME213B_HELPER_CUDA_SHIM_
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <iostream>   // the real helper_cuda.h drags this in; gpu_func.h relies on it
#define checkCudaErrors(call)                                                          \
    do {                                                                               \
        cudaError_t e__ = (call);                                                      \
        if (e__ != cudaSuccess) {                                                      \
            fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, __LINE__); \
            exit(1);                                                                   \
        }                                                                              \
    } while (0)
// gpu_func.cu calls cudaDeviceSetSharedMemConfig(cudaSharedMemBankSizeEightByte) before every
// GEMM launch (ref fpcode/gpu_func.cu:269,280,291).  The bank size has not been configurable since
// Kepler and the sm_100 driver answers cudaErrorInvalidValue, which the reference's own
// check_launch() then reports as a failed kernel and exits.  Neutralise the call here (it is a
// no-op on every GPU that can run this build) so the unmodified source runs on B200.
#define cudaDeviceSetSharedMemConfig(cfg) ((void)(cfg), cudaSuccess)
#endif
