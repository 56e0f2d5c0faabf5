// Host-side helpers shared by the tensor-core translation units: TMA tensor maps and the
// opt-in for large dynamic shared memory.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

namespace tc_host {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() { return reinterpret_cast<EncodeTiledFn>(cme::tensor_map_encoder()); }

// 2-D bf16 tensor map: dims {inner, outer}, row pitch ld elements, 128-byte swizzle
static inline int make_map(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint64_t ld,
             uint32_t box_inner, uint32_t box_outer) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        cme::set_error("cuTensorMapEncodeTiled is not available from this driver");
        return -2;
    }
    if (((uintptr_t)base & 15) || ((ld * 2) & 15)) {
        cme::set_error("tensor map: base and row pitch must be 16-byte aligned (ld=%llu)",
                       (unsigned long long)ld);
        return -1;
    }
    cuuint64_t dims[2] = {inner, outer};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {box_inner, box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims,
                    strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        cme::set_error("cuTensorMapEncodeTiled failed with CUresult %d (inner=%llu outer=%llu ld=%llu "
                       "box=%ux%u)", (int)r, (unsigned long long)inner, (unsigned long long)outer,
                       (unsigned long long)ld, box_inner, box_outer);
        return -3;
    }
    return 0;
}

template <typename K>
static inline int opt_in_smem(K kernel, int bytes) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, bool> done;  // per (kernel, device)
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    auto key = std::make_pair((const void*)kernel, dev);
    if (!done.count(key)) {
        CME_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        done[key] = true;
    }
    return 0;
}

}  // namespace tc_host
