// Host-side TMA descriptor (CUtensorMap) construction through the driver entry point, so the
// library links only against cudart. Descriptors are cheap (~1 us) and built per launch.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace fod {

typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                        CUtensorMapFloatOOBfill);

inline PFN_tmapEncodeTiled get_tmap_encode() {
    static PFN_tmapEncodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
            fprintf(stderr, "fod_b200: cuTensorMapEncodeTiled entry point unavailable\n");
            return nullptr;
        }
        fn = reinterpret_cast<PFN_tmapEncodeTiled>(p);
    }
    return fn;
}

// bf16 tensor of rank 2 or 3; dims[0] is the contiguous dimension. strides_bytes[i] is the byte stride of
// dims[i+1]. swizzle_bytes in {0, 32, 64, 128}; box[0]*2 must equal swizzle_bytes (or be <= it).
inline int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                          const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
    PFN_tmapEncodeTiled enc = get_tmap_encode();
    if (!enc) return 1;
    cuuint64_t gdim[5];
    cuuint64_t gstr[5];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        es[i] = 1;
    }
    for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
    CUtensorMapSwizzle sw = swizzle_bytes == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                            : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                            : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                  : CU_TENSOR_MAP_SWIZZLE_NONE;
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr,
                     bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "fod_b200: cuTensorMapEncodeTiled failed (%d) rank=%d dims=%llu,%llu box=%u,%u\n", (int)r,
                rank, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1]);
        return 2;
    }
    return 0;
}

}  // namespace fod
