// Internal helpers shared by the fod_b200 translation units (error plumbing only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/fod_b200.h"

namespace fod {
int fod_fail(int code, const char* msg);          // records msg for fod_last_error(), returns code
int fod_check_launch(const char* kernel_name);    // cudaGetLastError -> FOD_OK / FOD_E_LAUNCH
bool fod_pdl_enabled();
int fod_env_int(const char* name, int dflt);         // integer environment switch, read on every call (cheap)                           // env FOD_PDL (default on): programmatic dependent launch of the GEMMs

// Kernel launch with the programmatic-stream-serialization attribute (when FOD_PDL is on): the kernel may start its
// prologue while the previous kernel in the stream drains; it must execute griddepcontrol.wait (pdl_wait()) before its
// first global-memory access.
template <typename... KArgs, typename... Args>
inline void fod_launch(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = fod_pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// attn_bwd_fused_sm100.cu: the one-kernel attention backward (dQ resident in TMEM); arguments validated by the caller
bool attn_bwd_fused_supported(const fod_attn_bwd_args* g);
int launch_attn_bwd_fused(const fod_attn_bwd_args* g, cudaStream_t st);
}  // namespace fod
