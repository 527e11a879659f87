// Internal helpers shared by the fod_b200 translation units (error plumbing only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/fod_b200.h"

namespace fod {
int fod_fail(int code, const char* msg);          // records msg for fod_last_error(), returns code
int fod_check_launch(const char* kernel_name);    // cudaGetLastError -> FOD_OK / FOD_E_LAUNCH
}  // namespace fod
