// Error plumbing + version for the fod_b200 C ABI.
#include <cstdlib>
#include <stdio.h>
#include <string.h>
#include "fod_internal.h"

namespace fod {
static thread_local char g_err[512] = "";
int fod_fail(int code, const char* msg) {
    snprintf(g_err, sizeof(g_err), "%s", msg);
    return code;
}
int fod_check_launch(const char* kernel_name) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(g_err, sizeof(g_err), "%s: %s", kernel_name, cudaGetErrorString(e));
        return FOD_E_LAUNCH;
    }
    return FOD_OK;
}
bool fod_pdl_enabled() {
    static const int on = [] {
        const char* e = getenv("FOD_PDL");
        return e ? atoi(e) : 1;
    }();
    return on != 0;
}
int fod_env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
}  // namespace fod

extern "C" int fod_version(void) { return 13; }
extern "C" const char* fod_last_error(void) { return fod::g_err; }
