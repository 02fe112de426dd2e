#include "common.cuh"
extern "C" int vb_abi_version(void) { return 1; }
extern "C" int vb_compiled_sm(void) { return 100; }
extern "C" const char* vb_error_string(int status) {
    switch (status) {
        case VB_OK: return "ok";
        case VB_ERR_ARG: return "invalid argument";
        case VB_ERR_UNSUPPORTED: return "unsupported configuration";
        case VB_ERR_WORKSPACE: return "workspace too small";
        case VB_ERR_ALIGN: return "misaligned pointer";
        case VB_ERR_TOO_LARGE: return "problem larger than the fixed scratch of this entry point allows";
        default: return status > 0 ? cudaGetErrorString((cudaError_t)status) : "unknown error";
    }
}
