// version / error strings of the C ABI
#include "common.cuh"

extern "C" int dppo_version(void) { return 100; }   // 0.1.0

extern "C" const char* dppo_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case DPPO_E_BADARG: return "dppo: bad argument (null pointer, negative size or unknown flavour)";
        case DPPO_E_TOOLARGE: return "dppo: dimension larger than the kernels support";
        case DPPO_E_SCRATCH: return "dppo: scratch buffer too small (see *_scratch_bytes)";
        case DPPO_E_ALIGN: return "dppo: pointer not 16-byte aligned";
        case DPPO_E_COMM: return "dppo: communicator misuse";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "dppo: unknown error";
}
