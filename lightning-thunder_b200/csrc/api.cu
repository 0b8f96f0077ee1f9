// Library-level entry points: version, error strings, device queries.
#include "common.cuh"

extern "C" int b200k_version(void) { return 6; }

extern "C" int b200k_sm_count(void) { return b200k::sm_count(); }

extern "C" const char* b200k_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case B200K_EINVAL: return "b200k: invalid argument (null pointer, negative size or bad flag)";
        case B200K_EDTYPE: return "b200k: unsupported dtype code";
        case B200K_EALIGN: return "b200k: pointer or row pitch is not 16-byte aligned";
        case B200K_ESCRATCH: return "b200k: scratch buffer missing, misaligned or too small";
        case B200K_ELIMIT: return "b200k: size exceeds a documented maximum";
        default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "b200k: unknown error code";
}
