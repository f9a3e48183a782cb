// Library identity and error strings for the C ABI (include/ia_b200.h).
#include <cuda_runtime.h>

#include "../../include/ia_b200.h"

extern "C" int ia_version(void) { return IA_ABI_VERSION; }

extern "C" const char *ia_error_string(int code) {
    switch (code) {
    case 0: return "ok";
    case IA_ERR_NULL: return "a required pointer is NULL";
    case IA_ERR_SHAPE: return "a size argument is out of range";
    case IA_ERR_DTYPE: return "unknown or unsupported dtype / variant / mode";
    case IA_ERR_ALIGN: return "a pointer is not 16-byte aligned";
    case IA_ERR_WORKSPACE: return "workspace too small (see ia_meandiff_workspace_bytes)";
    case IA_ERR_TOO_LARGE: return "size beyond the documented maximum";
    default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "unknown error";
}
