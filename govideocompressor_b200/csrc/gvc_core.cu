// gvc_core.cu -- library life cycle and error state of libgvc.so.
#include <cstdlib>
#include <cstring>

#include "gvc_common.h"

namespace gvc {

static thread_local char t_err[512] = "";
static int               g_device   = -1;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

int  current_device() { return g_device; }
bool initialised() { return g_device >= 0; }

int enter() {
    if (g_device < 0) {
        set_error("gvc_init() has not been called");
        return GVC_ERR_NOT_INIT;
    }
    GVC_CUDA(cudaSetDevice(g_device));
    return GVC_OK;
}

}  // namespace gvc

extern "C" {

int gvc_init(int device_ordinal) {
    using namespace gvc;
    if (device_ordinal < 0) {
        const char *e  = getenv("GVC_DEVICE");
        device_ordinal = e && *e ? atoi(e) : 0;
    }
    int         n  = 0;
    cudaError_t ce = cudaGetDeviceCount(&n);
    if (ce != cudaSuccess || n == 0) {
        set_error("no CUDA device: %s", ce == cudaSuccess ? "device count is 0" : cudaGetErrorString(ce));
        return GVC_ERR_NO_DEVICE;
    }
    if (device_ordinal >= n) {
        set_error("device %d requested but only %d visible", device_ordinal, n);
        return GVC_ERR_INVALID_ARG;
    }
    GVC_CUDA(cudaSetDevice(device_ordinal));
    cudaDeviceProp prop;
    GVC_CUDA(cudaGetDeviceProperties(&prop, device_ordinal));
    if (prop.major != 10) {  // the fatbin holds sm_100a SASS only
        set_error("device %d is sm_%d%d; libgvc is built for sm_100a (B200) only", device_ordinal, prop.major, prop.minor);
        return GVC_ERR_UNSUPPORTED;
    }
    GVC_CUDA(cudaFree(0));
    // start-up is the time to reserve fixed device memory: the built-in scratch of gvc_resize_frames_device_static()
    int rc = resize_static_prepare(device_ordinal);
    if (rc) return rc;
    g_device = device_ordinal;
    return GVC_OK;
}

void gvc_shutdown(void) { gvc::g_device = -1; }

const char *gvc_strerror(int status) {
    switch (status) {
        case GVC_OK: return "ok";
        case GVC_ERR_INVALID_ARG: return "invalid argument";
        case GVC_ERR_NO_DEVICE: return "no CUDA device";
        case GVC_ERR_CUDA: return "CUDA error";
        case GVC_ERR_NOMEM: return "out of memory";
        case GVC_ERR_UNSUPPORTED: return "unsupported request (no CPU fallback)";
        case GVC_ERR_IO: return "I/O error";
        case GVC_ERR_NOT_INIT: return "library not initialised";
        case GVC_ERR_TOO_LARGE: return "size exceeds a fixed limit";
        default: return "unknown status";
    }
}

const char *gvc_last_error(void) { return gvc::t_err; }
void        gvc_free(void *p) { free(p); }
const char *gvc_version(void) { return "gvc-b200 0.1 sm_100a"; }

}  // extern "C"
