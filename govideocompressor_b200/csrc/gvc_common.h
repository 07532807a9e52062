// gvc_common.h -- shared host-side helpers of libgvc.so (error state, CUDA call checking).
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/gvc.h"

namespace gvc {

void set_error(const char *fmt, ...);
int  current_device();   // -1 until gvc_init succeeded
bool initialised();

// Makes the library's device current on the calling thread (cgo calls arrive on arbitrary OS threads,
// SURVEY.md 8b) and returns GVC_OK or an error status.
int enter();

// gvc_resize.cu: resolves the device address of the built-in scratch array of gvc_resize_frames_device_static() on
// device `dev` (which must be current).  gvc_init() calls it so the module load and the reservation happen at start-up.
int resize_static_prepare(int dev);

#define GVC_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            ::gvc::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            return e__ == cudaErrorMemoryAllocation ? GVC_ERR_NOMEM : GVC_ERR_CUDA;                 \
        }                                                                                           \
    } while (0)

}  // namespace gvc
