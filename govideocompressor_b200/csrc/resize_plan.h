// resize_plan.h -- host-side construction of the resize kernels' tables (pure C++, no CUDA calls).
// gvc_resize.cu uploads the result; tests/emu runs the kernel phases against it on the CPU.
#pragma once
#include <cstddef>
#include <vector>

#include "filter_bank.h"
#include "resize_types.h"

namespace gvc {

struct ResizeTables {
    int sw = 0, sh = 0, dw = 0, dh = 0, algo = 0;
    FilterBank hb[2], vb[2];                 // [0] luma, [1] chroma
    GenericClassDev gcls[2];                 // table pointers left null
    bool   fused_ok = false;
    int    ng = 1, nt = 1, vq = 1, vt = 1, s_chunks_max = 1;   // ng: groups of 4 horizontal taps (table layout); nt: tap pairs (kernel template)
    size_t smem_bytes = 0;
    PlaneClassDev cls[2];                    // table pointers left null
    std::vector<int32_t> htab[2], vtab[2];   // fused-kernel tables (see resize_types.h)
};

inline int chroma_dim(int v) { return (v + 1) >> 1; }   // AV_CEIL_RSHIFT(v, 1)

bool build_resize_tables(ResizeTables &t, int sw, int sh, int dw, int dh, int algo);

}  // namespace gvc
