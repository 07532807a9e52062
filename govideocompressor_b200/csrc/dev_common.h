// dev_common.h -- definitions shared by all kernel headers.  With GVC_EMU defined the same headers compile as
// plain C++ so the CPU tests can execute kernel bodies thread by thread (tests/emu, a test harness only).
#pragma once
#include <stdint.h>

#if defined(GVC_EMU)
#define GVC_HD inline
#define GVC_HDH inline
struct gvc_int4 { int x, y, z, w; };
struct gvc_uint4 { unsigned x, y, z, w; };
#else
#define GVC_HD __device__ __forceinline__
#define GVC_HDH __host__ __device__ __forceinline__   // also callable from host launch code
typedef int4  gvc_int4;
typedef uint4 gvc_uint4;
#endif

// After a lane-dependent if / else the compiler is free to keep the two groups of lanes apart (tail duplication) and run
// the code that follows once per group.  GVC_RECONVERGE(mask) names the lanes that must be back together there
// (ncu: the Intra16x16 kernel executed 1.8x the instructions without it).
#if defined(GVC_EMU)
#define GVC_RECONVERGE(mask) ((void)0)
#else
#define GVC_RECONVERGE(mask) __syncwarp(mask)
#endif

namespace gvc {
GVC_HD int clip_u8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }
GVC_HD int imin(int a, int b) { return a < b ? a : b; }
GVC_HD int imax(int a, int b) { return a > b ? a : b; }
GVC_HD int iclamp(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
GVC_HD int iabs(int v) { return v < 0 ? -v : v; }
GVC_HD unsigned pack_sat_u8x4(int v0, int v1, int v2, int v3) {   // bytes clip_u8(v0) | clip_u8(v1) << 8 | clip_u8(v2) << 16 | clip_u8(v3) << 24
#if defined(GVC_EMU)
    return (unsigned)clip_u8(v0) | ((unsigned)clip_u8(v1) << 8) | ((unsigned)clip_u8(v2) << 16) | ((unsigned)clip_u8(v3) << 24);
#else
    unsigned hi, d;                                      // cvt.pack: d = c << 16 | sat(a) << 8 | sat(b)
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(v3), "r"(v2), "r"(0));
    asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(v1), "r"(v0), "r"(hi));
    return d;
#endif
}

}  // namespace gvc
