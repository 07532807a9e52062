// resize_types.h -- plain structs shared by the resize kernels (device), the plan builder (host C++)
// and the host emulation harness used by the CPU tests.
#pragma once
#include <stdint.h>

namespace gvc {

constexpr int kTileW      = 128;  // output columns per tile (32 lanes x 4 pixels)
constexpr int kThreads    = 256;
constexpr int kWarps      = kThreads / 32;
constexpr int kRowGroups  = 2;    // source rows per tile <= 64 (lanes <-> rows, 2 groups)
constexpr int kTileRows   = kRowGroups * 32;
constexpr int kTPad       = kTileRows / 2 + 1;  // row PAIRS per quad of the intermediate tile, odd (bank spread)
constexpr int kHEntry     = 8;    // int32 words per horizontal table entry: q, r8, 6 packed tap pairs
constexpr int kMaxHGroups = 3;    // <= 12 horizontal taps in the fused kernel
constexpr int kMaxVTaps   = 15;   // vertical taps in the fused kernel
constexpr int kTSlack     = 16;   // int4 of slack behind the intermediate tile (zero-padded taps may read past a tile's rows)
constexpr int kVEntry     = 12;   // int32 words per vertical table entry: first row (even), 4 + 4 packed coefficient words, end row, pad
constexpr int kMaxVPairs  = (kMaxVTaps + 2) / 2;   // row pairs an output row may touch

struct PlaneClassDev {            // geometry + tables shared by all planes of one class (luma / chroma)
    int sw, sh, dw, dh;
    int hgroups;                  // 4-tap groups actually used (1..3)
    int vtaps;                    // vertical taps
    int vq;                       // int4 per vertical table entry (kVEntry / 4)
    int th;                       // output rows per tile
    int tiles_x, tiles_y;
    int tiles_x_magic;            // t / tiles_x == (t * magic) >> 20 for every tile index of the plane (0: divide)
    int s_chunks;                 // odd; 16-byte chunks per source-tile row in shared memory
    int s_chunks_magic;           // ceil(2^32 / s_chunks): i / s_chunks == umulhi(i, magic) for the small i used
    int vt;                       // row pairs the kernel is instantiated for: every output row's taps lie in that many pairs from its first (even) row
    const int32_t *htab;          // [dw][kHEntry]
    const int32_t *vtab;          // [dh][kVEntry]: {first row & ~1, lo0, hi0, lo1 | hi1, lo2, hi2, lo3 | hi3, first row + taps, 0, 0}; lo_k / hi_k hold the
                                  // low (unsigned) / high (signed) bytes of the coefficients of rows 4k .. 4k+3 counted from the first row & ~1
};

struct PlaneJob {
    const uint8_t *src; int sstride; int spix; unsigned long long sframe;
    uint8_t       *dst; int dstride; int dpix; unsigned long long dframe;
    int cls;                      // index into ResizeParams::cls
    int tile_base;                // first tile id of this plane within a frame
};

struct ResizeParams {
    PlaneClassDev cls[2];
    PlaneJob      job[3];
    int           njobs;
    int           tiles_per_frame;
};

struct GenericClassDev {
    int sw, sh, dw, dh, htaps, vtaps;
    const int16_t *hc; const int32_t *hp;   // [dw][htaps], [dw]
    const int16_t *vc; const int32_t *vp;   // [dh][vtaps], [dh]
};

}  // namespace gvc
