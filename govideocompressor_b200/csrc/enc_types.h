// enc_types.h -- plain structs shared by the encoder-side kernels (device), their host launchers and the
// host emulation harness of the CPU tests.
//
// Frame layout in HBM ("padded planes"): every luma plane of a frame (source, reconstruction and the three
// half-sample planes H, V, C) is stored with kEncPad replicated samples on all four sides of the CODED
// picture (w16 x h16, multiples of 16), rows `sy` bytes apart; pointers address sample (0,0).  Chroma
// planes carry kEncPad/2 samples of padding, rows `sc` bytes apart.  Motion vectors are limited so that
// every referenced sample lies inside the padded planes (x264 pads by 32; 64 covers merange 32 + sub-pel).
#pragma once
#include <stdint.h>

namespace gvc {

constexpr int kEncPad = 64;

struct EncGeom {
    int w, h;        // picture size
    int w16, h16;    // coded size
    int mbw, mbh;    // macroblocks
    int sy, sc;      // luma / chroma row pitch in bytes
    int sl;          // lowres (half resolution luma) row pitch; lowres planes carry kEncPad/2 padding
};

// One reference frame: full-sample luma + half-sample planes, chroma, lowres luma.
struct RefPlanes {
    const uint8_t *y, *hh, *hv, *hc;   // luma: full, H (x+1/2), V (y+1/2), C (both)
    const uint8_t *u, *v;
    const uint8_t *low;                // half-resolution luma (phase 0)
};

// Quantiser constants of one frame (x264 encoder/set.c semantics, flat matrices), by coefficient class
// (class = (row&1)+(col&1) inside a 4x4 block).
struct QuantParams {
    int qp, qpc;                 // luma / chroma QP
    int mf_y[3], bias_y[3];      // luma forward multipliers / dead-zone bias
    int mf_c[3], bias_c[3];      // chroma
    int dq_y[3], dq_c[3];        // LevelScale (dequant multipliers, 16*normAdjust)
    int lambda;                  // motion-cost lambda (x264_lambda_tab[qp])
    // smallest |coefficient| that quantises to a nonzero level, per class: (bias + |c|) * mf >= 65536; cdc = chroma DC (mf >> 1, bias << 1)
    int thr_y[3], thr_c[3], thr_cdc;
};

// 8x8 transform (High-profile transform_size_8x8): quantiser constants by coefficient class
// (class = table[(row&3)*4 + (col&3)], H.264 8.5.9 / x264 quant8_scan).
struct Quant8Params {
    int qp;
    int mf[6], bias[6], dq[6];
};

// Per-macroblock side information produced by motion search / mode decision.
struct MbMotion {
    int16_t mv[4][2];   // quarter-sample vectors of the four 8x8 partitions (all equal for 16x16)
    int32_t cost;       // cost of the chosen mode
    uint8_t mode;       // 0 = P_L0_16x16, 1 = P_8x8 (four 8x8 partitions)
    uint8_t pad[3];
};

constexpr int kMbCoefs = 26 * 16;

// Segments are independent (server.go:176-190 hands out whole pieces), so up to kMaxBatch of them advance picture by picture
// together: every stage is ONE launch whose grid carries the segment index, and the kernel picks that segment's argument
// record out of a Batch<> kernel parameter (parameter space: up to 32 KB on sm_70+ with CUDA >= 12.1).
constexpr int kMaxBatch = 16;
template <class A> struct Batch { A a[kMaxBatch]; };
   // int16 per macroblock: see mb_coef_layout in enc_kernels.cuh

}  // namespace gvc
