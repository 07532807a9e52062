// resize_kernels.cuh -- K1: swscale-equivalent separable resize of 8-bit 4:2:0 frames (sm_100a).
//
// Replaces libswscale as reached from client.go:115 (`ffmpeg ... -s WxH`).  Arithmetic contract
// (bit-exact with swscale's C path: hScale8To15_c + yuv2planeX_8_c):
//     t[r][x] = min( (sum_j src[r][hpos[x]+j] * hc[x][j]) >> 7, 32767 )
//     d[y][x] = clip_u8( ((64<<12) + sum_j t[vpos[y]+j][x] * vc[y][j]) >> 19 )
//
// Two device paths:
//   * fused (resize_fused_kernel): one CTA = one 128 x TH output tile of one plane of one frame.
//     phase 0  source tile -> shared memory: ONE TMA box per tile (cp.async.bulk.tensor.3d, completion on an
//              mbarrier); plain loads for planes TMA cannot address
//     phase 1  horizontal pass, lanes <-> output columns, each lane walks down the tile's rows with its taps in
//              registers; pixels are consumed as packed words through dp2a (2 taps / instruction)
//     phase 2  vertical pass from the int32 intermediate tile in shared memory, 4 pixels per lane,
//              32-bit stores.  No intermediate ever touches HBM: traffic = source + destination.
//   * two-pass (resize_hpass_kernel / resize_vpass_kernel): any tap count / ratio, int16 intermediate
//     plane in device scratch.  Used when a plan does not meet the fused kernel's limits.
//
// The phase functions are plain C++ over (tid, shared buffers) so tests can run them on the host
// (tests/emu/emu_resize.cpp, compiled with GVC_EMU) to check indexing before a GPU is involved.
#pragma once
#include <stdint.h>

#include "dev_common.h"
#include "resize_types.h"

namespace gvc {

// ---------------------------------------------------------------- small intrinsics ------------
GVC_HD int dp2a_lo_su(int coef_pair, unsigned pix, int acc) {   // acc + c.lo*pix.b0 + c.hi*pix.b1
#if defined(GVC_EMU)
    return acc + (int)(short)(coef_pair & 0xffff) * (int)(pix & 0xff) + (int)(short)(coef_pair >> 16) * (int)((pix >> 8) & 0xff);
#else
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef_pair), "r"(pix), "r"(acc));
    return d;
#endif
}
GVC_HD int dp2a_hi_su(int coef_pair, unsigned pix, int acc) {   // acc + c.lo*pix.b2 + c.hi*pix.b3
#if defined(GVC_EMU)
    return acc + (int)(short)(coef_pair & 0xffff) * (int)((pix >> 16) & 0xff) + (int)(short)(coef_pair >> 16) * (int)(pix >> 24);
#else
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(coef_pair), "r"(pix), "r"(acc));
    return d;
#endif
}
GVC_HD int dp2a_ss(unsigned pair, unsigned coef, int acc, bool hi) {   // acc + pair.lo16 * coef.b[2hi] + pair.hi16 * coef.b[2hi+1], all signed
#if defined(GVC_EMU)
    const int sh = hi ? 16 : 0;
    return acc + (int)(short)(pair & 0xffff) * (int)(signed char)(coef >> sh) + (int)(short)(pair >> 16) * (int)(signed char)(coef >> (sh + 8));
#else
    int d;
    if (hi) asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(coef), "r"(acc));
    else    asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(coef), "r"(acc));
    return d;
#endif
}
GVC_HD int dp2a_su(unsigned pair, unsigned coef, int acc, bool hi) {   // the same with unsigned coefficient bytes
    return hi ? dp2a_hi_su((int)pair, coef, acc) : dp2a_lo_su((int)pair, coef, acc);
}
GVC_HD unsigned funnel_r(unsigned lo, unsigned hi, int sh) {    // bytes of hi:lo shifted right by sh (0..24) bits
#if defined(GVC_EMU)
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#else
    return __funnelshift_r(lo, hi, sh);
#endif
}

struct TileGeom {
    int x0, ncols;   // first output column, number of output columns (<= kTileW)
    int y0, nrows;   // first output row, number of output rows (<= th)
    int r0, rows;    // first source row, number of source rows needed (<= kTileRows)
    int qa0;         // first source word (4 bytes) of the shared tile, multiple of 4 (16 B chunk aligned)
    int nchunks;     // 16-byte chunks per row that must be loaded
};

GVC_HD TileGeom tile_geom(const PlaneClassDev &pc, int tx, int ty) {
    TileGeom g;
    g.x0    = tx * kTileW;
    g.ncols = imin(kTileW, pc.dw - g.x0);
    g.y0    = ty * pc.th;
    g.nrows = imin(pc.th, pc.dh - g.y0);
    g.r0    = pc.vtab[(long long)g.y0 * kVEntry];                               // even: rows are kept in (even, odd) pairs
    g.rows  = pc.vtab[(long long)(g.y0 + g.nrows - 1) * kVEntry + 9] - g.r0;
    g.qa0   = pc.htab[(long long)g.x0 * kHEntry] & ~3;
    const int qlast = pc.htab[(long long)(g.x0 + g.ncols - 1) * kHEntry] + pc.hgroups;  // last word touched
    g.nchunks = ((qlast - g.qa0) >> 2) + 1;
    return g;
}

// ---------------------------------------------------------------- phase 0: source tile -> smem --
// S layout (row major): source row r of the tile occupies words [r * pitch_w, (r + 1) * pitch_w), byte 0 = sample 4 * qa0 of
// the plane (a multiple of 16 bytes), all kTileRows rows present.  On the GPU the tile normally arrives as ONE TMA box
// (gvc_resize.cu; rows / columns outside the plane come back as zeros); this is the same fill with plain loads, for planes
// TMA cannot address (NV12 chroma, unaligned pointers or pitches) and for the CPU emulation.  Bytes outside the plane are zero.
GVC_HD void resize_phase0(int tid, const PlaneClassDev &pc, const PlaneJob &pj, const uint8_t *src, const TileGeom &g, unsigned *S, int pitch_w) {
    const bool aligned = pj.spix == 1 && ((((unsigned long long)src) | (unsigned)pj.sstride) & 3) == 0;
    for (int i = tid; i < kTileRows * pitch_w; i += kThreads) {
        const int r = i / pitch_w, c = i - r * pitch_w, row = g.r0 + r, col = 4 * (g.qa0 + c);
        unsigned w = 0;
        if (row < pc.sh && col < pc.sw) {
            const uint8_t *p = src + (long long)row * pj.sstride;
            if (aligned && col + 4 <= pc.sw) {
                w = *reinterpret_cast<const unsigned *>(p + col);
            } else {
                for (int b = 0; b < 4; ++b)
                    if (col + b < pc.sw) w |= (unsigned)p[(long long)(col + b) * pj.spix] << (8 * b);
            }
        }
        S[i] = w;
    }
}

// ---------------------------------------------------------------- phase 1: horizontal ----------
// Lanes <-> output columns: a lane keeps its column's table entry (first source word, byte offset, packed tap pairs) in
// registers and walks down the tile's source rows -- per intermediate sample three (NG = 2) shared-memory words of the row,
// two funnel shifts to the column's byte offset, four dp2a (two taps each), shift, clamp, one store; nothing per sample is
// recomputed or reloaded.  A warp's 32 columns read at most a dozen neighbouring words of a row (no bank conflicts), and the
// stores of a warp fall into 32 different banks.  Warp w takes columns 32 * (w & 3) .. + 31 and rows 32 * (w >> 2) .. + 31.
// T layout: [quad (4 output columns)][kTPad row pairs][4] words, a word = the int16 samples of rows (2p, 2p + 1) of one column.
// NT = tap PAIRS of the plan (the larger of the luma / chroma banks): one dp2a each.
template <int NT>
GVC_HD void resize_phase1(int tid, const PlaneClassDev &pc, const TileGeom &g, const unsigned *S, int pitch_w, gvc_int4 *T) {
    const int warp = tid >> 5, lane = tid & 31;
    const int xl = (warp & 3) * 32 + lane;                 // tile-relative output column
    if ((warp & 3) * 32 >= g.ncols) return;                // warp-uniform: the tile has no such columns
    const gvc_int4 *htab = reinterpret_cast<const gvc_int4 *>(pc.htab);
    const int x = imin(g.x0 + xl, pc.dw - 1);
#if defined(GVC_EMU)
    const gvc_int4 e0 = htab[(long long)x * 2], e1 = htab[(long long)x * 2 + 1];
#else
    const gvc_int4 e0 = __ldg(htab + (long long)x * 2), e1 = __ldg(htab + (long long)x * 2 + 1);
#endif
    const int r8 = e0.y, r0 = (warp >> 2) * 32;
    const unsigned *w = S + r0 * pitch_w + (e0.x - g.qa0);
    unsigned *t = reinterpret_cast<unsigned *>(T) + ((xl >> 2) * kTPad + (r0 >> 1)) * 4 + (xl & 3);
    auto sample = [&](const unsigned *row) {
        const unsigned w0 = row[0], w1 = row[1];
        const unsigned u0 = funnel_r(w0, w1, r8);
        int a = dp2a_lo_su(e0.z, u0, 0);
        if (NT >= 2) a = dp2a_hi_su(e0.w, u0, a);
        if (NT >= 3) {
            const unsigned w2 = row[2];
            const unsigned u1 = funnel_r(w1, w2, r8);
            a = dp2a_lo_su(e1.x, u1, a);
            if (NT >= 4) a = dp2a_hi_su(e1.y, u1, a);
            if (NT >= 5) {
                const unsigned w3 = row[3];
                const unsigned u2 = funnel_r(w2, w3, r8);
                a = dp2a_lo_su(e1.z, u2, a);
                if (NT >= 6) a = dp2a_hi_su(e1.w, u2, a);
            }
        }
        return a >> 7;                                          // hScale8To15's upper clamp is part of the pack below
    };
#if !defined(GVC_EMU)
#pragma unroll 4
#endif
    for (int r = 0; r < 32; r += 2, w += 2 * pitch_w, t += 4) {
        const int t0 = sample(w), t1 = sample(w + pitch_w);
        // rows (r, r + 1) as one pair of int16, saturated: min(v, 32767) is swscale's clamp, the lower bound is never reached
        // (the negative lobes of a filter sum to far less than 2^15 / 255 * 128)
#if defined(GVC_EMU)
        *t = ((unsigned)imax(imin(t0, 32767), -32768) & 0xffffu) | ((unsigned)imax(imin(t1, 32767), -32768) << 16);
#else
        unsigned pr;
        asm("cvt.pack.sat.s16.s32 %0, %1, %2;" : "=r"(pr) : "r"(t1), "r"(t0));
        *t = pr;
#endif
    }
}

// ---------------------------------------------------------------- phase 2: vertical ------------
// NP = row pairs of the plan (every output row's taps lie within NP pairs from its first, even, row).  A vertical coefficient is
// split into its low byte (unsigned) and high byte (signed), c = 256 hi + lo, so that sum t c = sum t lo + 256 sum t hi and each of
// the two sums takes TWO rows per dp2a from a (row, row + 1) pair word: per output pixel 2 NP dp2a and a quarter of NP 128-bit
// shared-memory loads (the int32 tile of the earlier version needed one multiply and twice the bytes per tap).
template <int NP>
GVC_HD void resize_phase2(int tid, const PlaneClassDev &pc, const PlaneJob &pj, uint8_t *dst,
                          const TileGeom &g, const gvc_int4 *T) {
    constexpr int NE = NP <= 2 ? 1 : NP <= 6 ? 2 : 3;    // int4 of a table entry that carry coefficients of the first NP pairs
#if defined(GVC_EMU)
    const int warp = tid >> 5;
#else
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
#endif
    const int lane = tid & 31;
    const gvc_int4 *vtab = reinterpret_cast<const gvc_int4 *>(pc.vtab);
    const int  xl      = lane * 4;                   // tile-relative column of this lane's 4 pixels
    if (xl >= g.ncols) return;
    const bool store32 = pj.dpix == 1 && ((((unsigned long long)dst) | (unsigned)pj.dstride) & 3) == 0 && xl + 4 <= g.ncols;
    const gvc_int4 *colbase = T + lane * kTPad;
    // the lane's four pixels of row y, and the table entry of y: both advance by kWarps rows per iteration
    uint8_t        *p  = dst + (long long)(g.x0 + xl) * pj.dpix + (long long)(g.y0 + warp) * pj.dstride;
    const gvc_int4 *vt = vtab + (long long)(g.y0 + warp) * (kVEntry / 4);
    const long long pstep = (long long)kWarps * pj.dstride;
#pragma unroll 1
    for (int y = g.y0 + warp; y < g.y0 + g.nrows; y += kWarps, p += pstep, vt += kWarps * (kVEntry / 4)) {
        unsigned cf[4 * NE];
#pragma unroll
        for (int i = 0; i < NE; ++i) {
#if defined(GVC_EMU)
            const gvc_int4 e = vt[i];
#else
            const gvc_int4 e = __ldg(vt + i);
#endif
            cf[4 * i] = (unsigned)e.x; cf[4 * i + 1] = (unsigned)e.y; cf[4 * i + 2] = (unsigned)e.z; cf[4 * i + 3] = (unsigned)e.w;
        }
        const gvc_int4 *col = colbase + (((int)cf[0] - g.r0) >> 1);
        int l0 = 64 << 12, l1 = 64 << 12, l2 = 64 << 12, l3 = 64 << 12, h0 = 0, h1 = 0, h2 = 0, h3 = 0;      // the rounding term rides in the low sums
#pragma unroll
        for (int j = 0; j < NP; ++j) {
            const unsigned lo = cf[1 + 2 * (j >> 1)], hi = cf[2 + 2 * (j >> 1)];      // bytes 0, 1 for even j, bytes 2, 3 for odd j
            const bool     up = (j & 1) != 0;
            // an output row whose first tap sits on an even source row leaves the last pair empty (warp-uniform: y is)
            if (j == NP - 1 && NP > 1 && (((lo | hi) >> (up ? 16 : 0)) & 0xffffu) == 0) break;
            const gvc_int4 t = col[j];
            l0 = dp2a_su((unsigned)t.x, lo, l0, up); h0 = dp2a_ss((unsigned)t.x, hi, h0, up);
            l1 = dp2a_su((unsigned)t.y, lo, l1, up); h1 = dp2a_ss((unsigned)t.y, hi, h1, up);
            l2 = dp2a_su((unsigned)t.z, lo, l2, up); h2 = dp2a_ss((unsigned)t.z, hi, h2, up);
            l3 = dp2a_su((unsigned)t.w, lo, l3, up); h3 = dp2a_ss((unsigned)t.w, hi, h3, up);
        }
        const int a0 = l0 + h0 * 256, a1 = l1 + h1 * 256, a2 = l2 + h2 * 256, a3 = l3 + h3 * 256;
        const unsigned out = pack_sat_u8x4(a0 >> 19, a1 >> 19, a2 >> 19, a3 >> 19);      // two saturating pack instructions
        if (store32) {
            *reinterpret_cast<unsigned *>(p) = out;
        } else {
            const int n = g.ncols - xl;
            p[0] = (uint8_t)out;
            if (n > 1) p[pj.dpix] = (uint8_t)(out >> 8);
            if (n > 2) p[2 * pj.dpix] = (uint8_t)(out >> 16);
            if (n > 3) p[3 * pj.dpix] = (uint8_t)(out >> 24);
        }
    }
}

// Which plane job / tile does a frame-local tile id belong to?
GVC_HD int find_job(const ResizeParams &P, int tile) {
    int j = 0;
#pragma unroll
    for (int i = 1; i < 3; ++i)
        if (i < P.njobs && tile >= P.job[i].tile_base) j = i;
    return j;
}

// ---------------------------------------------------------------- two-pass generic path --------
GVC_HD int generic_h_sample(const GenericClassDev &pc, const uint8_t *row, int spix, int x) {
    const int16_t *c  = pc.hc + (long long)x * pc.htaps;
    const int      sp = pc.hp[x];
    int            a  = 0;
    for (int j = 0; j < pc.htaps; ++j) {
        const int col = sp + j;
        if (col < pc.sw) a += (int)row[(long long)col * spix] * c[j];
    }
    return imin(a >> 7, 32767);
}

GVC_HD int generic_v_sample(const GenericClassDev &pc, const int16_t *tmp, int y, int x) {
    const int16_t *c  = pc.vc + (long long)y * pc.vtaps;
    const int      sp = pc.vp[y];
    int            a  = 64 << 12;
    for (int j = 0; j < pc.vtaps; ++j) {
        const int r = sp + j;
        if (r < pc.sh) a += (int)tmp[(long long)r * pc.dw + x] * c[j];
    }
    return clip_u8(a >> 19);
}

}  // namespace gvc
