// enc_kernels.cuh -- encoder-side kernels K2..K6 (sm_100a): border extension, half-sample planes, lowres,
// hierarchical motion search, motion compensation, 4x4 transform / quant / dequant / reconstruction.
//
// These stand where the reference runs libx264 inside the spawned ffmpeg (client.go:115, args
// `-vcodec libx264` from server.go:69-70).  Arithmetic follows ITU-T H.264 and x264's routines
// (common/mc.c hpel_filter, frame_init_lowres_core, mc_luma/mc_chroma; common/pixel.c sad/satd;
// common/dct.c sub4x4_dct/add4x4_idct/dct2x2dc; common/quant.c quant/dequant); the search strategy is
// specified by oracle/h264_mb.c (every macroblock independent: exhaustive lowres window, fixed
// full-resolution candidate list, half- then quarter-sample SATD refinement, 16x16 vs 8x8 decision).
//
// Bodies are written per (block, thread/lane) with explicit phases so the CPU tests can execute them
// serially (GVC_EMU, tests/emu); warp-level exchange goes through a small shared-memory record.
#pragma once
#include "dev_common.h"
#include "enc_types.h"

namespace gvc {

// ------------------------------------------------------------------ byte-granular word loads ----
GVC_HD unsigned ld_u32_any(const uint8_t *p) {   // 4 bytes at any address (reads up to 7 bytes past p's word)
#if defined(GVC_EMU)
    return (unsigned)p[0] | ((unsigned)p[1] << 8) | ((unsigned)p[2] << 16) | ((unsigned)p[3] << 24);
#else
    const unsigned long long a = (unsigned long long)p;
    const unsigned *w = reinterpret_cast<const unsigned *>(a & ~3ull);
    return __funnelshift_r(__ldg(w), __ldg(w + 1), (unsigned)(a & 3) * 8);   // read-only planes: ld.global.nc
#endif
}
GVC_HD void ld_u64_any(const uint8_t *p, unsigned &lo, unsigned &hi) {   // 8 bytes at any address
#if defined(GVC_EMU)
    lo = ld_u32_any(p); hi = ld_u32_any(p + 4);
#else
    const unsigned long long a = (unsigned long long)p;
    const unsigned *w = reinterpret_cast<const unsigned *>(a & ~3ull);
    const unsigned  s = (unsigned)(a & 3) * 8, w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2);
    lo = __funnelshift_r(w0, w1, s); hi = __funnelshift_r(w1, w2, s);
#endif
}
GVC_HD void ld_u128_any(const uint8_t *p, unsigned w[4]) {   // 16 bytes at any address
#if defined(GVC_EMU)
    for (int i = 0; i < 4; ++i) w[i] = ld_u32_any(p + 4 * i);
#else
    const unsigned long long a = (unsigned long long)p;
    const unsigned *q = reinterpret_cast<const unsigned *>(a & ~3ull);
    const unsigned  s = (unsigned)(a & 3) * 8, w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2), w3 = __ldg(q + 3), w4 = __ldg(q + 4);
    w[0] = __funnelshift_r(w0, w1, s); w[1] = __funnelshift_r(w1, w2, s);
    w[2] = __funnelshift_r(w2, w3, s); w[3] = __funnelshift_r(w3, w4, s);
#endif
}
GVC_HD unsigned sad_u8x4(unsigned a, unsigned b, unsigned acc) {   // acc + sum |a.b[i] - b.b[i]|
#if defined(GVC_EMU)
    for (int i = 0; i < 4; ++i) { const int d = (int)((a >> (8 * i)) & 255) - (int)((b >> (8 * i)) & 255); acc += d < 0 ? -d : d; }
    return acc;
#else
    unsigned d;                    // VABSDIFF4.U8.ACC with the accumulator as its third operand: written as `__vsadu4(a, b) + acc` the
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));   // compiler sums pairs with an extra IADD3
    return d;
#endif
}

// byte that another CTA of the same launch may have written: must not be served from this SM's L1
GVC_HD int ld_u8_cg(const uint8_t *p) {
#if defined(GVC_EMU)
    return *p;
#else
    return (int)__ldcg(p);
#endif
}

GVC_HD unsigned avg_u8x4(unsigned a, unsigned b) {   // per byte (a + b + 1) >> 1
#if defined(GVC_EMU)
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) r |= ((((a >> (8 * i)) & 255) + ((b >> (8 * i)) & 255) + 1) >> 1) << (8 * i);
    return r;
#else
    return __vavgu4(a, b);
#endif
}
GVC_HD unsigned byte_sel(unsigned a, unsigned b, unsigned sel) {   // prmt.b32 without the replicate-sign modes
#if defined(GVC_EMU)
    const unsigned long long ab = ((unsigned long long)b << 32) | a;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) r |= (unsigned)((ab >> (8 * ((sel >> (4 * i)) & 7))) & 255) << (8 * i);
    return r;
#else
    return __byte_perm(a, b, sel);
#endif
}
GVC_HD int byte_of(unsigned w, int i) {
#if defined(GVC_EMU)
    return (int)((w >> (8 * i)) & 255);
#else
    return (int)__byte_perm(w, 0, 0x4440 + i);
#endif
}

// Two samples per 32-bit register, value = lo + (hi << 16) modulo 2^32 with both halves signed ("lo" may borrow from
// "hi"; sums and differences of such registers stay consistent, as in x264's sum2_t).  |lo|, |hi| < 2^15 throughout.
GVC_HD unsigned pack_even(unsigned w) {   // samples 0 and 2 of a 4-sample word
#if defined(GVC_EMU)
    return w & 0x00ff00ffu;
#else
    return __byte_perm(w, 0, 0x4240);
#endif
}
GVC_HD unsigned pack_odd(unsigned w) {    // samples 1 and 3
#if defined(GVC_EMU)
    return (w >> 8) & 0x00ff00ffu;
#else
    return __byte_perm(w, 0, 0x4341);
#endif
}

// ------------------------------------------------------------------ se(v) bit count ------------
GVC_HD int se_bits(int v) {   // 2 * floor(log2(codeNum + 1)) + 1 with codeNum = 2|v| - (v > 0): floor(log2) = bit length of |v|
#if defined(GVC_EMU)
    int n = 0;
    for (unsigned c = (unsigned)(v < 0 ? -v : v); c; c >>= 1) ++n;
#else
    const int n = 32 - __clz(v < 0 ? -v : v);
#endif
    return 2 * n + 1;
}
GVC_HD int mv_bits(int mx, int my, int px, int py) { return se_bits(mx - px) + se_bits(my - py); }
GVC_HD int median3(int a, int b, int c) { return imax(imin(a, b), imin(imax(a, b), c)); }

// ------------------------------------------------------------------ border extension -----------
GVC_HD void extend_border_px(uint8_t *p, int stride, int w, int h, int x, int y) {
    if (x >= 0 && x < w && y >= 0 && y < h) return;
    p[(long long)y * stride + x] = p[(long long)iclamp(y, 0, h - 1) * stride + iclamp(x, 0, w - 1)];
}

// Y, U and V planes of n equally laid out pictures in one launch (blockIdx.z = picture*3 + plane)
struct BorderBatchArgs {
    uint8_t *p[3];
    int stride[3], w[3], h[3], x_lo[3], x_hi[3], y_lo[3], y_hi[3];
    unsigned long long frame_stride[3];
    int n_frames;
};
// Work decomposition of the batched border kernel: piece `id` (256 samples) of plane k.  Pieces enumerate first the rows
// that lie completely outside the picture (full padded width each), then, for the picture rows, the strips left and right
// of the picture: nothing covers the picture interior.  Returns false when the (id, tid) pair maps to no sample.
GVC_HDH int border_pieces(const BorderBatchArgs &a, int k) {
    const int full_w = a.x_hi[k] - a.x_lo[k], strip_w = full_w - a.w[k];
    return ((a.y_hi[k] - a.y_lo[k]) - a.h[k]) * ((full_w + 255) >> 8) + a.h[k] * ((strip_w + 255) >> 8);
}
GVC_HD bool border_piece_xy(const BorderBatchArgs &a, int k, int id, int tid, int &x, int &y) {
    const int full_w = a.x_hi[k] - a.x_lo[k], strip_w = full_w - a.w[k];
    const int segs_a = (full_w + 255) >> 8, segs_b = (strip_w + 255) >> 8;
    const int rows_out = (a.y_hi[k] - a.y_lo[k]) - a.h[k], n_a = rows_out * segs_a;
    if (id < n_a) {
        const int r = id / segs_a, t = (id - r * segs_a) * 256 + tid;
        y = r < -a.y_lo[k] ? a.y_lo[k] + r : a.h[k] + (r + a.y_lo[k]);
        x = a.x_lo[k] + t;
    } else {
        id -= n_a;
        if (segs_b == 0) return false;
        y = id / segs_b;
        if (y >= a.h[k]) return false;
        const int t = (id - y * segs_b) * 256 + tid;
        x = t < -a.x_lo[k] ? a.x_lo[k] + t : a.w[k] + (t + a.x_lo[k]);
    }
    return x < a.x_hi[k] && y < a.y_hi[k];
}

// Vector variant: 16 samples per lane.  Usable for a plane whose picture width, padded extents, pitch and base address
// are multiples of 16 (every BASELINE size); a piece is one warp = 32 groups of 16 samples.
GVC_HDH bool border16_ok(const BorderBatchArgs &a) {
    for (int k = 0; k < 3; ++k)
        if (((a.w[k] | a.x_lo[k] | a.x_hi[k] | a.stride[k]) & 15) || (((unsigned long long)a.p[k] | a.frame_stride[k]) & 15)) return false;
    return true;
}
GVC_HDH int border16_pieces(const BorderBatchArgs &a, int k) {
    const int gf = (a.x_hi[k] - a.x_lo[k]) >> 4, gs = gf - (a.w[k] >> 4);
    return ((a.y_hi[k] - a.y_lo[k]) - a.h[k]) * ((gf + 31) >> 5) + a.h[k] * ((gs + 31) >> 5);
}
GVC_HD bool border16_piece_xy(const BorderBatchArgs &a, int k, int id, int lane, int &x, int &y) {
    const int gf = (a.x_hi[k] - a.x_lo[k]) >> 4, gs = gf - (a.w[k] >> 4);
    const int segs_a = (gf + 31) >> 5, segs_b = (gs + 31) >> 5;
    const int rows_out = (a.y_hi[k] - a.y_lo[k]) - a.h[k], n_a = rows_out * segs_a;
    int g;
    if (id < n_a) {
        const int r = id / segs_a;
        g = (id - r * segs_a) * 32 + lane;
        y = r < -a.y_lo[k] ? a.y_lo[k] + r : a.h[k] + (r + a.y_lo[k]);
        x = a.x_lo[k] + 16 * g;
    } else {
        id -= n_a;
        if (segs_b == 0) return false;
        y = id / segs_b;
        if (y >= a.h[k]) return false;
        g = (id - y * segs_b) * 32 + lane;
        const int left = (-a.x_lo[k]) >> 4;
        x = g < left ? a.x_lo[k] + 16 * g : a.w[k] + 16 * (g - left);
    }
    return x < a.x_hi[k] && y < a.y_hi[k];
}
GVC_HD void extend_border_16(uint8_t *p, int stride, int w, int h, int x, int y) {   // 16 samples at (x..x+15, y), x % 16 == 0
    const uint8_t *row = p + (long long)iclamp(y, 0, h - 1) * stride;
    gvc_uint4 v;
    if (x >= 0 && x < w) {
        v = *reinterpret_cast<const gvc_uint4 *>(row + x);
    } else {
        const unsigned b = (unsigned)row[x < 0 ? 0 : w - 1] * 0x01010101u;
        v.x = v.y = v.z = v.w = b;
    }
    *reinterpret_cast<gvc_uint4 *>(p + (long long)y * stride + x) = v;
}

struct LowresBatchArgs {
    const uint8_t *src; uint8_t *dst;
    int sstride, dstride, x_lo, x_hi, y_lo, y_hi, lx_lo, lx_hi, ly_lo, ly_hi;   // x_hi / y_hi inclusive maxima of readable source coordinates
    unsigned long long sframe, dframe;
    int n_frames;
};

// ------------------------------------------------------------------ K3 lowres ------------------
// x264 frame_init_lowres_core on the border-extended frame; source coordinates clamp to [lo, hi].
GVC_HD int avg4(int a, int b, int c, int d) { return (((a + b + 1) >> 1) + ((c + d + 1) >> 1) + 1) >> 1; }
GVC_HD void lowres_px(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh,
                      uint8_t *dv, uint8_t *dc, int dstride, int lx, int ly) {
    int v[3][3];
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            v[j][i] = src[(long long)iclamp(2 * ly + j, y_lo, y_hi) * sstride + iclamp(2 * lx + i, x_lo, x_hi)];
    const long long o = (long long)ly * dstride + lx;
    d0[o] = (uint8_t)avg4(v[0][0], v[1][0], v[0][1], v[1][1]);
    if (dh) dh[o] = (uint8_t)avg4(v[0][1], v[1][1], v[0][2], v[1][2]);
    if (dv) dv[o] = (uint8_t)avg4(v[1][0], v[2][0], v[1][1], v[2][1]);
    if (dc) dc[o] = (uint8_t)avg4(v[1][1], v[2][1], v[1][2], v[2][2]);
}

// Four phase-0 lowres samples lx .. lx+3 of row ly in one go.  avg4 is two nested rounding averages, so on packed bytes it
// is vavg(vavg(top, bottom) of the even columns, vavg(top, bottom) of the odd columns): 4 word loads, 4 byte selects, 3
// averages, 1 word store.  Word path when planes / pitches are multiples of four and the eight source columns need no
// clamping; otherwise sample by sample through lowres_px (same results).
GVC_HDH bool lowres_aligned(const LowresBatchArgs &a) {
    return (((unsigned long long)a.src | (unsigned long long)a.dst | (unsigned long long)(unsigned)a.sstride | (unsigned long long)(unsigned)a.dstride |
             a.sframe | a.dframe | (unsigned long long)(unsigned)a.lx_lo | (unsigned long long)(unsigned)a.x_lo |
             (unsigned long long)(unsigned)(a.x_hi + 1)) & 3) == 0;
}
GVC_HD void lowres_quad(const LowresBatchArgs &a, bool al, const uint8_t *src, uint8_t *dst, int lx, int ly) {
    // eight source columns 2lx .. 2lx+7: all readable, or all left of the readable columns (= the first one eight times), or
    // all right of them (= the last one)
    const int  sx = 2 * lx;
    const bool in = sx >= a.x_lo && sx + 7 <= a.x_hi, left = sx + 7 < a.x_lo, right = sx > a.x_hi;
    if (al && lx + 4 <= a.lx_hi && (in || left || right)) {
        const int      xa = in ? sx : (left ? a.x_lo : a.x_hi - 3), xb = in ? sx + 4 : xa;
        const unsigned sel = in ? 0x3210u : (left ? 0x0000u : 0x3333u);
        const uint8_t *r0 = src + (long long)iclamp(2 * ly, a.y_lo, a.y_hi) * a.sstride;
        const uint8_t *r1 = src + (long long)iclamp(2 * ly + 1, a.y_lo, a.y_hi) * a.sstride;
        const unsigned t0 = byte_sel(*reinterpret_cast<const unsigned *>(r0 + xa), 0, sel), t1 = byte_sel(*reinterpret_cast<const unsigned *>(r0 + xb), 0, sel);
        const unsigned b0 = byte_sel(*reinterpret_cast<const unsigned *>(r1 + xa), 0, sel), b1 = byte_sel(*reinterpret_cast<const unsigned *>(r1 + xb), 0, sel);
        const unsigned even = avg_u8x4(byte_sel(t0, t1, 0x6420), byte_sel(b0, b1, 0x6420));
        const unsigned odd  = avg_u8x4(byte_sel(t0, t1, 0x7531), byte_sel(b0, b1, 0x7531));
        *reinterpret_cast<unsigned *>(dst + ((long long)ly * a.dstride + lx)) = avg_u8x4(even, odd);
        return;
    }
    for (int i = 0; i < 4; ++i)
        if (lx + i < a.lx_hi) lowres_px(src, a.sstride, a.x_lo, a.x_hi, a.y_lo, a.y_hi, dst, nullptr, nullptr, nullptr, a.dstride, lx + i, ly);
}

// ------------------------------------------------------------------ K2 half-sample planes -------
// One CTA per 64x32 tile; four samples per work item.  Phase A stages the source tile (with a 4-sample apron, replicated
// at the region edge) in shared memory as words.  Phase B runs the vertical 6-tap on two samples per register (sum2
// form, see pack_even), keeps the unrounded sums as int16 for phase C and writes the V plane.  Phase C runs the horizontal
// 6-tap over source bytes (dp4a, H plane) and over the int16 sums (dp2a, C plane).  Planes, pitch and region bounds that
// are multiples of four move whole words to and from HBM (`al`); anything else goes byte by byte through the same
// arithmetic.
constexpr int kHpTileW = 64, kHpTileH = 32, kHpThreads = 256, kHpWords = kHpTileW / 4 + 2;
struct HpelSmem {
    unsigned src[kHpTileH + 5][kHpWords];      // rows y0-2 .. y0+TH+2, samples x0-4 .. x0+TW+3
    unsigned v[kHpTileH][2 * kHpWords];        // unrounded vertical sums as int16 pairs, same samples
};
GVC_HD int tap6(int a, int b, int c, int d, int e, int f) { return a + f - 5 * (b + e) + 20 * (c + d); }

GVC_HD unsigned funnel_r(unsigned lo, unsigned hi, int bits) {     // (hi:lo) >> bits, 0 <= bits < 32
#if defined(GVC_EMU)
    return (unsigned)((((unsigned long long)hi << 32) | lo) >> bits);
#else
    return __funnelshift_r(lo, hi, (unsigned)bits);
#endif
}
GVC_HD int dp4a_us(unsigned pix, int coef, int acc) {    // acc + sum of pix.b[i] (unsigned) * coef.b[i] (signed)
#if defined(GVC_EMU)
    for (int i = 0; i < 4; ++i) acc += (int)((pix >> (8 * i)) & 255) * (int)(int8_t)((unsigned)coef >> (8 * i));
    return acc;
#else
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pix), "r"(coef), "r"(acc));
    return d;
#endif
}
GVC_HD int dp2a_ss(unsigned pair, int coef, int acc, bool hi) {   // acc + pair.lo16 * coef.b[2hi] + pair.hi16 * coef.b[2hi+1], all signed
#if defined(GVC_EMU)
    const int sh = hi ? 16 : 0;
    return acc + (int)(int16_t)(pair & 0xffff) * (int)(int8_t)((unsigned)coef >> sh) + (int)(int16_t)(pair >> 16) * (int)(int8_t)((unsigned)coef >> (sh + 8));
#else
    int d;
    if (hi) asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(coef), "r"(acc));
    else    asm("dp2a.lo.s32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(pair), "r"(coef), "r"(acc));
    return d;
#endif
}
// (v + bias) clamped to [0, top] on both int16 halves of a clean pair (no borrow between the halves)
GVC_HD unsigned bias_clamp_s16x2(unsigned v, unsigned bias2, unsigned top2) {
#if defined(GVC_EMU)
    unsigned r = 0;
    for (int i = 0; i < 2; ++i) {
        int h = (int)(int16_t)(v >> (16 * i)) + (int)(int16_t)(bias2 >> (16 * i));
        const int t = (int)(int16_t)(top2 >> (16 * i));
        h = h < 0 ? 0 : (h > t ? t : h);
        r |= (unsigned)(h & 0xffff) << (16 * i);
    }
    return r;
#else
    return __vmins2(__vmaxs2(__vadd2(v, bias2), 0u), top2);
#endif
}
constexpr int kTapA = 0x1414fb01;    // bytes  1, -5, 20, 20
constexpr int kTapB = 0x000001fb;    // bytes -5,  1,  0,  0

struct HpelArgs {
    const uint8_t *src; uint8_t *dh, *dv, *dc;   // all at sample (0,0), same stride
    int stride, x_lo, x_hi, y_lo, y_hi;          // region to fill: [x_lo,x_hi) x [y_lo,y_hi)
    int cx_lo, cx_hi, cy_lo, cy_hi;              // readable part of src (reads clamp to it: edge replication); = the region, or the
                                                 // picture itself when the border has not been extended yet
};
GVC_HDH bool hpel_aligned(const HpelArgs &a) {
    return (((unsigned long long)a.src | (unsigned long long)a.dh | (unsigned long long)a.dv | (unsigned long long)a.dc |
             (unsigned long long)(unsigned)a.stride | (unsigned long long)(unsigned)a.x_lo | (unsigned long long)(unsigned)a.x_hi |
             (unsigned long long)(unsigned)a.cx_lo | (unsigned long long)(unsigned)a.cx_hi) & 3) == 0;
}
GVC_HD void hpel_store4(uint8_t *plane, const HpelArgs &a, bool al, int x, int y, unsigned w) {   // samples x..x+3 of row y
    if (y >= a.y_hi || x >= a.x_hi) return;
    uint8_t *p = plane + (long long)y * a.stride + x;
    if (al) { *reinterpret_cast<unsigned *>(p) = w; return; }
    for (int i = 0; i < 4; ++i)
        if (x + i < a.x_hi) p[i] = (uint8_t)(w >> (8 * i));
}
GVC_HD void hpel_phase_a(int tid, const HpelArgs &a, bool al, int x0, int y0, HpelSmem &sm) {
    for (int i = tid; i < (kHpTileH + 5) * kHpWords; i += kHpThreads) {
        const int r = i / kHpWords, c = i - r * kHpWords;
        const int x = x0 - 4 + 4 * c;
        const uint8_t *row = a.src + (long long)iclamp(y0 - 2 + r, a.cy_lo, a.cy_hi - 1) * a.stride;
        unsigned w;
        if (al && x >= a.cx_lo && x < a.cx_hi) {
            w = *reinterpret_cast<const unsigned *>(row + x);
        } else if (al) {
            w = 0x01010101u * row[x < a.cx_lo ? a.cx_lo : a.cx_hi - 1];
        } else {
            w = 0;
            for (int k = 0; k < 4; ++k) w |= (unsigned)row[iclamp(x + k, a.cx_lo, a.cx_hi - 1)] << (8 * k);
        }
        sm.src[r][c] = w;
    }
}
GVC_HD void hpel_phase_b(int tid, const HpelArgs &a, bool al, int x0, int y0, HpelSmem &sm) {
    for (int i = tid; i < kHpTileH * kHpWords; i += kHpThreads) {
        const int r = i / kHpWords, c = i - r * kHpWords;
        unsigned e[6], o[6];
        for (int k = 0; k < 6; ++k) { const unsigned w = sm.src[r + k][c]; e[k] = pack_even(w); o[k] = pack_odd(w); }
        // sums stay within int16 (-2550 .. 10710); a negative low half has borrowed one from the high half: give it back
        unsigned ve = (e[0] + e[5]) + 20u * (e[2] + e[3]) - 5u * (e[1] + e[4]);
        unsigned vo = (o[0] + o[5]) + 20u * (o[2] + o[3]) - 5u * (o[1] + o[4]);
        ve += (ve & 0x8000u) << 1;
        vo += (vo & 0x8000u) << 1;
        sm.v[r][2 * c]     = byte_sel(ve, vo, 0x5410);      // samples 0, 1
        sm.v[r][2 * c + 1] = byte_sel(ve, vo, 0x7632);      // samples 2, 3
        if (c >= 1 && c <= kHpTileW / 4) {
            const unsigned be = (bias_clamp_s16x2(ve, 0x00100010u, 0x1fff1fffu) >> 5) & 0x00ff00ffu;
            const unsigned bo = (bias_clamp_s16x2(vo, 0x00100010u, 0x1fff1fffu) >> 5) & 0x00ff00ffu;
            hpel_store4(a.dv, a, al, x0 - 4 + 4 * c, y0 + r, byte_sel(be, bo, 0x6240));
        }
    }
}
GVC_HD void hpel_phase_c(int tid, const HpelArgs &a, bool al, int x0, int y0, HpelSmem &sm) {
    for (int i = tid; i < kHpTileH * (kHpTileW / 4); i += kHpThreads) {
        const int r = i / (kHpTileW / 4), c = i - r * (kHpTileW / 4);
        const int x = x0 + 4 * c, y = y0 + r;
        if (x >= a.x_hi || y >= a.y_hi) continue;
        // H plane: samples x-2 .. x+6 sit in source words c, c+1, c+2 (word c+1 starts at x)
        const unsigned w0 = sm.src[r + 2][c], w1 = sm.src[r + 2][c + 1], w2 = sm.src[r + 2][c + 2];
        int h[4];
        h[0] = dp4a_us(funnel_r(w1, w2, 16), kTapB, dp4a_us(funnel_r(w0, w1, 16), kTapA, 16));
        h[1] = dp4a_us(funnel_r(w1, w2, 24), kTapB, dp4a_us(funnel_r(w0, w1, 24), kTapA, 16));
        h[2] = dp4a_us(w2, kTapB, dp4a_us(w1, kTapA, 16));
        h[3] = dp4a_us(w2 >> 8, kTapB, dp4a_us(funnel_r(w1, w2, 8), kTapA, 16));
        unsigned out = 0;
        for (int k = 0; k < 4; ++k) out |= (unsigned)clip_u8(h[k] >> 5) << (8 * k);
        hpel_store4(a.dh, a, al, x, y, out);
        // C plane: the same taps over the vertical sums of samples x-2 .. x+6: int16 pairs 2c+1 .. 2c+5 of the row
        const unsigned *vr = &sm.v[r][2 * c + 1];
        const unsigned v0 = vr[0], v1 = vr[1], v2 = vr[2], v3 = vr[3], v4 = vr[4];
        const unsigned s0 = funnel_r(v0, v1, 16), s1 = funnel_r(v1, v2, 16), s2 = funnel_r(v2, v3, 16), s3 = funnel_r(v3, v4, 16);
        int q[4];
        q[0] = dp2a_ss(v2, kTapB, dp2a_ss(v1, kTapA, dp2a_ss(v0, kTapA, 512, false), true), false);
        q[1] = dp2a_ss(s2, kTapB, dp2a_ss(s1, kTapA, dp2a_ss(s0, kTapA, 512, false), true), false);
        q[2] = dp2a_ss(v3, kTapB, dp2a_ss(v2, kTapA, dp2a_ss(v1, kTapA, 512, false), true), false);
        q[3] = dp2a_ss(s3, kTapB, dp2a_ss(s2, kTapA, dp2a_ss(s1, kTapA, 512, false), true), false);
        out = 0;
        for (int k = 0; k < 4; ++k) out |= (unsigned)clip_u8(q[k] >> 10) << (8 * k);
        hpel_store4(a.dc, a, al, x, y, out);
    }
}

// ------------------------------------------------------------------ K5 motion compensation -----
// H.264 8.4.2.2.1 quarter-sample rule / x264 mc_luma: which two planes are averaged for fraction q = fy*4+fx
// (2 bits per entry: 0 full, 1 H, 2 V, 3 C).
constexpr unsigned kQpelRef0 = 0u | (1u << 2) | (1u << 4) | (1u << 6) | (0u << 8) | (1u << 10) | (1u << 12) | (1u << 14) |
                               (2u << 16) | (3u << 18) | (3u << 20) | (3u << 22) | (0u << 24) | (1u << 26) | (1u << 28) | (1u << 30);
constexpr unsigned kQpelRef1 = 0u | (0u << 2) | (1u << 4) | (0u << 6) | (2u << 8) | (2u << 10) | (3u << 12) | (2u << 14) |
                               (2u << 16) | (2u << 18) | (3u << 20) | (2u << 22) | (2u << 24) | (2u << 26) | (3u << 28) | (2u << 30);

struct LumaSrc { const uint8_t *s1, *s2; };   // s2 == nullptr: plain copy
GVC_HD LumaSrc mc_luma_src(const RefPlanes &r, int stride, int x, int y, int mvx, int mvy) {
    const int fx = mvx & 3, fy = mvy & 3, q = fy * 4 + fx;
    const uint8_t *pl[4] = {r.y, r.hh, r.hv, r.hc};
    const long long off = (long long)(y + (mvy >> 2)) * stride + x + (mvx >> 2);
    LumaSrc o;
    o.s1 = pl[(kQpelRef0 >> (2 * q)) & 3] + off + (fy == 3 ? stride : 0);
    o.s2 = (q & 5) ? pl[(kQpelRef1 >> (2 * q)) & 3] + off + (fx == 3 ? 1 : 0) : nullptr;
    return o;
}
// ------------------------------------------------------------------ 4x4 transforms --------------
GVC_HD void dct4x4(int d[16]) {   // x264 sub4x4_dct butterflies, rows then columns, in place
    for (int y = 0; y < 4; ++y) {
        const int s03 = d[y * 4] + d[y * 4 + 3], s12 = d[y * 4 + 1] + d[y * 4 + 2];
        const int d03 = d[y * 4] - d[y * 4 + 3], d12 = d[y * 4 + 1] - d[y * 4 + 2];
        d[y * 4] = s03 + s12; d[y * 4 + 1] = 2 * d03 + d12; d[y * 4 + 2] = s03 - s12; d[y * 4 + 3] = d03 - 2 * d12;
    }
    for (int x = 0; x < 4; ++x) {
        const int s03 = d[x] + d[12 + x], s12 = d[4 + x] + d[8 + x];
        const int d03 = d[x] - d[12 + x], d12 = d[4 + x] - d[8 + x];
        d[x] = s03 + s12; d[4 + x] = 2 * d03 + d12; d[8 + x] = s03 - s12; d[12 + x] = d03 - 2 * d12;
    }
}
GVC_HD void idct4x4(int c[16]) {  // H.264 8.5.12.2, result includes the (x+32)>>6
    for (int y = 0; y < 4; ++y) {
        const int s02 = c[y * 4] + c[y * 4 + 2], d02 = c[y * 4] - c[y * 4 + 2];
        const int s13 = c[y * 4 + 1] + (c[y * 4 + 3] >> 1), d13 = (c[y * 4 + 1] >> 1) - c[y * 4 + 3];
        c[y * 4] = s02 + s13; c[y * 4 + 1] = d02 + d13; c[y * 4 + 2] = d02 - d13; c[y * 4 + 3] = s02 - s13;
    }
    for (int x = 0; x < 4; ++x) {
        const int s02 = c[x] + c[8 + x], d02 = c[x] - c[8 + x];
        const int s13 = c[4 + x] + (c[12 + x] >> 1), d13 = (c[4 + x] >> 1) - c[12 + x];
        c[x] = (s02 + s13 + 32) >> 6; c[4 + x] = (d02 + d13 + 32) >> 6; c[8 + x] = (d02 - d13 + 32) >> 6; c[12 + x] = (s02 - s13 + 32) >> 6;
    }
}
GVC_HD int quant_one(int c, int mf, int bias) {   // x264 QUANT_ONE: sign(c) * (((bias + |c|) * mf) >> 16)
    const int s = c >> 31;                                                   // 0 or -1
    const int q = (int)(((unsigned)(bias + ((c ^ s) - s)) * (unsigned)mf) >> 16);
    return (q ^ s) - s;
}
GVC_HD int dequant_one(int c, int scale, int qp) {   // H.264 8.5.12.1 / x264 dequant_4x4; wraps to int16 like the oracle
    const int qbits = qp / 6 - 4;
    const int v = qbits >= 0 ? (c * scale) * (1 << qbits) : (c * scale + (1 << (-qbits - 1))) >> (-qbits);
    return (int)(int16_t)v;
}
GVC_HD int coef_class(int i) { return ((i >> 2) & 1) + (i & 1); }
GVC_HD int zigzag4(int i) {   // raster index of the i-th coefficient in frame zig-zag order
    // 0,1,4,8,5,2,3,6,9,12,13,10,7,11,14,15 packed 4 bits each, LSB first
    const unsigned long long z = 0xFEB7ADC963258410ull;
    return (int)((z >> (4 * i)) & 15);
}

// ------------------------------------------------------------------ 8x8 transform path ----------
// x264 sub8x8_dct8 / add8x8_idct8 butterflies (H.264 8.5.13), one 1-D pass over 8 values with stride `s`.
GVC_HD void dct8_1d(int *d, int s) {
    const int s07 = d[0] + d[7 * s], s16 = d[s] + d[6 * s], s25 = d[2 * s] + d[5 * s], s34 = d[3 * s] + d[4 * s];
    const int a0 = s07 + s34, a1 = s16 + s25, a2 = s07 - s34, a3 = s16 - s25;
    const int d07 = d[0] - d[7 * s], d16 = d[s] - d[6 * s], d25 = d[2 * s] - d[5 * s], d34 = d[3 * s] - d[4 * s];
    const int a4 = d16 + d25 + (d07 + (d07 >> 1)), a5 = d07 - d34 - (d25 + (d25 >> 1));
    const int a6 = d07 + d34 - (d16 + (d16 >> 1)), a7 = d16 - d25 + (d34 + (d34 >> 1));
    d[0] = a0 + a1; d[s] = a4 + (a7 >> 2); d[2 * s] = a2 + (a3 >> 1); d[3 * s] = a5 + (a6 >> 2);
    d[4 * s] = a0 - a1; d[5 * s] = a6 - (a5 >> 2); d[6 * s] = (a2 >> 1) - a3; d[7 * s] = (a4 >> 2) - a7;
}
GVC_HD void idct8_1d(int *d, int s) {
    const int a0 = d[0] + d[4 * s], a2 = d[0] - d[4 * s];
    const int a4 = (d[2 * s] >> 1) - d[6 * s], a6 = (d[6 * s] >> 1) + d[2 * s];
    const int b0 = a0 + a6, b2 = a2 + a4, b4 = a2 - a4, b6 = a0 - a6;
    const int a1 = -d[3 * s] + d[5 * s] - d[7 * s] - (d[7 * s] >> 1), a3 = d[s] + d[7 * s] - d[3 * s] - (d[3 * s] >> 1);
    const int a5 = -d[s] + d[7 * s] + d[5 * s] + (d[5 * s] >> 1), a7 = d[3 * s] + d[5 * s] + d[s] + (d[s] >> 1);
    const int b1 = (a7 >> 2) + a1, b3 = a3 + (a5 >> 2), b5 = (a3 >> 2) - a5, b7 = a7 - (a1 >> 2);
    d[0] = b0 + b7; d[s] = b2 + b5; d[2 * s] = b4 + b3; d[3 * s] = b6 + b1;
    d[4 * s] = b6 - b1; d[5 * s] = b4 - b3; d[6 * s] = b2 - b5; d[7 * s] = b0 - b7;
}
GVC_HD int coef8_class(int i) {     // x264 quant8_scan by (row & 3, col & 3): 0,3,4,3, 3,1,5,1, 4,5,2,5, 3,1,5,1 packed 4 bits each
    return (int)((0x1513525415133430ull >> (4 * (((i >> 3) & 3) * 4 + (i & 3)))) & 15);
}
struct Transform8Args {
    const uint8_t *src, *pred;
    uint8_t       *rec;
    int            sstride, pstride, rstride, bw, bh;   // bw x bh blocks of 8x8
    Quant8Params   q;
    int16_t       *coefs;        // 64 levels per block, raster (row = vertical frequency)
    uint8_t       *nz;           // per block
};
// one thread = one 8x8 block: residual, forward transform, quant, levels out, dequant, inverse, reconstruction
GVC_HD void transform8x8_block(const Transform8Args &a, int bx, int by) {
    int d[64];
    const uint8_t *s = a.src + (long long)(8 * by) * a.sstride + 8 * bx, *p = a.pred + (long long)(8 * by) * a.pstride + 8 * bx;
    for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) d[y * 8 + x] = (int)s[(long long)y * a.sstride + x] - (int)p[(long long)y * a.pstride + x];
    for (int y = 0; y < 8; ++y) dct8_1d(d + 8 * y, 1);
    for (int x = 0; x < 8; ++x) dct8_1d(d + x, 8);
    int16_t *co = a.coefs + (long long)(by * a.bw + bx) * 64;
    int      nz = 0;
    const int qbits = a.q.qp / 6 - 6;
    for (int i = 0; i < 64; ++i) {
        const int cls = coef8_class(i);
        const int l = quant_one((int)(int16_t)d[i], a.q.mf[cls], a.q.bias[cls]);
        co[i] = (int16_t)l;
        nz |= l;
        const int v = qbits >= 0 ? (l * a.q.dq[cls]) * (1 << qbits) : (l * a.q.dq[cls] + (1 << (-qbits - 1))) >> (-qbits);
        d[i] = (int)(int16_t)v;
    }
    a.nz[by * a.bw + bx] = nz != 0;
    for (int y = 0; y < 8; ++y) idct8_1d(d + 8 * y, 1);
    for (int x = 0; x < 8; ++x) idct8_1d(d + x, 8);
    uint8_t *r = a.rec + (long long)(8 * by) * a.rstride + 8 * bx;
    for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) r[(long long)y * a.rstride + x] = (uint8_t)clip_u8((int)p[(long long)y * a.pstride + x] + ((d[y * 8 + x] + 32) >> 6));
}

// SATD contribution of one 4x4 block: sum |H D H| (always even; callers halve per x264's 8x4 pairing)
GVC_HD int hadamard4_abs(const int d[16]) {
    int t[16], s = 0;
    for (int y = 0; y < 4; ++y) {
        const int s01 = d[y * 4] + d[y * 4 + 1], d01 = d[y * 4] - d[y * 4 + 1];
        const int s23 = d[y * 4 + 2] + d[y * 4 + 3], d23 = d[y * 4 + 2] - d[y * 4 + 3];
        t[y * 4] = s01 + s23; t[y * 4 + 1] = s01 - s23; t[y * 4 + 2] = d01 + d23; t[y * 4 + 3] = d01 - d23;
    }
    for (int x = 0; x < 4; ++x) {
        const int s01 = t[x] + t[4 + x], d01 = t[x] - t[4 + x];
        const int s23 = t[8 + x] + t[12 + x], d23 = t[8 + x] - t[12 + x];
        s += iabs(s01 + s23) + iabs(s01 - s23) + iabs(d01 + d23) + iabs(d01 - d23);
    }
    return s;
}

// ------------------------------------------------------------------ K4a lowres search ----------
// One warp per macroblock (an 8x8 block of the half-resolution plane), exhaustive over (2r+1)^2 displacements.  The warp
// first stages the (8+2r)^2 search window in shared memory.  A work item is one column dx and kLowChunk consecutive dy:
// the lane shifts each window row to dx once and feeds it to every candidate of the chunk that uses that row, so a
// candidate costs its 16 four-byte SADs and a sixth of the row fetches.  key = cost<<11 | index; the warp minimum wins.
constexpr int kLowMaxR = 22, kLowRows = 8 + 2 * kLowMaxR, kLowPitch = kLowMaxR / 2 + 4;
constexpr int kLowChunk = 6, kLowChunkWide = 11;     // dy per work item: 6 fills two rounds of lanes at r = 8, 11 four rounds at r = 16
GVC_HDH int me_lowres_chunk(int r) { return r <= 10 ? kLowChunk : kLowChunkWide; }
struct MeLowArgs {
    const uint8_t *cur, *ref;   // lowres planes at (0,0)
    int stride, mbw, mbh, r, lambda;
    int16_t *out_mv;            // (dx,dy) per MB
};
struct MeLowSmem { unsigned win[kLowRows + kLowChunkWide][kLowPitch]; };   // per warp: window rows, byte 0 = column 8*mx - r; the
                                                                       // spare rows are read for dy past the range, never filled
// The zero displacement costs SAD0 + 2 lambda; every other one at least 8 lambda of vector bits alone (se(+-4) takes 7
// bits).  So SAD0 < 6 lambda decides the exhaustive search before it starts: same result, no window, no candidates.
GVC_HD unsigned me_lowres_zero_sad(int lane, const MeLowArgs &a, int mx, int my) {   // lanes 0..7: one row each
    if (lane >= 8) return 0;
    const int o = (8 * my + lane) * a.stride + 8 * mx;
    unsigned c0, c1, r0, r1;
    ld_u64_any(a.cur + o, c0, c1);
    ld_u64_any(a.ref + o, r0, r1);
    return sad_u8x4(c1, r1, sad_u8x4(c0, r0, 0));
}
GVC_HD bool me_lowres_zero_wins(const MeLowArgs &a, unsigned sad0) { return sad0 < 6u * (unsigned)a.lambda; }
GVC_HD void me_lowres_load(int lane, const MeLowArgs &a, int mx, int my, MeLowSmem &sm) {
    const int rows = 8 + 2 * a.r, words = (8 + 2 * a.r + 3) >> 2;
    const uint8_t *org = a.ref + ((8 * my - a.r) * a.stride + 8 * mx - a.r);
    const int cs = words <= 8 ? 3 : 4, c = lane & ((1 << cs) - 1);     // 8 or 16 lanes across a row
    if (c < words)
        for (int y = lane >> cs; y < rows; y += 32 >> cs) sm.win[y][c] = ld_u32_any(org + (y * a.stride + 4 * c));
}
template <int kChunk>
GVC_HD unsigned me_lowres_lane(int lane, const MeLowArgs &a, int mx, int my, const MeLowSmem &sm) {
    const int n = 2 * a.r + 1, items = n * ((n + kChunk - 1) / kChunk);
    unsigned  c[8][2];
    const uint8_t *cp = a.cur + (8 * my * a.stride + 8 * mx);
    for (int j = 0; j < 8; ++j) ld_u64_any(cp + j * a.stride, c[j][0], c[j][1]);
    unsigned best = 0xffffffffu;
    for (int it = lane; it < items; it += 32) {
        const int ch = it / n, o = it - ch * n, y0 = ch * kChunk;    // o = dx + r, y0 = first dy + r of the chunk
        const int wi = o >> 2, sh = (o & 3) * 8;
        unsigned acc[kChunk];
        for (int k = 0; k < kChunk; ++k) acc[k] = 0;
#if !defined(GVC_EMU)
#pragma unroll
#endif
        for (int y = 0; y < kChunk + 7; ++y) {
            const unsigned *row = sm.win[y0 + y];                        // rows past the window only feed discarded dy
            const unsigned w1 = row[wi + 1], lo = funnel_r(row[wi], w1, sh), hi = funnel_r(w1, row[wi + 2], sh);
#if !defined(GVC_EMU)
#pragma unroll
#endif
            for (int k = 0; k < kChunk; ++k) {
                const int j = y - k;
                if (j >= 0 && j < 8) acc[k] = sad_u8x4(c[j][1], hi, sad_u8x4(c[j][0], lo, acc[k]));
            }
        }
        const int bx = se_bits(4 * (o - a.r));
        for (int k = 0; k < kChunk; ++k) {
            const int q = y0 + k;                                        // dy + r
            if (q >= n) break;
            const unsigned cost = acc[k] + (unsigned)(a.lambda * (bx + se_bits(4 * (q - a.r))));
            const unsigned key  = (cost << 11) | (unsigned)(q * n + o);
            best = key < best ? key : best;
        }
    }
    return best;
}
GVC_HD void me_lowres_out(int lane, const MeLowArgs &a, int mx, int my, unsigned best) {
    if (lane != 0) return;
    const int n = 2 * a.r + 1, idx = (int)(best & 2047);
    a.out_mv[2 * (my * a.mbw + mx)]     = (int16_t)(idx - (idx / n) * n - a.r);
    a.out_mv[2 * (my * a.mbw + mx) + 1] = (int16_t)(idx / n - a.r);
}

}  // namespace gvc
#include "me_refine.cuh"
namespace gvc {

// ------------------------------------------------------------------ record compaction ------------
// Most 4x4 blocks quantise to nothing; the host entropy coder only needs the ones that did not.  Per macroblock the
// blocks with a nonzero level are packed back to back (32 bytes each, ascending block number, block 24 = both chroma
// DC sets, block 25 = luma DC) straight into host memory that is mapped into the device address space, so the copy back
// shrinks from 832 bytes per macroblock to what the picture actually holds.
GVC_HDH unsigned compact_mask(unsigned nz) {     // nzmask -> one bit per 16-coefficient block of the record (26 blocks)
    return (nz & 0x00ffffffu) | ((nz & 0x03000000u) ? (1u << 24) : 0u) | (((nz >> 26) & 1u) << 25);
}
GVC_HDH int popcount32(unsigned v) {
#if defined(__CUDA_ARCH__)
    return __popc(v);
#else
    return __builtin_popcount(v);
#endif
}
// Destination of the packed blocks of one picture (all three members null = packing off, e.g. in kernel-level tests).
struct PackOut {
    int16_t  *blocks;     // host-mapped: 16 coefficients per present block
    uint32_t *offsets;    // per MB: index (in blocks) of its first packed block
    unsigned *counter;    // blocks handed out so far in this picture (zeroed before the launch)
};
// lane 0 reserves room for the macroblock's present blocks (order between macroblocks is whatever the hardware
// schedules; the offset table makes it irrelevant), every lane then stores its block there if present.
GVC_HD void pack_reserve(int lane, const PackOut &po, int mb, uint32_t nz, int *flags) {
    if (lane != 0 || !po.blocks) return;
    const unsigned cm = compact_mask(nz);
#if defined(GVC_EMU)
    const unsigned base = *po.counter;
    *po.counter += (unsigned)popcount32(cm);
#else
    const unsigned base = atomicAdd(po.counter, (unsigned)popcount32(cm));
#endif
    po.offsets[mb] = base;
    flags[30] = (int)cm;
    flags[31] = (int)base;
}

// the small kernel that closes a picture: *count_out = *count_in; system-wide fence; *done = 1
struct SignalArgs { int *done; unsigned *count_out; const unsigned *count_in; };

// ------------------------------------------------------------------ K5+K6 inter MB encode -------
// coef layout per MB (int16[26][16]): blocks 0..15 luma 4x4 (raster block order, zig-zag coefficient order),
// 16..19 Cb AC, 20..23 Cr AC (entry 0 = 0), 24: chroma DC levels Cb[0..3] Cr[4..7], 25: luma DC (Intra16x16 only).
// nzmask: bit b = block b has a nonzero level; bits 24 / 25 = Cb / Cr DC.
struct EncMbArgs {
    const uint8_t *cur_y, *cur_u, *cur_v;   // source planes (0,0)
    RefPlanes      ref;
    uint8_t       *rec_y, *rec_u, *rec_v;   // reconstruction planes (0,0)
    int            sy, sc, mbw, mbh;
    QuantParams    q;
    const MbMotion *motion;
    int16_t       *coefs;
    uint32_t      *nzmask;
    PackOut        pack;
    uint8_t       *nnz_y, *nnz_u, *nnz_v;   // optional: nonzero levels per 4x4 block, [4*mbh][4*mbw] / [2*mbh][2*mbw] (for the entropy kernels)
};
struct EncMbSmem {               // per warp
    int      cdc[8];             // DC of the eight chroma blocks after the forward transform
    int      dclev[8];           // quantised chroma DC levels, Cb[0..3] Cr[4..7]
    unsigned mask;               // nzmask under construction
    int      flags[32];          // [30], [31]: compact mask and first packed block (pack_reserve)
    unsigned cpred[32];          // chroma prediction, one 4-sample row word per lane: [plane][row 0..7][half]
};
struct EncLane {                 // what a lane carries in registers across the phases
    int      d[16];              // residual -> coefficients (raster)
    unsigned pred[4];            // prediction rows, four samples each
    unsigned lev[8];             // quantised levels in zig-zag order, two int16 per word
};
GVC_HDH bool enc_mb_aligned(const EncMbArgs &a) {   // planes and pitches allow word loads / stores of 4-sample rows
    return (((unsigned long long)a.cur_y | (unsigned long long)a.cur_u | (unsigned long long)a.cur_v | (unsigned long long)a.rec_y |
             (unsigned long long)a.rec_u | (unsigned long long)a.rec_v | (unsigned long long)(unsigned)a.sy |
             (unsigned long long)(unsigned)a.sc) & 3) == 0 && ((unsigned long long)a.coefs & 15) == 0;
}
GVC_HD void smem_or(unsigned *p, unsigned v) {
#if defined(GVC_EMU)
    *p |= v;
#else
    atomicOr(p, v);
#endif
}
// lane -> its 4x4 block: lanes 0..15 luma (raster), 16..19 Cb, 20..23 Cr.  Position in the block's plane, partition index.
GVC_HD void enc_lane_block(int lane, int mx, int my, int &x, int &y, int &k) {
    if (lane < 16) { const int bx = 4 * (lane & 3), by = 4 * (lane >> 2); x = 16 * mx + bx; y = 16 * my + by; k = (by >> 3) * 2 + (bx >> 3); }
    else           { const int b = (lane - 16) & 3; x = 8 * mx + 4 * (b & 1); y = 8 * my + 4 * (b >> 1); k = b; }   // chroma block b lies under partition b
}
// number of nonzero levels of the lane's block -> the maps the entropy kernels derive nC from (H.264 9.2.1)
GVC_HD void store_block_nnz(int lane, int mx, int my, int mbw, uint8_t *nnz_y, uint8_t *nnz_u, uint8_t *nnz_v, const int d[16]) {
    if (!nnz_y) return;
    int cnt = 0;
    for (int i = 0; i < 16; ++i) cnt += d[i] != 0;
    if (lane < 16) nnz_y[(4 * my + (lane >> 2)) * (4 * mbw) + 4 * mx + (lane & 3)] = (uint8_t)cnt;
    else { const int b = (lane - 16) & 3; (lane < 20 ? nnz_u : nnz_v)[(2 * my + (b >> 1)) * (2 * mbw) + 2 * mx + (b & 1)] = (uint8_t)cnt; }
}
// Phase 0: chroma prediction (H.264 8.4.2.2.2, x264 mc_chroma) with the whole warp: the two 8x8 planes are 32 row words of
// four samples, one per lane.  Two samples per register: the four weights sum to 64, so a weighted sum stays below 2^16 and
// the halves never meet.
GVC_HD void enc_mb_p0(int lane, const EncMbArgs &a, int mx, int my, EncMbSmem &sm) {
    const int row = (lane >> 1) & 7, half = lane & 1, k = (row >> 2) * 2 + half;      // chroma block k lies under partition k
    const MbMotion &mo = a.motion[my * a.mbw + mx];
    const int mvx = mo.mv[k][0], mvy = mo.mv[k][1];
    const unsigned dx = (unsigned)(mvx & 7), dy = (unsigned)(mvy & 7);
    const unsigned cA = (8 - dx) * (8 - dy), cB = dx * (8 - dy), cC = (8 - dx) * dy, cD = dx * dy;
    const uint8_t *s = (lane < 16 ? a.ref.u : a.ref.v) + ((long long)(8 * my + row + (mvy >> 3)) * a.sc + 8 * mx + 4 * half + (mvx >> 3));
    unsigned t0, t1, b0, b1;
    ld_u64_any(s, t0, t1);
    ld_u64_any(s + a.sc, b0, b1);
    const unsigned tr = funnel_r(t0, t1, 8), br = funnel_r(b0, b1, 8);                 // the samples one to the right
    const unsigned e = cA * pack_even(t0) + cB * pack_even(tr) + cC * pack_even(b0) + cD * pack_even(br) + 0x00200020u;
    const unsigned o = cA * pack_odd(t0) + cB * pack_odd(tr) + cC * pack_odd(b0) + cD * pack_odd(br) + 0x00200020u;
    sm.cpred[lane] = byte_sel((e >> 6) & 0x00ff00ffu, (o >> 6) & 0x00ff00ffu, 0x6240);
    if (lane == 0) sm.mask = 0;
}
// Phase 1: prediction, residual, forward 4x4 transform of every block; chroma DCs go to the shared record.
GVC_HD void enc_mb_p1(int lane, const EncMbArgs &a, bool al, int mx, int my, EncMbSmem &sm, EncLane &L) {
    if (lane >= 24) return;
    int x, y, k;
    enc_lane_block(lane, mx, my, x, y, k);
    if (lane < 16) {
        const MbMotion &mo = a.motion[my * a.mbw + mx];
        const LumaSrc s = mc_luma_src(a.ref, a.sy, x, y, mo.mv[k][0], mo.mv[k][1]);
        for (int j = 0; j < 4; ++j) {
            L.pred[j] = ld_u32_any(s.s1 + j * a.sy);
            if (s.s2) L.pred[j] = avg_u8x4(L.pred[j], ld_u32_any(s.s2 + j * a.sy));
        }
    } else {
        const int b = (lane - 16) & 3, base = (lane < 20 ? 0 : 16) + 8 * (b >> 1) + (b & 1);   // rows 4(b>>1) .., half b&1 of phase 0's words
        for (int j = 0; j < 4; ++j) L.pred[j] = sm.cpred[base + 2 * j];
    }
    GVC_RECONVERGE(0x00ffffffu);
    const int st = lane < 16 ? a.sy : a.sc;
    const uint8_t *src = (lane < 16 ? a.cur_y : lane < 20 ? a.cur_u : a.cur_v) + (y * st + x);
    for (int j = 0; j < 4; ++j) {
        const unsigned c = al ? *reinterpret_cast<const unsigned *>(src + j * st) : ld_u32_any(src + j * st);
        for (int i = 0; i < 4; ++i) L.d[j * 4 + i] = byte_of(c, i) - byte_of(L.pred[j], i);
    }
    dct4x4(L.d);
    if (lane >= 16) sm.cdc[lane - 16] = L.d[0];
}
// Between phase 1 and phase 2: can ANY coefficient of the lane's block quantise to a nonzero level?  (bias + |c|) * mf >> 16 >= 1
// exactly when |c| >= thr of its class (QuantParams::thr_*), so three maxima per block decide it.  Static content leaves most
// macroblocks of a P picture without a single level: when no lane of the warp answers yes, phase 2 is replaced by
// enc_mb_p2_zero -- same outputs (all levels zero, reconstruction = prediction) without quantiser, inverse transform or packing.
GVC_HD bool enc_mb_any_level(int lane, const EncMbArgs &a, const EncMbSmem &sm, const EncLane &L) {
    if (lane >= 24) return false;
    const QuantParams &q = a.q;
    const bool luma = lane < 16;
    // maxima per coefficient class (0: both indices even, 2: both odd, 1: mixed); a chroma block's DC goes the 2x2 way.  Written out so
    // that every index is a compile-time constant (the block stays in registers).
    const int dc = luma ? iabs(L.d[0]) : 0;
    const int m0 = imax(imax(dc, iabs(L.d[2])), imax(iabs(L.d[8]), iabs(L.d[10])));
    const int m2 = imax(imax(iabs(L.d[5]), iabs(L.d[7])), imax(iabs(L.d[13]), iabs(L.d[15])));
    const int m1 = imax(imax(imax(iabs(L.d[1]), iabs(L.d[3])), imax(iabs(L.d[4]), iabs(L.d[6]))),
                        imax(imax(iabs(L.d[9]), iabs(L.d[11])), imax(iabs(L.d[12]), iabs(L.d[14]))));
    bool any = m0 >= (luma ? q.thr_y[0] : q.thr_c[0]) || m1 >= (luma ? q.thr_y[1] : q.thr_c[1]) || m2 >= (luma ? q.thr_y[2] : q.thr_c[2]);
    if (!luma) {
        const int p = (lane - 16) >> 2;
        const int d0 = sm.cdc[4 * p], d1 = sm.cdc[4 * p + 1], d2 = sm.cdc[4 * p + 2], d3 = sm.cdc[4 * p + 3];
        const int mdc = imax(imax(iabs(d0 + d1 + d2 + d3), iabs(d0 - d1 + d2 - d3)), imax(iabs(d0 + d1 - d2 - d3), iabs(d0 - d1 - d2 + d3)));
        any = any || mdc >= q.thr_cdc;
    }
    return any;
}
GVC_HD void enc_mb_p2_zero(int lane, const EncMbArgs &a, bool al, int mx, int my, EncMbSmem &sm, EncLane &L) {
    if (lane >= 24) return;
    const bool luma = lane < 16;
    int16_t *mbco = a.coefs ? a.coefs + (long long)(my * a.mbw + mx) * kMbCoefs : nullptr;
    for (int i = 0; i < 8; ++i) L.lev[i] = 0;
    if (!luma && ((lane - 16) & 3) == 0) {
        const int p = (lane - 16) >> 2;
        for (int i = 0; i < 4; ++i) sm.dclev[4 * p + i] = 0;
        if (mbco) for (int i = 0; i < 4; ++i) mbco[24 * 16 + 4 * p + i] = 0;
    }
    if (a.nnz_y) {
        if (luma) a.nnz_y[(4 * my + (lane >> 2)) * (4 * a.mbw) + 4 * mx + (lane & 3)] = 0;
        else { const int b = (lane - 16) & 3; (lane < 20 ? a.nnz_u : a.nnz_v)[(2 * my + (b >> 1)) * (2 * a.mbw) + 2 * mx + (b & 1)] = 0; }
    }
    if (mbco) {
        if (al) {
            gvc_uint4 *co = reinterpret_cast<gvc_uint4 *>(mbco + lane * 16);
            gvc_uint4 z; z.x = z.y = z.z = z.w = 0;
            co[0] = z; co[1] = z;
        } else {
            for (int i = 0; i < 16; ++i) mbco[lane * 16 + i] = 0;
        }
    }
    int x, y, k;
    enc_lane_block(lane, mx, my, x, y, k);
    const int st = luma ? a.sy : a.sc;
    uint8_t *rec = (luma ? a.rec_y : lane < 20 ? a.rec_u : a.rec_v) + (y * st + x);
    for (int j = 0; j < 4; ++j) {
        const unsigned w = L.pred[j];
        if (al) *reinterpret_cast<unsigned *>(rec + j * st) = w;
        else for (int i = 0; i < 4; ++i) rec[j * st + i] = (uint8_t)(w >> (8 * i));
    }
}
// Phase 2: chroma DC path (2x2 transform of the plane's four DCs; every lane of the plane recomputes it and keeps element
// b), quantisation, levels, dequantisation, inverse transform, reconstruction.
GVC_HD void enc_mb_p2(int lane, const EncMbArgs &a, bool al, int mx, int my, EncMbSmem &sm, EncLane &L) {
    if (lane >= 24) return;
    const QuantParams &q = a.q;
    const bool luma = lane < 16;
    int16_t *mbco = a.coefs ? a.coefs + (long long)(my * a.mbw + mx) * kMbCoefs : nullptr;
    int mydc = 0;
    if (!luma) {
        const int l = lane - 16, p = l >> 2, b = l & 3;
        const int d0 = sm.cdc[4 * p], d1 = sm.cdc[4 * p + 1], d2 = sm.cdc[4 * p + 2], d3 = sm.cdc[4 * p + 3];
        int dc[4] = {d0 + d1 + d2 + d3, d0 - d1 + d2 - d3, d0 + d1 - d2 - d3, d0 - d1 - d2 + d3};
        int nzdc = 0;
        for (int i = 0; i < 4; ++i) { dc[i] = quant_one(dc[i], q.mf_c[0] >> 1, q.bias_c[0] << 1); nzdc |= dc[i]; }
        if (b == 0) {
            for (int i = 0; i < 4; ++i) sm.dclev[4 * p + i] = dc[i];
            if (mbco) for (int i = 0; i < 4; ++i) mbco[24 * 16 + 4 * p + i] = (int16_t)dc[i];
            if (nzdc) smem_or(&sm.mask, 1u << (24 + p));
        }
        const int f = b == 0 ? dc[0] + dc[1] + dc[2] + dc[3] : b == 1 ? dc[0] - dc[1] + dc[2] - dc[3] : b == 2 ? dc[0] + dc[1] - dc[2] - dc[3] : dc[0] - dc[1] - dc[2] + dc[3];
        mydc = (int)(int16_t)((f * (q.dq_c[0] << (q.qpc / 6))) >> 5);
        L.d[0] = 0;
    }
    GVC_RECONVERGE(0x00ffffffu);
    int mf[3], bias[3], dq[3];
    for (int c = 0; c < 3; ++c) { mf[c] = luma ? q.mf_y[c] : q.mf_c[c]; bias[c] = luma ? q.bias_y[c] : q.bias_c[c]; dq[c] = luma ? q.dq_y[c] : q.dq_c[c]; }
    const int qp = luma ? q.qp : q.qpc;
    int nz = 0;
    for (int i = 0; i < 16; ++i) { L.d[i] = quant_one(L.d[i], mf[coef_class(i)], bias[coef_class(i)]); nz |= L.d[i]; }
    for (int i = 0; i < 8; ++i) L.lev[i] = ((unsigned)L.d[zigzag4(2 * i)] & 0xffffu) | ((unsigned)L.d[zigzag4(2 * i + 1)] << 16);
    if (nz) smem_or(&sm.mask, 1u << lane);
    store_block_nnz(lane, mx, my, a.mbw, a.nnz_y, a.nnz_u, a.nnz_v, L.d);
    if (mbco) {
        if (al) {
            gvc_uint4 *co = reinterpret_cast<gvc_uint4 *>(mbco + lane * 16);
            gvc_uint4 v0, v1;
            v0.x = L.lev[0]; v0.y = L.lev[1]; v0.z = L.lev[2]; v0.w = L.lev[3]; v1.x = L.lev[4]; v1.y = L.lev[5]; v1.z = L.lev[6]; v1.w = L.lev[7];
            co[0] = v0; co[1] = v1;
        } else {
            for (int i = 0; i < 16; ++i) mbco[lane * 16 + i] = (int16_t)(L.lev[i >> 1] >> (16 * (i & 1)));
        }
    }
    GVC_RECONVERGE(0x00ffffffu);
    for (int i = 0; i < 16; ++i) L.d[i] = dequant_one(L.d[i], dq[coef_class(i)], qp);
    if (!luma) L.d[0] = mydc;
    idct4x4(L.d);
    int x, y, k;
    enc_lane_block(lane, mx, my, x, y, k);
    const int st = luma ? a.sy : a.sc;
    uint8_t *rec = (luma ? a.rec_y : lane < 20 ? a.rec_u : a.rec_v) + (y * st + x);
    for (int j = 0; j < 4; ++j) {
        const unsigned w = pack_sat_u8x4(byte_of(L.pred[j], 0) + L.d[j * 4], byte_of(L.pred[j], 1) + L.d[j * 4 + 1],
                                         byte_of(L.pred[j], 2) + L.d[j * 4 + 2], byte_of(L.pred[j], 3) + L.d[j * 4 + 3]);
        if (al) *reinterpret_cast<unsigned *>(rec + j * st) = w;
        else for (int i = 0; i < 4; ++i) rec[j * st + i] = (uint8_t)(w >> (8 * i));
    }
}
// Phase 3: lane 0 publishes the mask and reserves room for the packed blocks; lanes 24..26 clear the unused tail of the record.
GVC_HD void enc_mb_p3(int lane, const EncMbArgs &a, int mx, int my, EncMbSmem &sm) {
    const int mb = my * a.mbw + mx;
    if (lane == 0) {
        a.nzmask[mb] = sm.mask;
        pack_reserve(lane, a.pack, mb, sm.mask, sm.flags);
    } else if (lane >= 24 && lane < 27 && a.coefs) {
        int16_t *z = a.coefs + (long long)mb * kMbCoefs + 24 * 16 + 8 * (lane - 23);   // 8 int16 each: [24][8..15], [25][0..7], [25][8..15]
        for (int i = 0; i < 8; ++i) z[i] = 0;
    }
}
// Phase 4: present blocks go straight from the lanes' registers to their packed place in host-mapped memory.
GVC_HD void enc_mb_p4(int lane, const EncMbArgs &a, int mx, int my, const EncMbSmem &sm, const EncLane &L) {
    if (lane >= 25 || !a.pack.blocks) return;
    const unsigned cm = (unsigned)sm.flags[30];
    if (!((cm >> lane) & 1u)) return;
    const int rank = popcount32(cm & ((1u << lane) - 1u));
    gvc_uint4 *dst = reinterpret_cast<gvc_uint4 *>(a.pack.blocks + ((long long)(unsigned)sm.flags[31] + rank) * 16);
    gvc_uint4 v0, v1;
    if (lane < 24) {
        v0.x = L.lev[0]; v0.y = L.lev[1]; v0.z = L.lev[2]; v0.w = L.lev[3]; v1.x = L.lev[4]; v1.y = L.lev[5]; v1.z = L.lev[6]; v1.w = L.lev[7];
    } else {
        v0.x = ((unsigned)sm.dclev[0] & 0xffffu) | ((unsigned)sm.dclev[1] << 16); v0.y = ((unsigned)sm.dclev[2] & 0xffffu) | ((unsigned)sm.dclev[3] << 16);
        v0.z = ((unsigned)sm.dclev[4] & 0xffffu) | ((unsigned)sm.dclev[5] << 16); v0.w = ((unsigned)sm.dclev[6] & 0xffffu) | ((unsigned)sm.dclev[7] << 16);
        v1.x = v1.y = v1.z = v1.w = 0;
    }
    dst[0] = v0; dst[1] = v1;
}
// ------------------------------------------------------------------ K7 all-Intra16x16 frame (IDR) -
// Specification: oracle/h264_mb.c gvc_oracle_encode_intra_frame (H.264 8.3.3, 8.3.4, 8.5.2, 8.5.10).  One warp
// per macroblock ROW walks its row left to right; a row may start macroblock mx once the row above has
// finished macroblock mx (vertical / DC prediction read its bottom samples).  The hand-over is a mailbox per macroblock
// in global memory (zeroed before the launch): eight 64-bit words, each carrying four bottom-row samples and a nonzero
// tag, stored by the lanes that own those samples.  A 64-bit store is single-copy atomic, so a consumer lane that
// sees the tag has the samples in the same load: no fence on either side and one L2 round trip per dependent step.  All
// rows are co-resident (<= 148 CTAs of one warp, dispatched in row order).
struct IntraArgs {
    const uint8_t *cur_y, *cur_u, *cur_v;
    uint8_t       *rec_y, *rec_u, *rec_v;
    int            sy, sc, mbw, mbh;
    QuantParams    q;          // intra dead zone
    int16_t       *coefs;
    uint32_t      *nzmask;
    uint8_t       *modes;
    unsigned long long *mail;  // [mbh*mbw][8] bottom-row mailboxes (zeroed before the launch), see intra_p0
    PackOut        pack;
    uint8_t       *nnz_y, *nnz_u, *nnz_v;   // optional, as in EncMbArgs (AC levels only)
};
struct IntraSmem {
    unsigned topw[4], leftw[4];        // reconstructed neighbours of the luma MB: the row above, the column to the left (4 samples per word)
    unsigned ctopw[2][2], cleftw[2][2];
    int      sad[3];                   // SAD of the three candidate modes
    int      ldc[16], t1[16], dcl[16], t2[16];   // luma DC path: DCs, half-transformed, levels, half-inverse
    int      cdc[8], dclev[8];         // chroma DCs / their levels (Cb[0..3], Cr[4..7])
    unsigned mask;
    int      flags[32];                // [30], [31]: compact mask and first packed block (pack_reserve)
};
struct IntraLane {                     // per-lane registers carried across the phases
    unsigned cur[4], pred[4];          // source / prediction rows of the lane's 4x4 block
    int      d[16];
    unsigned lev[8];
};
GVC_HDH bool intra_aligned(const IntraArgs &a) {
    return (((unsigned long long)a.cur_y | (unsigned long long)a.cur_u | (unsigned long long)a.cur_v | (unsigned long long)a.rec_y |
             (unsigned long long)a.rec_u | (unsigned long long)a.rec_v | (unsigned long long)(unsigned)a.sy |
             (unsigned long long)(unsigned)a.sc) & 3) == 0 && ((unsigned long long)a.coefs & 15) == 0;
}
GVC_HD void smem_add(int *p, int v) {
#if defined(GVC_EMU)
    *p += v;
#else
    atomicAdd(p, v);
#endif
}
GVC_HD int sum_u8x4(unsigned w) { return dp4a_us(w, 0x01010101, 0); }
GVC_HD unsigned rep_u8(int v) { return (unsigned)v * 0x01010101u; }
GVC_HD unsigned mail_wait(const unsigned long long *slot) {      // spins until the producer's tagged word has arrived
#if defined(GVC_EMU)
    return (unsigned)*slot;
#else
    unsigned long long v;
    do { v = *reinterpret_cast<const volatile unsigned long long *>(slot); } while ((v >> 32) == 0);
    return (unsigned)v;
#endif
}
GVC_HD void mail_post(unsigned long long *slot, unsigned data) {
#if defined(GVC_EMU)
    *slot = (1ull << 32) | data;
#else
    *reinterpret_cast<volatile unsigned long long *>(slot) = (1ull << 32) | data;
#endif
}
// the lane's four source rows of macroblock (mx, my)
GVC_HD void intra_load_cur(int lane, const IntraArgs &a, bool al, int mx, int my, unsigned out[4]) {
    if (lane >= 24) return;
    int x, y, k;
    enc_lane_block(lane, mx, my, x, y, k);
    const int st = lane < 16 ? a.sy : a.sc;
    const uint8_t *src = (lane < 16 ? a.cur_y : lane < 20 ? a.cur_u : a.cur_v) + (y * st + x);
    for (int j = 0; j < 4; ++j) out[j] = al ? *reinterpret_cast<const unsigned *>(src + j * st) : ld_u32_any(src + j * st);
}
// phase 0: the row above arrives through the mailbox of the macroblock above (the column to the left is still in shared
// memory from the previous macroblock of this row)
GVC_HD void intra_p0(int lane, const IntraArgs &a, int mx, int my, IntraSmem &sm) {
    if (lane == 0) { sm.sad[0] = sm.sad[1] = sm.sad[2] = 0; sm.mask = 0; }
    if (my > 0 && lane < 8) {
        const unsigned w = mail_wait(a.mail + ((long long)((my - 1) * a.mbw + mx) * 8 + lane));
        if (lane < 4) sm.topw[lane] = w;
        else          sm.ctopw[(lane - 4) >> 1][lane & 1] = w;
    }
}
GVC_HD int intra_dc_value(const IntraSmem &sm, bool has_top, bool has_left) {
    int sum = 0;
    for (int i = 0; i < 4; ++i) sum += (has_top ? sum_u8x4(sm.topw[i]) : 0) + (has_left ? sum_u8x4(sm.leftw[i]) : 0);
    return has_top && has_left ? (sum + 16) >> 5 : (has_top || has_left) ? (sum + 8) >> 4 : 128;
}
// prediction rows of luma block (bx, by) for a mode: 0 vertical, 1 horizontal, 2 DC
GVC_HD void intra_pred_rows(const IntraSmem &sm, int mode, int dcv, int bx, int by, unsigned pr[4]) {
    const unsigned lw = sm.leftw[by >> 2];
    for (int j = 0; j < 4; ++j) pr[j] = mode == 0 ? sm.topw[bx >> 2] : mode == 1 ? rep_u8(byte_of(lw, j)) : rep_u8(dcv);
}
// phase 1: SAD of the three candidate modes over the lane's block, summed per mode in the shared record
GVC_HD void intra_p1(int lane, const IntraArgs &a, int mx, int my, IntraSmem &sm, const IntraLane &L) {
    if (lane >= 16) return;
    const bool has_top = my > 0, has_left = mx > 0;
    const int  dcv = intra_dc_value(sm, has_top, has_left);
    const int  bx = 4 * (lane & 3), by = 4 * (lane >> 2);
    for (int m = 0; m < 3; ++m) {
        if ((m == 0 && !has_top) || (m == 1 && !has_left)) continue;
        unsigned pr[4], s = 0;
        intra_pred_rows(sm, m, dcv, bx, by, pr);
        for (int j = 0; j < 4; ++j) s = sad_u8x4(L.cur[j], pr[j], s);
        smem_add(&sm.sad[m], (int)s);
    }
}
GVC_HD int intra_pick_mode(const IntraSmem &sm, bool has_top, bool has_left) {
    int best = -1, mode = 2;
    for (int m = 0; m < 3; ++m) {
        if ((m == 0 && !has_top) || (m == 1 && !has_left)) continue;
        if (best < 0 || sm.sad[m] < best) { best = sm.sad[m]; mode = m; }
    }
    return mode;
}
// phase 2: prediction, residual, forward transform of every block; DCs to the shared record
GVC_HD void intra_p2(int lane, const IntraArgs &a, int mx, int my, IntraSmem &sm, IntraLane &L) {
    if (lane >= 24) return;
    const bool has_top = my > 0, has_left = mx > 0;
    if (lane < 16) {
        const int mode = intra_pick_mode(sm, has_top, has_left), dcv = intra_dc_value(sm, has_top, has_left);
        intra_pred_rows(sm, mode, dcv, 4 * (lane & 3), 4 * (lane >> 2), L.pred);
        if (lane == 0) a.modes[my * a.mbw + mx] = (uint8_t)mode;
    } else {
        const int l = lane - 16, p = l >> 2, b = l & 3, bx = b & 1, by = b >> 1;
        const int st[2] = {sum_u8x4(sm.ctopw[p][0]), sum_u8x4(sm.ctopw[p][1])}, sl[2] = {sum_u8x4(sm.cleftw[p][0]), sum_u8x4(sm.cleftw[p][1])};
        int v;
        if (bx == by)     v = has_top && has_left ? (st[bx] + sl[by] + 4) >> 3 : has_top ? (st[bx] + 2) >> 2 : has_left ? (sl[by] + 2) >> 2 : 128;
        else if (bx == 1) v = has_top ? (st[1] + 2) >> 2 : has_left ? (sl[0] + 2) >> 2 : 128;
        else              v = has_left ? (sl[1] + 2) >> 2 : has_top ? (st[0] + 2) >> 2 : 128;
        for (int j = 0; j < 4; ++j) L.pred[j] = rep_u8(v);
    }
    GVC_RECONVERGE(0x00ffffffu);
    for (int j = 0; j < 4; ++j)
        for (int i = 0; i < 4; ++i) L.d[j * 4 + i] = byte_of(L.cur[j], i) - byte_of(L.pred[j], i);
    dct4x4(L.d);
    if (lane < 16) { sm.ldc[lane] = L.d[0]; L.d[0] = 0; }
    else           { sm.cdc[lane - 16] = L.d[0]; }
}
// 4-point butterfly of (u0..u3), output k: the 1-D pass of the luma DC Hadamard (x264 dct4x4dc ordering)
GVC_HD int had4_out(int u0, int u1, int u2, int u3, int k) {
    const int s01 = u0 + u1, d01 = u0 - u1, s23 = u2 + u3, d23 = u2 - u3;
    return k == 0 ? s01 + s23 : k == 1 ? s01 - s23 : k == 2 ? d01 - d23 : d01 + d23;
}
// phases 3a-3c: the 4x4 Hadamard of the sixteen luma DCs, one element per lane and one dimension per phase
GVC_HD void intra_p3a(int lane, IntraSmem &sm) {
    if (lane >= 16) return;
    const int r = lane >> 2, c = lane & 3;
    sm.t1[lane] = had4_out(sm.ldc[4 * r], sm.ldc[4 * r + 1], sm.ldc[4 * r + 2], sm.ldc[4 * r + 3], c);
}
GVC_HD void intra_p3b(int lane, const IntraArgs &a, IntraSmem &sm) {
    if (lane >= 16) return;
    const int r = lane >> 2, c = lane & 3;
    const int o = (had4_out(sm.t1[c], sm.t1[4 + c], sm.t1[8 + c], sm.t1[12 + c], r) + 1) >> 1;
    const int lev = quant_one((int)(int16_t)o, a.q.mf_y[0] >> 1, a.q.bias_y[0] << 1);
    sm.dcl[lane] = lev;
    if (lev) smem_or(&sm.mask, 1u << 26);
}
GVC_HD void intra_p3c(int lane, IntraSmem &sm) {
    if (lane >= 16) return;
    const int r = lane >> 2, c = lane & 3;
    sm.t2[lane] = had4_out(sm.dcl[4 * r], sm.dcl[4 * r + 1], sm.dcl[4 * r + 2], sm.dcl[4 * r + 3], c);
}
// phase 3d: DC values back, quantisation, levels, dequantisation, inverse transform, reconstruction; the right-hand
// column of the macroblock stays in shared memory for the next macroblock of the row
GVC_HD void intra_p3d(int lane, const IntraArgs &a, bool al, int mx, int my, IntraSmem &sm, IntraLane &L) {
    if (lane >= 24) return;
    const QuantParams &q = a.q;
    const bool luma = lane < 16;
    int16_t *mbco = a.coefs ? a.coefs + (long long)(my * a.mbw + mx) * kMbCoefs : nullptr;
    int mydc;
    if (luma) {
        const int r = lane >> 2, c = lane & 3;
        const int mine = (int)(int16_t)had4_out(sm.t2[c], sm.t2[4 + c], sm.t2[8 + c], sm.t2[12 + c], r);
        const int qbits = q.qp / 6 - 6;
        mydc = (int)(int16_t)(qbits >= 0 ? mine * (q.dq_y[0] << qbits) : (mine * q.dq_y[0] + (1 << (-qbits - 1))) >> (-qbits));
        if (lane == 0 && mbco)
            for (int i = 0; i < 16; ++i) mbco[25 * 16 + i] = (int16_t)sm.dcl[zigzag4(i)];
    } else {
        const int l = lane - 16, p = l >> 2, b = l & 3;
        const int d0 = sm.cdc[4 * p], d1 = sm.cdc[4 * p + 1], d2 = sm.cdc[4 * p + 2], d3 = sm.cdc[4 * p + 3];
        int dc[4] = {d0 + d1 + d2 + d3, d0 - d1 + d2 - d3, d0 + d1 - d2 - d3, d0 - d1 - d2 + d3};
        int nzdc = 0;
        for (int i = 0; i < 4; ++i) { dc[i] = quant_one(dc[i], q.mf_c[0] >> 1, q.bias_c[0] << 1); nzdc |= dc[i]; }
        if (b == 0) {
            for (int i = 0; i < 4; ++i) sm.dclev[4 * p + i] = dc[i];
            if (mbco) for (int i = 0; i < 4; ++i) mbco[24 * 16 + 4 * p + i] = (int16_t)dc[i];
            if (nzdc) smem_or(&sm.mask, 1u << (24 + p));
        }
        const int f = b == 0 ? dc[0] + dc[1] + dc[2] + dc[3] : b == 1 ? dc[0] - dc[1] + dc[2] - dc[3] : b == 2 ? dc[0] + dc[1] - dc[2] - dc[3] : dc[0] - dc[1] - dc[2] + dc[3];
        mydc = (int)(int16_t)((f * (q.dq_c[0] << (q.qpc / 6))) >> 5);
        L.d[0] = 0;
    }
    GVC_RECONVERGE(0x00ffffffu);
    int mf[3], bias[3], dq[3];
    for (int c = 0; c < 3; ++c) { mf[c] = luma ? q.mf_y[c] : q.mf_c[c]; bias[c] = luma ? q.bias_y[c] : q.bias_c[c]; dq[c] = luma ? q.dq_y[c] : q.dq_c[c]; }
    const int qp = luma ? q.qp : q.qpc;
    int nz = 0;
    for (int i = 0; i < 16; ++i) { L.d[i] = quant_one(L.d[i], mf[coef_class(i)], bias[coef_class(i)]); nz |= L.d[i]; }
    for (int i = 0; i < 8; ++i) L.lev[i] = ((unsigned)L.d[zigzag4(2 * i)] & 0xffffu) | ((unsigned)L.d[zigzag4(2 * i + 1)] << 16);
    if (nz) smem_or(&sm.mask, 1u << lane);
    store_block_nnz(lane, mx, my, a.mbw, a.nnz_y, a.nnz_u, a.nnz_v, L.d);
    if (mbco) {
        if (al) {
            gvc_uint4 *co = reinterpret_cast<gvc_uint4 *>(mbco + lane * 16);
            gvc_uint4 v0, v1;
            v0.x = L.lev[0]; v0.y = L.lev[1]; v0.z = L.lev[2]; v0.w = L.lev[3]; v1.x = L.lev[4]; v1.y = L.lev[5]; v1.z = L.lev[6]; v1.w = L.lev[7];
            co[0] = v0; co[1] = v1;
        } else {
            for (int i = 0; i < 16; ++i) mbco[lane * 16 + i] = (int16_t)(L.lev[i >> 1] >> (16 * (i & 1)));
        }
        if (lane == 16) for (int i = 8; i < 16; ++i) mbco[24 * 16 + i] = 0;
    }
    GVC_RECONVERGE(0x00ffffffu);
    for (int i = 0; i < 16; ++i) L.d[i] = dequant_one(L.d[i], dq[coef_class(i)], qp);
    L.d[0] = mydc;
    idct4x4(L.d);
    int x, y, k;
    enc_lane_block(lane, mx, my, x, y, k);
    const int st = luma ? a.sy : a.sc;
    uint8_t *rec = (luma ? a.rec_y : lane < 20 ? a.rec_u : a.rec_v) + (y * st + x);
    unsigned right = 0, bottom = 0;
    for (int j = 0; j < 4; ++j) {
        const unsigned w = pack_sat_u8x4(byte_of(L.pred[j], 0) + L.d[j * 4], byte_of(L.pred[j], 1) + L.d[j * 4 + 1],
                                         byte_of(L.pred[j], 2) + L.d[j * 4 + 2], byte_of(L.pred[j], 3) + L.d[j * 4 + 3]);
        if (al) *reinterpret_cast<unsigned *>(rec + j * st) = w;
        else for (int i = 0; i < 4; ++i) rec[j * st + i] = (uint8_t)(w >> (8 * i));
        right |= (w >> 24) << (8 * j);
        bottom = w;
    }
    if (luma) { if ((lane & 3) == 3) sm.leftw[lane >> 2] = right; }
    else      { const int l = lane - 16, p = l >> 2, b = l & 3; if (b & 1) sm.cleftw[p][b >> 1] = right; }
    if (my + 1 < a.mbh) {       // the macroblock below waits for these
        unsigned long long *mail = a.mail + (long long)(my * a.mbw + mx) * 8;
        if (luma) { if (lane >= 12) mail_post(mail + (lane & 3), bottom); }
        else      { const int l = lane - 16, p = l >> 2, b = l & 3; if (b >= 2) mail_post(mail + 4 + 2 * p + (b & 1), bottom); }
    }
}
// phase 4: mask out; room for the packed blocks comes from the row's own region of the picture's block area (row my starts
// at block my * mbw * 26), so the wavefront needs no atomic: lane 0 keeps the row's running count
GVC_HD void intra_p4(int lane, const IntraArgs &a, int mx, int my, IntraSmem &sm, int &row_count) {
    if (lane != 0) return;
    const int mb = my * a.mbw + mx;
    a.nzmask[mb] = sm.mask;
    if (!a.pack.blocks) return;
    const unsigned cm = compact_mask(sm.mask), base = (unsigned)(my * a.mbw * 26 + row_count);
    row_count += popcount32(cm);
    a.pack.offsets[mb] = base;
    sm.flags[30] = (int)cm;
    sm.flags[31] = (int)base;
}
GVC_HD void intra_row_done(int lane, const IntraArgs &a, int row_count) {   // the picture's block count, once per row
    if (lane != 0 || !a.pack.blocks) return;
#if defined(GVC_EMU)
    *a.pack.counter += (unsigned)row_count;
#else
    atomicAdd(a.pack.counter, (unsigned)row_count);
#endif
}
// phase 5: present blocks from registers / the shared record to their packed place
GVC_HD void intra_p5(int lane, const IntraArgs &a, int mx, int my, const IntraSmem &sm, const IntraLane &L) {
    if (lane >= 26 || !a.pack.blocks) return;
    const unsigned cm = (unsigned)sm.flags[30];
    if (!((cm >> lane) & 1u)) return;
    const int rank = popcount32(cm & ((1u << lane) - 1u));
    gvc_uint4 *dst = reinterpret_cast<gvc_uint4 *>(a.pack.blocks + ((long long)(unsigned)sm.flags[31] + rank) * 16);
    unsigned w[8];
    if (lane < 24) {
        for (int i = 0; i < 8; ++i) w[i] = L.lev[i];
    } else if (lane == 24) {
        for (int i = 0; i < 4; ++i) w[i] = ((unsigned)sm.dclev[2 * i] & 0xffffu) | ((unsigned)sm.dclev[2 * i + 1] << 16);
        for (int i = 4; i < 8; ++i) w[i] = 0;
    } else {
        for (int i = 0; i < 8; ++i) w[i] = ((unsigned)sm.dcl[zigzag4(2 * i)] & 0xffffu) | ((unsigned)sm.dcl[zigzag4(2 * i + 1)] << 16);
    }
    gvc_uint4 v0, v1;
    v0.x = w[0]; v0.y = w[1]; v0.z = w[2]; v0.w = w[3]; v1.x = w[4]; v1.y = w[5]; v1.z = w[6]; v1.w = w[7];
    dst[0] = v0; dst[1] = v1;
}

}  // namespace gvc
