/*
 * h264_core.c -- CPU oracle (test infrastructure, not product): the per-macroblock arithmetic that
 * libx264 performs when the reference runs `ffmpeg ... -vcodec libx264` (client.go:115, args from
 * server.go:69-70).  libx264 is a third-party dependency that is NOT in /root/reference and not
 * pinned by it; no x264 source, library or binary exists in this environment.  This file restates
 * the published algorithms -- ITU-T H.264 (04/2017) clauses 8.5.10-8.5.13 (transforms, scaling),
 * 8.4.2.2 (fractional sample interpolation) -- in x264's function decomposition (common/dct.c,
 * common/quant.c, common/pixel.c, common/mc.c; encoder/set.c for the quantiser tables).
 *
 * PARITY UNPINNED against libx264 itself.  Pinned by: spec known-answer tests
 * (tests/test_oracle_h264.py) and, for everything a decoder can observe (dequant, inverse
 * transforms, interpolation, prediction), by bit-exact agreement of the encoder's reconstruction
 * with FFmpeg's h264 decoder on the streams we emit (tests/test_tier3_decode.py).
 *
 * Layout convention: coefficient blocks are spec raster order, c[row*N + col] with row = vertical
 * frequency (x264 stores the transpose and compensates in its scan tables).
 */
#include "gvc_oracle.h"

#include <stdlib.h>
#include <string.h>

static int clip_u8(int v) { return v < 0 ? 0 : v > 255 ? 255 : v; }

/* ------------------------------------------------------------------ tables ------------------ */
/* H.264 Table 8-? normAdjust4x4 / LevelScale (flat matrices): v[qp%6][class], class = (row&1)+(col&1) */
static const uint8_t  k_dequant4[6][3] = {{10, 13, 16}, {11, 14, 18}, {13, 16, 20}, {14, 18, 23}, {16, 20, 25}, {18, 23, 29}};
static const uint16_t k_quant4[6][3]   = {{13107, 8066, 5243}, {11916, 7490, 4660}, {10082, 6554, 4194},
                                          {9362, 5825, 3647},  {8192, 5243, 3355},  {7282, 4559, 2893}};
/* 8x8: class = k_scan8[(row&3)*4 + (col&3)] */
static const uint8_t  k_scan8[16]      = {0, 3, 4, 3, 3, 1, 5, 1, 4, 5, 2, 5, 3, 1, 5, 1};
static const uint8_t  k_dequant8[6][6] = {{20, 18, 32, 19, 25, 24}, {22, 19, 35, 21, 28, 26}, {26, 23, 42, 24, 33, 31},
                                          {28, 25, 45, 26, 35, 33}, {32, 28, 51, 30, 40, 38}, {36, 32, 58, 34, 46, 43}};
static const uint16_t k_quant8[6][6]   = {{13107, 11428, 20972, 12222, 16777, 15481}, {11916, 10826, 19174, 11058, 14980, 14290},
                                          {10082, 8943, 15978, 9675, 12710, 11985},   {9362, 8228, 14913, 8931, 11984, 11259},
                                          {8192, 7346, 13159, 7740, 10486, 9777},     {7282, 6428, 11570, 6830, 9118, 8640}};

static int shift_round(int x, int s) { return s <= 0 ? x << -s : (x + (1 << (s - 1))) >> s; }
static int div_round(int n, int d) { return (n + (d >> 1)) / d; }

/* x264 encoder/set.c x264_cqm_init, flat matrices: mf = SHIFT(scale, qp/6 - 1);
 * bias = min(DIV(deadzone<<10, mf), (1<<15)/mf); deadzone 21 intra / 11 inter (of 64ths after >>16). */
void gvc_oracle_quant4_tables(int qp, int intra, uint16_t mf[16], uint16_t bias[16])
{
    const int dz = intra ? 21 : 11;
    for (int i = 0; i < 16; i++) {
        const int cls = ((i >> 2) & 1) + (i & 1);
        const int m   = shift_round(k_quant4[qp % 6][cls], qp / 6 - 1);
        int       b   = div_round(dz << 10, m);
        if (b > (1 << 15) / m) b = (1 << 15) / m;
        mf[i] = (uint16_t)m; bias[i] = (uint16_t)b;
    }
}
void gvc_oracle_quant8_tables(int qp, int intra, uint16_t mf[64], uint16_t bias[64])
{
    const int dz = intra ? 21 : 11;
    for (int i = 0; i < 64; i++) {
        const int cls = k_scan8[((i >> 3) & 3) * 4 + (i & 3)];
        const int m   = shift_round(k_quant8[qp % 6][cls], qp / 6);
        int       b   = div_round(dz << 10, m);
        if (b > (1 << 15) / m) b = (1 << 15) / m;
        mf[i] = (uint16_t)m; bias[i] = (uint16_t)b;
    }
}
int gvc_oracle_dequant4_scale(int qp, int idx) { return k_dequant4[qp % 6][((idx >> 2) & 1) + (idx & 1)] * 16; }
int gvc_oracle_dequant8_scale(int qp, int idx) { return k_dequant8[qp % 6][k_scan8[((idx >> 3) & 3) * 4 + (idx & 3)]] * 16; }

/* H.264 Table 8-15: QPc as a function of qPI (chroma_qp_index_offset = 0) */
int gvc_oracle_chroma_qp(int qp)
{
    static const uint8_t t[22] = {29, 30, 31, 32, 32, 33, 34, 34, 35, 35, 36, 36, 37, 37, 37, 38, 38, 38, 39, 39, 39, 39};
    if (qp < 0) qp = 0;
    if (qp > 51) qp = 51;
    return qp < 30 ? qp : t[qp - 30];
}

/* ------------------------------------------------------------------ forward transforms ------- */
/* x264 common/dct.c sub4x4_dct: core transform of (src - pred), rows then columns */
void gvc_oracle_sub4x4_dct(int16_t c[16], const uint8_t *src, int sstride, const uint8_t *pred, int pstride)
{
    int d[16], t[16];
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) d[y * 4 + x] = src[y * sstride + x] - pred[y * pstride + x];
    for (int y = 0; y < 4; y++) { /* horizontal */
        const int s03 = d[y * 4 + 0] + d[y * 4 + 3], s12 = d[y * 4 + 1] + d[y * 4 + 2];
        const int d03 = d[y * 4 + 0] - d[y * 4 + 3], d12 = d[y * 4 + 1] - d[y * 4 + 2];
        t[y * 4 + 0] = s03 + s12; t[y * 4 + 1] = 2 * d03 + d12; t[y * 4 + 2] = s03 - s12; t[y * 4 + 3] = d03 - 2 * d12;
    }
    for (int x = 0; x < 4; x++) { /* vertical */
        const int s03 = t[0 + x] + t[12 + x], s12 = t[4 + x] + t[8 + x];
        const int d03 = t[0 + x] - t[12 + x], d12 = t[4 + x] - t[8 + x];
        c[0 + x] = (int16_t)(s03 + s12); c[4 + x] = (int16_t)(2 * d03 + d12);
        c[8 + x] = (int16_t)(s03 - s12); c[12 + x] = (int16_t)(d03 - 2 * d12);
    }
}

#define DCT8_1D(S, D)                                                                   \
    {                                                                                   \
        const int s07 = S(0) + S(7), s16 = S(1) + S(6), s25 = S(2) + S(5), s34 = S(3) + S(4); \
        const int a0 = s07 + s34, a1 = s16 + s25, a2 = s07 - s34, a3 = s16 - s25;       \
        const int d07 = S(0) - S(7), d16 = S(1) - S(6), d25 = S(2) - S(5), d34 = S(3) - S(4); \
        const int a4 = d16 + d25 + (d07 + (d07 >> 1));                                  \
        const int a5 = d07 - d34 - (d25 + (d25 >> 1));                                  \
        const int a6 = d07 + d34 - (d16 + (d16 >> 1));                                  \
        const int a7 = d16 - d25 + (d34 + (d34 >> 1));                                  \
        D(0) = a0 + a1; D(1) = a4 + (a7 >> 2); D(2) = a2 + (a3 >> 1); D(3) = a5 + (a6 >> 2); \
        D(4) = a0 - a1; D(5) = a6 - (a5 >> 2); D(6) = (a2 >> 1) - a3; D(7) = (a4 >> 2) - a7; \
    }

/* x264 sub8x8_dct8 */
void gvc_oracle_sub8x8_dct8(int16_t c[64], const uint8_t *src, int sstride, const uint8_t *pred, int pstride)
{
    int d[64];
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) d[y * 8 + x] = src[y * sstride + x] - pred[y * pstride + x];
    for (int y = 0; y < 8; y++) {
#define S(i) d[y * 8 + (i)]
#define D(i) d[y * 8 + (i)]
        DCT8_1D(S, D)
#undef S
#undef D
    }
    for (int x = 0; x < 8; x++) {
#define S(i) d[(i)*8 + x]
#define D(i) d[(i)*8 + x]
        DCT8_1D(S, D)
#undef S
#undef D
    }
    for (int i = 0; i < 64; i++) c[i] = (int16_t)d[i];
}

/* 4x4 Hadamard butterflies on rows then columns: out = H * in * H^T, H rows (1,1,1,1),(1,1,-1,-1),(1,-1,-1,1),(1,-1,1,-1) */
static void hadamard4x4(const int16_t d[16], int o[16])
{
    int t[16];
    for (int i = 0; i < 4; i++) {
        const int s01 = d[i * 4 + 0] + d[i * 4 + 1], d01 = d[i * 4 + 0] - d[i * 4 + 1];
        const int s23 = d[i * 4 + 2] + d[i * 4 + 3], d23 = d[i * 4 + 2] - d[i * 4 + 3];
        t[i * 4 + 0] = s01 + s23; t[i * 4 + 1] = s01 - s23; t[i * 4 + 2] = d01 - d23; t[i * 4 + 3] = d01 + d23;
    }
    for (int j = 0; j < 4; j++) {
        const int s01 = t[0 + j] + t[4 + j], d01 = t[0 + j] - t[4 + j];
        const int s23 = t[8 + j] + t[12 + j], d23 = t[8 + j] - t[12 + j];
        o[0 + j] = s01 + s23; o[4 + j] = s01 - s23; o[8 + j] = d01 - d23; o[12 + j] = d01 + d23;
    }
}
/* x264 dct4x4dc: 4x4 Hadamard of the luma DCs of an Intra16x16 MB, (x+1)>>1 on the result */
void gvc_oracle_dct4x4dc(int16_t d[16])
{
    int o[16];
    hadamard4x4(d, o);
    for (int i = 0; i < 16; i++) d[i] = (int16_t)((o[i] + 1) >> 1);
}
/* x264 idct4x4dc: same butterflies, no rounding (H.264 8.5.10) */
void gvc_oracle_idct4x4dc(int16_t d[16])
{
    int o[16];
    hadamard4x4(d, o);
    for (int i = 0; i < 16; i++) d[i] = (int16_t)o[i];
}
/* x264 dct2x2dc / idct2x2 (self-inverse up to scale): d = [[a,b],[c,e]] raster */
void gvc_oracle_dct2x2dc(int16_t d[4])
{
    const int a = d[0] + d[1], b = d[0] - d[1], c = d[2] + d[3], e = d[2] - d[3];
    d[0] = (int16_t)(a + c); d[1] = (int16_t)(b + e); d[2] = (int16_t)(a - c); d[3] = (int16_t)(b - e);
}

/* ------------------------------------------------------------------ quant / dequant ---------- */
/* x264 common/quant.c QUANT_ONE: sign(c) * ((bias + |c|) * mf >> 16); returns nonzero flag */
int gvc_oracle_quant(int16_t *c, int n, const uint16_t *mf, const uint16_t *bias)
{
    int nz = 0;
    for (int i = 0; i < n; i++) {
        if (c[i] > 0) c[i] = (int16_t)((((uint32_t)bias[i] + (uint32_t)c[i]) * mf[i]) >> 16);
        else          c[i] = (int16_t)(-(int32_t)((((uint32_t)bias[i] + (uint32_t)(-c[i])) * mf[i]) >> 16));
        nz |= c[i];
    }
    return nz != 0;
}
/* DC variants use position-0 tables: quant_4x4_dc(mf[0]>>1, bias[0]<<1), quant_2x2_dc likewise */
int gvc_oracle_quant_dc(int16_t *c, int n, int mf, int bias)
{
    int nz = 0;
    for (int i = 0; i < n; i++) {
        if (c[i] > 0) c[i] = (int16_t)((((uint32_t)bias + (uint32_t)c[i]) * (uint32_t)mf) >> 16);
        else          c[i] = (int16_t)(-(int32_t)((((uint32_t)bias + (uint32_t)(-c[i])) * (uint32_t)mf) >> 16));
        nz |= c[i];
    }
    return nz != 0;
}
/* H.264 8.5.12.1 (4x4, non-DC rule applied to all 16) / x264 dequant_4x4 */
void gvc_oracle_dequant4(int16_t c[16], int qp)
{
    const int qbits = qp / 6 - 4;
    for (int i = 0; i < 16; i++) {
        const int s = gvc_oracle_dequant4_scale(qp, i);
        if (qbits >= 0) c[i] = (int16_t)((c[i] * s) * (1 << qbits));
        else            c[i] = (int16_t)((c[i] * s + (1 << (-qbits - 1))) >> (-qbits));
    }
}
/* H.264 8.5.13 / x264 dequant_8x8 */
void gvc_oracle_dequant8(int16_t c[64], int qp)
{
    const int qbits = qp / 6 - 6;
    for (int i = 0; i < 64; i++) {
        const int s = gvc_oracle_dequant8_scale(qp, i);
        if (qbits >= 0) c[i] = (int16_t)((c[i] * s) * (1 << qbits));
        else            c[i] = (int16_t)((c[i] * s + (1 << (-qbits - 1))) >> (-qbits));
    }
}
/* H.264 8.5.10 luma DC of Intra16x16 / x264 dequant_4x4_dc (input: after idct4x4dc) */
void gvc_oracle_dequant4_dc(int16_t c[16], int qp)
{
    const int qbits = qp / 6 - 6, s = gvc_oracle_dequant4_scale(qp, 0);
    for (int i = 0; i < 16; i++) {
        if (qbits >= 0) c[i] = (int16_t)(c[i] * (s << qbits));
        else            c[i] = (int16_t)((c[i] * s + (1 << (-qbits - 1))) >> (-qbits));
    }
}
/* H.264 8.5.11.2 chroma DC (4:2:0): ((f * LevelScale(qp%6,0,0)) << (qp/6)) >> 5, input after the 2x2 transform */
void gvc_oracle_dequant2x2_dc(int16_t c[4], int qp)
{
    const int s = gvc_oracle_dequant4_scale(qp, 0) << (qp / 6);
    for (int i = 0; i < 4; i++) c[i] = (int16_t)((c[i] * s) >> 5);
}

/* ------------------------------------------------------------------ inverse transforms ------- */
/* H.264 8.5.12.2 / x264 add4x4_idct: rows first, then columns, (x+32)>>6, add to pred, clip */
void gvc_oracle_add4x4_idct(uint8_t *dst, int dstride, const uint8_t *pred, int pstride, const int16_t c[16])
{
    int t[16];
    for (int y = 0; y < 4; y++) {
        const int s02 = c[y * 4 + 0] + c[y * 4 + 2], d02 = c[y * 4 + 0] - c[y * 4 + 2];
        const int s13 = c[y * 4 + 1] + (c[y * 4 + 3] >> 1), d13 = (c[y * 4 + 1] >> 1) - c[y * 4 + 3];
        t[y * 4 + 0] = s02 + s13; t[y * 4 + 1] = d02 + d13; t[y * 4 + 2] = d02 - d13; t[y * 4 + 3] = s02 - s13;
    }
    for (int x = 0; x < 4; x++) {
        const int s02 = t[0 + x] + t[8 + x], d02 = t[0 + x] - t[8 + x];
        const int s13 = t[4 + x] + (t[12 + x] >> 1), d13 = (t[4 + x] >> 1) - t[12 + x];
        const int r[4] = {s02 + s13, d02 + d13, d02 - d13, s02 - s13};
        for (int y = 0; y < 4; y++) dst[y * dstride + x] = (uint8_t)clip_u8(pred[y * pstride + x] + ((r[y] + 32) >> 6));
    }
}

#define IDCT8_1D(S, D)                                                        \
    {                                                                         \
        const int a0 = S(0) + S(4), a2 = S(0) - S(4);                         \
        const int a4 = (S(2) >> 1) - S(6), a6 = (S(6) >> 1) + S(2);           \
        const int b0 = a0 + a6, b2 = a2 + a4, b4 = a2 - a4, b6 = a0 - a6;     \
        const int a1 = -S(3) + S(5) - S(7) - (S(7) >> 1);                     \
        const int a3 = S(1) + S(7) - S(3) - (S(3) >> 1);                      \
        const int a5 = -S(1) + S(7) + S(5) + (S(5) >> 1);                     \
        const int a7 = S(3) + S(5) + S(1) + (S(1) >> 1);                      \
        const int b1 = (a7 >> 2) + a1, b3 = a3 + (a5 >> 2);                   \
        const int b5 = (a3 >> 2) - a5, b7 = a7 - (a1 >> 2);                   \
        D(0) = b0 + b7; D(1) = b2 + b5; D(2) = b4 + b3; D(3) = b6 + b1;       \
        D(4) = b6 - b1; D(5) = b4 - b3; D(6) = b2 - b5; D(7) = b0 - b7;       \
    }

/* H.264 8.5.13 / x264 add8x8_idct8 */
void gvc_oracle_add8x8_idct8(uint8_t *dst, int dstride, const uint8_t *pred, int pstride, const int16_t c[64])
{
    int t[64];
    for (int i = 0; i < 64; i++) t[i] = c[i];
    for (int y = 0; y < 8; y++) {
#define S(i) t[y * 8 + (i)]
#define D(i) t[y * 8 + (i)]
        IDCT8_1D(S, D)
#undef S
#undef D
    }
    for (int x = 0; x < 8; x++) {
#define S(i) t[(i)*8 + x]
#define D(i) t[(i)*8 + x]
        IDCT8_1D(S, D)
#undef S
#undef D
    }
    for (int y = 0; y < 8; y++)
        for (int x = 0; x < 8; x++) dst[y * dstride + x] = (uint8_t)clip_u8(pred[y * pstride + x] + ((t[y * 8 + x] + 32) >> 6));
}

/* ------------------------------------------------------------------ SAD / SATD --------------- */
/* x264 common/pixel.c pixel_sad_WxH */
int gvc_oracle_sad(const uint8_t *a, int as, const uint8_t *b, int bs, int w, int h)
{
    int s = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) s += abs(a[y * as + x] - b[y * bs + x]);
    return s;
}
/* sum of |4x4 Hadamard| of the difference block (always even) */
static int hadamard4_abs(const uint8_t *a, int as, const uint8_t *b, int bs)
{
    int d[16], t[16], s = 0;
    for (int y = 0; y < 4; y++)
        for (int x = 0; x < 4; x++) d[y * 4 + x] = a[y * as + x] - b[y * bs + x];
    for (int y = 0; y < 4; y++) {
        const int s01 = d[y * 4] + d[y * 4 + 1], d01 = d[y * 4] - d[y * 4 + 1];
        const int s23 = d[y * 4 + 2] + d[y * 4 + 3], d23 = d[y * 4 + 2] - d[y * 4 + 3];
        t[y * 4] = s01 + s23; t[y * 4 + 1] = s01 - s23; t[y * 4 + 2] = d01 + d23; t[y * 4 + 3] = d01 - d23;
    }
    for (int x = 0; x < 4; x++) {
        const int s01 = t[x] + t[4 + x], d01 = t[x] - t[4 + x];
        const int s23 = t[8 + x] + t[12 + x], d23 = t[8 + x] - t[12 + x];
        s += abs(s01 + s23) + abs(s01 - s23) + abs(d01 + d23) + abs(d01 - d23);
    }
    return s;
}
/* x264 pixel_satd_WxH: 4x4 Hadamard blocks, each 8x4 pair (or lone 4x4) contributes (sum >> 1) */
int gvc_oracle_satd(const uint8_t *a, int as, const uint8_t *b, int bs, int w, int h)
{
    int s = 0;
    for (int y = 0; y < h; y += 4)
        for (int x = 0; x < w; x += 8) {
            int v = hadamard4_abs(a + y * as + x, as, b + y * bs + x, bs);
            if (x + 4 < w) v += hadamard4_abs(a + y * as + x + 4, as, b + y * bs + x + 4, bs);
            s += v >> 1;
        }
    return s;
}

/* ------------------------------------------------------------------ reference-plane filters -- */
#define TAP6(p, d) ((p)[-2 * (d)] + (p)[3 * (d)] - 5 * ((p)[-(d)] + (p)[2 * (d)]) + 20 * ((p)[0] + (p)[(d)]))

/* x264 common/mc.c hpel_filter / H.264 8.4.2.2.1: half-sample planes of a luma plane.
 *   H(x,y) = sample at (x+1/2, y), V(x,y) = (x, y+1/2), C(x,y) = (x+1/2, y+1/2) from unrounded vertical sums.
 * src points at sample (0,0) of a plane that is readable for [-pad, w+pad) x [-pad, h+pad); outputs cover
 * x in [-(pad-3), w+pad-3), y likewise (the filters need 2 samples before and 3 after). */
void gvc_oracle_hpel_filter(uint8_t *dh, uint8_t *dv, uint8_t *dc, const uint8_t *src, int stride, int w, int h, int pad)
{
    const int m = pad - 3;
    int *buf = (int *)malloc(sizeof(int) * (size_t)(w + 2 * pad));
    for (int y = -m; y < h + m; y++) {
        const uint8_t *s = src + (ptrdiff_t)y * stride;
        for (int x = -m - 2; x < w + m + 3; x++) {
            const int v = TAP6(s + x, stride);
            buf[x + pad] = v;
            if (x >= -m && x < w + m) dv[(ptrdiff_t)y * stride + x] = (uint8_t)clip_u8((v + 16) >> 5);
        }
        for (int x = -m; x < w + m; x++) {
            dc[(ptrdiff_t)y * stride + x] = (uint8_t)clip_u8((TAP6(buf + pad + x, 1) + 512) >> 10);
            dh[(ptrdiff_t)y * stride + x] = (uint8_t)clip_u8((TAP6(s + x, 1) + 16) >> 5);
        }
    }
    free(buf);
}

/* x264 frame_init_lowres_core: four half-resolution planes (phases 0, H, V, C) by nested rounding averages.
 * src must be readable one sample past w and two rows past... exactly as x264: rows 2y..2y+2, cols 2x..2x+2. */
void gvc_oracle_lowres(const uint8_t *src, int sstride, uint8_t *d0, uint8_t *dh, uint8_t *dv, uint8_t *dc, int dstride,
                       int lw, int lh)
{
#define F4(a, b, c, d) (((((a) + (b) + 1) >> 1) + (((c) + (d) + 1) >> 1) + 1) >> 1)
    for (int y = 0; y < lh; y++) {
        const uint8_t *s0 = src + (ptrdiff_t)(2 * y) * sstride, *s1 = s0 + sstride, *s2 = s1 + sstride;
        for (int x = 0; x < lw; x++) {
            d0[y * dstride + x] = (uint8_t)F4(s0[2 * x], s1[2 * x], s0[2 * x + 1], s1[2 * x + 1]);
            dh[y * dstride + x] = (uint8_t)F4(s0[2 * x + 1], s1[2 * x + 1], s0[2 * x + 2], s1[2 * x + 2]);
            dv[y * dstride + x] = (uint8_t)F4(s1[2 * x], s2[2 * x], s1[2 * x + 1], s2[2 * x + 1]);
            dc[y * dstride + x] = (uint8_t)F4(s1[2 * x + 1], s2[2 * x + 1], s1[2 * x + 2], s2[2 * x + 2]);
        }
    }
#undef F4
}

/* ------------------------------------------------------------------ motion compensation ------ */
/* x264 mc_luma / get_ref, H.264 8.4.2.2.1 quarter-sample rule: each quarter position is the rounded
 * average of the two nearest integer/half samples.  planes[4] = {full, H, V, C} at sample (0,0);
 * (mvx, mvy) in quarter samples relative to block origin (bx, by). */
void gvc_oracle_mc_luma(uint8_t *dst, int dstride, const uint8_t *const planes[4], int stride, int bx, int by,
                        int mvx, int mvy, int w, int h)
{
    static const uint8_t ref0[16] = {0, 1, 1, 1, 0, 1, 1, 1, 2, 3, 3, 3, 0, 1, 1, 1};
    static const uint8_t ref1[16] = {0, 0, 1, 0, 2, 2, 3, 2, 2, 2, 3, 2, 2, 2, 3, 2};
    const int fx = mvx & 3, fy = mvy & 3, q = fy * 4 + fx;
    const ptrdiff_t off = (ptrdiff_t)(by + (mvy >> 2)) * stride + bx + (mvx >> 2);
    const uint8_t *s1 = planes[ref0[q]] + off + (fy == 3 ? stride : 0);
    if (q & 5) {
        const uint8_t *s2 = planes[ref1[q]] + off + (fx == 3 ? 1 : 0);
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) dst[y * dstride + x] = (uint8_t)((s1[y * stride + x] + s2[y * stride + x] + 1) >> 1);
    } else {
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++) dst[y * dstride + x] = s1[y * stride + x];
    }
}
/* x264 mc_chroma / H.264 8.4.2.2.2: eighth-sample bilinear; (mvx,mvy) is the LUMA vector (quarter luma samples);
 * (bx,by), w, h in chroma samples */
void gvc_oracle_mc_chroma(uint8_t *dst, int dstride, const uint8_t *src, int stride, int bx, int by, int mvx, int mvy,
                          int w, int h)
{
    const int dx = mvx & 7, dy = mvy & 7;
    const int cA = (8 - dx) * (8 - dy), cB = dx * (8 - dy), cC = (8 - dx) * dy, cD = dx * dy;
    const uint8_t *s = src + (ptrdiff_t)(by + (mvy >> 3)) * stride + bx + (mvx >> 3);
    for (int y = 0; y < h; y++, s += stride)
        for (int x = 0; x < w; x++)
            dst[y * dstride + x] = (uint8_t)((cA * s[x] + cB * s[x + 1] + cC * s[stride + x] + cD * s[stride + x + 1] + 32) >> 6);
}
