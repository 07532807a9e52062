/*
 * h264_mb.c -- CPU oracle (test infrastructure, not product): frame-level composition of the primitives in
 * h264_core.c -- the hierarchical motion search and the inter macroblock encode/reconstruct loop that the
 * CUDA kernels K4/K5/K6 implement.  Sequential, obvious C; it is the executable specification the GPU
 * results are compared with bit for bit (vectors, costs, coefficient levels, reconstruction).
 *
 * Relation to the reference: the reference reaches this work through `ffmpeg -vcodec libx264`
 * (client.go:115).  libx264's own search (encoder/me.c: predictor candidates + diamond/hex/UMH with early
 * exits) is inherently sequential per macroblock; north_star asks for "x264-semantics hierarchical motion
 * search (SAD/SATD over 16x16/8x8 blocks, integer-pel then sub-pel via the 6-tap half-pel filter)".  What is
 * kept from x264: the metrics (pixel_sad, pixel_satd), cost = metric + lambda*bits(mvd) with
 * x264_lambda_tab, half-then-quarter refinement on the hpel planes, lowres pre-search.  What is ours: the
 * exhaustive lowres window and the fixed candidate lists, chosen so every macroblock is independent.
 * PARITY UNPINNED against libx264 (absent from this environment); see gvc_oracle.h.
 */
#include "gvc_oracle.h"

#include <stdlib.h>
#include <string.h>

/* x264 encoder/analyse.c x264_lambda_tab */
static const uint8_t k_lambda[52] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 4, 4, 4,
                                     5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 23, 25, 29, 32, 36, 40, 45, 51, 57, 64, 72, 81, 91};
int gvc_oracle_lambda(int qp) { return k_lambda[qp < 0 ? 0 : qp > 51 ? 51 : qp]; }

/* bits of the signed Exp-Golomb code se(v) (H.264 9.1) */
int gvc_oracle_se_bits(int v)
{
    unsigned code = v > 0 ? 2u * (unsigned)v - 1u : 2u * (unsigned)(-v);
    int n = 0;
    for (code += 1; code > 1; code >>= 1) n++;
    return 2 * n + 1;
}
static int mv_bits(int mx, int my, int px, int py) { return gvc_oracle_se_bits(mx - px) + gvc_oracle_se_bits(my - py); }

static int median3(int a, int b, int c)
{
    const int mn = a < b ? a : b, mx = a < b ? b : a;
    return c < mn ? mn : c > mx ? mx : c;
}

/* ---- K4a: exhaustive lowres search --------------------------------------------------------------
 * cur/ref: half-resolution luma at sample (0,0), padded by >= r+? samples; one 8x8 lowres block per MB.
 * Candidates (dx,dy) in [-r,r]^2, raster order (dy outer); cost = SAD8x8 + lambda*(se(4dx)+se(4dy));
 * lowest cost wins, first candidate wins ties.  out: int16 (dx,dy) per MB. */
void gvc_oracle_me_lowres(const uint8_t *cur, const uint8_t *ref, int stride, int mbw, int mbh, int r, int lambda,
                          int16_t *out_mv, int32_t *out_cost)
{
    for (int my = 0; my < mbh; my++)
        for (int mx = 0; mx < mbw; mx++) {
            const uint8_t *c = cur + (ptrdiff_t)(8 * my) * stride + 8 * mx;
            long best = -1;
            int  bdx = 0, bdy = 0;
            for (int dy = -r; dy <= r; dy++)
                for (int dx = -r; dx <= r; dx++) {
                    const uint8_t *p = ref + (ptrdiff_t)(8 * my + dy) * stride + 8 * mx + dx;
                    const long cost = gvc_oracle_sad(c, stride, p, stride, 8, 8) +
                                      lambda * (gvc_oracle_se_bits(4 * dx) + gvc_oracle_se_bits(4 * dy));
                    if (best < 0 || cost < best) { best = cost; bdx = dx; bdy = dy; }
                }
            out_mv[2 * (my * mbw + mx)]     = (int16_t)bdx;
            out_mv[2 * (my * mbw + mx) + 1] = (int16_t)bdy;
            if (out_cost) out_cost[my * mbw + mx] = (int32_t)best;
        }
}

/* ---- K4b: full-resolution refinement, sub-pel search, partition decision ------------------------- */
typedef struct {
    const uint8_t *cur;           /* source luma (0,0) */
    const uint8_t *planes[4];     /* reference full/H/V/C at (0,0) */
    int stride, w16, h16, pad, lambda;
} me_ctx;

static const int8_t k_sq8[8][2] = {{0, -1}, {0, 1}, {-1, 0}, {1, 0}, {-1, -1}, {-1, 1}, {1, -1}, {1, 1}};

static int subpel_refine(const me_ctx *c, int bx, int by, int size, int *mvx, int *mvy, int px, int py)
{
    uint8_t pred[256];
    const uint8_t *src = c->cur + (ptrdiff_t)by * c->stride + bx;
    int bmx = *mvx, bmy = *mvy;
    gvc_oracle_mc_luma(pred, 16, c->planes, c->stride, bx, by, bmx, bmy, size, size);
    int best = gvc_oracle_satd(src, c->stride, pred, 16, size, size) + c->lambda * mv_bits(bmx, bmy, px, py);
    for (int step = 2; step >= 1; step--) {
        const int cx = bmx, cy = bmy;
        for (int i = 0; i < 8; i++) {
            const int mx = cx + k_sq8[i][0] * step, my = cy + k_sq8[i][1] * step;
            gvc_oracle_mc_luma(pred, 16, c->planes, c->stride, bx, by, mx, my, size, size);
            const int cost = gvc_oracle_satd(src, c->stride, pred, 16, size, size) + c->lambda * mv_bits(mx, my, px, py);
            if (cost < best) { best = cost; bmx = mx; bmy = my; }
        }
    }
    *mvx = bmx; *mvy = bmy;
    return best;
}

/* lowres_mv: int16 (dx,dy) per MB from gvc_oracle_me_lowres.  out: per MB {mv[4][2], cost, mode} as int32[10]:
 * mv0x,mv0y,...,mv3x,mv3y,cost,mode. */
void gvc_oracle_me_refine(const uint8_t *cur, const uint8_t *const planes[4], int stride, int w16, int h16, int pad,
                          const int16_t *lowres_mv, int lambda, int32_t *out)
{
    const int mbw = w16 / 16, mbh = h16 / 16;
    me_ctx c;
    c.cur = cur; c.stride = stride; c.w16 = w16; c.h16 = h16; c.pad = pad; c.lambda = lambda;
    for (int i = 0; i < 4; i++) c.planes[i] = planes[i];
    for (int my = 0; my < mbh; my++)
        for (int mx = 0; mx < mbw; mx++) {
#define LMV(x, y, k) ((x) >= 0 && (x) < mbw && (y) >= 0 && (y) < mbh ? 8 * lowres_mv[2 * ((y)*mbw + (x)) + (k)] : 0)
            int cxn = mx + 1, cyn = my - 1;
            if (!(cxn < mbw && cyn >= 0)) { cxn = mx - 1; cyn = my - 1; }
            const int px = median3(LMV(mx - 1, my, 0), LMV(mx, my - 1, 0), LMV(cxn, cyn, 0));
            const int py = median3(LMV(mx - 1, my, 1), LMV(mx, my - 1, 1), LMV(cxn, cyn, 1));
#undef LMV
            const int x0 = 16 * mx, y0 = 16 * my;
            const int xmin = -x0 - (pad - 8), xmax = w16 - x0 - 16 + (pad - 8);
            const int ymin = -y0 - (pad - 8), ymax = h16 - y0 - 16 + (pad - 8);
            const int basex = 2 * lowres_mv[2 * (my * mbw + mx)], basey = 2 * lowres_mv[2 * (my * mbw + mx) + 1];
            long best16 = -1, best8[4] = {-1, -1, -1, -1}, zero_sad = 0;
            int  b16x = 0, b16y = 0, b8x[4] = {0, 0, 0, 0}, b8y[4] = {0, 0, 0, 0};
            for (int i = 0; i < 27; i++) {
                int cx, cy;
                if (i < 25)       { cx = basex + (i % 5) - 2; cy = basey + (i / 5) - 2; }
                else if (i == 25) { cx = (px + 2) >> 2; cy = (py + 2) >> 2; }
                else              { cx = 0; cy = 0; }
                cx = cx < xmin ? xmin : cx > xmax ? xmax : cx;
                cy = cy < ymin ? ymin : cy > ymax ? ymax : cy;
                const int bits = lambda * mv_bits(4 * cx, 4 * cy, px, py);
                long s16 = 0;
                for (int k = 0; k < 4; k++) {
                    const int ox = 8 * (k & 1), oy = 8 * (k >> 1);
                    const long s = gvc_oracle_sad(cur + (ptrdiff_t)(y0 + oy) * stride + x0 + ox, stride,
                                                  planes[0] + (ptrdiff_t)(y0 + oy + cy) * stride + x0 + ox + cx, stride, 8, 8);
                    s16 += s;
                    if (best8[k] < 0 || s + bits < best8[k]) { best8[k] = s + bits; b8x[k] = cx; b8y[k] = cy; }
                }
                if (best16 < 0 || s16 + bits < best16) { best16 = s16 + bits; b16x = cx; b16y = cy; }
                if (i == 26) zero_sad = s16;
            }
            int32_t *o = out + 10 * (my * mbw + mx);
            /* early termination (the idea of x264's early-skip probe, analyse.c): a macroblock the zero vector already
             * predicts to within GVC_ME_EARLY_MUL * lambda of SAD keeps it as one 16x16 partition, no sub-sample search */
            if (zero_sad <= (long)GVC_ME_EARLY_MUL * lambda) {
                for (int k = 0; k < 8; k++) o[k] = 0;
                o[8] = (int32_t)(zero_sad + lambda * mv_bits(0, 0, px, py)); o[9] = 0;
                continue;
            }
            int m16x = 4 * b16x, m16y = 4 * b16y;
            const int j16 = subpel_refine(&c, x0, y0, 16, &m16x, &m16y, px, py);
            /* one-partition rule: four 8x8 partitions that picked the 16x16 partition's integer vector are not refined */
            if (b8x[0] == b16x && b8x[1] == b16x && b8x[2] == b16x && b8x[3] == b16x &&
                b8y[0] == b16y && b8y[1] == b16y && b8y[2] == b16y && b8y[3] == b16y) {
                for (int k = 0; k < 4; k++) { o[2 * k] = m16x; o[2 * k + 1] = m16y; }
                o[8] = j16; o[9] = 0;
                continue;
            }
            int j8 = 8 * lambda, m8x[4], m8y[4];
            for (int k = 0; k < 4; k++) {
                m8x[k] = 4 * b8x[k]; m8y[k] = 4 * b8y[k];
                j8 += subpel_refine(&c, x0 + 8 * (k & 1), y0 + 8 * (k >> 1), 8, &m8x[k], &m8y[k], px, py);
            }
            if (j8 < j16) {
                for (int k = 0; k < 4; k++) { o[2 * k] = m8x[k]; o[2 * k + 1] = m8y[k]; }
                o[8] = j8; o[9] = 1;
            } else {
                for (int k = 0; k < 4; k++) { o[2 * k] = m16x; o[2 * k + 1] = m16y; }
                o[8] = j16; o[9] = 0;
            }
        }
}

/* ---- K5+K6: inter macroblock prediction, transform, quantisation, reconstruction ------------------ */
static const uint8_t k_zigzag4[16] = {0, 1, 4, 8, 5, 2, 3, 6, 9, 12, 13, 10, 7, 11, 14, 15};

/* coef layout per MB (int16[26][16]): blocks 0..15 luma 4x4 (raster block order, zigzag coefficient order),
 * 16..19 Cb AC, 20..23 Cr AC (zigzag, entry 0 = 0), 24: chroma DC levels Cb[0..3], Cr[4..7], 25: unused for
 * inter MBs (luma DC of Intra16x16).  nzmask bit b set = block b has a nonzero level; bit 24 / 25 = Cb / Cr DC. */
void gvc_oracle_encode_inter_frame(const uint8_t *const cur[3], const uint8_t *const ref_luma[4],
                                   const uint8_t *const ref_chroma[2], uint8_t *const recon[3], int sy, int sc,
                                   int w16, int h16, const int32_t *motion /* 10 int32 per MB */, int qp,
                                   int16_t *coefs, uint32_t *nzmask)
{
    const int mbw = w16 / 16, mbh = h16 / 16, qpc = gvc_oracle_chroma_qp(qp);
    uint16_t mfy[16], by_[16], mfc[16], bc[16];
    gvc_oracle_quant4_tables(qp, 0, mfy, by_);
    gvc_oracle_quant4_tables(qpc, 0, mfc, bc);
    for (int my = 0; my < mbh; my++)
        for (int mx = 0; mx < mbw; mx++) {
            const int32_t *mo = motion + 10 * (my * mbw + mx);
            int16_t       *co = coefs + (size_t)(my * mbw + mx) * 26 * 16;
            uint32_t       nz = 0;
            uint8_t        py[256], pc[2][64];
            memset(co, 0, sizeof(int16_t) * 26 * 16);
            for (int k = 0; k < 4; k++) {
                const int ox = 8 * (k & 1), oy = 8 * (k >> 1);
                gvc_oracle_mc_luma(py + oy * 16 + ox, 16, ref_luma, sy, 16 * mx + ox, 16 * my + oy, mo[2 * k], mo[2 * k + 1], 8, 8);
                for (int p = 0; p < 2; p++)
                    gvc_oracle_mc_chroma(pc[p] + (oy / 2) * 8 + ox / 2, 8, ref_chroma[p], sc, 8 * mx + ox / 2, 8 * my + oy / 2,
                                         mo[2 * k], mo[2 * k + 1], 4, 4);
            }
            for (int b = 0; b < 16; b++) { /* luma */
                const int bx = 4 * (b & 3), by = 4 * (b >> 2);
                int16_t c[16];
                const uint8_t *s = cur[0] + (ptrdiff_t)(16 * my + by) * sy + 16 * mx + bx;
                gvc_oracle_sub4x4_dct(c, s, sy, py + by * 16 + bx, 16);
                if (gvc_oracle_quant(c, 16, mfy, by_)) nz |= 1u << b;
                for (int i = 0; i < 16; i++) co[b * 16 + i] = c[k_zigzag4[i]];
                gvc_oracle_dequant4(c, qp);
                gvc_oracle_add4x4_idct(recon[0] + (ptrdiff_t)(16 * my + by) * sy + 16 * mx + bx, sy, py + by * 16 + bx, 16, c);
            }
            for (int p = 0; p < 2; p++) { /* chroma */
                int16_t c[4][16], dc[4];
                for (int b = 0; b < 4; b++) {
                    const int bx = 4 * (b & 1), by = 4 * (b >> 1);
                    gvc_oracle_sub4x4_dct(c[b], cur[1 + p] + (ptrdiff_t)(8 * my + by) * sc + 8 * mx + bx, sc, pc[p] + by * 8 + bx, 8);
                    dc[b] = c[b][0];
                    c[b][0] = 0;
                }
                gvc_oracle_dct2x2dc(dc);
                if (gvc_oracle_quant_dc(dc, 4, mfc[0] >> 1, bc[0] << 1)) nz |= 1u << (24 + p);
                for (int i = 0; i < 4; i++) co[24 * 16 + 4 * p + i] = dc[i];
                gvc_oracle_dct2x2dc(dc);
                gvc_oracle_dequant2x2_dc(dc, qpc);
                for (int b = 0; b < 4; b++) {
                    const int bx = 4 * (b & 1), by = 4 * (b >> 1);
                    if (gvc_oracle_quant(c[b], 16, mfc, bc)) nz |= 1u << (16 + 4 * p + b);
                    for (int i = 0; i < 16; i++) co[(16 + 4 * p + b) * 16 + i] = c[b][k_zigzag4[i]];
                    gvc_oracle_dequant4(c[b], qpc);
                    c[b][0] = dc[b];
                    gvc_oracle_add4x4_idct(recon[1 + p] + (ptrdiff_t)(8 * my + by) * sc + 8 * mx + bx, sc, pc[p] + by * 8 + bx, 8, c[b]);
                }
            }
            nzmask[my * mbw + mx] = nz;
        }
}

/* replicate the picture edges of a plane stored with padding: fills [x_lo,x_hi) x [y_lo,y_hi) outside [0,w) x [0,h) */
void gvc_oracle_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi)
{
    for (int y = y_lo; y < y_hi; y++)
        for (int x = x_lo; x < x_hi; x++) {
            if (x >= 0 && x < w && y >= 0 && y < h) continue;
            const int cx = x < 0 ? 0 : x >= w ? w - 1 : x, cy = y < 0 ? 0 : y >= h ? h - 1 : y;
            p[(ptrdiff_t)y * stride + x] = p[(ptrdiff_t)cy * stride + cx];
        }
}

/* ---- K7: all-Intra16x16 frame (IDR) ----------------------------------------------------------------
 * Every macroblock is Intra16x16 with the cheapest (SAD against the source) of the available modes
 * vertical (0, needs the MB above), horizontal (1, needs the MB to the left), DC (2); candidates are tried in
 * that order and a later one must be strictly cheaper.  Chroma always uses DC prediction (mode 0).
 * H.264 8.3.3 (Intra_16x16 prediction), 8.3.4 (chroma DC rules per 4x4 block), 8.5.2 / 8.5.10 (DC path);
 * x264 mb_encode_i16x16 for the order of operations.  Neighbours are RECONSTRUCTED samples.
 * coef layout as for inter MBs; luma blocks carry AC only (entry 0 = 0), block 25 holds the 16 luma DC levels
 * (zig-zag), nzmask bit 26 = luma DC nonzero.  modes[mb] = chosen luma mode. */
static void chroma_dc_pred(const uint8_t *rec, int sc, int x0, int y0, int has_top, int has_left, uint8_t pred[64])
{
    int st[2] = {0, 0}, sl[2] = {0, 0};
    for (int i = 0; i < 8; i++) {
        if (has_top) st[i >> 2] += rec[(ptrdiff_t)(y0 - 1) * sc + x0 + i];
        if (has_left) sl[i >> 2] += rec[(ptrdiff_t)(y0 + i) * sc + x0 - 1];
    }
    for (int b = 0; b < 4; b++) {
        const int bx = b & 1, by = b >> 1;
        int v;
        if (bx == by) { /* (0,0) and (4,4): both neighbours when present */
            if (has_top && has_left) v = (st[bx] + sl[by] + 4) >> 3;
            else if (has_top) v = (st[bx] + 2) >> 2;
            else if (has_left) v = (sl[by] + 2) >> 2;
            else v = 128;
        } else if (bx == 1) { /* (4,0): top preferred */
            if (has_top) v = (st[1] + 2) >> 2;
            else if (has_left) v = (sl[0] + 2) >> 2;
            else v = 128;
        } else { /* (0,4): left preferred */
            if (has_left) v = (sl[1] + 2) >> 2;
            else if (has_top) v = (st[0] + 2) >> 2;
            else v = 128;
        }
        for (int j = 0; j < 4; j++)
            for (int i = 0; i < 4; i++) pred[(4 * by + j) * 8 + 4 * bx + i] = (uint8_t)v;
    }
}

void gvc_oracle_encode_intra_frame(const uint8_t *const cur[3], uint8_t *const recon[3], int sy, int sc, int w16, int h16,
                                   int qp, int16_t *coefs, uint32_t *nzmask, uint8_t *modes)
{
    const int mbw = w16 / 16, mbh = h16 / 16, qpc = gvc_oracle_chroma_qp(qp);
    uint16_t mfy[16], by_[16], mfc[16], bc[16];
    gvc_oracle_quant4_tables(qp, 1, mfy, by_);
    gvc_oracle_quant4_tables(qpc, 1, mfc, bc);
    for (int my = 0; my < mbh; my++)
        for (int mx = 0; mx < mbw; mx++) {
            const int has_top = my > 0, has_left = mx > 0, x0 = 16 * mx, y0 = 16 * my;
            int16_t *co = coefs + (size_t)(my * mbw + mx) * 26 * 16;
            uint32_t nz = 0;
            uint8_t  pred[3][256], py[256], pc[64];
            const uint8_t *src = cur[0] + (ptrdiff_t)y0 * sy + x0;
            uint8_t       *rec = recon[0] + (ptrdiff_t)y0 * sy + x0;
            memset(co, 0, sizeof(int16_t) * 26 * 16);
            int sum = 0, n = 0;
            for (int i = 0; i < 16; i++) {
                if (has_top) { sum += rec[-sy + i]; n++; }
                if (has_left) { sum += rec[(ptrdiff_t)i * sy - 1]; n++; }
            }
            const int dcv = n == 32 ? (sum + 16) >> 5 : n == 16 ? (sum + 8) >> 4 : 128;
            for (int y = 0; y < 16; y++)
                for (int x = 0; x < 16; x++) {
                    pred[0][y * 16 + x] = has_top ? rec[-sy + x] : 0;
                    pred[1][y * 16 + x] = has_left ? rec[(ptrdiff_t)y * sy - 1] : 0;
                    pred[2][y * 16 + x] = (uint8_t)dcv;
                }
            int best = -1, mode = 2;
            for (int m = 0; m < 3; m++) {
                if ((m == 0 && !has_top) || (m == 1 && !has_left)) continue;
                const int s = gvc_oracle_sad(src, sy, pred[m], 16, 16, 16);
                if (best < 0 || s < best) { best = s; mode = m; }
            }
            memcpy(py, pred[mode], 256);
            modes[my * mbw + mx] = (uint8_t)mode;
            int16_t c[16][16], dc[16];
            for (int b = 0; b < 16; b++) {
                const int bx = 4 * (b & 3), by = 4 * (b >> 2);
                gvc_oracle_sub4x4_dct(c[b], src + (ptrdiff_t)by * sy + bx, sy, py + by * 16 + bx, 16);
                dc[b] = c[b][0];
                c[b][0] = 0;
                if (gvc_oracle_quant(c[b], 16, mfy, by_)) nz |= 1u << b;
                for (int i = 0; i < 16; i++) co[b * 16 + i] = c[b][k_zigzag4[i]];
                gvc_oracle_dequant4(c[b], qp);
            }
            gvc_oracle_dct4x4dc(dc);
            if (gvc_oracle_quant_dc(dc, 16, mfy[0] >> 1, by_[0] << 1)) nz |= 1u << 26;
            for (int i = 0; i < 16; i++) co[25 * 16 + i] = dc[k_zigzag4[i]];
            gvc_oracle_idct4x4dc(dc);
            gvc_oracle_dequant4_dc(dc, qp);
            for (int b = 0; b < 16; b++) {
                const int bx = 4 * (b & 3), by = 4 * (b >> 2);
                c[b][0] = dc[b];
                gvc_oracle_add4x4_idct(rec + (ptrdiff_t)by * sy + bx, sy, py + by * 16 + bx, 16, c[b]);
            }
            for (int p = 0; p < 2; p++) {
                int16_t cc[4][16], cdc[4];
                uint8_t *crec = recon[1 + p];
                chroma_dc_pred(crec, sc, 8 * mx, 8 * my, has_top, has_left, pc);
                for (int b = 0; b < 4; b++) {
                    const int bx = 4 * (b & 1), by = 4 * (b >> 1);
                    gvc_oracle_sub4x4_dct(cc[b], cur[1 + p] + (ptrdiff_t)(8 * my + by) * sc + 8 * mx + bx, sc, pc + by * 8 + bx, 8);
                    cdc[b] = cc[b][0];
                    cc[b][0] = 0;
                }
                gvc_oracle_dct2x2dc(cdc);
                if (gvc_oracle_quant_dc(cdc, 4, mfc[0] >> 1, bc[0] << 1)) nz |= 1u << (24 + p);
                for (int i = 0; i < 4; i++) co[24 * 16 + 4 * p + i] = cdc[i];
                gvc_oracle_dct2x2dc(cdc);
                gvc_oracle_dequant2x2_dc(cdc, qpc);
                for (int b = 0; b < 4; b++) {
                    const int bx = 4 * (b & 1), by = 4 * (b >> 1);
                    if (gvc_oracle_quant(cc[b], 16, mfc, bc)) nz |= 1u << (16 + 4 * p + b);
                    for (int i = 0; i < 16; i++) co[(16 + 4 * p + b) * 16 + i] = cc[b][k_zigzag4[i]];
                    gvc_oracle_dequant4(cc[b], qpc);
                    cc[b][0] = cdc[b];
                    gvc_oracle_add4x4_idct(crec + (ptrdiff_t)(8 * my + by) * sc + 8 * mx + bx, sc, pc + by * 8 + bx, 8, cc[b]);
                }
            }
            nzmask[my * mbw + mx] = nz;
        }
}
