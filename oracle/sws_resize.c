/*
 * sws_resize.c -- CPU oracle (test infrastructure, not product): restatement of the libswscale
 * legacy scaler's C path for 8-bit 4:2:0 content.  See gvc_oracle.h for the pin.
 *
 * Reached in the reference through client.go:115 (`ffmpeg -i in <args> out`, args carrying
 * `-s WxH` / `-vf scale=`), which auto-inserts libswscale.  Algorithm source (FFmpeg, not in
 * /root/reference): libswscale/utils.c initFilter(); swscale.c hScale8To15_c();
 * output.c yuv2planeX_8_c(), yuv2nv12cX_c(); input.c nv12ToUV_c().
 *
 * Fixed-point pipeline per plane:
 *   horizontal  t[y][x] = min( (sum_j src[y][hpos[x]+j] * hcoef[x][j]) >> 7 , 32767 )    int16 (15 bit)
 *   vertical    d[y][x] = clip_u8( ( (64<<12) + sum_j t[vpos[y]+j][x] * vcoef[y][j] ) >> 19 )
 * with sum(hcoef[x][.]) == 1<<14 and sum(vcoef[y][.]) == 1<<12.
 */
#include "gvc_oracle.h"

#include <stdlib.h>
#include <string.h>

#define MAX_REDUCE_CUTOFF 0.002 /* libswscale SWS_MAX_REDUCE_CUTOFF */

static int64_t i64abs(int64_t v) { return v < 0 ? -v : v; }

static int ilog2(unsigned v) /* av_log2: floor(log2(v)), 0 for v==0 */
{
    int n = 0;
    while (v >>= 1) n++;
    return n;
}

static int64_t rounded_div(int64_t a, int64_t b) /* FFmpeg ROUNDED_DIV */
{
    return (a >= 0 ? a + (b >> 1) : a - (b >> 1)) / b;
}

/* Kernel weight at distance d (2.30 fixed point, already divided by the downscale ratio). */
static int64_t tap_weight(int algo, int64_t d, int64_t fone)
{
    if (algo == GVC_ORACLE_BICUBIC) {
        const int64_t B = 0;                                /* param[0] default */
        const int64_t C = (int64_t)(0.6 * (1 << 24));       /* param[1] default */
        int64_t w;
        if (d >= 1LL << 31) return 0;
        {
            int64_t dd  = (d * d) >> 30;
            int64_t ddd = (dd * d) >> 30;
            if (d < 1LL << 30)
                w = (12 * (1 << 24) - 9 * B - 6 * C) * ddd + (-18 * (1 << 24) + 12 * B + 6 * C) * dd +
                    (6 * (1 << 24) - 2 * B) * (1LL << 30);
            else
                w = (-B - 6 * C) * ddd + (6 * B + 30 * C) * dd + (-12 * B - 48 * C) * d +
                    (8 * B + 24 * C) * (1LL << 30);
        }
        return w / ((1LL << 54) / fone);
    } else { /* bilinear */
        int64_t w = (1LL << 30) - d;
        if (w < 0) w = 0;
        return w * (fone >> 30);
    }
}

int gvc_oracle_sws_init_filter(gvc_oracle_filter *f, int src_len, int dst_len, int algo, int one,
                               int src_pos128, int dst_pos128)
{
    const int     inc  = (int)((((int64_t)src_len << 16) + (dst_len >> 1)) / dst_len);
    int           r    = src_len / dst_len;
    int           lg   = ilog2((unsigned)r);
    const int64_t fone = 1LL << (54 - (lg < 8 ? lg : 8));
    int           n, i, j, min_n = 0;
    int64_t      *w   = NULL;
    int32_t      *pos = NULL;
    int16_t      *out = NULL;

    memset(f, 0, sizeof(*f));
    if (src_len < 1 || dst_len < 1 || (algo != GVC_ORACLE_BICUBIC && algo != GVC_ORACLE_BILINEAR)) return -1;
    pos = (int32_t *)malloc(sizeof(int32_t) * (size_t)dst_len);
    if (!pos) return -1;

    if (abs(inc - 0x10000) < 10 && src_pos128 == dst_pos128) { /* unscaled axis */
        n = 1;
        w = (int64_t *)calloc((size_t)dst_len, sizeof(int64_t));
        if (!w) goto fail;
        for (i = 0; i < dst_len; i++) { w[i] = fone; pos[i] = i; }
    } else {
        const int support = algo == GVC_ORACLE_BICUBIC ? 4 : 2; /* scale_algorithms[].size_factor */
        int64_t   center;
        if (inc <= 1 << 16) n = 1 + support;
        else                n = 1 + (support * src_len + dst_len - 1) / dst_len;
        if (n > src_len - 2) n = src_len - 2;
        if (n < 1) n = 1;
        w = (int64_t *)malloc(sizeof(int64_t) * (size_t)dst_len * (size_t)n);
        if (!w) goto fail;
        center = ((dst_pos128 * (int64_t)inc) >> 7) - ((src_pos128 * 0x10000LL) >> 7);
        for (i = 0; i < dst_len; i++) {
            int xx = (int)((center - (n - 2) * (1LL << 16)) / (1 << 17));
            pos[i] = xx;
            for (j = 0; j < n; j++) {
                int64_t d = i64abs(((int64_t)xx * (1 << 17)) - center) << 13;
                if (inc > 1 << 16) d = d * dst_len / src_len;
                w[(size_t)i * n + j] = tap_weight(algo, d, fone);
                xx++;
            }
            center += 2 * inc;
        }
    }

    /* trim: drop near-zero leading taps (shifting the row left) and count trailing ones */
    for (i = dst_len - 1; i >= 0; i--) {
        int64_t *row  = w + (size_t)i * n;
        int      keep = n;
        int64_t  acc  = 0;
        for (j = 0; j < n; j++) {
            int k;
            acc += i64abs(row[0]);
            if (acc > MAX_REDUCE_CUTOFF * fone) break;
            if (i < dst_len - 1 && pos[i] >= pos[i + 1]) break; /* keep positions monotone */
            for (k = 1; k < n; k++) row[k - 1] = row[k];
            row[n - 1] = 0;
            pos[i]++;
        }
        acc = 0;
        for (j = n - 1; j > 0; j--) {
            acc += i64abs(row[j]);
            if (acc > MAX_REDUCE_CUTOFF * fone) break;
            keep--;
        }
        if (keep > min_n) min_n = keep;
    }

    /* compact rows to min_n taps (SWS_BITEXACT zeroes anything past min_n; align 1 drops it) */
    {
        int64_t *c = (int64_t *)malloc(sizeof(int64_t) * (size_t)dst_len * (size_t)min_n);
        if (!c) goto fail;
        for (i = 0; i < dst_len; i++)
            for (j = 0; j < min_n; j++) c[(size_t)i * min_n + j] = w[(size_t)i * n + j];
        free(w);
        w = c;
        n = min_n;
    }

    /* fold taps that fall outside [0, src_len) onto the edge samples */
    for (i = 0; i < dst_len; i++) {
        int64_t *row = w + (size_t)i * n;
        if (pos[i] < 0) {
            for (j = 1; j < n; j++) {
                int left = j + pos[i] > 0 ? j + pos[i] : 0;
                row[left] += row[j];
                row[j] = 0;
            }
            pos[i] = 0;
        }
        if (pos[i] + n > src_len) {
            int     shift = pos[i] + (n - src_len < 0 ? n - src_len : 0);
            int64_t acc   = 0;
            for (j = n - 1; j >= 0; j--)
                if (pos[i] + j >= src_len) { acc += row[j]; row[j] = 0; }
            for (j = n - 1; j >= 0; j--) row[j] = j < shift ? 0 : row[j - shift];
            pos[i] -= shift;
            row[src_len - 1 - pos[i]] += acc;
        }
    }

    /* normalise each row to `one` with error feedback, store int16 */
    out = (int16_t *)malloc(sizeof(int16_t) * (size_t)dst_len * (size_t)n);
    if (!out) goto fail;
    for (i = 0; i < dst_len; i++) {
        int64_t *row = w + (size_t)i * n;
        int64_t  sum = 0, err = 0;
        for (j = 0; j < n; j++) sum += row[j];
        sum = (sum + one / 2) / one;
        if (!sum) sum = 1;
        for (j = 0; j < n; j++) {
            int64_t v  = row[j] + err;
            int     iv = (int)rounded_div(v, sum);
            out[(size_t)i * n + j] = (int16_t)iv;
            err = v - iv * sum;
        }
    }
    free(w);
    f->size = n; f->dst_len = dst_len; f->coeff = out; f->pos = pos;
    return 0;
fail:
    free(w); free(pos); free(out);
    return -1;
}

void gvc_oracle_sws_free_filter(gvc_oracle_filter *f)
{
    free(f->coeff); free(f->pos);
    memset(f, 0, sizeof(*f));
}

/* utils.c get_local_pos() with the context's default chroma positions (-513) */
static int local_pos(int subsample) { return (((128 << subsample) - 128) + 128) >> subsample; }

int gvc_oracle_sws_plane(const uint8_t *src, int sw, int sh, int sstride, int spixstep,
                         uint8_t *dst, int dw, int dh, int dstride, int dpixstep,
                         int algo, int is_chroma)
{
    gvc_oracle_filter hf, vf;
    int16_t          *tmp = NULL;
    int               x, y, j, rc = -1;
    const int         p = local_pos(is_chroma ? 1 : 0);

    if (gvc_oracle_sws_init_filter(&hf, sw, dw, algo, 1 << 14, p, p)) return -1;
    if (gvc_oracle_sws_init_filter(&vf, sh, dh, algo, 1 << 12, p, p)) { gvc_oracle_sws_free_filter(&hf); return -1; }
    tmp = (int16_t *)malloc(sizeof(int16_t) * (size_t)sh * (size_t)dw);
    if (!tmp) goto done;

    for (y = 0; y < sh; y++) { /* hScale8To15_c */
        const uint8_t *s = src + (size_t)y * sstride;
        int16_t       *t = tmp + (size_t)y * dw;
        for (x = 0; x < dw; x++) {
            const int16_t *c = hf.coeff + (size_t)x * hf.size;
            int            v = 0, sp = hf.pos[x];
            for (j = 0; j < hf.size; j++) v += (int)s[(size_t)(sp + j) * spixstep] * c[j];
            v >>= 7;
            t[x] = (int16_t)(v < 32767 ? v : 32767);
        }
    }
    for (y = 0; y < dh; y++) { /* yuv2planeX_8_c, dither row = 64 (8-bit source) */
        const int16_t *c  = vf.coeff + (size_t)y * vf.size;
        const int      sp = vf.pos[y];
        uint8_t       *d  = dst + (size_t)y * dstride;
        for (x = 0; x < dw; x++) {
            int v = 64 << 12;
            for (j = 0; j < vf.size; j++) v += tmp[(size_t)(sp + j) * dw + x] * c[j];
            v >>= 19;
            d[(size_t)x * dpixstep] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
    rc = 0;
done:
    free(tmp);
    gvc_oracle_sws_free_filter(&hf);
    gvc_oracle_sws_free_filter(&vf);
    return rc;
}

int gvc_oracle_sws_yuv420p(const uint8_t *const src[3], const int sstride[3], int sw, int sh,
                           uint8_t *const dst[3], const int dstride[3], int dw, int dh, int algo)
{
    const int csw = (sw + 1) >> 1, csh = (sh + 1) >> 1, cdw = (dw + 1) >> 1, cdh = (dh + 1) >> 1;
    if (gvc_oracle_sws_plane(src[0], sw, sh, sstride[0], 1, dst[0], dw, dh, dstride[0], 1, algo, 0)) return -1;
    if (gvc_oracle_sws_plane(src[1], csw, csh, sstride[1], 1, dst[1], cdw, cdh, dstride[1], 1, algo, 1)) return -1;
    if (gvc_oracle_sws_plane(src[2], csw, csh, sstride[2], 1, dst[2], cdw, cdh, dstride[2], 1, algo, 1)) return -1;
    return 0;
}

int gvc_oracle_sws_nv12(const uint8_t *const src[2], const int sstride[2], int sw, int sh,
                        uint8_t *const dst[2], const int dstride[2], int dw, int dh, int algo)
{
    const int csw = (sw + 1) >> 1, csh = (sh + 1) >> 1, cdw = (dw + 1) >> 1, cdh = (dh + 1) >> 1;
    if (gvc_oracle_sws_plane(src[0], sw, sh, sstride[0], 1, dst[0], dw, dh, dstride[0], 1, algo, 0)) return -1;
    /* nv12ToUV_c de-interleaves; yuv2nv12cX_c re-interleaves: same filter on the even / odd bytes */
    if (gvc_oracle_sws_plane(src[1], csw, csh, sstride[1], 2, dst[1], cdw, cdh, dstride[1], 2, algo, 1)) return -1;
    if (gvc_oracle_sws_plane(src[1] + 1, csw, csh, sstride[1], 2, dst[1] + 1, cdw, cdh, dstride[1], 2, algo, 1)) return -1;
    return 0;
}
