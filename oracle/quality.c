/*
 * quality.c -- CPU oracle (test infrastructure, not product) of the PSNR / SSIM measurement kernel (K9, csrc/gvc_quality.cu).
 *
 * Restates x264's common/pixel.c (third-party, not under /root/reference; PARITY UNPINNED against libx264 itself, the formulas
 * are the published ones): pixel_ssd_wxh -- the sum of squared differences over every sample -- and pixel_ssim_wxh, i.e.
 * ssim_4x4x2_core's integer sums per 4x4 block (s1, s2, ss = sum a^2 + b^2, s12) added over the four blocks of each 8x8
 * window (windows step by four samples) and turned into a value by ssim_end1 in float arithmetic.  One deviation, shared with
 * the kernel: the windows' values are accumulated in double (x264 uses float), so the mean does not depend on summation order.
 */
#include "gvc_oracle.h"

static float ssim_end1(int s1, int s2, int ss, int s12)
{
    const int c1 = 416, c2 = 235963;  /* (int)(.01*.01*255*255*64 + .5), (int)(.03*.03*255*255*64*63 + .5) */
    const int vars = ss * 64 - s1 * s1 - s2 * s2, covar = s12 * 64 - s1 * s2;
    volatile float num = (float)(2 * s1 * s2 + c1) * (float)(2 * covar + c2);    /* volatile: each product rounded to float on its own */
    volatile float den = (float)(s1 * s1 + s2 * s2 + c1) * (float)(vars + c2);
    return num / den;
}

void gvc_oracle_psnr_ssim(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h, uint64_t *ssd, double *ssim)
{
    uint64_t s = 0;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            const int d = (int)a[(size_t)y * sa + x] - (int)b[(size_t)y * sb + x];
            s += (uint64_t)(d * d);
        }
    *ssd = s;
    const int bw = w >> 2, bh = h >> 2;
    double sum = 0.0;
    for (int by = 0; by + 1 < bh; by++)
        for (int bx = 0; bx + 1 < bw; bx++) {
            int s1 = 0, s2 = 0, ss = 0, s12 = 0;
            for (int y = 4 * by; y < 4 * by + 8; y++)
                for (int x = 4 * bx; x < 4 * bx + 8; x++) {
                    const int p = a[(size_t)y * sa + x], q = b[(size_t)y * sb + x];
                    s1 += p; s2 += q; ss += p * p + q * q; s12 += p * q;
                }
            sum += (double)ssim_end1(s1, s2, ss, s12);
        }
    *ssim = (bw > 1 && bh > 1) ? sum / ((double)(bw - 1) * (bh - 1)) : 1.0;
}
