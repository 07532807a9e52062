/*
 * gvc_oracle.h -- CPU oracle for the goVideoCompressor segment compress path.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  libgvc.so never links it.
 *
 * The reference (/root/reference, Go) holds no arithmetic: client.go:115 spawns an `ffmpeg`
 * binary and every pixel operation happens in third-party code that is neither vendored nor
 * pinned (libswscale, libx264).  This oracle is a plain-C restatement of those published
 * algorithms:
 *
 *   sws_*   : FFmpeg libswscale legacy scaler, C path (libswscale/utils.c initFilter,
 *             libswscale/swscale.c hScale8To15_c, libswscale/output.c yuv2planeX_8_c /
 *             yuv2nv12cX_c, libswscale/input.c nv12ToUV_c).
 *             PINNED: bit-exact against FFmpeg 8.0.1 libswscale 9.1.100 (bundled in the
 *             opencv-python-headless wheel) with SWS_BITEXACT|SWS_ACCURATE_RND -- see
 *             tests/golden/make_golden.py and tests/test_oracle_resize.py.
 *   x264_* / h264_* : ITU-T H.264 (transform/quant 8.5, interpolation 8.4.2.2, deblocking 8.7,
 *             intra 8.3) in x264's function decomposition (common/dct.c, quant.c, pixel.c,
 *             mc.c).  PARITY UNPINNED against libx264 itself (no x264 source, library or binary
 *             exists in this environment); pinned instead by spec known-answer vectors
 *             (tests/test_oracle_h264.py) and, for the decoder-visible half (dequant, IDCT,
 *             motion compensation, intra prediction, deblocking), by bit-exact agreement of
 *             reconstructed frames with FFmpeg's h264 decoder (tests/test_tier3_decode.py).
 */
#ifndef GVC_ORACLE_H
#define GVC_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* same numeric values as libswscale's SWS_BILINEAR / SWS_BICUBIC flags */
#define GVC_ORACLE_BILINEAR 2
#define GVC_ORACLE_BICUBIC  4

/* One polyphase filter bank: dst_len rows of `size` int16 taps; row i reads source samples
 * pos[i] .. pos[i]+size-1.  Taps of a row sum to `one` (1<<14 horizontal, 1<<12 vertical). */
typedef struct {
    int      size;
    int      dst_len;
    int16_t *coeff; /* dst_len * size */
    int32_t *pos;   /* dst_len */
} gvc_oracle_filter;

/* libswscale/utils.c initFilter with srcFilter=dstFilter=NULL, filterAlign=1, SWS_BITEXACT.
 * chroma_pos128: phase in 1/128 source-sample units as produced by utils.c get_local_pos
 * (128 for luma and for default-sited 4:2:0 chroma).  Returns 0 or -1. */
int  gvc_oracle_sws_init_filter(gvc_oracle_filter *f, int src_len, int dst_len, int algo, int one,
                                int src_pos128, int dst_pos128);
void gvc_oracle_sws_free_filter(gvc_oracle_filter *f);

/* One 8-bit plane: hScale8To15_c per source row, then yuv2planeX_8_c per output row. */
int gvc_oracle_sws_plane(const uint8_t *src, int sw, int sh, int sstride, int spixstep,
                         uint8_t *dst, int dw, int dh, int dstride, int dpixstep,
                         int algo, int is_chroma);

/* Whole frame, planar 4:2:0 (three planes) */
int gvc_oracle_sws_yuv420p(const uint8_t *const src[3], const int sstride[3], int sw, int sh,
                           uint8_t *const dst[3], const int dstride[3], int dw, int dh, int algo);
/* Whole frame, NV12 (Y plane + interleaved UV plane) */
int gvc_oracle_sws_nv12(const uint8_t *const src[2], const int sstride[2], int sw, int sh,
                        uint8_t *const dst[2], const int dstride[2], int dw, int dh, int algo);

#include <stddef.h>

/* ---------------------------------------------------------------- x264-semantics routines (h264_core.c) --- */
void gvc_oracle_quant4_tables(int qp, int intra, uint16_t mf[16], uint16_t bias[16]);
void gvc_oracle_quant8_tables(int qp, int intra, uint16_t mf[64], uint16_t bias[64]);
int  gvc_oracle_dequant4_scale(int qp, int idx);
int  gvc_oracle_dequant8_scale(int qp, int idx);
int  gvc_oracle_chroma_qp(int qp);
void gvc_oracle_sub4x4_dct(int16_t c[16], const uint8_t *src, int sstride, const uint8_t *pred, int pstride);
void gvc_oracle_sub8x8_dct8(int16_t c[64], const uint8_t *src, int sstride, const uint8_t *pred, int pstride);
void gvc_oracle_dct4x4dc(int16_t d[16]);
void gvc_oracle_idct4x4dc(int16_t d[16]);
void gvc_oracle_dct2x2dc(int16_t d[4]);
int  gvc_oracle_quant(int16_t *c, int n, const uint16_t *mf, const uint16_t *bias);
int  gvc_oracle_quant_dc(int16_t *c, int n, int mf, int bias);
void gvc_oracle_dequant4(int16_t c[16], int qp);
void gvc_oracle_dequant8(int16_t c[64], int qp);
void gvc_oracle_dequant4_dc(int16_t c[16], int qp);
void gvc_oracle_dequant2x2_dc(int16_t c[4], int qp);
void gvc_oracle_add4x4_idct(uint8_t *dst, int dstride, const uint8_t *pred, int pstride, const int16_t c[16]);
void gvc_oracle_add8x8_idct8(uint8_t *dst, int dstride, const uint8_t *pred, int pstride, const int16_t c[64]);
int  gvc_oracle_sad(const uint8_t *a, int as, const uint8_t *b, int bs, int w, int h);
int  gvc_oracle_satd(const uint8_t *a, int as, const uint8_t *b, int bs, int w, int h);
void gvc_oracle_hpel_filter(uint8_t *dh, uint8_t *dv, uint8_t *dc, const uint8_t *src, int stride, int w, int h, int pad);
void gvc_oracle_lowres(const uint8_t *src, int sstride, uint8_t *d0, uint8_t *dh, uint8_t *dv, uint8_t *dc, int dstride,
                       int lw, int lh);
void gvc_oracle_mc_luma(uint8_t *dst, int dstride, const uint8_t *const planes[4], int stride, int bx, int by,
                        int mvx, int mvy, int w, int h);
void gvc_oracle_mc_chroma(uint8_t *dst, int dstride, const uint8_t *src, int stride, int bx, int by, int mvx, int mvy,
                          int w, int h);

/* ---------------------------------------------------------------- frame-level composition (h264_mb.c) ------ */
int  gvc_oracle_lambda(int qp);
/* quality.c: x264 pixel_ssd_wxh / pixel_ssim_wxh (mean SSIM over the 8x8 windows, accumulated in double) of two w x h planes */
void gvc_oracle_psnr_ssim(const uint8_t *a, int sa, const uint8_t *b, int sb, int w, int h, uint64_t *ssd, double *ssim);
int  gvc_oracle_se_bits(int v);
void gvc_oracle_me_lowres(const uint8_t *cur, const uint8_t *ref, int stride, int mbw, int mbh, int r, int lambda,
                          int16_t *out_mv, int32_t *out_cost);
#ifndef GVC_ME_EARLY_MUL
#define GVC_ME_EARLY_MUL 32   /* SAD threshold of the zero-vector early termination, in units of lambda (product: enc_kernels.cuh kMeEarlyMul) */
#endif
void gvc_oracle_me_refine(const uint8_t *cur, const uint8_t *const planes[4], int stride, int w16, int h16, int pad,
                          const int16_t *lowres_mv, int lambda, int32_t *out);
void gvc_oracle_encode_inter_frame(const uint8_t *const cur[3], const uint8_t *const ref_luma[4],
                                   const uint8_t *const ref_chroma[2], uint8_t *const recon[3], int sy, int sc,
                                   int w16, int h16, const int32_t *motion, int qp, int16_t *coefs, uint32_t *nzmask);
void gvc_oracle_encode_intra_frame(const uint8_t *const cur[3], uint8_t *const recon[3], int sy, int sc, int w16, int h16,
                                   int qp, int16_t *coefs, uint32_t *nzmask, uint8_t *modes);
void gvc_oracle_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi);

#ifdef __cplusplus
}
#endif
#endif
