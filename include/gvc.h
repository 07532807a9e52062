/*
 * gvc.h -- C ABI of libgvc.so, the B200 (sm_100a) implementation of goVideoCompressor's
 * per-segment compress path.
 *
 * What it replaces.  The reference worker transcodes one segment by spawning an external
 * process: client.go:101-130 `convert(path, args)` builds `ffmpeg -i <id>.mp4 <args...> c<id>.mp4`
 * (client.go:103-112), runs it (client.go:115-118) and writes its stderr to c<id>.mp4.log
 * (client.go:120-123).  There is no FFI in the reference; this header is the boundary a cgo
 * build of client.go binds instead of exec.Command (see INTEGRATION.md for the stub).
 *
 * Conventions: every function returns GVC_OK (0) or a negative gvc_status; nothing throws or
 * aborts; no torch / C++ types cross the boundary; pointers are plain host or device addresses
 * as each function states; `stream` is a cudaStream_t passed as void* (NULL = default stream).
 * There is NO CPU fallback: without a usable CUDA device every compute entry point fails with
 * GVC_ERR_NO_DEVICE / GVC_ERR_CUDA.
 */
#ifndef GVC_H
#define GVC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum gvc_status {
    GVC_OK              = 0,
    GVC_ERR_INVALID_ARG = -1,
    GVC_ERR_NO_DEVICE   = -2,
    GVC_ERR_CUDA        = -3,
    GVC_ERR_NOMEM       = -4,
    GVC_ERR_UNSUPPORTED = -5, /* e.g. -vcodec other than libx264: no fallback (north_star) */
    GVC_ERR_IO          = -6,
    GVC_ERR_NOT_INIT    = -7,
    GVC_ERR_TOO_LARGE   = -8
} gvc_status;

/* pixel layouts (values follow libavutil's AV_PIX_FMT_* so ffmpeg-side code can pass them through) */
#define GVC_PIX_YUV420P 0  /* three planes Y, U, V           (AV_PIX_FMT_YUV420P) */
#define GVC_PIX_NV12    23 /* Y plane + interleaved UV plane (AV_PIX_FMT_NV12)    */

/* scaler kernels (values follow libswscale's SWS_* flags; ffmpeg's default for -s / -vf scale is bicubic) */
#define GVC_SCALE_BILINEAR 2
#define GVC_SCALE_BICUBIC  4

/* ---------------------------------------------------------------- library life cycle ---------- */

/* Replaces: process start-up of the spawned ffmpeg (client.go:115).  Selects the CUDA device for
 * this worker process (one worker per GPU).  device_ordinal < 0 reads env GVC_DEVICE
 * (default 0) -- env vars are the worker's configuration mechanism (client.go:22-36). */
int gvc_init(int device_ordinal);
void gvc_shutdown(void);

const char *gvc_strerror(int status);
/* Detail text of the most recent failure on the calling thread ("" if none). */
const char *gvc_last_error(void);
/* Frees buffers handed out by this library (the log returned by gvc_transcode_*). */
void gvc_free(void *p);
/* "gvc-b200 <version> sm_100a" */
const char *gvc_version(void);

/* ---------------------------------------------------------------- resize (K1) ----------------- */
/* Replaces: libswscale as reached by `-s WxH` / `-vf scale=W:H` in the args string that
 * server.go:179 sends and client.go:105-115 splices into the ffmpeg command line.  Results are
 * bit-identical to sws_scale() of FFmpeg's C path (flags algo|SWS_BITEXACT|SWS_ACCURATE_RND).   */

typedef struct gvc_resize_plan gvc_resize_plan;

/* Builds the polyphase filter banks (libswscale utils.c initFilter semantics) and uploads the
 * device tables.  pix_fmt: GVC_PIX_*; algo: GVC_SCALE_*.  Needs gvc_init(). */
int  gvc_resize_plan_create(gvc_resize_plan **plan, int src_w, int src_h, int dst_w, int dst_h,
                            int pix_fmt, int algo);
void gvc_resize_plan_destroy(gvc_resize_plan *plan);

/* A batch of n_frames equally laid out frames.  plane[k] points at frame 0's plane k; frame f's
 * plane k starts at plane[k] + f*frame_stride[k]; rows are stride[k] bytes apart.  YUV420P uses
 * k = 0..2, NV12 k = 0..1 (entry 2 ignored). */
typedef struct gvc_frames {
    uint8_t *plane[3];
    int      stride[3];
    size_t   frame_stride[3];
} gvc_frames;

/* src/dst are DEVICE pointers.  Asynchronous on `stream`. */
int gvc_resize_frames_device(const gvc_resize_plan *plan, const gvc_frames *src, const gvc_frames *dst,
                             int n_frames, void *stream);
/* src/dst are HOST pointers (pinned or pageable): copies in, runs the kernels, copies out,
 * returns when dst is complete.  Staging buffers are cached inside the plan. */
int gvc_resize_frames_host(gvc_resize_plan *plan, const gvc_frames *src, const gvc_frames *dst, int n_frames);

/* No counterpart in the reference: an extra entry point for callers that must not allocate.
 * Same result as gvc_resize_frames_device() (src/dst are DEVICE pointers, asynchronous on `stream`), but
 * it always runs the two-pass kernels and keeps their int16 intermediate plane in a fixed 16 MiB array
 * built into the library, so it never calls cudaMalloc/cudaFree or any other allocating routine.  Meant
 * for a worker in its steady state (cudaMalloc would synchronise the device and stall the copy streams of
 * neighbouring segments) and for GPUs whose memory is fully budgeted at start-up (a late allocation could
 * fail after the job was accepted; server.go:276-283 never re-queues a failed piece).
 *   - Limit: one frame needs 2*(src_h*dst_w + 2*ceil(src_h/2)*ceil(dst_w/2)) bytes of the array (4.1 MB for
 *     1080p->720p, 12.4 MB for 2160p->1080p); larger plans return GVC_ERR_TOO_LARGE.
 *   - The 16 MiB are reserved by gvc_init() on the GPU it selects (it looks the array's device address up, which
 *     also makes CUDA load the code that holds it) and stay reserved until the process exits; if they cannot be
 *     had, gvc_init() fails.  The entry point itself does no set-up on its first call.
 *   - All calls share the one array: calls whose kernels can overlap on the GPU (different streams, or
 *     several host threads) race on it.  Serialise them on one stream.  Nothing in the array needs clearing
 *     between calls; every element read has been written earlier in the same call. */
int gvc_resize_frames_device_static(const gvc_resize_plan *plan, const gvc_frames *src, const gvc_frames *dst,
                                    int n_frames, void *stream);

/* Introspection for tests / bench: which kernel path the plan uses (1 = fused smem-tiled kernel,
 * 0 = two-pass kernel with an int16 intermediate plane in device scratch) and tap counts. */
int gvc_resize_plan_info(const gvc_resize_plan *plan, int *fused, int *h_taps_luma, int *v_taps_luma);
/* Forces the two-pass path (tests exercise both). */
int gvc_resize_plan_force_two_pass(gvc_resize_plan *plan, int on);

/* ---------------------------------------------------------------- encoder stages (K2..K6) ----- */
/* Replace: the per-frame work libx264 does inside the spawned ffmpeg (`-vcodec libx264`, server.go:69-70 ->
 * client.go:115).  All pointers are DEVICE pointers addressing sample (0,0) of planes that carry padding
 * (see DESIGN.md "Data layout"); calls are asynchronous on `stream`.  Arithmetic: ITU-T H.264 8.4.2.2 / 8.5
 * in x264's decomposition; results are bit-identical to oracle/h264_core.c + oracle/h264_mb.c. */

typedef struct gvc_ref_planes {     /* one reference frame */
    const uint8_t *y, *h, *v, *c;   /* luma: full sample, and half-sample planes H (x+1/2), V (y+1/2), C (both) */
    const uint8_t *cb, *cr;         /* chroma */
    const uint8_t *low;             /* half-resolution luma */
} gvc_ref_planes;

typedef struct gvc_mb_motion {      /* per macroblock result of the motion search */
    int16_t mv[4][2];               /* quarter-sample vectors of the four 8x8 partitions (equal for 16x16) */
    int32_t cost;
    uint8_t mode;                   /* 0 = P_L0_16x16, 1 = P_8x8 with four 8x8 partitions */
    uint8_t pad[3];
} gvc_mb_motion;

#define GVC_MB_COEFS (26 * 16)      /* int16 per macroblock written by gvc_encode_inter (layout: DESIGN.md) */

/* x264 plane_expand_border: replicate the w x h picture into [x_lo,x_hi) x [y_lo,y_hi) around it. */
int gvc_extend_border(uint8_t *plane, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi, void *stream);
/* x264 frame_init_lowres_core (K3): lowres samples [lx_lo,lx_hi) x [ly_lo,ly_hi); source readable in
 * [x_lo,x_hi) x [y_lo,y_hi) (reads clamp to it).  dh/dv/dc may be NULL. */
int gvc_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh,
               uint8_t *dv, uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi, void *stream);
/* x264 hpel_filter (K2): H, V, C planes over [x_lo,x_hi) x [y_lo,y_hi), which is also the readable region of src. */
int gvc_hpel_filter(const uint8_t *src, uint8_t *dh, uint8_t *dv, uint8_t *dc, int stride, int x_lo, int x_hi, int y_lo,
                    int y_hi, void *stream);
/* K4a: exhaustive lowres search, one 8x8 lowres block per macroblock, window +-range (<= 22); out_mv: (dx,dy) per MB. */
int gvc_me_lowres(const uint8_t *cur_low, const uint8_t *ref_low, int stride, int mbw, int mbh, int range, int lambda,
                  int16_t *out_mv, void *stream);
/* K4b: full-resolution SAD candidates, half/quarter-sample SATD refinement, 16x16 vs 8x8 decision. */
int gvc_me_refine(const uint8_t *cur_y, const gvc_ref_planes *ref, int stride, int w16, int h16, const int16_t *lowres_mv,
                  int lambda, gvc_mb_motion *out, void *stream);
/* K5+K6: motion compensation (x264 mc_luma / mc_chroma), 4x4 transform, quant, dequant, reconstruction of an
 * all-inter frame.  coefs: GVC_MB_COEFS int16 per MB; nzmask: one uint32 per MB. */
int gvc_encode_inter(const uint8_t *cur_y, const uint8_t *cur_u, const uint8_t *cur_v, const gvc_ref_planes *ref,
                     uint8_t *rec_y, uint8_t *rec_u, uint8_t *rec_v, int sy, int sc, int w16, int h16,
                     const gvc_mb_motion *motion, int qp, int16_t *coefs, uint32_t *nzmask, void *stream);
/* K7: an all-Intra16x16 frame (the IDR that opens every segment): prediction from reconstructed neighbours
 * (H.264 8.3.3 vertical / horizontal / DC by SAD, chroma DC), transform with the luma DC Hadamard path, quant,
 * reconstruction.  Macroblock rows run as a wavefront inside one launch; scratch is 8-byte aligned device memory of
 * GVC_INTRA_SCRATCH_BYTES(w16, h16) bytes (cleared by the call: one 64-byte mailbox per macroblock hands the bottom row
 * of samples to the macroblock below).  modes: chosen Intra16x16 prediction mode per MB. */
#define GVC_INTRA_SCRATCH_BYTES(w16, h16) ((size_t)((w16) / 16) * ((h16) / 16) * 64)
int gvc_encode_intra(const uint8_t *cur_y, const uint8_t *cur_u, const uint8_t *cur_v, uint8_t *rec_y, uint8_t *rec_u,
                     uint8_t *rec_v, int sy, int sc, int w16, int h16, int qp, int16_t *coefs, uint32_t *nzmask,
                     uint8_t *modes, void *scratch, void *stream);
/* 8x8 transform path (x264 sub8x8_dct8, quant_8x8, dequant_8x8, add8x8_idct8; H.264 8.5.13): residual of src - pred in 8x8
 * blocks -> levels (64 per block, raster, row = vertical frequency), nz flag per block, reconstruction.  Kernel-level
 * entry point with oracle parity; the segment encoder writes Constrained-Baseline streams and uses the 4x4 path. */
int gvc_transform8x8(const uint8_t *src, int sstride, const uint8_t *pred, int pstride, uint8_t *rec, int rstride, int w,
                     int h, int qp, int intra, int16_t *coefs, uint8_t *nz, void *stream);
/* x264_lambda_tab[qp] */
int gvc_lambda_for_qp(int qp);

/* ---------------------------------------------------------------- quality (K9) --------------- */
/* Measures a coded plane against its source the way north_star's tier 2 states the tolerance (PSNR, SSIM): x264's
 * pixel_ssd_wxh and pixel_ssim_wxh (common/pixel.c; what `ffmpeg -psnr -ssim` with libx264 reports for the spawned process of
 * client.go:115).  a, b: DEVICE pointers to two w x h planes.  *ssd = sum of squared differences (PSNR = 10 log10(255^2 w h /
 * ssd)); *ssim = mean SSIM over the 8x8 windows stepping by 4 samples.  Synchronous on `stream`. */
int gvc_psnr_ssim(const uint8_t *a, int stride_a, const uint8_t *b, int stride_b, int w, int h, unsigned long long *ssd,
                  double *ssim, void *stream);

/* ---------------------------------------------------------------- segment encoder ------------- */
/* Replaces: what one spawned `ffmpeg -i <id>.mp4 <args> c<id>.mp4` does for one piece (client.go:101-130):
 * resize, H.264 encode (IDR + P pictures, closed GOP per call), container.  Kernels run on the library's own
 * streams for the encoder; the CAVLC slice data is written by the entropy kernels by default or by
 * host threads with cfg.entropy = GVC_ENTROPY_HOST (north_star's first wording); NAL wrapping and muxing run on the host. */
typedef struct gvc_encoder gvc_encoder;

typedef struct gvc_encoder_config {
    int src_w, src_h;          /* incoming pictures: 8-bit I420, planes tight */
    int dst_w, dst_h;          /* encoded size; 0 = same as source (`-s WxH` / `-vf scale=`) */
    int fps_num, fps_den;      /* 0 = 30/1 */
    int crf;                   /* `-crf`, 0..51; P pictures use QP = crf, the IDR QP = max(crf - 3, 4) (levels stay codable in Baseline CAVLC) */
    int merange;               /* `-me_range`, 0 = 16 */
    int scale_algo;            /* GVC_SCALE_*, 0 = bicubic */
    int keep_recon;            /* keep the reconstruction of every picture (tests compare it with a decoder) */
    int host_threads;          /* entropy-coding threads, 0 = all host cores */
    int disable_graph;         /* 1 = launch every kernel individually instead of replaying a captured CUDA graph */
    int entropy;               /* GVC_ENTROPY_*: where the CAVLC slice data is written; the streams are bit-identical */
    int max_batch;             /* segments one call may encode together (gvc_encoder_*_batch), 0/1 = one; <= GVC_MAX_BATCH */
} gvc_encoder_config;
#define GVC_MAX_BATCH 16
#define GVC_ENTROPY_DEFAULT 0  /* = GVC_ENTROPY_DEVICE */
#define GVC_ENTROPY_HOST    1  /* host threads code the kernels' per-macroblock records (h264_stream.cpp) */
#define GVC_ENTROPY_DEVICE  2  /* entropy kernels: the finished slices cross PCIe, not the records */

int  gvc_encoder_create(gvc_encoder **enc, const gvc_encoder_config *cfg);
void gvc_encoder_destroy(gvc_encoder *enc);
/* One segment from HOST memory: n_frames I420 pictures.  Copies in, runs every kernel (entropy coding included unless
 * cfg.entropy = GVC_ENTROPY_HOST, in which case host threads code the records the kernels send back).  If out != NULL it receives a malloc'ed stream (gvc_free): Annex B when annexb != 0,
 * else 4-byte length-prefixed NAL units (the MP4 sample format). */
int gvc_encoder_encode_host(gvc_encoder *enc, const uint8_t *frames_i420, int n_frames, int annexb, uint8_t **out, size_t *out_len);
/* The three steps of gvc_encoder_encode_host, separately, so a caller can upload segment n+1 while segment n is being
 * encoded: upload_host starts an asynchronous copy (pinned host memory) into one of two staging buffers on the encoder's
 * copy stream; encode_uploaded issues resize + encode kernels for the oldest uploaded segment (after gvc_encoder_finish
 * of the previous one); gvc_encoder_finish collects the stream. */
int gvc_encoder_upload_host(gvc_encoder *enc, const uint8_t *frames_i420, int n_frames);
int gvc_encoder_encode_uploaded(gvc_encoder *enc);
/* Batches.  Segments share nothing (server.go:176-190 hands whole pieces to workers), so up to cfg.max_batch segments of the
 * encoder's shape are encoded TOGETHER: they advance picture by picture in lockstep and every stage is one kernel launch for
 * all of them (lengths may differ; a shorter segment drops out of the later launches).  The streams are byte-identical to the
 * ones the segments give when encoded one at a time.  frames[s] / n_frames[s] describe segment s; finish_batch fills out[s] /
 * out_len[s] (each gvc_free'd by the caller; both arrays may be NULL).  The single-segment calls above are batches of one.
 * gvc_encoder_select() chooses which segment of the last finished batch gvc_encoder_recon / _stats / _record_bytes /
 * _write_mp4 refer to (0 after every finish). */
int gvc_encoder_upload_host_batch(gvc_encoder *enc, const uint8_t *const *frames_i420, const int *n_frames, int count);
int gvc_encoder_encode_device_batch(gvc_encoder *enc, const uint8_t *const *d_frames_i420, const int *n_frames, int count);
int gvc_encoder_finish_batch(gvc_encoder *enc, int annexb, uint8_t **out, size_t *out_len);
int gvc_encoder_select(gvc_encoder *enc, int segment);
/* Same, pictures already in DEVICE memory; only issues the kernels (asynchronous on gvc_encoder_stream()). */
int gvc_encoder_encode_device(gvc_encoder *enc, const uint8_t *d_frames_i420, int n_frames);
/* Changes `-crf` between segments (quantiser constants are kernel arguments: the next call captures a new graph). */
int gvc_encoder_set_crf(gvc_encoder *enc, int crf);
int gvc_encoder_sync(gvc_encoder *enc);
void *gvc_encoder_stream(gvc_encoder *enc);      /* the cudaStream_t the encoder launches on */
/* Completes a segment started with gvc_encoder_encode_device / _encode_uploaded: waits for the slices (or entropy-codes the
 * records on host threads), wraps them into NAL units.  GVC_ERR_TOO_LARGE if the slices outgrew twice the raw pictures. */
int gvc_encoder_finish(gvc_encoder *enc, int annexb, uint8_t **out, size_t *out_len);
/* keep_recon: the coded-size (multiples of 16) I420 reconstruction of every picture of the last segment. */
int gvc_encoder_recon(gvc_encoder *enc, uint8_t *dst, size_t cap);
/* per picture: skipped MBs, P16x16 MBs, P8x8 MBs, bytes */
int gvc_encoder_stats(gvc_encoder *enc, int *stats4_per_picture, int n_pictures);
/* Per-stage device times (CUDA events on the encoder's stream) for measurement: stages in order resize, border,
 * lowres, hpel, me_lowres, me_refine, encode_inter, encode_intra, entropy (the four entropy kernels of a picture timed as
 * one entry).  set_timing(on) also clears the totals;
 * stage_times synchronises the stream and returns accumulated milliseconds and launch counts (GVC_STAGES entries each). */
#define GVC_STAGES 9
int gvc_encoder_set_timing(gvc_encoder *enc, int on);
int gvc_encoder_stage_times(gvc_encoder *enc, double *ms, int *launches);
/* Bytes the last finished segment sent device -> host (finished slices, or packed nonzero blocks + per-macroblock tables). */
long long gvc_encoder_record_bytes(gvc_encoder *enc);
/* Muxes the last segment into an MP4 file (temp name + rename). */
int gvc_encoder_write_mp4(gvc_encoder *enc, const char *path);

/* THE drop-in for client.go:115 `exec.Command("ffmpeg", "-i", in_path, args..., out_path)`: ffmpeg_args is the
 * third field of the job string (server.go:179) verbatim.  On success out_path exists; on ANY failure it does not
 * (the reference detects failure by the missing file, client.go:163-167).  *log_out (gvc_free) is the text the
 * reference writes to c<id>.mp4.log (client.go:120-123).  Input containers: YUV4MPEG2 natively, anything libavformat reads
 * (MP4 / H.264 pieces as the reference's splitter makes them) when FFmpeg's libraries can be dlopen'ed (GVC_FFMPEG_LIBDIR). */
int gvc_transcode_file(const char *in_path, const char *ffmpeg_args, const char *out_path, char **log_out, size_t *log_len);
/* Diagnostic: how many times gvc_transcode_file had to create its encoder or allocate its pinned staging memory.  The worker
 * converts piece after piece (client.go:37); pieces of one shape reuse both, so the count stays put from the second piece on. */
long gvc_transcode_allocations(void);

#ifdef __cplusplus
}
#endif
#endif /* GVC_H */
