// gvc_segment.cu -- the segment encoder behind the C ABI: gvc_encoder_* and gvc_transcode_file.
//
// gvc_transcode_file is the drop-in for the reference's `exec.Command("ffmpeg", "-i", path, args..., "c"+path)`
// (client.go:103-118): it reads the segment, runs resize (K1) and the encoder kernels (K2..K7) on the GPU, writes the
// H.264 stream and the MP4 container on the host, and returns the text the reference stores as c<id>.mp4.log
// (client.go:120-123).  On any failure it leaves NO output file, because the reference detects failure only by the
// missing file (client.go:163-167).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "enc_launch.h"
#include "ff_args.h"
#include "gvc_common.h"
#include "media_input.h"
#include "mp4_mux.h"
#include "segment_encoder.h"

using namespace gvc;

namespace {

// Optional per-stage timing with CUDA events on the launching stream (bench.py's roofline figures).
static_assert(GVC_STAGES == 9, "Stage below must list GVC_STAGES entries");
enum Stage { kStResize = 0, kStBorder, kStLowres, kStHpel, kStMeLowres, kStMeRefine, kStEncInter, kStEncIntra, kStEntropy, kStCount };
struct StageTimer {
    bool on = false;
    struct Rec { int stage; cudaEvent_t a, b; };
    std::vector<cudaEvent_t> pool;
    size_t                   used = 0;
    std::vector<Rec>         recs;
    double ms[kStCount] = {};
    int    launches[kStCount] = {};
    cudaEvent_t get() {
        if (used == pool.size()) { cudaEvent_t e; cudaEventCreate(&e); pool.push_back(e); }
        return pool[used++];
    }
    void begin(int stage, cudaStream_t st) {
        if (!on) return;
        Rec r{stage, get(), get()};
        cudaEventRecord(r.a, st);
        recs.push_back(r);
    }
    void end(cudaStream_t st) { if (on) cudaEventRecord(recs.back().b, st); }
    void reset() { recs.clear(); used = 0; for (int i = 0; i < kStCount; ++i) { ms[i] = 0; launches[i] = 0; } }
    void collect() {      // the stream must have been synchronised
        for (const Rec &r : recs) {
            float t = 0;
            if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ms[r.stage] += t; ++launches[r.stage]; }
        }
        recs.clear(); used = 0;
    }
    ~StageTimer() { for (cudaEvent_t e : pool) cudaEventDestroy(e); }
};

struct CudaOps {
    cudaStream_t st = nullptr;       // the pictures' dependent kernel chain
    cudaStream_t side = nullptr;     // per-picture work nothing on the chain waits for (entropy pass, arrival flag)
    cudaStream_t aux[2] = {nullptr, nullptr};   // branches of the reference plane preparation that run beside the chain
    cudaEvent_t  ev_fork = nullptr, ev_side[2] = {nullptr, nullptr}, ev_prep = nullptr, ev_aux[2] = {nullptr, nullptr};
    bool         on_side = false, prep_open = false;
    int          branch = 0;
    StageTimer   tm;
    cudaStream_t cur() const { return on_side ? side : branch ? aux[branch - 1] : st; }
    // prep_fork: the two auxiliary branches start after everything issued on the chain so far; prep_on(b) selects where the
    // next launches go (0 = the chain); prep_join makes the chain wait for both branches.
    int prep_fork() {
        if (tm.on) return 0;
        GVC_CUDA(cudaEventRecord(ev_prep, st));
        GVC_CUDA(cudaStreamWaitEvent(aux[0], ev_prep, 0));
        GVC_CUDA(cudaStreamWaitEvent(aux[1], ev_prep, 0));
        prep_open = true;
        return 0;
    }
    int prep_on(int b) { if (!tm.on) branch = b; return 0; }
    int prep_join() {
        branch = 0;
        if (!prep_open) return 0;
        for (int b = 0; b < 2; ++b) {
            GVC_CUDA(cudaEventRecord(ev_aux[b], aux[b]));
            GVC_CUDA(cudaStreamWaitEvent(st, ev_aux[b], 0));
        }
        prep_open = false;
        return 0;
    }
    // side_begin: what follows runs on the side stream, after everything issued on the chain so far.  side_end(slot):
    // back to the chain; wait_side(slot) makes the chain wait for the side work closed with that slot.  While per-stage
    // timing is on everything stays on the chain (its events bracket single launches).
    int side_begin() {
        if (tm.on) return 0;
        GVC_CUDA(cudaEventRecord(ev_fork, st));
        GVC_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        on_side = true;
        return 0;
    }
    int side_end(int slot) {
        if (!on_side) return 0;
        GVC_CUDA(cudaEventRecord(ev_side[slot], side));
        on_side = false;
        return 0;
    }
    int wait_side(int slot) {
        if (tm.on) return 0;
        GVC_CUDA(cudaStreamWaitEvent(st, ev_side[slot], 0));
        return 0;
    }
#if defined(GVC_MEASURE)
    // GVC_DUP_STAGE=<stage index>: measurement aid, issues that stage's launches twice (what does the stage cost in the mix?)
    static int dup_stage() { static const int s = getenv("GVC_DUP_STAGE") ? atoi(getenv("GVC_DUP_STAGE")) : -1; return s; }
#else
    static int dup_stage() { return -1; }
#endif
    template <class F> int timed(int stage, F f) {
        tm.begin(stage, cur());
        int rc = f();
        if (!rc && stage == dup_stage()) rc = f();
        tm.end(cur());
        return rc;
    }
    void *alloc(size_t n) { void *p = nullptr; return cudaMalloc(&p, n + 64) == cudaSuccess ? p : nullptr; }
    void  release(void *p) { cudaFree(p); }
    void *alloc_host(size_t n) { void *p = nullptr; return cudaHostAlloc(&p, n + 64, cudaHostAllocMapped) == cudaSuccess ? p : nullptr; }
    void *device_alias(void *h) { void *d = nullptr; return cudaHostGetDevicePointer(&d, h, 0) == cudaSuccess ? d : nullptr; }
    int clear(void *dev, size_t n) { GVC_CUDA(cudaMemsetAsync(dev, 0, n, st)); return 0; }
    void  release_host(void *p) { cudaFreeHost(p); }
    int extend_border3(const BorderBatchArgs *a, int n) { return timed(kStBorder, [&] { return launch_extend_border_batch(a, n, cur()); }); }
    int lowres_batch(const LowresBatchArgs *a, int n) { return timed(kStLowres, [&] { return launch_lowres_batch(a, n, cur()); }); }
    int hpel(const HpelArgs *a, int n) { return timed(kStHpel, [&] { return launch_hpel(a, n, cur()); }); }
    int me_lowres(const MeLowArgs *a, int n) { return timed(kStMeLowres, [&] { return launch_me_lowres(a, n, cur()); }); }
    int me_refine(const MeRefArgs *a, int n) { return timed(kStMeRefine, [&] { return launch_me_refine(a, n, st); }); }
    int encode_inter(const EncMbArgs *a, int n) { return timed(kStEncInter, [&] { return launch_encode_inter(a, n, st); }); }
    int encode_intra(const IntraArgs *a, int n) { return timed(kStEncIntra, [&] { return launch_encode_intra(a, n, st); }); }
    int signal_picture(const SignalArgs *a, int n) { return launch_signal_picture(a, n, cur()); }
    int entropy_picture(const EntropyPicture *p, int n) { return timed(kStEntropy, [&] { return launch_entropy_picture(p, n, cur()); }); }
    int copy_plane(uint8_t *dst, int dpitch, const uint8_t *src, int spitch, int w, int h) {
        GVC_CUDA(cudaMemcpy2DAsync(dst, dpitch, src, spitch, w, h, cudaMemcpyDeviceToHost, st));
        return 0;
    }
    int sync() { GVC_CUDA(cudaStreamSynchronize(st)); return 0; }
};

}  // namespace

struct gvc_encoder {
    gvc_encoder_config cfg;
    CudaOps            ops;
    SegmentEncoder<CudaOps> *seg = nullptr;
    gvc_resize_plan   *plan = nullptr;
    int                max_batch = 1;
    // staging for host-supplied source pictures: two buffers (each holds a whole batch) so the next batch can upload (on its
    // own copy stream) while the current one is being encoded
    uint8_t           *d_src[2] = {nullptr, nullptr};
    size_t             d_src_cap[2] = {0, 0};
    std::vector<size_t> d_src_off[2];          // per segment of the batch: where its pictures start in the buffer
    std::vector<int>   d_src_frames[2];
    cudaStream_t       copy_stream = nullptr;
    cudaEvent_t        uploaded[2] = {nullptr, nullptr};
    cudaEvent_t        consumed[2] = {nullptr, nullptr};     // the resize that read staging buffer b has run: b may be overwritten
    bool               consumed_valid[2] = {false, false};
    int                up_next = 0, up_pending = 0;     // next buffer to fill, number of uploaded-but-not-encoded batches
    std::vector<SegmentOutput> outs;           // one per segment of the last finished batch
    int                lanes_last = 0, selected = 0;    // segments in the last batch; the one stats / recon / write_mp4 refer to
    bool               pending = false;
    double             ms_kernels = 0, ms_entropy = 0;
    // the per-picture kernel chain of a batch (everything after the resize) captured once per batch shape and replayed
    cudaGraph_t        graph = nullptr;
    cudaGraphExec_t    graph_exec = nullptr;
    std::vector<int>   graph_n;
};

namespace {

double now_ms() {
    static const auto t0 = std::chrono::steady_clock::now();
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
#if defined(GVC_MEASURE)
bool tracing() { static const bool on = getenv("GVC_TRACE") != nullptr; return on; }     // timeline of issue / finish on stderr
#else
constexpr bool tracing() { return false; }
#endif

// QP of the IDR picture for `-crf`: 3 below the P pictures' (x264's default ip_factor 1.4 is about 3 QP), but not below 4.  The
// streams declare Constrained Baseline, whose CAVLC has no level_prefix above 15 (H.264 9.2.2.1): a level must stay below 2064.
// Only the luma DC of an Intra16x16 macroblock can get larger -- up to 16 * 16 * 255 / 2 before the quantiser -- and from QP 4 on
// (multiplier 8192 / 2) its largest level is 2040 (2331 at QP 3); every other coefficient stays below 820 even at QP 0.
int idr_qp(int crf) { return std::max(crf - 3, 4); }

size_t i420_bytes(int w, int h) { return size_t(w) * h + 2 * size_t((w + 1) / 2) * ((h + 1) / 2); }

// issue resize (source I420 on the device -> padded current pictures) and every encoder kernel for a batch of segments;
// staging >= 0: the pictures sit in that staging buffer, which may be refilled once the resizes have read it
int issue_batch(gvc_encoder *e, const uint8_t *const *d_frames, const int *n, int count, int staging) {
    if (count < 1 || count > e->max_batch) { set_error("batch of %d segments outside 1..%d (gvc_encoder_config.max_batch)", count, e->max_batch); return GVC_ERR_INVALID_ARG; }
    int nmax = 0;
    for (int s = 0; s < count; ++s) {
        if (!d_frames[s] || n[s] < 1) { set_error("segment %d of the batch: null pictures or no pictures", s); return GVC_ERR_INVALID_ARG; }
        nmax = std::max(nmax, n[s]);
    }
    int rc = e->seg->reserve(count, nmax);
    if (rc) { set_error("out of device memory reserving %d segments of %d pictures", count, nmax); return rc; }
    const EncGeom   &g  = e->seg->geom();
    const PlaneSizes ps = e->seg->planes();
    const int sw = e->cfg.src_w, sh = e->cfg.src_h, scw = (sw + 1) / 2, sch = (sh + 1) / 2;
    rc = e->ops.timed(kStResize, [&] {
        for (int s = 0; s < count; ++s) {
            auto cur = e->seg->cur_frames(s);
            gvc_frames src, dst;
            src.plane[0] = const_cast<uint8_t *>(d_frames[s]);
            src.plane[1] = src.plane[0] + size_t(sw) * sh;
            src.plane[2] = src.plane[1] + size_t(scw) * sch;
            src.stride[0] = sw; src.stride[1] = src.stride[2] = scw;
            src.frame_stride[0] = src.frame_stride[1] = src.frame_stride[2] = i420_bytes(sw, sh);
            dst.plane[0] = cur.y + ps.oy; dst.plane[1] = cur.u + ps.oc; dst.plane[2] = cur.v + ps.oc;
            dst.stride[0] = g.sy; dst.stride[1] = dst.stride[2] = g.sc;
            dst.frame_stride[0] = ps.y; dst.frame_stride[1] = dst.frame_stride[2] = ps.c;
            const int r = gvc_resize_frames_device(e->plan, &src, &dst, n[s], e->ops.st);
            if (r) return r;
        }
        return 0;
    });
    if (rc) return rc;
    if (staging >= 0) {
        GVC_CUDA(cudaEventRecord(e->consumed[staging], e->ops.st));
        e->consumed_valid[staging] = true;
    }
    e->seg->begin_segment(n, count);
    const bool use_graph = !e->cfg.disable_graph && !e->cfg.keep_recon && !e->ops.tm.on;
    if (!use_graph) {
        if ((rc = e->seg->encode_device(n, count))) return rc;
    } else {
        const std::vector<int> shape(n, n + count);
        if (!e->graph_exec || e->graph_n != shape) {       // ~11 launches per picture: capture once per batch shape
            if (e->graph_exec) { cudaGraphExecDestroy(e->graph_exec); e->graph_exec = nullptr; }
            if (e->graph) { cudaGraphDestroy(e->graph); e->graph = nullptr; }
            GVC_CUDA(cudaStreamBeginCapture(e->ops.st, cudaStreamCaptureModeThreadLocal));
            rc = e->seg->encode_device(n, count);
            cudaError_t ce = cudaStreamEndCapture(e->ops.st, &e->graph);
            if (rc) return rc;
            if (ce != cudaSuccess) { set_error("stream capture failed: %s", cudaGetErrorString(ce)); return GVC_ERR_CUDA; }
            GVC_CUDA(cudaGraphInstantiate(&e->graph_exec, e->graph, 0));
            e->graph_n = shape;
        } else {
            e->seg->set_frame_counts(n, count);
        }
        GVC_CUDA(cudaGraphLaunch(e->graph_exec, e->ops.st));
    }
    e->lanes_last = count;
    e->pending = true;
    return GVC_OK;
}

uint8_t *to_stream(const SegmentOutput &o, int annexb, size_t *len) {
    uint8_t *buf = static_cast<uint8_t *>(malloc(o.payload.size() + 16));
    if (!buf) return nullptr;
    memcpy(buf, o.payload.data(), o.payload.size());
    if (annexb) {   // 4-byte lengths -> 00 00 00 01 start codes, in place
        size_t i = 0;
        while (i + 4 <= o.payload.size()) {
            const size_t n = (size_t(buf[i]) << 24) | (size_t(buf[i + 1]) << 16) | (size_t(buf[i + 2]) << 8) | buf[i + 3];
            buf[i] = 0; buf[i + 1] = 0; buf[i + 2] = 0; buf[i + 3] = 1;
            i += 4 + n;
        }
    }
    *len = o.payload.size();
    return buf;
}

}  // namespace

extern "C" {

int gvc_encoder_create(gvc_encoder **enc, const gvc_encoder_config *cfg) {
    if (!enc || !cfg) { set_error("gvc_encoder_create: NULL argument"); return GVC_ERR_INVALID_ARG; }
    *enc = nullptr;
    int rc = enter();
    if (rc) return rc;
    gvc_encoder_config c = *cfg;
    if (c.dst_w <= 0) c.dst_w = c.src_w;
    if (c.dst_h <= 0) c.dst_h = c.src_h;
    if (c.fps_num <= 0) { c.fps_num = 30; c.fps_den = 1; }
    if (c.fps_den <= 0) c.fps_den = 1;
    if (c.merange <= 0) c.merange = 16;
    if (c.scale_algo == 0) c.scale_algo = GVC_SCALE_BICUBIC;
    if (c.src_w < 16 || c.src_h < 16 || c.dst_w < 16 || c.dst_h < 16 || c.src_w > 8192 || c.src_h > 8192 || c.dst_w > 8192 || c.dst_h > 8192) {
        set_error("picture size %dx%d -> %dx%d outside 16..8192", c.src_w, c.src_h, c.dst_w, c.dst_h);
        return GVC_ERR_INVALID_ARG;
    }
    if ((c.src_w | c.src_h | c.dst_w | c.dst_h) & 1) { set_error("odd picture sizes are not supported by the 4:2:0 encoder"); return GVC_ERR_UNSUPPORTED; }
    if (c.crf < 0 || c.crf > 51) { set_error("crf %d outside 0..51", c.crf); return GVC_ERR_INVALID_ARG; }
    if (c.merange < 4 || c.merange > 44) { set_error("merange %d outside 4..44", c.merange); return GVC_ERR_INVALID_ARG; }
    if (c.max_batch <= 0) c.max_batch = 1;
    if (c.max_batch > kMaxBatch) { set_error("max_batch %d above the limit of %d segments per batch", c.max_batch, kMaxBatch); return GVC_ERR_INVALID_ARG; }
    gvc_encoder *e = new (std::nothrow) gvc_encoder;
    if (!e) return GVC_ERR_NOMEM;
    e->cfg = c;
    e->max_batch = c.max_batch;
    if (cudaStreamCreateWithFlags(&e->ops.st, cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&e->ops.side, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ops.ev_fork, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&e->ops.ev_side[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ops.ev_side[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&e->ops.aux[0], cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&e->ops.aux[1], cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ops.ev_prep, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&e->ops.ev_aux[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ops.ev_aux[1], cudaEventDisableTiming) != cudaSuccess) {
        gvc_encoder_destroy(e);
        set_error("cudaStreamCreate / cudaEventCreate failed");
        return GVC_ERR_CUDA;
    }
    if ((rc = gvc_resize_plan_create(&e->plan, c.src_w, c.src_h, c.dst_w, c.dst_h, GVC_PIX_YUV420P, c.scale_algo))) { gvc_encoder_destroy(e); return rc; }
    SegmentConfig sc;
    sc.w = c.dst_w; sc.h = c.dst_h; sc.fps_num = c.fps_num; sc.fps_den = c.fps_den;
    sc.qp_p = c.crf; sc.qp_i = idr_qp(c.crf);
    sc.merange = c.merange; sc.keep_recon = c.keep_recon != 0;
    sc.device_entropy = c.entropy != GVC_ENTROPY_HOST;
    sc.max_batch = c.max_batch;
    sc.host_threads = c.host_threads > 0 ? c.host_threads : int(std::max(1u, std::thread::hardware_concurrency()));
    e->seg = new (std::nothrow) SegmentEncoder<CudaOps>(e->ops, sc);
    if (!e->seg) { gvc_encoder_destroy(e); return GVC_ERR_NOMEM; }
    *enc = e;
    return GVC_OK;
}

void gvc_encoder_destroy(gvc_encoder *e) {
    if (!e) return;
    if (gvc::initialised()) cudaSetDevice(gvc::current_device());
    if (e->ops.st) cudaStreamSynchronize(e->ops.st);
    if (e->graph_exec) cudaGraphExecDestroy(e->graph_exec);
    if (e->graph) cudaGraphDestroy(e->graph);
    delete e->seg;
    if (e->plan) gvc_resize_plan_destroy(e->plan);
    for (int b = 0; b < 2; ++b) { if (e->d_src[b]) cudaFree(e->d_src[b]); if (e->uploaded[b]) cudaEventDestroy(e->uploaded[b]); if (e->consumed[b]) cudaEventDestroy(e->consumed[b]); }
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->ops.side) cudaStreamDestroy(e->ops.side);
    for (int b = 0; b < 2; ++b) { if (e->ops.aux[b]) cudaStreamDestroy(e->ops.aux[b]); if (e->ops.ev_aux[b]) cudaEventDestroy(e->ops.ev_aux[b]); }
    if (e->ops.ev_prep) cudaEventDestroy(e->ops.ev_prep);
    if (e->ops.st) cudaStreamDestroy(e->ops.st);
    if (e->ops.ev_fork) cudaEventDestroy(e->ops.ev_fork);
    for (int b = 0; b < 2; ++b) if (e->ops.ev_side[b]) cudaEventDestroy(e->ops.ev_side[b]);
    delete e;
}

int gvc_encoder_encode_device_batch(gvc_encoder *e, const uint8_t *const *d_frames_i420, const int *n_frames, int count) {
    int rc = enter();
    if (rc) return rc;
    if (!e || !d_frames_i420 || !n_frames) { set_error("gvc_encoder_encode_device_batch: bad arguments"); return GVC_ERR_INVALID_ARG; }
    return issue_batch(e, d_frames_i420, n_frames, count, -1);
}
int gvc_encoder_encode_device(gvc_encoder *e, const uint8_t *d_frames_i420, int n_frames) {
    return gvc_encoder_encode_device_batch(e, &d_frames_i420, &n_frames, 1);
}

int gvc_encoder_sync(gvc_encoder *e) {
    int rc = enter();
    if (rc) return rc;
    if (!e) { set_error("encoder is NULL"); return GVC_ERR_INVALID_ARG; }
    GVC_CUDA(cudaStreamSynchronize(e->ops.st));
    return GVC_OK;
}

int gvc_encoder_set_timing(gvc_encoder *e, int on) {
    if (!e) { set_error("encoder is NULL"); return GVC_ERR_INVALID_ARG; }
    e->ops.tm.on = on != 0;
    e->ops.tm.reset();
    return GVC_OK;
}

int gvc_encoder_stage_times(gvc_encoder *e, double *ms8, int *launches8) {   // GVC_STAGES entries each
    int rc = enter();
    if (rc) return rc;
    if (!e || !ms8 || !launches8) { set_error("gvc_encoder_stage_times: bad arguments"); return GVC_ERR_INVALID_ARG; }
    GVC_CUDA(cudaStreamSynchronize(e->ops.st));
    e->ops.tm.collect();
    for (int i = 0; i < kStCount; ++i) { ms8[i] = e->ops.tm.ms[i]; launches8[i] = e->ops.tm.launches[i]; }   // 9 entries
    return GVC_OK;
}

void *gvc_encoder_stream(gvc_encoder *e) { return e ? static_cast<void *>(e->ops.st) : nullptr; }

int gvc_encoder_finish_batch(gvc_encoder *e, int annexb, uint8_t **out, size_t *out_len) {
    int rc = enter();
    if (rc) return rc;
    if (!e || !e->pending) { set_error("gvc_encoder_finish: no segment in flight"); return GVC_ERR_INVALID_ARG; }
    const auto t0 = std::chrono::steady_clock::now();
    if (tracing()) fprintf(stderr, "[gvc %p] %.2f finish begin\n", (void *)e, now_ms());
    if (e->outs.size() < size_t(e->lanes_last)) e->outs.resize(size_t(e->lanes_last));
    if ((rc = e->seg->finish(e->outs.data()))) { set_error("the coded pictures outgrew the slice buffers (twice the raw pictures)"); return rc; }
    if (tracing()) fprintf(stderr, "[gvc %p] %.2f finish end\n", (void *)e, now_ms());
    e->ms_entropy = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    e->pending = false;
    e->selected = 0;
    if (out && out_len)
        for (int s = 0; s < e->lanes_last; ++s) {
            out[s] = to_stream(e->outs[size_t(s)], annexb, &out_len[s]);
            if (!out[s]) { for (int k = 0; k < s; ++k) { free(out[k]); out[k] = nullptr; } return GVC_ERR_NOMEM; }
        }
    return GVC_OK;
}
int gvc_encoder_finish(gvc_encoder *e, int annexb, uint8_t **out, size_t *out_len) {
    if (e && e->pending && e->lanes_last != 1 && out) { set_error("gvc_encoder_finish: %d segments in flight, use gvc_encoder_finish_batch", e->lanes_last); return GVC_ERR_INVALID_ARG; }
    return gvc_encoder_finish_batch(e, annexb, out, out_len);
}

int gvc_encoder_select(gvc_encoder *e, int segment) {
    if (!e || segment < 0 || segment >= e->lanes_last) { set_error("gvc_encoder_select: segment %d outside the last batch", segment); return GVC_ERR_INVALID_ARG; }
    e->selected = segment;
    return GVC_OK;
}

int gvc_encoder_upload_host_batch(gvc_encoder *e, const uint8_t *const *frames_i420, const int *n_frames, int count) {
    int rc = enter();
    if (rc) return rc;
    if (!e || !frames_i420 || !n_frames || count < 1 || count > e->max_batch) { set_error("gvc_encoder_upload_host: bad arguments"); return GVC_ERR_INVALID_ARG; }
    if (e->up_pending >= 2) { set_error("gvc_encoder_upload_host: two batches are already waiting to be encoded"); return GVC_ERR_INVALID_ARG; }
    if (!e->copy_stream) GVC_CUDA(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
    const int    b = e->up_next;
    const size_t fb = i420_bytes(e->cfg.src_w, e->cfg.src_h);
    size_t bytes = 0;
    e->d_src_off[b].assign(size_t(count), 0);
    e->d_src_frames[b].assign(n_frames, n_frames + count);
    for (int s = 0; s < count; ++s) {
        if (!frames_i420[s] || n_frames[s] < 1) { set_error("gvc_encoder_upload_host: segment %d has no pictures", s); return GVC_ERR_INVALID_ARG; }
        e->d_src_off[b][size_t(s)] = bytes;
        bytes += (fb * size_t(n_frames[s]) + 255) & ~size_t(255);
    }
    if (!e->uploaded[b]) GVC_CUDA(cudaEventCreateWithFlags(&e->uploaded[b], cudaEventDisableTiming));
    if (!e->consumed[b]) GVC_CUDA(cudaEventCreateWithFlags(&e->consumed[b], cudaEventDisableTiming));
    // the resize kernels of the batch that last used this buffer must have read it before it is overwritten (or freed)
    if (e->consumed_valid[b]) {
        if (e->d_src_cap[b] < bytes) GVC_CUDA(cudaEventSynchronize(e->consumed[b]));
        else GVC_CUDA(cudaStreamWaitEvent(e->copy_stream, e->consumed[b], 0));
    }
    if (e->d_src_cap[b] < bytes) {
        if (e->d_src[b]) cudaFree(e->d_src[b]);
        e->d_src[b] = nullptr; e->d_src_cap[b] = 0;
        GVC_CUDA(cudaMalloc((void **)&e->d_src[b], bytes + 64));
        e->d_src_cap[b] = bytes;
    }
    if (tracing()) fprintf(stderr, "[gvc %p] %.2f upload issue buf %d\n", (void *)e, now_ms(), b);
    for (int s = 0; s < count; ++s)
        GVC_CUDA(cudaMemcpyAsync(e->d_src[b] + e->d_src_off[b][size_t(s)], frames_i420[s], fb * size_t(n_frames[s]), cudaMemcpyHostToDevice, e->copy_stream));
    GVC_CUDA(cudaEventRecord(e->uploaded[b], e->copy_stream));
    e->up_next ^= 1;
    ++e->up_pending;
    return GVC_OK;
}
int gvc_encoder_upload_host(gvc_encoder *e, const uint8_t *frames_i420, int n_frames) {
    return gvc_encoder_upload_host_batch(e, &frames_i420, &n_frames, 1);
}

int gvc_encoder_encode_uploaded(gvc_encoder *e) {
    int rc = enter();
    if (rc) return rc;
    if (!e || e->up_pending < 1) { set_error("gvc_encoder_encode_uploaded: nothing uploaded"); return GVC_ERR_INVALID_ARG; }
    const int b = e->up_pending == 2 ? e->up_next : e->up_next ^ 1;     // oldest uploaded buffer
    GVC_CUDA(cudaStreamWaitEvent(e->ops.st, e->uploaded[b], 0));
    --e->up_pending;
    const double t0 = now_ms();
    const int count = int(e->d_src_frames[b].size());
    const uint8_t *ptr[kMaxBatch];
    for (int s = 0; s < count; ++s) ptr[s] = e->d_src[b] + e->d_src_off[b][size_t(s)];
    rc = issue_batch(e, ptr, e->d_src_frames[b].data(), count, b);
    if (tracing()) fprintf(stderr, "[gvc %p] %.2f encode issued (%.2f ms on the host)\n", (void *)e, now_ms(), now_ms() - t0);
    return rc;
}

int gvc_encoder_encode_host(gvc_encoder *e, const uint8_t *frames_i420, int n_frames, int annexb, uint8_t **out, size_t *out_len) {
    int rc = gvc_encoder_upload_host(e, frames_i420, n_frames);
    if (rc) return rc;
    if ((rc = gvc_encoder_encode_uploaded(e))) return rc;
    return gvc_encoder_finish(e, annexb, out, out_len);
}

static const SegmentOutput *selected_output(gvc_encoder *e) {
    return e && size_t(e->selected) < e->outs.size() ? &e->outs[size_t(e->selected)] : nullptr;
}

int gvc_encoder_recon(gvc_encoder *e, uint8_t *dst, size_t cap) {
    const SegmentOutput *o = selected_output(e);
    if (!o || !dst) { set_error("gvc_encoder_recon: bad arguments"); return GVC_ERR_INVALID_ARG; }
    if (o->recon.empty()) { set_error("the encoder was created without keep_recon"); return GVC_ERR_INVALID_ARG; }
    if (cap < o->recon.size()) { set_error("recon buffer too small: need %zu bytes", o->recon.size()); return GVC_ERR_TOO_LARGE; }
    memcpy(dst, o->recon.data(), o->recon.size());
    return GVC_OK;
}

int gvc_encoder_stats(gvc_encoder *e, int *stats4_per_picture, int n_pictures) {
    const SegmentOutput *o = selected_output(e);
    if (!o || !stats4_per_picture) { set_error("gvc_encoder_stats: bad arguments"); return GVC_ERR_INVALID_ARG; }
    for (int f = 0; f < n_pictures && f < (int)o->stats.size(); ++f) {
        stats4_per_picture[4 * f] = o->stats[f].skipped; stats4_per_picture[4 * f + 1] = o->stats[f].p16;
        stats4_per_picture[4 * f + 2] = o->stats[f].p8; stats4_per_picture[4 * f + 3] = int(o->au_size[f]);
    }
    return GVC_OK;
}

long long gvc_encoder_record_bytes(gvc_encoder *e) { return e && e->seg && e->selected < e->lanes_last ? (long long)e->seg->record_bytes(e->selected) : -1; }

int gvc_encoder_write_mp4(gvc_encoder *e, const char *path) {
    const SegmentOutput *o = selected_output(e);
    if (!o || !path || o->au_size.empty()) { set_error("gvc_encoder_write_mp4: nothing encoded"); return GVC_ERR_INVALID_ARG; }
    std::vector<Mp4Sample> samples;
    for (size_t f = 0; f < o->au_size.size(); ++f) samples.push_back(Mp4Sample{o->au_offset[f], o->au_size[f], f == 0});
    std::string err, tmp = std::string(path) + ".gvc.tmp";
    if (!write_mp4(tmp, e->cfg.dst_w, e->cfg.dst_h, e->cfg.fps_num, e->cfg.fps_den, o->sps_nal, o->pps_nal, o->payload, samples, err)) {
        set_error("%s", err.c_str());
        return GVC_ERR_IO;
    }
    if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); set_error("cannot rename %s to %s", tmp.c_str(), path); return GVC_ERR_IO; }
    return GVC_OK;
}

int gvc_encoder_set_crf(gvc_encoder *e, int crf) {
    if (!e || crf < 0 || crf > 51) { set_error("gvc_encoder_set_crf: crf %d outside 0..51", crf); return GVC_ERR_INVALID_ARG; }
    if (e->pending) { set_error("gvc_encoder_set_crf: a segment is in flight"); return GVC_ERR_INVALID_ARG; }
    if (crf == e->cfg.crf) return GVC_OK;
    e->cfg.crf = crf;
    e->seg->set_qp(idr_qp(crf), crf);
    e->graph_n.clear();          // quantiser constants are kernel arguments: the captured graph is for the old ones
    return GVC_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------- transcode
namespace {

// `-r`: ffmpeg's output frame rate drops or repeats pictures so that the duration stays; output picture j shows the input
// picture that is on screen at time j / out_fps.
class RateAdapter : public FrameSource {
public:
    RateAdapter(std::unique_ptr<FrameSource> src, int out_num, int out_den) : src_(std::move(src)) {
        w = src_->w; h = src_->h; format_name = src_->format_name;
        in_num_ = src_->fps_num; in_den_ = src_->fps_den;
        fps_num = out_num; fps_den = out_den;
        has_audio = src_->has_audio; audio_copyable = src_->audio_copyable; audio_codec_id = src_->audio_codec_id;
        last_.resize(size_t(w) * h + 2 * size_t((w + 1) / 2) * ((h + 1) / 2));
    }
    int read(uint8_t *dst, int max_frames, std::string &err) override {
        int n = 0;
        while (n < max_frames && !eof_) {
            // input index shown at output picture out_idx_: floor(out_idx * in_fps / out_fps)
            const long long want = (long long)out_idx_ * in_num_ * fps_den / ((long long)in_den_ * fps_num);
            while (in_idx_ < want && !eof_) {
                const int r = src_->read(last_.data(), 1, err);
                if (r < 0) return -1;
                if (r == 0) eof_ = true; else ++in_idx_;
            }
            if (eof_ && in_idx_ < want) break;
            if (in_idx_ < 0) break;
            memcpy(dst + last_.size() * size_t(n), last_.data(), last_.size());
            ++n; ++out_idx_;
        }
        if (eof_) audio = src_->audio;
        return n;
    }
private:
    std::unique_ptr<FrameSource> src_;
    std::vector<uint8_t> last_;
    long long in_idx_ = -1;
    long      out_idx_ = 0;
    int       in_num_ = 30, in_den_ = 1;
    bool      eof_ = false;
};

// The worker process converts piece after piece (client.go:37 loops forever): the encoder, its device buffers and the pinned
// staging memory of the previous piece are kept and reused when the next piece has the same shape.
struct TranscodeCache {
    std::mutex   mu;
    gvc_encoder *enc = nullptr;
    int          key[8] = {0, 0, 0, 0, 0, 0, 0, 0};      // src w, h, dst w, h, fps num, den, merange, scaler
    uint8_t     *pinned[2] = {nullptr, nullptr};
    size_t       pinned_bytes = 0;
    int          device = -1;
    long         allocations = 0;                        // encoder creations + pinned allocations so far (gvc_transcode_allocations)
    void drop() {
        if (enc) { gvc_encoder_destroy(enc); enc = nullptr; }
        for (int i = 0; i < 2; ++i) { if (pinned[i]) cudaFreeHost(pinned[i]); pinned[i] = nullptr; }
        pinned_bytes = 0;
    }
};
TranscodeCache &transcode_cache() { static TranscodeCache c; return c; }

constexpr size_t kChunkBytes = size_t(512) << 20;        // raw pictures of one GOP held in pinned memory (two such buffers)

}  // namespace

extern "C" long gvc_transcode_allocations(void) { return transcode_cache().allocations; }

extern "C" {

int gvc_transcode_file(const char *in_path, const char *ffmpeg_args, const char *out_path, char **log_out, size_t *log_len) {
    std::string log;
    auto finish = [&](int rc) {
        if (rc) log += std::string("gvc: error: ") + gvc_last_error() + "\n";
        if (log_out && log_len) {
            *log_out = static_cast<char *>(malloc(log.size() + 1));
            if (*log_out) { memcpy(*log_out, log.c_str(), log.size() + 1); *log_len = log.size(); }
        }
        return rc;
    };
    if (log_out) *log_out = nullptr;
    if (log_len) *log_len = 0;
    if (!in_path || !ffmpeg_args || !out_path) { set_error("gvc_transcode_file: NULL argument"); return finish(GVC_ERR_INVALID_ARG); }
    int rc = enter();
    if (rc) return finish(rc);
    log += std::string(gvc_version()) + "\n";
    TranscodeOptions opt;
    std::string      err;
    rc = parse_ffmpeg_args(ffmpeg_args, opt, err);
    if (rc != 0) { set_error("%s", err.c_str()); return finish(rc); }
    std::unique_ptr<FrameSource> srcf = open_frame_source(in_path, err);
    if (!srcf) { set_error("%s: %s", in_path, err.c_str()); return finish(err == "No such file or directory" ? GVC_ERR_IO : GVC_ERR_UNSUPPORTED); }
    const int w = srcf->w, h = srcf->h;
    if ((w | h) & 1) { set_error("%s: odd picture size %dx%d is not supported", in_path, w, h); return finish(GVC_ERR_UNSUPPORTED); }
    // audio: an AAC stream is carried over unchanged; anything else cannot be, and silently dropping it would hand the operator's
    // concat step (server.go:357) a file without sound -- so it is an error unless the job says -an
    const bool want_audio = srcf->has_audio && !opt.drop_audio;
    if (want_audio && !srcf->audio_copyable) {
        set_error("%s: the audio stream (codec id 0x%x) cannot be carried over (only AAC is stream-copied); add -an to drop it", in_path, srcf->audio_codec_id);
        return finish(GVC_ERR_UNSUPPORTED);
    }
    const int in_fn = srcf->fps_num, in_fd = srcf->fps_den;
    if (opt.fps_num > 0 && (long long)opt.fps_num * in_fd != (long long)in_fn * opt.fps_den)
        srcf.reset(new RateAdapter(std::move(srcf), opt.fps_num, opt.fps_den));
    const int fn = srcf->fps_num, fd = srcf->fps_den;
    const int dw = opt.dst_w ? opt.dst_w : w, dh = opt.dst_h ? opt.dst_h : h;

    TranscodeCache &C = transcode_cache();
    std::lock_guard<std::mutex> lock(C.mu);
    const size_t fb = i420_bytes(w, h);
    const int    gop = int(std::max<size_t>(1, std::min<size_t>(size_t(std::min(opt.gop, 250)), std::max<size_t>(16, kChunkBytes / fb))));
    const int    key[8] = {w, h, dw, dh, fn, fd, opt.merange, opt.scale_algo};
    if (C.enc && (memcmp(key, C.key, sizeof(key)) != 0 || C.device != gvc::current_device())) C.drop();
    const int    start_crf = opt.bitrate > 0 ? (C.enc ? C.enc->cfg.crf : 26) : opt.crf;     // bitrate mode starts from the last piece's QP
    if (!C.enc) {
        gvc_encoder_config c;
        memset(&c, 0, sizeof(c));
        c.src_w = w; c.src_h = h; c.dst_w = dw; c.dst_h = dh; c.fps_num = fn; c.fps_den = fd; c.crf = start_crf; c.merange = opt.merange; c.scale_algo = opt.scale_algo;
        if ((rc = gvc_encoder_create(&C.enc, &c))) return finish(rc);
        memcpy(C.key, key, sizeof(key));
        C.device = gvc::current_device();
        ++C.allocations;
    }
    if (C.pinned_bytes < fb * size_t(gop)) {
        for (int i = 0; i < 2; ++i) { if (C.pinned[i]) cudaFreeHost(C.pinned[i]); C.pinned[i] = nullptr; }
        C.pinned_bytes = 0;
        for (int i = 0; i < 2; ++i)
            if (cudaHostAlloc((void **)&C.pinned[i], fb * size_t(gop), cudaHostAllocDefault) != cudaSuccess) { C.drop(); set_error("cannot allocate pinned staging"); return finish(GVC_ERR_NOMEM); }
        C.pinned_bytes = fb * size_t(gop);
        ++C.allocations;
    }
    gvc_encoder *e = C.enc;
    if ((rc = gvc_encoder_set_crf(e, start_crf))) return finish(rc);
    char buf[640];
    snprintf(buf, sizeof(buf), "Input #0, %s, from '%s': %dx%d, %d/%d fps%s\nOutput #0, mp4, to '%s': h264 (gvc sm_100a), yuv420p, %dx%d, %d/%d fps, ",
             srcf->format_name.c_str(), in_path, w, h, in_fn, in_fd, srcf->has_audio ? (want_audio ? ", audio: aac (copy)" : ", audio dropped (-an)") : "", out_path, dw, dh, fn, fd);
    log += buf;
    if (opt.bitrate > 0) snprintf(buf, sizeof(buf), "bitrate %ld kb/s (constant QP per GOP, chosen by re-encoding)\n", opt.bitrate / 1000);
    else snprintf(buf, sizeof(buf), "crf %d\n", opt.crf);
    log += buf;

    std::vector<uint8_t>   payload;
    std::vector<Mp4Sample> samples;
    std::vector<uint8_t>   sps, pps;
    int       total = 0, gops = 0, passes = 0;
    long      n_skip = 0, n_p16 = 0, n_p8 = 0, n_intra = 0;
    double    budget_carry = 0;                  // bits the earlier GOPs left over (+) or overspent (-)
    const auto t0 = std::chrono::steady_clock::now();
    auto encode_gop = [&](const uint8_t *host, int n) -> int {      // upload + kernels, asynchronous
        int r = gvc_encoder_upload_host(e, host, n);
        return r ? r : gvc_encoder_encode_uploaded(e);
    };
    auto gop_bits = [&]() { double b = 0; for (size_t sz : e->outs[0].au_size) b += 8.0 * double(sz); return b; };
    int cur = 0;
    int n_cur = srcf->read(C.pinned[0], gop, err);
    if (n_cur < 0) { set_error("%s: %s (picture %d)", in_path, err.c_str(), total); rc = GVC_ERR_IO; }
    while (rc == 0 && n_cur > 0) {
        if ((rc = encode_gop(C.pinned[cur], n_cur))) break;
        // the host decodes the next GOP while the GPU encodes this one
        int n_next = n_cur == gop ? srcf->read(C.pinned[cur ^ 1], gop, err) : 0;
        if (n_next < 0) { set_error("%s: %s (picture %d)", in_path, err.c_str(), total + n_cur); rc = GVC_ERR_IO; }
        int frc = gvc_encoder_finish(e, 0, nullptr, nullptr);
        if (rc == 0) rc = frc;
        if (rc) break;
        ++passes;
        if (opt.bitrate > 0) {
            // one constant QP per GOP, found by re-encoding it (milliseconds each): bits fall by about 2^(-1/6) per QP step
            const double target = std::max(1000.0, double(opt.bitrate) * n_cur * fd / fn + budget_carry);
            int    qp = e->cfg.crf, best_qp = qp;
            double bits = gop_bits(), best_err = fabs(bits - target);
            int    lo = -1, hi = 52;                 // QPs known to give too many / too few bits
            for (int pass = 0; pass < 6 && fabs(bits - target) > 0.05 * target; ++pass) {
                if (bits > target) lo = std::max(lo, qp); else hi = std::min(hi, qp);
                int next = qp + int(lround(6.0 * log2(bits / target)));
                next = std::max(lo + 1, std::min(hi - 1, std::max(0, std::min(51, next))));
                if (next <= lo || next >= hi || next == qp) break;
                qp = next;
                if ((rc = gvc_encoder_set_crf(e, qp)) || (rc = encode_gop(C.pinned[cur], n_cur)) || (rc = gvc_encoder_finish(e, 0, nullptr, nullptr))) break;
                ++passes;
                bits = gop_bits();
                if (fabs(bits - target) < best_err) { best_err = fabs(bits - target); best_qp = qp; }
            }
            if (rc) break;
            if (best_qp != e->cfg.crf) {
                if ((rc = gvc_encoder_set_crf(e, best_qp)) || (rc = encode_gop(C.pinned[cur], n_cur)) || (rc = gvc_encoder_finish(e, 0, nullptr, nullptr))) break;
                ++passes;
                bits = gop_bits();
            }
            budget_carry = std::max(-0.5 * target, std::min(0.5 * target, target - bits));
            snprintf(buf, sizeof(buf), "[gvc] gop %d: %d pictures, qp %d, %.0f bits (target %.0f)\n", gops, n_cur, e->cfg.crf, bits, target);
            log += buf;
        }
        const SegmentOutput &so = e->outs[0];
        if (sps.empty()) { sps = so.sps_nal; pps = so.pps_nal; }
        for (int i = 0; i < n_cur; ++i) {
            samples.push_back(Mp4Sample{payload.size() + so.au_offset[i], so.au_size[i], i == 0});
            n_skip += so.stats[i].skipped; n_p16 += so.stats[i].p16; n_p8 += so.stats[i].p8; n_intra += so.stats[i].intra;
        }
        payload.insert(payload.end(), so.payload.begin(), so.payload.end());
        total += n_cur;
        ++gops;
        cur ^= 1;
        n_cur = n_next;
    }
    if (rc) {                    // leave the cached encoder in a clean state for the next piece
        if (e->pending) gvc_encoder_finish(e, 0, nullptr, nullptr);
        while (e->up_pending > 0) { if (gvc_encoder_encode_uploaded(e) || gvc_encoder_finish(e, 0, nullptr, nullptr)) { C.drop(); break; } }
        return finish(rc);
    }
    if (total == 0) { set_error("%s: no frames", in_path); return finish(GVC_ERR_IO); }
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::string tmp = std::string(out_path) + ".gvc.tmp";
    const Mp4Audio *audio = want_audio && !srcf->audio.packets.empty() ? &srcf->audio : nullptr;
    if (!write_mp4(tmp, dw, dh, fn, fd, sps, pps, payload, samples, err, audio)) { set_error("%s", err.c_str()); return finish(GVC_ERR_IO); }
    if (rename(tmp.c_str(), out_path) != 0) { remove(tmp.c_str()); set_error("cannot rename %s to %s", tmp.c_str(), out_path); return finish(GVC_ERR_IO); }
    const double dur = double(total) * fd / fn, kbps = dur > 0 ? payload.size() * 8.0 / 1000.0 / dur : 0;
    snprintf(buf, sizeof(buf), "frame=%5d fps=%.0f size=%8zukB time=%.2fs bitrate=%.1fkbits/s\n[gvc] mb I:%ld P16:%ld P8:%ld skip:%ld gops:%d passes:%d audio:%zu packets\n",
             total, secs > 0 ? total / secs : 0.0, payload.size() / 1024, dur, kbps, n_intra, n_p16, n_p8, n_skip, gops, passes, audio ? audio->packets.size() : size_t(0));
    log += buf;
    return finish(GVC_OK);
}

}  // extern "C"
