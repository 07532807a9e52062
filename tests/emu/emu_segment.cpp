// emu_segment.cpp -- TEST HARNESS: the product's SegmentEncoder (segment_encoder.h) instantiated over the host
// emulation of the kernel bodies, so the CPU suite can check the orchestration, the bitstream writer and the MP4
// muxer against a real H.264 decoder without a GPU.  Not part of libgvc.so.
#define GVC_EMU 1
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../govideocompressor_b200/csrc/mp4_mux.h"
#include "../../govideocompressor_b200/csrc/segment_encoder.h"

using namespace gvc;

extern "C" {
void emu_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi);
void emu_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh, uint8_t *dv,
                uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi);
void emu_hpel(const uint8_t *src, uint8_t *dh, uint8_t *dv, uint8_t *dc, int stride, int x_lo, int x_hi, int y_lo, int y_hi);
}
void emu_run_me_lowres(const MeLowArgs &a);
void emu_run_hpel(const HpelArgs &a);
void emu_run_me_refine(const MeRefArgs &a);
void emu_run_encode_inter(const EncMbArgs &a);
void emu_run_encode_intra(const IntraArgs &a);

namespace {
struct EmuOps {
    void *alloc(size_t n) { return calloc(1, n + 64); }
    void  release(void *p) { free(p); }
    void *alloc_host(size_t n) { return calloc(1, n + 64); }
    void  release_host(void *p) { free(p); }
    void *device_alias(void *h) { return h; }
    int clear(void *p, size_t n) { memset(p, 0, n); return 0; }
    // the batched stage launchers: one argument record per segment, run one after the other
    int extend_border3(const BorderBatchArgs *b, int count) { for (int i = 0; i < count; ++i) extend_border3_one(b[i]); return 0; }
    int lowres_batch(const LowresBatchArgs *b, int count) { for (int i = 0; i < count; ++i) lowres_batch_one(b[i]); return 0; }
    int extend_border3_one(const BorderBatchArgs &a) {      // same (piece, thread) decompositions as the CUDA kernels
        const bool vec = border16_ok(a);
        for (int f = 0; f < a.n_frames; ++f)
            for (int k = 0; k < 3; ++k) {
                uint8_t *p = a.p[k] + size_t(f) * a.frame_stride[k];
                int x, y;
                if (vec) {
                    for (int id = 0; id < border16_pieces(a, k) + 9; ++id)
                        for (int lane = 0; lane < 32; ++lane)
                            if (border16_piece_xy(a, k, id, lane, x, y)) extend_border_16(p, a.stride[k], a.w[k], a.h[k], x, y);
                } else {
                    for (int id = 0; id < border_pieces(a, k) + 3; ++id)
                        for (int tid = 0; tid < 256; ++tid)
                            if (border_piece_xy(a, k, id, tid, x, y)) extend_border_px(p, a.stride[k], a.w[k], a.h[k], x, y);
                }
            }
        return 0;
    }
    int lowres_batch_one(const LowresBatchArgs &a) {        // the kernel's thread decomposition: four samples per thread
        const bool al = lowres_aligned(a);
        for (int f = 0; f < a.n_frames; ++f)
            for (int ly = a.ly_lo; ly < a.ly_hi; ++ly)
                for (int lx = a.lx_lo; lx < a.lx_hi; lx += 4)
                    lowres_quad(a, al, a.src + size_t(f) * a.sframe, a.dst + size_t(f) * a.dframe, lx, ly);
        return 0;
    }
    int hpel(const HpelArgs *a, int count) { for (int i = 0; i < count; ++i) emu_run_hpel(a[i]); return 0; }
    int me_lowres(const MeLowArgs *a, int count) { for (int i = 0; i < count; ++i) emu_run_me_lowres(a[i]); return 0; }
    int me_refine(const MeRefArgs *a, int count) { for (int i = 0; i < count; ++i) emu_run_me_refine(a[i]); return 0; }
    int encode_inter(const EncMbArgs *a, int count) { for (int i = 0; i < count; ++i) emu_run_encode_inter(a[i]); return 0; }
    int encode_intra(const IntraArgs *a, int count) { for (int i = 0; i < count; ++i) emu_run_encode_intra(a[i]); return 0; }
    int signal_picture(const SignalArgs *a, int count) { for (int i = 0; i < count; ++i) { *a[i].count_out = *a[i].count_in; *a[i].done = 1; } return 0; }
    int entropy_picture(const EntropyPicture *p, int count) { for (int i = 0; i < count; ++i) entropy_picture_one(p[i]); return 0; }
    int entropy_picture_one(const EntropyPicture &p) {      // the entropy kernels' per-thread bodies, serially; the scan as a plain loop
        const int nmb = p.ent.mbw * p.ent.mbh;
        for (int mb = 0; mb < nmb; ++mb) {                  // the warp of ent_mb_kernel, lane by lane
            static EntSmem sm;
            sm.kind = 0; sm.overflow = 0;
            int bits[32] = {0}, off[32] = {0}, total = 0;
            for (int l = 0; l < 32; ++l) bits[l] = ent_lane_encode(l, p.ent, mb, sm);
            for (int pos = 0; pos < kEntLanes; ++pos) { off[ent_lane_at(pos)] = total; total += bits[ent_lane_at(pos)]; }
            const int words = (total + 31) >> 5;
            for (int j = 0; j < words; ++j) sm.mb[j] = 0;
            for (int l = 0; l < kEntLanes; ++l)
                if (bits[l]) ent_or_bits(sm.mb, kMbBufWords, (unsigned long long)off[l], sm.lane_buf[l], bits[l]);
            memcpy(p.ent.mb_buf + size_t(mb) * kMbBufWords, sm.mb, size_t(words) * 4);
            p.ent.mb_bits[mb] = uint32_t(total);
            p.ent.mb_kind[mb] = uint8_t(sm.kind | (sm.overflow ? 128 : 0));
        }
        const SliceArgs &a = p.slice;
        int prev = -1;
        unsigned long long bits = 0;
        unsigned counts[4] = {0, 0, 0, 0};
        bool ovf = false;
        for (int i = 0; i < nmb; ++i) {
            const int k = a.mb_kind[i];
            ++counts[k & 3];
            ovf = ovf || (k & 128);
            if (!(k & 127)) continue;
            const unsigned run = unsigned(i - 1 - prev);
            a.mb_run[i] = run;
            a.mb_off[i] = uint32_t(a.header_bits + bits);
            bits += (a.p_slice ? ue_len(run) : 0) + a.mb_bits[i];
            prev = i;
        }
        const long long need = std::min<long long>(a.slice_words, ((long long)a.header_bits + (long long)bits + 64 + 1 + 31) / 32 + 4);
        memset(a.slice, 0, size_t(need) * 4);
        ent_finish_slice(a, bits, unsigned(nmb - 1 - prev), counts, ovf);
        for (int mb = 0; mb < nmb; ++mb) ent_place(a, mb);
        const unsigned bytes = a.info[1], base = p.used[0], padded = (bytes + 15u) & ~15u;
        const bool fits = a.info[6] == 0 && base + padded <= p.host_cap;
        if (fits)
            for (unsigned j = 0; j < padded / 4; ++j) {
                const uint32_t w = a.slice[j];
                uint8_t *d = p.host_buf + base + 4 * size_t(j);
                d[0] = uint8_t(w >> 24); d[1] = uint8_t(w >> 16); d[2] = uint8_t(w >> 8); d[3] = uint8_t(w);
            }
        p.used[1] = fits ? base + padded : base;
        p.pos[0] = base; p.pos[1] = bytes; p.pos[2] = a.info[0]; p.pos[3] = a.info[2]; p.pos[4] = a.info[3]; p.pos[5] = a.info[4];
        p.pos[6] = a.info[5]; p.pos[7] = fits ? 0u : 1u;
        return 0;
    }
    int copy_plane(uint8_t *dst, int dpitch, const uint8_t *src, int spitch, int w, int h) {
        for (int y = 0; y < h; ++y) memcpy(dst + size_t(y) * dpitch, src + size_t(y) * spitch, w);
        return 0;
    }
    int prep_fork() { return 0; }
    int prep_on(int) { return 0; }
    int prep_join() { return 0; }
    int side_begin() { return 0; }
    int side_end(int) { return 0; }
    int wait_side(int) { return 0; }
    int sync() { return 0; }
};
}  // namespace

// Test input with sound: emu_set_audio(n, seed) makes the next MP4 files written by emu_encode_segment carry an AAC track of n
// pseudo-random packets (AudioSpecificConfig 0x12 0x10: AAC-LC, 44.1 kHz, stereo).  The payload is never decoded by anything here.
static Mp4Audio g_emu_audio;
extern "C" void emu_set_audio(int n_packets, unsigned seed) {
    g_emu_audio = Mp4Audio();
    if (n_packets <= 0) return;
    g_emu_audio.asc = {0x12, 0x10};
    parse_aac_config(g_emu_audio);
    unsigned x = seed * 2654435761u + 12345u;
    for (int i = 0; i < n_packets; ++i) {
        x = x * 1664525u + 1013904223u;
        std::vector<uint8_t> p(size_t(40 + (x >> 24) % 200));
        for (uint8_t &b : p) { x = x * 1664525u + 1013904223u; b = uint8_t(x >> 24); }
        g_emu_audio.packets.push_back(p);
    }
}
static unsigned long long fnv(unsigned long long h, const uint8_t *p, size_t n) { for (size_t i = 0; i < n; ++i) { h ^= p[i]; h *= 1099511628211ull; } return h; }
static unsigned long long audio_hash(const Mp4Audio &a) {
    unsigned long long h = 1469598103934665603ull;
    h = fnv(h, a.asc.data(), a.asc.size());
    for (const auto &p : a.packets) { const unsigned n = unsigned(p.size()); h = fnv(h, reinterpret_cast<const uint8_t *>(&n), 4); h = fnv(h, p.data(), p.size()); }
    return h;
}
extern "C" unsigned long long emu_audio_hash(void) { return audio_hash(g_emu_audio); }

extern "C" {

// frames: n pictures, I420 tight (w*h*3/2 each).  Writes an MP4 to mp4_path (if not NULL), the Annex B stream to
// annexb (caller buffer of annexb_cap bytes; *annexb_len receives the size), the encoder's reconstruction to recon
// (n * w16*h16*3/2 bytes, or NULL) and per-picture stats (4 ints each: skipped, p16, p8, bytes).  Returns 0 or <0.
int emu_encode_segment(const uint8_t *frames, int n, int w, int h, int fps, int qp_i, int qp_p, int merange, const char *mp4_path,
                       uint8_t *annexb, size_t annexb_cap, size_t *annexb_len, uint8_t *recon, int *stats, int device_entropy) {
    EmuOps        ops;
    SegmentConfig cfg;
    cfg.w = w; cfg.h = h; cfg.fps_num = fps; cfg.fps_den = 1; cfg.qp_i = qp_i; cfg.qp_p = qp_p; cfg.merange = merange;
    cfg.keep_recon = recon != nullptr;
    cfg.device_entropy = device_entropy != 0;
    SegmentEncoder<EmuOps> enc(ops, cfg);
    if (enc.reserve(n)) return -4;
    const EncGeom   &g  = enc.geom();
    const PlaneSizes ps = enc.planes();
    auto cur = enc.cur_frames();
    const int cw = (w + 1) / 2, ch = (h + 1) / 2;
    for (int f = 0; f < n; ++f) {
        const uint8_t *src = frames + size_t(w) * h * 3 / 2 * f;   // tests use even sizes
        for (int y = 0; y < h; ++y) memcpy(cur.y + ps.y * f + ps.oy + size_t(y) * g.sy, src + size_t(y) * w, w);
        for (int y = 0; y < ch; ++y) {
            memcpy(cur.u + ps.c * f + ps.oc + size_t(y) * g.sc, src + size_t(w) * h + size_t(y) * cw, cw);
            memcpy(cur.v + ps.c * f + ps.oc + size_t(y) * g.sc, src + size_t(w) * h + size_t(cw) * ch + size_t(y) * cw, cw);
        }
    }
    enc.begin_segment(n);
    int rc = enc.encode_device(n);
    if (rc) return rc;
    SegmentOutput out;
    if ((rc = enc.finish(out))) return rc;
    if (const char *reps = getenv("GVC_EMU_TIME_STREAM")) {      // developer aid: time the host bitstream writer alone
        const int n_rep = atoi(reps);
        const auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < n_rep; ++i) { SegmentOutput tmp; enc.write_stream(tmp); }
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        fprintf(stderr, "write_stream: %.3f ms per picture (%d pictures x %d reps)\n", ms / (double(n) * n_rep), n, n_rep);
    }
    if (recon) memcpy(recon, out.recon.data(), out.recon.size());
    if (stats)
        for (int f = 0; f < n; ++f) {
            stats[4 * f] = out.stats[f].skipped; stats[4 * f + 1] = out.stats[f].p16; stats[4 * f + 2] = out.stats[f].p8;
            stats[4 * f + 3] = int(out.au_size[f]);
        }
    if (annexb_len) {   // length-prefixed -> start codes
        size_t o = 0, i = 0;
        while (i + 4 <= out.payload.size()) {
            const size_t len = (size_t(out.payload[i]) << 24) | (size_t(out.payload[i + 1]) << 16) | (size_t(out.payload[i + 2]) << 8) | out.payload[i + 3];
            if (o + 4 + len > annexb_cap) return -8;
            annexb[o] = 0; annexb[o + 1] = 0; annexb[o + 2] = 0; annexb[o + 3] = 1;
            memcpy(annexb + o + 4, out.payload.data() + i + 4, len);
            o += 4 + len; i += 4 + len;
        }
        *annexb_len = o;
    }
    if (mp4_path) {
        std::vector<Mp4Sample> samples;
        for (int f = 0; f < n; ++f) samples.push_back(Mp4Sample{out.au_offset[f], out.au_size[f], f == 0});
        std::string err;
        if (!write_mp4(mp4_path, w, h, fps, 1, out.sps_nal, out.pps_nal, out.payload, samples, err, g_emu_audio.packets.empty() ? nullptr : &g_emu_audio)) return -6;
    }
    return 0;
}
}

// A batch of `count` segments of one shape through ONE SegmentEncoder (they advance together, one emulated launch per stage):
// frames = all segments' pictures back to back, n[s] pictures each.  annexb receives the streams back to back, annexb_len[s] bytes each.
extern "C" int emu_encode_batch(const uint8_t *frames, const int *n, int count, int w, int h, int fps, int qp_i, int qp_p, int merange,
                                uint8_t *annexb, size_t annexb_cap, size_t *annexb_len, int device_entropy) {
    EmuOps        ops;
    SegmentConfig cfg;
    cfg.w = w; cfg.h = h; cfg.fps_num = fps; cfg.fps_den = 1; cfg.qp_i = qp_i; cfg.qp_p = qp_p; cfg.merange = merange;
    cfg.device_entropy = device_entropy != 0;
    cfg.max_batch = count;
    SegmentEncoder<EmuOps> enc(ops, cfg);
    int nmax = 0;
    for (int s = 0; s < count; ++s) nmax = std::max(nmax, n[s]);
    if (enc.reserve(count, nmax)) return -4;
    const EncGeom   &g  = enc.geom();
    const PlaneSizes ps = enc.planes();
    const int cw = (w + 1) / 2, ch = (h + 1) / 2;
    const uint8_t *src = frames;
    for (int s = 0; s < count; ++s) {
        auto cur = enc.cur_frames(s);
        for (int f = 0; f < n[s]; ++f, src += size_t(w) * h * 3 / 2) {
            for (int y = 0; y < h; ++y) memcpy(cur.y + ps.y * f + ps.oy + size_t(y) * g.sy, src + size_t(y) * w, w);
            for (int y = 0; y < ch; ++y) {
                memcpy(cur.u + ps.c * f + ps.oc + size_t(y) * g.sc, src + size_t(w) * h + size_t(y) * cw, cw);
                memcpy(cur.v + ps.c * f + ps.oc + size_t(y) * g.sc, src + size_t(w) * h + size_t(cw) * ch + size_t(y) * cw, cw);
            }
        }
    }
    enc.begin_segment(n, count);
    int rc = enc.encode_device(n, count);
    if (rc) return rc;
    std::vector<SegmentOutput> outs{size_t(count)};
    if ((rc = enc.finish(outs.data()))) return rc;
    size_t o = 0;
    for (int s = 0; s < count; ++s) {
        const std::vector<uint8_t> &pl = outs[size_t(s)].payload;
        size_t i = 0, o0 = o;
        while (i + 4 <= pl.size()) {
            const size_t len = (size_t(pl[i]) << 24) | (size_t(pl[i + 1]) << 16) | (size_t(pl[i + 2]) << 8) | pl[i + 3];
            if (o + 4 + len > annexb_cap) return -8;
            annexb[o] = 0; annexb[o + 1] = 0; annexb[o + 2] = 0; annexb[o + 3] = 1;
            memcpy(annexb + o + 4, pl.data() + i + 4, len);
            o += 4 + len; i += 4 + len;
        }
        annexb_len[s] = o - o0;
    }
    return 0;
}

// ---- media input (product code media_input.cpp) exercised on the host -------------------------------------
#include "../../govideocompressor_b200/csrc/media_input.h"
extern "C" int emu_read_media(const char *path, uint8_t *dst, size_t cap, int *info /* w, h, fps_num, fps_den, frames, has_audio, copyable, packets, rate, channels, hash lo, hash hi */, char *err, int err_cap) {
    std::string e;
    auto src = open_frame_source(path, e);
    if (!src) { snprintf(err, err_cap, "%s", e.c_str()); return -1; }
    info[0] = src->w; info[1] = src->h; info[2] = src->fps_num; info[3] = src->fps_den;
    const size_t fb = size_t(src->w) * src->h + 2 * size_t((src->w + 1) / 2) * ((src->h + 1) / 2);
    int total = 0;
    for (;;) {
        const int room = int((cap - fb * total) / fb);
        if (room <= 0) break;
        const int n = src->read(dst + fb * total, room < 7 ? room : 7, e);     // odd chunk size: reads span calls
        if (n < 0) { snprintf(err, err_cap, "%s", e.c_str()); return -2; }
        if (n == 0) break;
        total += n;
    }
    info[4] = total;
    info[5] = src->has_audio; info[6] = src->audio_copyable; info[7] = int(src->audio.packets.size()); info[8] = src->audio.sample_rate; info[9] = src->audio.channels;
    *reinterpret_cast<unsigned long long *>(info + 10) = audio_hash(src->audio);
    return 0;
}

// ---- the job's argument string (product code ff_args.cpp) parsed on the host ---------------------------------------------------
#include "../../govideocompressor_b200/csrc/ff_args.h"
extern "C" int emu_parse_args(const char *args, int *out /* dst_w, dst_h, crf, scale_algo, gop, merange, fps_num, fps_den, drop_audio, bitrate */, char *err, int err_cap) {
    TranscodeOptions o;
    std::string e;
    const int rc = parse_ffmpeg_args(args, o, e);
    snprintf(err, err_cap, "%s", e.c_str());
    out[0] = o.dst_w; out[1] = o.dst_h; out[2] = o.crf; out[3] = o.scale_algo; out[4] = o.gop; out[5] = o.merange; out[6] = o.fps_num; out[7] = o.fps_den;
    out[8] = o.drop_audio; out[9] = int(o.bitrate);
    return rc;
}

// the product's NAL wrapper (h264_stream.cpp append_nal) on an arbitrary payload: returns the bytes written (Annex B framing)
extern "C" long emu_append_nal(const uint8_t *rbsp, long n, int ref_idc, int type, uint8_t *out, long cap) {
    std::vector<uint8_t> v, in(rbsp, rbsp + n);
    gvc::append_nal(v, ref_idc, type, in, true);
    if (long(v.size()) > cap) return -1;
    memcpy(out, v.data(), v.size());
    return long(v.size());
}
