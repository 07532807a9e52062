// segment_encoder.h -- the per-segment encode loop (product code): orchestrates kernels K2..K7 over one closed-GOP
// segment and turns their per-macroblock records into an H.264 elementary stream.
//
// This is what the reference obtains from `ffmpeg -i <id>.mp4 -vcodec libx264 ... c<id>.mp4` for one piece
// (client.go:101-130): one IDR followed by P pictures, no state shared with other segments (server.go:176-190 hands
// whole pieces to workers), so segments are independent units of work per GPU.
//
// Up to kMaxBatch segments of one shape are encoded TOGETHER: they advance picture by picture in lockstep and every stage
// is one launch for all of them (the launchers take an array of per-segment argument records), so a launch is S times
// larger than a 720p picture's single wave of warps and a step of S segments costs the launches of one.
//
// The loop is a template over `Ops`, the set of stage launchers.  libgvc.so instantiates it with CUDA launchers
// (gvc_segment.cu); the CPU test-suite instantiates it with the host emulation of the same kernel bodies
// (tests/emu) to check the orchestration and the bitstream against a real decoder without a GPU.
#pragma once
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/gvc.h"
#include "enc_host.h"
#include "enc_kernels.cuh"
#include "entropy_kernels.cuh"
#include "h264_stream.h"

namespace gvc {

struct SegmentConfig {
    int  w = 0, h = 0;                  // picture size being encoded (after any resize)
    int  fps_num = 30, fps_den = 1;
    int  qp_i = 20, qp_p = 23;
    int  merange = 16;                  // full-resolution search range; the lowres window is merange/2
    bool keep_recon = false;            // keep every reconstructed picture (tests compare them with a decoder)
    int  host_threads = 1;              // worker threads for entropy coding (pictures are independent)
    bool device_entropy = false;        // CAVLC on the GPU (entropy_kernels.cuh): finished slices cross PCIe instead of records
    int  max_batch = 1;                 // segments encoded together, 1..kMaxBatch
};

struct PlaneSizes { size_t y, c, l; int oy, oc, ol; };   // bytes per plane and offset of sample (0,0)

inline PlaneSizes plane_sizes(const EncGeom &g) {
    PlaneSizes s;
    s.y = size_t(g.sy) * (g.h16 + 2 * kEncPad + 2);
    s.c = size_t(g.sc) * (g.h16 / 2 + kEncPad + 2);
    s.l = size_t(g.sl) * (g.h16 / 2 + kEncPad + 2);
    s.oy = kEncPad * g.sy + kEncPad;
    s.oc = (kEncPad / 2) * g.sc + kEncPad / 2;
    s.ol = (kEncPad / 2) * g.sl + kEncPad / 2;
    return s;
}

struct SegmentOutput {
    std::vector<uint8_t>  sps_nal, pps_nal;       // NAL units without start code / length prefix
    std::vector<uint8_t>  payload;                // length-prefixed NAL units, access unit after access unit
    std::vector<size_t>   au_offset, au_size;     // one access unit per picture
    std::vector<SliceStats> stats;
    std::vector<uint8_t>  recon;                  // keep_recon: w16*h16*3/2 bytes per picture (Y, U, V tight)
};

// Ops concept (all calls asynchronous and in order; sync() waits for everything issued so far).  Stage launchers take
// `count` argument records, one per segment of the batch, and cover them with one launch:
//   void *alloc(size_t bytes);  void release(void *p);   void *alloc_host(size_t bytes);  void release_host(void *p);   // pinned
//   void *device_alias(void *pinned_host);   // the device address of mapped pinned memory
//   int clear(void *dev, size_t bytes);
//   int extend_border3(const BorderBatchArgs *a, int count);   int lowres_batch(const LowresBatchArgs *a, int count);
//   int hpel(const HpelArgs *a, int count);  int me_lowres(const MeLowArgs *a, int count);  int me_refine(const MeRefArgs *a, int count);
//   int encode_inter(const EncMbArgs *a, int count);  int encode_intra(const IntraArgs *a, int count);
//   int signal_picture(const SignalArgs *a, int count);                              // *count_out = *count_in; fence; *done = 1
//   int entropy_picture(const EntropyPicture *p, int count);                         // device_entropy: the pictures' slices -> host buffers
//   int side_begin();  int side_end(int slot);  int wait_side(int slot);            // fork / join of work the picture chain does not wait for
//   int prep_fork();  int prep_on(int branch);  int prep_join();                     // three concurrent branches (0 = the chain itself) and their join
//   int copy_plane(uint8_t *host_or_dev_dst, int dpitch, const uint8_t *dev_src, int spitch, int w, int h);   // device -> recon store
//   int sync();
template <class Ops>
class SegmentEncoder {
public:
    SegmentEncoder(Ops &ops, const SegmentConfig &cfg) : ops_(ops), cfg_(cfg), g_(make_geom(cfg.w, cfg.h)), ps_(plane_sizes(g_)) {
        cfg_.max_batch = std::max(1, std::min(cfg_.max_batch, kMaxBatch));
        lanes_.resize(size_t(cfg_.max_batch));
        // How much of the planes' padding anything ever reads: a vector reaches merange (twice the lowres window) + 2 (the 5x5 grid)
        // + 1 (the sample right of / below a quarter position) samples beyond the picture, chroma half of that + 1 (bilinear), the
        // lowres search its window.  Border extension, half-sample planes and lowres planes are produced only that far (rounded up to
        // the kernels' vector widths) instead of over the whole kEncPad.
        const int reach = cfg_.merange + 3;
        pad_y_ = std::min(kEncPad, (reach + 1 + 15) / 16 * 16);
        pad_c_ = std::min(kEncPad / 2, (reach / 2 + 2 + 15) / 16 * 16);
        pad_l_ = std::min(kEncPad / 2, (cfg_.merange / 2 + 4 + 3) / 4 * 4);
    }
    ~SegmentEncoder() { free_all(); }

    const EncGeom &geom() const { return g_; }
    const PlaneSizes &planes() const { return ps_; }
    int max_batch() const { return cfg_.max_batch; }
    void set_qp(int qp_i, int qp_p) { cfg_.qp_i = qp_i; cfg_.qp_p = qp_p; }      // takes effect with the next encode_device

    // Source pictures in the padded layout (interiors filled, borders not yet extended): device arrays holding
    // n planes each, ps_.y / ps_.c bytes apart.  The resize stage writes straight into these.
    struct CurFrames { uint8_t *y, *u, *v; };

    // Room for `lanes` segments of up to n_frames pictures each.
    int reserve(int lanes, int n_frames) {
        if (lanes < 1 || lanes > cfg_.max_batch) return GVC_ERR_INVALID_ARG;
        if (n_frames > cap_) { free_all(); cap_ = n_frames; }
        for (int s = 0; s < lanes; ++s)
            if (!lanes_[size_t(s)].ready) {
                const int rc = reserve_lane(lanes_[size_t(s)], cap_);
                if (rc) { free_all(); return rc; }
            }
        return 0;
    }
    int reserve(int n_frames) { return reserve(1, n_frames); }
    CurFrames cur_frames(int lane = 0) const { return lanes_[size_t(lane)].cur; }
    void set_frame_counts(const int *n, int lanes) { for (int s = 0; s < lanes; ++s) lanes_[size_t(s)].n = n[s]; active_ = lanes; }
    int lanes_in_flight() const { return active_; }
    // bytes that crossed PCIe towards the host for the last finished segment of a lane: packed blocks + per-macroblock tables + flags
    size_t record_bytes(int lane = 0) const {
        const Lane &L = lanes_[size_t(lane)];
        if (cfg_.device_entropy) {
            size_t bytes = 0;
            for (int f = 0; f < L.n; ++f) bytes += ((L.h_pos[8 * f + 1] + 15u) & ~15u) + 32 + 8;
            return bytes;
        }
        size_t blocks = 0;
        for (int f = 0; f < L.n; ++f) blocks += L.h_counters[f];
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        return blocks * 32 + size_t(L.n) * (nmb * (4 + 4) + 8) + size_t(L.n - 1) * nmb * sizeof(MbMotion) + nmb;
    }

    // Host-side reset of the per-picture arrival flags; call before issuing (or replaying) a batch's kernels.
    void begin_segment(const int *n, int lanes) {
        for (int s = 0; s < lanes; ++s)
            for (int f = 0; f < n[s]; ++f) __atomic_store_n(&lanes_[size_t(s)].h_done[f], 0, __ATOMIC_RELEASE);
    }
    void begin_segment(int n) { begin_segment(&n, 1); }

private:
    struct Lane {                 // everything one segment of the batch owns
        bool      ready = false;
        int       n = 0, idr_count = 0;
        bool      failed = false;
        CurFrames cur{nullptr, nullptr, nullptr};
        uint8_t  *cur_low = nullptr;
        uint8_t  *rec_y[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
        uint8_t  *rec_u[2] = {nullptr, nullptr}, *rec_v[2] = {nullptr, nullptr}, *rec_low[2] = {nullptr, nullptr};
        MbMotion *d_motion = nullptr;
        int16_t  *d_lowmv = nullptr;
        unsigned long long *d_mail = nullptr;   // intra wavefront mailboxes
        // Records for the host entropy coder live in pinned memory mapped into the device address space; the kernels store
        // into it directly (no copy engine on the pictures' critical path).
        int16_t  *h_blocks = nullptr;    // the nonzero blocks of every picture, packed
        uint32_t *h_offsets = nullptr;   // per MB: first packed block
        unsigned *d_counters = nullptr, *h_counters = nullptr;   // per picture: blocks handed out
        uint32_t *h_nz = nullptr;
        MbMotion *h_motion = nullptr;
        uint8_t  *h_modes = nullptr;
        int      *h_done = nullptr;      // per picture: set to 1 by the kernel that closes the picture
        // device_entropy: one picture's records and the entropy pass's scratch on the device, finished slices on the host
        int16_t  *e_blocks = nullptr;
        uint32_t *e_offsets = nullptr, *e_nz = nullptr, *e_mb_buf = nullptr, *e_mb_bits = nullptr, *e_slice = nullptr, *e_info = nullptr;
        uint8_t  *e_modes = nullptr, *e_nnz = nullptr, *e_mb_kind = nullptr;
        unsigned *e_used = nullptr;
        uint8_t  *h_slices = nullptr;    // the segment's slices, back to back (each padded to 16 bytes)
        uint32_t *h_pos = nullptr;       // per picture: base, bytes, bits, skipped, p16, p8, intra, error
        std::vector<uint8_t> recon_store;
    };

    int reserve_lane(Lane &L, int n_frames) {
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        L.cur.y = static_cast<uint8_t *>(ops_.alloc(ps_.y * n_frames));
        L.cur.u = static_cast<uint8_t *>(ops_.alloc(ps_.c * n_frames));
        L.cur.v = static_cast<uint8_t *>(ops_.alloc(ps_.c * n_frames));
        L.cur_low = static_cast<uint8_t *>(ops_.alloc(ps_.l * n_frames));
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) L.rec_y[i][k] = static_cast<uint8_t *>(ops_.alloc(ps_.y));
            L.rec_u[i] = static_cast<uint8_t *>(ops_.alloc(ps_.c));
            L.rec_v[i] = static_cast<uint8_t *>(ops_.alloc(ps_.c));
            L.rec_low[i] = static_cast<uint8_t *>(ops_.alloc(ps_.l));
        }
        L.d_motion = static_cast<MbMotion *>(ops_.alloc(sizeof(MbMotion) * nmb * n_frames));
        L.d_lowmv  = static_cast<int16_t *>(ops_.alloc(sizeof(int16_t) * 2 * nmb));
        L.d_mail   = static_cast<unsigned long long *>(ops_.alloc(sizeof(unsigned long long) * 8 * nmb));
        L.d_counters = static_cast<unsigned *>(ops_.alloc(sizeof(unsigned) * n_frames));
        L.h_done   = static_cast<int *>(ops_.alloc_host(sizeof(int) * n_frames));
        L.h_counters = static_cast<unsigned *>(ops_.alloc_host(sizeof(unsigned) * n_frames));
        bool ok = L.cur.y && L.cur.u && L.cur.v && L.cur_low && L.d_motion && L.d_lowmv && L.d_mail && L.d_counters && L.h_counters && L.h_done;
        if (!cfg_.device_entropy) {     // records for the host entropy coder
            L.h_blocks = static_cast<int16_t *>(ops_.alloc_host(sizeof(int16_t) * kMbCoefs * nmb * n_frames));   // worst case: every block nonzero
            L.h_offsets = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * nmb * n_frames));
            L.h_nz     = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * nmb * n_frames));
            L.h_motion = static_cast<MbMotion *>(ops_.alloc_host(sizeof(MbMotion) * nmb * n_frames));
            L.h_modes  = static_cast<uint8_t *>(ops_.alloc_host(nmb));
            ok = ok && L.h_blocks && L.h_offsets && L.h_nz && L.h_motion && L.h_modes;
        } else {                        // one picture's records stay on the device; finished slices go to the host
            // two sets of records: the entropy pass of picture f runs beside the kernels of picture f + 1
            L.e_blocks  = static_cast<int16_t *>(ops_.alloc(sizeof(int16_t) * kMbCoefs * nmb * 2));
            L.e_offsets = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 2));
            L.e_nz      = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 2));
            L.e_modes   = static_cast<uint8_t *>(ops_.alloc(nmb));
            L.e_nnz     = static_cast<uint8_t *>(ops_.alloc(24 * nmb * 2));
            L.e_mb_buf  = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * kMbBufWords * nmb));
            L.e_mb_bits = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 3));      // bits, offsets, runs
            L.e_mb_kind = static_cast<uint8_t *>(ops_.alloc(nmb));
            slice_words_ = int(std::min<size_t>(size_t(kMbBufWords) * nmb + 64, size_t(1) << 28));
            L.e_slice   = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * slice_words_));
            L.e_info    = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * 8));
            L.e_used    = static_cast<unsigned *>(ops_.alloc(sizeof(unsigned) * (n_frames + 1)));
            slices_cap_ = size_t(g_.w16) * g_.h16 * 3 * n_frames + 65536;   // twice the raw pictures (noise at QP 0 reaches 1.8x)
            L.h_slices  = static_cast<uint8_t *>(ops_.alloc_host(slices_cap_));
            L.h_pos     = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * 8 * n_frames));
            ok = ok && L.e_blocks && L.e_offsets && L.e_nz && L.e_modes && L.e_nnz && L.e_mb_buf && L.e_mb_bits && L.e_mb_kind && L.e_slice && L.e_info &&
                 L.e_used && L.h_slices && L.h_pos;
        }
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) ok = ok && L.rec_y[i][k];
            ok = ok && L.rec_u[i] && L.rec_v[i] && L.rec_low[i];
        }
        if (!ok) return GVC_ERR_NOMEM;
        L.ready = true;
        return 0;
    }
    void free_lane(Lane &L) {
        auto rel = [&](auto *&p) { if (p) ops_.release(p); p = nullptr; };
        rel(L.cur.y); rel(L.cur.u); rel(L.cur.v); rel(L.cur_low);
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) rel(L.rec_y[i][k]);
            rel(L.rec_u[i]); rel(L.rec_v[i]); rel(L.rec_low[i]);
        }
        rel(L.d_motion); rel(L.d_lowmv); rel(L.d_mail);
        auto relh = [&](auto *&p) { if (p) ops_.release_host(p); p = nullptr; };
        rel(L.d_counters); relh(L.h_offsets); relh(L.h_counters);
        relh(L.h_blocks); relh(L.h_nz); relh(L.h_motion); relh(L.h_modes); relh(L.h_done);
        rel(L.e_blocks); rel(L.e_offsets); rel(L.e_nz); rel(L.e_modes); rel(L.e_nnz); rel(L.e_mb_buf); rel(L.e_mb_bits); rel(L.e_mb_kind);
        rel(L.e_slice); rel(L.e_info); rel(L.e_used); relh(L.h_slices); relh(L.h_pos);
        L.ready = false;
    }
    void free_all() {
        for (Lane &L : lanes_) free_lane(L);
        cap_ = 0;
    }

public:
    // Issues every kernel of the batch: pictures 0..n[s]-1 of cur_frames(s) for each segment s < lanes, all segments advancing
    // together (a segment shorter than the longest simply drops out of the later launches).  Behind each picture's encode kernel
    // come its entropy pass / record hand-over and an arrival flag.  Nothing is waited for.
    int encode_device(const int *n, int lanes) {
        if (lanes < 1 || lanes > cfg_.max_batch) return GVC_ERR_INVALID_ARG;
        int rc, nmax = 0;
        for (int s = 0; s < lanes; ++s) {
            if (n[s] < 1 || n[s] > cap_ || !lanes_[size_t(s)].ready) return GVC_ERR_INVALID_ARG;
            nmax = std::max(nmax, n[s]);
        }
        set_frame_counts(n, lanes);
        const int P = pad_y_;
        const int lw = g_.w16 / 2, lh = g_.h16 / 2;
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        const bool de = cfg_.device_entropy;
        // records of picture f: its own area of the host tables, or device set f & 1
        auto rset = [&](int f) { return size_t(de ? (f & 1) : f); };
        {   // source pictures: extend the picture edges into the padding and build lowres, one launch each for the whole batch
            BorderBatchArgs b[kMaxBatch];
            LowresBatchArgs l[kMaxBatch];
            for (int s = 0; s < lanes; ++s) {
                Lane &L = lanes_[size_t(s)];
                // source pictures are only read inside the coded picture: fill what lies between the picture and the coded size,
                // lowres of the coded area only
                b[s] = border_args(L.cur.y + ps_.oy, L.cur.u + ps_.oc, L.cur.v + ps_.oc, cfg_.w, cfg_.h);
                b[s].x_lo[0] = b[s].y_lo[0] = 0; b[s].x_hi[0] = g_.w16; b[s].y_hi[0] = g_.h16;
                for (int k = 1; k < 3; ++k) { b[s].x_lo[k] = b[s].y_lo[k] = 0; b[s].x_hi[k] = lw; b[s].y_hi[k] = lh; }
                b[s].frame_stride[0] = ps_.y; b[s].frame_stride[1] = b[s].frame_stride[2] = ps_.c; b[s].n_frames = n[s];
                l[s] = lowres_args(L.cur.y + ps_.oy, L.cur_low + ps_.ol);
                l[s].x_lo = 0; l[s].x_hi = g_.w16 - 1; l[s].y_lo = 0; l[s].y_hi = g_.h16 - 1;
                l[s].lx_lo = 0; l[s].lx_hi = lw; l[s].ly_lo = 0; l[s].ly_hi = lh;
                l[s].sframe = ps_.y; l[s].dframe = ps_.l; l[s].n_frames = n[s];
            }
            if ((rc = ops_.extend_border3(b, lanes))) return rc;
            if ((rc = ops_.lowres_batch(l, lanes))) return rc;
        }
        for (int s = 0; s < lanes; ++s) {
            Lane &L = lanes_[size_t(s)];
            if ((rc = ops_.clear(L.d_counters, sizeof(unsigned) * n[s]))) return rc;
            if (de && (rc = ops_.clear(L.e_used, sizeof(unsigned)))) return rc;
            if (cfg_.keep_recon) L.recon_store.assign(size_t(g_.w16) * g_.h16 * 3 / 2 * n[s], 0);
        }
        QuantParams qi, qp;
        fill_quant_params(qi, cfg_.qp_i, true);
        fill_quant_params(qp, cfg_.qp_p, false);
        for (int f = 0; f < nmax; ++f) {
            const int r = f & 1, p = r ^ 1;
            int act[kMaxBatch], na = 0;          // segments that have a picture f
            for (int s = 0; s < lanes; ++s)
                if (f < n[s]) act[na++] = s;
            if (de && f >= 2 && (rc = ops_.wait_side(f & 1))) return rc;      // the entropy pass of picture f - 2 has read this record set
            if (f == 0) {
                IntraArgs ia[kMaxBatch];
                for (int i = 0; i < na; ++i) {
                    Lane &L = lanes_[size_t(act[i])];
                    IntraArgs &a = ia[i];
                    a.cur_y = L.cur.y + ps_.oy; a.cur_u = L.cur.u + ps_.oc; a.cur_v = L.cur.v + ps_.oc;
                    a.rec_y = L.rec_y[r][0] + ps_.oy; a.rec_u = L.rec_u[r] + ps_.oc; a.rec_v = L.rec_v[r] + ps_.oc;
                    a.sy = g_.sy; a.sc = g_.sc; a.mbw = g_.mbw; a.mbh = g_.mbh;
                    a.q = qi;
                    a.coefs = nullptr; a.mail = L.d_mail;
                    a.nzmask = de ? L.e_nz : static_cast<uint32_t *>(ops_.device_alias(L.h_nz));
                    a.modes = de ? L.e_modes : static_cast<uint8_t *>(ops_.device_alias(L.h_modes));
                    a.pack.blocks = de ? L.e_blocks : static_cast<int16_t *>(ops_.device_alias(L.h_blocks));
                    a.pack.offsets = de ? L.e_offsets : static_cast<uint32_t *>(ops_.device_alias(L.h_offsets));
                    a.pack.counter = L.d_counters;
                    a.nnz_y = de ? L.e_nnz : nullptr; a.nnz_u = de ? a.nnz_y + 16 * nmb : nullptr; a.nnz_v = de ? a.nnz_y + 20 * nmb : nullptr;
                }
                if ((rc = ops_.encode_intra(ia, na))) return rc;
            } else {
                // the lowres search of this picture was issued with the reference's plane preparation (below, previous
                // iteration), beside the half-sample filter and the border extension: wait for all three branches
                if ((rc = ops_.prep_join())) return rc;
                MeRefArgs mr[kMaxBatch];
                EncMbArgs ea[kMaxBatch];
                for (int i = 0; i < na; ++i) {
                    Lane &L = lanes_[size_t(act[i])];
                    RefPlanes ref;
                    ref.y = L.rec_y[p][0] + ps_.oy; ref.hh = L.rec_y[p][1] + ps_.oy; ref.hv = L.rec_y[p][2] + ps_.oy; ref.hc = L.rec_y[p][3] + ps_.oy;
                    ref.u = L.rec_u[p] + ps_.oc; ref.v = L.rec_v[p] + ps_.oc; ref.low = L.rec_low[p] + ps_.ol;
                    uint8_t *cy = L.cur.y + ps_.y * f + ps_.oy, *cu = L.cur.u + ps_.c * f + ps_.oc, *cv = L.cur.v + ps_.c * f + ps_.oc;
                    MeRefArgs &m = mr[i];
                    m.cur = cy; m.ref = ref; m.stride = g_.sy; m.w16 = g_.w16; m.h16 = g_.h16; m.mbw = g_.mbw; m.mbh = g_.mbh;
                    m.lambda = qp.lambda; m.lowres_mv = L.d_lowmv; m.out = L.d_motion + nmb * f;
                    m.out_host = de ? nullptr : static_cast<MbMotion *>(ops_.device_alias(L.h_motion)) + nmb * f;
                    EncMbArgs &e = ea[i];
                    e.cur_y = cy; e.cur_u = cu; e.cur_v = cv; e.ref = ref;
                    e.rec_y = L.rec_y[r][0] + ps_.oy; e.rec_u = L.rec_u[r] + ps_.oc; e.rec_v = L.rec_v[r] + ps_.oc;
                    e.sy = g_.sy; e.sc = g_.sc; e.mbw = g_.mbw; e.mbh = g_.mbh; e.q = qp;
                    e.motion = L.d_motion + nmb * f; e.coefs = nullptr;      // levels go straight to the packed blocks
                    e.nzmask = (de ? L.e_nz : static_cast<uint32_t *>(ops_.device_alias(L.h_nz))) + nmb * rset(f);
                    e.pack.blocks = (de ? L.e_blocks : static_cast<int16_t *>(ops_.device_alias(L.h_blocks))) + size_t(kMbCoefs) * nmb * rset(f);
                    e.pack.offsets = (de ? L.e_offsets : static_cast<uint32_t *>(ops_.device_alias(L.h_offsets))) + nmb * rset(f);
                    e.pack.counter = L.d_counters + f;
                    e.nnz_y = de ? L.e_nnz + 24 * nmb * rset(f) : nullptr; e.nnz_u = de ? e.nnz_y + 16 * nmb : nullptr; e.nnz_v = de ? e.nnz_y + 20 * nmb : nullptr;
                }
                if ((rc = ops_.me_refine(mr, na))) return rc;
                if ((rc = ops_.encode_inter(ea, na))) return rc;
            }
            // Nothing below is waited for by the next picture's kernels: it runs beside them.  device_entropy: the pictures' slice
            // data (macroblock bits, scan, placement, copy to the host buffer).  Then a small kernel copies the block counts and
            // raises the arrival flags (stream order + a system-wide fence: everything has landed by then).
            if ((rc = ops_.side_begin())) return rc;
            if (de) {
                EntropyPicture ep[kMaxBatch];
                for (int i = 0; i < na; ++i) {
                    Lane &L = lanes_[size_t(act[i])];
                    EntropyPicture &e = ep[i];
                    uint8_t *nnz_y = L.e_nnz + 24 * nmb * rset(f);
                    e.ent.mbw = g_.mbw; e.ent.mbh = g_.mbh; e.ent.idr = f == 0;
                    e.ent.nzmask = L.e_nz + nmb * rset(f); e.ent.motion = L.d_motion + nmb * f; e.ent.modes = L.e_modes;
                    e.ent.blocks = L.e_blocks + size_t(kMbCoefs) * nmb * rset(f); e.ent.offsets = L.e_offsets + nmb * rset(f);
                    e.ent.nnz_y = nnz_y; e.ent.nnz_c[0] = nnz_y + 16 * nmb; e.ent.nnz_c[1] = nnz_y + 20 * nmb;
                    e.ent.mb_buf = L.e_mb_buf; e.ent.mb_bits = L.e_mb_bits; e.ent.mb_kind = L.e_mb_kind;
                    e.slice.nmb = int(nmb); e.slice.mb_buf = L.e_mb_buf; e.slice.mb_bits = L.e_mb_bits; e.slice.mb_kind = L.e_mb_kind;
                    e.slice.mb_off = L.e_mb_bits + nmb; e.slice.mb_run = L.e_mb_bits + 2 * nmb;
                    e.slice.slice = L.e_slice; e.slice.slice_words = slice_words_;
                    e.slice.header_bits = int(slice_header_bits(picture_header(L, f)));
                    e.slice.p_slice = f != 0; e.slice.info = L.e_info;
                    e.host_buf = static_cast<uint8_t *>(ops_.device_alias(L.h_slices)); e.host_cap = unsigned(std::min<size_t>(slices_cap_, 0xfffffff0u));
                    e.used = L.e_used + f; e.pos = static_cast<uint32_t *>(ops_.device_alias(L.h_pos)) + 8 * f;
                }
                if ((rc = ops_.entropy_picture(ep, na))) return rc;
            }
            {
                SignalArgs sg[kMaxBatch];
                for (int i = 0; i < na; ++i) {
                    Lane &L = lanes_[size_t(act[i])];
                    sg[i].done = static_cast<int *>(ops_.device_alias(L.h_done)) + f;
                    sg[i].count_out = static_cast<unsigned *>(ops_.device_alias(L.h_counters)) + f;
                    sg[i].count_in = L.d_counters + f;
                }
                if ((rc = ops_.signal_picture(sg, na))) return rc;
            }
            if ((rc = ops_.side_end(f & 1))) return rc;
            if (cfg_.keep_recon)
                for (int i = 0; i < na; ++i) {
                    Lane &L = lanes_[size_t(act[i])];
                    uint8_t *dst = L.recon_store.data() + size_t(g_.w16) * g_.h16 * 3 / 2 * f;
                    if ((rc = ops_.copy_plane(dst, g_.w16, L.rec_y[r][0] + ps_.oy, g_.sy, g_.w16, g_.h16))) return rc;
                    if ((rc = ops_.copy_plane(dst + size_t(g_.w16) * g_.h16, lw, L.rec_u[r] + ps_.oc, g_.sc, lw, lh))) return rc;
                    if ((rc = ops_.copy_plane(dst + size_t(g_.w16) * g_.h16 * 5 / 4, lw, L.rec_v[r] + ps_.oc, g_.sc, lw, lh))) return rc;
                }
            int nxt[kMaxBatch], nn = 0;          // segments that have a picture f + 1
            for (int s = 0; s < lanes; ++s)
                if (f + 1 < n[s]) nxt[nn++] = s;
            if (nn) {
                // The reconstruction becomes the next picture's reference.  Three independent branches (the half-sample and
                // lowres kernels replicate the picture edge themselves instead of reading the extended border): border
                // extension | lowres planes, then the next picture's lowres search | half-sample planes.
                LowresBatchArgs la[kMaxBatch];
                MeLowArgs       ml[kMaxBatch];
                HpelArgs        ha[kMaxBatch];
                BorderBatchArgs bb[kMaxBatch];
                for (int i = 0; i < nn; ++i) {
                    Lane &L = lanes_[size_t(nxt[i])];
                    uint8_t *ry = L.rec_y[r][0] + ps_.oy;
                    la[i] = lowres_args(ry, L.rec_low[r] + ps_.ol);
                    la[i].x_lo = 0; la[i].x_hi = g_.w16 - 1; la[i].y_lo = 0; la[i].y_hi = g_.h16 - 1;
                    ml[i].cur = L.cur_low + ps_.l * (f + 1) + ps_.ol; ml[i].ref = L.rec_low[r] + ps_.ol; ml[i].stride = g_.sl; ml[i].mbw = g_.mbw; ml[i].mbh = g_.mbh;
                    ml[i].r = cfg_.merange / 2; ml[i].lambda = qp.lambda; ml[i].out_mv = L.d_lowmv;
                    ha[i].src = ry; ha[i].dh = L.rec_y[r][1] + ps_.oy; ha[i].dv = L.rec_y[r][2] + ps_.oy; ha[i].dc = L.rec_y[r][3] + ps_.oy;
                    ha[i].stride = g_.sy; ha[i].x_lo = -P; ha[i].x_hi = g_.w16 + P; ha[i].y_lo = -P; ha[i].y_hi = g_.h16 + P;
                    ha[i].cx_lo = 0; ha[i].cx_hi = g_.w16; ha[i].cy_lo = 0; ha[i].cy_hi = g_.h16;
                    bb[i] = border_args(ry, L.rec_u[r] + ps_.oc, L.rec_v[r] + ps_.oc, g_.w16, g_.h16);
                }
                if ((rc = ops_.prep_fork())) return rc;
                if ((rc = ops_.prep_on(1))) return rc;
                if ((rc = ops_.lowres_batch(la, nn))) return rc;
                if ((rc = ops_.me_lowres(ml, nn))) return rc;
                if ((rc = ops_.prep_on(2))) return rc;
                if ((rc = ops_.hpel(ha, nn))) return rc;
                if ((rc = ops_.prep_on(0))) return rc;
                if ((rc = ops_.extend_border3(bb, nn))) return rc;
            }
        }
        // the chain ends where the side work ends (one stream to wait on; a captured graph must rejoin its origin)
        if ((rc = ops_.wait_side(0))) return rc;
        if (nmax >= 2 && (rc = ops_.wait_side(1))) return rc;
        return 0;
    }
    int encode_device(int n) { return encode_device(&n, 1); }

    // Completes the batch: outs[s] receives segment s.  Host entropy coding: worker threads code the pictures as their records
    // arrive (they wait on the arrival flags while the GPU is still working on later pictures), then the stream is waited for.
    int finish(SegmentOutput *outs) {
        int rc;
        for (int s = 0; s < active_; ++s) lanes_[size_t(s)].failed = false;
        if (cfg_.device_entropy) {      // the slices arrive finished: wait once (no polling thread), then wrap them
            if ((rc = ops_.sync())) return rc;
            if (active_ == 1) {
                write_stream(0, outs[0]);
            } else {                    // the segments of a batch share nothing: wrap them side by side
                std::vector<std::thread> pool;
                for (int s = 0; s < active_; ++s) pool.emplace_back([this, s, outs] { write_stream(s, outs[s]); });
                for (std::thread &th : pool) th.join();
            }
        } else {                        // host entropy coding overlaps the GPU: workers start on a picture when its flag is up
            for (int s = 0; s < active_; ++s) write_stream(s, outs[s]);
            if ((rc = ops_.sync())) return rc;
        }
        for (int s = 0; s < active_; ++s) {
            if (lanes_[size_t(s)].failed) return GVC_ERR_TOO_LARGE;     // the slices outgrew the buffers sized from the raw pictures
            if (cfg_.keep_recon) outs[s].recon.swap(lanes_[size_t(s)].recon_store);
        }
        return 0;
    }
    int finish(SegmentOutput &out) { return finish(&out); }

    void write_stream(int lane, SegmentOutput &out) {
        Lane &L = lanes_[size_t(lane)];
        const int n_ = L.n;
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        StreamParams sp;
        sp.w = cfg_.w; sp.h = cfg_.h; sp.w16 = g_.w16; sp.h16 = g_.h16; sp.fps_num = cfg_.fps_num; sp.fps_den = cfg_.fps_den;
        sp.level_idc = pick_level(g_.mbw, g_.mbh, cfg_.fps_num, cfg_.fps_den);
        {
            BitWriter s, p;
            write_sps(s, sp);
            write_pps(p);
            std::vector<uint8_t> tmp;
            append_nal(tmp, 3, 7, s.bytes(), false);
            out.sps_nal.assign(tmp.begin() + 4, tmp.end());
            tmp.clear();
            append_nal(tmp, 3, 8, p.bytes(), false);
            out.pps_nal.assign(tmp.begin() + 4, tmp.end());
        }
        out.payload.clear(); out.au_offset.clear(); out.au_size.clear(); out.stats.clear();
        // pictures are independent for entropy coding: one worker thread per stripe of pictures
        std::vector<std::vector<uint8_t>> au(n_);
        out.stats.assign(n_, SliceStats());
        auto code_picture = [&](int f) {
            while (__atomic_load_n(&L.h_done[f], __ATOMIC_ACQUIRE) == 0) std::this_thread::yield();
            PictureData pic = picture_header(L, f);
            if (cfg_.device_entropy) {      // the GPU wrote slice_data() behind room for the header: fill the header in, wrap the NAL
                const uint32_t *pos = L.h_pos + 8 * f;
                if (pos[7]) { L.failed = true; return; }
                std::vector<uint8_t> rbsp(L.h_slices + pos[0], L.h_slices + pos[0] + pos[1]);
                const std::vector<uint8_t> hdr = slice_header_bytes(pic);
                for (size_t i = 0; i < hdr.size() && i < rbsp.size(); ++i) rbsp[i] |= hdr[i];
                SliceStats st;
                st.skipped = int(pos[3]); st.p16 = int(pos[4]); st.p8 = int(pos[5]); st.intra = int(pos[6]); st.bits = size_t(pos[1]) * 8;
                out.stats[f] = st;
                if (f == 0) {
                    BitWriter s, p;
                    write_sps(s, sp);
                    write_pps(p);
                    append_nal(au[f], 3, 7, s.bytes(), false);
                    append_nal(au[f], 3, 8, p.bytes(), false);
                }
                append_nal(au[f], f == 0 ? 3 : 2, f == 0 ? 5 : 1, rbsp, false);
                return;
            }
            pic.coefs = L.h_blocks + size_t(kMbCoefs) * nmb * f;
            pic.compact = true;
            pic.block_offsets = L.h_offsets + nmb * f;
            pic.nzmask = L.h_nz + nmb * f;
            pic.motion = L.h_motion + nmb * f;
            pic.modes = L.h_modes;
            BitWriter bw;
#if defined(GVC_MEASURE)
            static const bool skip_entropy = getenv("GVC_SKIP_ENTROPY") != nullptr;   // measurement aid: stream is not decodable
            if (skip_entropy) bw.rbsp_trailing();
            else
#endif
            write_slice(bw, pic, &out.stats[f]);
            if (f == 0) {   // parameter sets travel in-band with the IDR too (Annex B extraction stays decodable)
                BitWriter s, p;
                write_sps(s, sp);
                write_pps(p);
                append_nal(au[f], 3, 7, s.bytes(), false);
                append_nal(au[f], 3, 8, p.bytes(), false);
            }
            append_nal(au[f], f == 0 ? 3 : 2, f == 0 ? 5 : 1, bw.bytes(), false);
        };
        const int workers = cfg_.device_entropy ? 1 : std::max(1, std::min(n_, std::min(cfg_.host_threads, 64)));
        if (workers == 1) {
            for (int f = 0; f < n_; ++f) code_picture(f);
        } else {
            std::vector<std::thread> pool;
            for (int t = 0; t < workers; ++t)
                pool.emplace_back([&, t] { for (int f = t; f < n_; f += workers) code_picture(f); });
            for (std::thread &th : pool) th.join();
        }
        for (int f = 0; f < n_; ++f) {
            out.au_offset.push_back(out.payload.size());
            out.au_size.push_back(au[f].size());
            out.payload.insert(out.payload.end(), au[f].begin(), au[f].end());
        }
        ++L.idr_count;
    }
    void write_stream(SegmentOutput &out) { write_stream(0, out); }

private:
    // the fields of picture f's slice header.  idr_pic_id runs through 255..510: consecutive segments differ and ue(v) always
    // takes 17 bits, so the header length is a constant of the picture type (the GPU entropy pass is replayed from a graph).
    PictureData picture_header(const Lane &L, int f) const {
        PictureData pic;
        pic.mbw = g_.mbw; pic.mbh = g_.mbh; pic.idr = f == 0; pic.frame_num = f; pic.idr_pic_id = 255 + (L.idr_count & 255);
        pic.qp = f == 0 ? cfg_.qp_i : cfg_.qp_p;
        return pic;
    }
    // Y/U/V planes at sample (0,0) of one picture whose valid area is w x h luma samples
    BorderBatchArgs border_args(uint8_t *y, uint8_t *u, uint8_t *v, int w, int h) const {
        const int P = pad_y_, Pc = pad_c_, lw = g_.w16 / 2, lh = g_.h16 / 2;
        BorderBatchArgs b;
        b.p[0] = y; b.p[1] = u; b.p[2] = v;
        b.stride[0] = g_.sy; b.stride[1] = b.stride[2] = g_.sc;
        b.w[0] = w; b.h[0] = h; b.w[1] = b.w[2] = (w + 1) / 2; b.h[1] = b.h[2] = (h + 1) / 2;
        b.x_lo[0] = -P; b.x_hi[0] = g_.w16 + P; b.y_lo[0] = -P; b.y_hi[0] = g_.h16 + P;
        for (int k = 1; k < 3; ++k) { b.x_lo[k] = -Pc; b.x_hi[k] = lw + Pc; b.y_lo[k] = -Pc; b.y_hi[k] = lh + Pc; }
        b.frame_stride[0] = b.frame_stride[1] = b.frame_stride[2] = 0;
        b.n_frames = 1;
        return b;
    }
    LowresBatchArgs lowres_args(const uint8_t *y, uint8_t *low) const {
        const int P = pad_y_, Pc = pad_l_, lw = g_.w16 / 2, lh = g_.h16 / 2;
        LowresBatchArgs l;
        l.src = y; l.dst = low; l.sstride = g_.sy; l.dstride = g_.sl;
        l.x_lo = -P; l.x_hi = g_.w16 + P - 1; l.y_lo = -P; l.y_hi = g_.h16 + P - 1;
        l.lx_lo = -Pc; l.lx_hi = lw + Pc; l.ly_lo = -Pc; l.ly_hi = lh + Pc;
        l.sframe = l.dframe = 0; l.n_frames = 1;
        return l;
    }

    Ops          &ops_;
    SegmentConfig cfg_;
    EncGeom       g_;
    PlaneSizes    ps_;
    int           cap_ = 0, active_ = 0;
    int           pad_y_ = kEncPad, pad_c_ = kEncPad / 2, pad_l_ = kEncPad / 2;   // padding actually produced (luma, chroma, lowres)
    int           slice_words_ = 0;
    size_t        slices_cap_ = 0;
    std::vector<Lane> lanes_;
};

}  // namespace gvc
