// segment_encoder.h -- the per-segment encode loop (product code): orchestrates kernels K2..K7 over one closed-GOP
// segment and turns their per-macroblock records into an H.264 elementary stream.
//
// This is what the reference obtains from `ffmpeg -i <id>.mp4 -vcodec libx264 ... c<id>.mp4` for one piece
// (client.go:101-130): one IDR followed by P pictures, no state shared with other segments (server.go:176-190 hands
// whole pieces to workers), so segments are independent units of work per GPU.
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

// Ops concept (all calls asynchronous and in order; sync() waits for everything issued so far):
//   void *alloc(size_t bytes);  void release(void *p);   void *alloc_host(size_t bytes);  void release_host(void *p);   // pinned
//   void *device_alias(void *pinned_host);   // the device address of mapped pinned memory
//   int clear(void *dev, size_t bytes);
//   int extend_border3(const BorderBatchArgs &a);   int lowres_batch(const LowresBatchArgs &a);
//   int hpel(const HpelArgs &a);  int me_lowres(const MeLowArgs &a);  int me_refine(const MeRefArgs &a);
//   int encode_inter(const EncMbArgs &a);  int encode_intra(const IntraArgs &a);
//   int signal_picture(int *done, unsigned *count_out, const unsigned *count_in);   // *count_out = *count_in; fence; *done = 1
//   int entropy_picture(const EntropyPicture &p);                                    // device_entropy: the picture's slice -> host buffer
//   int side_begin();  int side_end(int slot);  int wait_side(int slot);            // fork / join of work the picture chain does not wait for
//   int prep_fork();  int prep_on(int branch);  int prep_join();                     // three concurrent branches (0 = the chain itself) and their join
//   int copy_plane(uint8_t *host_or_dev_dst, int dpitch, const uint8_t *dev_src, int spitch, int w, int h);   // device -> recon store
//   int sync();
template <class Ops>
class SegmentEncoder {
public:
    SegmentEncoder(Ops &ops, const SegmentConfig &cfg) : ops_(ops), cfg_(cfg), g_(make_geom(cfg.w, cfg.h)), ps_(plane_sizes(g_)) {}
    ~SegmentEncoder() { free_all(); }

    const EncGeom &geom() const { return g_; }
    const PlaneSizes &planes() const { return ps_; }

    // Source pictures in the padded layout (interiors filled, borders not yet extended): device arrays holding
    // n planes each, ps_.y / ps_.c bytes apart.  The resize stage writes straight into these.
    struct CurFrames { uint8_t *y, *u, *v; };
    int reserve(int n_frames) {
        if (n_frames <= cap_) return 0;
        free_all();
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        cur_.y = static_cast<uint8_t *>(ops_.alloc(ps_.y * n_frames));
        cur_.u = static_cast<uint8_t *>(ops_.alloc(ps_.c * n_frames));
        cur_.v = static_cast<uint8_t *>(ops_.alloc(ps_.c * n_frames));
        cur_low_ = static_cast<uint8_t *>(ops_.alloc(ps_.l * n_frames));
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) rec_y_[i][k] = static_cast<uint8_t *>(ops_.alloc(ps_.y));
            rec_u_[i] = static_cast<uint8_t *>(ops_.alloc(ps_.c));
            rec_v_[i] = static_cast<uint8_t *>(ops_.alloc(ps_.c));
            rec_low_[i] = static_cast<uint8_t *>(ops_.alloc(ps_.l));
        }
        d_motion_ = static_cast<MbMotion *>(ops_.alloc(sizeof(MbMotion) * nmb * n_frames));
        d_lowmv_  = static_cast<int16_t *>(ops_.alloc(sizeof(int16_t) * 2 * nmb));
        d_mail_   = static_cast<unsigned long long *>(ops_.alloc(sizeof(unsigned long long) * 8 * nmb));
        d_counters_ = static_cast<unsigned *>(ops_.alloc(sizeof(unsigned) * n_frames));
        h_done_   = static_cast<int *>(ops_.alloc_host(sizeof(int) * n_frames));
        h_counters_ = static_cast<unsigned *>(ops_.alloc_host(sizeof(unsigned) * n_frames));
        bool ok = cur_.y && cur_.u && cur_.v && cur_low_ && d_motion_ && d_lowmv_ && d_mail_ && d_counters_ && h_counters_ && h_done_;
        if (!cfg_.device_entropy) {     // records for the host entropy coder
            h_blocks_ = static_cast<int16_t *>(ops_.alloc_host(sizeof(int16_t) * kMbCoefs * nmb * n_frames));   // worst case: every block nonzero
            h_offsets_ = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * nmb * n_frames));
            h_nz_     = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * nmb * n_frames));
            h_motion_ = static_cast<MbMotion *>(ops_.alloc_host(sizeof(MbMotion) * nmb * n_frames));
            h_modes_  = static_cast<uint8_t *>(ops_.alloc_host(nmb));
            ok = ok && h_blocks_ && h_offsets_ && h_nz_ && h_motion_ && h_modes_;
        } else {                        // one picture's records stay on the device; finished slices go to the host
            // two sets of records: the entropy pass of picture f runs beside the kernels of picture f + 1
            e_blocks_  = static_cast<int16_t *>(ops_.alloc(sizeof(int16_t) * kMbCoefs * nmb * 2));
            e_offsets_ = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 2));
            e_nz_      = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 2));
            e_modes_   = static_cast<uint8_t *>(ops_.alloc(nmb));
            e_nnz_     = static_cast<uint8_t *>(ops_.alloc(24 * nmb * 2));
            e_mb_buf_  = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * kMbBufWords * nmb));
            e_mb_bits_ = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * nmb * 3));      // bits, offsets, runs
            e_mb_kind_ = static_cast<uint8_t *>(ops_.alloc(nmb));
            slice_words_ = int(std::min<size_t>(size_t(kMbBufWords) * nmb + 64, size_t(1) << 28));
            e_slice_   = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * slice_words_));
            e_info_    = static_cast<uint32_t *>(ops_.alloc(sizeof(uint32_t) * 8));
            e_used_    = static_cast<unsigned *>(ops_.alloc(sizeof(unsigned) * (n_frames + 1)));
            slices_cap_ = size_t(g_.w16) * g_.h16 * 3 * n_frames + 65536;   // twice the raw pictures (noise at QP 0 reaches 1.8x)
            h_slices_  = static_cast<uint8_t *>(ops_.alloc_host(slices_cap_));
            h_pos_     = static_cast<uint32_t *>(ops_.alloc_host(sizeof(uint32_t) * 8 * n_frames));
            ok = ok && e_blocks_ && e_offsets_ && e_nz_ && e_modes_ && e_nnz_ && e_mb_buf_ && e_mb_bits_ && e_mb_kind_ && e_slice_ && e_info_ &&
                 e_used_ && h_slices_ && h_pos_;
        }
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) ok = ok && rec_y_[i][k];
            ok = ok && rec_u_[i] && rec_v_[i] && rec_low_[i];
        }
        if (!ok) { free_all(); return GVC_ERR_NOMEM; }
        cap_ = n_frames;
        return 0;
    }
    CurFrames cur_frames() const { return cur_; }
    void set_frame_count(int n) { n_ = n; }
    // bytes that crossed PCIe towards the host for the last finished segment: packed blocks + per-macroblock tables + flags
    size_t record_bytes() const {
        if (cfg_.device_entropy) {
            size_t bytes = 0;
            for (int f = 0; f < n_; ++f) bytes += ((h_pos_[8 * f + 1] + 15u) & ~15u) + 32 + 8;
            return bytes;
        }
        size_t blocks = 0;
        for (int f = 0; f < n_; ++f) blocks += h_counters_[f];
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        return blocks * 32 + size_t(n_) * (nmb * (4 + 4) + 8) + size_t(n_ - 1) * nmb * sizeof(MbMotion) + nmb;
    }   // when the kernels of an n-picture segment are replayed from a captured graph

    // Host-side reset of the per-picture arrival flags; call before issuing (or replaying) a segment's kernels.
    void begin_segment(int n) {
        for (int f = 0; f < n; ++f) __atomic_store_n(&h_done_[f], 0, __ATOMIC_RELEASE);
    }

    // Issues every kernel of the segment (pictures 0..n-1 of cur_frames()) and, behind each picture's encode kernel, the
    // copies of its macroblock records to pinned host memory plus an arrival flag.  Nothing is waited for.
    int encode_device(int n) {
        int rc;
        const int P = kEncPad;
        const int lw = g_.w16 / 2, lh = g_.h16 / 2;
        n_ = n;
        {   // source pictures: extend the picture edges into the padding and build lowres, one launch each for the segment
            BorderBatchArgs b = border_args(cur_.y + ps_.oy, cur_.u + ps_.oc, cur_.v + ps_.oc, cfg_.w, cfg_.h);
            b.frame_stride[0] = ps_.y; b.frame_stride[1] = b.frame_stride[2] = ps_.c; b.n_frames = n;
            if ((rc = ops_.extend_border3(b))) return rc;
            LowresBatchArgs l = lowres_args(cur_.y + ps_.oy, cur_low_ + ps_.ol);
            l.sframe = ps_.y; l.dframe = ps_.l; l.n_frames = n;
            if ((rc = ops_.lowres_batch(l))) return rc;
        }
        const size_t nmb = size_t(g_.mbw) * g_.mbh;
        if ((rc = ops_.clear(d_counters_, sizeof(unsigned) * n))) return rc;
        const bool de = cfg_.device_entropy;
        // where the kernels put a picture's records: mapped host memory for the host entropy coder (one area per picture),
        // or one device-resident set that the entropy kernels consume before the next picture overwrites it
        int16_t  *blocks_dev  = de ? e_blocks_ : static_cast<int16_t *>(ops_.device_alias(h_blocks_));
        uint32_t *offsets_dev = de ? e_offsets_ : static_cast<uint32_t *>(ops_.device_alias(h_offsets_));
        uint32_t *nz_dev      = de ? e_nz_ : static_cast<uint32_t *>(ops_.device_alias(h_nz_));
        MbMotion *motion_dev  = de ? nullptr : static_cast<MbMotion *>(ops_.device_alias(h_motion_));
        uint8_t  *modes_dev   = de ? e_modes_ : static_cast<uint8_t *>(ops_.device_alias(h_modes_));
        // records of picture f: its own area of the host tables, or device set f & 1
        auto rset = [&](int f) { return size_t(de ? (f & 1) : f); };
        if (de && (rc = ops_.clear(e_used_, sizeof(unsigned)))) return rc;
        int      *done_dev    = static_cast<int *>(ops_.device_alias(h_done_));
        unsigned *counters_dev = static_cast<unsigned *>(ops_.device_alias(h_counters_));
        if (cfg_.keep_recon) recon_store_.assign(size_t(g_.w16) * g_.h16 * 3 / 2 * n, 0);
        for (int f = 0; f < n; ++f) {
            const int r = f & 1, p = r ^ 1;
            uint8_t *cy = cur_.y + ps_.y * f + ps_.oy, *cu = cur_.u + ps_.c * f + ps_.oc, *cv = cur_.v + ps_.c * f + ps_.oc;
            uint8_t *ry = rec_y_[r][0] + ps_.oy, *ru = rec_u_[r] + ps_.oc, *rv = rec_v_[r] + ps_.oc;
            uint8_t *nnz_y = de ? e_nnz_ + 24 * nmb * rset(f) : nullptr, *nnz_u = de ? nnz_y + 16 * nmb : nullptr, *nnz_v = de ? nnz_y + 20 * nmb : nullptr;
            if (de && f >= 2 && (rc = ops_.wait_side(f & 1))) return rc;      // the entropy pass of picture f - 2 has read this set
            if (f == 0) {
                IntraArgs a;
                a.cur_y = cy; a.cur_u = cu; a.cur_v = cv; a.rec_y = ry; a.rec_u = ru; a.rec_v = rv;
                a.sy = g_.sy; a.sc = g_.sc; a.mbw = g_.mbw; a.mbh = g_.mbh;
                fill_quant_params(a.q, cfg_.qp_i, true);
                a.coefs = nullptr; a.nzmask = nz_dev; a.modes = modes_dev; a.mail = d_mail_;
                a.pack.blocks = blocks_dev; a.pack.offsets = offsets_dev; a.pack.counter = d_counters_;
                a.nnz_y = nnz_y; a.nnz_u = nnz_u; a.nnz_v = nnz_v;
                if ((rc = ops_.encode_intra(a))) return rc;
            } else {
                RefPlanes ref;
                ref.y = rec_y_[p][0] + ps_.oy; ref.hh = rec_y_[p][1] + ps_.oy; ref.hv = rec_y_[p][2] + ps_.oy; ref.hc = rec_y_[p][3] + ps_.oy;
                ref.u = rec_u_[p] + ps_.oc; ref.v = rec_v_[p] + ps_.oc; ref.low = rec_low_[p] + ps_.ol;
                QuantParams q;
                fill_quant_params(q, cfg_.qp_p, false);
                // the lowres search of this picture was issued with the reference's plane preparation (below, previous
                // iteration), beside the half-sample filter and the border extension: wait for all three branches
                if ((rc = ops_.prep_join())) return rc;
                MeRefArgs mr;
                mr.cur = cy; mr.ref = ref; mr.stride = g_.sy; mr.w16 = g_.w16; mr.h16 = g_.h16; mr.mbw = g_.mbw; mr.mbh = g_.mbh;
                mr.lambda = q.lambda; mr.lowres_mv = d_lowmv_; mr.out = d_motion_ + nmb * f; mr.out_host = de ? nullptr : motion_dev + nmb * f;
                if ((rc = ops_.me_refine(mr))) return rc;
                EncMbArgs e;
                e.cur_y = cy; e.cur_u = cu; e.cur_v = cv; e.ref = ref; e.rec_y = ry; e.rec_u = ru; e.rec_v = rv;
                e.sy = g_.sy; e.sc = g_.sc; e.mbw = g_.mbw; e.mbh = g_.mbh; e.q = q;
                e.motion = d_motion_ + nmb * f; e.coefs = nullptr; e.nzmask = nz_dev + nmb * rset(f);   // levels go straight to the packed blocks
                e.pack.blocks = blocks_dev + size_t(kMbCoefs) * nmb * rset(f); e.pack.offsets = offsets_dev + nmb * rset(f); e.pack.counter = d_counters_ + f;
                e.nnz_y = nnz_y; e.nnz_u = nnz_u; e.nnz_v = nnz_v;
                if ((rc = ops_.encode_inter(e))) return rc;
            }
            // the picture's records were stored into host memory by the kernels above; a last small kernel copies the block
            // count and raises the arrival flag (stream order + a system-wide fence: everything has landed by then)
            // nothing below is waited for by the next picture's kernels: it runs beside them
            if ((rc = ops_.side_begin())) return rc;
            if (de) {     // the picture's slice data: macroblock bits, scan, placement, copy to the host buffer
                EntropyPicture ep;
                ep.ent.mbw = g_.mbw; ep.ent.mbh = g_.mbh; ep.ent.idr = f == 0;
                ep.ent.nzmask = e_nz_ + nmb * rset(f); ep.ent.motion = d_motion_ + nmb * f; ep.ent.modes = e_modes_;
                ep.ent.blocks = e_blocks_ + size_t(kMbCoefs) * nmb * rset(f); ep.ent.offsets = e_offsets_ + nmb * rset(f);
                ep.ent.nnz_y = nnz_y; ep.ent.nnz_c[0] = nnz_u; ep.ent.nnz_c[1] = nnz_v;
                ep.ent.mb_buf = e_mb_buf_; ep.ent.mb_bits = e_mb_bits_; ep.ent.mb_kind = e_mb_kind_;
                ep.slice.nmb = int(nmb); ep.slice.mb_buf = e_mb_buf_; ep.slice.mb_bits = e_mb_bits_; ep.slice.mb_kind = e_mb_kind_;
                ep.slice.mb_off = e_mb_bits_ + nmb; ep.slice.mb_run = e_mb_bits_ + 2 * nmb;
                ep.slice.slice = e_slice_; ep.slice.slice_words = slice_words_;
                ep.slice.header_bits = int(slice_header_bits(picture_header(f)));
                ep.slice.p_slice = f != 0; ep.slice.info = e_info_;
                ep.host_buf = static_cast<uint8_t *>(ops_.device_alias(h_slices_)); ep.host_cap = unsigned(std::min<size_t>(slices_cap_, 0xfffffff0u));
                ep.used = e_used_ + f; ep.pos = static_cast<uint32_t *>(ops_.device_alias(h_pos_)) + 8 * f;
                if ((rc = ops_.entropy_picture(ep))) return rc;
            }
            if ((rc = ops_.signal_picture(done_dev + f, counters_dev + f, d_counters_ + f))) return rc;
            if ((rc = ops_.side_end(f & 1))) return rc;
            if (cfg_.keep_recon) {
                uint8_t *dst = recon_store_.data() + size_t(g_.w16) * g_.h16 * 3 / 2 * f;
                if ((rc = ops_.copy_plane(dst, g_.w16, ry, g_.sy, g_.w16, g_.h16))) return rc;
                if ((rc = ops_.copy_plane(dst + size_t(g_.w16) * g_.h16, lw, ru, g_.sc, lw, lh))) return rc;
                if ((rc = ops_.copy_plane(dst + size_t(g_.w16) * g_.h16 * 5 / 4, lw, rv, g_.sc, lw, lh))) return rc;
            }
            if (f + 1 < n) {
                // The reconstruction becomes the next picture's reference.  Three independent branches (the half-sample and
                // lowres kernels replicate the picture edge themselves instead of reading the extended border): border
                // extension | lowres planes, then the next picture's lowres search | half-sample planes.
                if ((rc = ops_.prep_fork())) return rc;
                if ((rc = ops_.prep_on(1))) return rc;
                LowresBatchArgs la = lowres_args(ry, rec_low_[r] + ps_.ol);
                la.x_lo = 0; la.x_hi = g_.w16 - 1; la.y_lo = 0; la.y_hi = g_.h16 - 1;
                if ((rc = ops_.lowres_batch(la))) return rc;
                QuantParams qn;
                fill_quant_params(qn, cfg_.qp_p, false);
                MeLowArgs ml;
                ml.cur = cur_low_ + ps_.l * (f + 1) + ps_.ol; ml.ref = rec_low_[r] + ps_.ol; ml.stride = g_.sl; ml.mbw = g_.mbw; ml.mbh = g_.mbh;
                ml.r = cfg_.merange / 2; ml.lambda = qn.lambda; ml.out_mv = d_lowmv_;
                if ((rc = ops_.me_lowres(ml))) return rc;
                if ((rc = ops_.prep_on(2))) return rc;
                HpelArgs ha;
                ha.src = ry; ha.dh = rec_y_[r][1] + ps_.oy; ha.dv = rec_y_[r][2] + ps_.oy; ha.dc = rec_y_[r][3] + ps_.oy;
                ha.stride = g_.sy; ha.x_lo = -P; ha.x_hi = g_.w16 + P; ha.y_lo = -P; ha.y_hi = g_.h16 + P;
                ha.cx_lo = 0; ha.cx_hi = g_.w16; ha.cy_lo = 0; ha.cy_hi = g_.h16;
                if ((rc = ops_.hpel(ha))) return rc;
                if ((rc = ops_.prep_on(0))) return rc;
                BorderBatchArgs b = border_args(ry, ru, rv, g_.w16, g_.h16);
                if ((rc = ops_.extend_border3(b))) return rc;
            }
        }
        // the chain ends where the side work ends (one stream to wait on; a captured graph must rejoin its origin)
        if ((rc = ops_.wait_side(0))) return rc;
        if (n >= 2 && (rc = ops_.wait_side(1))) return rc;
        return 0;
    }

    // Entropy-codes the pictures as their records arrive (worker threads wait on the arrival flags while the GPU is still
    // working on later pictures), then waits for the stream.
    int finish(SegmentOutput &out) {
        failed_ = false;
        int rc;
        if (cfg_.device_entropy) {      // the slices arrive finished: wait once (no polling thread), then wrap them
            if ((rc = ops_.sync())) return rc;
            write_stream(out);
        } else {                        // host entropy coding overlaps the GPU: workers start on a picture when its flag is up
            write_stream(out);
            if ((rc = ops_.sync())) return rc;
        }
        if (failed_) return GVC_ERR_TOO_LARGE;     // the slices outgrew the buffers sized from the raw pictures
        if (cfg_.keep_recon) out.recon.swap(recon_store_);
        return 0;
    }

    void write_stream(SegmentOutput &out) {
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
            while (__atomic_load_n(&h_done_[f], __ATOMIC_ACQUIRE) == 0) std::this_thread::yield();
            PictureData pic = picture_header(f);
            if (cfg_.device_entropy) {      // the GPU wrote slice_data() behind room for the header: fill the header in, wrap the NAL
                const uint32_t *pos = h_pos_ + 8 * f;
                if (pos[7]) { failed_ = true; return; }
                std::vector<uint8_t> rbsp(h_slices_ + pos[0], h_slices_ + pos[0] + pos[1]);
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
            pic.coefs = h_blocks_ + size_t(kMbCoefs) * nmb * f;
            pic.compact = true;
            pic.block_offsets = h_offsets_ + nmb * f;
            pic.nzmask = h_nz_ + nmb * f;
            pic.motion = h_motion_ + nmb * f;
            pic.modes = h_modes_;
            BitWriter bw;
            static const bool skip_entropy = getenv("GVC_SKIP_ENTROPY") != nullptr;   // measurement aid: stream is not decodable
            if (!skip_entropy) write_slice(bw, pic, &out.stats[f]);
            else bw.rbsp_trailing();
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
        ++idr_count_;
    }
    bool failed() const { return failed_; }

private:
    // the fields of picture f's slice header.  idr_pic_id runs through 255..510: consecutive segments differ and ue(v) always
    // takes 17 bits, so the header length is a constant of the picture type (the GPU entropy pass is replayed from a graph).
    PictureData picture_header(int f) const {
        PictureData pic;
        pic.mbw = g_.mbw; pic.mbh = g_.mbh; pic.idr = f == 0; pic.frame_num = f; pic.idr_pic_id = 255 + (idr_count_ & 255);
        pic.qp = f == 0 ? cfg_.qp_i : cfg_.qp_p;
        return pic;
    }
    // Y/U/V planes at sample (0,0) of one picture whose valid area is w x h luma samples
    BorderBatchArgs border_args(uint8_t *y, uint8_t *u, uint8_t *v, int w, int h) const {
        const int P = kEncPad, Pc = kEncPad / 2, lw = g_.w16 / 2, lh = g_.h16 / 2;
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
        const int P = kEncPad, Pc = kEncPad / 2, lw = g_.w16 / 2, lh = g_.h16 / 2;
        LowresBatchArgs l;
        l.src = y; l.dst = low; l.sstride = g_.sy; l.dstride = g_.sl;
        l.x_lo = -P; l.x_hi = g_.w16 + P - 1; l.y_lo = -P; l.y_hi = g_.h16 + P - 1;
        l.lx_lo = -Pc; l.lx_hi = lw + Pc; l.ly_lo = -Pc; l.ly_hi = lh + Pc;
        l.sframe = l.dframe = 0; l.n_frames = 1;
        return l;
    }
    void free_all() {
        auto rel = [&](auto *&p) { if (p) ops_.release(p); p = nullptr; };
        rel(cur_.y); rel(cur_.u); rel(cur_.v); rel(cur_low_);
        for (int i = 0; i < 2; ++i) {
            for (int k = 0; k < 4; ++k) rel(rec_y_[i][k]);
            rel(rec_u_[i]); rel(rec_v_[i]); rel(rec_low_[i]);
        }
        rel(d_motion_); rel(d_lowmv_); rel(d_mail_);
        auto relh = [&](auto *&p) { if (p) ops_.release_host(p); p = nullptr; };
        rel(d_counters_); relh(h_offsets_); relh(h_counters_);
        relh(h_blocks_); relh(h_nz_); relh(h_motion_); relh(h_modes_); relh(h_done_);
        rel(e_blocks_); rel(e_offsets_); rel(e_nz_); rel(e_modes_); rel(e_nnz_); rel(e_mb_buf_); rel(e_mb_bits_); rel(e_mb_kind_);
        rel(e_slice_); rel(e_info_); rel(e_used_); relh(h_slices_); relh(h_pos_);
        cap_ = 0;
    }

    Ops          &ops_;
    SegmentConfig cfg_;
    EncGeom       g_;
    PlaneSizes    ps_;
    int           cap_ = 0, n_ = 0, idr_count_ = 0;
    bool          failed_ = false;
    CurFrames     cur_{nullptr, nullptr, nullptr};
    uint8_t      *cur_low_ = nullptr;
    uint8_t      *rec_y_[2][4] = {{nullptr, nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr, nullptr}};
    uint8_t      *rec_u_[2] = {nullptr, nullptr}, *rec_v_[2] = {nullptr, nullptr}, *rec_low_[2] = {nullptr, nullptr};
    MbMotion     *d_motion_ = nullptr;
    int16_t      *d_lowmv_ = nullptr;
    unsigned long long *d_mail_ = nullptr;   // intra wavefront mailboxes
    // Records for the host entropy coder live in pinned memory mapped into the device address space; the kernels store
    // into it directly (no copy engine on the pictures' critical path).
    int16_t      *h_blocks_ = nullptr;    // the nonzero blocks of every picture, packed
    uint32_t     *h_offsets_ = nullptr;   // per MB: first packed block
    unsigned     *d_counters_ = nullptr, *h_counters_ = nullptr;   // per picture: blocks handed out
    uint32_t     *h_nz_ = nullptr;
    MbMotion     *h_motion_ = nullptr;
    uint8_t      *h_modes_ = nullptr;
    int          *h_done_ = nullptr;      // per picture: set to 1 by the kernel that closes the picture
    // device_entropy: one picture's records and the entropy pass's scratch on the device, finished slices on the host
    int16_t      *e_blocks_ = nullptr;
    uint32_t     *e_offsets_ = nullptr, *e_nz_ = nullptr, *e_mb_buf_ = nullptr, *e_mb_bits_ = nullptr, *e_slice_ = nullptr, *e_info_ = nullptr;
    uint8_t      *e_modes_ = nullptr, *e_nnz_ = nullptr, *e_mb_kind_ = nullptr;
    unsigned     *e_used_ = nullptr;
    int           slice_words_ = 0;
    size_t        slices_cap_ = 0;
    uint8_t      *h_slices_ = nullptr;    // the segment's slices, back to back (each padded to 16 bytes)
    uint32_t     *h_pos_ = nullptr;       // per picture: base, bytes, bits, skipped, p16, p8, intra, error
    std::vector<uint8_t> recon_store_;
};

}  // namespace gvc
