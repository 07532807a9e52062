// h264_stream.cpp -- see h264_stream.h.  Tables are those of ITU-T H.264 (Table 9-4 coded_block_pattern
// mapping, Table 9-5 coeff_token, Tables 9-7/9-8/9-9 total_zeros, Table 9-10 run_before), stored as
// (length, bits) pairs; tests/test_emu_entropy.py checks that the GPU writer's phases give the same bytes, and
// tests/test_tier3_decode.py that a real decoder (FFmpeg's h264) reproduces the kernels' reconstruction from the
// streams written here (noise at QP 0..30 exercises every table).
#include "h264_stream.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace gvc {

// ------------------------------------------------------------------------------- bit writer ----
void BitWriter::rbsp_trailing() {
    put(1, 1);
    if (n_ & 7) put(0, 8 - (n_ & 7));
    out_.assign(buf_.begin(), buf_.begin() + len_);
    for (int k = n_ - 8; k >= 0; k -= 8) out_.push_back(uint8_t(acc_ >> k));
}

void append_nal(std::vector<uint8_t> &out, int nal_ref_idc, int nal_unit_type, const std::vector<uint8_t> &rbsp, bool annexb) {
    std::vector<uint8_t> nal;
    nal.reserve(rbsp.size() + rbsp.size() / 64 + 8);
    nal.push_back(uint8_t((nal_ref_idc << 5) | nal_unit_type));
    // emulation prevention (7.4.1): a byte <= 3 after two zero bytes gets a 0x03 in front.  Zero bytes are rare in entropy-coded
    // data, so the stretches between them are appended whole and only the bytes after a zero go through the state machine.
    const uint8_t *p = rbsp.data();
    const size_t   n = rbsp.size();
    size_t i = 0;
    int    zeros = 0;
    while (i < n) {
        if (zeros == 0) {
            const void *z = memchr(p + i, 0, n - i);
            const size_t j = z ? size_t(static_cast<const uint8_t *>(z) - p) : n;
            nal.insert(nal.end(), p + i, p + j);
            i = j;
            if (i == n) break;
        }
        const uint8_t b = p[i++];
        if (zeros >= 2 && b <= 3) { nal.push_back(3); zeros = 0; }   // emulation_prevention_three_byte
        nal.push_back(b);
        zeros = b == 0 ? zeros + 1 : 0;
    }
    if (annexb) {
        static const uint8_t sc[4] = {0, 0, 0, 1};
        out.insert(out.end(), sc, sc + 4);
    } else {
        const uint32_t n = uint32_t(nal.size());
        const uint8_t  len[4] = {uint8_t(n >> 24), uint8_t(n >> 16), uint8_t(n >> 8), uint8_t(n)};
        out.insert(out.end(), len, len + 4);
    }
    out.insert(out.end(), nal.begin(), nal.end());
}

// ------------------------------------------------------------------------------- parameter sets -
int pick_level(int mbw, int mbh, int fps_num, int fps_den) {
    struct L { int idc, max_fs; long max_mbps; };
    static const L levels[] = {{30, 1620, 40500},   {31, 3600, 108000},  {32, 5120, 216000},  {40, 8192, 245760},
                               {42, 8704, 522240},  {50, 22080, 589824}, {51, 36864, 983040}, {52, 36864, 2073600}};
    const long fs = long(mbw) * mbh, mbps = fs * fps_num / std::max(1, fps_den);
    for (const L &l : levels)
        if (fs <= l.max_fs && mbps <= l.max_mbps) return l.idc;
    return 52;
}

void write_sps(BitWriter &bw, const StreamParams &p) {
    bw.put(66, 8);            // profile_idc: Baseline
    bw.put(1, 1);             // constraint_set0_flag
    bw.put(1, 1);             // constraint_set1_flag (Constrained Baseline: also decodable as Main)
    bw.put(0, 6);             // constraint_set2..5, reserved_zero_2bits
    bw.put(uint32_t(p.level_idc), 8);
    bw.ue(0);                 // seq_parameter_set_id
    bw.ue(4);                 // log2_max_frame_num_minus4 -> frame_num is 8 bits
    bw.ue(2);                 // pic_order_cnt_type 2: output order = decoding order (no B pictures)
    bw.ue(1);                 // max_num_ref_frames
    bw.put(0, 1);             // gaps_in_frame_num_value_allowed_flag
    bw.ue(uint32_t(p.w16 / 16 - 1));
    bw.ue(uint32_t(p.h16 / 16 - 1));
    bw.put(1, 1);             // frame_mbs_only_flag
    bw.put(1, 1);             // direct_8x8_inference_flag
    const int crop_r = (p.w16 - p.w) / 2, crop_b = (p.h16 - p.h) / 2;
    if (crop_r || crop_b) {
        bw.put(1, 1);
        bw.ue(0); bw.ue(uint32_t(crop_r)); bw.ue(0); bw.ue(uint32_t(crop_b));
    } else {
        bw.put(0, 1);
    }
    bw.put(1, 1);             // vui_parameters_present_flag
    bw.put(0, 1);             //   aspect_ratio_info_present_flag
    bw.put(0, 1);             //   overscan_info_present_flag
    bw.put(0, 1);             //   video_signal_type_present_flag
    bw.put(0, 1);             //   chroma_loc_info_present_flag
    bw.put(1, 1);             //   timing_info_present_flag
    bw.put(uint32_t(p.fps_den), 32);       // num_units_in_tick
    bw.put(uint32_t(2 * p.fps_num), 32);   // time_scale (field rate)
    bw.put(1, 1);             //   fixed_frame_rate_flag
    bw.put(0, 1);             //   nal_hrd_parameters_present_flag
    bw.put(0, 1);             //   vcl_hrd_parameters_present_flag
    bw.put(0, 1);             //   pic_struct_present_flag
    bw.put(1, 1);             //   bitstream_restriction_flag
    bw.put(1, 1);             //     motion_vectors_over_pic_boundaries_flag
    bw.ue(0);                 //     max_bytes_per_pic_denom
    bw.ue(0);                 //     max_bits_per_mb_denom
    bw.ue(10);                //     log2_max_mv_length_horizontal
    bw.ue(10);                //     log2_max_mv_length_vertical
    bw.ue(0);                 //     max_num_reorder_frames
    bw.ue(1);                 //     max_dec_frame_buffering
    bw.rbsp_trailing();
}

void write_pps(BitWriter &bw) {
    bw.ue(0);      // pic_parameter_set_id
    bw.ue(0);      // seq_parameter_set_id
    bw.put(0, 1);  // entropy_coding_mode_flag: CAVLC
    bw.put(0, 1);  // bottom_field_pic_order_in_frame_present_flag
    bw.ue(0);      // num_slice_groups_minus1
    bw.ue(0);      // num_ref_idx_l0_default_active_minus1
    bw.ue(0);      // num_ref_idx_l1_default_active_minus1
    bw.put(0, 1);  // weighted_pred_flag
    bw.put(0, 2);  // weighted_bipred_idc
    bw.se(0);      // pic_init_qp_minus26
    bw.se(0);      // pic_init_qs_minus26
    bw.se(0);      // chroma_qp_index_offset
    bw.put(1, 1);  // deblocking_filter_control_present_flag
    bw.put(0, 1);  // constrained_intra_pred_flag
    bw.put(0, 1);  // redundant_pic_cnt_present_flag
    bw.rbsp_trailing();
}

// ------------------------------------------------------------------------------- CAVLC tables --
namespace {

#define GVC_TAB const
#include "cavlc_tables.h"
#undef GVC_TAB

struct CbpCode {
    uint8_t inter[48], intra[48];
    CbpCode() {
        for (int c = 0; c < 48; ++c) { inter[kCbpInterFromCode[c]] = uint8_t(c); intra[kCbpIntraFromCode[c]] = uint8_t(c); }
    }
};
const CbpCode kCbpCode;

// blkIdx (coding order, 8x8-wise Z scan) -> raster 4x4 block inside the MB
const uint8_t kBlkToRaster[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15};

// residual_block_cavlc(): levels in scan order l[0..n), nC as derived in 9.2.1.  Returns TotalCoeff.
int write_residual(BitWriter &bw, const int16_t *l, int n, int nC) {
    int level[16], run[16], total = 0, last = -1;
    for (int i = n - 1; i >= 0; --i)
        if (l[i]) { last = i; break; }
    if (last >= 0) {
        int zeros = 0;
        for (int i = last; i >= 0; --i) {
            if (l[i]) {
                if (total) run[total - 1] = zeros;
                level[total++] = l[i];
                zeros = 0;
            } else {
                ++zeros;
            }
        }
        run[total - 1] = zeros;   // zeros below the lowest coefficient (not coded as run_before)
    }
    int t1 = 0;
    while (t1 < total && t1 < 3 && std::abs(level[t1]) == 1) ++t1;
    // coeff_token
    if (nC == -1) {
        bw.put(kChromaDcTokenBits[4 * total + t1], kChromaDcTokenLen[4 * total + t1]);
    } else {
        const int tab = nC < 2 ? 0 : nC < 4 ? 1 : nC < 8 ? 2 : 3;
        bw.put(kCoeffTokenBits[tab][4 * total + t1], kCoeffTokenLen[tab][4 * total + t1]);
    }
    if (!total) return 0;
    for (int k = 0; k < t1; ++k) bw.put(level[k] < 0, 1);   // trailing_ones_sign_flag
    int suffix_len = (total > 10 && t1 < 3) ? 1 : 0;
    for (int k = t1; k < total; ++k) {
        const int lv = level[k];
        int code = lv > 0 ? 2 * lv - 2 : -2 * lv - 1;       // levelCode
        if (k == t1 && t1 < 3) code -= 2;
        if (suffix_len == 0) {
            if (code < 14) {
                bw.put(1, code + 1);
            } else if (code < 30) {
                bw.put(1, 15);
                bw.put(uint32_t(code - 14), 4);
            } else {
                int prefix = 15, base = 30;                  // level_prefix >= 15: suffix of (prefix - 3) bits
                while (code - base >= (1 << (prefix - 3))) { base += 1 << (prefix - 3); ++prefix; }
                bw.put(0, prefix > 31 ? 31 : prefix);        // (prefix never approaches 31 for 16-bit levels)
                bw.put(1, 1);
                bw.put(uint32_t(code - base), prefix - 3);
            }
        } else {
            if (code < (15 << suffix_len)) {
                bw.put(1, (code >> suffix_len) + 1);
                bw.put(uint32_t(code & ((1 << suffix_len) - 1)), suffix_len);
            } else {
                int prefix = 15, base = 15 << suffix_len;
                while (code - base >= (1 << (prefix - 3))) { base += 1 << (prefix - 3); ++prefix; }
                bw.put(0, prefix);
                bw.put(1, 1);
                bw.put(uint32_t(code - base), prefix - 3);
            }
        }
        if (suffix_len == 0) suffix_len = 1;
        if (std::abs(lv) > (3 << (suffix_len - 1)) && suffix_len < 6) ++suffix_len;
    }
    int total_zeros = last + 1 - total;
    if (total < n) {
        if (nC == -1) bw.put(kChromaDcTotalZerosBits[total - 1][total_zeros], kChromaDcTotalZerosLen[total - 1][total_zeros]);
        else          bw.put(kTotalZerosBits[total - 1][total_zeros], kTotalZerosLen[total - 1][total_zeros]);
    }
    int zeros_left = total_zeros;
    for (int k = 0; k < total - 1 && zeros_left > 0; ++k) {
        const int t = std::min(zeros_left, 7) - 1;
        bw.put(kRunBits[t][run[k]], kRunLen[t][run[k]]);
        zeros_left -= run[k];
    }
    return total;
}

struct PicState {
    int mbw, mbh, w4, h4;
    std::vector<uint8_t> nnz_y;        // TotalCoeff per luma 4x4 block, [h4][w4]
    std::vector<uint8_t> nnz_c[2];     // per chroma 4x4 block, [2*mbh][2*mbw]
    std::vector<int16_t> mv;           // per luma 4x4 block (x,y)
    std::vector<uint8_t> done;         // block already decoded (availability for MV prediction)
};

int luma_nc(const PicState &s, int bx, int by) {
    const bool a = bx > 0, b = by > 0;
    const int  na = a ? s.nnz_y[by * s.w4 + bx - 1] : 0, nb = b ? s.nnz_y[(by - 1) * s.w4 + bx] : 0;
    return a && b ? (na + nb + 1) >> 1 : a ? na : b ? nb : 0;
}
int chroma_nc(const PicState &s, int p, int bx, int by) {
    const int  w = 2 * s.mbw;
    const bool a = bx > 0, b = by > 0;
    const int  na = a ? s.nnz_c[p][by * w + bx - 1] : 0, nb = b ? s.nnz_c[p][(by - 1) * w + bx] : 0;
    return a && b ? (na + nb + 1) >> 1 : a ? na : b ? nb : 0;
}

struct Mv { int x, y; };
int median3(int a, int b, int c) { return std::max(std::min(a, b), std::min(std::max(a, b), c)); }

// 8.4.1.3 for a partition whose top-left 4x4 block is (bx,by) and which is `w` blocks wide; single reference.
Mv predict_mv(const PicState &s, int bx, int by, int w, bool *avail_a = nullptr, bool *avail_b = nullptr, Mv *mva = nullptr, Mv *mvb = nullptr) {
    auto ok = [&](int x, int y) { return x >= 0 && y >= 0 && x < s.w4 && y < s.h4 && s.done[y * s.w4 + x]; };
    auto at = [&](int x, int y) { return Mv{s.mv[2 * (y * s.w4 + x)], s.mv[2 * (y * s.w4 + x) + 1]}; };
    const bool A = ok(bx - 1, by), B = ok(bx, by - 1);
    bool       C = ok(bx + w, by - 1);
    Mv a = A ? at(bx - 1, by) : Mv{0, 0}, b = B ? at(bx, by - 1) : Mv{0, 0}, c{0, 0};
    if (C) c = at(bx + w, by - 1);
    else if (ok(bx - 1, by - 1)) { C = true; c = at(bx - 1, by - 1); }
    if (avail_a) *avail_a = A;
    if (avail_b) *avail_b = B;
    if (mva) *mva = a;
    if (mvb) *mvb = b;
    if (!B && !C && A) return a;
    const int n = int(A) + int(B) + int(C);
    if (n == 1) return A ? a : B ? b : c;
    return Mv{median3(a.x, b.x, c.x), median3(a.y, b.y, c.y)};
}

void set_mv(PicState &s, int bx, int by, int w, int h, Mv v) {
    for (int y = by; y < by + h; ++y)
        for (int x = bx; x < bx + w; ++x) {
            s.mv[2 * (y * s.w4 + x)] = int16_t(v.x);
            s.mv[2 * (y * s.w4 + x) + 1] = int16_t(v.y);
            s.done[y * s.w4 + x] = 1;
        }
}

inline void write_empty_block(BitWriter &bw, int nC) {     // coeff_token with TotalCoeff = 0
    if (nC == -1) { bw.put(kChromaDcTokenBits[0], kChromaDcTokenLen[0]); return; }
    const int tab = nC < 2 ? 0 : nC < 4 ? 1 : nC < 8 ? 2 : 3;
    bw.put(kCoeffTokenBits[tab][0], kCoeffTokenLen[tab][0]);
}

// residual( ) of one macroblock; updates the nnz maps.  i16: Intra16x16 (DC block + 15-coefficient AC blocks).
// nz is the kernels' per-block "has a nonzero level" mask: empty blocks are coded without touching their coefficients.
void write_mb_residual(BitWriter &bw, PicState &s, const int16_t *mbco, bool compact, int mx, int my, int cbp, bool i16, uint32_t nz) {
    // block b of this macroblock's record: fixed slot in the full layout, rank among the present blocks when packed
    const unsigned cm = (nz & 0x00ffffffu) | ((nz & 0x03000000u) ? (1u << 24) : 0u) | (((nz >> 26) & 1u) << 25);
    auto block_at = [&](int b) { return compact ? mbco + 16 * __builtin_popcount(cm & ((1u << b) - 1u)) : mbco + 16 * b; };
    if (i16) {
        if (nz & (1u << 26)) write_residual(bw, block_at(25), 16, luma_nc(s, 4 * mx, 4 * my));
        else write_empty_block(bw, luma_nc(s, 4 * mx, 4 * my));
    }
    for (int blk = 0; blk < 16; ++blk) {
        const int r = kBlkToRaster[blk], bx = 4 * mx + (r & 3), by = 4 * my + (r >> 2);
        int       total = 0;
        if (cbp & (1 << (blk >> 2))) {
            if (!(nz & (1u << r))) write_empty_block(bw, luma_nc(s, bx, by));
            else if (i16) total = write_residual(bw, block_at(r) + 1, 15, luma_nc(s, bx, by));
            else          total = write_residual(bw, block_at(r), 16, luma_nc(s, bx, by));
        }
        s.nnz_y[by * s.w4 + bx] = uint8_t(total);
    }
    if (cbp & 0x30)
        for (int p = 0; p < 2; ++p) {
            if (nz & (1u << (24 + p))) write_residual(bw, block_at(24) + 4 * p, 4, -1);
            else write_empty_block(bw, -1);
        }
    for (int p = 0; p < 2; ++p)
        for (int b = 0; b < 4; ++b) {
            const int bx = 2 * mx + (b & 1), by = 2 * my + (b >> 1);
            int       total = 0;
            if (cbp & 0x20) {
                if (nz & (1u << (16 + 4 * p + b))) total = write_residual(bw, block_at(16 + 4 * p + b) + 1, 15, chroma_nc(s, p, bx, by));
                else write_empty_block(bw, chroma_nc(s, p, bx, by));
            }
            s.nnz_c[p][by * 2 * s.mbw + bx] = uint8_t(total);
        }
}

int mb_cbp(uint32_t nz, bool i16) {
    int luma = 0;
    for (int k = 0; k < 4; ++k) {
        const int x = (k & 1) * 2, y = (k >> 1) * 2;
        const uint32_t m = (1u << (y * 4 + x)) | (1u << (y * 4 + x + 1)) | (1u << ((y + 1) * 4 + x)) | (1u << ((y + 1) * 4 + x + 1));
        if (nz & m) luma |= 1 << k;
    }
    if (i16 && luma) luma = 15;
    const int chroma = (nz & 0x00ff0000u) ? 2 : (nz & 0x03000000u) ? 1 : 0;
    return luma | (chroma << 4);
}

void clear_mb_nnz(PicState &s, int mx, int my) {
    for (int y = 0; y < 4; ++y)
        for (int x = 0; x < 4; ++x) s.nnz_y[(4 * my + y) * s.w4 + 4 * mx + x] = 0;
    for (int p = 0; p < 2; ++p)
        for (int b = 0; b < 4; ++b) s.nnz_c[p][(2 * my + (b >> 1)) * 2 * s.mbw + 2 * mx + (b & 1)] = 0;
}

}  // namespace

// ------------------------------------------------------------------------------- slice ---------
void write_slice_header(BitWriter &bw, const PictureData &pic) {
    bw.ue(0);                              // first_mb_in_slice
    bw.ue(pic.idr ? 7 : 5);                // slice_type: all slices of the picture are I / P
    bw.ue(0);                              // pic_parameter_set_id
    bw.put(uint32_t(pic.frame_num & 255), 8);
    if (pic.idr) bw.ue(uint32_t(pic.idr_pic_id));
    if (!pic.idr) {
        bw.put(0, 1);                      // num_ref_idx_active_override_flag
        bw.put(0, 1);                      // ref_pic_list_modification_flag_l0
        bw.put(0, 1);                      // adaptive_ref_pic_marking_mode_flag (sliding window)
    } else {
        bw.put(0, 1);                      // no_output_of_prior_pics_flag
        bw.put(0, 1);                      // long_term_reference_flag
    }
    bw.se(pic.qp - 26);                    // slice_qp_delta
    bw.ue(1);                              // disable_deblocking_filter_idc = 1: no in-loop filter
}
size_t slice_header_bits(const PictureData &pic) {
    BitWriter bw;
    write_slice_header(bw, pic);
    return bw.bit_count();
}
std::vector<uint8_t> slice_header_bytes(const PictureData &pic) {
    BitWriter bw;
    write_slice_header(bw, pic);
    return bw.peek();
}

void write_slice(BitWriter &bw, const PictureData &pic, SliceStats *stats) {
    PicState s;
    s.mbw = pic.mbw; s.mbh = pic.mbh; s.w4 = 4 * pic.mbw; s.h4 = 4 * pic.mbh;
    s.nnz_y.assign(size_t(s.w4) * s.h4, 0);
    s.nnz_c[0].assign(size_t(4) * pic.mbw * pic.mbh, 0);
    s.nnz_c[1].assign(size_t(4) * pic.mbw * pic.mbh, 0);
    s.mv.assign(size_t(2) * s.w4 * s.h4, 0);
    s.done.assign(size_t(s.w4) * s.h4, 0);
    SliceStats st;

    write_slice_header(bw, pic);

    // slice_data()
    int skip_run = 0;
    for (int my = 0; my < pic.mbh; ++my)
        for (int mx = 0; mx < pic.mbw; ++mx) {
            const int       mb = my * pic.mbw + mx;
            const uint32_t  nz = pic.nzmask[mb];
            const int16_t  *co = pic.compact ? pic.coefs + size_t(16) * pic.block_offsets[mb] : pic.coefs + size_t(mb) * kMbCoefs;
            if (pic.idr) {
                const int cbp = mb_cbp(nz, true);
                bw.ue(uint32_t(1 + pic.modes[mb] + 4 * (cbp >> 4) + ((cbp & 15) ? 12 : 0)));   // mb_type I_16x16_<mode>_<cbpc>_<cbpl>
                bw.ue(0);                  // intra_chroma_pred_mode: DC
                bw.se(0);                  // mb_qp_delta
                write_mb_residual(bw, s, co, pic.compact, mx, my, cbp, true, nz);
                ++st.intra;
                continue;
            }
            const MbMotion &mo  = pic.motion[mb];
            const int       cbp = mb_cbp(nz, false);
            if (mo.mode == 0) {
                const Mv mv{mo.mv[0][0], mo.mv[0][1]};
                bool A, B;
                Mv   a, b;
                const Mv mvp = predict_mv(s, 4 * mx, 4 * my, 4, &A, &B, &a, &b);
                Mv skip = mvp;
                if (!A || !B || (a.x == 0 && a.y == 0) || (b.x == 0 && b.y == 0)) skip = Mv{0, 0};
                if (cbp == 0 && mv.x == skip.x && mv.y == skip.y) {      // P_Skip
                    ++skip_run;
                    ++st.skipped;
                    set_mv(s, 4 * mx, 4 * my, 4, 4, mv);
                    clear_mb_nnz(s, mx, my);
                    continue;
                }
                bw.ue(uint32_t(skip_run));
                skip_run = 0;
                bw.ue(0);                  // mb_type P_L0_16x16
                bw.se(mv.x - mvp.x);
                bw.se(mv.y - mvp.y);
                set_mv(s, 4 * mx, 4 * my, 4, 4, mv);
                ++st.p16;
            } else {
                bw.ue(uint32_t(skip_run));
                skip_run = 0;
                bw.ue(3);                  // mb_type P_8x8
                for (int k = 0; k < 4; ++k) bw.ue(0);   // sub_mb_type P_L0_8x8
                for (int k = 0; k < 4; ++k) {
                    const int bx = 4 * mx + 2 * (k & 1), by = 4 * my + 2 * (k >> 1);
                    const Mv  mv{mo.mv[k][0], mo.mv[k][1]};
                    const Mv  mvp = predict_mv(s, bx, by, 2);
                    bw.se(mv.x - mvp.x);
                    bw.se(mv.y - mvp.y);
                    set_mv(s, bx, by, 2, 2, mv);
                }
                ++st.p8;
            }
            bw.ue(kCbpCode.inter[cbp]);    // coded_block_pattern me(v)
            if (cbp) {
                bw.se(0);                  // mb_qp_delta
                write_mb_residual(bw, s, co, pic.compact, mx, my, cbp, false, nz);
            } else {
                clear_mb_nnz(s, mx, my);
            }
        }
    if (skip_run) bw.ue(uint32_t(skip_run));
    bw.rbsp_trailing();
    st.bits = bw.bit_count();
    if (stats) *stats = st;
}

}  // namespace gvc
