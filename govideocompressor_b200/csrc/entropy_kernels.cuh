// entropy_kernels.cuh -- K8: CAVLC slice data written on the GPU (sm_100a), so that what crosses PCIe per picture is the
// finished slice instead of its coefficient records, and the host cores do nothing per macroblock.
//
// This is SURVEY.md 8(f) row 3 ("entropy coding offload"): in the reference the bits are produced by libx264 inside the
// spawned ffmpeg (client.go:115).  Syntax and tables: ITU-T H.264 7.3.4 / 7.3.5 (slice data, macroblock layer), 8.4.1.3
// (motion vector prediction), 9.1 (Exp-Golomb), 9.2 (CAVLC); the stream is bit-identical to the host writer in
// h264_stream.cpp (same tables: cavlc_tables.h), which the tests use as the checker.
//
// A macroblock's bits depend on its neighbours only through data that is final before entropy coding starts (their
// motion vectors and per-block coefficient counts), never on bit positions, so:
//   ent_mb     one warp per macroblock, one lane per residual block: its bits (without mb_skip_run), its length and kind;
//   ent_scan   one CTA: skip runs and bit offsets of all macroblocks by two block-wide scans, slice header, trailing bits;
//   ent_place  one thread per coded macroblock: ue(mb_skip_run) + its bits OR-ed into the slice at its offset;
//   ent_copy   slice bytes to mapped host memory (byte order fixed on the way), appended to the segment's buffer.
// Bodies are per-thread functions so the CPU tests run them serially (GVC_EMU).
#pragma once
#include "dev_common.h"
#include "enc_types.h"
#include "enc_kernels.cuh"   // compact_mask, popcount32

namespace gvc {

namespace etab {
#if defined(GVC_EMU)
#define GVC_TAB static const
#else
#define GVC_TAB __device__ const
#endif
#include "cavlc_tables.h"
#undef GVC_TAB
}  // namespace etab

// Upper bound of one macroblock's bits: 27 residual blocks x (coeff_token 16 + 16 levels x 36 + total_zeros 9 + runs 45)
// + header < 18,944 bits.  The writers still check and flag an overflow instead of writing past the end.
constexpr int kMbBufWords = 592;

struct EntropyArgs {
    int mbw, mbh, idr;
    const uint32_t *nzmask;      // per MB
    const MbMotion *motion;      // per MB (P pictures)
    const uint8_t  *modes;       // per MB (IDR pictures)
    const int16_t  *blocks;      // packed nonzero blocks of the picture, 16 levels each
    const uint32_t *offsets;     // per MB: index of its first packed block
    const uint8_t  *nnz_y;       // [4*mbh][4*mbw] nonzero levels per luma 4x4 block (AC only for Intra16x16)
    const uint8_t  *nnz_c[2];    // [2*mbh][2*mbw] per chroma 4x4 block (AC)
    uint32_t       *mb_buf;      // [nmb][kMbBufWords]
    uint32_t       *mb_bits;     // [nmb] bits in mb_buf (0 for P_Skip)
    uint8_t        *mb_kind;     // [nmb] 0 P_Skip, 1 P_L0_16x16, 2 P_8x8, 3 I_16x16, +128: buffer overflow
};

// ---- bit writer into 32-bit words, most significant bit first ----------------------------------------
struct BitOut { uint32_t *w; unsigned long long acc; int n, len, cap; bool overflow; };
GVC_HD void bo_init(BitOut &b, uint32_t *w, int cap) { b.w = w; b.acc = 0; b.n = 0; b.len = 0; b.cap = cap; b.overflow = false; }
GVC_HD void bo_put(BitOut &b, uint32_t v, int bits) {          // 0 <= bits <= 32
    if (bits < 32) v &= (1u << bits) - 1u;
    b.acc = (b.acc << bits) | v;
    b.n += bits;
    if (b.n >= 32) {
        if (b.len < b.cap) b.w[b.len++] = (uint32_t)(b.acc >> (b.n - 32));
        else b.overflow = true;
        b.n -= 32;
    }
}
GVC_HD int bo_finish(BitOut &b) {                              // pending bits left-aligned in a last word; returns the bit count
    const int bits = 32 * b.len + b.n;
    if (b.n) {
        if (b.len < b.cap) b.w[b.len] = (uint32_t)(b.acc << (32 - b.n));
        else b.overflow = true;
    }
    return bits;
}
GVC_HD int ent_clz(unsigned v) {
#if defined(GVC_EMU)
    return v ? __builtin_clz(v) : 32;
#else
    return __clz((int)v);
#endif
}
GVC_HD int ue_len(unsigned v) { return 2 * (31 - ent_clz(v + 1)) + 1; }
GVC_HD void bo_ue(BitOut &b, unsigned v) {
    const int len = 31 - ent_clz(v + 1);
    bo_put(b, 0, len);
    bo_put(b, v + 1, len + 1);
}
GVC_HD void bo_se(BitOut &b, int v) { bo_ue(b, v > 0 ? (unsigned)(2 * v - 1) : (unsigned)(-2 * v)); }

// ---- residual_block_cavlc(): levels l[0..n) in scan order, nC per 9.2.1; returns TotalCoeff ----------
GVC_HD int ent_residual(BitOut &bw, const int16_t *l, int n, int nC) {
    unsigned nzb = 0;
    for (int i = 0; i < n; ++i) nzb |= (unsigned)(l[i] != 0) << i;
    const int total = popcount32(nzb);
    int t1 = 0;
    for (unsigned m = nzb; m && t1 < 3; ++t1) {                // trailing ones: the last (highest-frequency) levels, while +-1
        const int p = 31 - ent_clz(m), lv = l[p];
        if (lv != 1 && lv != -1) break;
        m &= ~(1u << p);
    }
    if (nC == -1) {
        bo_put(bw, etab::kChromaDcTokenBits[4 * total + t1], etab::kChromaDcTokenLen[4 * total + t1]);
    } else {
        const int tab = nC < 2 ? 0 : nC < 4 ? 1 : nC < 8 ? 2 : 3;
        bo_put(bw, etab::kCoeffTokenBits[tab][4 * total + t1], etab::kCoeffTokenLen[tab][4 * total + t1]);
    }
    if (!total) return 0;
    unsigned m = nzb;
    for (int k = 0; k < t1; ++k) {                             // trailing_ones_sign_flag
        const int p = 31 - ent_clz(m);
        bo_put(bw, l[p] < 0, 1);
        m &= ~(1u << p);
    }
    int suffix_len = (total > 10 && t1 < 3) ? 1 : 0;
    for (int k = t1; k < total; ++k) {
        const int p = 31 - ent_clz(m), lv = l[p];
        m &= ~(1u << p);
        int code = lv > 0 ? 2 * lv - 2 : -2 * lv - 1;          // levelCode
        if (k == t1 && t1 < 3) code -= 2;
        if (suffix_len == 0) {
            if (code < 14) {
                bo_put(bw, 1, code + 1);
            } else if (code < 30) {
                bo_put(bw, 1, 15);
                bo_put(bw, (uint32_t)(code - 14), 4);
            } else {
                int prefix = 15, base = 30;                    // level_prefix >= 15: suffix of (prefix - 3) bits
                while (code - base >= (1 << (prefix - 3))) { base += 1 << (prefix - 3); ++prefix; }
                bo_put(bw, 0, prefix > 31 ? 31 : prefix);
                bo_put(bw, 1, 1);
                bo_put(bw, (uint32_t)(code - base), prefix - 3);
            }
        } else {
            if (code < (15 << suffix_len)) {
                bo_put(bw, 1, (code >> suffix_len) + 1);
                bo_put(bw, (uint32_t)(code & ((1 << suffix_len) - 1)), suffix_len);
            } else {
                int prefix = 15, base = 15 << suffix_len;
                while (code - base >= (1 << (prefix - 3))) { base += 1 << (prefix - 3); ++prefix; }
                bo_put(bw, 0, prefix);
                bo_put(bw, 1, 1);
                bo_put(bw, (uint32_t)(code - base), prefix - 3);
            }
        }
        if (suffix_len == 0) suffix_len = 1;
        if ((lv < 0 ? -lv : lv) > (3 << (suffix_len - 1)) && suffix_len < 6) ++suffix_len;
    }
    const int last = 31 - ent_clz(nzb), total_zeros = last + 1 - total;
    if (total < n) {
        if (nC == -1) bo_put(bw, etab::kChromaDcTotalZerosBits[total - 1][total_zeros], etab::kChromaDcTotalZerosLen[total - 1][total_zeros]);
        else          bo_put(bw, etab::kTotalZerosBits[total - 1][total_zeros], etab::kTotalZerosLen[total - 1][total_zeros]);
    }
    int zeros_left = total_zeros;
    m = nzb;
    for (int k = 0; k < total - 1 && zeros_left > 0; ++k) {    // run_before of every level but the lowest, highest first
        const int p = 31 - ent_clz(m);
        m &= ~(1u << p);
        const int run = p - (31 - ent_clz(m)) - 1, t = (zeros_left < 7 ? zeros_left : 7) - 1;
        bo_put(bw, etab::kRunBits[t][run], etab::kRunLen[t][run]);
        zeros_left -= run;
    }
    return total;
}
GVC_HD void ent_empty_block(BitOut &bw, int nC) {              // coeff_token with TotalCoeff = 0
    if (nC == -1) { bo_put(bw, etab::kChromaDcTokenBits[0], etab::kChromaDcTokenLen[0]); return; }
    const int tab = nC < 2 ? 0 : nC < 4 ? 1 : nC < 8 ? 2 : 3;
    bo_put(bw, etab::kCoeffTokenBits[tab][0], etab::kCoeffTokenLen[tab][0]);
}
GVC_HD int ent_nc(const uint8_t *map, int pitch, int bx, int by) {   // 9.2.1: from the blocks to the left and above (one slice per picture)
    const bool a = bx > 0, b = by > 0;
    const int  na = a ? map[by * pitch + bx - 1] : 0, nb = b ? map[(by - 1) * pitch + bx] : 0;
    return a && b ? (na + nb + 1) >> 1 : a ? na : b ? nb : 0;
}
GVC_HD int ent_cbp(uint32_t nz, bool i16) {
    int luma = 0;
    if (nz & 0x0033u) luma |= 1;
    if (nz & 0x00ccu) luma |= 2;
    if (nz & 0x3300u) luma |= 4;
    if (nz & 0xcc00u) luma |= 8;
    if (i16 && luma) luma = 15;
    const int chroma = (nz & 0x00ff0000u) ? 2 : (nz & 0x03000000u) ? 1 : 0;
    return luma | (chroma << 4);
}
// ---- motion vector prediction (8.4.1.3), single reference ---------------------------------------------
struct EntMv { int x, y; };
// vector of the 4x4 block (x, y) if it precedes partition cur_k of macroblock cur_mb in decoding order
GVC_HD bool ent_mv_at(const EntropyArgs &a, int cur_mb, int cur_k, int x, int y, EntMv &out) {
    if (x < 0 || y < 0 || x >= 4 * a.mbw || y >= 4 * a.mbh) return false;
    const int mb = (y >> 2) * a.mbw + (x >> 2), k = ((y & 3) >> 1) * 2 + ((x & 3) >> 1);
    if (mb > cur_mb || (mb == cur_mb && k >= cur_k)) return false;
    out.x = a.motion[mb].mv[k][0];
    out.y = a.motion[mb].mv[k][1];
    return true;
}
GVC_HD int ent_median3(int a, int b, int c) { return imax(imin(a, b), imin(imax(a, b), c)); }
GVC_HD EntMv ent_predict_mv(const EntropyArgs &a, int cur_mb, int cur_k, int bx, int by, int w, bool &A, bool &B, EntMv &mva, EntMv &mvb) {
    EntMv c;
    mva.x = mva.y = mvb.x = mvb.y = c.x = c.y = 0;
    A = ent_mv_at(a, cur_mb, cur_k, bx - 1, by, mva);
    B = ent_mv_at(a, cur_mb, cur_k, bx, by - 1, mvb);
    bool C = ent_mv_at(a, cur_mb, cur_k, bx + w, by - 1, c);
    if (!C) C = ent_mv_at(a, cur_mb, cur_k, bx - 1, by - 1, c);
    if (!B && !C && A) return mva;
    const int n = (int)A + (int)B + (int)C;
    if (n == 1) return A ? mva : B ? mvb : c;
    EntMv r;
    r.x = ent_median3(mva.x, mvb.x, c.x);
    r.y = ent_median3(mva.y, mvb.y, c.y);
    return r;
}

// ---- one macroblock: macroblock_layer() without the preceding mb_skip_run ------------------------------
// One warp per macroblock, one lane per syntax element group: lanes 0..15 the luma blocks (lane = blkIdx, coding order),
// 16..23 the chroma AC blocks (Cb 0..3, Cr 0..3), 24 / 25 the chroma DC blocks, 26 the Intra16x16 luma DC block, 27 the
// header (mb_type, sub types, mvds, coded_block_pattern, mb_qp_delta).  Each lane writes its bits into its own small
// buffer; the warp then lays them end to end in coding order (header, luma DC, luma, chroma DC, chroma AC).
constexpr int kEntLanes = 28, kEntLaneWords = 21;      // a residual block is at most 16 + 16*36 + 9 + 45 = 646 bits
struct EntSmem {                                       // per warp
    uint32_t lane_buf[kEntLanes][kEntLaneWords];
    uint32_t mb[kMbBufWords];
    int      kind, overflow;
};
GVC_HD int ent_lane_order(int lane) { return lane == 27 ? 0 : lane == 26 ? 1 : lane < 16 ? 2 + lane : lane < 24 ? 4 + lane : lane - 6; }
GVC_HD int ent_lane_at(int pos) { return pos == 0 ? 27 : pos == 1 ? 26 : pos < 18 ? pos - 2 : pos < 20 ? pos + 6 : pos < 28 ? pos - 4 : pos; }

// the lane's bits into sm.lane_buf[lane]; returns their number.  Lane 27 also decides P_Skip and sets sm.kind.
GVC_HD int ent_lane_encode(int lane, const EntropyArgs &a, int mb, EntSmem &sm) {
    if (lane >= kEntLanes) return 0;
    const int      my = mb / a.mbw, mx = mb - my * a.mbw;
    const uint32_t nz = a.nzmask[mb];
    const int16_t *co = a.blocks + (size_t)16 * a.offsets[mb];
    const bool     i16 = a.idr != 0;
    const int      cbp = ent_cbp(nz, i16);
    const unsigned cm = compact_mask(nz);
    BitOut bw;
    bo_init(bw, sm.lane_buf[lane], kEntLaneWords);
    if (lane == 27) {
        if (i16) {
            bo_ue(bw, (unsigned)(1 + a.modes[mb] + 4 * (cbp >> 4) + ((cbp & 15) ? 12 : 0)));   // mb_type I_16x16_<mode>_<cbpc>_<cbpl>
            bo_ue(bw, 0);                  // intra_chroma_pred_mode: DC
            bo_se(bw, 0);                  // mb_qp_delta
            sm.kind = 3;
        } else {
            const MbMotion &mo = a.motion[mb];
            if (mo.mode == 0) {
                bool  A, B;
                EntMv va, vb;
                const EntMv mvp = ent_predict_mv(a, mb, 0, 4 * mx, 4 * my, 4, A, B, va, vb);
                EntMv skip = mvp;
                if (!A || !B || (va.x == 0 && va.y == 0) || (vb.x == 0 && vb.y == 0)) skip.x = skip.y = 0;
                if (cbp == 0 && mo.mv[0][0] == skip.x && mo.mv[0][1] == skip.y) { sm.kind = 0; return 0; }   // P_Skip: no bits of its own
                bo_ue(bw, 0);              // mb_type P_L0_16x16
                bo_se(bw, mo.mv[0][0] - mvp.x);
                bo_se(bw, mo.mv[0][1] - mvp.y);
                sm.kind = 1;
            } else {
                bo_ue(bw, 3);              // mb_type P_8x8
                for (int k = 0; k < 4; ++k) bo_ue(bw, 0);   // sub_mb_type P_L0_8x8
                for (int k = 0; k < 4; ++k) {
                    bool  A, B;
                    EntMv va, vb;
                    const EntMv mvp = ent_predict_mv(a, mb, k, 4 * mx + 2 * (k & 1), 4 * my + 2 * (k >> 1), 2, A, B, va, vb);
                    bo_se(bw, mo.mv[k][0] - mvp.x);
                    bo_se(bw, mo.mv[k][1] - mvp.y);
                }
                sm.kind = 2;
            }
            bo_ue(bw, etab::kCbpInterCode[cbp]);   // coded_block_pattern me(v)
            if (cbp) bo_se(bw, 0);                 // mb_qp_delta
        }
    } else {
        // Residual lanes.  Which block, how many coefficients and which nC differ per lane; the coder itself is called from ONE
        // place, so luma, chroma AC and DC lanes run it together instead of one group after the other.
        const int16_t *l = co;
        const uint8_t *map = a.nnz_y;
        int  n = 16, pitch = 4 * a.mbw, bx = 0, by = 0, bit = 0;
        bool coded = false, dc = false;
        if (lane == 26) {                                   // Intra16x16 luma DC
            coded = i16; bit = 26; bx = 4 * mx; by = 4 * my;
            l = co + 16 * popcount32(cm & ((1u << 25) - 1u));
        } else if (lane < 16) {
            const int r = ((lane & 8) | ((lane & 4) >> 1) | ((lane & 2) << 1) | (lane & 1));    // blkIdx (8x8-wise Z scan) -> raster 4x4 block
            coded = (cbp & (1 << (lane >> 2))) != 0; bit = r; bx = 4 * mx + (r & 3); by = 4 * my + (r >> 2);
            l = co + 16 * popcount32(cm & ((1u << r) - 1u)) + (i16 ? 1 : 0);
            n = i16 ? 15 : 16;
        } else if (lane < 24) {
            const int p = (lane - 16) >> 2, k = (lane - 16) & 3;
            coded = (cbp & 0x20) != 0; bit = lane; map = a.nnz_c[p]; pitch = 2 * a.mbw; bx = 2 * mx + (k & 1); by = 2 * my + (k >> 1);
            l = co + 16 * popcount32(cm & ((1u << lane) - 1u)) + 1;
            n = 15;
        } else {                                            // lanes 24, 25: chroma DC
            const int p = lane - 24;
            coded = (cbp & 0x30) != 0; bit = 24 + p; dc = true;
            l = co + 16 * popcount32(cm & ((1u << 24) - 1u)) + 4 * p;
            n = 4;
        }
        if (coded) {
            const int nC = dc ? -1 : ent_nc(map, pitch, bx, by);
            if (nz & (1u << bit)) ent_residual(bw, l, n, nC);
            else ent_empty_block(bw, nC);
        }
    }
    const int bits = bo_finish(bw);
    if (bw.overflow) sm.overflow = 1;
    return bits;
}

// ---- slice assembly -------------------------------------------------------------------------------------
struct SliceArgs {
    int             nmb;
    const uint32_t *mb_buf, *mb_bits;
    const uint8_t  *mb_kind;
    uint32_t       *mb_off;        // [nmb] bit offset of the macroblock's mb_skip_run (coded macroblocks)
    uint32_t       *mb_run;        // [nmb] its mb_skip_run
    uint32_t       *slice;         // slice RBSP under construction, 32-bit words, MSB first (zeroed by ent_scan)
    int             slice_words;   // capacity
    int             header_bits;   // room left at the front for slice_header(), which the host fills in (its length is fixed
                                   // per picture type, its idr_pic_id changes from segment to segment: the launch stays replayable)
    int             p_slice;       // P slices carry mb_skip_run
    uint32_t       *info;          // [8]: bits, bytes, skipped, p16, p8, intra, error, -
};
// OR `bits` bits (left-aligned in words w[]) into dst at bit offset `off`
GVC_HD void ent_or_word(uint32_t *p, uint32_t v) {
    if (!v) return;
#if defined(GVC_EMU)
    *p |= v;
#else
    atomicOr(p, v);
#endif
}
GVC_HD void ent_or_bits(uint32_t *dst, long long cap_words, unsigned long long off, const uint32_t *w, int bits) {
    const int       s = (int)(off & 31);
    long long       j = (long long)(off >> 5);
    const int       words = (bits + 31) >> 5;
    for (int i = 0; i < words; ++i, ++j) {
        uint32_t v = w[i];
        if (i == words - 1 && (bits & 31)) v &= ~0u << (32 - (bits & 31));    // ignore what lies past the last bit
        if (j < cap_words) ent_or_word(dst + j, v >> s);
        if (s && j + 1 < cap_words) ent_or_word(dst + j + 1, v << (32 - s));
    }
}
// coded macroblock mb: ue(mb_skip_run) (P slices) followed by its own bits
GVC_HD void ent_place(const SliceArgs &a, int mb) {
    if ((a.mb_kind[mb] & 127) == 0) return;
    unsigned long long off = a.mb_off[mb];
    if (a.p_slice) {
        const unsigned run = a.mb_run[mb], code = run + 1;
        const int      len = ue_len(run);
        uint32_t       w[2];
        // ue(v): (len-1)/2 zeros then code in (len+1)/2 bits = `code` right-aligned in a len-bit field
        const unsigned long long field = (unsigned long long)code << (64 - len);
        w[0] = (uint32_t)(field >> 32);
        w[1] = (uint32_t)field;
        ent_or_bits(a.slice, a.slice_words, off, w, len);
        off += (unsigned)len;
    }
    ent_or_bits(a.slice, a.slice_words, off, a.mb_buf + (size_t)mb * kMbBufWords, (int)a.mb_bits[mb]);
}
// after the scan: the trailing mb_skip_run of a P slice, rbsp_trailing_bits(); fills info[]
GVC_HD void ent_finish_slice(const SliceArgs &a, unsigned long long data_bits, unsigned trailing_run, const unsigned counts[4], bool overflow) {
    unsigned long long bits = (unsigned long long)a.header_bits + data_bits;
    if (a.p_slice && trailing_run) {
        const int len = ue_len(trailing_run);
        const unsigned long long field = (unsigned long long)(trailing_run + 1) << (64 - len);
        uint32_t w[2] = {(uint32_t)(field >> 32), (uint32_t)field};
        ent_or_bits(a.slice, a.slice_words, bits, w, len);
        bits += (unsigned)len;
    }
    const uint32_t one = 0x80000000u;
    ent_or_bits(a.slice, a.slice_words, bits, &one, 1);          // rbsp_stop_one_bit; alignment zeros are already there
    bits += 1;
    const unsigned long long bytes = (bits + 7) >> 3;
    a.info[0] = (uint32_t)bits;
    a.info[1] = (uint32_t)bytes;
    a.info[2] = counts[0]; a.info[3] = counts[1]; a.info[4] = counts[2]; a.info[5] = counts[3];
    a.info[6] = (overflow || bytes > (unsigned long long)a.slice_words * 4) ? 1u : 0u;
}

// everything one picture's entropy pass needs (launch_entropy_picture / the emulation in the CPU tests)
struct EntropyPicture {
    EntropyArgs ent;
    SliceArgs   slice;
    uint8_t    *host_buf;      // the segment's slice buffer (mapped host memory, device address)
    unsigned    host_cap;
    unsigned   *used;          // device: used[0] = bytes taken by earlier pictures, used[1] receives the new total
    uint32_t   *pos;           // host (mapped), 8 words: base, bytes, bits, skipped, p16, p8, intra, error
};

}  // namespace gvc
