// me_refine.cuh -- K4b: full-resolution motion refinement of one picture (included by enc_kernels.cuh).
//
// Specification: oracle/h264_mb.c gvc_oracle_me_refine (x264's metrics -- pixel_sad, pixel_satd -- and cost = metric +
// lambda * bits(mvd); 27 integer candidates, then half- and quarter-sample SATD refinement of the 16x16 partition and of
// the four 8x8 partitions, early termination and one-partition rule as stated there).
//
// One warp per macroblock.  The reference samples every candidate reads are staged ONCE in shared memory:
//   * four windows of 48 bytes x 22 rows, one per sample plane (full, H, V, C).  They hold columns wbx-3 .. wbx+18 and rows
//     wby-3 .. wby+18 relative to the macroblock, (wbx, wby) = twice the lowres vector (kept inside the legal vector range):
//     the 5x5 integer candidates around it and every sub-sample position within 3/4 sample of any of them;
//   * one window of the full plane at the predictor candidate (16 rows used).
// On the GPU the windows arrive by TMA (cp.async.bulk.tensor.2d: one box per window, completion on an mbarrier), so no
// address arithmetic or L1 traffic is spent on them and candidates read shared memory (LDS + one funnel shift).  A TMA box
// must start on a 16-byte boundary of the innermost dimension (measured: tools/tma_probe.cu -- any other start raises
// "illegal instruction" on B200), so a window's first column is the wanted one rounded down to a multiple of 16 and the
// remainder (ax, 0..15) is added to every column index; 22 + 15 columns fit the 48-byte box.  Planes that do not meet
// TMA's alignment rules (kernel-level C-ABI callers) and the CPU emulation fill the same windows with plain loads.
// A partition whose integer vector is the predictor or the zero vector AND lies outside the 5x5 grid (under 1 % of the
// macroblocks of the bench clip) reads the planes directly, as the first version of this kernel did for everything.
#pragma once

namespace gvc {

struct MeRefArgs {
    const uint8_t *cur;          // source luma (0,0)
    RefPlanes      ref;
    int            stride, w16, h16, mbw, mbh, lambda;
    const int16_t *lowres_mv;
    MbMotion      *out;
    MbMotion      *out_host;     // optional second copy of the result (host-mapped memory), or null
};
#ifndef GVC_ME_EARLY_MUL
#define GVC_ME_EARLY_MUL 32
#endif
constexpr int kMeEarlyMul = GVC_ME_EARLY_MUL;   // early termination: zero-vector SAD <= kMeEarlyMul * lambda (oracle/h264_mb.c)

constexpr int kMeWinPitch = 48, kMeWinPitchW = kMeWinPitch / 4, kMeWinRows = 22;
constexpr int kMeWinWords = 288;                // 1056 bytes used of 1152 (keeps every window 128-byte aligned; the tail absorbs the
                                                // second word of a two-word fetch at the end of the last row)
constexpr int kMeWinX = 3, kMeWinY = 3;         // columns / rows of margin left of and above integer vector (wbx, wby) - (0, 0)
constexpr int kMePredRows = 16;

struct
#if !defined(GVC_EMU)
    __align__(128)
#endif
    MeRefSmem {                  // per warp
    unsigned           win[4][kMeWinWords];   // full, H, V, C sample planes (mc_luma_src order)
    unsigned           pred[kMeWinWords];     // full plane at the predictor candidate
    gvc_uint4          cur[16];               // the macroblock's 16 source rows
    unsigned long long bar[2];                // mbarriers: [0] full + predictor windows, [1] H, V, C windows
    int                mvp[2], low[2];
};

GVC_HD void sq8(int i, int &dx, int &dy) {   // (0,-1),(0,1),(-1,0),(1,0),(-1,-1),(-1,1),(1,-1),(1,1)
    dx = i < 2 ? 0 : ((i == 2 || i == 4 || i == 5) ? -1 : 1);
    dy = i == 0 ? -1 : i == 1 ? 1 : (i < 4 ? 0 : ((i & 1) ? 1 : -1));
}
GVC_HD void me_mvp(const MeRefArgs &a, int mx, int my, int &px, int &py) {
    int v[3][2];
    int nx[3] = {mx - 1, mx, mx + 1}, ny[3] = {my, my - 1, my - 1};
    if (!(nx[2] < a.mbw && ny[2] >= 0)) { nx[2] = mx - 1; ny[2] = my - 1; }
    for (int k = 0; k < 3; ++k) {
        const bool ok = nx[k] >= 0 && nx[k] < a.mbw && ny[k] >= 0 && ny[k] < a.mbh;
        v[k][0] = ok ? 8 * a.lowres_mv[2 * (ny[k] * a.mbw + nx[k])] : 0;
        v[k][1] = ok ? 8 * a.lowres_mv[2 * (ny[k] * a.mbw + nx[k]) + 1] : 0;
    }
    px = median3(v[0][0], v[1][0], v[2][0]);
    py = median3(v[0][1], v[1][1], v[2][1]);
}
GVC_HDH bool me_cur_aligned(const MeRefArgs &a) { return ((((unsigned long long)a.cur) | (unsigned)a.stride) & 15) == 0; }
GVC_HD void me_ref_load(int lane, const MeRefArgs &a, int mx, int my, MeRefSmem &sm) {
    if (lane < 16) {
        const uint8_t *p = a.cur + (long long)(16 * my + lane) * a.stride + 16 * mx;
        gvc_uint4 r;
        if (me_cur_aligned(a)) {
            r = *reinterpret_cast<const gvc_uint4 *>(p);
        } else {
            unsigned w[4];
            ld_u128_any(p, w);
            r.x = w[0]; r.y = w[1]; r.z = w[2]; r.w = w[3];
        }
        sm.cur[lane] = r;
    } else if (lane == 16) {
        int px, py;
        me_mvp(a, mx, my, px, py);
        sm.mvp[0] = px; sm.mvp[1] = py;
    } else if (lane == 17) {
        sm.low[0] = a.lowres_mv[2 * (my * a.mbw + mx)];
        sm.low[1] = a.lowres_mv[2 * (my * a.mbw + mx) + 1];
    }
}
// the lane's share of the 16x16 SAD at the zero vector (two lanes per row: lane = 2 * row + half).  The sum over all
// lanes decides the early termination; the sums per 8x8 quadrant (rows 0..7 / 8..15 x halves) are candidate 26's SADs.
GVC_HD unsigned me_ref_zero_part(int lane, const MeRefArgs &a, int mx, int my, const MeRefSmem &sm) {
    const int row = lane >> 1, half = lane & 1;
    unsigned r0, r1;
    ld_u64_any(a.ref.y + ((16 * my + row) * a.stride + 16 * mx + 8 * half), r0, r1);
    const gvc_uint4 c = sm.cur[row];
    return half ? sad_u8x4(c.w, r1, sad_u8x4(c.z, r0, 0)) : sad_u8x4(c.y, r1, sad_u8x4(c.x, r0, 0));
}
GVC_HD int me_zero_quadrant(int lane) { return ((lane >> 4) << 1) | (lane & 1); }   // which 8x8 the lane's share belongs to

// Geometry every lane derives from the shared record: legal vector range, window centre, predictor candidate, and where
// the two kinds of window start in the planes (x rounded down to a multiple of 16 bytes of the padded plane: TMA's rule).
struct MeGeom {
    int xmin, xmax, ymin, ymax, wbx, wby, pcx, pcy;
    int wox, woy, ax;        // windows: first sample (picture coordinates), ax = columns between it and column wbx - kMeWinX
    int pox, poy, pax;       // predictor window likewise (pax = columns between its first sample and the candidate)
};
GVC_HD MeGeom me_geom(const MeRefArgs &a, const MeRefSmem &sm, int mx, int my) {
    MeGeom g;
    g.xmin = -16 * mx - (kEncPad - 8); g.xmax = a.w16 - 16 * mx - 16 + (kEncPad - 8);
    g.ymin = -16 * my - (kEncPad - 8); g.ymax = a.h16 - 16 * my - 16 + (kEncPad - 8);
    // the clamped 5x5 candidates around twice the lowres vector all lie within +-2 of this centre
    g.wbx = iclamp(2 * sm.low[0], g.xmin + 2, g.xmax - 2);
    g.wby = iclamp(2 * sm.low[1], g.ymin + 2, g.ymax - 2);
    g.pcx = iclamp((sm.mvp[0] + 2) >> 2, g.xmin, g.xmax);
    g.pcy = iclamp((sm.mvp[1] + 2) >> 2, g.ymin, g.ymax);
    const int wx = 16 * mx + g.wbx - kMeWinX, px = 16 * mx + g.pcx;     // kEncPad is a multiple of 16: rounding in picture or plane coordinates agrees
    g.ax = wx & 15; g.wox = wx - g.ax; g.woy = 16 * my + g.wby - kMeWinY;
    g.pax = px & 15; g.pox = px - g.pax; g.poy = 16 * my + g.pcy;
    return g;
}
// integer candidate i of the macroblock: 5x5 around twice the lowres vector, the rounded predictor, zero
GVC_HD void me_int_cand(const MeRefSmem &sm, const MeGeom &g, int i, int &cx, int &cy) {
    if (i < 25)       { cx = iclamp(2 * sm.low[0] + (i % 5) - 2, g.xmin, g.xmax); cy = iclamp(2 * sm.low[1] + (i / 5) - 2, g.ymin, g.ymax); }
    else if (i == 25) { cx = g.pcx; cy = g.pcy; }
    else              { cx = iclamp(0, g.xmin, g.xmax); cy = iclamp(0, g.ymin, g.ymax); }
}
GVC_HD bool me_in_window(const MeGeom &g, int cx, int cy) {   // integer vector whose sub-sample neighbourhood the windows hold
    return cx >= g.wbx - 2 && cx <= g.wbx + 2 && cy >= g.wby - 2 && cy <= g.wby + 2;
}

// Plain staging of the windows (CPU emulation; GPU callers whose planes TMA cannot address): same bytes as the TMA boxes.
GVC_HD void me_stage_plain(int lane, const MeRefArgs &a, const MeGeom &g, MeRefSmem &sm) {
    const uint8_t *pl[4] = {a.ref.y, a.ref.hh, a.ref.hv, a.ref.hc};
    const long long org = (long long)g.woy * a.stride + g.wox;
    for (int i = lane; i < 4 * kMeWinRows * kMeWinPitchW; i += 32) {
        const int p = i / (kMeWinRows * kMeWinPitchW), j = i - p * (kMeWinRows * kMeWinPitchW), r = j / kMeWinPitchW, c = j - r * kMeWinPitchW;
        sm.win[p][j] = ld_u32_any(pl[p] + org + (long long)r * a.stride + 4 * c);
    }
    const long long porg = (long long)g.poy * a.stride + g.pox;
    for (int i = lane; i < kMePredRows * kMeWinPitchW; i += 32) {
        const int r = i / kMeWinPitchW, c = i - r * kMeWinPitchW;
        sm.pred[i] = ld_u32_any(a.ref.y + porg + (long long)r * a.stride + 4 * c);
    }
}

GVC_HD unsigned funnel_rc(unsigned lo, unsigned hi, int bits) {    // (hi:lo) >> bits for 0 <= bits <= 32
#if defined(GVC_EMU)
    return (unsigned)((((unsigned long long)hi << 32) | lo) >> bits);
#else
    return __funnelshift_rc(lo, hi, (unsigned)bits);
#endif
}
// Keys of integer candidate `lane` for the 16x16 partition (key[0]) and the four 8x8 partitions: (SAD + vector cost) << 5 | lane.
// Lanes 0..24 read the full-plane window, lane 25 the predictor window (same code, different base), lane 26 takes the
// quadrant SADs of the zero vector that the early-termination test already produced.
GVC_HD void me_ref_int(int lane, const MeRefArgs &a, const MeRefSmem &sm, const MeGeom &g, const unsigned zq[4], unsigned key[5]) {
    for (int k = 0; k < 5; ++k) key[k] = 0xffffffffu;
    if (lane >= 27) return;
    int cx, cy;
    me_int_cand(sm, g, lane, cx, cy);
    unsigned s[4] = {zq[0], zq[1], zq[2], zq[3]};
    if (lane < 26) {
        // first sample of the candidate inside its window
        const int col = lane < 25 ? g.ax + kMeWinX + (cx - g.wbx) : g.pax, row0 = lane < 25 ? kMeWinY + (cy - g.wby) : 0;
        const unsigned *row = (lane < 25 ? sm.win[0] : sm.pred) + row0 * kMeWinPitchW + (col >> 2);
        const int sh = 8 * (col & 3);
        s[0] = s[1] = s[2] = s[3] = 0;
        for (int j = 0; j < 16; ++j, row += kMeWinPitchW) {
            const unsigned w0 = row[0], w1 = row[1], w2 = row[2], w3 = row[3], w4 = row[4];
            const unsigned r0 = funnel_r(w0, w1, sh), r1 = funnel_r(w1, w2, sh), r2 = funnel_r(w2, w3, sh), r3 = funnel_r(w3, w4, sh);
            const gvc_uint4 c = sm.cur[j];
            const int k = (j >> 3) * 2;
            s[k]     = sad_u8x4(c.y, r1, sad_u8x4(c.x, r0, s[k]));
            s[k + 1] = sad_u8x4(c.w, r3, sad_u8x4(c.z, r2, s[k + 1]));
        }
    }
    const unsigned bits = (unsigned)(a.lambda * mv_bits(4 * cx, 4 * cy, sm.mvp[0], sm.mvp[1]));
    key[0] = ((s[0] + s[1] + s[2] + s[3] + bits) << 5) | (unsigned)lane;
    for (int k = 0; k < 4; ++k) key[1 + k] = ((s[k] + bits) << 5) | (unsigned)lane;
}
// early termination: the zero vector predicts the macroblock well enough; one 16x16 partition, no sub-sample search
GVC_HD bool me_ref_early(const MeRefArgs &a, int zero_sad) { return zero_sad <= kMeEarlyMul * a.lambda; }
GVC_HD void me_ref_early_out(int lane, const MeRefArgs &a, int mx, int my, const MeRefSmem &sm, int zero_sad) {
    if (lane != 0) return;
    MbMotion m;
    m.mode = 0;
    m.cost = zero_sad + a.lambda * mv_bits(0, 0, sm.mvp[0], sm.mvp[1]);
    for (int k = 0; k < 4; ++k) m.mv[k][0] = m.mv[k][1] = 0;
    m.pad[0] = m.pad[1] = m.pad[2] = 0;
    a.out[my * a.mbw + mx] = m;
    if (a.out_host) a.out_host[my * a.mbw + mx] = m;
}
