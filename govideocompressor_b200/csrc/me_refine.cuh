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
    int                mvp[2], low[2], nb[6];
};

GVC_HD void sq8(int i, int &dx, int &dy) {   // (0,-1),(0,1),(-1,0),(1,0),(-1,-1),(-1,1),(1,-1),(1,1): d + 1 in two bits per entry
    dx = (int)((0xA085u >> (2 * i)) & 3u) - 1;
    dy = (int)((0x8858u >> (2 * i)) & 3u) - 1;
}
// ---- candidate descriptors ---------------------------------------------------------------------------------------------
// Where a sub-sample candidate's samples come from depends only on the fractional part of the search centre (0 or 2 per
// component) and on the candidate: mc_luma_src's plane pair, the full-sample displacement of each plane's block relative to the
// centre's, and whether a second plane is averaged in.  All 9 + 4 x 8 cases are worked out at compile time into one word each:
// bits 0-1 plane 1, 2-3 column + 1, 4-5 row + 1, 6 second plane used, 7-8 plane 2, 9-10 column + 1, 11-12 row + 1, 13-14 dx + 1,
// 15-16 dy + 1 ((dx, dy) = the candidate's direction, sq8 order; the centre itself is direction (0, 0)).
constexpr int me_dx8(int i) { return (int)((0xA085u >> (2 * i)) & 3u) - 1; }
constexpr int me_dy8(int i) { return (int)((0x8858u >> (2 * i)) & 3u) - 1; }
constexpr unsigned me_desc_make(int fcx, int fcy, int dx, int dy, int step) {
    const int mvx = fcx + dx * step, mvy = fcy + dy * step;          // relative to the centre's full-sample position
    const int X = mvx >> 2, Y = mvy >> 2, fx = mvx & 3, fy = mvy & 3, q = fy * 4 + fx;
    const unsigned p1 = (kQpelRef0 >> (2 * q)) & 3u, p2 = (kQpelRef1 >> (2 * q)) & 3u, two = (q & 5) ? 1u : 0u;
    return p1 | (unsigned)(X + 1) << 2 | (unsigned)(Y + (fy == 3 ? 1 : 0) + 1) << 4 | two << 6 | p2 << 7 | (unsigned)(X + (fx == 3 ? 1 : 0) + 1) << 9 |
           (unsigned)(Y + 1) << 11 | (unsigned)(dx + 1) << 13 | (unsigned)(dy + 1) << 15;
}
struct MeDescTable { unsigned v[41]; };
constexpr MeDescTable me_desc_table() {
    MeDescTable t{};
    t.v[0] = me_desc_make(0, 0, 0, 0, 2);
    for (int i = 0; i < 8; ++i) t.v[1 + i] = me_desc_make(0, 0, me_dx8(i), me_dy8(i), 2);
    for (int c = 0; c < 4; ++c)
        for (int i = 0; i < 8; ++i) t.v[9 + 8 * c + i] = me_desc_make((c & 1) * 2, (c >> 1) * 2, me_dx8(i), me_dy8(i), 1);
    return t;
}
#if defined(GVC_EMU)
static const MeDescTable kMeDesc = me_desc_table();
#else
__constant__ const MeDescTable kMeDesc = me_desc_table();
#endif
// descriptor of candidate ci (-1 = the centre, step 2 only) of a step around centre (cx, cy) in quarter samples
GVC_HD unsigned me_desc(int cx, int cy, int ci, int step) {
    return kMeDesc.v[step == 2 ? ci + 1 : 9 + 8 * (((cx >> 1) & 1) | (cy & 2)) + ci];
}

// component c of neighbour k's lowres vector in quarter samples (k: left, above, above right -- or above left when there is none), 0 outside
GVC_HD int me_neighbour(const MeRefArgs &a, int mx, int my, int k, int c) {
    int nx = k == 0 ? mx - 1 : k == 1 ? mx : mx + 1, ny = k == 0 ? my : my - 1;
    if (k == 2 && !(nx < a.mbw && ny >= 0)) nx = mx - 1;
    const bool ok = nx >= 0 && nx < a.mbw && ny >= 0 && ny < a.mbh;
    return ok ? 8 * a.lowres_mv[2 * (ny * a.mbw + nx) + c] : 0;
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
    } else if (lane < 22) {          // six lanes fetch the neighbours' vectors; every lane takes the medians itself (me_mvp)
        const int i = lane - 16;
        sm.nb[i] = me_neighbour(a, mx, my, i >> 1, i & 1);
    } else if (lane < 24) {
        sm.low[lane - 22] = a.lowres_mv[2 * (my * a.mbw + mx) + (lane - 22)];
    }
}
GVC_HD void me_mvp(MeRefSmem &sm, int lane) {      // after me_ref_load: the median predictor into the shared record (same value from every lane)
    const int px = median3(sm.nb[0], sm.nb[2], sm.nb[4]), py = median3(sm.nb[1], sm.nb[3], sm.nb[5]);
    if (lane == 0) { sm.mvp[0] = px; sm.mvp[1] = py; }
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

// ------------------------------------------------------------------ sub-sample refinement ------
// lane -> (partition P, 4x4 block offset inside the MB): lanes 0..15 the 16x16 partition, 16..31 the four 8x8s
GVC_HD void me_lane_block(int lane, int &P, int &ox, int &oy) {
    if (lane < 16) { P = 0; ox = 4 * (lane & 3); oy = 4 * (lane >> 2); }
    else { const int l = lane - 16, k = l >> 2, j = l & 3; P = 1 + k; ox = 8 * (k & 1) + 4 * (j & 1); oy = 8 * (k >> 1) + 4 * (j >> 1); }
}
// Half the 4x4 Hadamard magnitude sum of (w - source).  The horizontal transform of a row of four samples is four dp4a
// with +-1 patterns (FMA pipe: the integer ALU pipe is the busy one in these kernels), started from the negated transform
// of the source row (hc, computed once per macroblock), so the subtraction is free.  Vertical: one butterfly stage, the last
// one through |p + q| + |p - q| = 2 max(|p|, |q|).
GVC_HD int absdiff(int a, int b) {
#if defined(GVC_EMU)
    return a > b ? a - b : b - a;
#else
    return (int)__sad(a, b, 0u);
#endif
}
constexpr int kHad0 = 0x01010101, kHad1 = (int)0xffff0101, kHad2 = 0x01ffff01, kHad3 = (int)0xff01ff01;   // (1,1,1,1) (1,1,-1,-1) (1,-1,-1,1) (1,-1,1,-1)
GVC_HD void had_row_neg(unsigned w, int out[4]) {     // minus the horizontal Hadamard transform of four samples
    out[0] = -dp4a_us(w, kHad0, 0); out[1] = -dp4a_us(w, kHad1, 0); out[2] = -dp4a_us(w, kHad2, 0); out[3] = -dp4a_us(w, kHad3, 0);
}
GVC_HD int satd4x4_rows(const int hc[16], unsigned w0, unsigned w1, unsigned w2, unsigned w3) {
    const unsigned w[4] = {w0, w1, w2, w3};
    int t[16];
    for (int j = 0; j < 4; ++j) {
        t[4 * j]     = dp4a_us(w[j], kHad0, hc[4 * j]);
        t[4 * j + 1] = dp4a_us(w[j], kHad1, hc[4 * j + 1]);
        t[4 * j + 2] = dp4a_us(w[j], kHad2, hc[4 * j + 2]);
        t[4 * j + 3] = dp4a_us(w[j], kHad3, hc[4 * j + 3]);
    }
    // vertical: |t0 + t1| = |t0 - (-t1)| and |t0 - t1| are one absolute-difference instruction each (the negation is a multiply-add
    // pipe move), then |p + q| + |p - q| = 2 max(|p|, |q|) replaces the second butterfly stage
    int acc = 0;
    for (int k = 0; k < 4; ++k) {
        const int n1 = -t[4 + k], n3 = -t[12 + k];
        acc += imax(absdiff(t[k], n1), absdiff(t[8 + k], n3)) + imax(absdiff(t[k], t[4 + k]), absdiff(t[8 + k], t[12 + k]));
    }
    return acc;
}
// Per-lane state that stays in registers across the 17 sub-sample candidates: the lane's 4x4 source block (as the negated
// horizontal transform of its rows), where the block sits in the windows / planes, the predictor, and the search centre of
// the lane's partition (every lane of a partition carries the same centre and best: no shared-memory round trip or leader
// election per candidate).  The emulation harness keeps one record per lane.
struct MeLane {
    int  P, px, py;
    int  wx, wy;      // window column / row of the block's sample displaced by (X, Y) full samples: (X + wx, Y + wy)
    int  base;        // the same for the planes themselves: offset (16*my + oy) * stride + 16*mx + ox
    int  hc[16];      // minus the horizontal Hadamard transform of the lane's four source rows
    int  cx, cy;      // search centre, quarter samples
    bool far;         // the partition's integer vector lies outside the windows: read the planes
};
GVC_HD void me_lane_init(int lane, const MeRefArgs &a, int mx, int my, const MeRefSmem &sm, const MeGeom &g, int ccx, int ccy, MeLane &L) {
    int ox, oy;
    me_lane_block(lane, L.P, ox, oy);
    L.px = sm.mvp[0]; L.py = sm.mvp[1];
    L.wx = ox + g.ax + kMeWinX - g.wbx; L.wy = oy + kMeWinY - g.wby;
    L.base = (16 * my + oy) * a.stride + 16 * mx + ox;
    for (int j = 0; j < 4; ++j) {
        const gvc_uint4 r = sm.cur[oy + j];
        const unsigned w = ox == 0 ? r.x : ox == 4 ? r.y : ox == 8 ? r.z : r.w;
        had_row_neg(w, L.hc + 4 * j);
    }
    L.cx = 4 * ccx; L.cy = 4 * ccy;
    L.far = !me_in_window(g, ccx, ccy);
}
GVC_HD unsigned me_win4(const unsigned *win, int col, int row) {     // four samples at window (col .. col+3, row), any col
    const unsigned *r = win + row * kMeWinPitchW + (col >> 2);
    return funnel_r(r[0], r[1], (col & 3) * 8);
}
GVC_HD const uint8_t *me_plane(const MeRefArgs &a, int p) { return p == 0 ? a.ref.y : p == 1 ? a.ref.hh : p == 2 ? a.ref.hv : a.ref.hc; }
// SATD share of the lane's 4x4 block for the vector (mvx, mvy), any quarter-sample position (mc_luma_src's rule: one
// plane, or the average of two), from the windows or -- far partitions -- from the planes.  Already halved.
GVC_HD int me_satd_at(const MeRefArgs &a, const MeRefSmem &sm, const MeLane &L, int mvx, int mvy) {
    const int fx = mvx & 3, fy = mvy & 3, q = fy * 4 + fx;
    const int p1 = (kQpelRef0 >> (2 * q)) & 3, p2 = (kQpelRef1 >> (2 * q)) & 3;
    const bool two = (q & 5) != 0;
    unsigned w[4];
    if (!L.far) {
        const int col = (mvx >> 2) + L.wx, row = (mvy >> 2) + L.wy;
        for (int j = 0; j < 4; ++j) w[j] = me_win4(sm.win[p1], col, row + (fy == 3 ? 1 : 0) + j);
        if (two) for (int j = 0; j < 4; ++j) w[j] = avg_u8x4(w[j], me_win4(sm.win[p2], col + (fx == 3 ? 1 : 0), row + j));
    } else {
        const int off = L.base + (mvy >> 2) * a.stride + (mvx >> 2);
        const uint8_t *s1 = me_plane(a, p1) + (off + (fy == 3 ? a.stride : 0)), *s2 = me_plane(a, p2) + (off + (fx == 3 ? 1 : 0));
        for (int j = 0; j < 4; ++j) w[j] = ld_u32_any(s1 + j * a.stride);
        if (two) for (int j = 0; j < 4; ++j) w[j] = avg_u8x4(w[j], ld_u32_any(s2 + j * a.stride));
    }
    return satd4x4_rows(L.hc, w[0], w[1], w[2], w[3]);
}
// The same for the candidate a descriptor names, relative to the lane's search centre: no fraction arithmetic, no plane table.
GVC_HD int me_satd_desc(const MeRefArgs &a, const MeRefSmem &sm, const MeLane &L, unsigned d) {
    const int p1 = (int)(d & 3u), p2 = (int)((d >> 7) & 3u);
    const int c1 = (int)((d >> 2) & 3u) - 1, r1 = (int)((d >> 4) & 3u) - 1, c2 = (int)((d >> 9) & 3u) - 1, r2 = (int)((d >> 11) & 3u) - 1;
    const bool two = (d & 64u) != 0;
    unsigned w[4];
    if (!L.far) {
        const int col = (L.cx >> 2) + L.wx, row = (L.cy >> 2) + L.wy;
        for (int j = 0; j < 4; ++j) w[j] = me_win4(sm.win[p1], col + c1, row + r1 + j);
        if (two) for (int j = 0; j < 4; ++j) w[j] = avg_u8x4(w[j], me_win4(sm.win[p2], col + c2, row + r2 + j));
    } else {
        const int off = L.base + (L.cy >> 2) * a.stride + (L.cx >> 2);
        const uint8_t *s1 = me_plane(a, p1) + (off + r1 * a.stride + c1), *s2 = me_plane(a, p2) + (off + r2 * a.stride + c2);
        for (int j = 0; j < 4; ++j) w[j] = ld_u32_any(s1 + j * a.stride);
        if (two) for (int j = 0; j < 4; ++j) w[j] = avg_u8x4(w[j], ld_u32_any(s2 + j * a.stride));
    }
    return satd4x4_rows(L.hc, w[0], w[1], w[2], w[3]);
}
// The nine half-sample candidates around a FULL-sample centre, v[ci + 1] for ci = -1 (the centre) .. 7 (sq8 order), with the
// window rows shared between the candidates that read the same plane: full plane 4 rows; V plane 5 rows for (0,-1),(0,+1);
// H plane 4 rows, two shifts, for (-1,0),(+1,0); C plane 5 rows, two shifts, for the four diagonal ones.
GVC_HD void me_half_all(const MeRefArgs &a, const MeRefSmem &sm, const MeLane &L, int v[9]) {
    if (L.far) {
        for (int ci = -1; ci < 8; ++ci) {
            int dx = 0, dy = 0;
            if (ci >= 0) sq8(ci, dx, dy);
            v[ci + 1] = me_satd_at(a, sm, L, L.cx + 2 * dx, L.cy + 2 * dy);
        }
        return;
    }
    constexpr int W = kMeWinPitchW;
    const int col = (L.cx >> 2) + L.wx, row = (L.cy >> 2) + L.wy;
    {
        const unsigned *r = sm.win[0] + row * W + (col >> 2);
        const int sh = (col & 3) * 8;
        v[0] = satd4x4_rows(L.hc, funnel_r(r[0], r[1], sh), funnel_r(r[W], r[W + 1], sh), funnel_r(r[2 * W], r[2 * W + 1], sh), funnel_r(r[3 * W], r[3 * W + 1], sh));
    }
    {
        const unsigned *r = sm.win[2] + (row - 1) * W + (col >> 2);
        const int sh = (col & 3) * 8;
        unsigned w[5];
        for (int j = 0; j < 5; ++j) w[j] = funnel_r(r[W * j], r[W * j + 1], sh);
        v[1] = satd4x4_rows(L.hc, w[0], w[1], w[2], w[3]);      // (0,-1)
        v[2] = satd4x4_rows(L.hc, w[1], w[2], w[3], w[4]);      // (0,+1)
    }
    const int cm = col - 1, shm = (cm & 3) * 8;                 // five samples from column col-1 sit in two words
    {
        const unsigned *r = sm.win[1] + row * W + (cm >> 2);
        unsigned m[4], p[4];
        for (int j = 0; j < 4; ++j) { const unsigned lo = r[W * j], hi = r[W * j + 1]; m[j] = funnel_r(lo, hi, shm); p[j] = funnel_rc(lo, hi, shm + 8); }
        v[3] = satd4x4_rows(L.hc, m[0], m[1], m[2], m[3]);      // (-1,0)
        v[4] = satd4x4_rows(L.hc, p[0], p[1], p[2], p[3]);      // (+1,0)
    }
    {
        const unsigned *r = sm.win[3] + (row - 1) * W + (cm >> 2);
        unsigned m[5], p[5];
        for (int j = 0; j < 5; ++j) { const unsigned lo = r[W * j], hi = r[W * j + 1]; m[j] = funnel_r(lo, hi, shm); p[j] = funnel_rc(lo, hi, shm + 8); }
        v[5] = satd4x4_rows(L.hc, m[0], m[1], m[2], m[3]);      // (-1,-1)
        v[6] = satd4x4_rows(L.hc, m[1], m[2], m[3], m[4]);      // (-1,+1)
        v[7] = satd4x4_rows(L.hc, p[0], p[1], p[2], p[3]);      // (+1,-1)
        v[8] = satd4x4_rows(L.hc, p[1], p[2], p[3], p[4]);      // (+1,+1)
    }
}
// the eight quarter-sample candidates around the centre (sq8 order)
GVC_HD void me_quarter_all(const MeRefArgs &a, const MeRefSmem &sm, const MeLane &L, int v[8]) {
    for (int ci = 0; ci < 8; ++ci) v[ci] = me_satd_desc(a, sm, L, me_desc(L.cx, L.cy, ci, 1));
}
// Best candidate so far of a partition as key = cost << 4 | order (order = position in the sequential candidate list, 0 = the
// one carried over from the previous step: the sequential search replaces its best only by a strictly lower cost).
struct MeBest { unsigned key; int x, y; };
// lambda * se(v) bits of the three x and three y components a step's candidates use: [0] centre - step, [1] centre, [2] centre + step
struct MeStepCost { int x[3], y[3]; };
GVC_HD MeStepCost me_step_cost(const MeRefArgs &a, const MeLane &L, int step) {
    MeStepCost c;
    for (int i = 0; i < 3; ++i) {
        c.x[i] = a.lambda * se_bits(L.cx + (i - 1) * step - L.px);
        c.y[i] = a.lambda * se_bits(L.cy + (i - 1) * step - L.py);
    }
    return c;
}
GVC_HD void me_take(const MeLane &L, const MeStepCost &c, int total, int ci, int step, MeBest &b) {
    int dx = 0, dy = 0;
    if (ci >= 0) sq8(ci, dx, dy);
    const int bits = (dx < 0 ? c.x[0] : dx > 0 ? c.x[2] : c.x[1]) + (dy < 0 ? c.y[0] : dy > 0 ? c.y[2] : c.y[1]);
    const unsigned k = ((unsigned)(total + bits) << 4) | (unsigned)(ci + 1);
    if (k < b.key) { b.key = k; b.x = L.cx + dx * step; b.y = L.cy + dy * step; }
}
// the same with the direction taken from the candidate's descriptor (the one-partition path: the candidate differs per half-warp)
GVC_HD void me_take_desc(const MeLane &L, const MeStepCost &c, int total, unsigned d, int ci, int step, MeBest &b) {
    const int dx = (int)((d >> 13) & 3u) - 1, dy = (int)((d >> 15) & 3u) - 1;
    const int bits = (dx < 0 ? c.x[0] : dx > 0 ? c.x[2] : c.x[1]) + (dy < 0 ? c.y[0] : dy > 0 ? c.y[2] : c.y[1]);
    const unsigned k = ((unsigned)(total + bits) << 4) | (unsigned)(ci + 1);
    if (k < b.key) { b.key = k; b.x = L.cx + dx * step; b.y = L.cy + dy * step; }
}
GVC_HD void me_next_step(MeLane &L, MeBest &b) { b.key &= ~15u; L.cx = b.x; L.cy = b.y; }   // the best becomes the centre, order 0
// first lane of the partition the lane belongs to, and the number of lanes in it
GVC_HD void me_group(int lane, int &first, int &count) {
    if (lane < 16) { first = 0; count = 16; } else { first = lane & ~3; count = 4; }
}
#if !defined(GVC_EMU)
GVC_HD int me_group_sum(int lane, int v) {     // sum over the lanes of the partition: 16-lane or 4-lane butterfly
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    const int t4 = __shfl_xor_sync(0xffffffffu, v, 4);
    if (lane < 16) v += t4;
    const int t8 = __shfl_xor_sync(0xffffffffu, v, 8);
    if (lane < 16) v += t8;
    return v;
}
#endif
// ---- one-partition path --------------------------------------------------------------------------------
// Part of the search specification (oracle/h264_mb.c): when the four 8x8 partitions pick the same integer vector as the
// 16x16 partition, the macroblock is coded as one 16x16 partition and the 8x8 refinements are not run (on the test
// clips: bytes within 0.01 %, PSNR unchanged).  The warp then puts both half-warps on the 16x16 partition, two sub-sample
// candidates per round: 5 + 4 rounds instead of 9 + 8.
// candidate of the lane's half-warp in round r of `step` (step 2: -1, 0 | 1, 2 | ... | 7, none; step 1: 0, 1 | ... | 6, 7); 8 = none
GVC_HD int me_fast_cand(int lane, int r, int step) { return step == 2 ? 2 * r - 1 + (lane >> 4) : 2 * r + (lane >> 4); }
GVC_HD int me_fast_satd(const MeRefArgs &a, const MeRefSmem &sm, const MeLane &L, int ci, int step) {
    return me_satd_desc(a, sm, L, me_desc(L.cx, L.cy, ci, step));
}
GVC_HD void me_write_out(const MeRefArgs &a, int mx, int my, int mode, int cost, const int mvx[4], const int mvy[4]) {
    MbMotion m;
    m.mode = (uint8_t)mode;
    m.cost = cost;
    for (int k = 0; k < 4; ++k) { m.mv[k][0] = (int16_t)mvx[k]; m.mv[k][1] = (int16_t)mvy[k]; }
    m.pad[0] = m.pad[1] = m.pad[2] = 0;
    a.out[my * a.mbw + mx] = m;
    if (a.out_host) a.out_host[my * a.mbw + mx] = m;
}
GVC_HD void me_fast_out(int lane, const MeRefArgs &a, int mx, int my, const MeBest &b) {
    if (lane != 0) return;
    const int x[4] = {b.x, b.x, b.x, b.x}, y[4] = {b.y, b.y, b.y, b.y};
    me_write_out(a, mx, my, 0, (int)(b.key >> 4), x, y);
}
// lane 0 with the bests of the five partitions (its own = the 16x16's, and those of lanes 16, 20, 24, 28)
GVC_HD void me_ref_out(int lane, const MeRefArgs &a, int mx, int my, const MeBest b[5]) {
    if (lane != 0) return;
    const int j16 = (int)(b[0].key >> 4);
    const int j8  = (int)(b[1].key >> 4) + (int)(b[2].key >> 4) + (int)(b[3].key >> 4) + (int)(b[4].key >> 4) + 8 * a.lambda;
    const int mode = j8 < j16 ? 1 : 0;
    int x[4], y[4];
    for (int k = 0; k < 4; ++k) { x[k] = mode ? b[1 + k].x : b[0].x; y[k] = mode ? b[1 + k].y : b[0].y; }
    me_write_out(a, mx, my, mode, mode ? j8 : j16, x, y);
}

}  // namespace gvc
