// emu_enc.cpp -- TEST HARNESS: executes the encoder kernels' bodies (enc_kernels.cuh built with GVC_EMU) lane by
// lane / thread by thread on the host, phase by phase, so the CPU suite can compare their logic with the oracle.
// Not part of libgvc.so.
#define GVC_EMU 1
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../govideocompressor_b200/csrc/enc_host.h"
#include "../../govideocompressor_b200/csrc/enc_kernels.cuh"

using namespace gvc;

void emu_run_hpel(const HpelArgs &a) {
    for (int y0 = a.y_lo; y0 < a.y_hi; y0 += kHpTileH)
        for (int x0 = a.x_lo; x0 < a.x_hi; x0 += kHpTileW) {
            static HpelSmem sm;
            memset(&sm, 0xAB, sizeof(sm));
            const bool al = hpel_aligned(a);
            for (int t = 0; t < kHpThreads; ++t) hpel_phase_a(t, a, al, x0, y0, sm);
            for (int t = 0; t < kHpThreads; ++t) hpel_phase_b(t, a, al, x0, y0, sm);
            for (int t = 0; t < kHpThreads; ++t) hpel_phase_c(t, a, al, x0, y0, sm);
        }
}
void emu_run_me_lowres(const MeLowArgs &a) {
    for (int my = 0; my < a.mbh; ++my)
        for (int mx = 0; mx < a.mbw; ++mx) {
            unsigned sad0 = 0;
            for (int l = 0; l < 32; ++l) sad0 += me_lowres_zero_sad(l, a, mx, my);
            if (me_lowres_zero_wins(a, sad0)) {
                for (int l = 0; l < 32; ++l) me_lowres_out(l, a, mx, my, unsigned(a.r * (2 * a.r + 1) + a.r));
                continue;
            }
            MeLowSmem sm;
            memset(&sm, 0xAB, sizeof(sm));
            for (int l = 0; l < 32; ++l) me_lowres_load(l, a, mx, my, sm);
            unsigned best = 0xffffffffu;
            for (int l = 0; l < 32; ++l) { const unsigned k = me_lowres_chunk(a.r) == kLowChunk ? me_lowres_lane<kLowChunk>(l, a, mx, my, sm) : me_lowres_lane<kLowChunkWide>(l, a, mx, my, sm); best = k < best ? k : best; }
            for (int l = 0; l < 32; ++l) me_lowres_out(l, a, mx, my, best);
        }
}
void emu_run_me_refine(const MeRefArgs &a) {
    for (int my = 0; my < a.mbh; ++my)
        for (int mx = 0; mx < a.mbw; ++mx) {
            MeRefSmem sm;
            memset(&sm, 0xAB, sizeof(sm));
            for (int l = 0; l < 32; ++l) me_ref_load(l, a, mx, my, sm);
            for (int l = 0; l < 32; ++l) me_mvp(sm, l);
            unsigned zq[4] = {0, 0, 0, 0};
            for (int l = 0; l < 32; ++l) zq[me_zero_quadrant(l)] += me_ref_zero_part(l, a, mx, my, sm);
            const int zero_sad = int(zq[0] + zq[1] + zq[2] + zq[3]);
            if (me_ref_early(a, zero_sad)) {
                for (int l = 0; l < 32; ++l) me_ref_early_out(l, a, mx, my, sm, zero_sad);
                continue;
            }
            const MeGeom g = me_geom(a, sm, mx, my);
            for (int l = 0; l < 32; ++l) me_stage_plain(l, a, g, sm);
            // integer candidates: the warp minimum of each partition's keys, the vector of the winning lane
            unsigned best[5] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
            for (int l = 0; l < 32; ++l) {
                unsigned key[5];
                me_ref_int(l, a, sm, g, zq, key);
                for (int k = 0; k < 5; ++k) best[k] = key[k] < best[k] ? key[k] : best[k];
            }
            int ccx[5], ccy[5];
            for (int k = 0; k < 5; ++k) me_int_cand(sm, g, int(best[k] & 31u), ccx[k], ccy[k]);
            bool one = true;
            for (int k = 1; k < 5; ++k) one = one && ccx[k] == ccx[0] && ccy[k] == ccy[0];
            MeLane L[32];
            MeBest b[32];
            if (one) {                 // the warp's one-partition path, shuffles spelled out
                for (int l = 0; l < 32; ++l) { me_lane_init(l & 15, a, mx, my, sm, g, ccx[0], ccy[0], L[l]); b[l].key = 0xffffffffu; b[l].x = L[l].cx; b[l].y = L[l].cy; }
                for (int step = 2; step >= 1; --step) {
                    MeStepCost sc[32];
                    for (int l = 0; l < 32; ++l) sc[l] = me_step_cost(a, L[l], step);
                    for (int r = 0; r < (step == 2 ? 5 : 4); ++r) {
                        int part[32];
                        for (int l = 0; l < 32; ++l) { const int ci = me_fast_cand(l, r, step); part[l] = ci < 8 ? me_fast_satd(a, sm, L[l], ci, step) : 0; }
                        for (int l = 0; l < 32; ++l) {
                            int total = 0;
                            for (int i = 0; i < 16; ++i) total += part[(l & 16) + i];
                            const int ci = me_fast_cand(l, r, step);
                            if (ci < 8) me_take_desc(L[l], sc[l], total, me_desc(L[l].cx, L[l].cy, ci, step), ci, step, b[l]);
                        }
                    }
                    for (int l = 0; l < 16; ++l) {
                        const MeBest w = b[l + 16].key < b[l].key ? b[l + 16] : b[l];
                        b[l] = b[l + 16] = w;
                    }
                    for (int l = 0; l < 32; ++l) me_next_step(L[l], b[l]);
                }
                for (int l = 0; l < 32; ++l) me_fast_out(l, a, mx, my, b[l]);
                continue;
            }
            for (int l = 0; l < 32; ++l) {
                int P, ox, oy;
                me_lane_block(l, P, ox, oy);
                me_lane_init(l, a, mx, my, sm, g, ccx[P], ccy[P], L[l]);
                b[l].key = 0xffffffffu; b[l].x = L[l].cx; b[l].y = L[l].cy;
            }
            for (int step = 2; step >= 1; --step) {
                int v[32][9];
                for (int l = 0; l < 32; ++l) {
                    if (step == 2) me_half_all(a, sm, L[l], v[l]);
                    else           me_quarter_all(a, sm, L[l], v[l] + 1);
                }
                MeStepCost sc[32];
                for (int l = 0; l < 32; ++l) sc[l] = me_step_cost(a, L[l], step);
                for (int ci = (step == 2 ? -1 : 0); ci < 8; ++ci)
                    for (int l = 0; l < 32; ++l) {          // what the shuffle butterfly computes on the GPU
                        int first, count, total = 0;
                        me_group(l, first, count);
                        for (int i = 0; i < count; ++i) total += v[first + i][ci + 1];
                        me_take(L[l], sc[l], total, ci, step, b[l]);
                    }
                if (step == 2) for (int l = 0; l < 32; ++l) me_next_step(L[l], b[l]);
            }
            const MeBest all[5] = {b[0], b[16], b[20], b[24], b[28]};
            for (int l = 0; l < 32; ++l) me_ref_out(l, a, mx, my, all);
        }
}
void emu_run_encode_inter(const EncMbArgs &a) {
    for (int my = 0; my < a.mbh; ++my)
        for (int mx = 0; mx < a.mbw; ++mx) {
            EncMbSmem sm;
            memset(&sm, 0xAB, sizeof(sm));
            const bool al = enc_mb_aligned(a);
            EncLane L[32];
            memset(L, 0xAB, sizeof(L));
            for (int l = 0; l < 32; ++l) enc_mb_p0(l, a, mx, my, sm);
            for (int l = 0; l < 32; ++l) enc_mb_p1(l, a, al, mx, my, sm, L[l]);
            bool any = false;
            for (int l = 0; l < 32; ++l) any = enc_mb_any_level(l, a, sm, L[l]) || any;
            for (int l = 0; l < 32; ++l) { if (any) enc_mb_p2(l, a, al, mx, my, sm, L[l]); else enc_mb_p2_zero(l, a, al, mx, my, sm, L[l]); }
            for (int l = 0; l < 32; ++l) enc_mb_p3(l, a, mx, my, sm);
            for (int l = 0; l < 32; ++l) enc_mb_p4(l, a, mx, my, sm, L[l]);
        }
}
void emu_run_encode_intra(const IntraArgs &a) {
    const bool al = intra_aligned(a);
    for (int my = 0; my < a.mbh; ++my) {        // rows in order: every wavefront dependency is already satisfied
        IntraSmem sm;                           // lives for the whole row: carries the left neighbours
        int row_count = 0;
        memset(&sm, 0xAB, sizeof(sm));
        for (int mx = 0; mx < a.mbw; ++mx) {
            IntraLane L[32];
            memset(L, 0xAB, sizeof(L));
            for (int l = 0; l < 32; ++l) intra_load_cur(l, a, al, mx, my, L[l].cur);
            for (int l = 0; l < 32; ++l) intra_p0(l, a, mx, my, sm);
            for (int l = 0; l < 32; ++l) intra_p1(l, a, mx, my, sm, L[l]);
            for (int l = 0; l < 32; ++l) intra_p2(l, a, mx, my, sm, L[l]);
            for (int l = 0; l < 32; ++l) intra_p3a(l, sm);
            for (int l = 0; l < 32; ++l) intra_p3b(l, a, sm);
            for (int l = 0; l < 32; ++l) intra_p3c(l, sm);
            for (int l = 0; l < 32; ++l) intra_p3d(l, a, al, mx, my, sm, L[l]);
            for (int l = 0; l < 32; ++l) intra_p4(l, a, mx, my, sm, row_count);
            for (int l = 0; l < 32; ++l) intra_p5(l, a, mx, my, sm, L[l]);
        }
        intra_row_done(0, a, row_count);
    }
}

extern "C" {

void emu_extend_border(uint8_t *p, int stride, int w, int h, int x_lo, int x_hi, int y_lo, int y_hi) {
    for (int y = y_lo; y < y_hi; ++y)
        for (int x = x_lo; x < x_hi; ++x) extend_border_px(p, stride, w, h, x, y);
}

void emu_lowres(const uint8_t *src, int sstride, int x_lo, int x_hi, int y_lo, int y_hi, uint8_t *d0, uint8_t *dh, uint8_t *dv,
                uint8_t *dc, int dstride, int lx_lo, int lx_hi, int ly_lo, int ly_hi) {
    for (int ly = ly_lo; ly < ly_hi; ++ly)
        for (int lx = lx_lo; lx < lx_hi; ++lx) lowres_px(src, sstride, x_lo, x_hi - 1, y_lo, y_hi - 1, d0, dh, dv, dc, dstride, lx, ly);
}

void emu_hpel(const uint8_t *src, uint8_t *dh, uint8_t *dv, uint8_t *dc, int stride, int x_lo, int x_hi, int y_lo, int y_hi) {
    HpelArgs a;
    a.src = src; a.dh = dh; a.dv = dv; a.dc = dc; a.stride = stride; a.x_lo = x_lo; a.x_hi = x_hi; a.y_lo = y_lo; a.y_hi = y_hi;
    a.cx_lo = x_lo; a.cx_hi = x_hi; a.cy_lo = y_lo; a.cy_hi = y_hi;
    emu_run_hpel(a);
}

void emu_me_lowres(const uint8_t *cur, const uint8_t *ref, int stride, int mbw, int mbh, int r, int lambda, int16_t *out_mv) {
    MeLowArgs a;
    a.cur = cur; a.ref = ref; a.stride = stride; a.mbw = mbw; a.mbh = mbh; a.r = r; a.lambda = lambda; a.out_mv = out_mv;
    emu_run_me_lowres(a);
}

static RefPlanes mkref(const uint8_t *const p[7]) {
    RefPlanes r;
    r.y = p[0]; r.hh = p[1]; r.hv = p[2]; r.hc = p[3]; r.u = p[4]; r.v = p[5]; r.low = p[6];
    return r;
}

// out: MbMotion records (24 bytes each)
void emu_me_refine(const uint8_t *cur, const uint8_t *const ref[7], int stride, int w16, int h16, const int16_t *lowres_mv,
                   int lambda, void *out) {
    MeRefArgs a;
    a.cur = cur; a.ref = mkref(ref); a.stride = stride; a.w16 = w16; a.h16 = h16; a.mbw = w16 / 16; a.mbh = h16 / 16;
    a.lambda = lambda; a.lowres_mv = lowres_mv; a.out = static_cast<MbMotion *>(out); a.out_host = nullptr;
    emu_run_me_refine(a);
}

void emu_encode_inter(const uint8_t *const cur[3], const uint8_t *const ref[7], uint8_t *const rec[3], int sy, int sc, int w16,
                      int h16, const void *motion, int qp, int16_t *coefs, uint32_t *nzmask) {
    EncMbArgs a;
    a.cur_y = cur[0]; a.cur_u = cur[1]; a.cur_v = cur[2]; a.ref = mkref(ref);
    a.rec_y = rec[0]; a.rec_u = rec[1]; a.rec_v = rec[2]; a.sy = sy; a.sc = sc; a.mbw = w16 / 16; a.mbh = h16 / 16;
    fill_quant_params(a.q, qp, false);
    a.motion = static_cast<const MbMotion *>(motion); a.coefs = coefs; a.nzmask = nzmask;
    a.pack.blocks = nullptr; a.pack.offsets = nullptr; a.pack.counter = nullptr; a.nnz_y = a.nnz_u = a.nnz_v = nullptr;
    emu_run_encode_inter(a);
}

void emu_encode_intra(const uint8_t *const cur[3], uint8_t *const rec[3], int sy, int sc, int w16, int h16, int qp, int16_t *coefs,
                      uint32_t *nzmask, uint8_t *modes) {
    IntraArgs a;
    a.cur_y = cur[0]; a.cur_u = cur[1]; a.cur_v = cur[2]; a.rec_y = rec[0]; a.rec_u = rec[1]; a.rec_v = rec[2];
    a.sy = sy; a.sc = sc; a.mbw = w16 / 16; a.mbh = h16 / 16;
    fill_quant_params(a.q, qp, true);
    a.coefs = coefs; a.nzmask = nzmask; a.modes = modes;
    std::vector<unsigned long long> mail(size_t(8) * a.mbw * a.mbh, 0);
    a.mail = mail.data();
    a.pack.blocks = nullptr; a.pack.offsets = nullptr; a.pack.counter = nullptr; a.nnz_y = a.nnz_u = a.nnz_v = nullptr;
    emu_run_encode_intra(a);
}

void emu_transform8x8(const uint8_t *src, int sstride, const uint8_t *pred, int pstride, uint8_t *rec, int rstride, int w, int h, int qp,
                      int intra, int16_t *coefs, uint8_t *nz) {
    Transform8Args a;
    a.src = src; a.pred = pred; a.rec = rec; a.sstride = sstride; a.pstride = pstride; a.rstride = rstride; a.bw = w / 8; a.bh = h / 8;
    fill_quant8_params(a.q, qp, intra != 0);
    a.coefs = coefs; a.nz = nz;
    for (int by = 0; by < a.bh; ++by)
        for (int bx = 0; bx < a.bw; ++bx) transform8x8_block(a, bx, by);
}

// quantiser constants as the product computes them: out[0..2]=mf_y, [3..5]=bias_y, [6..8]=dq_y, [9..11]=mf_c, [12..14]=bias_c,
// [15..17]=dq_c, [18]=qpc, [19]=lambda
void emu_quant_params(int qp, int intra, int out[20]) {
    QuantParams q;
    fill_quant_params(q, qp, intra != 0);
    for (int i = 0; i < 3; ++i) {
        out[i] = q.mf_y[i]; out[3 + i] = q.bias_y[i]; out[6 + i] = q.dq_y[i];
        out[9 + i] = q.mf_c[i]; out[12 + i] = q.bias_c[i]; out[15 + i] = q.dq_c[i];
    }
    out[18] = q.qpc; out[19] = q.lambda;
}

void emu_geom(int w, int h, int out[9]) {
    const EncGeom g = make_geom(w, h);
    out[0] = g.w16; out[1] = g.h16; out[2] = g.mbw; out[3] = g.mbh; out[4] = g.sy; out[5] = g.sc; out[6] = g.sl; out[7] = kEncPad; out[8] = kMbCoefs;
}
}
