// emu_resize.cpp -- TEST HARNESS: runs the resize kernels' phase functions (resize_kernels.cuh, built
// with GVC_EMU) serially on the host so the CPU test-suite can check tile geometry, table packing and
// indexing against the oracle without a GPU.  Not part of libgvc.so; never used by the product path.
#define GVC_EMU 1
#include <cstring>
#include <vector>

#include "../../govideocompressor_b200/csrc/resize_kernels.cuh"
#include "../../govideocompressor_b200/csrc/resize_plan.h"

using namespace gvc;

namespace {

template <int NG, int VT>
void run_tile(const ResizeParams &P, int tile, int frame, int s_chunks_max) {
    const int            j  = find_job(P, tile);
    const PlaneJob      &pj = P.job[j];
    const PlaneClassDev &pc = P.cls[pj.cls];
    const int            t  = tile - pj.tile_base;
    const int            ty = t / pc.tiles_x, tx = t - ty * pc.tiles_x;
    const TileGeom       g  = tile_geom(pc, tx, ty);
    std::vector<gvc_uint4> S(size_t(kTileRows) * s_chunks_max);
    std::vector<gvc_int4>  T(size_t(kTileW / 4) * kTPad + kTSlack);
    memset(S.data(), 0xAB, S.size() * sizeof(gvc_uint4));   // poison: stale smem must never matter
    memset(T.data(), 0xCD, T.size() * sizeof(gvc_int4));
    const uint8_t *src = pj.src + size_t(frame) * pj.sframe;
    uint8_t       *dst = pj.dst + size_t(frame) * pj.dframe;
    for (int tid = 0; tid < kThreads; ++tid) resize_phase0(tid, pc, pj, src, g, reinterpret_cast<unsigned *>(S.data()), 4 * s_chunks_max);
    for (int tid = 0; tid < kThreads; ++tid) resize_phase1<NG>(tid, pc, g, reinterpret_cast<const unsigned *>(S.data()), 4 * s_chunks_max, T.data());
    for (int tid = 0; tid < kThreads; ++tid) resize_phase2<VT>(tid, pc, pj, dst, g, T.data());
}

typedef void (*tile_fn)(const ResizeParams &, int, int, int);
template <int NG, int VT>
struct PickVt {
    static tile_fn get(int vt) { return vt >= VT ? static_cast<tile_fn>(run_tile<NG, VT>) : PickVt<NG, VT - 1>::get(vt); }
};
template <int NG>
struct PickVt<NG, 1> {
    static tile_fn get(int) { return run_tile<NG, 1>; }
};
tile_fn pick(int nt, int vt) {
    switch (nt) {
        case 1: return PickVt<1, kMaxVPairs>::get(vt);
        case 2: return PickVt<2, kMaxVPairs>::get(vt);
        case 3: return PickVt<3, kMaxVPairs>::get(vt);
        case 4: return PickVt<4, kMaxVPairs>::get(vt);
        case 5: return PickVt<5, kMaxVPairs>::get(vt);
        default: return PickVt<6, kMaxVPairs>::get(vt);
    }
}

}  // namespace

extern "C" {

// planes as in gvc_frames (host pointers).  fmt: 0 yuv420p, 23 nv12.  mode: 0 fused, 1 two-pass.
// returns 0 ok, 1 fused path not eligible, -1 error
int emu_resize(const uint8_t *const sp[3], const int sstride[3], const size_t sframe[3], uint8_t *const dp[3],
               const int dstride[3], const size_t dframe[3], int n_frames, int sw, int sh, int dw, int dh, int fmt,
               int algo, int mode) {
    ResizeTables t;
    if (!build_resize_tables(t, sw, sh, dw, dh, algo)) return -1;
    const bool nv12 = fmt == 23;
    if (mode == 0) {
        if (!t.fused_ok) return 1;
        ResizeParams P;
        memset(&P, 0, sizeof(P));
        for (int c = 0; c < 2; ++c) {
            P.cls[c]      = t.cls[c];
            P.cls[c].htab = t.htab[c].data();
            P.cls[c].vtab = t.vtab[c].data();
        }
        P.njobs = 3;
        const int lt = t.cls[0].tiles_x * t.cls[0].tiles_y, ct = t.cls[1].tiles_x * t.cls[1].tiles_y;
        for (int k = 0; k < 3; ++k) {
            PlaneJob &j  = P.job[k];
            const int sk = nv12 && k == 2 ? 1 : k;
            j.src = sp[sk] + (nv12 && k == 2 ? 1 : 0);
            j.dst = dp[sk] + (nv12 && k == 2 ? 1 : 0);
            j.sstride = sstride[sk]; j.dstride = dstride[sk];
            j.sframe = sframe[sk]; j.dframe = dframe[sk];
            j.spix = j.dpix = nv12 && k > 0 ? 2 : 1;
            j.cls = k == 0 ? 0 : 1;
            j.tile_base = k == 0 ? 0 : k == 1 ? lt : lt + ct;
        }
        P.tiles_per_frame = lt + 2 * ct;
        tile_fn fn = pick(t.nt, t.vt);
        for (int f = 0; f < n_frames; ++f)
            for (int tile = 0; tile < P.tiles_per_frame; ++tile) fn(P, tile, f, t.s_chunks_max);
        return 0;
    }
    for (int c = 0; c < 2; ++c) {
        t.gcls[c].hc = t.hb[c].coef.data(); t.gcls[c].hp = t.hb[c].pos.data();
        t.gcls[c].vc = t.vb[c].coef.data(); t.gcls[c].vp = t.vb[c].pos.data();
    }
    for (int f = 0; f < n_frames; ++f)
        for (int k = 0; k < 3; ++k) {
            const GenericClassDev &pc = t.gcls[k == 0 ? 0 : 1];
            const int sk = nv12 && k == 2 ? 1 : k, pix = nv12 && k > 0 ? 2 : 1;
            const uint8_t *s = sp[sk] + (nv12 && k == 2 ? 1 : 0) + size_t(f) * sframe[sk];
            uint8_t       *d = dp[sk] + (nv12 && k == 2 ? 1 : 0) + size_t(f) * dframe[sk];
            std::vector<int16_t> tmp(size_t(pc.sh) * pc.dw);
            for (int r = 0; r < pc.sh; ++r)
                for (int x = 0; x < pc.dw; ++x)
                    tmp[size_t(r) * pc.dw + x] = (int16_t)generic_h_sample(pc, s + size_t(r) * sstride[sk], pix, x);
            for (int y = 0; y < pc.dh; ++y)
                for (int x = 0; x < pc.dw; ++x)
                    d[size_t(y) * dstride[sk] + size_t(x) * pix] = (uint8_t)generic_v_sample(pc, tmp.data(), y, x);
        }
    return 0;
}

// product filter bank -> caller buffers (coef: dst_len*max_taps int16, pos: dst_len int32); returns taps or -1
int emu_filter_bank(int src_len, int dst_len, int algo, int one, int16_t *coef, int32_t *pos, int max_taps) {
    FilterBank b;
    if (!build_filter_bank(b, src_len, dst_len, algo, one)) return -1;
    if (b.taps > max_taps) return -2;
    memcpy(coef, b.coef.data(), b.coef.size() * sizeof(int16_t));
    memcpy(pos, b.pos.data(), b.pos.size() * sizeof(int32_t));
    return b.taps;
}

// plan facts: out[0]=fused_ok out[1]=ng out[2]=vq out[3]=th luma out[4]=th chroma out[5]=smem bytes
int emu_plan_info(int sw, int sh, int dw, int dh, int algo, int out[8]) {
    ResizeTables t;
    if (!build_resize_tables(t, sw, sh, dw, dh, algo)) return -1;
    out[0] = t.fused_ok; out[1] = t.ng; out[2] = t.vq;
    out[3] = t.fused_ok ? t.cls[0].th : 0; out[4] = t.fused_ok ? t.cls[1].th : 0;
    out[5] = (int)t.smem_bytes; out[6] = t.hb[0].taps; out[7] = t.vb[0].taps;
    return 0;
}
}
