// resize_plan.cpp -- see resize_plan.h.
#include "resize_plan.h"

#include <algorithm>

namespace gvc {
namespace {

inline int32_t pack_pair(int lo, int hi) {
    return int32_t(uint32_t(uint16_t(int16_t(lo))) | (uint32_t(uint16_t(int16_t(hi))) << 16));
}

// Largest tile height whose every tile needs <= kTileRows source rows (0 if none).
int pick_tile_height(const FilterBank &vb, int dh) {
    for (int th = std::min(dh, 128); th >= 1; --th) {
        bool ok = true;
        for (int y0 = 0; y0 < dh && ok; y0 += th) {
            const int y1 = std::min(y0 + th, dh) - 1;
            ok = vb.pos[y1] + vb.taps - (vb.pos[y0] & ~1) <= kTileRows;      // a tile starts on an even source row (rows are stored in pairs)
        }
        if (ok) return th;
    }
    return 0;
}

bool monotone(const std::vector<int32_t> &v) {
    for (size_t i = 1; i < v.size(); ++i)
        if (v[i] < v[i - 1]) return false;
    return true;
}

}  // namespace

bool build_resize_tables(ResizeTables &t, int sw, int sh, int dw, int dh, int algo) {
    t.sw = sw; t.sh = sh; t.dw = dw; t.dh = dh; t.algo = algo;
    for (int c = 0; c < 2; ++c) {
        const int psw = c ? chroma_dim(sw) : sw, psh = c ? chroma_dim(sh) : sh;
        const int pdw = c ? chroma_dim(dw) : dw, pdh = c ? chroma_dim(dh) : dh;
        if (!build_filter_bank(t.hb[c], psw, pdw, algo, 1 << 14)) return false;
        if (!build_filter_bank(t.vb[c], psh, pdh, algo, 1 << 12)) return false;
        GenericClassDev &g = t.gcls[c];
        g.sw = psw; g.sh = psh; g.dw = pdw; g.dh = pdh;
        g.htaps = t.hb[c].taps; g.vtaps = t.vb[c].taps;
        g.hc = g.vc = nullptr; g.hp = g.vp = nullptr;
    }
    // ---- fused-kernel eligibility
    int  ng = 1, vq = 1, th[2] = {0, 0};
    bool ok = true;
    for (int c = 0; c < 2; ++c) {
        ng = std::max(ng, (t.hb[c].taps + 3) / 4);
        vq = std::max(vq, (t.vb[c].taps + 1 + 3) / 4);
        ok = ok && monotone(t.hb[c].pos) && monotone(t.vb[c].pos);
        th[c] = pick_tile_height(t.vb[c], t.gcls[c].dh);
        ok = ok && th[c] >= std::min(8, t.gcls[c].dh);
    }
    const int vtaps = std::max(t.vb[0].taps, t.vb[1].taps);
    ok = ok && ng <= kMaxHGroups && vtaps <= kMaxVTaps;
    int vt = (vtaps + 2) / 2;          // row pairs: taps + 1 rows when the first row is odd
    vq = kVEntry / 4;
    t.fused_ok = ok;
    if (!ok) return true;
    t.ng = ng; t.vq = vq; t.vt = vt;
    t.nt = (std::max(t.hb[0].taps, t.hb[1].taps) + 1) / 2;
    int s_chunks_max = 1;
    for (int c = 0; c < 2; ++c) {
        PlaneClassDev    &d  = t.cls[c];
        const FilterBank &hb = t.hb[c], &vb = t.vb[c];
        d.sw = t.gcls[c].sw; d.sh = t.gcls[c].sh; d.dw = t.gcls[c].dw; d.dh = t.gcls[c].dh;
        d.hgroups = ng; d.vtaps = vb.taps; d.vq = vq; d.th = th[c];
        d.tiles_x = (d.dw + kTileW - 1) / kTileW;
        d.tiles_y = (d.dh + d.th - 1) / d.th;
        d.tiles_x_magic = (1 << 20) / d.tiles_x + 1;
        for (long long i = 0; i < (long long)d.tiles_x * d.tiles_y && d.tiles_x_magic; ++i)
            if (i * d.tiles_x_magic >= (1ll << 31) || int((i * d.tiles_x_magic) >> 20) != int(i / d.tiles_x)) d.tiles_x_magic = 0;
        d.htab = nullptr; d.vtab = nullptr;
        t.htab[c].assign(size_t(d.dw) * kHEntry, 0);
        t.vtab[c].assign(size_t(d.dh) * kVEntry, 0);
        for (int x = 0; x < d.dw; ++x) {
            int32_t *e = &t.htab[c][size_t(x) * kHEntry];
            e[0] = hb.pos[x] >> 2;            // aligned source word holding the first tap
            e[1] = (hb.pos[x] & 3) * 8;       // bit offset of the first tap inside that word
            for (int j = 0; j < 4 * ng; j += 2) {
                const int a = j < hb.taps ? hb.coef[size_t(x) * hb.taps + j] : 0;
                const int b = j + 1 < hb.taps ? hb.coef[size_t(x) * hb.taps + j + 1] : 0;
                e[2 + j / 2] = pack_pair(a, b);
            }
        }
        for (int y = 0; y < d.dh; ++y) {
            // coefficient c = 256 * hi + lo (lo = c & 255 unsigned, hi = c >> 8 signed): sum t * c = sum t * lo + 256 * sum t * hi, each
            // sum two rows per dp2a on (row, row + 1) pairs of 16-bit intermediates
            int32_t *e = &t.vtab[c][size_t(y) * kVEntry];
            const int first = vb.pos[y] & ~1;
            uint32_t lo[4] = {0, 0, 0, 0}, hi[4] = {0, 0, 0, 0};
            for (int j = 0; j < vb.taps; ++j) {
                const int cf = vb.coef[size_t(y) * vb.taps + j], slot = vb.pos[y] - first + j;
                lo[slot >> 2] |= uint32_t(cf & 255) << (8 * (slot & 3));
                hi[slot >> 2] |= uint32_t((cf >> 8) & 255) << (8 * (slot & 3));
            }
            e[0] = first; e[1] = int32_t(lo[0]); e[2] = int32_t(hi[0]); e[3] = int32_t(lo[1]);
            e[4] = int32_t(hi[1]); e[5] = int32_t(lo[2]); e[6] = int32_t(hi[2]); e[7] = int32_t(lo[3]);
            e[8] = int32_t(hi[3]); e[9] = vb.pos[y] + vb.taps;
        }
        int chunks = 1;   // widest tile in 16-byte chunks, +1 look-ahead chunk for the sliding window
        for (int x0 = 0; x0 < d.dw; x0 += kTileW) {
            const int x1  = std::min(x0 + kTileW, d.dw) - 1;
            const int qa0 = (hb.pos[x0] >> 2) & ~3, ql = (hb.pos[x1] >> 2) + ng;
            chunks = std::max(chunks, ((ql - qa0) >> 2) + 1 + 1);
        }
        d.s_chunks   = chunks | 1;   // odd pitch: lanes <-> rows reads hit distinct banks
        d.s_chunks_magic = int(unsigned((0x100000000ull + d.s_chunks - 1) / d.s_chunks));
        d.vt = vt;
        s_chunks_max = std::max(s_chunks_max, d.s_chunks);
    }
    t.s_chunks_max = s_chunks_max;
    t.smem_bytes   = size_t(kTileRows) * s_chunks_max * 16 + (size_t(kTileW / 4) * kTPad + kTSlack) * 16;
    if (t.smem_bytes > 200 * 1024) t.fused_ok = false;
    return true;
}

}  // namespace gvc
