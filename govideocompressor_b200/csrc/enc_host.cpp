// enc_host.cpp -- see enc_host.h.
#include "enc_host.h"

namespace gvc {
namespace {
const int kQuantScale[6][3]   = {{13107, 8066, 5243}, {11916, 7490, 4660}, {10082, 6554, 4194},
                                 {9362, 5825, 3647},  {8192, 5243, 3355},  {7282, 4559, 2893}};
const int kDequantScale[6][3] = {{10, 13, 16}, {11, 14, 18}, {13, 16, 20}, {14, 18, 23}, {16, 20, 25}, {18, 23, 29}};
const unsigned char kChromaQp[22] = {29, 30, 31, 32, 32, 33, 34, 34, 35, 35, 36, 36, 37, 37, 37, 38, 38, 38, 39, 39, 39, 39};
const unsigned char kLambda[52]   = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 4, 4, 4,
                                     5, 6, 6, 7, 8, 9, 10, 11, 13, 14, 16, 18, 20, 23, 25, 29, 32, 36, 40, 45, 51, 57, 64, 72, 81, 91};

const int kQuant8Scale[6][6]   = {{13107, 11428, 20972, 12222, 16777, 15481}, {11916, 10826, 19174, 11058, 14980, 14290},
                                  {10082, 8943, 15978, 9675, 12710, 11985},   {9362, 8228, 14913, 8931, 11984, 11259},
                                  {8192, 7346, 13159, 7740, 10486, 9777},     {7282, 6428, 11570, 6830, 9118, 8640}};
const int kDequant8Scale[6][6] = {{20, 18, 32, 19, 25, 24}, {22, 19, 35, 21, 28, 26}, {26, 23, 42, 24, 33, 31},
                                  {28, 25, 45, 26, 35, 33}, {32, 28, 51, 30, 40, 38}, {36, 32, 58, 34, 46, 43}};

void one_qp(int qp, int deadzone, int mf[3], int bias[3], int dq[3]) {
    for (int c = 0; c < 3; ++c) {
        const int s = qp / 6 - 1, v = kQuantScale[qp % 6][c];
        const int m = s <= 0 ? v << -s : (v + (1 << (s - 1))) >> s;
        int       b = ((deadzone << 10) + (m >> 1)) / m;
        if (b > (1 << 15) / m) b = (1 << 15) / m;
        mf[c] = m; bias[c] = b; dq[c] = kDequantScale[qp % 6][c] * 16;
    }
}
}  // namespace

int chroma_qp(int qp) { qp = qp < 0 ? 0 : qp > 51 ? 51 : qp; return qp < 30 ? qp : kChromaQp[qp - 30]; }
int lambda_for_qp(int qp) { return kLambda[qp < 0 ? 0 : qp > 51 ? 51 : qp]; }

void fill_quant_params(QuantParams &q, int qp, bool intra) {
    qp = qp < 0 ? 0 : qp > 51 ? 51 : qp;
    q.qp = qp; q.qpc = chroma_qp(qp);
    const int dz = intra ? 21 : 11;
    one_qp(q.qp, dz, q.mf_y, q.bias_y, q.dq_y);
    one_qp(q.qpc, dz, q.mf_c, q.bias_c, q.dq_c);
    q.lambda = lambda_for_qp(qp);
    for (int c = 0; c < 3; ++c) {
        q.thr_y[c] = (65536 + q.mf_y[c] - 1) / q.mf_y[c] - q.bias_y[c];
        q.thr_c[c] = (65536 + q.mf_c[c] - 1) / q.mf_c[c] - q.bias_c[c];
    }
    const int mdc = q.mf_c[0] >> 1;
    q.thr_cdc = (65536 + mdc - 1) / mdc - (q.bias_c[0] << 1);
}

void fill_quant8_params(Quant8Params &q, int qp, bool intra) {
    qp = qp < 0 ? 0 : qp > 51 ? 51 : qp;
    q.qp = qp;
    const int dz = intra ? 21 : 11;
    for (int c = 0; c < 6; ++c) {
        const int s = qp / 6, v = kQuant8Scale[qp % 6][c];
        const int m = s <= 0 ? v : (v + (1 << (s - 1))) >> s;      // x264: SHIFT(quant8_scale, qp/6)
        int       b = ((dz << 10) + (m >> 1)) / m;
        if (b > (1 << 15) / m) b = (1 << 15) / m;
        q.mf[c] = m; q.bias[c] = b; q.dq[c] = kDequant8Scale[qp % 6][c] * 16;
    }
}

EncGeom make_geom(int w, int h) {
    EncGeom g;
    g.w = w; g.h = h;
    g.w16 = align_up(w, 16); g.h16 = align_up(h, 16);
    g.mbw = g.w16 / 16; g.mbh = g.h16 / 16;
    g.sy = align_up(g.w16 + 2 * kEncPad, 128);
    g.sc = align_up(g.w16 / 2 + kEncPad, 64);
    g.sl = align_up(g.w16 / 2 + kEncPad, 64);
    return g;
}

}  // namespace gvc
