"""CPU: the encoder kernels' bodies executed on the host (tests/emu) against the oracle -- border extension,
lowres, half-sample planes, lowres search, refinement + partition decision, inter MB encode/reconstruct."""
import ctypes

import numpy as np
import pytest

import emu_lib
import enc_util as E
import oracle_lib

BACKEND = "emu"


def _vp(x):
    return ctypes.c_void_p(x)


@pytest.fixture(scope="module", params=["pitch128", "odd-pitch"])
def pair(request):
    g = E.geom(96, 64)
    if request.param == "odd-pitch":      # pitches / addresses that are not multiples of four: the byte-wise load and store paths
        g = dict(g, sy=g["sy"] + 3, sc=g["sc"] + 1, sl=g["sl"] + 1)
    ref_yuv, cur_yuv = E.make_pair(g["w16"], g["h16"], seed=3)
    return g, E.Frame(g, ref_yuv), E.Frame(g, cur_yuv, with_hpel=False)


def test_quant_params_equal_oracle_tables():
    out = (ctypes.c_int * 20)()
    for qp in range(52):
        for intra in (0, 1):
            emu_lib.load().emu_quant_params(qp, intra, out)
            mf, bias = oracle_lib.quant4_tables(qp, intra)
            qpc = oracle_lib.load().gvc_oracle_chroma_qp(qp)
            mfc, biasc = oracle_lib.quant4_tables(qpc, intra)
            for cls, idx in ((0, 0), (1, 1), (2, 5)):
                assert out[cls] == mf[idx] and out[3 + cls] == bias[idx]
                assert out[9 + cls] == mfc[idx] and out[12 + cls] == biasc[idx]
                assert out[6 + cls] == oracle_lib.load().gvc_oracle_dequant4_scale(qp, idx)
                assert out[15 + cls] == oracle_lib.load().gvc_oracle_dequant4_scale(qpc, idx)
            assert out[18] == qpc and out[19] == oracle_lib.load().gvc_oracle_lambda(qp)


def test_border_extension(pair):
    g, ref, cur = pair
    w, h = 90, 60                      # picture smaller than the coded size: right/bottom strips get filled too
    pl = E.luma_plane(g, 0xEE)
    img = np.random.Generator(np.random.PCG64(5)).integers(0, 256, (h, w), dtype=np.uint8)
    pl.arr[E.PAD:E.PAD + h, E.PAD:E.PAD + w] = img
    want = np.pad(img, ((E.PAD, E.PAD + g["h16"] - h), (E.PAD, E.PAD + g["w16"] - w)), mode="edge")
    emu_lib.load().emu_extend_border(_vp(pl.ptr()), g["sy"], w, h, -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD)
    np.testing.assert_array_equal(pl.full(), want)


def test_lowres_all_phases(pair):
    g, ref, cur = pair
    want = E.oracle_lowres_full(ref.Y.full())
    outs = [E.low_plane(g, 0xEE) for _ in range(4)]
    lp = E.PAD // 2
    emu_lib.load().emu_lowres(_vp(ref.Y.ptr()), g["sy"], -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD,
                              *[_vp(o.ptr()) for o in outs], g["sl"], -lp, g["w16"] // 2 + lp, -lp, g["h16"] // 2 + lp)
    for o, wnt in zip(outs, want):
        np.testing.assert_array_equal(o.full(), wnt)


def test_hpel_planes(pair):
    g, ref, cur = pair
    outs = [E.luma_plane(g, 0xEE) for _ in range(3)]
    emu_lib.load().emu_hpel(_vp(ref.Y.ptr()), *[_vp(o.ptr()) for o in outs], g["sy"], -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD)
    for o, wnt in zip(outs, (ref.H, ref.Vp, ref.C)):
        np.testing.assert_array_equal(o.full(), wnt.full())


@pytest.mark.parametrize("w,h", [(131, 77), (132, 70), (64, 32), (5, 3), (200, 33)])
def test_hpel_any_size_and_alignment(w, h):
    """Odd widths / pitches take the byte-wise load and store path, multiples of four the word path; partial tiles in both."""
    rng = np.random.Generator(np.random.PCG64(w * 1000 + h))
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    src[: h // 2, : w // 2] = np.where(rng.integers(0, 2, (h // 2, w // 2)) > 0, 255, 0)     # saturating edges
    outs = [np.full((h, w), 0xEE, np.uint8) for _ in range(3)]
    emu_lib.load().emu_hpel(_vp(src.ctypes.data), *[_vp(o.ctypes.data) for o in outs], w, 0, w, 0, h)
    for o, wnt in zip(outs, E.oracle_hpel_full(src)):
        np.testing.assert_array_equal(o, wnt)


@pytest.mark.parametrize("r,qp", [(8, 23), (16, 36), (3, 0)])
def test_me_lowres(pair, r, qp):
    g, ref, cur = pair
    lam = oracle_lib.load().gvc_oracle_lambda(qp)
    a = E.me_lowres("oracle", cur, ref, r, lam)
    b = E.me_lowres(BACKEND, cur, ref, r, lam)
    np.testing.assert_array_equal(a, b)
    if r >= 8:   # planted motion: top-left quadrant moved by (3,-2) full samples => lowres (1 or 2, -1)
        assert abs(int(a[0, 0, 0]) * 2 - 3) <= 1 and abs(int(a[0, 0, 1]) * 2 + 2) <= 1


@pytest.mark.parametrize("qp", [12, 23, 40])
def test_me_refine_and_encode_inter(pair, qp):
    g, ref, cur = pair
    lam = oracle_lib.load().gvc_oracle_lambda(qp)
    low = E.me_lowres("oracle", cur, ref, 8, lam)
    mo = E.me_refine("oracle", cur, ref, low, lam)
    mb = E.me_refine(BACKEND, cur, ref, low, lam)
    np.testing.assert_array_equal(mo["mv"], mb["mv"])
    np.testing.assert_array_equal(mo["cost"], mb["cost"])
    np.testing.assert_array_equal(mo["mode"], mb["mode"])
    # known answer: MB (0,0) follows the planted (3,-2) full-sample shift = (12,-8) quarter samples
    assert tuple(mo["mv"][0, 0]) == (12, -8)
    co, nzo, reco = E.encode_inter("oracle", cur, ref, mo, qp)
    cb, nzb, recb = E.encode_inter(BACKEND, cur, ref, mo, qp)
    np.testing.assert_array_equal(co, cb)
    np.testing.assert_array_equal(nzo, nzb)
    for x, y in zip(reco, recb):
        np.testing.assert_array_equal(x, y)
    # reconstruction quality sanity: PSNR of recon vs source rises as qp falls
    err = np.mean((reco[0].astype(float) - cur.Y.inner()) ** 2)
    assert err < {12: 3.0, 23: 12.0, 40: 120.0}[qp]


def test_encode_inter_with_8x8_partitions_and_odd_vectors(pair):
    g, ref, cur = pair
    n = g["mbw"] * g["mbh"]
    rng = np.random.Generator(np.random.PCG64(9))
    mo = np.zeros(n, E.MOTION_DT)
    mo["mv"] = rng.integers(-37, 38, (n, 4, 2))          # every fractional phase, different per partition
    mo["mode"] = 1
    co, nzo, reco = E.encode_inter("oracle", cur, ref, mo, 26)
    cb, nzb, recb = E.encode_inter(BACKEND, cur, ref, mo, 26)
    np.testing.assert_array_equal(co, cb)
    np.testing.assert_array_equal(nzo, nzb)
    for x, y in zip(reco, recb):
        np.testing.assert_array_equal(x, y)


def test_me_refine_early_termination_on_static_content():
    """Macroblocks the zero vector predicts to within 32*lambda of SAD stop before the sub-sample search: a frame whose left
    half repeats the reference exactly (plus a faint flicker on a few macroblocks) exercises both exits side by side."""
    g = E.geom(160, 96)
    ref_yuv, cur_yuv = E.make_pair(g["w16"], g["h16"], seed=21)
    cur_yuv = {k: v.copy() for k, v in cur_yuv.items()}
    for k, v in cur_yuv.items():
        v[:, : v.shape[1] // 2] = ref_yuv[k][:, : v.shape[1] // 2]
    cur_yuv["Y"][5:9, 20:30] ^= 1                      # SAD 40: still below the threshold at lambda 4 (128)
    cur_yuv["Y"][40:44, 3:60] = 255 - cur_yuv["Y"][40:44, 3:60]      # far above it
    ref, cur = E.Frame(g, ref_yuv), E.Frame(g, cur_yuv, with_hpel=False)
    lam = oracle_lib.load().gvc_oracle_lambda(23)
    low = E.me_lowres("oracle", cur, ref, 8, lam)
    np.testing.assert_array_equal(low, E.me_lowres(BACKEND, cur, ref, 8, lam))     # static blocks: decided by the zero displacement alone
    assert not low.reshape(g["mbh"], g["mbw"], 2)[0, :4].any()
    mo = E.me_refine("oracle", cur, ref, low, lam)
    mb = E.me_refine(BACKEND, cur, ref, low, lam)
    for f in ("mv", "cost", "mode"):
        np.testing.assert_array_equal(mo[f], mb[f])
    mv = mo["mv"].reshape(g["mbh"], g["mbw"], 4, 2)
    assert not mv[0, :4].any() and not mv[5, :4].any()          # static macroblocks keep the zero vector
    assert mo["cost"].reshape(g["mbh"], g["mbw"])[0, 0] <= 32 * lam + 64
    assert mv[:, g["mbw"] // 2 + 1:].any()                      # the moving half went through the full search


@pytest.mark.parametrize("qp", [10, 23, 38])
def test_encode_intra_frame(pair, qp):
    g, ref, cur = pair
    co, nzo, mo, reco = E.encode_intra("oracle", cur, qp)
    cb, nzb, mb, recb = E.encode_intra(BACKEND, cur, qp)
    np.testing.assert_array_equal(mo, mb)
    np.testing.assert_array_equal(co, cb)
    np.testing.assert_array_equal(nzo, nzb)
    for x, y in zip(reco, recb):
        np.testing.assert_array_equal(x, y)
    assert mo[0] == 2 and set(mo[1:g["mbw"]]) <= {1, 2}          # first MB: DC only; first row: no vertical
    err = np.mean((reco[0].astype(float) - cur.Y.inner()) ** 2)
    assert err < {10: 2.5, 23: 14.0, 38: 140.0}[qp]


@pytest.mark.parametrize("qp,intra", [(0, 1), (18, 0), (23, 1), (37, 0), (51, 0)])
def test_transform8x8_path(qp, intra):
    """8x8 transform / quant / dequant / reconstruction kernel body against the oracle (x264 sub8x8_dct8 ... add8x8_idct8)."""
    rng = np.random.Generator(np.random.PCG64(qp))
    w, h = 48, 32
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    pred = np.clip(src.astype(int) + rng.integers(-40, 41, (h, w)), 0, 255).astype(np.uint8)
    rec = np.zeros((h, w), np.uint8)
    coefs = np.zeros((h // 8) * (w // 8) * 64, np.int16)
    nz = np.zeros((h // 8) * (w // 8), np.uint8)
    emu_lib.load().emu_transform8x8(_vp(src.ctypes.data), w, _vp(pred.ctypes.data), w, _vp(rec.ctypes.data), w, w, h, qp, intra,
                                    _vp(coefs.ctypes.data), _vp(nz.ctypes.data))
    mf, bias = oracle_lib.quant8_tables(qp, intra)
    k = 0
    for by in range(h // 8):
        for bx in range(w // 8):
            s, p = src[8 * by:8 * by + 8, 8 * bx:8 * bx + 8], pred[8 * by:8 * by + 8, 8 * bx:8 * bx + 8]
            q, flag = oracle_lib.quant(oracle_lib.sub_dct(s, p, 8), mf, bias)
            np.testing.assert_array_equal(coefs[64 * k:64 * k + 64], q)
            assert nz[k] == flag
            want = oracle_lib.add_idct(p, oracle_lib.dequant(q.reshape(8, 8), qp, 8), 8)
            np.testing.assert_array_equal(rec[8 * by:8 * by + 8, 8 * bx:8 * bx + 8], want)
            k += 1
