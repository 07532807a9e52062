"""GPU parity (-m gpu): encoder-side kernels of libgvc.so through the C ABI against the oracle, at a small size
and at the BASELINE size (1280x720 coded frames = 3,600 macroblocks)."""
import ctypes

import numpy as np
import pytest

import enc_util as E
import oracle_lib

pytestmark = pytest.mark.gpu


def _vp(x):
    return ctypes.c_void_p(x)


@pytest.fixture(scope="module", params=[(96, 64, 3, 0), (1280, 720, 11, 0), (200, 120, 5, 0), (96, 64, 3, 1)], ids=["96x64", "720p", "200x120", "96x64-odd-pitch"])
def pair(request):
    w, h, seed, odd = request.param
    g = E.geom(w, h)
    if odd:                               # pitches / addresses that are not multiples of four: the byte-wise load and store paths
        g = dict(g, sy=g["sy"] + 3, sc=g["sc"] + 1, sl=g["sl"] + 1)
    ref_yuv, cur_yuv = E.make_pair(g["w16"], g["h16"], seed=seed)
    return g, E.Frame(g, ref_yuv), E.Frame(g, cur_yuv, with_hpel=False)


@pytest.fixture(scope="module")
def dev():
    from gpu_enc import Dev
    return Dev()


def _down(devplane, plane):
    out = type(plane)(plane.w, plane.h, plane.pad, plane.stride)
    out.arr[...] = devplane.t.cpu().numpy().reshape(plane.arr.shape)
    return out


def test_border_lowres_hpel(pair, dev):
    import torch
    g, ref, cur = pair
    # border: start from a plane whose padding is junk
    w, h = g["w16"] - 6, g["h16"] - 4
    pl = E.luma_plane(g, 0xEE)
    pl.arr[E.PAD:E.PAD + h, E.PAD:E.PAD + w] = ref.Y.inner()[:h, :w]
    want = np.pad(ref.Y.inner()[:h, :w], ((E.PAD, E.PAD + g["h16"] - h), (E.PAD, E.PAD + g["w16"] - w)), mode="edge")
    d = dev.up(pl)
    dev.check(dev.lib.gvc_extend_border(_vp(d.ptr), g["sy"], w, h, -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD, None))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(_down(d, pl).full(), want)
    # lowres, all four phases, over the padded lowres area
    dy = dev.up(ref.Y)
    outs = [E.low_plane(g, 0xEE) for _ in range(4)]
    douts = [dev.up(o) for o in outs]
    lp = E.PAD // 2
    dev.check(dev.lib.gvc_lowres(_vp(dy.ptr), g["sy"], -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD, *[_vp(o.ptr) for o in douts],
                                 g["sl"], -lp, g["w16"] // 2 + lp, -lp, g["h16"] // 2 + lp, None))
    torch.cuda.synchronize()
    for dd, o, wnt in zip(douts, outs, E.oracle_lowres_full(ref.Y.full())):
        np.testing.assert_array_equal(_down(dd, o).full(), wnt)
    # half-sample planes over the whole padded area
    outs = [E.luma_plane(g, 0xEE) for _ in range(3)]
    douts = [dev.up(o) for o in outs]
    dev.check(dev.lib.gvc_hpel_filter(_vp(dy.ptr), *[_vp(o.ptr) for o in douts], g["sy"], -E.PAD, g["w16"] + E.PAD, -E.PAD, g["h16"] + E.PAD, None))
    torch.cuda.synchronize()
    for dd, o, wnt in zip(douts, outs, (ref.H, ref.Vp, ref.C)):
        np.testing.assert_array_equal(_down(dd, o).full(), wnt.full())


@pytest.mark.parametrize("w,h", [(131, 77), (132, 70), (64, 32), (5, 3), (200, 33), (1408, 848)])
def test_hpel_any_size_and_alignment(dev, w, h):
    """Odd widths / pitches take the byte-wise load and store path, multiples of four the word path; partial tiles in both."""
    import torch
    rng = np.random.Generator(np.random.PCG64(w * 1000 + h))
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    src[: h // 2, : w // 2] = np.where(rng.integers(0, 2, (h // 2, w // 2)) > 0, 255, 0)     # saturating edges
    dsrc = torch.from_numpy(src).cuda()
    douts = [torch.full((h, w), 0xEE, dtype=torch.uint8, device="cuda") for _ in range(3)]
    dev.check(dev.lib.gvc_hpel_filter(_vp(dsrc.data_ptr()), *[_vp(o.data_ptr()) for o in douts], w, 0, w, 0, h, None))
    torch.cuda.synchronize()
    for o, wnt in zip(douts, E.oracle_hpel_full(src)):
        np.testing.assert_array_equal(o.cpu().numpy(), wnt)


@pytest.mark.parametrize("r,qp", [(8, 23), (16, 36)])
def test_me_lowres(pair, r, qp):
    g, ref, cur = pair
    lam = oracle_lib.load().gvc_oracle_lambda(qp)
    np.testing.assert_array_equal(E.me_lowres("oracle", cur, ref, r, lam), E.me_lowres("gpu", cur, ref, r, lam))


@pytest.mark.parametrize("qp", [12, 23, 40])
def test_me_refine_and_encode_inter(pair, qp):
    g, ref, cur = pair
    lam = oracle_lib.load().gvc_oracle_lambda(qp)
    low = E.me_lowres("gpu", cur, ref, 8, lam)
    mo = E.me_refine("oracle", cur, ref, low, lam)
    mg = E.me_refine("gpu", cur, ref, low, lam)
    for f in ("mv", "cost", "mode"):
        np.testing.assert_array_equal(mo[f], mg[f])
    assert tuple(mo["mv"][0, 0]) == (12, -8)          # planted shift of the top-left quadrant
    co, nzo, reco = E.encode_inter("oracle", cur, ref, mo, qp)
    cg, nzg, recg = E.encode_inter("gpu", cur, ref, mg, qp)
    np.testing.assert_array_equal(co, cg)
    np.testing.assert_array_equal(nzo, nzg)
    for x, y in zip(reco, recg):
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
    np.testing.assert_array_equal(low, E.me_lowres("gpu", cur, ref, 8, lam))     # static blocks: decided by the zero displacement alone
    assert not low.reshape(g["mbh"], g["mbw"], 2)[0, :4].any()
    mo = E.me_refine("oracle", cur, ref, low, lam)
    mb = E.me_refine("gpu", cur, ref, low, lam)
    for f in ("mv", "cost", "mode"):
        np.testing.assert_array_equal(mo[f], mb[f])
    mv = mo["mv"].reshape(g["mbh"], g["mbw"], 4, 2)
    assert not mv[0, :4].any() and not mv[5, :4].any()          # static macroblocks keep the zero vector
    assert mo["cost"].reshape(g["mbh"], g["mbw"])[0, 0] <= 32 * lam + 64
    assert mv[:, g["mbw"] // 2 + 1:].any()                      # the moving half went through the full search


def test_encode_inter_all_fractional_phases(pair):
    g, ref, cur = pair
    n = g["mbw"] * g["mbh"]
    rng = np.random.Generator(np.random.PCG64(9))
    mo = np.zeros(n, E.MOTION_DT)
    mo["mv"] = rng.integers(-37, 38, (n, 4, 2))
    mo["mode"] = 1
    co, nzo, reco = E.encode_inter("oracle", cur, ref, mo, 26)
    cg, nzg, recg = E.encode_inter("gpu", cur, ref, mo, 26)
    np.testing.assert_array_equal(co, cg)
    np.testing.assert_array_equal(nzo, nzg)
    for x, y in zip(reco, recg):
        np.testing.assert_array_equal(x, y)


def test_bad_arguments(dev):
    assert dev.lib.gvc_me_lowres(None, None, 64, 4, 4, 8, 4, None, None) == -1
    assert dev.lib.gvc_encode_inter(None, None, None, None, None, None, None, 0, 0, 16, 16, None, 23, None, None, None) == -1


@pytest.mark.parametrize("qp", [10, 23, 38])
def test_encode_intra_frame(pair, qp):
    g, ref, cur = pair
    co, nzo, mo, reco = E.encode_intra("oracle", cur, qp)
    cg, nzg, mg, recg = E.encode_intra("gpu", cur, qp)
    np.testing.assert_array_equal(mo, mg)
    np.testing.assert_array_equal(co, cg)
    np.testing.assert_array_equal(nzo, nzg)
    for x, y in zip(reco, recg):
        np.testing.assert_array_equal(x, y)


@pytest.mark.parametrize("qp,intra", [(0, 1), (23, 0), (38, 1), (51, 0)])
def test_transform8x8_path(dev, qp, intra):
    import torch
    rng = np.random.Generator(np.random.PCG64(100 + qp))
    w, h = 1280, 720
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    pred = np.clip(src.astype(int) + rng.integers(-40, 41, (h, w)), 0, 255).astype(np.uint8)
    ds, dp = torch.from_numpy(src).cuda(), torch.from_numpy(pred).cuda()
    dr = torch.zeros((h, w), dtype=torch.uint8, device="cuda")
    nb = (h // 8) * (w // 8)
    dco = torch.zeros(nb * 64, dtype=torch.int16, device="cuda")
    dnz = torch.zeros(nb, dtype=torch.uint8, device="cuda")
    dev.check(dev.lib.gvc_transform8x8(_vp(ds.data_ptr()), w, _vp(dp.data_ptr()), w, _vp(dr.data_ptr()), w, w, h, qp, intra,
                                       _vp(dco.data_ptr()), _vp(dnz.data_ptr()), None))
    torch.cuda.synchronize()
    rec, coefs, nz = dr.cpu().numpy(), dco.cpu().numpy(), dnz.cpu().numpy()
    # EVERY block against the oracle's x264-semantics routines (sub8x8_dct8, quant_8x8, dequant_8x8, add8x8_idct8)
    mf, bias = oracle_lib.quant8_tables(qp, intra)
    for k in range(nb):
        by, bx = divmod(k, w // 8)
        s, p = src[8 * by:8 * by + 8, 8 * bx:8 * bx + 8], pred[8 * by:8 * by + 8, 8 * bx:8 * bx + 8]
        q, qnz = oracle_lib.quant(oracle_lib.sub_dct(s, p, 8), mf, bias)
        assert np.array_equal(coefs[64 * k:64 * k + 64], q), f"levels of block {k}"
        assert bool(nz[k]) == bool(qnz), f"nz flag of block {k}"
        assert np.array_equal(rec[8 * by:8 * by + 8, 8 * bx:8 * bx + 8], oracle_lib.add_idct(p, oracle_lib.dequant(q.reshape(8, 8), qp, 8), 8)), f"reconstruction of block {k}"
