"""Test helpers for the encoder-side kernels: padded-plane containers in the kernels' HBM layout, synthetic
sequences with planted motion, and three interchangeable back ends over the same host arrays:
  'oracle' -- oracle/libgvc_oracle.so (the CPU specification),
  'emu'    -- tests/emu/libgvc_emu.so (the CUDA kernel bodies executed serially on the host),
  'gpu'    -- libgvc.so through the C ABI (device buffers via torch)."""
import ctypes

import numpy as np

import emu_lib
import oracle_lib

PAD = 64
MOTION_DT = np.dtype([("mv", np.int16, (4, 2)), ("cost", np.int32), ("mode", np.uint8), ("pad", np.uint8, 3)])
MB_COEFS = 26 * 16


def geom(w, h):
    out = (ctypes.c_int * 9)()
    emu_lib.load().emu_geom(w, h, out)
    return dict(w=w, h=h, w16=out[0], h16=out[1], mbw=out[2], mbh=out[3], sy=out[4], sc=out[5], sl=out[6])


class Plane:
    """A padded plane: array rows are `stride` bytes, sample (0,0) sits at [pad, pad]."""

    def __init__(self, w, h, pad, stride, fill=0):
        self.w, self.h, self.pad, self.stride = w, h, pad, stride
        self.arr = np.full((h + 2 * pad + 2, stride), fill, np.uint8)   # 2 rows of slack for word loads

    @property
    def origin(self):
        return self.pad * self.stride + self.pad

    def ptr(self):
        return self.arr.ctypes.data + self.origin

    def inner(self):
        return self.arr[self.pad:self.pad + self.h, self.pad:self.pad + self.w]

    def full(self):
        return self.arr[0:self.h + 2 * self.pad, 0:self.w + 2 * self.pad]

    def set_full(self, a):
        self.full()[...] = a
        return self


def luma_plane(g, fill=0):
    return Plane(g["w16"], g["h16"], PAD, g["sy"], fill)


def chroma_plane(g, fill=0):
    return Plane(g["w16"] // 2, g["h16"] // 2, PAD // 2, g["sc"], fill)


def low_plane(g, fill=0):
    return Plane(g["w16"] // 2, g["h16"] // 2, PAD // 2, g["sl"], fill)


# ------------------------------------------------------------------ synthetic content ---------------
def smooth_field(h, w, seed, scale=8):
    """Band-limited texture: noise blown up by `scale` and box filtered, 8-bit."""
    rng = np.random.Generator(np.random.PCG64(seed))
    small = rng.integers(0, 256, (h // scale + 3, w // scale + 3)).astype(np.float64)
    big = np.kron(small, np.ones((scale, scale)))
    k = scale
    c = np.cumsum(np.cumsum(np.pad(big, ((k, 0), (k, 0))), 0), 1)
    box = (c[k:, k:] - c[:-k, k:] - c[k:, :-k] + c[:-k, :-k]) / (k * k)
    fine = rng.integers(-12, 13, box.shape)
    return np.clip(box + fine, 0, 255).astype(np.uint8)[: h + 2 * scale, : w + 2 * scale]


def make_pair(w16, h16, seed, shift=(3, -2), noise=2):
    """(ref, cur) frames as dicts of Y/U/V coded-size arrays; cur is ref displaced by `shift` (x,y) with one quadrant
    moving differently and a little noise: a known-answer motion field."""
    rng = np.random.Generator(np.random.PCG64(seed + 1000))
    m = 32
    out = []
    base = {}
    for name, (hh, ww, s) in dict(Y=(h16, w16, 8), U=(h16 // 2, w16 // 2, 4), V=(h16 // 2, w16 // 2, 4)).items():
        base[name] = smooth_field(hh + 2 * m, ww + 2 * m, seed * 7 + len(name) + (0 if name == "Y" else ord(name)), s)
    for (dx, dy) in ((0, 0), shift):
        f = {}
        for name in "YUV":
            div = 1 if name == "Y" else 2
            hh, ww = (h16 // div, w16 // div)
            mm = m // div if name != "Y" else m
            a = base[name][mm + dy // div: mm + dy // div + hh, mm + dx // div: mm + dx // div + ww].copy()
            if (dx, dy) != (0, 0):
                q = base[name][mm - 2 // div: mm - 2 // div + hh, mm + 5 // div: mm + 5 // div + ww]
                a[hh // 2:, ww // 2:] = q[hh // 2:, ww // 2:]
                a = np.clip(a.astype(int) + rng.integers(-noise, noise + 1, a.shape), 0, 255).astype(np.uint8)
            f[name] = np.ascontiguousarray(a)
        out.append(f)
    return out[0], out[1]


# ------------------------------------------------------------------ oracle-built frames -------------
def oracle_hpel_full(y_padded):
    """y_padded: (h+2P, w+2P) border-extended luma.  Returns H, V, C of the same shape under infinite replication."""
    big = np.ascontiguousarray(np.pad(y_padded, 3, mode="edge"))
    outs = [np.zeros_like(big) for _ in range(3)]
    h, w = y_padded.shape
    off = 3 * big.strides[0] + 3
    P = ctypes.c_void_p
    # treat the whole padded plane as the "picture" with pad 3: margin m = 0 -> outputs exactly [0,w) x [0,h)
    oracle_lib.load().gvc_oracle_hpel_filter(P(outs[0].ctypes.data + off), P(outs[1].ctypes.data + off), P(outs[2].ctypes.data + off),
                                             P(big.ctypes.data + off), big.strides[0], w, h, 3)
    return [o[3:-3, 3:-3] for o in outs]


def oracle_lowres_full(y_padded):
    """lowres (4 phases) of a border-extended plane, reads clamped at its far edge."""
    h, w = y_padded.shape
    src = np.ascontiguousarray(np.pad(y_padded, ((0, 2), (0, 2)), mode="edge"))
    outs = [np.zeros((h // 2, w // 2), np.uint8) for _ in range(4)]
    P = ctypes.c_void_p
    oracle_lib.load().gvc_oracle_lowres(P(src.ctypes.data), src.strides[0], *[P(o.ctypes.data) for o in outs], w // 2, w // 2, h // 2)
    return outs


class Frame:
    """All planes of one frame in kernel layout, built with the oracle (border extension, hpel, lowres)."""

    def __init__(self, g, yuv, with_hpel=True):
        self.g = g
        self.Y = luma_plane(g).set_full(np.pad(yuv["Y"], PAD, mode="edge"))
        self.U = chroma_plane(g).set_full(np.pad(yuv["U"], PAD // 2, mode="edge"))
        self.V = chroma_plane(g).set_full(np.pad(yuv["V"], PAD // 2, mode="edge"))
        self.L = low_plane(g).set_full(oracle_lowres_full(self.Y.full())[0])
        self.H = self.Vp = self.C = None
        if with_hpel:
            hh, vv, cc = oracle_hpel_full(self.Y.full())
            self.H, self.Vp, self.C = luma_plane(g).set_full(hh), luma_plane(g).set_full(vv), luma_plane(g).set_full(cc)

    def ref_ptrs(self):
        return [self.Y.ptr(), self.H.ptr(), self.Vp.ptr(), self.C.ptr(), self.U.ptr(), self.V.ptr(), self.L.ptr()]


# ------------------------------------------------------------------ back ends ------------------------
def _vp(x):
    return ctypes.c_void_p(x)


def me_lowres(backend, cur, ref, r, lam):
    g = cur.g
    out = np.zeros((g["mbh"], g["mbw"], 2), np.int16)
    if backend == "oracle":
        oracle_lib.load().gvc_oracle_me_lowres(_vp(cur.L.ptr()), _vp(ref.L.ptr()), g["sl"], g["mbw"], g["mbh"], r, lam, _vp(out.ctypes.data), None)
    elif backend == "emu":
        emu_lib.load().emu_me_lowres(_vp(cur.L.ptr()), _vp(ref.L.ptr()), g["sl"], g["mbw"], g["mbh"], r, lam, _vp(out.ctypes.data))
    else:
        import torch
        from gpu_enc import Dev
        d = Dev()
        dc, dr = d.up(cur.L), d.up(ref.L)
        o = torch.zeros(out.size, dtype=torch.int16, device="cuda")
        d.check(d.lib.gvc_me_lowres(_vp(dc.ptr), _vp(dr.ptr), g["sl"], g["mbw"], g["mbh"], r, lam, _vp(o.data_ptr()), None))
        torch.cuda.synchronize()
        out = o.cpu().numpy().reshape(out.shape)
    return out


def me_refine(backend, cur, ref, lowmv, lam):
    g = cur.g
    n = g["mbw"] * g["mbh"]
    lowmv = np.ascontiguousarray(lowmv, np.int16)
    if backend == "oracle":
        raw = np.zeros((n, 10), np.int32)
        pl = (ctypes.c_void_p * 4)(*ref.ref_ptrs()[:4])
        oracle_lib.load().gvc_oracle_me_refine(_vp(cur.Y.ptr()), pl, g["sy"], g["w16"], g["h16"], PAD, _vp(lowmv.ctypes.data), lam, _vp(raw.ctypes.data))
        out = np.zeros(n, MOTION_DT)
        out["mv"] = raw[:, :8].reshape(n, 4, 2)
        out["cost"] = raw[:, 8]
        out["mode"] = raw[:, 9]
        return out
    out = np.zeros(n, MOTION_DT)
    if backend == "emu":
        pl = (ctypes.c_void_p * 7)(*ref.ref_ptrs())
        emu_lib.load().emu_me_refine(_vp(cur.Y.ptr()), pl, g["sy"], g["w16"], g["h16"], _vp(lowmv.ctypes.data), lam, _vp(out.ctypes.data))
        return out
    import torch
    from gpu_enc import Dev
    d = Dev()
    dcur = d.up(cur.Y)
    rp = d.ref_planes(ref)
    dl = torch.from_numpy(lowmv.reshape(-1)).cuda()
    o = torch.zeros(n * MOTION_DT.itemsize, dtype=torch.uint8, device="cuda")
    d.check(d.lib.gvc_me_refine(_vp(dcur.ptr), ctypes.byref(rp), g["sy"], g["w16"], g["h16"], _vp(dl.data_ptr()), lam, _vp(o.data_ptr()), None))
    torch.cuda.synchronize()
    return np.frombuffer(o.cpu().numpy().tobytes(), MOTION_DT).copy()


def encode_inter(backend, cur, ref, motion, qp):
    """Returns (coefs[n,26,16] int16, nzmask[n] uint32, recon Y/U/V inner arrays)."""
    g = cur.g
    n = g["mbw"] * g["mbh"]
    coefs = np.zeros((n, 26, 16), np.int16)
    nz = np.zeros(n, np.uint32)
    rec = [luma_plane(g, 0x11), chroma_plane(g, 0x11), chroma_plane(g, 0x11)]
    motion = np.ascontiguousarray(motion)
    if backend == "oracle":
        raw = np.zeros((n, 10), np.int32)
        raw[:, :8] = motion["mv"].reshape(n, 8)
        raw[:, 8] = motion["cost"]
        raw[:, 9] = motion["mode"]
        P3, P4, P2 = ctypes.c_void_p * 3, ctypes.c_void_p * 4, ctypes.c_void_p * 2
        oracle_lib.load().gvc_oracle_encode_inter_frame(P3(cur.Y.ptr(), cur.U.ptr(), cur.V.ptr()), P4(*ref.ref_ptrs()[:4]),
                                                        P2(ref.U.ptr(), ref.V.ptr()), P3(*[r.ptr() for r in rec]), g["sy"], g["sc"],
                                                        g["w16"], g["h16"], _vp(raw.ctypes.data), qp, _vp(coefs.ctypes.data), _vp(nz.ctypes.data))
    elif backend == "emu":
        P3, P7 = ctypes.c_void_p * 3, ctypes.c_void_p * 7
        emu_lib.load().emu_encode_inter(P3(cur.Y.ptr(), cur.U.ptr(), cur.V.ptr()), P7(*ref.ref_ptrs()), P3(*[r.ptr() for r in rec]),
                                        g["sy"], g["sc"], g["w16"], g["h16"], _vp(motion.ctypes.data), qp, _vp(coefs.ctypes.data), _vp(nz.ctypes.data))
    else:
        import torch
        from gpu_enc import Dev
        d = Dev()
        cy, cu, cv = d.up(cur.Y), d.up(cur.U), d.up(cur.V)
        rp = d.ref_planes(ref)
        ry, ru, rv = d.up(rec[0]), d.up(rec[1]), d.up(rec[2])
        dm = torch.from_numpy(np.frombuffer(motion.tobytes(), np.uint8).copy()).cuda()
        dco = torch.zeros(coefs.size, dtype=torch.int16, device="cuda")
        dnz = torch.zeros(n, dtype=torch.int32, device="cuda")
        d.check(d.lib.gvc_encode_inter(_vp(cy.ptr), _vp(cu.ptr), _vp(cv.ptr), ctypes.byref(rp), _vp(ry.ptr), _vp(ru.ptr), _vp(rv.ptr),
                                       g["sy"], g["sc"], g["w16"], g["h16"], _vp(dm.data_ptr()), qp, _vp(dco.data_ptr()), _vp(dnz.data_ptr()), None))
        torch.cuda.synchronize()
        coefs = dco.cpu().numpy().reshape(coefs.shape)
        nz = dnz.cpu().numpy().view(np.uint32)
        for r, dev in zip(rec, (ry, ru, rv)):
            r.arr[...] = dev.t.cpu().numpy().reshape(r.arr.shape)
    return coefs, nz, [r.inner().copy() for r in rec]


def encode_intra(backend, cur, qp):
    """Returns (coefs[n,26,16], nzmask[n], modes[n], recon Y/U/V inner arrays)."""
    g = cur.g
    n = g["mbw"] * g["mbh"]
    coefs = np.zeros((n, 26, 16), np.int16)
    nz = np.zeros(n, np.uint32)
    modes = np.zeros(n, np.uint8)
    rec = [luma_plane(g, 0x11), chroma_plane(g, 0x11), chroma_plane(g, 0x11)]
    P3 = ctypes.c_void_p * 3
    if backend == "oracle":
        oracle_lib.load().gvc_oracle_encode_intra_frame(P3(cur.Y.ptr(), cur.U.ptr(), cur.V.ptr()), P3(*[r.ptr() for r in rec]), g["sy"], g["sc"],
                                                        g["w16"], g["h16"], qp, _vp(coefs.ctypes.data), _vp(nz.ctypes.data), _vp(modes.ctypes.data))
    elif backend == "emu":
        emu_lib.load().emu_encode_intra(P3(cur.Y.ptr(), cur.U.ptr(), cur.V.ptr()), P3(*[r.ptr() for r in rec]), g["sy"], g["sc"],
                                        g["w16"], g["h16"], qp, _vp(coefs.ctypes.data), _vp(nz.ctypes.data), _vp(modes.ctypes.data))
    else:
        import torch
        from gpu_enc import Dev
        d = Dev()
        cy, cu, cv = d.up(cur.Y), d.up(cur.U), d.up(cur.V)
        ry, ru, rv = d.up(rec[0]), d.up(rec[1]), d.up(rec[2])
        dco = torch.zeros(coefs.size, dtype=torch.int16, device="cuda")
        dnz = torch.zeros(n, dtype=torch.int32, device="cuda")
        dmo = torch.zeros(n, dtype=torch.uint8, device="cuda")
        prog = torch.full((n * 16,), 77, dtype=torch.int32, device="cuda")     # GVC_INTRA_SCRATCH_BYTES, dirty on purpose
        d.check(d.lib.gvc_encode_intra(_vp(cy.ptr), _vp(cu.ptr), _vp(cv.ptr), _vp(ry.ptr), _vp(ru.ptr), _vp(rv.ptr), g["sy"], g["sc"],
                                       g["w16"], g["h16"], qp, _vp(dco.data_ptr()), _vp(dnz.data_ptr()), _vp(dmo.data_ptr()),
                                       _vp(prog.data_ptr()), None))
        torch.cuda.synchronize()
        coefs = dco.cpu().numpy().reshape(coefs.shape)
        nz = dnz.cpu().numpy().view(np.uint32)
        modes = dmo.cpu().numpy()
        for r, dev in zip(rec, (ry, ru, rv)):
            r.arr[...] = dev.t.cpu().numpy().reshape(r.arr.shape)
    return coefs, nz, modes, [r.inner().copy() for r in rec]
