"""The whole per-segment path composed from the CPU oracle (test infrastructure): resize (oracle port, or the real
libswscale when asked), border extension, lowres, motion search, inter / intra macroblock encode, reconstruction,
half-sample planes.  Used as the segment-level parity reference for the GPU pipeline and as bench.py's CPU arm.
No entropy coding here: the oracle covers the kernels' arithmetic, the bitstream writer is host product code."""
import ctypes

import numpy as np

import enc_util as E
import oracle_lib


def _resize(planes, sw, sh, dw, dh, use_libswscale):
    if (sw, sh) == (dw, dh):
        return [p.copy() for p in planes]
    if use_libswscale:
        import sws_ref
        return [p.copy() for p in sws_ref.scale(planes, sw, sh, dw, dh)]
    return oracle_lib.sws_scale(planes, sw, sh, dw, dh, "yuv420p", 4)


def _pad_to_coded(planes, w, h, w16, h16):
    y, u, v = planes
    return dict(Y=np.pad(y, ((0, h16 - h), (0, w16 - w)), mode="edge"),
                U=np.pad(u, ((0, (h16 - h) // 2), (0, (w16 - w) // 2)), mode="edge"),
                V=np.pad(v, ((0, (h16 - h) // 2), (0, (w16 - w) // 2)), mode="edge"))


def encode_segment(frames, sw, sh, dw, dh, qp_i, qp_p, merange=16, use_libswscale=False):
    """frames: (n, sw*sh*3/2) I420.  Returns dict(recon=[(Y,U,V) coded size], motion=[...], coefs=[...], nz=[...], modes)."""
    lib = oracle_lib.load()
    g = E.geom(dw, dh)
    n = frames.shape[0]
    out = dict(recon=[], motion=[None], coefs=[], nz=[], modes=None)
    ref = None
    for f in range(n):
        y = frames[f][: sw * sh].reshape(sh, sw)
        u = frames[f][sw * sh: sw * sh * 5 // 4].reshape(sh // 2, sw // 2)
        v = frames[f][sw * sh * 5 // 4:].reshape(sh // 2, sw // 2)
        cur = E.Frame(g, _pad_to_coded(_resize([y, u, v], sw, sh, dw, dh, use_libswscale), dw, dh, g["w16"], g["h16"]), with_hpel=False)
        if f == 0:
            co, nz, modes, rec = E.encode_intra("oracle", cur, qp_i)
            out["modes"] = modes
        else:
            lam = lib.gvc_oracle_lambda(qp_p)
            low = E.me_lowres("oracle", cur, ref, merange // 2, lam)
            mo = E.me_refine("oracle", cur, ref, low, lam)
            co, nz, rec = E.encode_inter("oracle", cur, ref, mo, qp_p)
            out["motion"].append(mo)
        out["coefs"].append(co)
        out["nz"].append(nz)
        out["recon"].append(rec)
        if f + 1 < n:
            ref = E.Frame(g, dict(Y=rec[0], U=rec[1], V=rec[2]), with_hpel=True)
    return out
