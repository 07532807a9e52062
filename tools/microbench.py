#!/usr/bin/env python
"""cfg5: kernel microbenchmark over picture sizes, each stage alone through the C ABI.

    python tools/microbench.py [--iters 50] [--out gpurun_out/microbench.md]

Inputs: a smooth random field as reference, the current picture = the reference displaced by a planted vector (one quadrant
moving differently) + noise (tests/enc_util.make_pair); planes built with the oracle.  Timing: CUDA events around `iters`
back-to-back launches after 3 warm-up launches, on the stream the kernels run on.  GB/s = SURVEY 8(d) algorithmic bytes /
time; the last column is the fraction of the measured copy peak (MEASURED_PEAKS.json, else the recipe's fallback).
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch

import enc_util as E
import oracle_lib
from gpu_enc import Dev

vp = ctypes.c_void_p


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3      # microseconds per launch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "microbench.md"))
    args = ap.parse_args()
    import govideocompressor_b200 as gvc
    dev = Dev()
    lib = dev.lib
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = float(json.load(open(pk))["hbm_gbs"]) if os.path.exists(pk) else 6650.0
    rows = []

    def add(kernel, size, us, nbytes):
        gbs = nbytes / us / 1e3
        rows.append((kernel, size, us, nbytes, gbs, gbs / peak))
        print("%-14s %-22s %9.1f us %8.1f GB/s  %5.2f %%" % (kernel, size, us, gbs, 100 * gbs / peak), flush=True)

    # ---- K1 resize, 8 pictures per launch
    for (sw, sh, dw, dh) in ((1920, 1080, 1280, 720), (3840, 2160, 1920, 1080), (1280, 720, 854, 480), (1920, 1080, 854, 480)):
        n = 8
        plan = gvc.ResizePlan(sw, sh, dw, dh)
        src = torch.randint(0, 256, (n, plan.frame_bytes(sw, sh)), dtype=torch.uint8, device="cuda")
        dst = torch.empty((n, plan.frame_bytes(dw, dh)), dtype=torch.uint8, device="cuda")
        us = timed(lambda: plan.resize_device(src, dst, n), args.iters)
        add("resize x8", "%dx%d->%dx%d" % (sw, sh, dw, dh), us, int(1.5 * (sw * sh + dw * dh)) * n)

    # ---- K1 resize, one 38-picture segment per launch (the pipeline's launch), and the allocation-free two-pass entry point
    import govideocompressor_b200.lib as L
    for (sw, sh, dw, dh) in ((1920, 1080, 1280, 720), (3840, 2160, 1920, 1080)):
        n = 38
        plan = gvc.ResizePlan(sw, sh, dw, dh)
        src = torch.randint(0, 256, (n, plan.frame_bytes(sw, sh)), dtype=torch.uint8, device="cuda")
        dst = torch.empty((n, plan.frame_bytes(dw, dh)), dtype=torch.uint8, device="cuda")
        us = timed(lambda: plan.resize_device(src, dst, n), args.iters)
        add("resize x38", "%dx%d->%dx%d" % (sw, sh, dw, dh), us, int(1.5 * (sw * sh + dw * dh)) * n)
        s_, d_ = plan._frames_packed(src.data_ptr(), sw, sh), plan._frames_packed(dst.data_ptr(), dw, dh)
        st = torch.cuda.current_stream().cuda_stream
        n2 = 8
        us = timed(lambda: plan._lib.gvc_resize_frames_device_static(plan._h, ctypes.byref(s_), ctypes.byref(d_), n2, ctypes.c_void_p(st)), max(5, args.iters // 5))
        add("resize static x8", "%dx%d->%dx%d (two-pass, built-in scratch)" % (sw, sh, dw, dh), us, int(1.5 * (sw * sh + dw * dh)) * n2)

    # ---- encoder stages per coded size
    P = E.PAD
    for (w, h, r) in ((854, 480, 8), (1280, 720, 8), (1920, 1080, 8), (3840, 2160, 16)):
        g = E.geom(w, h)
        W, H, nmb = g["w16"], g["h16"], g["mbw"] * g["mbh"]
        size = "%dx%d (%d macroblocks)" % (W, H, nmb)
        ref_yuv, cur_yuv = E.make_pair(W, H, seed=w)
        ref, cur = E.Frame(g, ref_yuv), E.Frame(g, cur_yuv, with_hpel=False)
        rp = dev.ref_planes(ref)
        cy, cu, cv, cl = dev.up(cur.Y), dev.up(cur.U), dev.up(cur.V), dev.up(cur.L)
        ry, ru, rv = dev.up(E.luma_plane(g)), dev.up(E.chroma_plane(g)), dev.up(E.chroma_plane(g))
        outs = [dev.up(E.luma_plane(g)) for _ in range(3)]
        louts = [dev.up(E.low_plane(g)) for _ in range(4)]
        lam, qp = oracle_lib.load().gvc_oracle_lambda(23), 23
        lp = P // 2
        add("lowres", size, timed(lambda: lib.gvc_lowres(vp(cy.ptr), g["sy"], -P, W + P, -P, H + P, vp(louts[0].ptr), None, None, None, g["sl"],
                                                          -lp, W // 2 + lp, -lp, H // 2 + lp, None), args.iters), W * H + W * H // 4)
        add("hpel", size, timed(lambda: lib.gvc_hpel_filter(vp(cy.ptr), *[vp(o.ptr) for o in outs], g["sy"], -P, W + P, -P, H + P, None), args.iters),
            4 * W * H)
        low = torch.zeros(2 * nmb, dtype=torch.int16, device="cuda")
        add("me_lowres r=%d" % r, size, timed(lambda: lib.gvc_me_lowres(vp(cl.ptr), rp.low, g["sl"], g["mbw"], g["mbh"], r, lam, vp(low.data_ptr()), None),
                                             args.iters), W * H // 2 + 4 * nmb)
        mot = torch.zeros(nmb * E.MOTION_DT.itemsize, dtype=torch.uint8, device="cuda")
        add("me_refine", size, timed(lambda: lib.gvc_me_refine(vp(cy.ptr), ctypes.byref(rp), g["sy"], W, H, vp(low.data_ptr()), lam, vp(mot.data_ptr()), None),
                                      args.iters), 5 * W * H + 24 * nmb)
        co = torch.zeros(nmb * 26 * 16, dtype=torch.int16, device="cuda")
        nz = torch.zeros(nmb, dtype=torch.int32, device="cuda")
        add("encode_inter", size, timed(lambda: lib.gvc_encode_inter(vp(cy.ptr), vp(cu.ptr), vp(cv.ptr), ctypes.byref(rp), vp(ry.ptr), vp(ru.ptr), vp(rv.ptr),
                                                                      g["sy"], g["sc"], W, H, vp(mot.data_ptr()), qp, vp(co.data_ptr()), vp(nz.data_ptr()), None),
                                         args.iters), int(7.5 * W * H))
        modes = torch.zeros(nmb, dtype=torch.uint8, device="cuda")
        scratch = torch.zeros(nmb * 16, dtype=torch.int32, device="cuda")
        add("encode_intra", size, timed(lambda: lib.gvc_encode_intra(vp(cy.ptr), vp(cu.ptr), vp(cv.ptr), vp(ry.ptr), vp(ru.ptr), vp(rv.ptr), g["sy"], g["sc"], W, H,
                                                                      qp - 3, vp(co.data_ptr()), vp(nz.data_ptr()), vp(modes.data_ptr()), vp(scratch.data_ptr()), None),
                                         max(5, args.iters // 5)), 6 * W * H)

    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        f.write("| kernel | size | us / launch | algorithmic bytes | GB/s | of measured peak (%.1f GB/s) |\n|---|---|---|---|---|---|\n" % peak)
        for k, s, us, nb, gbs, fr in rows:
            f.write("| %s | %s | %.1f | %d | %.1f | %.2f %% |\n" % (k, s, us, nb, gbs, 100 * fr))
    print("wrote", args.out)


if __name__ == "__main__":
    main()
