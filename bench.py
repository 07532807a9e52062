#!/usr/bin/env python
"""bench.py -- headline benchmark of the segment compress hot path (BASELINE.json: 1080p->720p frames/sec).

A "step" is one pass of the hot path over one segment of the cfg2 workload (10 s 1080p30 clip split into 8
segments => 38 frames of 1920x1080 YUV420P, resized to 1280x720 bicubic).  One process per GPU; with N GPUs
every rank processes its own segment (segments share nothing: weak scaling, no collective on the data path).

  value   frames/s over all ranks, inputs resident in HBM, CUDA events on the launching stream
  e2e     the same metric through the C ABI with pinned HOST buffers (gvc_resize_frames_host: H2D + kernels + D2H)
  roofline / cpu_baseline / clocks: see DESIGN.md "Measurement"

`--impl reference` times the reference's CPU implementation of the same path on the host cores (FFmpeg's
libswscale bundled in the image when loadable, else the oracle port), on the same config/metric.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

SW, SH, DW, DH = 1920, 1080, 1280, 720
SEG_FRAMES = 38                      # ceil(300 frames / 8 segments)
ALGO_BYTES_PER_FRAME = int(1.5 * (SW * SH + DW * DH))   # 4,492,800 B (read source + write destination)
NBUF = 3                             # rotate segment buffers so no step finds its input in the 126 MB L2
METRIC = "1080p->720p H.264-path frames/sec (resize stage)"
CONFIG = {"workload": "cfg2: 1080p30 synthetic clip, 38-frame segment per GPU, 1920x1080->1280x720 bicubic YUV420P",
          "stages": ["resize"], "segment_frames": SEG_FRAMES, "l2": "inputs rotate over %d segment buffers (>L2)" % NBUF,
          "parallelism": "segment-per-GPU"}


def synth_segment(seed):
    import numpy as np
    from util_inputs import make_frame
    fb = SW * SH * 3 // 2
    seg = np.empty((SEG_FRAMES, fb), np.uint8)
    base = make_frame(SW, SH, "yuv420p", "random", seed)
    for f in range(SEG_FRAMES):
        # testsrc-style motion on top of a noise floor keeps every frame distinct without 38 RNG passes
        pl = make_frame(SW, SH, "yuv420p", "testsrc", seed, t=f)
        pl[0][::2] = base[0][::2] ^ (f & 255)
        seg[f] = np.concatenate([p.reshape(-1) for p in pl])
    return seg


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_resize_fn():
    """(callable(planes)->planes, kind, description) for the CPU arm; one scaler context per worker thread."""
    tl = threading.local()
    try:
        import sws_ref
        v = sws_ref.load().swscale_version()

        def fn(pl):
            if not hasattr(tl, "s"):
                tl.s = sws_ref.Scaler(SW, SH, DW, DH)
            return tl.s(pl)
        return (fn, "reference", "FFmpeg libswscale %d.%d.%d (image's opencv wheel), SWS_BICUBIC|BITEXACT|ACCURATE_RND, "
                "context reused per thread" % (v >> 16, (v >> 8) & 255, v & 255))
    except Exception:
        import oracle_lib
        oracle_lib.load()
        return (lambda pl: oracle_lib.sws_scale(pl, SW, SH, DW, DH, "yuv420p", 4), "port", "oracle/sws_resize.c (plain C port)")


def cpu_frames_per_sec(n_frames, threads):
    from concurrent.futures import ThreadPoolExecutor
    from util_inputs import make_frame
    fn, kind, desc = cpu_resize_fn()
    frames = [make_frame(SW, SH, "yuv420p", "testsrc", 0, t=i) for i in range(min(n_frames, 8))]
    fn(frames[0])
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:          # ctypes releases the GIL inside the C call
        list(ex.map(lambda i: fn(frames[i % len(frames)]), range(n_frames)))
    dt = time.perf_counter() - t0
    return n_frames / dt, kind, desc, dt


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n = max(cores * 4, 32)
    for _ in range(args.warmup):
        cpu_frames_per_sec(cores, cores)
    t_total, frames = 0.0, 0
    kind = desc = None
    for _ in range(args.steps):
        fps, kind, desc, dt = cpu_frames_per_sec(n, cores)
        t_total += dt
        frames += n
    v = frames / t_total
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                      "config": CONFIG,
                      "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind,
                                       "sample": "%d frames per step of the cfg2 clip, %d threads; %s" % (n, cores, desc)},
                      "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gvc", choices=["gvc", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "gvc" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import numpy as np
    import torch
    import torch.distributed as dist
    import govideocompressor_b200 as gvc

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gvc.init(local)
    plan = gvc.ResizePlan(SW, SH, DW, DH, gvc.YUV420P, gvc.BICUBIC)
    assert plan.info()["fused"], "headline config must run the fused kernel"
    fb_s, fb_d = plan.frame_bytes(SW, SH), plan.frame_bytes(DW, DH)

    seg0 = synth_segment(rank)
    # further buffers: same clip with every byte XORed by a per-buffer constant (distinct data, cheap to make)
    host_src = [torch.from_numpy(seg0 if b == 0 else seg0 ^ np.uint8(37 * b + 1)).pin_memory() for b in range(NBUF)]
    host_dst = torch.empty((SEG_FRAMES, fb_d), dtype=torch.uint8).pin_memory()
    dev_src = [h.cuda() for h in host_src]
    dev_dst = [torch.empty((SEG_FRAMES, fb_d), dtype=torch.uint8, device="cuda") for _ in range(NBUF)]
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        plan.resize_device(dev_src[i % NBUF], dev_dst[i % NBUF], SEG_FRAMES, stream)

    # ---- device-resident timing ------------------------------------------------------------
    for i in range(args.warmup):
        step(i)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    for i in range(args.steps):
        step(i)
        ev[i + 1].record(stream)
    barrier()
    clocks = sampler.stop()
    ms_total = ev[0].elapsed_time(ev[-1])
    launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * SEG_FRAMES * args.steps / (ms_total * 1e-3)

    # ---- end to end through the C ABI with host buffers -------------------------------------
    for i in range(2):
        plan.resize_host_packed(host_src[i % NBUF].data_ptr(), host_dst.data_ptr(), SEG_FRAMES)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        plan.resize_host_packed(host_src[i % NBUF].data_ptr(), host_dst.data_ptr(), SEG_FRAMES)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e = world * SEG_FRAMES * args.steps / float(t.item())
    checksum = int(host_dst[-1].to(torch.int64).sum().item())   # the step's result is read on the host

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        med = sorted(launch_ms)[len(launch_ms) // 2]
        avg = sum(launch_ms) / len(launch_ms)
        achieved = ALGO_BYTES_PER_FRAME * SEG_FRAMES / (avg * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get("resize_fused_kernel_bytes_per_launch")
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": CONFIG,
                "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": SEG_FRAMES * fb_s,
                        "d2h_bytes_per_step": SEG_FRAMES * fb_d, "result_checksum": checksum},
                "gpu_launches": args.steps, "clocks": clocks,
                "roofline": {"kernel": "resize_fused_kernel<2,2>", "bound": "hbm", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                             "algorithmic_bytes_per_launch": ALGO_BYTES_PER_FRAME * SEG_FRAMES,
                             "launch_ms_avg": avg, "launch_ms_median": med,
                             "frac_of_8TBps_nominal": achieved / 8000.0}}
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            n = max(cores * 8, 64)
            fps, kind, desc, dt = cpu_frames_per_sec(n, cores)
            line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                                    "sample": "%d frames of the cfg2 clip in %.1f s, %d threads; %s" % (n, dt, cores, desc)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
