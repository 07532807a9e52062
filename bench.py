#!/usr/bin/env python
"""bench.py -- headline benchmark of the segment compress hot path (BASELINE.json: 1080p->720p H.264 frames/sec).

A "step" is one pass of the hot path over the cfg2 workload on one GPU: the 10 s 1080p30 clip split into 8 segments of 38
pictures (1920x1080 I420), all 8 in flight on the GPU at once (one encoder object and CUDA stream per segment, the way 8
reference workers would share a box): resize to 1280x720 (bicubic), then the H.264 encode kernels (IDR + 37 P pictures
per segment: lowres, motion search, motion compensation + transform/quant/reconstruction, border extension, half-sample
planes).  One process per GPU; with N GPUs every rank encodes its own clip (segments share nothing: weak scaling, no
collective on the data path).

  value   pictures/s over all ranks, source pictures resident in HBM, every GPU stage, CUDA events on the encoder's stream
  e2e     the same metric through the C ABI with pinned HOST pictures in and the H.264 stream out
          (gvc_encoder_encode_host: H2D + kernels + D2H of the macroblock records + CAVLC on host threads)
  roofline / stages / cpu_baseline / clocks: see DESIGN.md "Measurement"

`--impl reference` times the CPU implementation of the same path on the host cores: FFmpeg's libswscale (the image's
bundled build) for the resize and the oracle port (oracle/*.c) for the x264-semantics stages, one short segment per core.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

SW, SH, DW, DH = 1920, 1080, 1280, 720
SEG_FRAMES = 38                      # ceil(300 pictures / 8 segments)
NSEG = 8                             # segments of the clip, all in flight on one GPU
CRF = 23
NMB = (DW // 16) * (DH // 16)
METRIC = "1080p->720p H.264 frames/sec"
CONFIG = {"workload": "cfg2: 10 s 1080p30 synthetic testsrc-style clip (SURVEY.md 8d) in 8 segments of 38 pictures, the whole clip per GPU per step (8 segments "
                      "in flight), 1920x1080 -> 1280x720 bicubic + H.264 (IDR+P, crf 23, me_range 16)",
          "segment_frames": SEG_FRAMES, "segments_per_step": NSEG, "crf": CRF,
          "l2": "inputs larger than L2: 8 source segments = 945 MB per step, each read once",
          "parallelism": "clip-per-GPU (8 independent segments each), no collective"}
# ALGORITHMIC bytes per launch (SURVEY.md 8(d) formulas at the coded size 1280x720; DESIGN.md "Measurement")
WH = DW * DH
ALGO_BYTES = {"resize": int(1.5 * (SW * SH + WH)) * SEG_FRAMES,     # one launch covers the segment
              "hpel": 4 * WH, "lowres": WH + WH // 4, "me_lowres": WH // 2 + 4 * NMB, "me_refine": 5 * WH + 24 * NMB,
              "encode_inter": int(7.5 * WH), "encode_intra": 6 * WH, "border": None, "entropy": None}


def synth_segment(seed, n=SEG_FRAMES):
    """SURVEY.md 8(d) synthetic testsrc-style clip (colour bars, translating gradient, orbiting box, LCG-noise patch), I420."""
    import numpy as np
    from util_inputs import make_frame
    seg = np.empty((n, SW * SH * 3 // 2), np.uint8)
    for f in range(n):
        seg[f] = np.concatenate([p.reshape(-1) for p in make_frame(SW, SH, "yuv420p", "testsrc", seed, t=f)])
    return seg


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML every ~2 ms while running."""

    def __init__(self, index):
        self.index, self.rows, self._stop, self._t, self.nv = index, [], False, None, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
            return
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def _run(self):
        nv = self.nv
        while not self._stop:
            try:
                self.rows.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                  nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:
                pass
            time.sleep(0.002)

    def stop(self):
        self._stop = True
        if self._t:
            self._t.join()
        if not self.nv or not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        nv = self.nv
        sm = sorted(r[0] for r in self.rows)
        bits = 0
        for r in self.rows:
            bits |= r[1]
        names = [("hw_slowdown", getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8)),
                 ("hw_thermal_slowdown", getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40)),
                 ("sw_thermal_slowdown", getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20)),
                 ("sw_power_cap", getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4))]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.max, "reasons": [n for n, b in names if bits & b],
                "samples": len(sm)}


_CPU_SEG = {}


def _cpu_worker(args):
    """One whole segment of the workload through the CPU path (runs in a worker process; the segment's source pictures are
    made once per process and kept, so that only the path itself is timed)."""
    seed, n = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import oracle_segment
    if (seed, n) not in _CPU_SEG:
        _CPU_SEG.clear()
        _CPU_SEG[(seed, n)] = synth_segment(seed, n)
    frames = _CPU_SEG[(seed, n)]
    try:
        import sws_ref
        sws_ref.load()
        real = True
    except Exception:
        real = False
    t0 = time.perf_counter()
    oracle_segment.encode_segment(frames, SW, SH, DW, DH, qp_i=CRF - 3, qp_p=CRF, merange=16, use_libswscale=real)
    return time.perf_counter() - t0, real


class CpuArm:
    """`cores` worker processes, each encoding its own whole segment (SEG_FRAMES pictures: 1 IDR + P) per step, side by side:
    pictures/s of the whole host.  The pool lives across steps so every step times the path only."""

    def __init__(self, cores, frames_per_segment=SEG_FRAMES):
        from concurrent.futures import ProcessPoolExecutor
        import multiprocessing as mp
        self.cores, self.n = cores, frames_per_segment
        self.ex = ProcessPoolExecutor(cores, mp_context=mp.get_context("spawn"))
        self.real = True

    def step(self):
        t0 = time.perf_counter()
        res = list(self.ex.map(_cpu_worker, [(s, self.n) for s in range(self.cores)], chunksize=1))
        wall = time.perf_counter() - t0
        busy = max(r[0] for r in res)              # workers run side by side; the slowest one bounds the throughput
        self.real = all(r[1] for r in res)
        return self.cores * self.n / busy, busy, wall

    def describe(self):
        return ("%d processes x 1 segment of %d pictures (1 IDR + %d P) of the cfg2 clip per step; resize by %s, all other stages by "
                "the oracle port (oracle/*.c, -O3, no hand-written SIMD); no entropy coding on this arm" %
                (self.cores, self.n, self.n - 1, "FFmpeg libswscale 9.1.100 (bit-exact flags)" if self.real else "the oracle port"))

    def close(self):
        self.ex.shutdown()


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    arm = CpuArm(cores)
    arm.step()                                      # builds every worker's source segment; also the first warm-up step
    for _ in range(max(0, args.warmup - 1)):
        arm.step()
    vals, busy = [], []
    for _ in range(args.steps):
        v, b, _ = arm.step()
        vals.append(v)
        busy.append(b)
    desc = arm.describe()
    arm.close()
    v = cores * SEG_FRAMES * len(busy) / sum(busy)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(busy) / len(busy),
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                      "config": CONFIG,
                      "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": desc},
                      "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gvc", choices=["gvc", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch kernels one by one instead of replaying CUDA graphs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:     # before CUDA is initialised in this process
        cores = os.cpu_count() or 1
        arm = CpuArm(cores)
        arm.step()                                  # makes the source segments (untimed), warms the workers
        v, busy, wall = arm.step()
        cpu = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": arm.describe() + "; one step, %.1f s" % busy}
        arm.close()

    import numpy as np
    import torch
    import torch.distributed as dist
    import govideocompressor_b200 as gvc

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # the version banner goes to stdout: rank 0 prints ONE JSON line and nothing else
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gvc.init(local)
    cores = os.cpu_count() or 1
    # The clip's NSEG segments are encoded as NENC batches of NSEG / NENC segments: within a batch the segments advance picture by
    # picture together, one kernel launch per stage for all of them (one CUDA graph per batch, replayed per step); the batches run
    # on their own streams, each issued by its own host thread, so one batch's dependent chain fills the other's gaps.
    NENC = int(os.environ.get("GVC_BENCH_ENCODERS", "2"))
    assert NSEG % NENC == 0
    PER = NSEG // NENC
    encs = [gvc.Encoder(SW, SH, DW, DH, fps=(30, 1), crf=CRF, merange=16, host_threads=max(1, cores // NENC), disable_graph=args.no_graph, max_batch=PER)
            for _ in range(NENC)]
    enc = encs[0]
    fb = enc.frame_bytes()

    seg0 = synth_segment(rank)
    host_src = [torch.from_numpy(seg0 if b == 0 else np.roll(seg0, 5 * b, axis=0)).pin_memory() for b in range(NSEG)]
    dev_src = [h.cuda() for h in host_src]
    counts = [SEG_FRAMES] * PER
    dev_ptrs = [[d.data_ptr() for d in dev_src[k * PER:(k + 1) * PER]] for k in range(NENC)]
    host_ptrs = [[h.data_ptr() for h in host_src[k * PER:(k + 1) * PER]] for k in range(NENC)]
    from concurrent.futures import ThreadPoolExecutor
    pool = ThreadPoolExecutor(NENC)                  # one issuing thread per batch (ctypes releases the GIL)

    def barrier():
        torch.cuda.synchronize()
        for e in encs:
            e.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_device(steps):
        list(pool.map(lambda k: [encs[k].encode_device_batch(dev_ptrs[k], counts) for _ in range(steps)], range(NENC)))

    # end to end: the upload of a batch's next step (copy stream, second staging buffer) overlaps the kernels of its current one;
    # finish waits for the step's slices and wraps them into NAL units on the host
    def slot_host(k, steps):
        e = encs[k]
        e.upload_host_batch(host_ptrs[k], counts)
        for i in range(steps):
            e.encode_uploaded()
            if i + 1 < steps:
                e.upload_host_batch(host_ptrs[k], counts)
            e.finish_batch(keep=False)

    def run_host(steps):
        list(pool.map(lambda k: slot_host(k, steps), range(NENC)))

    # ---- device-resident timing: every GPU stage of the path, the whole clip per step -------
    run_device(args.warmup)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(NENC)]
    exts = [torch.cuda.ExternalStream(e.stream_handle()) for e in encs]
    ext, (e0, e1) = exts[0], ev[0]
    t0 = time.perf_counter()
    for k in range(NENC):
        ev[k][0].record(exts[k])
    run_device(args.steps)
    for k in range(NENC):
        ev[k][1].record(exts[k])
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    # device time of the step = the slowest stream's span (all streams start together); wall clock is reported beside it
    ms_total = max(ev[k][0].elapsed_time(ev[k][1]) for k in range(NENC))
    from govideocompressor_b200 import farm
    ms_total = farm.max_over_ranks(ms_total, "cuda")
    value = world * NSEG * SEG_FRAMES * args.steps / (ms_total * 1e-3)

    # ---- stage pass: the same batched launches with a CUDA-event pair around every one (roofline figures) -------
    enc.set_timing(True)
    for i in range(args.steps):
        enc.encode_device_batch(dev_ptrs[0], counts)
    stages = enc.stage_times()
    enc.set_timing(False)
    # one segment alone on the GPU (the reference's one-piece-per-worker model)
    enc.encode_device(dev_ptrs[0][0], SEG_FRAMES)
    enc.sync()
    e0.record(ext)
    for i in range(args.steps):
        enc.encode_device(dev_src[i % NSEG].data_ptr(), SEG_FRAMES)
    e1.record(ext)
    enc.sync()
    one_segment_ms = e0.elapsed_time(e1) / args.steps

    # ---- end to end through the C ABI: host pictures in, H.264 streams out ---
    run_host(1)
    barrier()
    t0 = time.perf_counter()
    run_host(args.steps)
    e2e_s = time.perf_counter() - t0
    stream_bytes, d2h_bytes = 0, 0
    for e in encs:
        for k in range(PER):
            e.select(k)
            stream_bytes += int(e.stats()[:, 3].sum())     # the step's result as read on the host
            d2h_bytes += e.record_bytes()
    e2e = world * NSEG * SEG_FRAMES * args.steps / farm.max_over_ranks(e2e_s, "cuda")

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        table, launches = {}, 0
        for name, (ms, n) in stages.items():
            launches += n * (4 if name == "entropy" else PER if name == "resize" else 1)    # entropy: four kernels per entry; resize: one launch per segment
            if n == 0:
                continue
            avg = ms / n
            b = ALGO_BYTES.get(name)
            if name == "resize":
                b = b * PER                         # the entry brackets the batch's resize launches (one per segment)
            elif b:
                b = b * PER                         # one launch covers one picture of every segment of the batch
            table[name] = {"launches_per_step": n / args.steps, "ms_per_step": ms / args.steps, "avg_launch_ms": avg,
                           "share_of_kernel_time": None, "algorithmic_bytes_per_launch": b,
                           "achieved_gbs": (b / (avg * 1e-3) / 1e9) if b else None}
        ktotal = sum(s["ms_per_step"] for s in table.values())
        for s in table.values():
            s["share_of_kernel_time"] = s["ms_per_step"] / ktotal
            if s["achieved_gbs"] is not None:
                s["frac_of_peak"] = s["achieved_gbs"] / peak
        # dominant single kernel ("entropy" is four small kernels timed as one entry, "border" has no byte model)
        dom = max((k for k in table if table[k]["algorithmic_bytes_per_launch"]), key=lambda k: table[k]["ms_per_step"])
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get(dom)
        d = table[dom]
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": CONFIG,
                "wall_ms_per_step": wall_ms / args.steps,
                "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": NSEG * SEG_FRAMES * fb,
                        "d2h_bytes_per_step": d2h_bytes,
                        "stream_bytes_per_step": stream_bytes, "host_threads": cores},
                # stage pass counts + the one-thread kernel that raises each picture's arrival flag (not a timed stage)
                "gpu_launches": (launches + SEG_FRAMES * args.steps) * NENC, "batches": NENC, "segments_per_batch": PER, "clocks": clocks, "one_segment_alone_ms": one_segment_ms,
                "roofline": {"kernel": dom, "bound": "hbm", "achieved": d["achieved_gbs"], "peak": peak, "unit": "GB/s",
                             "frac": (d["achieved_gbs"] / peak) if d["achieved_gbs"] else None, "traffic": traffic,
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": d["algorithmic_bytes_per_launch"],
                             "launch_ms_avg": d["avg_launch_ms"], "share_of_step": d["share_of_kernel_time"],
                             "note": "dominant single kernel by accumulated CUDA-event time in the stage pass: the same batched launches as the timed "
                                     "region (one launch per stage for the 8 segments), each bracketed by its own CUDA-event pair; the timed "
                                     "region replays them from one CUDA graph.  The kernel is instruction-issue bound, not HBM bound: see "
                                     "profiles/ and DESIGN.md"},
                "stages": table}
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    for e in encs:
        e.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
