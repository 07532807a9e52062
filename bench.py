#!/usr/bin/env python
"""bench.py -- headline benchmark of the segment compress hot path (BASELINE.json: 1080p->720p H.264 frames/sec).

A "step" is one pass of the hot path over one GPU's share of the configured workload (default cfg2: the 10 s 1080p30 clip
split into 8 segments of 38 pictures, 1920x1080 I420): resize (bicubic) and H.264 encode (IDR + P pictures per segment:
lowres, motion search, motion compensation + transform/quant/reconstruction, border extension, half-sample planes, CAVLC).
The segments of a step are encoded as batches: within a batch they advance picture by picture together, ONE kernel launch
per stage for all of them.  One process per GPU; with N GPUs every rank encodes its own clip (segments share nothing: weak
scaling, no collective on the data path).

  value     pictures/s over all ranks, source pictures resident in HBM, every GPU stage, CUDA events on the encoders' streams
  e2e       the same metric through the C ABI with pinned HOST pictures in and the H.264 streams out
  e2e_file  the call the Go worker makes (gvc_transcode_file) on the workload's pieces as .mp4 files: host demux + H.264
            decode by FFmpeg's libraries, H2D, kernels, MP4 out (cfg2 only; one piece at a time, like one worker)
  quality   PSNR / SSIM (x264's pixel_ssd / pixel_ssim, on the device) of the reconstruction against the resized source,
            and the bitrate, for the benchmarked clip
  roofline / stages / cpu_baseline / clocks: see DESIGN.md "Measurement"

--config cfg1|cfg2|cfg3|cfg4 selects BASELINE.json's configurations (the metric is quoted on cfg2, the default);
--clip testsrc|pan the content (SURVEY 8(d) testsrc-style clip, or a globally translating texture where no macroblock
leaves the motion search early).
`--impl reference` times the CPU implementation of the same path on the host cores: FFmpeg's libswscale (the image's
bundled build) for the resize and the oracle port (oracle/*.c) for the x264-semantics stages, one whole segment per core.
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

CRF = 23
METRIC = "1080p->720p H.264 frames/sec"
# BASELINE.json configs made concrete; per GPU and step: `nseg` segments of `seg` pictures
CONFIGS = {
    "cfg1": dict(sw=1920, sh=1080, dw=1920, dh=1080, seg=300, nseg=1, merange=16,
                 what="cfg1: 10 s 1080p30 synthetic clip as ONE segment of 300 pictures, no resize, H.264 (IDR+P, crf 23, me_range 16)"),
    "cfg2": dict(sw=1920, sh=1080, dw=1280, dh=720, seg=38, nseg=8, merange=16,
                 what="cfg2: 10 s 1080p30 synthetic clip (SURVEY.md 8d) in 8 segments of 38 pictures, the whole clip per GPU per step, "
                      "1920x1080 -> 1280x720 bicubic + H.264 (IDR+P, crf 23, me_range 16)"),
    "cfg3": dict(sw=3840, sh=2160, dw=1920, dh=1080, seg=56, nseg=8, merange=32,
                 what="cfg3: 60 s 4K60 synthetic clip in 64 segments of 56 pictures, 8 segments per GPU per step, 3840x2160 -> 1920x1080 "
                      "bicubic + H.264 (IDR+P, crf 23, me_range 32)"),
    "cfg4": dict(sw=1280, sh=720, dw=854, dh=480, seg=60, nseg=32, merange=16,
                 what="cfg4: 256 clips of 2 s 720p30 (60 pictures), 32 clips per GPU per step, 1280x720 -> 854x480 bicubic + H.264 "
                      "(IDR+P, crf 23, me_range 16)"),
}


def workload(name, clip):
    c = dict(CONFIGS[name])
    c["name"], c["clip"] = name, clip
    c["w16"], c["h16"] = (c["dw"] + 15) // 16 * 16, (c["dh"] + 15) // 16 * 16
    c["nmb"] = (c["w16"] // 16) * (c["h16"] // 16)
    c["frame_bytes"] = c["sw"] * c["sh"] * 3 // 2
    return c


def config_json(c, batches=None):
    src_mb = c["nseg"] * c["seg"] * c["frame_bytes"] / 1e6
    out = {"workload": c["what"], "config": c["name"], "clip": c["clip"], "segment_frames": c["seg"], "segments_per_step": c["nseg"], "crf": CRF,
           "l2": "inputs larger than L2: %.0f MB of source pictures per step, each read once" % src_mb,
           "parallelism": "clip-per-GPU (%d independent segments each), no collective" % c["nseg"]}
    return out


def algo_bytes(c):
    """ALGORITHMIC bytes per picture and stage (SURVEY.md 8(d) formulas at the coded size; DESIGN.md "Measurement")."""
    wh, nmb = c["w16"] * c["h16"], c["nmb"]
    resize = int(1.5 * (c["sw"] * c["sh"] + c["dw"] * c["dh"])) if (c["sw"], c["sh"]) != (c["dw"], c["dh"]) else None
    return {"resize": resize, "hpel": 4 * wh, "lowres": wh + wh // 4, "me_lowres": wh // 2 + 4 * nmb, "me_refine": 5 * wh + 24 * nmb,
            "encode_inter": int(7.5 * wh), "encode_intra": 6 * wh, "border": None, "entropy": None}


def synth_segment(c, seed, n=None):
    """One segment of the configured clip, I420: SURVEY.md 8(d) testsrc-style content (colour bars, translating gradient,
    orbiting box, LCG-noise patch) or the translating texture."""
    import numpy as np
    from util_inputs import make_frame
    n = n or c["seg"]
    seg = np.empty((n, c["frame_bytes"]), np.uint8)
    for f in range(n):
        seg[f] = np.concatenate([p.reshape(-1) for p in make_frame(c["sw"], c["sh"], "yuv420p", c["clip"], seed, t=f)])
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
    name, clip, seed, n = args
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import oracle_segment
    c = workload(name, clip)
    if args not in _CPU_SEG:
        _CPU_SEG.clear()
        _CPU_SEG[args] = synth_segment(c, seed, n)
    frames = _CPU_SEG[args]
    try:
        import sws_ref
        sws_ref.load()
        real = True
    except Exception:
        real = False
    t0 = time.perf_counter()
    oracle_segment.encode_segment(frames, c["sw"], c["sh"], c["dw"], c["dh"], qp_i=CRF - 3, qp_p=CRF, merange=c["merange"], use_libswscale=real)
    return time.perf_counter() - t0, real


class CpuArm:
    """`cores` worker processes, each encoding its own whole segment of the workload (1 IDR + P pictures) per step, side by
    side: pictures/s of the whole host.  The pool lives across steps so every step times the path only.  Segments longer than
    `max_frames` pictures are cut to that many (the rate per picture of a P-picture run does not depend on its length)."""

    def __init__(self, c, cores, max_frames=38):
        from concurrent.futures import ProcessPoolExecutor
        import multiprocessing as mp
        self.c, self.cores, self.n = c, cores, min(c["seg"], max_frames)
        self.ex = ProcessPoolExecutor(cores, mp_context=mp.get_context("spawn"))
        self.real = True

    def step(self):
        t0 = time.perf_counter()
        res = list(self.ex.map(_cpu_worker, [(self.c["name"], self.c["clip"], s, self.n) for s in range(self.cores)], chunksize=1))
        wall = time.perf_counter() - t0
        busy = max(r[0] for r in res)              # workers run side by side; the slowest one bounds the throughput
        self.real = all(r[1] for r in res)
        return self.cores * self.n / busy, busy, wall

    def describe(self):
        return ("%d processes x 1 segment of %d pictures (1 IDR + %d P) of the %s %s clip per step; resize by %s, all other stages by "
                "the oracle port (oracle/*.c, -O3, no hand-written SIMD); no entropy coding on this arm" %
                (self.cores, self.n, self.n - 1, self.c["name"], self.c["clip"],
                 "FFmpeg libswscale 9.1.100 (bit-exact flags)" if self.real else "the oracle port"))

    def close(self):
        self.ex.shutdown()


def run_reference(args, rank):
    if rank != 0:
        return
    c = workload(args.config, args.clip)
    cores = os.cpu_count() or 1
    arm = CpuArm(c, cores)
    arm.step()                                      # builds every worker's source segment; also the first warm-up step
    for _ in range(max(0, args.warmup - 1)):
        arm.step()
    busy = []
    for _ in range(args.steps):
        busy.append(arm.step()[1])
    desc = arm.describe()
    n = arm.n
    arm.close()
    v = cores * n * len(busy) / sum(busy)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(busy) / len(busy),
                      "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                      "config": config_json(c),
                      "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port", "sample": desc},
                      "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def measure_quality(gvc, torch, np, c, seg_frames, n_pictures=None):
    """PSNR-Y/U/V and SSIM-Y (mean over the pictures, as x264 reports them) of one segment's reconstruction against the resized
    source, and its bitrate at 30 pictures/s -- all pictures of the segment unless n_pictures says fewer."""
    n = min(n_pictures or seg_frames.shape[0], seg_frames.shape[0])
    frames = seg_frames[:n]
    enc = gvc.Encoder(c["sw"], c["sh"], c["dw"], c["dh"], fps=(30, 1), crf=CRF, merange=c["merange"], keep_recon=True)
    stream = enc.encode_host(frames, annexb=False)
    recon = torch.from_numpy(enc.recon()).cuda()
    w, h, w16, h16 = c["dw"], c["dh"], enc.w16, enc.h16
    src = torch.from_numpy(frames).cuda()
    if (c["sw"], c["sh"]) != (w, h):
        plan = gvc.ResizePlan(c["sw"], c["sh"], w, h)
        ref = torch.zeros((n, plan.frame_bytes(w, h)), dtype=torch.uint8, device="cuda")
        plan.resize_device(src, ref, n)
        torch.cuda.synchronize()
        plan.close()
    else:
        ref = src
    cw, ch = w // 2, h // 2
    acc = {"psnr_y": [], "psnr_u": [], "psnr_v": [], "ssim_y": []}
    for f in range(n):
        a, b = ref[f].data_ptr(), recon[f].data_ptr()
        _, ssim, py = gvc.psnr_ssim(a, w, b, w16, w, h)
        _, _, pu = gvc.psnr_ssim(a + w * h, cw, b + w16 * h16, w16 // 2, cw, ch)
        _, _, pv = gvc.psnr_ssim(a + w * h + cw * ch, cw, b + w16 * h16 * 5 // 4, w16 // 2, cw, ch)
        acc["psnr_y"].append(py); acc["psnr_u"].append(pu); acc["psnr_v"].append(pv); acc["ssim_y"].append(ssim)
    enc.close()
    out = {k: float(np.mean(v)) for k, v in acc.items()}
    out["kbit_per_s"] = len(stream) * 8 / 1000.0 / (n / 30.0)
    out["pictures"] = n
    out["note"] = ("mean over the pictures of one segment (1 IDR + P), reconstruction vs the resized source, x264's pixel_ssd / pixel_ssim "
                   "on the device (gvc_psnr_ssim); no libx264 exists on this box to compare with (tier 2 unmeasured)")
    return out


def measure_transcode_file(gvc, c, segments, workdir):
    """gvc_transcode_file -- the one call the Go worker makes -- on the step's pieces as <id>.mp4 files (H.264, made here with
    this encoder at crf 12), one piece at a time like one worker: pictures/s including host demux + H.264 decode."""
    import shutil
    os.makedirs(workdir, exist_ok=True)
    try:
        enc = gvc.Encoder(c["sw"], c["sh"], fps=(30, 1), crf=12)
        paths = []
        for i, seg in enumerate(segments):
            enc.encode_host(seg)
            p = os.path.join(workdir, "%d.mp4" % i)
            enc.write_mp4(p)
            paths.append(p)
        enc.close()
        args = "-threads 4 -vcodec libx264 -crf %d" % CRF + (" -s %dx%d" % (c["dw"], c["dh"]) if (c["sw"], c["sh"]) != (c["dw"], c["dh"]) else "")
        rc, log = gvc.transcode_file(paths[0], args, os.path.join(workdir, "c0.mp4"))      # warm-up: encoder + pinned memory are created here
        if rc != 0:
            return {"unavailable": log.strip().splitlines()[-1] if log else "status %d" % rc}
        t0 = time.perf_counter()
        n = 0
        for i, p in enumerate(paths):
            rc, log = gvc.transcode_file(p, args, os.path.join(workdir, "c%d.mp4" % i))
            if rc != 0:
                return {"unavailable": log.strip().splitlines()[-1]}
            n += segments[i].shape[0]
        dt = time.perf_counter() - t0
        return {"value": n / dt, "unit": "frames/s", "pieces": len(paths), "args": args,
                "note": "gvc_transcode_file on .mp4 pieces, one at a time (one worker process): host demux + H.264 decode (FFmpeg's libavcodec, "
                        "frame threads) + H2D + kernels + MP4 out"}
    finally:
        shutil.rmtree(workdir, ignore_errors=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gvc", choices=["gvc", "reference"])
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--clip", default="testsrc", choices=["testsrc", "pan"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the quality and e2e_file measurements")
    ap.add_argument("--no-graph", action="store_true", help="launch kernels one by one instead of replaying CUDA graphs")
    ap.add_argument("--strong", action="store_true", help="strong scaling: ONE clip per step for the whole job, its segments dealt out to the ranks "
                                                           "(govideocompressor_b200.farm.segments_for_rank: highest index first, like the dispatcher)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)
    c = workload(args.config, args.clip)
    cores = os.cpu_count() or 1
    # every rank gets its own share of the host cores BEFORE it allocates pinned memory or starts threads (first-touch placement
    # of the staging buffers, no migration of the issuing threads between ranks' cores)
    if world > 1 and hasattr(os, "sched_setaffinity"):
        try:
            avail = sorted(os.sched_getaffinity(0))
            per = max(1, len(avail) // world)
            os.sched_setaffinity(0, set(avail[local * per:(local + 1) * per]) or set(avail))
            cores = per
        except OSError:
            pass

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:     # before CUDA is initialised in this process
        arm = CpuArm(c, cores)
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
        # NCCL prints its version banner (NCCL_DEBUG=VERSION) on the C-level stdout when the communicator comes up; stdout carries
        # the ONE JSON line, so file descriptor 1 points at stderr while that happens (the environment is left as the driver set it)
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)
    gvc.init(local)
    SW, SH, DW, DH, SEG_FRAMES, NSEG = c["sw"], c["sh"], c["dw"], c["dh"], c["seg"], c["nseg"]
    from govideocompressor_b200 import farm
    TOTAL_SEG = world * NSEG                           # segments of a step over all ranks (weak scaling: every rank has its own clip)
    if args.strong:
        mine = farm.segments_for_rank(NSEG, rank, world)      # this rank's share of the ONE clip
        if not mine:
            raise SystemExit("--strong: more ranks than segments")
        TOTAL_SEG, NSEG = NSEG, len(mine)
    # The step's NSEG segments are encoded as NENC batches: within a batch the segments advance picture by picture together, one
    # kernel launch per stage for all of them (one CUDA graph per batch, replayed per step); the batches run on their own streams,
    # each issued by its own host thread, so one batch's dependent chain fills the other's gaps.
    NENC = int(os.environ.get("GVC_BENCH_ENCODERS", "2" if NSEG % 2 == 0 and NSEG >= 4 else "1"))
    assert NSEG % NENC == 0 and NSEG // NENC <= 16
    PER = NSEG // NENC
    encs = [gvc.Encoder(SW, SH, DW, DH, fps=(30, 1), crf=CRF, merange=c["merange"], host_threads=max(1, cores // NENC), disable_graph=args.no_graph, max_batch=PER)
            for _ in range(NENC)]
    enc = encs[0]
    fb = enc.frame_bytes()

    seg0 = synth_segment(c, rank)
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

    # ---- device-resident timing: every GPU stage of the path, the whole share per step -------
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
    ms_total = farm.max_over_ranks(ms_total, "cuda")
    value = TOTAL_SEG * SEG_FRAMES * args.steps / (ms_total * 1e-3)

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
    e2e = TOTAL_SEG * SEG_FRAMES * args.steps / farm.max_over_ranks(e2e_s, "cuda")
    for e in encs:
        e.close()

    quality = e2e_file = None
    if rank == 0 and not args.no_extras:
        quality = measure_quality(gvc, torch, np, c, seg0, 38)
        if args.config == "cfg2":
            e2e_file = measure_transcode_file(gvc, c, [seg0 if b == 0 else np.roll(seg0, 5 * b, axis=0) for b in range(NSEG)],
                                              os.path.join(os.environ.get("TMPDIR", "/tmp"), "gvc_bench_pieces_%d" % os.getpid()))

    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        else:
            peak, peak_src = 6650.0, "B200_PROFILING.md fallback (of fallback)"
        AB = algo_bytes(c)
        table, launches = {}, 0
        for name, (ms, n) in stages.items():
            launches += n * (4 if name == "entropy" else PER if name == "resize" else 1)    # entropy: four kernels per entry; resize: one launch per segment
            if n == 0:
                continue
            avg = ms / n
            b = AB.get(name)
            if b:
                # one launch covers one picture of every segment of the batch; a resize entry brackets the batch's launches (one per
                # segment, a segment's pictures each)
                b = b * PER * (SEG_FRAMES if name == "resize" else 1)
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
        if os.path.exists(tpath) and args.config == "cfg2":
            t = json.load(open(tpath)).get(dom)
            traffic = t * PER if isinstance(t, (int, float)) else t       # the file holds bytes per picture; a launch covers PER pictures
        d = table[dom]
        cfg = config_json(c)
        cfg.update({"batches": NENC, "segments_per_batch": PER})
        if args.strong:
            cfg["parallelism"] = "ONE clip per step, its %d segments dealt out to the %d ranks highest index first (no collective)" % (TOTAL_SEG, world)
            cfg["segments_per_step"] = TOTAL_SEG
        line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong" if args.strong else "weak",
                "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
                "wall_ms_per_step": wall_ms / args.steps,
                "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": NSEG * SEG_FRAMES * fb,
                        "d2h_bytes_per_step": d2h_bytes,
                        "stream_bytes_per_step": stream_bytes, "host_threads": cores},
                # stage pass counts + the small kernel that raises each picture's arrival flags (not a timed stage)
                "gpu_launches": (launches + SEG_FRAMES * args.steps) * NENC, "clocks": clocks, "one_segment_alone_ms": one_segment_ms,
                "roofline": {"kernel": dom, "bound": "hbm", "achieved": d["achieved_gbs"], "peak": peak, "unit": "GB/s",
                             "frac": (d["achieved_gbs"] / peak) if d["achieved_gbs"] else None, "traffic": traffic,
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": d["algorithmic_bytes_per_launch"],
                             "launch_ms_avg": d["avg_launch_ms"], "share_of_step": d["share_of_kernel_time"],
                             "note": "dominant single kernel by accumulated CUDA-event time in the stage pass: the same batched launches as the timed "
                                     "region (one launch per stage for the batch's segments), each bracketed by its own CUDA-event pair; the timed "
                                     "region replays them from one CUDA graph per batch.  The kernel is bound by the integer pipes, not by HBM: "
                                     "see DESIGN.md 5 and profiles/"},
                "stages": table}
        if cpu:
            line["cpu_baseline"] = cpu
        if quality:
            line["quality"] = quality
        if e2e_file:
            line["e2e_file"] = e2e_file
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
