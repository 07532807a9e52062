#!/usr/bin/env python
"""gvc_transcode_file on cfg2's .mp4 pieces with different libavcodec thread counts (GVC_DECODE_THREADS is read when a
piece is opened, so one process can sweep it).  Prints pictures/s per setting; pan = the textured clip (bigger pieces)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
import govideocompressor_b200 as gvc  # noqa: E402


def main():
    clip = sys.argv[1] if len(sys.argv) > 1 else "testsrc"
    c = bench.workload("cfg2", clip)
    gvc.init(0)
    segs = [bench.synth_segment(c, 100 + i) for i in range(8)]
    work = os.path.join(ROOT, "gpurun_out", "exp_decode")
    os.makedirs(work, exist_ok=True)
    enc = gvc.Encoder(c["sw"], c["sh"], fps=(30, 1), crf=12)
    paths = []
    for i, seg in enumerate(segs):
        enc.encode_host(seg)
        p = os.path.join(work, "%d.mp4" % i)
        enc.write_mp4(p)
        paths.append(p)
    enc.close()
    print("piece bytes", [os.path.getsize(p) for p in paths][:3])
    args = "-threads 4 -vcodec libx264 -crf 23 -s %dx%d" % (c["dw"], c["dh"])
    gvc.transcode_file(paths[0], args, os.path.join(work, "c.mp4"))
    for thr in ("1", "2", "4", "8", "12", "0"):
        os.environ["GVC_DECODE_THREADS"] = thr
        best = 0.0
        for _ in range(2):
            t0 = time.perf_counter()
            n = 0
            for i, p in enumerate(paths):
                rc, log = gvc.transcode_file(p, args, os.path.join(work, "c%d.mp4" % i))
                assert rc == 0, log
                n += segs[i].shape[0]
            best = max(best, n / (time.perf_counter() - t0))
        print("threads", thr, "pictures/s", round(best))


if __name__ == "__main__":
    main()
