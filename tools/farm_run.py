#!/usr/bin/env python
"""farm_run.py -- the reference's farm on one box: the dispatcher (server.go `c`: JobAlloc 161-191, NewConnect 267-308; here its
byte-level twin tests/protocol_twin.py, because `go` is not installed) hands the pieces of a clip, highest index first, to N
worker processes (client.go's loop; here govideocompressor_b200/worker.py), one per GPU (GVC_DEVICE = 0..N-1), each
converting its piece with gvc_transcode_file and storing the result through the local data plane.

    python tools/farm_run.py --gpus 8 [--pieces 8] [--config cfg2] [--workers-per-gpu 1]

Prints one JSON line: wall clock from the first job handed out to the last `success`, pictures/s of the farm, per-worker job
counts.  The pieces are <id>.mp4 files (H.264, made with this encoder at crf 12) of --frames pictures each."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKER = r'''
import os, sys, time
sys.path.insert(0, %(root)r)
import govideocompressor_b200 as gvc
from govideocompressor_b200 import worker
gvc.init(-1)                                    # GVC_DEVICE picks the GPU (client.go:22-36: env vars configure the worker)
done = worker.run_worker((%(host)r, %(port)d), gvc.convert, worker.local_fetch(%(src)r), worker.local_store(%(dst)r), %(work)r, max_idle=2)
print("\n".join(done))
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--workers-per-gpu", type=int, default=1)
    ap.add_argument("--pieces", type=int, default=8)
    ap.add_argument("--frames", type=int, default=38)
    ap.add_argument("--config", default="cfg2")
    ap.add_argument("--clip", default="testsrc")
    args = ap.parse_args()
    import numpy as np
    import bench
    import govideocompressor_b200 as gvc
    from protocol_twin import FakeServer
    c = bench.workload(args.config, args.clip)
    tmp = tempfile.mkdtemp(prefix="gvc_farm_")
    src, dst = os.path.join(tmp, "www"), os.path.join(tmp, "ftp")
    os.makedirs(os.path.join(src, "12clip"))
    # the pieces (made in a child process so that this one never initialises CUDA)
    maker = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import numpy as np, bench, govideocompressor_b200 as gvc\n"
             "c = bench.workload(%r, %r); gvc.init(0); seg = bench.synth_segment(c, 0, %d)\n"
             "enc = gvc.Encoder(c['sw'], c['sh'], fps=(30, 1), crf=12)\n"
             "for i in range(%d):\n    enc.encode_host(np.roll(seg, 5 * i, axis=0)); enc.write_mp4(%r %% i)\n"
             % (ROOT, os.path.join(ROOT, "tests"), args.config, args.clip, args.frames, args.pieces, os.path.join(src, "12clip", "%d.mp4")))
    subprocess.check_call([sys.executable, "-c", maker])
    ffargs = "-threads 4 -vcodec libx264" + (" -s %dx%d" % (c["dw"], c["dh"]) if (c["sw"], c["sh"]) != (c["dw"], c["dh"]) else "")
    server = FakeServer("12clip", args.pieces, ffargs)
    procs = []
    t0 = time.perf_counter()
    for g in range(args.gpus):
        for k in range(args.workers_per_gpu):
            work = os.path.join(tmp, "work%d_%d" % (g, k))
            os.makedirs(work)
            env = dict(os.environ, GVC_DEVICE=str(g))
            code = WORKER % dict(root=ROOT, host=server.addr[0], port=server.addr[1], src=src, dst=dst, work=work)
            procs.append(subprocess.Popen([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, text=True))
    ok = server.join(timeout=600)
    wall = time.perf_counter() - t0
    jobs = [p.communicate()[0].split() for p in procs]
    done = sum(len([j for j in js if j.startswith("success")]) for js in jobs)
    outs = [f for f in os.listdir(os.path.join(dst, "12clip"))] if os.path.isdir(os.path.join(dst, "12clip")) else []
    ts = sorted(server.t_success)
    k = len(procs)                                  # the first job of every worker pays its start-up (python, CUDA context, library load)
    steady = (len(ts) - k) * args.frames / (ts[-1] - ts[k - 1]) if len(ts) > k and ts[-1] > ts[k - 1] else None
    print(json.dumps({"farm": "dispatcher twin + %d worker processes on %d GPU(s)" % (len(procs), args.gpus), "config": args.config, "pieces": args.pieces,
                      "frames_per_piece": args.frames, "args": ffargs, "all_done": bool(ok) and done == args.pieces, "wall_s": wall,
                      "frames_per_s": args.pieces * args.frames / wall, "frames_per_s_after_every_worker_s_first_job": steady, "jobs_per_worker": [len(j) for j in jobs],
                      "outputs": len([f for f in outs if not f.endswith(".log")]),
                      "note": "wall clock includes the workers' start-up (python + CUDA context + library load), host H.264 decode of the pieces and MP4 writing"}))
    import shutil
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
