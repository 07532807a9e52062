"""Scratch: ncu stall breakdown of the intra kernel with and without a (never taken) ABI call."""
import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = os.path.join(ROOT, "govideocompressor_b200", "csrc", "gvc_enc.cu")
orig = open(src).read()
END = "    intra_row_done(lane, a, row_count);"
NOINL = "__device__ __noinline__ void gvc_never(int *p, int v) { if (p) atomicAdd(p, v); }\n// One warp per macroblock row; rows advance as a wavefront"
v5 = orig.replace("// One warp per macroblock row; rows advance as a wavefront", NOINL, 1).replace(END, "    if (my == 100000) gvc_never(reinterpret_cast<int *>(a.mail), lane);\n" + END)
try:
    for name, s in (("base", orig), ("v5", v5)):
        open(src, "w").write(s)
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        subprocess.run(["ncu", "--set", "full", "--clock-control", "none", "-k", "regex:intra_frame", "-s", "1", "-c", "1", "-f", "-o",
                        os.path.join(ROOT, "gpurun_out", "intra_" + name), sys.executable, os.path.join(ROOT, "tools", "profile_stage.py"), "segment", "2"],
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        print(name, "done", flush=True)
finally:
    open(src, "w").write(orig)
    subprocess.call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
