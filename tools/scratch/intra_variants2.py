"""Scratch experiment: find the source change that reproduces the fast behaviour of the printf build."""
import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = os.path.join(ROOT, "govideocompressor_b200", "csrc", "gvc_enc.cu")
orig = open(src).read()
MY = "    const int my = blockIdx.x, lane = threadIdx.x;\n    const bool al = intra_aligned(a);"
assert MY in orig
END = "    intra_row_done(lane, a, row_count);"
NOINL = "__device__ __noinline__ void gvc_never(int *p, int v) { if (p) atomicAdd(p, v); }\n// One warp per macroblock row; rows advance as a wavefront"
variants = {
    "V5-noinline-call": lambda s: s.replace("// One warp per macroblock row; rows advance as a wavefront", NOINL, 1).replace(END, "    if (my == 100000) gvc_never(reinterpret_cast<int *>(a.mail), lane);\n" + END),
    "V6-printf-in-other-fn": lambda s: s.replace("// One warp per macroblock row; rows advance as a wavefront", "__device__ __noinline__ void gvc_say(int v) { printf(\"%d\\n\", v); }\n// One warp per macroblock row; rows advance as a wavefront", 1).replace(END, "    if (my == 100000) gvc_say(lane);\n" + END).replace("#include <cuda_runtime.h>", "#include <cuda_runtime.h>\n#include <cstdio>", 1),
    "V7-trap": lambda s: s.replace(END, "    if (my == 100000) __trap();\n" + END),
}
try:
    for name, fn in variants.items():
        s = fn(orig)
        open(src, "w").write(s)
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "exp_intra_time.py")], capture_output=True, text=True).stdout
        print(name, [l for l in out.splitlines() if "intra" in l], flush=True)
finally:
    open(src, "w").write(orig)
    subprocess.call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
