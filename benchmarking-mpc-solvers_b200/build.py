"""Builds libmpcb200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.  No torch involved.

kernels.cu is compiled twice: float32 instantiation with FMA contraction, float64 instantiation with -fmad=false so
that the float64 (parity) path keeps the reference's rounding sequence operation by operation."""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libmpcb200.so")
HEADERS = ["kernels.cuh", "envs.cuh", "philox.cuh", "vec.cuh", "mppi.cuh", os.path.join("..", "..", "include", "mpc_b200.h")]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
# (object name, source, extra flags)
UNITS = [
    ("kernels_f32.o", "kernels.cu", ["-DMPC_REAL=float", "-DMPC_SUFFIX=_f32"]),
    ("kernels_f64.o", "kernels.cu", ["-DMPC_REAL=double", "-DMPC_SUFFIX=_f64", "-DMPC_IS_F64", "-fmad=false"]),
    ("dispatch.o", "dispatch.cu", []),
    ("capi.o", "capi.cu", []),
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libmpcb200.so cannot be built (there is no CPU fallback)")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in {u[1] for u in UNITS} | set(HEADERS)] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    extra = ["-Xptxas", "-v"] if verbose else []

    def compile_unit(u):
        obj, src, flags = u
        cmd = [nvcc] + COMMON + flags + extra + ["-c", os.path.join(CSRC, src), "-o", os.path.join(OBJ, obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return obj, r.returncode, r.stdout + r.stderr

    with ThreadPoolExecutor(len(UNITS)) as ex:
        results = list(ex.map(compile_unit, UNITS))
    for obj, rc, out in results:
        if verbose or rc != 0:
            print("==", obj, "\n" + out)
        if rc != 0:
            raise RuntimeError("nvcc failed on %s" % obj)
    subprocess.check_call([nvcc] + ARCH + ["-shared", "-o", LIB] + [os.path.join(OBJ, u[0]) for u in UNITS] + ["-ldl"])
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
