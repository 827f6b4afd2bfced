"""Builds libmpcb200.so (hand-written sm_100a CUDA + the C ABI) in-tree with nvcc.  No torch involved.

kernels.cu is compiled once per (dtype, env): float32 instantiations with FMA contraction, float64 instantiations with
-fmad=false so that the float64 (parity) path keeps the reference's rounding sequence operation by operation; the
Acrobot float32 unit is split once more (per-stage trig splits).  The units build in parallel."""
import glob
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libmpcb200.so")
# plain-C latency probe of the C ABI (tools/cabi_latency.c): bench.py runs it beside the Python wrapper's p50
CABI_SRC = os.path.join(HERE, "..", "tools", "cabi_latency.c")
CABI_BIN = os.path.join(OBJ, "cabi_latency")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
# (object name, source, extra flags)
F64 = ["-DMPC_REAL=double", "-DMPC_IS_F64", "-fmad=false"]
UNITS = [("kernels_f32_e%d.o" % e, "kernels.cu", ["-DMPC_REAL=float", "-DMPC_ENV_ONLY=%d" % e, "-DMPC_SUFFIX=_f32_e%d" % e])
         for e in (2, 1, 0)]
UNITS += [("kernels_f32_e2s.o", "kernels.cu", ["-DMPC_REAL=float", "-DMPC_ENV_ONLY=2", "-DMPC_TRIG_GROUP=1",
                                               "-DMPC_SUFFIX=_f32_e2s"])]
UNITS += [("kernels_f64_e%d.o" % e, "kernels.cu", F64 + ["-DMPC_ENV_ONLY=%d" % e, "-DMPC_SUFFIX=_f64_e%d" % e])
          for e in (2, 1, 0)]
UNITS += [("dispatch.o", "dispatch.cu", []), ("capi.o", "capi.cu", []), ("noalloc.o", "noalloc.cu", [])]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libmpcb200.so cannot be built (there is no CPU fallback)")


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    # every source and header of the library (globbed: a new .cuh can never be forgotten here)
    deps = glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [os.path.abspath(__file__)]
    if shutil.which("gcc") and os.path.exists(CABI_SRC):
        if not os.path.exists(CABI_BIN):
            return True
        deps.append(CABI_SRC)
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
        if verbose:
            with open(os.path.join(OBJ, obj + ".log"), "w") as f:
                f.write(r.stdout + r.stderr)
        return obj, r.returncode, r.stdout + r.stderr

    with ThreadPoolExecutor(min(len(UNITS), os.cpu_count() or 4)) as ex:
        results = list(ex.map(compile_unit, UNITS))
    for obj, rc, out in results:
        if rc != 0:
            print("==", obj, "\n" + out)
        if rc != 0:
            raise RuntimeError("nvcc failed on %s" % obj)
    subprocess.check_call([nvcc] + ARCH + ["-shared", "-o", LIB] + [os.path.join(OBJ, u[0]) for u in UNITS] + ["-ldl"])
    gcc = shutil.which("gcc")
    if gcc and os.path.exists(CABI_SRC):  # a consumer of the C ABI in plain C; found next to the library by $ORIGIN
        subprocess.check_call([gcc, "-O2", "-std=c99", "-I", os.path.join(HERE, "..", "include"), CABI_SRC, "-o", CABI_BIN,
                               "-L", HERE, "-lmpcb200", "-Wl,-rpath,$ORIGIN/..", "-lm"])
    return LIB


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
