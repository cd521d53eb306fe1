"""Build recipe of libstanmath_cuda.so: explicit nvcc for sm_100a, in-tree output.

The library is plain CUDA C++ behind the C ABI of include/stan_cuda.h: no torch, no Stan, no
Eigen headers.  cudart is linked statically; NCCL is dlopen'ed at run time (csrc/comm.cu).
"""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "math_b200", "csrc")
LIB_DIR = os.path.join(ROOT, "math_b200", "lib")
LIB = os.path.join(LIB_DIR, "libstanmath_cuda.so")
SOURCES = ["glm_fused.cu", "glm_warp.cu", "glm_categorical.cu", "glm_categorical_mma.cu", "glm_categorical_wide.cu", "glm_ordered_wide.cu", "normal_lpdf.cu", "student_t.cu", "density.cu", "gemm.cu", "cholesky.cu", "gp.cu", "data.cu", "context.cu", "comm.cu", "api.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-diag-suppress", "177"]


def nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libstanmath_cuda.so")
    return exe


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def cta_kernel_flags():
    """The superseded CTA-synchronous kernel (csrc/glm_fused.cu) is compiled in only on request."""
    return ("-DSTAN_CUDA_WITH_CTA_KERNEL",) if os.environ.get("STAN_CUDA_WITH_CTA_KERNEL") == "1" else ()


def build(force=False, verbose=False, variant="", extra_flags=()):
    """Compile every .cu into an object (parallel) and link the shared library.  `variant` names a
    debugging build (e.g. "timers" with -DSTAN_CUDA_PHASE_TIMERS, used by tools/phase_timers.py);
    it goes to libstanmath_cuda_<variant>.so and never replaces the product library."""
    os.makedirs(LIB_DIR, exist_ok=True)
    LIB = os.path.join(LIB_DIR, "libstanmath_cuda%s.so" % ("_" + variant if variant else ""))
    obj_dir = os.path.join(LIB_DIR, "obj" + ("_" + variant if variant else ""))
    os.makedirs(obj_dir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "stan_cuda.h"))
    procs, objs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(obj_dir, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + headers):
            cmd = [nvcc()] + NVCC_FLAGS + list(cta_kernel_flags()) + list(extra_flags) + ["-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd))
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (src, out.decode()))
        if verbose and out:
            print(out.decode())
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-ldl"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s" % r.stdout.decode())
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in os.sys.argv, verbose=True))
