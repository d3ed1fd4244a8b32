"""In-tree build of the CUDA library: ``python -m yat_b200.build``.

Plain ``nvcc`` for sm_100a only (no multi-arch fat binary, no JIT cache): the resulting
``yat_b200/libyat_b200.so`` exports the C-ABI declared in ``include/yat_b200.h`` and is what
``yat_b200._lib`` loads with ctypes.  The CUDA runtime is linked statically and the driver's
tensor-map encoder is resolved at run time, so the library loads on a box without a GPU.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libyat_b200.so")
SOURCES = ["api.cu", "noise.cu", "loss.cu", "repa.cu", "pad.cu", "linattn_simt.cu", "linattn_sm100.cu",
           "linattn_bwd_sm100.cu", "block_ops.cu", "glumb.cu", "optim.cu", "gemm_sm100.cu", "qkv_linattn_sm100.cu", "xattn_sm100.cu", "xattn_bwd_sm100.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; cannot build yat_b200")
    return nvcc


def _stale(srcs):
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps.append(os.path.join(HERE, "..", "include", "yat_b200.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    if not force and not _stale(srcs):
        return OUT
    tmp = f"{OUT}.{os.getpid()}.tmp"          # link to a private name, then rename: readers never see a partial file
    cmd = [_nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", "-shared", "-Xcompiler", "-fPIC",
           "--threads", "0", "-o", tmp, *srcs]
    extra = os.environ.get("YAT_NVCC_FLAGS")     # experiments only (e.g. -DYAT_BWD_PROBE); the release build sets nothing
    if extra:
        cmd[1:1] = extra.split()
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose:
        sys.stderr.write(r.stderr)
    if r.returncode != 0:
        if os.path.exists(tmp):
            os.remove(tmp)
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    os.replace(tmp, OUT)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
