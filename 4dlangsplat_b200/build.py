"""Build 4dlangsplat_b200/libfdls.so (the C-ABI library of include/fdls_abi.h) with nvcc for sm_100a ONLY.

No PTX fallback, no other architecture, no fast-math (the reference builds without it, setup.py:33, and the parity
contract on sort keys / contributor counts depends on identical expf / FMA contraction).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("FDLS_LIB", os.path.join(HERE, "libfdls.so"))  # FDLS_LIB / FDLS_NVCC_EXTRA: A/B builds
SOURCES = ["api.cu", "preprocess_fwd.cu", "binning.cu", "binning_partition.cu", "radix_sort.cu", "render_fwd.cu", "render_bwd.cu", "render_bwd_shfl.cu",
           "preprocess_bwd.cu", "knn.cu", "hexplane.cu", "deform_mlp.cu", "train_ops.cu"]
NVCC_FLAGS = ["-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler",
              "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _stale(out, deps):
    return not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps)


def build(force=False, verbose=False):
    """Compile each translation unit to an object (in parallel) and link the shared library.  Returns its path."""
    extra = os.environ.get("FDLS_NVCC_EXTRA", "").split()
    objdir = os.path.join(HERE, "build" + ("_" + "".join(c for c in "".join(extra) if c.isalnum()) if extra else ""))
    os.makedirs(objdir, exist_ok=True)
    hdrs = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "umma.cuh"), os.path.join(HERE, "..", "include", "fdls_abi.h"), os.path.abspath(__file__)]
    procs, objs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + hdrs):
            cmd = ["nvcc", "-c", "-o", obj, src] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else [])
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write("[%s]\n%s\n" % (s, out))
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(LIB, objs):
        subprocess.check_call(["nvcc", "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                                       "-lcudart"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
