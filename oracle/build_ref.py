"""TEST INFRASTRUCTURE ONLY: build the UNMODIFIED reference CUDA rasterizer for sm_100a into oracle/_ref/.

Recipe (no reference build system is run, no reference source is copied): nvcc compiles the reference's three
translation units where they lie under /root/reference/submodules/4d-langsplat-rasterization/cuda_rasterizer/
plus oracle/ref_shim.cu (a C ABI in place of the torch glue) into one shared library per language-feature width
F (the reference bakes F in as a macro: setup.py:18,32).  The flags are the reference's own (setup.py:33 passes
only the glm include path; no fast-math, default -fmad) plus `-include cstdint`, without which
rasterizer_impl.h:24 does not compile on gcc 13.

/root/reference only exists in the build container; on the GPU box the prebuilt oracle/_ref/*.so are used.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/submodules/4d-langsplat-rasterization"
OUT = os.path.join(HERE, "_ref")


def lib_path(F):
    return os.path.join(OUT, "libref_f%d.so" % F)


def build(force=False, widths=(3, 6, 1, 2, 4, 5)):
    """Returns the list of built library paths ([] when the reference tree is not present)."""
    if not os.path.isdir(REF):
        return [lib_path(F) for F in widths if os.path.exists(lib_path(F))]
    os.makedirs(OUT, exist_ok=True)
    srcs = [os.path.join(REF, "cuda_rasterizer", n) for n in ("forward.cu", "backward.cu", "rasterizer_impl.cu")]
    shim = os.path.join(HERE, "ref_shim.cu")
    built, procs = [], []
    for F in widths:
        out = lib_path(F)
        newest = max(os.path.getmtime(s) for s in srcs + [shim, os.path.abspath(__file__)])
        if force or not os.path.exists(out) or os.path.getmtime(out) < newest:
            cmd = ["nvcc", "-std=c++17", "-O3", "-shared", "-Xcompiler", "-fPIC", "-lineinfo",
                   "-gencode", "arch=compute_100a,code=sm_100a", "-include", "cstdint", "-diag-suppress", "177",
                   "-DNUM_CHANNELS_language_feature=%d" % F,
                   "-I", os.path.join(REF, "third_party", "glm"), "-I", os.path.join(REF, "cuda_rasterizer"),
                   "-o", out, shim] + srcs
            procs.append((cmd, subprocess.Popen(cmd)))  # the widths build side by side
        built.append(out)
    for cmd, p in procs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    return built


if __name__ == "__main__":
    print("\n".join(build(force="--force" in sys.argv)))
