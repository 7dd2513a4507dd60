"""Generate tests/golden/*.npz from the UNMODIFIED reference CUDA rasterizer (oracle/_ref/libref_f{3,6}.so, built by
oracle/build_ref.py from /root/reference/submodules/4d-langsplat-rasterization/cuda_rasterizer/*.cu) on a B200:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'     # then copy gpurun_out/golden/*.npz here

The reference repository holds no golden vectors of its own, so these files -- outputs of the
reference itself on the seeded synthetic scenes of 4dlangsplat_b200/synth.py -- are what pins the CPU oracle
(tests/test_oracle_golden.py) and, through it, CPU-only CI.  Everything the reference keeps in its geometry / binning /
image buffers after a forward, and the 9 API gradients (+ conic, depth) after a backward, is stored.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

# (scene, F, background, camera pose of synth.POSES or None = identity).  The rotated case pins the expression trees of
# transformPoint4x3 / computeCov2D, which are exact whatever the FMA contraction when the view matrix is the identity.
CASES = [("tiny", 3, (1.0, 1.0, 1.0), None), ("tiny", 6, (0.0, 0.0, 0.0), None), ("tiny", 3, (1.0, 1.0, 1.0), "rot2")]


def main(outdir):
    import refcuda
    synth = importlib.import_module("4dlangsplat_b200.synth")
    os.makedirs(outdir, exist_ok=True)
    for name, F, bg, pose in CASES:
        s = synth.scene(name, F=F, bg=bg, **({"pose": pose} if pose else {}))
        dcol, dlang, ddepth = synth.upstream_grads(s, depth=True)
        ref = refcuda.Ref(F)
        r = ref.forward(s)
        g = ref.backward(dcol, dlang, ddepth)
        n = lambda t: t.detach().cpu().numpy()
        vis = n(r.radii) > 0
        clean = lambda t: np.where(vis.reshape(-1, *([1] * (t.ndim - 1))), n(t), 0).astype(n(t).dtype)  # culled rows are uninitialised
        out = dict(
            R=np.int64(r.R), radii=n(r.radii), tiles_touched=n(r.tiles_touched), point_offsets=n(r.point_offsets),
            depths=clean(r.depths), means2D=clean(r.means2D), cov3D=clean(r.cov3D), conic_opacity=clean(r.conic_opacity),
            rgb=clean(r.rgb), clamped=(clean(r.clamped) != 0).astype(np.uint8),
            keys=n(r.keys).astype(np.uint64), point_list=n(r.point_list).astype(np.uint32),
            keys_unsorted=n(r.keys_unsorted).astype(np.uint64), point_list_unsorted=n(r.point_list_unsorted).astype(np.uint32),
            ranges=n(r.ranges).astype(np.uint32), n_contrib=n(r.n_contrib).astype(np.uint32), final_T=n(r.final_T),
            color=n(r.color), lang=n(r.lang), depth=n(r.depth),
            g_means2D=n(g.means2D), g_conic=n(g.conic), g_opacity=n(g.opacity), g_colors=n(g.colors), g_lang=n(g.lang),
            g_depths=n(g.depths), g_means3D=n(g.means3D), g_cov3D=n(g.cov3D), g_sh=n(g.sh), g_scales=n(g.scales),
            g_rots=n(g.rots))
        path = os.path.join(outdir, "ref_%s%s_f%d.npz" % (name, "-" + pose if pose else "", F))
        np.savez_compressed(path, **out)
        print(path, "R=%d visible=%d" % (r.R, int(vis.sum())), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.dirname(os.path.abspath(__file__)))
