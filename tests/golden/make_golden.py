#!/usr/bin/env python
"""Generates the golden vectors under tests/golden/ by running the UNMODIFIED reference CUDA
code (oracle/_ref/libref_dgr.so, libref_knn.so — compiled in place from /root/reference by
oracle/build_ref.sh) on a B200. The reference ships no fixtures of its own, so
these pin both the CPU oracle (oracle/rast_oracle.c, oracle/knn_oracle.py) and the product.

Run on the GPU box:   python tests/golden/make_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.npz into tests/golden/ and commit them.

The reference's backward uses float atomics whose order is not deterministic; each gradient is
therefore stored as the MEAN of 5 runs in float64->float32 together with the max spread seen
(`grad_spread_<name>`), which tests add to their tolerance."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from soar_b200 import synthetic  # noqa: E402
import refdgr  # noqa: E402

CASES = {
    # name: (scene kwargs, lrn_cam)
    "rast_vanilla_sh3": dict(P=1500, W=80, H=48, sh_degree=3, seed=11),
    "rast_surfel_soar": dict(P=2000, W=96, H=64, surfel=True, seed=12, bg=(1.0, 1.0, 1.0),
                             sort_descending=True),
    "rast_surfel_front_sh1": dict(P=2000, W=70, H=50, sh_degree=1, surfel=True, seed=13,
                                  render_front=True, config=(1.0, 0.0, 1.0, 0.0)),
    "rast_lrn_cam": dict(P=800, W=64, H=64, sh_degree=2, seed=14, config=(1.0, 1.0, 0.0, 1.0),
                         bg=(0.2, 0.5, 0.7)),
}
SCENE_KEYS = ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "viewmatrix",
              "projmatrix", "campos", "bg", "patch_bbox", "prcppoint", "config")
SCALARS = ("tanfovx", "tanfovy", "image_height", "image_width", "scale_modifier", "sh_degree",
           "prefiltered", "render_front", "sort_descending")


def rast_case(name, kw, out_dir):
    kw = dict(kw)
    P, W, H = kw.pop("P"), kw.pop("W"), kw.pop("H")
    s_cpu = synthetic.scene(P, W, H, **kw)
    cot_cpu = synthetic.cotangents(W, H, seed=kw.get("seed", 0))
    s = {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in s_cpu.items()}
    cot = {k: v.cuda() for k, v in cot_cpu.items()}
    ref = refdgr.RefRasterizer()
    o = ref.forward(s)
    d = {}
    for k in SCENE_KEYS:
        if s_cpu.get(k) is not None:
            d["in_" + k] = s_cpu[k].numpy()
    for k in SCALARS:
        d["in_" + k] = np.asarray(s_cpu[k])
    for k in ("color", "normal", "depth", "opac"):
        d["cot_" + k] = cot_cpu[k].numpy()
        d["out_" + k] = o[k].cpu().numpy()
    d["out_radii"] = o["radii"].cpu().numpy()
    d["out_num_rendered"] = np.asarray(o["num_rendered"])
    d["out_keys"] = ref.sorted_keys().cpu().numpy().view(np.uint64)
    d["out_point_list"] = ref.point_list().cpu().numpy().view(np.uint32)
    d["out_ranges"] = ref.ranges().cpu().numpy().view(np.uint32)
    d["out_n_contrib"] = ref.n_contrib().cpu().numpy().view(np.uint32)
    runs = [ref.backward(cot) for _ in range(5)]
    for k in runs[0]:
        if runs[0][k].numel() == 0:
            continue
        st = torch.stack([r[k].double() for r in runs])
        d["grad_" + k] = st.mean(0).float().cpu().numpy()
        d["grad_spread_" + k] = np.asarray((st.max(0).values - st.min(0).values).max().item())
    ref.close()
    np.savez_compressed(os.path.join(out_dir, name + ".npz"), **d)
    print(name, "R =", int(d["out_num_rendered"]), "visible =", int((d["out_radii"] > 0).sum()),
          {k[12:]: float(v) for k, v in d.items() if k.startswith("grad_spread_")})


def knn_case(name, P, kind, seed, out_dir):
    rng = np.random.Generator(np.random.PCG64(seed))
    if kind == "uniform":
        pts = rng.uniform(-1, 1, size=(P, 3))
    else:  # points near a sphere shell with duplicates
        v = rng.normal(size=(P, 3))
        pts = v / np.linalg.norm(v, axis=1, keepdims=True) * (1 + 0.01 * rng.normal(size=(P, 1)))
        pts[P // 2: P // 2 + 50] = pts[:50]
    pts = pts.astype(np.float32)
    out = refdgr.ref_dist2(torch.from_numpy(pts).cuda()).cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, name + ".npz"), points=pts, dist2=out)
    print(name, P, float(out.mean()))


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out_dir, exist_ok=True)
    assert refdgr.available(), "oracle/_ref libraries missing: run oracle/build_ref.sh where /root/reference is mounted"
    for name, kw in CASES.items():
        rast_case(name, kw, out_dir)
    knn_case("knn_uniform_5000", 5000, "uniform", 21, out_dir)
    knn_case("knn_shell_dups_3000", 3000, "shell", 22, out_dir)


if __name__ == "__main__":
    main()
