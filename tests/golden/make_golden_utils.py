"""Golden vectors of SURVEY §8 row f4 (gaussian2occgrid / contour_padding), recorded from the
UNMODIFIED reference kernels (oracle/_ref/libref_utils.so = TS/utils/cuda_utils.cu compiled in place
by oracle/build_ref.sh). Needs a GPU:

    gpurun -- 'python tests/golden/make_golden_utils.py gpurun_out/golden_utils'

then copy gpurun_out/golden_utils/utils_*.npz into tests/golden/. The occupancy grid is stored sparse
(flat indices + values); `spread` is the largest difference between two runs of the reference itself
(its float atomics are unordered)."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import refutils  # noqa: E402
import utils_cases as U  # noqa: E402


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    for name in U.OCC_GOLDEN:
        c = U.occgrid_case(name)
        t = {k: torch.from_numpy(v).cuda() for k, v in c.items()}
        args = [t[k] for k in ("pos_min", "pos_max", "grid_len", "grid_dim", "pos", "rot", "scale", "opac", "cutoff")]
        g1 = refutils.gaussian2occgrid(*args).cpu().numpy()
        g2 = refutils.gaussian2occgrid(*args).cpu().numpy()
        nz = np.flatnonzero(g1)
        np.savez_compressed(os.path.join(out_dir, f"utils_occgrid_{name}.npz"),
                            **{"in_" + k: v for k, v in c.items()}, grid_shape=np.array(g1.shape, np.int32),
                            grid_index=nz.astype(np.int64), grid_value=g1.reshape(-1)[nz],
                            spread=np.float32(np.abs(g1 - g2).max()))
        print(name, g1.shape, "non-zero cells", nz.size, "sum", float(g1.sum()), "spread", float(np.abs(g1 - g2).max()))
    for name in U.CONTOUR_GOLDEN:
        c = U.contour_case(name)
        res = refutils.contour_padding(torch.from_numpy(c["image"]).cuda(), torch.from_numpy(c["mask"]).cuda(),
                                       c["size"]).cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, f"utils_contour_{name}.npz"), in_image=c["image"],
                            in_mask=c["mask"], in_size=np.int32(c["size"]), res=res)
        print(name, res.shape, "non-zero", int(np.count_nonzero(res)))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden_utils")
