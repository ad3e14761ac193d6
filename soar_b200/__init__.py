"""soar_b200 — B200-native kernels for SOAR's differentiable Gaussian-splat render loop.

The product surface is the reference's own operator API:
  * `diff_gaussian_rasterization` (GaussianRasterizationSettings, GaussianRasterizer)
  * `simple_knn._C.distCUDA2`
  * `soar_b200.lbs` (skinning of canonical Gaussians; fused stand-in for the torch op
    chain in TS/utils/smpl.py:609-613 and TS/renderer/diff_gaussian_rasterizer.py:103-114)
  * `soar_b200.views` (view-sharded multi-GPU rendering with one gradient allreduce)
all of them thin hosts over the C-ABI library built from soar_b200/csrc.
"""
__version__ = "0.1.0"
