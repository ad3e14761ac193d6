"""SMPL-X body-model forward for one pose on the sm_100a kernels (soar_b200/csrc/smplx.cu).

Replaces, for the batch-1 calls the SOAR render loop makes (TS/utils/smpl.py:592-609):

    live_smpl = smpl_model.forward(betas, body_pose, global_orient, transl, ...)   # smplx.lbs.lbs()
    cano2live_jnt_mats = torch.matmul(live_smpl.A, self.inv_mats)

`lbs()` is TS/utils/smplx/lbs.py:147-246; SMPLX.forward folds `transl` into the vertices, the
joints and A[:, :3, 3] (TS/utils/smplx/body_models.py:1380-1383). The pose tensors are plain
buffers in the reference (TS/utils/smpl.py:543-545), so this op is forward-only.
There is no CPU path.
"""
import ctypes

import torch

from . import _capi

_lib = _capi.load()


class SmplxPoser:
    """Holds the (constant) body-model assets on the device and poses them.

    v_template [V,3], shapedirs [V,3,NB], posedirs [(J-1)*9, V*3], J_regressor [J,V],
    parents [J] (int), lbs_weights [V,J] — the buffers of the reference's SMPLX module."""

    def __init__(self, v_template, shapedirs, posedirs, J_regressor, parents, lbs_weights):
        if not v_template.is_cuda:
            raise RuntimeError("SmplxPoser: CUDA tensors required (there is no CPU path)")
        dev = v_template.device
        f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()
        self.v_template, self.shapedirs, self.posedirs = f(v_template), f(shapedirs), f(posedirs)
        self.J_regressor, self.lbs_weights = f(J_regressor), f(lbs_weights)
        self.parents = parents.detach().to(device=dev, dtype=torch.int32).contiguous()
        self.V, self.J = self.v_template.shape[0], self.J_regressor.shape[0]
        self.NB = self.shapedirs.shape[-1] if self.shapedirs.numel() else 0
        if self.posedirs.shape != ((self.J - 1) * 9, self.V * 3):
            raise RuntimeError(f"SmplxPoser: posedirs must be {((self.J - 1) * 9, self.V * 3)}, "
                               f"got {tuple(self.posedirs.shape)}")
        if self.lbs_weights.shape != (self.V, self.J) or self.J_regressor.shape != (self.J, self.V):
            raise RuntimeError("SmplxPoser: lbs_weights must be [V,J] and J_regressor [J,V]")
        self.device = dev
        self._model = _capi.SoarSmplxModel()
        m = self._model
        m.V, m.J, m.NB = self.V, self.J, self.NB
        m.v_template, m.shapedirs = self.v_template.data_ptr(), _capi.ptr(self.shapedirs)
        m.posedirs, m.J_regressor = self.posedirs.data_ptr(), self.J_regressor.data_ptr()
        m.parents, m.lbs_weights = self.parents.data_ptr(), self.lbs_weights.data_ptr()
        self._ws_bytes = _lib.soar_smplx_workspace_bytes(self.V, self.J)

    def __call__(self, betas, full_pose, transl=None, inv_mats=None, want_verts=True):
        """betas [NB] or [1,NB]; full_pose [J*3] / [1,J*3] / [J,3] axis-angle; transl [3] or None;
        inv_mats [J,4,4] or None -> (verts [V,3] or None, joints [J,3], A [J,4,4], bones or None)."""
        dev = self.device
        f = lambda t: None if t is None else t.detach().to(device=dev, dtype=torch.float32).reshape(-1).contiguous()
        betas, full_pose, transl, inv = f(betas), f(full_pose), f(transl), f(inv_mats)
        if full_pose.numel() != self.J * 3:
            raise RuntimeError(f"SmplxPoser: full_pose must hold {self.J * 3} values")
        if self.NB and (betas is None or betas.numel() != self.NB):
            raise RuntimeError(f"SmplxPoser: betas must hold {self.NB} values")
        if inv is not None and inv.numel() != self.J * 16:
            raise RuntimeError(f"SmplxPoser: inv_mats must be [{self.J},4,4]")
        e = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        verts = e(self.V, 3) if want_verts else None
        joints, A = e(self.J, 3), e(self.J, 4, 4)
        bones = e(self.J, 4, 4) if inv is not None else None
        with torch.cuda.device(dev):
            stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
            # scratch per call, from the caching allocator (stream-ordered reuse): poses enqueued on
            # different streams never share it
            ws = torch.empty(self._ws_bytes, dtype=torch.uint8, device=dev)
            _capi.check(_lib.soar_smplx_forward(
                ctypes.byref(self._model), _capi.ptr(betas), full_pose.data_ptr(), _capi.ptr(transl),
                _capi.ptr(inv), _capi.ptr(verts), joints.data_ptr(), A.data_ptr(), _capi.ptr(bones),
                ws.data_ptr(), ws.numel(), stream), "soar_smplx_forward")
        return verts, joints, A, bones
