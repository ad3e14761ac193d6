import sys, numpy as np, torch
sys.path.insert(0, '.')
from soar_b200 import synthetic, postops
out = {}
for name, (W, H, seed, pp, border) in {"postops_centered": (96, 64, 0, (0.5, 0.5), False), "postops_border_offcenter": (75, 53, 1, (0.45, 0.58), True)}.items():
    pb = synthetic.postops_problem(W, H, seed=seed, prcppoint=pp, touch_border=border)
    mask = (pb["opac"] > 1e-5).cuda()
    d = pb["depth"].cuda().requires_grad_(True)
    dn = postops.depth2normal(d, mask, pb["camera"])
    (dn * pb["g_depth_normal"].cuda()).sum().backward()
    n = pb["normal"].cuda().requires_grad_(True)
    cv = postops.normal2curv(n, mask)
    (cv * pb["g_curv"].cuda()).sum().backward()
    out[name + "_d2n"] = dn.detach().cpu().numpy(); out[name + "_gd"] = d.grad.cpu().numpy()
    out[name + "_n2c"] = cv.detach().cpu().numpy(); out[name + "_gn"] = n.grad.cpu().numpy()
np.savez_compressed("gpurun_out/post_dump.npz", **out)
