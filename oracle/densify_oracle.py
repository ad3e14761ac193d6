"""TEST INFRASTRUCTURE ONLY — CPU restatement (torch) of the reference's optimizer step and
densify / clone / split / prune logic over the per-Gaussian parameter groups
(TS/geometry/surfel_base.py:596-681, :847-1135, :1207-1216). Only tests/ and bench.py's baseline
legs may import it.

Pinned: tests/golden/densify_*.npz hold the complete state (parameters, both Adam moments, the six
statistics) after every stage of a scripted sequence run through the reference's OWN methods and a
real torch.optim.Adam, extracted from the reference file with `ast` and executed unchanged by
tests/golden/make_golden_densify.py; tests/test_densify_oracle.py replays the sequence here.

State = plain dict: p / m / v : {group: tensor}, stats: {name: [N] tensor}, step: {group: int}.
Unlike the reference it never touches an optimizer object: the moments are ordinary tensors.
"""
import torch

GROUPS = ("xyz", "f_dc", "f_rest", "color", "opacity", "scaling", "rotation")
STATS = ("xyz_gradient_accum", "scale_gradient_accum", "rot_gradient_accum", "opac_gradient_accum",
         "denom", "max_radii2D")


def new_state(params):
    p = {k: params[k].clone().float() for k in GROUPS}
    N = p["xyz"].shape[0]
    return dict(p=p, m={k: torch.zeros_like(v) for k, v in p.items()},
                v={k: torch.zeros_like(v) for k, v in p.items()},
                stats={k: torch.zeros(N, dtype=torch.float32, device=p["xyz"].device) for k in STATS},
                step={k: 0 for k in GROUPS})


def adam_step(st, grads, lrs, betas=(0.9, 0.999), eps=1e-15):
    """torch.optim.Adam's single-tensor update (torch/optim/adam.py), group by group (:596-681)."""
    b1, b2 = betas
    for k in GROUPS:
        g = grads.get(k)
        if g is None:
            continue
        st["step"][k] += 1
        t = st["step"][k]
        m, v, p = st["m"][k], st["v"][k], st["p"][k]
        m.lerp_(g, 1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** t, 1 - b2 ** t
        denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
        p.addcdiv_(m, denom, value=-(lrs.get(k, 0.0) / bc1))


def add_densification_stats(st, viewspace_grad, radii, scaling_grad):
    """update_states' per-view body (:1207-1216) + add_densification_stats (:1113-1135)."""
    s, f = st["stats"], radii > 0
    s["max_radii2D"] = torch.maximum(s["max_radii2D"], radii.float())
    s["xyz_gradient_accum"][f] += viewspace_grad[f, :2].norm(dim=-1)
    s["scale_gradient_accum"][f] += scaling_grad[f, :2].sum(1)
    s["rot_gradient_accum"][f] += st["p"]["rotation"][f].norm(dim=-1)
    s["opac_gradient_accum"][f] += st["p"]["opacity"][f].reshape(-1)
    s["denom"][f] += 1


def _select_rows(st, idx):
    for grp in ("p", "m", "v"):
        st[grp] = {k: t[idx] for k, t in st[grp].items()}
    st["stats"] = {k: t[idx] for k, t in st["stats"].items()}


def prune_points(st, mask):
    """:862-903 — keep the rows where mask is False, everywhere."""
    _select_rows(st, torch.nonzero(~mask.bool().reshape(-1)).reshape(-1))


def densification_postfix(st, new):
    """:905-977 — append rows; their moments are zero; ALL statistics are reset."""
    for k in GROUPS:
        add = new[k].reshape((new["xyz"].shape[0],) + tuple(st["p"][k].shape[1:]))
        st["p"][k] = torch.cat([st["p"][k], add], 0)
        st["m"][k] = torch.cat([st["m"][k], torch.zeros_like(add)], 0)
        st["v"][k] = torch.cat([st["v"][k], torch.zeros_like(add)], 0)
    N = st["p"]["xyz"].shape[0]
    st["stats"] = {k: torch.zeros(N, dtype=torch.float32, device=st["p"]["xyz"].device) for k in STATS}


def densify_masks(st, max_grad, extent, percent_dense=0.01):
    """adaptive_densify's averaged gradients and pre-mask (:1083-1106) and the two selection rules
    (:987-994, :1038-1046). Returns (clone_mask, split_mask)."""
    s = st["stats"]
    avg = lambda a: torch.nan_to_num(a / s["denom"], nan=0.0, posinf=float("inf"), neginf=float("-inf"))
    grad_pos, grad_scale, grad_opac = (avg(s[k]) for k in ("xyz_gradient_accum", "scale_gradient_accum",
                                                            "opac_gradient_accum"))
    pre = (grad_opac <= 2) & (grad_scale <= 1e-7)
    smax = torch.exp(st["p"]["scaling"]).max(1).values
    clone = (grad_pos.abs() >= max_grad) & (smax <= percent_dense * extent) & pre
    split = (grad_pos >= max_grad) & (smax > percent_dense * extent)
    return clone, split


def build_rotation(r):
    """utils/general_utils.py:100-123."""
    q = r / r.norm(dim=1, keepdim=True)
    w, x, y, z = q.unbind(1)
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)


def adaptive_densify(st, max_grad, extent, samples, N=2, surfel=True, percent_dense=0.01):
    """:1083-1111. `samples` [N·n_split, 3] are the torch.normal draws of :998 (recorded in the golden
    run; drawn by the caller otherwise)."""
    clone, split = densify_masks(st, max_grad, extent, percent_dense)
    n0 = st["p"]["xyz"].shape[0]
    densification_postfix(st, {k: st["p"][k][clone] for k in GROUPS})          # densify_and_clone
    p = st["p"]
    sel = torch.nonzero(split).reshape(-1)
    rots = build_rotation(p["rotation"][sel]).repeat(N, 1, 1)
    new = {k: p[k][sel].repeat((N,) + (1,) * (p[k].dim() - 1)) for k in GROUPS}
    new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + p["xyz"][sel].repeat(N, 1)
    new["scaling"] = torch.log(torch.exp(p["scaling"][sel]).repeat(N, 1) / (0.8 * N))
    if surfel:
        new["scaling"][:, -1] = -1e10
    n1 = p["xyz"].shape[0]
    densification_postfix(st, new)
    prune = torch.zeros(n1 + sel.numel() * N, dtype=torch.bool, device=p["xyz"].device)
    prune[:n0] = split
    prune_points(st, prune)
    return clone, split


def prune_mask(st, min_opacity, extent):
    """adaptive_prune's rule (:1069-1081)."""
    s = torch.exp(st["p"]["scaling"])[:, :2]
    smin, smax = s.min(1).values, s.max(1).values
    return (torch.sigmoid(st["p"]["opacity"]).reshape(-1) < min_opacity) | (st["stats"]["denom"] == 0) \
        | (smax > 0.5 * extent) | ((smin * smax) < (1e-8 * extent ** 2))


def adaptive_prune(st, min_opacity, extent):
    m = prune_mask(st, min_opacity, extent)
    prune_points(st, m)
    return m
