"""GPU tests of the multi-view extensions (no reference analogue — the reference renders one
view per blocking call): GaussianRasterizer.begin(), soar_b200.views.iter_views and in-place
gradient accumulation (FlatGrads.direct). Each is checked against the plain per-view path.
Forward outputs must be bit-identical; accumulated gradients are compared at 1e-3 relative to
the tensor's max-abs (the blend backward's global reductions are unordered)."""
import pytest
import torch

from soar_b200 import synthetic
from soar_b200.views import FlatGrads, iter_views
from util import rel_err, settings_of, to_device

pytestmark = pytest.mark.gpu

KEYS = ("means3D", "shs", "opacities", "scales", "rotations")


def _views(P, W, H, n, sh_degree=2, seed=21, surfel=False):
    scenes = [to_device(synthetic.scene(P, W, H, sh_degree=sh_degree, seed=seed, view=v, n_views=n,
                                        surfel=surfel)) for v in range(n)]
    cot = to_device(synthetic.cotangents(W, H, seed=seed))
    return scenes, cot


def _kwargs(params):
    return dict(means3D=params["means3D"], means2D=torch.zeros_like(params["means3D"], requires_grad=True),
                opacities=params["opacities"], shs=params["shs"], colors_precomp=None,
                scales=params["scales"], rotations=params["rotations"], cov3D_precomp=None)


def _loss(out, cot):
    color, normal, depth, opac, _ = out
    return ((color * cot["color"]).sum() + (normal * cot["normal"]).sum()
            + (depth * cot["depth"]).sum() + (opac * cot["opac"]).sum())


def test_begin_then_forward_is_bit_identical():
    from diff_gaussian_rasterization import GaussianRasterizer
    scenes, _ = _views(30_000, 320, 200, 1)
    s = scenes[0]
    params = {k: s[k].clone() for k in KEYS}
    plain = GaussianRasterizer(settings_of(s))(**_kwargs(params))
    r = GaussianRasterizer(settings_of(s))
    kw = _kwargs(params)
    assert r.begin(**kw) is r
    split = r(**kw)
    for a, b in zip(plain, split):
        assert torch.equal(a, b)
    assert r._pending is None
    # a forward with other arguments after begin() is refused
    r.begin(**kw)
    other = {k: v.clone() for k, v in params.items()}
    with pytest.raises(RuntimeError):
        r(**_kwargs(other))


@pytest.mark.parametrize("surfel", [False, True])
def test_iter_views_with_direct_accumulation_matches_autograd(surfel):
    from diff_gaussian_rasterization import GaussianRasterizer
    n = 3
    scenes, cot = _views(40_000, 256, 256, n, sh_degree=1 if surfel else 3, surfel=surfel)
    base = {k: scenes[0][k] for k in KEYS}

    # reference path of this test: one blocking call per view, autograd accumulates .grad
    ref = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    outs_ref = []
    for s in scenes:
        out = GaussianRasterizer(settings_of(s))(**_kwargs(ref))
        outs_ref.append([t.detach().clone() for t in out])
        _loss(out, cot).backward()

    got = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    flat = FlatGrads(got)
    flat.zero_()
    rs = [GaussianRasterizer(settings_of(s)) for s in scenes]
    kws = [_kwargs(got) for _ in scenes]
    with flat.direct():
        for v, out in enumerate(iter_views(rs, kws)):
            for a, b in zip(out, outs_ref[v]):
                assert torch.equal(a.detach(), b)
            _loss(out, cot).backward()
    for k in KEYS:
        assert got[k].grad.data_ptr() == flat.grad(k).data_ptr()  # still the flat buffer
        assert rel_err(flat.grad(k), ref[k].grad) <= 1e-3, k
    # means2D is not a registered leaf: it takes the ordinary autograd route
    assert all(kw["means2D"].grad is not None for kw in kws)


def test_direct_accumulation_ignores_unregistered_and_derived_tensors():
    """A tensor computed from a registered leaf (new storage) must keep flowing through autograd."""
    from diff_gaussian_rasterization import GaussianRasterizer
    scenes, cot = _views(10_000, 128, 128, 1, sh_degree=0)
    s = scenes[0]
    raw = {k: s[k].clone().requires_grad_(True) for k in KEYS}
    flat = FlatGrads(raw)
    flat.zero_()
    kw = _kwargs(raw)
    kw["opacities"] = raw["opacities"] * 1.0  # derived: not the leaf's storage
    with flat.direct():
        _loss(GaussianRasterizer(settings_of(s))(**kw), cot).backward()
    plain = {k: s[k].clone().requires_grad_(True) for k in KEYS}
    _loss(GaussianRasterizer(settings_of(s))(**_kwargs(plain)), cot).backward()
    for k in KEYS:
        assert rel_err(flat.grad(k), plain[k].grad) <= 1e-3, k


def test_begin_rejects_changed_or_modified_arguments():
    """forward() after begin() must see the SAME arguments: a different scales / opacities / SH
    tensor, an in-place update in between, or swapped settings would otherwise render the geometry
    begin() preprocessed."""
    from diff_gaussian_rasterization import GaussianRasterizer
    scenes, _ = _views(5_000, 128, 96, 2, sh_degree=1)
    s = scenes[0]
    params = {k: s[k].clone() for k in KEYS}
    for key in ("scales", "rotations", "opacities", "shs"):
        r = GaussianRasterizer(settings_of(s))
        kw = _kwargs(params)
        r.begin(**kw)
        kw2 = dict(kw)
        kw2[key] = params[key].clone()
        with pytest.raises(RuntimeError):
            r(**kw2)
    r = GaussianRasterizer(settings_of(s))
    kw = _kwargs(params)
    r.begin(**kw)
    params["scales"].mul_(1.5)   # same storage, new content
    with pytest.raises(RuntimeError):
        r(**kw)
    r = GaussianRasterizer(settings_of(s))
    kw = _kwargs(params)
    r.begin(**kw)
    r.raster_settings = settings_of(scenes[1])
    with pytest.raises(RuntimeError):
        r(**kw)


@pytest.mark.parametrize("P", [10_001, 10_002, 10_003])
def test_direct_accumulation_with_P_not_a_multiple_of_4(P):
    """with P % 4 != 0 an unpadded flat buffer put the rotations sink at an odd
    float offset and the 16-byte red.global.add faulted. FlatGrads now pads every slice to 16 bytes;
    overwrite=True covers the store path."""
    from diff_gaussian_rasterization import GaussianRasterizer
    scenes, cot = _views(P, 160, 128, 2, sh_degree=3)
    base = {k: scenes[0][k] for k in KEYS}
    ref = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    for s in scenes:
        _loss(GaussianRasterizer(settings_of(s))(**_kwargs(ref)), cot).backward()
    got = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    flat = FlatGrads(got)
    for k in KEYS:
        assert flat.grad(k).data_ptr() % 16 == 0, f"slice {k} is not 16-byte aligned"
    flat.flat.fill_(7.0)
    with flat.direct(overwrite=True):
        for s in scenes:
            _loss(GaussianRasterizer(settings_of(s))(**_kwargs(got)), cot).backward()
    torch.cuda.synchronize()
    for k in KEYS:
        assert rel_err(flat.grad(k), ref[k].grad) <= 1e-3, k


def test_misaligned_rotation_pointers_take_the_scalar_path():
    """Caller-owned [P,4] tensors at an odd float offset (a slice of somebody else's flat buffer): the
    launchers must notice and fall back to scalar loads / stores / reductions instead of faulting."""
    from diff_gaussian_rasterization import _C
    P = 9_001
    s = to_device(synthetic.scene(P, 160, 128, sh_degree=1, seed=5))
    cot = to_device(synthetic.cotangents(160, 128, seed=5))
    e = torch.Tensor([])

    def run(rot, sink):
        out = _C.rasterize_gaussians(s["bg"], s["means3D"], e, s["opacities"], s["scales"], rot, 1.0, e,
                                     s["viewmatrix"], s["projmatrix"], s["prcppoint"], s["patch_bbox"],
                                     s["tanfovx"], s["tanfovy"], 128, 160, s["shs"], 1, s["campos"], False,
                                     False, False, False, s["config"])
        R, color, normal, depth, opac, radii, geom, binning, img = out
        g = _C.rasterize_gaussians_backward(
            s["bg"], s["means3D"], radii, e, s["scales"], rot, 1.0, e, s["viewmatrix"], s["projmatrix"],
            s["prcppoint"], s["patch_bbox"], s["tanfovx"], s["tanfovy"], cot["color"], cot["normal"],
            cot["depth"], cot["opac"], s["shs"], 1, s["campos"], geom, R, binning, img, False, s["config"],
            {"rotations": sink} if sink is not None else None)
        torch.cuda.synchronize()
        return color, g[7]

    color0, grot0 = run(s["rotations"], None)
    store = torch.zeros(4 * P + 1, device="cuda")
    store[1:] = s["rotations"].reshape(-1)
    rot_odd = store[1:].view(P, 4)
    sink_store = torch.zeros(4 * P + 3, device="cuda")
    sink = sink_store[3:].view(P, 4)
    assert rot_odd.data_ptr() % 16 != 0 and sink.data_ptr() % 16 != 0
    color1, grot1 = run(rot_odd, sink)
    assert torch.equal(color0, color1)
    assert grot1.data_ptr() == sink.data_ptr()
    assert rel_err(sink, grot0) <= 1e-3
    color2, _ = run(rot_odd, sink)          # accumulates a second time
    assert rel_err(sink, 2 * grot0) <= 1e-3


@pytest.mark.parametrize("n_streams", [2, 3])
@pytest.mark.parametrize("surfel", [False, True])
def test_views_on_several_streams_match_the_single_stream_loop(n_streams, surfel):
    """iter_views(streams=...): the views of a step overlap on the device. Images must be bit-identical
    to the one-stream loop; the accumulated gradients are compared at 1e-3 relative (the blend
    backward's per-Gaussian reductions are unordered within a view, as on one stream). The additions
    of the views are chained through events (soar_rast_backward_ordered), and the first view stores
    into a buffer that held garbage."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from soar_b200.views import view_streams
    n = 5
    scenes, cot = _views(40_000, 320, 240, n, sh_degree=2, surfel=surfel)
    base = {k: scenes[0][k] for k in KEYS}

    def step(streams, fill):
        p = {k: v.clone().requires_grad_(True) for k, v in base.items()}
        flat = FlatGrads(p)
        flat.flat.fill_(fill)
        images = []
        rs = [GaussianRasterizer(settings_of(s)) for s in scenes]
        kws = [_kwargs(p) for _ in scenes]
        with flat.direct(overwrite=True):
            for out in iter_views(rs, kws, streams):
                _loss(out, cot).backward()
                images.append(out[0].detach())
        torch.cuda.synchronize()
        return flat, images

    one, img_one = step(None, 3.0)
    streams = view_streams(n_streams)
    many, img_many = step(streams, -5.0)
    for a, b in zip(img_one, img_many):
        assert torch.equal(a, b)
    for k in KEYS:
        assert torch.isfinite(many.grad(k)).all()
        assert rel_err(many.grad(k), one.grad(k)) <= 1e-3, k
    # and again on the same streams (event pool reuse, stream re-join)
    again, _ = step(streams, 11.0)
    for k in KEYS:
        assert rel_err(again.grad(k), one.grad(k)) <= 1e-3, k


@pytest.mark.parametrize("n_streams", [0, 3])
@pytest.mark.parametrize("surfel", [False, True])
def test_rasterize_views_matches_the_per_view_rasterizer(n_streams, surfel):
    """soar_b200.views.rasterize_views (all views of a step as one autograd node) against one
    GaussianRasterizer call per view + torch.stack: images and radii bit-identical, summed parameter
    gradients and per-view screen-space gradients at 1e-3 relative."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from soar_b200.views import rasterize_views, view_streams
    n = 4
    scenes, cot = _views(30_000, 256, 208, n, sh_degree=2, surfel=surfel)
    base = {k: scenes[0][k] for k in KEYS}
    # per-view cotangents that differ from view to view
    cots = [{k: v * (1.0 + 0.25 * i) for k, v in cot.items()} for i in range(n)]

    ref = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    ref_out, ref_m2 = [], []
    for s, c in zip(scenes, cots):
        kw = _kwargs(ref)
        out = GaussianRasterizer(settings_of(s))(**kw)
        _loss(out, c).backward()
        ref_out.append([t.detach() for t in out])
        ref_m2.append(kw["means2D"].grad.clone())

    got = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    m2 = torch.zeros((n, base["means3D"].size(0), 3), device="cuda", requires_grad=True)
    streams = view_streams(n_streams) if n_streams else None
    color, normal, depth, opac, radii = rasterize_views(
        [settings_of(s) for s in scenes], means3D=got["means3D"], opacities=got["opacities"], shs=got["shs"],
        scales=got["scales"], rotations=got["rotations"], means2D=m2, streams=streams)
    for v in range(n):
        assert torch.equal(color[v], ref_out[v][0]) and torch.equal(normal[v], ref_out[v][1])
        assert torch.equal(depth[v], ref_out[v][2]) and torch.equal(opac[v], ref_out[v][3])
        assert torch.equal(radii[v], ref_out[v][4])
    stack = lambda key: torch.stack([c[key] for c in cots])
    torch.autograd.backward([color, normal, depth, opac],
                            [stack("color"), stack("normal"), stack("depth"), stack("opac")])
    torch.cuda.synchronize()
    for k in KEYS:
        assert rel_err(got[k].grad, ref[k].grad) <= 1e-3, k
    for v in range(n):
        assert rel_err(m2.grad[v], ref_m2[v]) <= 1e-3, f"means2D view {v}"


@pytest.mark.parametrize("n_streams", [0, 2])
def test_rasterize_views_under_direct_accumulation(n_streams):
    """Under FlatGrads.direct(overwrite=True) the batched node stores its first view into the flat
    buffer (which held garbage) and adds the others; a second call in the same context adds on top."""
    from soar_b200.views import rasterize_views, view_streams
    n = 3
    scenes, cot = _views(20_000, 192, 160, n, sh_degree=1)
    base = {k: scenes[0][k] for k in KEYS}
    sts = [settings_of(s) for s in scenes]
    stack = lambda key: torch.stack([cot[key]] * n)
    cots = [stack("color"), stack("normal"), stack("depth"), stack("opac")]

    def call(p, streams):
        out = rasterize_views(sts, means3D=p["means3D"], opacities=p["opacities"], shs=p["shs"],
                              scales=p["scales"], rotations=p["rotations"], streams=streams)
        torch.autograd.backward(list(out[:4]), cots)

    ref = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    call(ref, None)
    got = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    flat = FlatGrads(got)
    flat.flat.fill_(9.0)
    streams = view_streams(n_streams) if n_streams else None
    with flat.direct(overwrite=True):
        call(got, streams)
        torch.cuda.synchronize()
        for k in KEYS:
            assert rel_err(flat.grad(k), ref[k].grad) <= 1e-3, k
        call(got, streams)
    torch.cuda.synchronize()
    for k in KEYS:
        assert rel_err(flat.grad(k), 2 * ref[k].grad) <= 1e-3, k


def test_sinkless_node_does_not_use_up_the_first_store():
    """A batched node whose inputs are NOT registered leaves (detached copies, like an occupancy pass)
    may run its backward before the node that owns the sinks. It must leave the step's one STORE to the
    owner: the flat buffer starts as NaN here, so an owner that adds instead of storing shows at once."""
    from soar_b200.views import rasterize_views
    n = 2
    scenes, cot = _views(10_000, 128, 96, n, sh_degree=1)
    base = {k: scenes[0][k] for k in KEYS}
    sts = [settings_of(s) for s in scenes]
    stack = lambda key: torch.stack([cot[key]] * n)
    cots = [stack("color"), stack("normal"), stack("depth"), stack("opac")]

    def render(p):
        return rasterize_views(sts, means3D=p["means3D"], opacities=p["opacities"], shs=p["shs"],
                               scales=p["scales"], rotations=p["rotations"])

    ref = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    torch.autograd.backward(list(render(ref)[:4]), cots)

    got = {k: v.clone().requires_grad_(True) for k, v in base.items()}
    side = {k: v.clone().requires_grad_(True) for k, v in base.items()}   # leaves, but not in the buffer
    flat = FlatGrads(got)
    flat.flat.fill_(float("nan"))
    with flat.direct(overwrite=True):
        owner = render(got)
        loose = render(side)
        torch.autograd.backward(list(loose[:4]), cots)     # the sink-less node goes first
        torch.autograd.backward(list(owner[:4]), cots)
    torch.cuda.synchronize()
    for k in KEYS:
        assert bool(torch.isfinite(flat.grad(k)).all()), k
        assert rel_err(flat.grad(k), ref[k].grad) <= 1e-3, k
        assert rel_err(side[k].grad, ref[k].grad) <= 1e-3, k


def test_rasterize_views_refuses_learnable_camera_and_second_backward():
    from soar_b200.views import rasterize_views
    scenes, cot = _views(2_000, 64, 64, 2, sh_degree=0)
    s = scenes[0]
    sts = [settings_of(x) for x in scenes]
    cam = sts[1]._replace(config=torch.tensor([1.0, 0.0, 0.0, 1.0], device="cuda"))
    with pytest.raises(RuntimeError, match="learnable camera"):
        rasterize_views([sts[0], cam], means3D=s["means3D"], opacities=s["opacities"], shs=s["shs"],
                        scales=s["scales"], rotations=s["rotations"])
    p = {k: s[k].clone().requires_grad_(True) for k in KEYS}
    out = rasterize_views(sts, means3D=p["means3D"], opacities=p["opacities"], shs=p["shs"],
                          scales=p["scales"], rotations=p["rotations"])
    loss = sum(t.sum() for t in out[:4])
    loss.backward(retain_graph=True)
    with pytest.raises(RuntimeError, match="second time"):
        loss.backward()


def test_rasterize_views_argument_errors_and_empty_cloud():
    from soar_b200.views import rasterize_views
    scenes, _ = _views(1_000, 64, 64, 2, sh_degree=0)
    s = scenes[0]
    sts = [settings_of(x) for x in scenes]
    with pytest.raises(Exception):
        rasterize_views(sts, means3D=s["means3D"], opacities=s["opacities"], scales=s["scales"],
                        rotations=s["rotations"])                      # neither SHs nor colours
    other = settings_of(s)._replace(image_height=32)
    with pytest.raises(RuntimeError):
        rasterize_views([sts[0], other], means3D=s["means3D"], opacities=s["opacities"], shs=s["shs"],
                        scales=s["scales"], rotations=s["rotations"])
    e = lambda *shape: torch.zeros(shape, device="cuda")
    out = rasterize_views(sts, means3D=e(0, 3), opacities=e(0, 1), shs=e(0, 1, 3), scales=e(0, 3),
                          rotations=e(0, 4))
    assert out[0].shape == (2, 3, 64, 64) and float(out[0].abs().max()) == 0.0 and out[4].shape == (2, 0)
