"""CPU suite, part 3: host-side mirror of the reference interface (no device work): API names and
field order, argument validation with the reference's messages, failure without a CUDA device,
view sharding, and the N>1 gradient allreduce on gloo with world_size 2."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

# DGR/diff_gaussian_rasterization/__init__.py:267-284
REFERENCE_SETTINGS_FIELDS = (
    "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
    "projmatrix", "patch_bbox", "prcppoint", "sh_degree", "campos", "prefiltered", "render_front",
    "sort_descending", "debug", "config")


def _settings(**over):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    kw = dict(image_height=32, image_width=32, tanfovx=0.5, tanfovy=0.5, bg=torch.zeros(3),
              scale_modifier=1.0, viewmatrix=torch.eye(4), projmatrix=torch.eye(4),
              patch_bbox=torch.tensor([0.0, 0.0, 32.0, 32.0]), prcppoint=torch.tensor([0.5, 0.5]),
              sh_degree=0, campos=torch.zeros(3), prefiltered=False, render_front=False,
              sort_descending=False, debug=False, config=torch.zeros(4))
    kw.update(over)
    return GaussianRasterizationSettings(**kw)


def test_public_names_and_field_order():
    import diff_gaussian_rasterization as dgr
    assert dgr.GaussianRasterizationSettings._fields == REFERENCE_SETTINGS_FIELDS
    assert callable(dgr.GaussianRasterizer) and callable(dgr.rasterize_gaussians)
    from simple_knn._C import distCUDA2  # noqa: F401  (KNN/ext.cpp:15-17)
    import inspect
    sig = inspect.signature(dgr.GaussianRasterizer.forward)
    assert list(sig.parameters)[1:] == ["means3D", "means2D", "opacities", "shs", "colors_precomp",
                                        "scales", "rotations", "cov3D_precomp"]


def test_argument_validation_messages():
    from diff_gaussian_rasterization import GaussianRasterizer
    r = GaussianRasterizer(_settings())
    x, z = torch.zeros(4, 3), torch.zeros(4, 3)
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(x, z, torch.ones(4, 1), scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(x, z, torch.ones(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=torch.zeros(4, 3),
          scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(x, z, torch.ones(4, 1), shs=torch.zeros(4, 1, 3))
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair or precomputed 3D covariance"):
        r(x, z, torch.ones(4, 1), shs=torch.zeros(4, 1, 3), scales=torch.ones(4, 3),
          rotations=torch.ones(4, 4), cov3D_precomp=torch.zeros(4, 6))


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    """CPU tensors must be rejected loudly: the product has no CPU path."""
    from diff_gaussian_rasterization import GaussianRasterizer
    from simple_knn._C import distCUDA2
    from soar_b200.lbs import skin_gaussians
    r = GaussianRasterizer(_settings())
    with pytest.raises(RuntimeError, match="CUDA"):
        r(torch.zeros(4, 3), torch.zeros(4, 3), torch.ones(4, 1), shs=torch.zeros(4, 1, 3),
          scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="means3D must have dimensions"):
        r(torch.zeros(4, 2), torch.zeros(4, 3), torch.ones(4, 1), shs=torch.zeros(4, 1, 3),
          scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        distCUDA2(torch.zeros(10, 3))
    with pytest.raises(RuntimeError, match="CUDA"):
        skin_gaussians(torch.zeros(4, 3), torch.zeros(4, 4), torch.zeros(4, 55), torch.zeros(55, 4, 4))


def test_rasterize_views_rejects_cpu_tensors_and_mixed_sizes():
    """The batched multi-view node has no CPU path either, and checks its arguments before any launch."""
    from soar_b200.views import rasterize_views, shard_views
    s = _settings()
    kw = dict(means3D=torch.zeros(4, 3), opacities=torch.ones(4, 1), shs=torch.zeros(4, 1, 3),
              scales=torch.ones(4, 3), rotations=torch.ones(4, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        rasterize_views([s, s], **kw)
    with pytest.raises(RuntimeError, match="one image size"):
        rasterize_views([s, s._replace(image_height=s.image_height + 16)], **kw)
    with pytest.raises(RuntimeError, match="no views"):
        rasterize_views([], **kw)
    with pytest.raises(RuntimeError, match="num_views"):
        rasterize_views([s, s], means2D=torch.zeros(3, 4, 3), **kw)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        rasterize_views([s], means3D=kw["means3D"], opacities=kw["opacities"], scales=kw["scales"],
                        rotations=kw["rotations"])
    assert shard_views(32, 3, 8) == [3, 11, 19, 27]


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under the product packages may name it."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for pkg in ("soar_b200", "diff_gaussian_rasterization", "simple_knn"):
        for dirpath, _, files in os.walk(os.path.join(root, pkg)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                    src = open(os.path.join(dirpath, f)).read()
                    assert "import oracle" not in src and "from oracle" not in src, f
                    assert "oracle/" not in src.replace("oracle/_ref", "") or f.endswith(".md"), f


def test_shard_views_partitions_and_balances():
    from soar_b200.views import shard_views
    for n, w in ((32, 8), (7, 2), (5, 4), (3, 8)):
        parts = [shard_views(n, r, w) for r in range(w)]
        assert sorted(v for p in parts for v in p) == list(range(n))
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
    costs = [9, 1, 1, 1, 1, 1, 1, 1, 8]
    parts = [shard_views(len(costs), r, 2, costs) for r in range(2)]
    assert sorted(v for p in parts for v in p) == list(range(len(costs)))
    loads = [sum(costs[v] for v in p) for p in parts]
    assert abs(loads[0] - loads[1]) <= 1
    assert shard_views(4, 0, 1) == [0, 1, 2, 3]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_view_grads(view, P):
    """Deterministic stand-in for one view's parameter gradients."""
    g = torch.Generator().manual_seed(1000 + view)
    return {"means3D": torch.randn(P, 3, generator=g), "opacities": torch.randn(P, 1, generator=g),
            "shs": torch.randn(P, 4, 3, generator=g)}


def _rank_main(rank, world, port, n_views, P, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from soar_b200.views import FlatGrads, shard_views
    params = {k: torch.zeros_like(v).requires_grad_(True) for k, v in _fake_view_grads(0, P).items()}
    flat = FlatGrads(params)
    flat.zero_()
    for v in shard_views(n_views, rank, world):
        # autograd accumulates into .grad in place, i.e. straight into the flat buffer
        loss = sum((params[k] * g).sum() for k, g in _fake_view_grads(v, P).items())
        loss.backward()
    assert params["means3D"].grad.data_ptr() == flat.grad("means3D").data_ptr()
    flat.allreduce()
    torch.save({k: flat.grad(k).clone() for k in params}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.destroy_process_group()


def test_view_sharded_gradients_equal_single_rank_gloo(tmp_path):
    """N-rank view-sharded gradient == 1-rank gradient (SURVEY.md §4 iv), world_size 2 on gloo."""
    world, n_views, P = 2, 7, 257
    mp.spawn(_rank_main, args=(world, _free_port(), n_views, P, str(tmp_path)), nprocs=world, join=True)
    want = None
    for v in range(n_views):
        g = _fake_view_grads(v, P)
        want = g if want is None else {k: want[k] + g[k] for k in g}
    for r in range(world):
        got = torch.load(os.path.join(tmp_path, f"rank{r}.pt"))
        for k in want:
            assert torch.allclose(got[k], want[k], rtol=1e-5, atol=1e-5), (r, k)


def test_bench_loss_helper_matches_autograd():
    """bench.py's e2e step feeds the rasterizer backward with explicit cotangents: they must be the
    gradients autograd would produce for the same sum of mean squared errors."""
    import importlib
    import torch
    bench = importlib.import_module("bench")
    torch.manual_seed(0)
    outs = [torch.randn(3, 8, 8, requires_grad=True), torch.randn(3, 8, 8, requires_grad=True),
            torch.randn(1, 8, 8, requires_grad=True), torch.randn(1, 8, 8, requires_grad=True)]
    tgt = torch.randn(8, 8, 8)
    loss, cots = bench.loss_and_cotangents(*outs, tgt)
    ref = sum(((o - t) ** 2).mean() for o, t in zip(outs, (tgt[0:3], tgt[3:6], tgt[6:7], tgt[7:8])))
    grads = torch.autograd.grad(ref, outs)
    assert torch.allclose(loss, ref.detach(), rtol=1e-6, atol=0)
    for c, g in zip(cots, grads):
        assert torch.allclose(c, g, rtol=1e-6, atol=1e-9)


def test_tile_id_bits_is_the_bit_length_of_the_largest_id(tmp_path):
    """common.cuh's tile_id_bits(): bits of n_tiles - 1, not the reference's getHigherMsb(n_tiles)
    (DGR/cuda_rasterizer/rasterizer_impl.cu:35-48) — one bit fewer at powers of two. Built as a
    host-only program with nvcc (no device needed)."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "bits.cu"
    src.write_text('#include "%s/soar_b200/csrc/common.cuh"\n'
                   'namespace soar { void set_error(const char*, ...) {} }\n'
                   'int main() { const unsigned n[] = {1u, 2u, 3u, 4096u, 4097u, 8160u, 16384u};\n'
                   '  for (unsigned v : n) printf("%%u:%%d ", v, soar::tile_id_bits(v)); return 0; }\n' % root)
    exe = tmp_path / "bits"
    subprocess.run([nvcc, "-std=c++17", "-o", str(exe), str(src)], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    got = dict((int(a), int(b)) for a, b in (x.split(":") for x in out))
    assert got == {1: 1, 2: 1, 3: 2, 4096: 12, 4097: 13, 8160: 13, 16384: 14}
    for n, b in got.items():
        assert (n - 1) < (1 << b) and (n <= 2 or (n - 1) >= (1 << (b - 1)))


def _stats_rank_main(rank, world, port, N, out_dir):
    from soar_b200.densify import STATS, allreduce_densify_stats
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(100 + rank)          # every rank saw different views
    stats = {k: torch.rand(N, generator=g) for k in STATS}
    stats["max_radii2D"] = torch.randint(0, 50, (N,), generator=g).float()
    allreduce_densify_stats(stats)
    torch.save(stats, os.path.join(out_dir, f"stats{rank}.pt"))
    dist.destroy_process_group()


def test_densify_statistics_are_identical_on_every_rank_after_the_allreduce_gloo(tmp_path):
    """clone / split / prune decisions must come from statistics that are the same on
    every replica — sums of the per-rank accumulators, maximum of max_radii2D. world_size 2, gloo."""
    from soar_b200.densify import STATS
    world, N = 2, 1001
    mp.spawn(_stats_rank_main, args=(world, _free_port(), N, str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(os.path.join(tmp_path, f"stats{r}.pt")) for r in range(world)]
    local = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        st = {k: torch.rand(N, generator=g) for k in STATS}
        st["max_radii2D"] = torch.randint(0, 50, (N,), generator=g).float()
        local.append(st)
    for k in STATS:
        assert torch.equal(got[0][k], got[1][k]), k           # bit-identical across ranks
        want = (torch.maximum(local[0][k], local[1][k]) if k == "max_radii2D" else local[0][k] + local[1][k])
        assert torch.equal(got[0][k], want), k


def test_replica_generator_is_a_function_of_seed_and_cycle_only():
    from soar_b200.densify import replica_generator
    std = torch.full((64, 3), 0.1)
    draw = lambda seed, cycle: torch.normal(mean=torch.zeros_like(std), std=std,
                                            generator=replica_generator("cpu", seed, cycle))
    assert torch.equal(draw(7, 1), draw(7, 1))
    assert not torch.equal(draw(7, 1), draw(7, 2))
    assert not torch.equal(draw(7, 1), draw(8, 1))
