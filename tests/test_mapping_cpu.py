"""Host logic of the view-sharded mapping step on the CPU (gloo, world_size 2).

The renderer and the image loss are injected from the oracle (the CUDA kernels need a
GPU); what is tested is the sharding arithmetic: window[r::p] partition, per-rank loss
share, ONE all-reduce of the flat [15 N] gradient, union-visibility isotropic term, and
that every rank ends the step with bit-identical parameters.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gslam_b200.mapping import (FlatSplats, MappingStep, add_isotropic_grad_, edge_aware_tv, shard_views)
from gslam_b200.synthetic import make_scene, target_images
from oracle import torch_ref as R

N_VIEWS, N_GAUSS = 4, 600


def _oracle_render(*a, **kw):
    out = R.rasterization(*a, **kw)
    out.render = torch.cat([out.rgbs, out.depthmaps[..., None], out.betas[..., None]], dim=-1)
    return out


def _oracle_loss(render, gt, exposure, w, mode):
    photo = R.mapping_photometric_loss(R.exposure_affine(render[..., :3], exposure), gt, render[..., 4])
    ssim = 1.0 - R.fused_ssim(render[..., :3].permute(0, 3, 1, 2), gt.permute(0, 3, 1, 2), padding="valid")
    return (1 - w) * photo + w * ssim, photo.detach()


class _Adam:
    """torch Adam + the opacity decay, with FusedAdam's step() signature."""
    def __init__(self, params):
        self.o = torch.optim.Adam(params, lr=1e-2)

    def step(self, decay_param=None, vis_count=None, decay=0.995):
        self.o.step()
        with torch.no_grad():
            decay_param[vis_count > 1] *= decay


def _run(rank, world, port, q):
    if world > 1:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    sc = make_scene(N_GAUSS, N_VIEWS, "tum", 0.1)
    gt = target_images(N_VIEWS, sc.height, sc.width)
    exposure = torch.zeros(N_VIEWS, 2)
    splats = FlatSplats(sc.params())
    step = MappingStep(splats, _Adam(splats.params), render_fn=_oracle_render, loss_fn=_oracle_loss,
                       depth_tv_weight=1e-6)
    mine = shard_views(N_VIEWS, rank, world)
    res = step(sc.viewmats[mine], sc.Ks[mine], gt[mine], exposure[mine], sc.width, sc.height, N_VIEWS)
    # numpy payloads are pickled by value (tensors would be shared through fds that die with the child)
    q.put((rank, splats.flat_grad().detach().numpy().copy(), splats.flat.detach().numpy().copy(),
           float(res.total), res.vis_count.numpy().copy()))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _launch(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_run, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=600) for _ in range(world)], key=lambda t: t[0])
    out = [(r, torch.from_numpy(g), torch.from_numpy(p), l, torch.from_numpy(v)) for r, g, p, l, v in out]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def test_shard_views_partition():
    for n in (1, 5, 8, 16):
        for w in (1, 2, 4, 8):
            parts = [shard_views(n, r, w) for r in range(w)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


@pytest.mark.timeout(900)
def test_two_ranks_match_one_rank():
    one = _launch(1)[0]
    two = _launch(2)
    _, g1, p1, l1, v1 = one
    for _, g, p, l, v in two:
        torch.testing.assert_close(g, g1, rtol=2e-4, atol=1e-7)      # summed per-view grads == all-view grads
        torch.testing.assert_close(p, p1, rtol=1e-5, atol=1e-6)      # identical update everywhere
        assert abs(l - l1) < 1e-5 * max(1.0, abs(l1))
        assert torch.equal(v, v1)
    assert torch.equal(two[0][2], two[1][2])                          # ranks stay bit-identical


def test_isotropic_grad_matches_autograd():
    g = torch.Generator().manual_seed(0)
    s = torch.randn(50, 3, generator=g, requires_grad=True)
    radii = torch.randint(0, 3, (2, 50), generator=g)
    ref = 5e-4 * R.isotropic_loss(s, radii)
    ref.backward()
    grad = torch.zeros(50, 3)
    val = add_isotropic_grad_(grad, s, radii.sum(0) > 0, 5e-4)
    torch.testing.assert_close(grad, s.grad)
    assert abs(5e-4 * float(val) - float(ref)) < 1e-7


def test_edge_aware_tv_matches_oracle():
    g = torch.Generator().manual_seed(1)
    depth, rgb = torch.rand(2, 12, 17, generator=g), torch.rand(2, 12, 17, 3, generator=g)
    mask = torch.rand(2, 12, 17, generator=g) > 0.4
    torch.testing.assert_close(edge_aware_tv(depth, rgb, mask), R.edge_aware_tv(depth, rgb, mask))


def _run_local(rank, world, port, q):
    """Rank 0 runs a whole LOCAL mapping step while rank 1 goes straight to the barrier (what bench.py's
    sequence leg does under torchrun): no collective may be issued by the step."""
    from gslam_b200.mapping import LOCAL
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    val = None
    if rank == 0:
        sc = make_scene(N_GAUSS, N_VIEWS, "tum", 0.1)
        gt = target_images(N_VIEWS, sc.height, sc.width)
        splats = FlatSplats(sc.params())
        step = MappingStep(splats, _Adam(splats.params), render_fn=_oracle_render, loss_fn=_oracle_loss,
                           depth_tv_weight=1e-6, group=LOCAL)
        res = step(sc.viewmats, sc.Ks, gt, torch.zeros(N_VIEWS, 2), sc.width, sc.height, N_VIEWS)
        val = float(res.total)
    q.put((rank, val))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_local_group_issues_no_collective_inside_a_two_rank_job():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=_run_local, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    one = _launch(1)[0]
    assert out[1] is None and abs(out[0] - one[3]) < 1e-4 * max(1.0, abs(one[3]))
