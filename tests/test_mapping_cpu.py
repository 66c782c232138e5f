"""Host logic of the view-sharded mapping step on the CPU (gloo, world_size 2).

The renderer, the image loss, the packing of the gradient rows and the fused update are injected from
the oracle (tests/cpu_stubs.py; the CUDA kernels need a GPU); what is tested is the sharding arithmetic:
window[r::p] partition, per-rank loss share, the reduce-scatter of the gradient rows, the update of each
rank's own Gaussian range, the all-gather of the updated parameter blocks, and that every rank ends every
step with bit-identical parameters equal to the one-rank result.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gslam_b200.mapping import (DEFAULT_LRS, PARAM_NAMES, FlatSplats, MappingStep, edge_aware_tv, shard_views)
from gslam_b200.synthetic import make_scene, target_images
from oracle import torch_ref as R

import cpu_stubs as S

N_VIEWS, N_GAUSS, N_STEPS = 4, 601, 3        # 601: not a multiple of the rank count (padding rows in play)


def _make_step(splats, **kw):
    adam = S.TorchRowsAdam(splats.params, [DEFAULT_LRS[n] for n in PARAM_NAMES])
    return MappingStep(splats, None, render_fn=S.oracle_render, loss_fn=S.oracle_loss, update_fn=adam.update,
                       collect_grads=S.collect_grads, depth_tv_weight=1e-6, **kw)


def _run(rank, world, port, q, n_views=None):
    N_VIEWS = n_views or globals()["N_VIEWS"]
    if world > 1:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    sc = make_scene(N_GAUSS, N_VIEWS, "tum", 0.1)
    gt = target_images(N_VIEWS, sc.height, sc.width)
    exposure = torch.zeros(N_VIEWS, 2)
    splats = FlatSplats(sc.params(), align=world)
    step = _make_step(splats)
    mine = shard_views(N_VIEWS, rank, world)
    totals = []
    for _ in range(N_STEPS):
        res = step(sc.viewmats[mine], sc.Ks[mine], gt[mine], exposure[mine], sc.width, sc.height, N_VIEWS)
        totals.append(float(res.total))
    # numpy payloads are pickled by value (tensors would be shared through fds that die with the child)
    params = torch.cat([p.detach().reshape(-1) for p in splats.params])
    q.put((rank, params.numpy().copy(), totals, res.vis_count.numpy().copy(), splats.cap))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _launch(world, target=_run):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = [ctx.Process(target=target, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted([q.get(timeout=900) for _ in range(world)], key=lambda t: t[0])
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


def test_flat_splats_blocks_and_padding():
    sc = make_scene(10, 1, "tum", 0.1)
    sp = FlatSplats(sc.params(), align=4)
    assert sp.n == 10 and sp.cap == 12 and sp.flat.numel() == 15 * 12
    for p, t in zip(sp.params, sc.params()):
        assert torch.equal(p.detach(), t) and p.is_contiguous()
    blocks = sp.blocks()
    assert [b.numel() for b in blocks] == [w * 12 for w in (3, 4, 3, 1, 3, 1)]
    assert all(float(b[w * 10:].abs().sum()) == 0.0 for b, w in zip(blocks, sp.widths))
    # SH coefficients [N,K,3] widen the colour block only
    tensors = list(sc.params())
    tensors[4] = torch.randn(10, 16, 3)
    sh = FlatSplats(tensors, align=8)
    assert sh.K == 16 and sh.widths == (3, 4, 3, 1, 48, 1) and sh.cap == 16 and sh.params[4].shape == (10, 16, 3)


@pytest.mark.timeout(1500)
def test_two_ranks_match_one_rank_over_three_steps():
    one = _launch(1)[0]
    two = _launch(2)
    _, p1, l1, v1, cap1 = one
    assert cap1 == N_GAUSS and two[0][4] == N_GAUSS + 1
    for _, p, l, v, _ in two:
        # summed per-view gradients == all-view gradients, so the Adam trajectories agree to rounding
        torch.testing.assert_close(torch.from_numpy(p), torch.from_numpy(p1), rtol=2e-4, atol=2e-6)
        for a, b in zip(l, l1):
            assert abs(a - b) < 1e-5 * max(1.0, abs(b))
        assert (v == v1).all()
    # the ranks stay bit-identical: every row is updated by exactly one rank and all-gathered
    assert (two[0][1] == two[1][1]).all() and two[0][2] == two[1][2]


def _run_three_views(rank, world, port, q):
    _run(rank, world, port, q, n_views=3)


@pytest.mark.timeout(1500)
def test_eight_ranks_with_five_viewless_ranks_match_one_rank():
    """More ranks than views (the first keyframes of a sequence on 8 GPUs): the ranks without a view contribute
    zero rows, still join every collective, update their own Gaussian range and end with the same parameters."""
    one = _launch(1, _run_three_views)[0]
    eight = _launch(8, _run_three_views)
    p1 = torch.from_numpy(one[1])
    for _, p, l, v, cap in eight:
        assert cap == 608                       # 601 rounded up to the 8 ranks
        torch.testing.assert_close(torch.from_numpy(p), p1, rtol=2e-4, atol=2e-6)
        for a, b in zip(l, one[2]):
            assert abs(a - b) < 1e-5 * max(1.0, abs(b))
        assert (v == one[3]).all()
        assert (p == eight[0][1]).all() and l == eight[0][2]


def _run_gather_moments(rank, world, port, q):
    """Every rank holds the true Adam moments only inside its own Gaussian range (garbage elsewhere: what a sharded
    step leaves behind); after gather_adam_state every rank must hold all of them."""
    from types import SimpleNamespace
    from gslam_b200.mapping import gather_adam_state
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = make_scene(N_GAUSS, 1, "tum", 0.1)
    tensors = list(sc.params())
    tensors[4] = torch.randn(N_GAUSS, 4, 3)                    # SH coefficients: a 12-float block
    splats = FlatSplats(tensors, align=world)
    g = torch.Generator().manual_seed(11)
    truth = {k: (torch.randn(p.shape, generator=g), torch.rand(p.shape, generator=g)) for k, p in enumerate(splats.params)}
    per = splats.cap // world
    lo, hi = min(rank * per, N_GAUSS), min((rank + 1) * per, N_GAUSS)
    state = {}
    for k, p in enumerate(splats.params):
        m, v = torch.full(p.shape, float("nan")), torch.full(p.shape, float("nan"))
        m[lo:hi], v[lo:hi] = truth[k][0][lo:hi], truth[k][1][lo:hi]
        state[p] = {"exp_avg": m, "exp_avg_sq": v}
    opt = SimpleNamespace(state=state)
    gather_adam_state(splats, opt, None)
    ok = all(torch.equal(state[p]["exp_avg"], truth[k][0]) and torch.equal(state[p]["exp_avg_sq"], truth[k][1])
             for k, p in enumerate(splats.params))
    q.put((rank, ok, hi - lo))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("world", [3, 8])
def test_adam_moments_are_complete_on_every_rank_before_a_repartition(world):
    out = _launch(world, _run_gather_moments)
    assert all(ok for _, ok, _ in out)
    assert sum(n for _, _, n in out) == N_GAUSS


def test_stub_update_matches_torch_adam_and_autograd_regulariser():
    """The torch restatement the CPU tests inject (and the GPU test compares gsb_map_update with) against
    torch.optim.Adam + autograd of the reference's isotropic loss + the masked opacity decay."""
    from gslam_b200.ops import RowLayout
    g = torch.Generator().manual_seed(0)
    n = 50
    sc = make_scene(n, 1, "tum", 0.1)
    sp = FlatSplats(sc.params())
    ref = [p.detach().clone().requires_grad_(True) for p in sp.params]
    opts = [torch.optim.Adam([p], lr=DEFAULT_LRS[nm]) for p, nm in zip(ref, PARAM_NAMES)]
    lay = RowLayout(1)
    step = _make_step(sp)
    vis_out = torch.zeros(n, dtype=torch.int32)
    for it in range(3):
        rows = torch.zeros(n, lay.stride)
        rows[:, :lay.vis] = torch.randn(n, lay.vis, generator=g) * 1e-2
        radii = torch.randint(0, 3, (3, n), generator=g)
        rows[:, lay.vis] = (radii > 0).sum(0).float()
        iso = 5e-4 * R.isotropic_loss(ref[2], radii)
        iso.backward()
        for p, c, w in zip(ref, lay.cols, lay.widths):
            gr = rows[:, c:c + w].reshape(p.shape)
            p.grad = gr.clone() if p.grad is None else p.grad + gr
        for o in opts:
            o.step()
            o.zero_grad()
        with torch.no_grad():
            ref[3][(radii > 0).sum(0) > 1] *= 0.995
        val = step.update_fn(step, rows, lay, 0, n, vis_out)
        assert abs(5e-4 * float(val) - float(iso)) < 1e-7
        for p, q in zip(sp.params, ref):
            torch.testing.assert_close(p.detach(), q.detach(), rtol=1e-5, atol=1e-7)
    assert torch.equal(vis_out, (radii > 0).sum(0).to(torch.int32))


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
        step = _make_step(splats, group=LOCAL)
        res = step(sc.viewmats, sc.Ks, gt, torch.zeros(N_VIEWS, 2), sc.width, sc.height, N_VIEWS)
        val = float(res.total)
    q.put((rank, val))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(900)
def test_local_group_issues_no_collective_inside_a_two_rank_job():
    out = dict(_launch(2, _run_local))
    one = _launch(1)[0]
    assert out[1] is None and abs(out[0] - one[2][0]) < 1e-4 * max(1.0, abs(one[2][0]))


def _run_prune(rank, world, port, q):
    """Each rank holds the radii / n_touched of ITS views; the reductions of prune_remove_mask must give
    every rank the mask one process computes from all views (MAX of radii, SUM of counts)."""
    from gslam_b200.pruning import prune_remove_mask
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(3)
    C, n = 6, 400
    radii = (torch.randint(0, 400, (C, n), generator=g) * (torch.rand(C, n, generator=g) > 0.3)).to(torch.int32)
    n_touched = torch.randint(0, 3, (C, n), generator=g) * (torch.rand(C, n, generator=g) > 0.5)
    opac = torch.randn(n, generator=g)
    full = prune_remove_mask(opac, radii, n_touched, 0.2, 256, 3, 8, group="local", n_views_total=C)
    mine = shard_views(C, rank, world)
    part = prune_remove_mask(opac, radii[mine], n_touched[mine], 0.2, 256, 3, 8, group=None, n_views_total=C)
    q.put((rank, bool(torch.equal(full, part)), int(full.sum())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_prune_mask_reductions_over_two_ranks():
    out = _launch(2, _run_prune)
    assert all(ok for _, ok, _ in out) and out[0][2] > 0
