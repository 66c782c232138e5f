"""Backend -> Frontend map hand-off across a process boundary (gslam_b200.handoff): a spawned
consumer process attaches to the publisher's shared slots through CUDA IPC, snapshots every published
version into a Tracker and renders it; the render must be bit-identical to the publisher's own render of
the same map, versions must arrive in order and untorn, and the tracker's graphs must survive the syncs."""
import pytest
import torch
import torch.multiprocessing as mp

from gslam_b200.mapping import FlatSplats
from gslam_b200.synthetic import make_scene

pytestmark = pytest.mark.gpu

N, CAP, SCALE = 30_000, 40_000, 0.5


def _maps():
    """Three versions of a map: different sizes and contents; row 0 carries the version in every block."""
    out = []
    for v, n in ((1, N), (2, N - 5_000), (3, N + 7_000)):
        sc = make_scene(n, 1, "tum", SCALE, seed=10 + v)
        ps = [p.clone() for p in sc.params()]
        ps[0][0, 0] = float(v)
        ps[5][0] = float(v)
        out.append((sc, ps))
    return out


def _render(splats_like, sc, dev):
    from gslam_b200.primitives import Camera
    cam = Camera(sc.Ks[0].clone().to(dev), sc.height, sc.width)
    with torch.no_grad():
        return splats_like.render([cam], [sc.viewmats[0].to(dev)]).render.clone()


def _consumer(handle_q, result_q, go_q):
    import numpy as np  # noqa: F401
    from gslam_b200.handoff import MapSubscriber
    from gslam_b200.primitives import Camera
    from gslam_b200.tracking import Tracker, TrackingConfig
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    handle = handle_q.get(timeout=120)
    sub = MapSubscriber(handle)
    sc0 = make_scene(8, 1, "tum", SCALE)
    tracker = Tracker(Camera(sc0.Ks[0].clone().to(dev), sc0.height, sc0.width), TrackingConfig(), dev)
    result_q.put(("ready", None))
    seen = []
    for want in (1, 2, 3):
        go_q.get(timeout=120)
        sub.wait_for(want)
        v = sub.snapshot(tracker)
        assert v is not None and v >= want
        cam = tracker.camera
        with torch.no_grad():
            img = tracker.splats.render([cam], [torch.eye(4, device=dev)]).render
        # untorn: every block of the snapshot carries the same version stamp
        stamp = (float(tracker.splats.means[0, 0]), float(tracker.splats.log_uncertainties[0]))
        seen.append(v)
        result_q.put(("snap", (v, tracker.n_live, stamp, img.cpu().numpy(), tracker.capacity)))
    assert sub.snapshot(tracker) is None          # nothing newer: a poll, no copy
    result_q.put(("done", seen))


@pytest.mark.timeout(600)
def test_published_maps_render_bit_identically_in_the_consumer_process(cuda):
    from gslam_b200.handoff import MapPublisher
    from gslam_b200.map import GaussianSplattingData
    ctx = mp.get_context("spawn")
    handle_q, result_q, go_q = ctx.Queue(), ctx.Queue(), ctx.Queue()
    proc = ctx.Process(target=_consumer, args=(handle_q, result_q, go_q))
    proc.start()
    try:
        pub = MapPublisher(CAP, cuda)
        handle_q.put(pub.handle())
        assert result_q.get(timeout=300)[0] == "ready"
        caps = set()
        for v, (sc, ps) in enumerate(_maps(), start=1):
            splats = FlatSplats([p.to(cuda) for p in ps])
            assert pub.publish(splats) and pub.version == v
            go_q.put(v)
            kind, (got_v, n_live, stamp, img, cap) = result_q.get(timeout=300)
            assert kind == "snap" and got_v == v and n_live == splats.n and stamp == (float(v), float(v))
            caps.add(cap)
            # the publisher's own render of the same map (identity pose, same camera)
            mine = GaussianSplattingData(*[p.detach() for p in splats.params],
                                         torch.zeros(splats.n, dtype=torch.long, device=cuda))
            from gslam_b200.primitives import Camera
            cam = Camera(sc.Ks[0].clone().to(cuda), sc.height, sc.width)
            with torch.no_grad():
                ref = mine.render([cam], [torch.eye(4, device=cuda)]).render
            assert torch.equal(torch.from_numpy(img), ref.cpu())
        assert len(caps) == 1                     # the tracker never re-allocated: its graphs survive the syncs
        kind, seen = result_q.get(timeout=120)
        assert kind == "done" and seen == [1, 2, 3] and pub.skipped == 0
    finally:
        proc.join(timeout=60)
        if proc.is_alive():
            proc.kill()
    assert proc.exitcode == 0


def test_publisher_refuses_to_tear_a_snapshot_in_progress(cuda):
    """Protocol only (one process): a consumer that has claimed version v but not confirmed it blocks the
    publish of v + 2 (same slot); once confirmed the publish goes through."""
    from gslam_b200.handoff import MapPublisher, MapSubscriber
    pub = MapPublisher(2_000, cuda)
    sub = MapSubscriber(pub.handle(), event=pub.event)
    sc = make_scene(1_500, 1, "tum", 0.25)
    splats = FlatSplats([p.to(cuda) for p in sc.params()])
    assert pub.publish(splats)                                  # v1 -> slot 1
    assert sub.poll() == (1, 1, 1_500)
    sub.claim[0] = 1                                            # the consumer starts copying v1 ...
    assert pub.publish(splats)                                  # v2 -> slot 0: fine
    assert not pub.publish(splats) and pub.skipped == 1         # v3 -> slot 1: would tear v1
    sub.hdr[1].fill_(1)                                         # ... and finishes
    assert pub.publish(splats) and pub.version == 3
    with pytest.raises(RuntimeError):
        pub.publish(FlatSplats([p.to(cuda) for p in make_scene(2_001, 1, "tum", 0.25).params()]))
