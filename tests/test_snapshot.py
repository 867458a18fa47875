"""Debug-snapshot tuples (reference diff_gaussian_rasterization/__init__.py:49-80, :91-120) as a fixture format."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rain_b200 import snapshot  # noqa: E402
from rain_b200.synthetic import make_camera, make_grad, make_scene  # noqa: E402


def _fw_args(P=2000, W=96, H=64, D=2):
    sc, cam = make_scene(P, D, 3), make_camera(1, 4, W, H)
    e = torch.empty(0)
    return (torch.zeros(3), sc.means3D, e, sc.opacities, sc.scales, sc.rotations, 1.0, e, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy,
            H, W, sc.shs, D, cam.campos, False, False, 0.3), sc


def test_forward_snapshot_round_trips_and_identifies_itself(tmp_path):
    args, _ = _fw_args()
    p = str(tmp_path / "snapshot_fw.dump")
    snapshot.save_forward_snapshot(p, args)
    kind, back = snapshot.load_snapshot(p)
    assert kind == "forward" and len(back) == len(snapshot.FORWARD_FIELDS)
    for a, b in zip(args, back):
        assert torch.equal(a, b) if isinstance(a, torch.Tensor) else a == b
    # an upstream tuple (no low_pass) gets the upstream constant
    torch.save(tuple(args[:-1]), p)
    kind, back = snapshot.load_snapshot(p)
    assert kind == "forward" and back[-1] == 0.3
    torch.save((1, 2, 3), p)
    with pytest.raises(RuntimeError):
        snapshot.load_snapshot(p)


@pytest.mark.gpu
def test_replay_matches_the_direct_call(tmp_path):
    from rain_b200 import _C
    dev = torch.device("cuda:0")
    args, sc = _fw_args(P=20000, W=160, H=96, D=3)
    pf, pb = str(tmp_path / "snapshot_fw.dump"), str(tmp_path / "snapshot_bw.dump")
    snapshot.save_forward_snapshot(pf, args)
    direct = _C.rasterize_gaussians(*(a.to(dev) if isinstance(a, torch.Tensor) else a for a in args))
    replay = snapshot.replay_forward(pf, dev)
    assert replay[0] == direct[0] and all(torch.equal(a, b) for a, b in zip(replay[1:4], direct[1:4]))
    N, color, radii, depth, geom, binning, img = direct
    dpix = make_grad(160, 96, 5).to(dev)
    bw = (args[0], args[1], radii, args[2], args[4], args[5], args[6], args[7], args[8], args[9], args[10], args[11], dpix, args[14], args[15], args[16],
          geom, N, binning, img, False, 0.3)
    snapshot.save_backward_snapshot(pb, bw)
    g_direct = _C.rasterize_gaussians_backward(*(a.to(dev) if isinstance(a, torch.Tensor) else a for a in bw))
    for g_replay in (snapshot.replay_backward(pb, dev), snapshot.replay_backward(pb, dev, rerun_forward=True, opacities=sc.opacities)):
        for a, b in zip(g_replay, g_direct):
            assert (a - b).abs().max() <= 2e-5 * b.abs().max() + 1e-12          # run-to-run spread of the float atomics only
