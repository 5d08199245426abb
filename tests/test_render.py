"""Headless picture of the crowd: lc_render (CUDA disc splat, learn_comms_b200/csrc/splat.cu) against
the CPU restatement of render::draw's picture and comm_vis colour maps (oracle/canon.c canon_render / canon_comm_vis;
render.cpp:87-172,352-355,364-415).  The reference's renderer needs an OpenGL context and ffmpeg, neither of which exists
on this box, so this oracle is a restatement only ("unpinned")."""
import numpy as np
import pytest

import _oracle


def test_comm_vis_known_answers():
    """Hand-derived values of the colour maps (render.cpp:87-172)."""
    # X2RG / Y2RG / Z2RG: one channel, red below zero, green above (render.cpp:157-170)
    assert np.array_equal(_oracle.comm_vis(1, [0.25, -1, 9]), np.array([0, 0.25, 0], np.float32))
    assert np.array_equal(_oracle.comm_vis(2, [0.25, -0.5, 9]), np.array([0.5, 0, 0], np.float32))
    assert np.array_equal(_oracle.comm_vis(3, [0, 0, 0.75]), np.array([0, 0.75, 0], np.float32))
    # xyz2 of a grey message (all channels equal): A = B = 0 up to rounding, L = sqrt(|c| / sqrt 3) -> Luv map gives (L, L, L)
    luv = _oracle.comm_vis(4, [0.5, 0.5, 0.5])
    assert np.allclose(luv, np.sqrt(0.5), atol=1e-6)
    lab = _oracle.comm_vis(0, [0.5, 0.5, 0.5])   # Lab map of a colourless input is grey too
    assert np.ptp(lab) < 2e-4 and 0.3 < lab[0] < 0.6
    # the zero message (every agent before its first comm::run) is black in both perceptual maps
    assert not _oracle.comm_vis(4, [0, 0, 0]).any() and _oracle.comm_vis(0, [0, 0, 0]).max() < 0.02


@pytest.mark.gpu
@pytest.mark.parametrize("scene,frames,cam", [(3, 200, 0.85), (9, 350, 0.6), (12, 60, 1.0)])
def test_lc_render_matches_the_cpu_picture(scene, frames, cam):
    import learn_comms_b200 as lcb
    from learn_comms_b200 import workloads as wl
    pos, goal, md, rng, walls = lcb.scene_full(scene, 0)
    m_cv, m_f = wl.weights(scene + 3)
    g = lcb.Sim(pos, goal, m_cv, m_f, max_duration=md, rng_state=rng, walls=walls, order="reference")
    g.step(frames)
    st = g.state()
    for vis in range(5):
        img = g.render(vis, cam_dist=cam)
        ref = _oracle.render(st["pos"], st["c"], walls, vis, cam_dist=cam)
        assert img.shape == (500, 500, 4) and (img[..., 3] == 255).all()
        d = np.abs(img.astype(np.int16) - ref.astype(np.int16))
        assert d.max() <= 1, f"vis {vis}: a channel differs by {d.max()}"          # pow() may differ in the last ulp
        assert (d > 0).any(axis=2).mean() < 1e-3
        assert tuple(img[0, 0]) == (253, 246, 227, 255) or len(walls)               # render.cpp:365 background
        painted = (img[..., :3] != np.array([253, 246, 227], np.uint8)).any(axis=2).sum()
        assert painted > 50
    small = g.render(0, cam_dist=cam, width=64, height=48)
    assert small.shape == (48, 64, 4)
    assert np.abs(small.astype(np.int16) - _oracle.render(st["pos"], st["c"], walls, 0, cam, 64, 48).astype(np.int16)).max() <= 1
