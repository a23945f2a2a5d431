"""The eye-camera ray caster (csrc/fly_eyes.cuh through flyb_render_eyes) against the brute-force float64 oracle
(oracle/eyes_ref.py).  Run with -m gpu."""
import numpy as np
import pytest


def _small_terrain(n=41, seed=0):
    rng = np.random.RandomState(seed)
    x = np.linspace(-20, 20, n)
    X, Y = np.meshgrid(x, x)
    return (1.0 + 0.8 * np.sin(0.4 * X + 1.0) * np.cos(0.3 * Y) + 2.0 * (np.hypot(X, Y) > 15) + 0.05 * rng.rand(n, n)).astype(np.float32)


def test_eye_camera_poses_follow_the_model():
    """eye_right / eye_left sit 0.0219 cm to either side of the head's mid-plane and look outward and forward (fruitfly.xml:335-336)."""
    from flybody_b200.engine import eye_cameras
    cams = eye_cameras([1.0, 2.0, 0.7], np.eye(3))
    assert np.allclose(cams[0, :3], [1.0219, 2.0131, 0.7]) and np.allclose(cams[1, :3], [0.9781, 2.0131, 0.7])
    fr, fl = -cams[0, 3:].reshape(3, 3)[:, 2], -cams[1, 3:].reshape(3, 3)[:, 2]     # viewing directions
    assert fr[0] > 0.3 and fl[0] < -0.3 and abs(fr[0] + fl[0]) < 1e-3 and np.allclose(fr[1:], fl[1:], atol=1e-3)
    for c in cams:
        R = c[3:].reshape(3, 3)
        assert np.allclose(R @ R.T, np.eye(3), atol=2e-3)      # (the XML quaternions are given to three digits)


@pytest.mark.gpu
def test_gpu_eye_images_match_the_bruteforce_oracle():
    from flybody_b200.engine import eye_cameras, render_eyes
    from oracle.eyes_ref import render_bruteforce
    h = _small_terrain()
    rng = np.random.RandomState(1)
    cams = []
    for k in range(3):
        yaw, pitch = rng.uniform(-np.pi, np.pi), np.deg2rad(47.5) * rng.uniform(0.5, 1.2)
        cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
        head = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]]) @ np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
        cams.append(eye_cameras([rng.uniform(-5, 5), rng.uniform(-5, 5), 2.5 + rng.uniform(0, 1)], head))
    cams = np.concatenate(cams)
    img = render_eyes(h, np.zeros(len(cams), np.int32), cams, sky=(0.1, 0.2, 0.3))
    assert img.shape == (6, 32, 32, 3)
    worst = 0
    for k in range(len(cams)):
        ref = render_bruteforce(h, cams[k], sky=(0.1, 0.2, 0.3))
        diff = np.abs(img[k].astype(int) - ref.astype(int)).max(-1)
        # float32 rays against float64 ones: a pixel on a silhouette or a triangle edge may pick the neighbouring surface
        assert (diff > 1).mean() < 0.02, (k, (diff > 1).mean())
        worst = max(worst, (diff > 1).mean())
        assert len(np.unique(ref.reshape(-1, 3), axis=0)) > 20      # a real image: shaded slopes and background
    print(f"\neye images: worst share of pixels off by more than 1/255: {worst:.4f}")


@pytest.mark.gpu
def test_gpu_eye_images_over_the_full_size_terrains():
    """16 cameras over a pool of 401 x 401 trench terrains: sanity of the full-size path (the brute-force oracle is for small grids)."""
    from flybody_b200 import terrain as T
    from flybody_b200.engine import eye_cameras, render_eyes
    tab = T.terrain_table("trench", 2, seed=0)
    pitch = np.deg2rad(47.5)
    head = np.array([[np.cos(pitch), 0, np.sin(pitch)], [0, 1, 0], [-np.sin(pitch), 0, np.cos(pitch)]])
    cams = np.concatenate([eye_cameras([-5.0 + 0.3 * k, 0.0, 0.7 + T.nearest_height(tab["height"][k % 2], (20, 20), -5.0 + 0.3 * k, 0.0)], head)
                           for k in range(8)])
    img = render_eyes(tab["height"], np.repeat(np.arange(8) % 2, 2), cams)
    assert img.shape == (16, 32, 32, 3) and img.max() > 60 and (img.reshape(16, -1).std(1) > 1).all()
    assert np.array_equal(img, render_eyes(tab["height"], np.repeat(np.arange(8) % 2, 2), cams))      # deterministic
