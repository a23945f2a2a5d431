"""CPU oracle of the eye-camera ray caster (test infrastructure only, like the rest of oracle/): the camera and shading model of
flybody_b200/csrc/fly_eyes.cuh restated in float64 numpy with a DIFFERENT visibility algorithm - every ray is intersected with
every triangle of the terrain (brute force) instead of walking the grid - so that the kernel's traversal is checked, not copied.
Follows the MuJoCo camera convention (x right, y up, looking along -z; vnl_ray/fruitfly/fruitfly.py:676-708 for the observables,
fruitfly.xml:335-336 for the eye poses, hills.py:186-225 for the scene).  Meant for small grids (cost: rays x triangles)."""
import numpy as np


def render_bruteforce(height, cam12, half=20.0, zoff=-0.01, size=32, fovy=150.0, rgb=(0.2, 0.3, 0.4), sky=(0.0, 0.0, 0.0),
                      ambient=0.4, diffuse=0.8):
    h = np.maximum(np.asarray(height, np.float64) + zoff, 0.0)
    nrow, ncol = h.shape
    xs, ys = np.linspace(-half, half, ncol), np.linspace(-half, half, nrow)
    X, Y = np.meshgrid(xs, ys)
    V = np.stack([X, Y, h], -1)
    A, B, C, D = V[:-1, :-1], V[1:, :-1], V[:-1, 1:], V[1:, 1:]
    tri = np.concatenate([np.stack([A, B, C], -2).reshape(-1, 3, 3), np.stack([B, C, D], -2).reshape(-1, 3, 3)])
    e1, e2 = tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
    nrm = np.cross(e1, e2)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    pos, mat = np.asarray(cam12[:3], np.float64), np.asarray(cam12[3:], np.float64).reshape(3, 3)
    th = np.tan(np.deg2rad(fovy) / 2)
    j, i = np.meshgrid(np.arange(size), np.arange(size))
    dc = np.stack([th * (2 * (j + 0.5) / size - 1), th * (1 - 2 * (i + 0.5) / size), -np.ones_like(j, float)], -1).reshape(-1, 3)
    d = dc @ mat.T
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    fwd = -mat[:, 2]
    out = np.zeros((size * size, 3))
    for k in range(size * size):                         # one ray against all triangles
        p = np.cross(d[k], e2)
        det = (e1 * p).sum(1)
        ok = np.abs(det) > 1e-12
        inv = np.where(ok, 1.0 / np.where(ok, det, 1.0), 0.0)
        s = pos - tri[:, 0]
        u = (s * p).sum(1) * inv
        q = np.cross(s, e1)
        v = (q @ d[k]) * inv
        t = (e2 * q).sum(1) * inv
        hit = ok & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 0)
        if hit.any():
            f = np.where(hit, t, np.inf).argmin()
            lam = abs(nrm[f] @ fwd)
            out[k] = np.minimum(1.0, np.asarray(rgb) * (ambient + diffuse * lam))
        else:
            out[k] = sky
    return np.floor(out * 255 + 0.5).astype(np.uint8).reshape(size, size, 3)
