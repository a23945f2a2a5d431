"""Mesh loading and mass properties for the host model compiler.

The fly MJCF (`vnl_ray/fruitfly/assets/fruitfly.xml:209-306`) references 85 OBJ
meshes; six of them are absent from the checkout (`.MISSING_LARGE_BLOBS`) but
five have legacy-binary `.msh` copies under `fruitfly/build_fruitfly/assets/`.
`head_red` exists in neither form; `compile_fly.py` documents the stand-in.

Mass properties are the exact signed-tetrahedron integrals of the closed
triangle surface (MuJoCo's `inertia="exact"` semantics).  The result is
returned in the mesh's own coordinates; the caller applies the geom pose.
"""
from __future__ import annotations

import os
import numpy as np


def load_obj(path: str) -> tuple[np.ndarray, np.ndarray]:
    """Returns (verts [V,3] float64, faces [F,3] int64) of a triangulated OBJ."""
    verts = []
    faces = []
    with open(path, "r") as f:
        for line in f:
            if line.startswith("v "):
                p = line.split()
                verts.append((float(p[1]), float(p[2]), float(p[3])))
            elif line.startswith("f "):
                idx = [int(tok.split("/")[0]) - 1 for tok in line.split()[1:]]
                for k in range(1, len(idx) - 1):  # fan-triangulate polygons
                    faces.append((idx[0], idx[k], idx[k + 1]))
    return np.asarray(verts, dtype=np.float64), np.asarray(faces, dtype=np.int64)


def load_msh(path: str) -> tuple[np.ndarray, np.ndarray]:
    """Legacy MuJoCo binary mesh: int32 (nvert, nnormal, ntexcoord, nface),
    float32 verts / normals / texcoords, int32 faces."""
    raw = open(path, "rb").read()
    nv, nn, nt, nf = np.frombuffer(raw[:16], dtype=np.int32)
    off = 16
    verts = np.frombuffer(raw, dtype=np.float32, count=3 * nv, offset=off).reshape(nv, 3)
    off += 4 * (3 * nv + 3 * nn + 2 * nt)
    faces = np.frombuffer(raw, dtype=np.int32, count=3 * nf, offset=off).reshape(nf, 3)
    return verts.astype(np.float64), faces.astype(np.int64)


def load_mesh(asset_dir: str, filename: str, fallback_dir: str | None = None):
    """Loads `filename` (an .obj name from the MJCF) or its .msh fallback.
    Returns None when neither exists."""
    p = os.path.join(asset_dir, filename)
    if os.path.exists(p):
        return load_obj(p)
    stem = os.path.splitext(filename)[0]
    for d in (asset_dir, fallback_dir):
        if d is None:
            continue
        q = os.path.join(d, stem + ".msh")
        if os.path.exists(q):
            return load_msh(q)
    return None


def mesh_mass_properties(verts: np.ndarray, faces: np.ndarray, scale=(1.0, 1.0, 1.0), mode: str = "legacy"):
    """Unit-density volume, centre of mass and inertia tensor about the centre
    of mass (3x3, mesh axes) of the triangle surface.

    mode="legacy": tetrahedra from the area-weighted surface centroid to every
    face, each counted with |volume| (MuJoCo's original mesh-inertia rule).  The
    fly model was built against this rule: with it the abdomen comes out at
    0.381 mg and a leg at 0.01629 mg, the empirical targets of
    `build_fruitfly/make_fruitfly.py:24` (0.38 / 0.0162); the abdomen tergite
    and sternite meshes are open shells, for which signed volumes give 0.705 mg.
    mode="exact": signed tetrahedra (closed, consistently wound surfaces only).
    """
    v = verts * np.asarray(scale, dtype=np.float64)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    if mode == "legacy":
        area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1)
        cen = ((a + b + c) / 3.0 * area[:, None]).sum(0) / area.sum()
    else:
        cen = np.zeros(3)
    a, b, c = a - cen, b - cen, c - cen
    det = np.einsum("ij,ij->i", a, np.cross(b, c))  # 6 * signed tet volume
    if mode == "legacy":
        det = np.abs(det)
    vol = det.sum() / 6.0
    if vol < 0:  # inward-facing winding
        det = -det
        vol = -vol
    com = ((a + b + c) * det[:, None]).sum(0) / (24.0 * vol)
    # second moments  integral x_i x_j dV  over each tet (0,a,b,c)
    s = a + b + c
    C = (
        np.einsum("n,ni,nj->ij", det, a, a)
        + np.einsum("n,ni,nj->ij", det, b, b)
        + np.einsum("n,ni,nj->ij", det, c, c)
        + np.einsum("n,ni,nj->ij", det, s, s)
    ) / 120.0
    C -= vol * np.outer(com, com)  # shift to centre of mass
    inertia = np.trace(C) * np.eye(3) - C
    return vol, com + cen, inertia
