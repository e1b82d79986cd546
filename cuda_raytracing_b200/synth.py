"""Procedural, seeded synthetic inputs shared by bench.py and tests/.

Pure numpy; produces triangle arrays in the reference's 72-byte TrianglePrimitive layout
(TrianglePrimitive.hpp:8-11) with face normals computed the reference's way
(OBJLoader.hpp:143-147: normalize(cross(v_i - v_0, v_{i+1} - v_0)) with the Q_rsqrt normalize of
utils.hpp:12-47, 32-bit pun).  All arithmetic is float32 without fused multiply-add.
"""
import math

import numpy as np

TRI = np.dtype([("v", "<f4", (3, 3)), ("n", "<f4", (3,)), ("uv", "<f4", (3, 2))])
assert TRI.itemsize == 72

# the reference's demo intrinsics / fisheye coefficients (kernel.cu:158-164)
REF_K = np.array([[862.097835972576, 0.0, 998.1702383680802],
                  [0.0, 862.1368447300727, 569.6759403225842],
                  [0.0, 0.0, 1.0]], np.float32)
REF_D = np.array([0.016233999489849514, -0.013875757716177956, 0.03264329940126211, -0.019561619947134234], np.float32)
C1_K = np.array([[400.0, 0.0, 320.37], [0.0, 400.0, 240.21], [0.0, 0.0, 1.0]], np.float32)
ZERO_D = np.zeros(4, np.float32)


def q_rsqrt(x):
    """utils.hpp:12-27 on float32 arrays (32-bit pun)."""
    x = np.asarray(x, np.float32)
    x2 = x * np.float32(0.5)
    i = x.view(np.int32)
    i = np.int32(0x5F3759DF) - (i >> 1)
    y = i.view(np.float32)
    return y * (np.float32(1.5) - (x2 * y * y))


def normalize(v):
    """utils.hpp:37-47."""
    v = np.asarray(v, np.float32)
    inv = q_rsqrt(v[..., 0] * v[..., 0] + v[..., 1] * v[..., 1] + v[..., 2] * v[..., 2])
    return v * inv[..., None]


def cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1],
                     a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], -1).astype(np.float32)


def make_triangles(verts, faces, uvs=None):
    """verts (V,3) float32, faces (F,3) int -> TRI array with reference face normals."""
    verts = np.asarray(verts, np.float32)
    faces = np.asarray(faces)
    t = np.zeros(len(faces), TRI)
    a, b, c = verts[faces[:, 0]], verts[faces[:, 1]], verts[faces[:, 2]]
    t["v"][:, 0], t["v"][:, 1], t["v"][:, 2] = a, b, c
    with np.errstate(all="ignore"):
        t["n"] = normalize(cross(b - a, c - a))
    if uvs is not None:
        t["uv"] = np.asarray(uvs, np.float32)
    return t


def bumpy_sphere_mesh(N, M, phase=0.0, amp=0.05):
    """Appendix A.6 generator G(N,M): vertices (rounded to 6 decimals like the OBJ text) and faces.

    2*N*M triangles.  r = 1 + amp*sin(7*theta + phase)*cos(5*phi), centre (0,0,1).
    """
    i = np.arange(N + 1, dtype=np.float64)[:, None]
    j = np.arange(M, dtype=np.float64)[None, :]
    th = math.pi * i / N
    ph = 2.0 * math.pi * j / M
    r = 1.0 + amp * np.sin(7.0 * th + phase) * np.cos(5.0 * ph)
    v = np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), 1.0 + r * np.cos(th)], -1).reshape(-1, 3)
    v = np.round(v, 6) + 0.0           # "%.6f"; +0.0 folds -0.0
    ii = np.arange(N)[:, None]
    jj = np.arange(M)[None, :]
    a = ii * M + jj
    b = (ii + 1) * M + jj
    c = (ii + 1) * M + (jj + 1) % M
    d = ii * M + (jj + 1) % M
    f = np.stack([np.stack([a, b, c], -1), np.stack([a, c, d], -1)], 2).reshape(-1, 3)
    return v.astype(np.float32), f.astype(np.int64)


def bumpy_sphere(N, M, phase=0.0, ground=False):
    v, f = bumpy_sphere_mesh(N, M, phase)
    t = make_triangles(v, f)
    if ground:
        t = np.concatenate([t, ground_quad()])
    return t


def ground_quad(z=-0.1, half=30.0):
    """Two triangles facing +z, below the sphere's lowest point (-0.05) so nothing interpenetrates."""
    v = np.array([[-half, -half, z], [half, -half, z], [half, half, z], [-half, half, z]], np.float32)
    return make_triangles(v, np.array([[0, 1, 2], [0, 2, 3]]))


def write_obj(path, verts, faces, vts=None, face_vts=None):
    """OBJ text with %.6f coordinates; faces 1-based; optional vt indices (v/vt tokens)."""
    with open(path, "w") as fh:
        fh.write("# procedural mesh\n")
        fh.write("".join("v %.6f %.6f %.6f\n" % (float(x), float(y), float(z)) for x, y, z in verts))
        if vts is not None:
            fh.write("".join("vt %.6f %.6f\n" % (float(u), float(w)) for u, w in vts))
        if face_vts is None:
            fh.write("".join("f %d %d %d\n" % (a + 1, b + 1, c + 1) for a, b, c in faces))
        else:
            fh.write("".join("f %d/%d %d/%d %d/%d\n" % (a + 1, ta + 1, b + 1, tb + 1, c + 1, tc + 1)
                             for (a, b, c), (ta, tb, tc) in zip(faces, face_vts)))


def lcg_yaws(n, seed=1234):
    """C3 instance yaws: LCG(seed) mapped to [-pi, pi)."""
    out = []
    s = seed & 0xFFFFFFFF
    for _ in range(n):
        s = (1664525 * s + 1013904223) & 0xFFFFFFFF
        out.append((s / 4294967296.0) * 2.0 * math.pi - math.pi)
    return np.array(out, np.float32)


def instance_grid(nx=10, ny=10, spacing=2.5, seed=1234):
    """C3 layout: nx*ny instances on a grid centred at the origin, random yaw, scale 1. -> list of pose6."""
    yaws = lcg_yaws(nx * ny, seed)
    poses = []
    k = 0
    for iy in range(ny):
        for ix in range(nx):
            x = (ix - (nx - 1) / 2.0) * spacing
            y = (iy - (ny - 1) / 2.0) * spacing
            poses.append((x, y, 0.0, float(yaws[k]), 0.0, 0.0))
            k += 1
    return poses


def random_soup(n, seed=0, extent=1.0, size=0.15):
    """Seeded random triangle soup for small parity cases (includes ragged / sliver shapes)."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, (n, 1, 3))
    v = (c + rng.normal(0, size, (n, 3, 3))).astype(np.float32)
    v[..., 2] += 1.0
    t = np.zeros(n, TRI)
    t["v"] = v
    with np.errstate(all="ignore"):
        t["n"] = normalize(cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]))
    t["uv"] = rng.uniform(0, 1, (n, 3, 2)).astype(np.float32)
    return t
