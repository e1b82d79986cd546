#!/usr/bin/env python
"""CPU design tool (not product code): where do C3's node visits go?  The C3 scene (100 yaw-rotated instances of the bumpy sphere
on a 2.5 grid) is traced on the host with the device pipeline's BLAS (tools/bvhlab.c): per ray, the instances whose world box
it crosses are visited front to back with the closest hit carried along, exactly what the TLAS does, and BLAS node visits are
split into those spent in instances the ray HITS and in instances it only passes through (near misses).
   python tools/bvh_lab_c3.py [N M] [width] [root|cut]     default 500 1000, 384x216 pixels, primary + reflection rays to depth 4"""
import ctypes as C, math, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytracing_b200 import synth

so = "/tmp/bvhlab.so"
subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tools", "bvhlab.c"), "-lm"])
L = C.CDLL(so)
L.lab_build.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
L.lab_trace_t.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]

N, M = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (500, 1000)
WIDTH = int(sys.argv[3]) if len(sys.argv) > 3 else 4
BOXES = sys.argv[4] if len(sys.argv) > 4 else "cut"
mesh = synth.bumpy_sphere(N, M)
tris = np.ascontiguousarray(mesh["v"].reshape(-1, 9), np.float32)
nrm = np.ascontiguousarray(mesh["n"], np.float32)
stats = np.zeros(2, np.float32)
L.lab_build(tris.ctypes.data, len(tris), 0, 10, 8, WIDTH, stats.ctypes.data)
poses = synth.instance_grid(10, 10)
lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)

def rot(yaw):                       # mesh -> world is R^T (.) + t with R = world -> mesh yaw rotation (transforms.hpp:129-144, pitch = roll = 0)
    c, s = math.cos(yaw), math.sin(yaw)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], np.float64)

Rs = [rot(p[3]) for p in poses]
Ts = [np.array(p[:3], np.float64) for p in poses]
# world boxes: transformed root box (the loose choice) or the box of the transformed vertices (what a fine BLAS cut converges to)
wlo, whi = [], []
corners = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float64)
sample = tris.reshape(-1, 3)[::37].astype(np.float64)
for R, t in zip(Rs, Ts):
    pts = (corners if BOXES == "root" else sample) @ R + t          # row vectors: p R = R^T p
    wlo.append(pts.min(0) - 1e-3); whi.append(pts.max(0) + 1e-3)
wlo, whi = np.array(wlo), np.array(whi)

W, H = 384, 216
K = synth.REF_K.astype(np.float64) * np.array([[2.0], [2.0], [1.0]]) * np.array([[W / 3840.0], [H / 2160.0], [1.0]])
a = 0.0
cam = np.array([19.0 * math.sin(a), -19.0 * math.cos(a), 9.0])
pitch = -0.42
xs, ys = np.meshgrid(np.arange(W) + 0.0, np.arange(H) + 0.0)
dc = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1).reshape(-1, 3)
dc /= np.linalg.norm(dc, axis=1, keepdims=True)
d = np.stack([dc[:, 0], dc[:, 2], -dc[:, 1]], -1)                   # raycast.cu:182 swizzle: camera looks along +y, z up
cp, sp_ = math.cos(pitch), math.sin(pitch)
d = np.stack([d[:, 0], cp * d[:, 1] - sp_ * d[:, 2], sp_ * d[:, 1] + cp * d[:, 2]], -1)
o = np.tile(cam, (len(d), 1))

def trace(o, d, prev):
    """closest hit over all instances, front to back; returns t, normal, and the visit split (prev = instance a ray starts on, -1 = camera)"""
    n = len(o)
    inv = 1.0 / np.where(np.abs(d) < 1e-20, 1e-20, d)
    t0 = (wlo[None] - o[:, None]) * inv[:, None]; t1 = (whi[None] - o[:, None]) * inv[:, None]
    tn = np.minimum(t0, t1).max(2).clip(min=0); tf = np.maximum(t0, t1).min(2)
    crossed = tn <= tf                                               # n x 100
    best = np.full(n, np.inf); bn = np.zeros((n, 3)); binst = np.full(n, -1)
    visits_hit = np.zeros(n); visits_miss = np.zeros(n); visits_self = np.zeros(n); entries = np.zeros(n); tri_tests = np.zeros(n)
    order = np.argsort(np.where(crossed, tn, np.inf), axis=1)
    for rank in range(100):
        inst = order[:, rank]
        ent = np.take_along_axis(tn, order[:, rank:rank + 1], 1)[:, 0]
        act = np.take_along_axis(crossed, order[:, rank:rank + 1], 1)[:, 0] & (ent < best)
        if not act.any():
            break
        idx = np.nonzero(act)[0]
        for i in np.unique(inst[idx]):
            sel = idx[inst[idx] == i]
            R, t = Rs[i], Ts[i]
            ro = (o[sel] - t) @ R.T; rd = d[sel] @ R.T               # world -> mesh: R (w - t)
            rays = np.ascontiguousarray(np.concatenate([ro, rd], 1), np.float32)
            out = np.zeros((len(sel), 3), np.int32)
            tio = np.zeros((len(sel), 2), np.float32); tio[:, 0] = np.where(np.isinf(best[sel]), 3e38, best[sel])
            L.lab_trace_t(rays.ctypes.data, len(sel), out.ctypes.data, tio.ctypes.data)
            h = out[:, 2] > 0
            entries[sel] += 1; tri_tests[sel] += out[:, 1]
            own = prev[sel] == i
            visits_hit[sel[h]] += out[h, 0]; visits_miss[sel[~h & ~own]] += out[~h & ~own, 0]; visits_self[sel[~h & own]] += out[~h & own, 0]
            best[sel[h]] = tio[h, 0]; binst[sel[h]] = i
            bn[sel[h]] = nrm[tio[h, 1].astype(int)] @ R                # mesh normal -> world: R^T n
    return best, bn, binst, visits_hit, visits_miss, visits_self, entries, tri_tests, crossed.sum(1)

print("C3 on the host: %d triangles/BLAS, %d-wide nodes, instance boxes: %s, %dx%d pixels" % (len(tris), WIDTH, BOXES, W, H))
print("%-10s %9s %7s | %9s %12s %12s %12s %9s %9s" % ("bounce", "rays", "hit %", "boxes/ray", "visits: hit", "own inst.", "other miss", "entries", "tris"))
tot = np.zeros(6); nrays = 0
prev = np.full(len(o), -1)
for depth in range(5):
    t, nw, inst, vh, vm, vs, ent, tt, nb = trace(o, d, prev)
    n = len(o)
    print("%-10d %9d %7.1f | %9.2f %12.2f %12.2f %12.2f %9.2f %9.2f" % (depth, n, 100 * np.isfinite(t).mean(), nb.mean(), vh.mean(), vs.mean(), vm.mean(), ent.mean(), tt.mean()), flush=True)
    tot += np.array([vh.sum(), vs.sum(), vm.sum(), ent.sum(), tt.sum(), nb.sum()]); nrays += n
    hit = np.isfinite(t)
    if not hit.any():
        break
    p = o[hit] + d[hit] * t[hit, None]
    nn = nw[hit] / np.linalg.norm(nw[hit], axis=1, keepdims=True)
    dd = d[hit]
    nn = np.where((dd * nn).sum(1, keepdims=True) > 0, -nn, nn)
    d = dd - 2 * (dd * nn).sum(1, keepdims=True) * nn
    o = p + 1e-4 * nn
    prev = inst[hit]
print("all closest-hit rays: %.2f BLAS visits per ray in instances that are hit, %.2f leaving the instance the ray starts on, %.2f in other instances it misses, "
      "%.2f instance entries, %.2f triangle tests, %.2f boxes crossed" % tuple(tot / nrays))
