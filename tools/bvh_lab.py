#!/usr/bin/env python
"""CPU design tool (not product code): wide-node visits / triangle tests per ray on the C3 BLAS for different builders and
node widths, with the device pipeline's own leaf decision, collapse and traversal order (tools/bvhlab.c).
   python tools/bvh_lab.py [N M] [rays]        default 500 1000 (the 1 M-triangle bumpy sphere), 200000 rays
Rays are random chords through the mesh's bounding box (what a ray that entered the instance's TLAS box looks like)."""
import ctypes as C, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytracing_b200 import synth

so = "/tmp/bvhlab.so"
subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tools", "bvhlab.c"), "-lm"])
L = C.CDLL(so)
L.lab_build.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
L.lab_trace.argtypes = [C.c_void_p, C.c_int, C.c_void_p]

N, M = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (500, 1000)
R = int(sys.argv[3]) if len(sys.argv) > 3 else 200000
tris = np.ascontiguousarray(synth.bumpy_sphere(N, M)["v"].reshape(-1, 9), np.float32)
lo, hi = tris.reshape(-1, 3).min(0), tris.reshape(-1, 3).max(0)
rng = np.random.default_rng(7)
d = rng.normal(size=(R, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
target = rng.uniform(lo, hi, (R, 3))
org = target - d * 4.0
rays = np.ascontiguousarray(np.concatenate([org, d], axis=1), np.float32)
print("mesh %d triangles, %d rays through its box" % (len(tris), R))
print("%-22s %5s %9s %8s | %14s %14s | %14s %14s" % ("builder", "width", "nodes", "leaves", "hit: nodes", "tris", "miss: nodes", "tris"))
ref = None
L.lab_set_quant.argtypes = [C.c_int]
L.lab_set_order.argtypes = [C.c_int]
for name, builder, bits, quant, order in (("LBVH 10 bits/axis", 0, 10, 0, 0), ("LBVH 16 bits/axis", 0, 16, 0, 0), ("binned SAH (16 bins)", 1, 0, 0, 0),
                                          ("LBVH, 8-bit child boxes", 0, 10, 8, 0), ("LBVH, 6-bit child boxes", 0, 10, 6, 0),
                                          ("LBVH, nearest-first only", 0, 10, 0, 1), ("LBVH, slot order", 0, 10, 0, 2),
                                          ("LBVH, octant-diagonal order", 0, 10, 0, 3), ("LBVH, 8-bit + octant order", 0, 10, 8, 3)):
    L.lab_set_quant(quant)
    L.lab_set_order(order)
    for width in (4, 8):
        stats = np.zeros(2, np.float32)
        t0 = time.time()
        nw = L.lab_build(tris.ctypes.data, len(tris), builder, bits, 8, width, stats.ctypes.data)
        out = np.zeros((R, 3), np.int32)
        L.lab_trace(rays.ctypes.data, R, out.ctypes.data)
        hit = out[:, 2] > 0
        if ref is None:
            ref = hit
        assert np.array_equal(hit, ref), "builders disagree on which rays hit"
        print("%-22s %5d %9d %8d | %14.2f %14.2f | %14.2f %14.2f   (hit %.1f%%, sah %.1f, %.1f s)" % (
            name, width, nw, int(stats[0]), out[hit, 0].mean(), out[hit, 1].mean(), out[~hit, 0].mean(), out[~hit, 1].mean(),
            100 * hit.mean(), stats[1], time.time() - t0), flush=True)
