#!/usr/bin/env python
"""Rebuild the BLAS of an N x M bumpy sphere a few times (for an ncu launch list of the builder kernels).
   python tools/build_prof.py [N M] [repeats]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from cuda_raytracing_b200 import api, synth

if os.environ.get("RTB_VARIANT"):                           # A/B a differently compiled library (tools/build_variant.sh)
    api.LIB_PATH = os.path.join(ROOT, "tools", "variants", os.environ["RTB_VARIANT"])
N, M = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1582, 3161)
rep = int(sys.argv[3]) if len(sys.argv) > 3 else 3
tris = synth.bumpy_sphere(N, M)
t = torch.from_numpy(tris.view(np.uint8).reshape(-1, 72)).cuda()
scene = api.Scene(0)
if os.environ.get("RTB_BUILD_FUSED"):                       # A/B the fused fit + chunk-local collapse against the one-pass collapse
    scene.set_option("build_fused", int(os.environ["RTB_BUILD_FUSED"]))
for kv in os.environ.get("RTB_OPTIONS", "").split(","):
    if "=" in kv:
        scene.set_option(kv.split("=")[0], int(kv.split("=")[1]))
scene.add_material(api.Material())
mesh = api.MeshPrimitive(tris[:0])
mesh.device_ptr, mesh.num_triangles = t.data_ptr(), len(tris)
scene.add_mesh(mesh)
scene.add_mesh_instance(api.MeshInstance(0, 0))
scene.upload_to_device()
for i in range(rep):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    scene.rebuild_mesh(0, device_ptr=t.data_ptr(), n=len(tris))
    torch.cuda.synchronize()
    info = scene.mesh_info(0)
    print("rebuild %d: wall %.2f ms, info build_ms %.2f, nodes %d leaves %d depth %d" % (i, 1e3 * (time.perf_counter() - t0), info["build_ms"], info["num_nodes"], info["num_leaves"], info["depth"]), flush=True)
scene.close()
