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
    print("rebuild %d: wall %.2f ms, info build_ms %.2f, nodes %d" % (i, 1e3 * (time.perf_counter() - t0), scene.mesh_info(0)["build_ms"], scene.mesh_info(0)["num_nodes"]), flush=True)
scene.close()
