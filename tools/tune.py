#!/usr/bin/env python
"""Knob sweep on a GPU box: frame ms of a bench config for each setting of the launch-bounds / refill knobs.
   python tools/tune.py c3 trace_minb=4,5,6,8 refill_idle=4,8,16     (cartesian product; CUDA events, L2 flushed)"""
import itertools, os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from cuda_raytracing_b200 import api

def main():
    if os.environ.get("RTB_VARIANT"):                       # A/B a differently compiled library (tools/variants/*.so)
        api.LIB_PATH = os.path.join(ROOT, "tools", "variants", os.environ["RTB_VARIANT"])
    name = sys.argv[1]
    knobs = [(a.split("=")[0], [int(v) for v in a.split("=")[1].split(",")]) for a in sys.argv[2:]]
    cfg = bench.CONFIGS[name]
    W, H = cfg["w"], cfg["h"]
    K, D = bench.config_K(cfg)
    tris = bench.build_tris(cfg)
    scene = api.Scene(0)
    for kv in os.environ.get("PRE", "").split():           # options that must be set before the upload, e.g. PRE="leaf_size=8"
        scene.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    for m in bench.scene_materials(cfg):
        scene.add_material(api.Material(albedo=m["albedo"], metallic=m["metallic"]))
    scene.add_mesh(api.MeshPrimitive(tris))
    for (mi, ma, pose, scale) in bench.scene_instances(cfg):
        scene.add_mesh_instance(api.MeshInstance(mi, ma, pose, scale))
    scene.upload_to_device()
    cam = api.Camera(W, H, K, D)
    cam.mode, cam.max_depth, cam.spp = cfg["mode"], cfg["depth"], min(cfg["spp"], 2)
    frame = torch.zeros((H, W * 3), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ref = None
    for combo in itertools.product(*[v for _, v in knobs]):
        for (k, _), v in zip(knobs, combo):
            if k == "tile_count":                           # emulate one rank of an N-GPU run: render 1/N of the bands
                cam.tile_rank, cam.tile_count = 0, v
            else:
                scene.set_option(k, v)
        ms = []
        for i in range(7):
            cam.pose = np.array(bench.camera_pose(cfg, 3 if os.environ.get("SAMEPOSE") else i), np.float32)
            if not os.environ.get("NOFLUSH"): flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            cam.render_scene(scene, frame.data_ptr(), W * 3)
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                ms.append(e0.elapsed_time(e1))
        h = int(frame.to(torch.int64).sum().item())
        if ref is None:
            ref = h
        st = scene.frame_stats()
        rays = st["rays"]
        extra = {}
        if os.environ.get("COUNT"):                          # COUNT=1: one more frame with the device counters on (slower kernel variant)
            scene.set_option("counters", 1)
            cam.render_scene(scene, frame.data_ptr(), W * 3)
            torch.cuda.synchronize()
            c = scene.frame_stats()
            extra = dict(nodes_per_ray=round(c["nodes"] / max(1, c["rays"]), 2), tris_per_ray=round(c["tris"] / max(1, c["rays"]), 2))
            scene.set_option("counters", 0)
        print(json.dumps(dict(config=name, knobs=dict(zip([k for k, _ in knobs], combo)), ms=round(float(np.mean(ms)), 3), ms_min=round(min(ms), 3),
                              mrays=round(rays / np.mean(ms) / 1e3, 1), same_image=(h == ref), hash=h, **extra)), flush=True)
    scene.close()

if __name__ == "__main__":
    main()
