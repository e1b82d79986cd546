#!/usr/bin/env python
"""bench.py -- Mrays/s of the B200 ray-casting core on the BASELINE.json configs (default: C3, the 4K
instanced scene the metric is quoted on: 100 MeshInstances of a 1M-triangle BLAS, reflections to depth 4).

    python bench.py --gpus N --steps K --warmup W [--config c1|c2|c3|c4|c5] [--impl reference]

One process per GPU (torchrun for N>1).  A step = one frame of the hot path (camera rays -> TLAS/BLAS
traversal -> shading -> framebuffer; multi-GPU: + NCCL gather of the image bands to rank 0).  The camera
orbits so every step is a different frame; L2 is flushed (untimed) between steps.  Prints ONE JSON line.
torch is used for device buffers, streams, events and torch.distributed only.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from cuda_raytracing_b200 import synth  # noqa: E402

CONFIGS = {
    # name: (width, height, mesh (N, M), ground, instances, mode, max_depth, spp, K scale, description)
    "c1": dict(w=640, h=480, mesh=(132, 264), ground=False, grid=None, mode=0, depth=0, spp=1, K="c1",
               desc="C1 69,696-tri OBJ mesh, 640x480, primary closest hit"),
    "c2": dict(w=1920, h=1080, mesh=(500, 1000), ground=True, grid=None, mode=1, depth=0, spp=1, K="ref",
               desc="C2 1M-tri mesh + ground, 1080p, primary + shadow rays"),
    "c3": dict(w=3840, h=2160, mesh=(500, 1000), ground=False, grid=(10, 10), mode=2, depth=4, spp=1, K="ref2",
               desc="C3 4K, 100 instances of a 1M-tri BLAS, mirror reflections to depth 4"),
    "c4": dict(w=7680, h=4320, mesh=(500, 1000), ground=False, grid=(10, 10), mode=3, depth=8, spp=64, K="ref4",
               desc="C4 8K diffuse path tracing, 64 spp, 8 bounces, 100 instances of a 1M-tri BLAS"),
    "c5": dict(w=3840, h=2160, mesh=(1582, 3161), ground=False, grid=None, mode=1, depth=0, spp=1, K="ref2",
               desc="C5 per-frame LBVH rebuild of an animated 10M-tri mesh + 4K primary+shadow render"),
}


def config_K(cfg):
    if cfg["K"] == "c1":
        return synth.C1_K.copy(), synth.ZERO_D.copy()
    s = {"ref": 1.0, "ref2": 2.0, "ref4": 4.0}[cfg["K"]]
    K = synth.REF_K.copy()
    K[:2, :] *= np.float32(s)
    return K, synth.REF_D.copy()


def camera_pose(cfg, step):
    """Orbiting camera that keeps the geometry in frame (benchmark cameras must frame the mesh)."""
    a = 0.05 * step
    if cfg["grid"] is None:
        r = 2.6
        x, y, z = r * math.sin(a), -r * math.cos(a), 1.6
        return (x, y, z, -a, -0.22, 0.0)
    r = 19.0
    x, y, z = r * math.sin(a), -r * math.cos(a), 9.0
    return (x, y, z, -a, -0.42, 0.0)


def scene_materials(cfg):
    if cfg["mode"] == 2:
        return [dict(albedo=(0.9, 0.6, 0.3), metallic=0.6), dict(albedo=(0.3, 0.8, 0.4), metallic=0.35)]
    return [dict(albedo=(0.1, 0.2, 0.9), metallic=0.0), dict(albedo=(0.7, 0.7, 0.7), metallic=0.0)]


def scene_instances(cfg):
    if cfg["grid"] is None:
        return [(0, 0, (0.0, 0.0, 0.0, 0.0, 0.0, 0.0), (1.0, 1.0, 1.0))]
    poses = synth.instance_grid(*cfg["grid"])
    return [(0, i % 2, p, (1.0, 1.0, 1.0)) for i, p in enumerate(poses)]


def config_dict(cfg, name, n_tris):
    """The workload description both arms print verbatim (the driver compares the two `config` objects)."""
    return dict(workload=cfg["desc"], config=name, width=cfg["w"], height=cfg["h"], triangles=int(n_tris),
                instances=len(scene_instances(cfg)), mode=cfg["mode"], max_depth=cfg["depth"], spp=cfg["spp"])


def flops_per_ray(nodes, tris):
    """SURVEY 8(d)'s arithmetic model: per wide-node visit 4 children x 24 flops (6 fma = 12, 6 min/max, compares and the key),
    per triangle test 30, 80 per ray for setup and the accept rule."""
    return nodes * 4 * 24 + tris * 30 + 80


class ClockSampler(threading.Thread):
    """SM clock, max clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe), through NVML
    every ~2 ms so that even a sub-millisecond-per-step config is covered; nvidia-smi -lms is the fallback."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc, self.stop_flag = index, [], None, False
        self.sm, self.mx, self.reasons, self.how = [], [], set(), "nvml"
        self.ready = threading.Event()                      # set once the first sample is in: the timed region starts after it

    def run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            while not self.stop_flag:
                self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                self.mx.append(mx)
                try:
                    bits = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    bits = int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                for name, bit in self.REASONS:
                    if bits & bit:
                        self.reasons.add(name)
                self.ready.set()
                time.sleep(0.002)
            return
        except Exception:
            self.how = "nvidia-smi"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                r = [c.strip() for c in line.split(",")]
                if len(r) >= 9 and r[1].replace(".", "").isdigit():
                    self.sm.append(float(r[1]))
                    self.mx.append(float(r[2]))
                    self.ready.set()
                    for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                        if v.lower().startswith("active"):
                            self.reasons.add(name)
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        return dict(sm_mhz=float(np.median(self.sm)) if self.sm else None, sm_max_mhz=max(self.mx) if self.mx else None,
                    reasons=sorted(self.reasons), samples=len(self.sm), how=self.how)


def load_peaks():
    """(HBM GB/s, SM max MHz, kind): the driver-written measured peaks, else the profiling recipe's fallback."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return j.get("hbm_gbs", 6650.0), j.get("sm_max_mhz", 1965.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


def build_tris(cfg, phase=0.0):
    return synth.bumpy_sphere(cfg["mesh"][0], cfg["mesh"][1], phase=phase, ground=cfg["ground"])


# ------------------------------------------------------------------------------------------- reference arm
def run_reference(args, cfg, rank):
    """The reference's own CPU implementation of the path on the host cores (oracle/_ref when present, else the
    oracle port), bounded sample per step: a centred band of rows of the same frame."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    K, D = config_K(cfg)
    cores = os.cpu_count() or 1
    tris = build_tris(cfg)
    mats, insts = scene_materials(cfg), scene_instances(cfg)
    use_ref = orc.RefHost.available() and cfg["mode"] == 0 and len(tris) <= 200000
    O = orc.Oracle()
    if use_ref:
        R = orc.RefHost()
        mh = R.mesh_create(tris)
        sc = R.scene([mh], mats, insts)
        kind = "reference"
    else:
        om = O.mesh_create(tris)
        sc = orc.OracleScene(O, [om], mats, [O.build_instance(m, a, p, s) for (m, a, p, s) in insts])
        kind = "port"
    W, H = cfg["w"], cfg["h"]
    spp = min(cfg["spp"], 2)

    def one(step, rows):
        y0 = max(0, H // 2 - rows // 2)
        pose = camera_pose(cfg, step)
        t0 = time.perf_counter()
        if use_ref:
            R.render(sc, W, H, K, D, pose, nthreads=cores, y0=y0, y1=y0 + rows)
            rays = rows * W
        else:
            out = sc.render(W, H, K, D, pose, mode=cfg["mode"], max_depth=cfg["depth"], spp=spp, y0=y0, y1=y0 + rows,
                            nthreads=cores, aovs=False, stats=True)
            rays = out["stats"]["rays"]
        return rays, time.perf_counter() - t0

    rows = 8
    rays, dt = one(0, rows)                                  # calibrate: aim at ~3 s of CPU work per step
    rows = int(max(8, min(H, rows * 3.0 / max(dt, 1e-3))))
    for s in range(args.warmup):
        one(s, rows)
    tot_r, tot_t = 0, 0.0
    for s in range(args.steps):
        r, dt = one(args.warmup + s, rows)
        tot_r += r
        tot_t += dt
    val = tot_r / tot_t / 1e6
    sample = "%d centred rows of %dx%d per step (%s), %d spp" % (rows, W, H, cfg["desc"], spp)
    line = dict(impl="reference", metric="Mrays/s", value=val, unit="Mrays/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * tot_t / max(1, args.steps), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=config_dict(cfg, args.config, len(tris)),
                cpu_baseline=dict(value=val, unit="Mrays/s", cores=cores, kind=kind, sample=sample),
                e2e=dict(value=val, unit="Mrays/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------- our arm
def main():
    # stdout carries exactly ONE line, the JSON result: libraries that print banners on fd 1 (NCCL's version line) go to stderr
    out_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(out_fd, "w")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", default="c3")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--ref-cuda-as-shipped", action="store_true", help="also time the reference kernel exactly as shipped (UB Q_rsqrt), +25 s")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--completion", default="nccl", choices=["peer", "nccl"],
                    help="N>1, peer exchange: how rank 0 learns that every band has landed (flag words in peer memory, or a 4-byte all-reduce)")
    ap.add_argument("--exchange", default="direct", choices=["direct", "push", "nccl"],
                    help="N>1: how the bands reach rank 0 (tiles.PeerFrame direct stores / peer copy, or the NCCL gather)")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return
    args.warmup = max(args.warmup, 3)

    import ctypes as C
    import torch
    import torch.distributed as dist
    from cuda_raytracing_b200 import api

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, H = cfg["w"], cfg["h"]
    K, D = config_K(cfg)

    # ---- scene (replicated on every rank) -------------------------------------------------------
    tris = build_tris(cfg)
    scene = api.Scene(local_rank)
    for kv in os.environ.get("RTB_OPTIONS", "").split(","):                      # experiments only: "key_bits=32,build_fused=0"
        if "=" in kv:
            scene.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    for m in scene_materials(cfg):
        scene.add_material(api.Material(albedo=m["albedo"], metallic=m["metallic"]))
    tris_t = torch.from_numpy(tris.view(np.uint8).reshape(-1, 72)).to(dev)      # triangles resident in HBM
    mesh = api.MeshPrimitive(tris[:0])
    mesh.device_ptr, mesh.num_triangles = tris_t.data_ptr(), len(tris)
    scene.add_mesh(mesh)
    for (mi, ma, pose, scale) in scene_instances(cfg):
        scene.add_mesh_instance(api.MeshInstance(mi, ma, pose, scale))
    scene.upload_to_device()
    info = scene.mesh_info(0)

    cam = api.Camera(W, H, K, D)
    cam.mode, cam.max_depth, cam.spp = cfg["mode"], cfg["depth"], cfg["spp"]
    cam.tile_rank, cam.tile_count = rank, world
    pitch = W * 3
    from cuda_raytracing_b200 import tiles
    if world > 1 and args.exchange != "nccl":
        # frame exchange over NVLink peer memory: ranks > 0 write their bands into rank 0's frame (tiles.PeerFrame)
        fg = tiles.PeerFrame(scene, H, pitch, rank, world, dev, mode=args.exchange, signal=args.completion)
        frame, target = fg.frame, fg.target
    else:
        frame = torch.zeros((H, pitch), dtype=torch.uint8, device=dev)
        fg, target = tiles.FrameGather(H, pitch, rank, world, dev), frame.data_ptr()
    stream = torch.cuda.current_stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    host_frame = torch.empty((H, pitch), dtype=torch.uint8).pin_memory()

    def step(i, e2e=False):
        """one frame: (C5: rebuild) -> render own bands -> gather to rank 0 (-> D2H when e2e)."""
        if args.config == "c5":
            # animated mesh: new vertex positions arrive in HBM and the BLAS of the NEXT frame is built on the library's build stream
            # while this frame traverses the one built during the previous step (rtb_mesh_rebuild_async / rtb_mesh_commit:
            # double-buffered BLAS).  Every step = one full rebuild of the 10 M-triangle mesh + one frame.
            scene.rebuild_mesh_async(0, anim[(i + 1) % len(anim)].data_ptr(), len(tris))
        cam.pose = np.array(camera_pose(cfg, i), np.float32)
        cam.render_scene(scene, target, pitch, stream=C.c_void_p(stream.cuda_stream))
        if args.config == "c5":
            scene.commit_mesh(0)                            # the host waits for the build only; the frame may still be running
        if isinstance(fg, tiles.PeerFrame):
            fg.gather(stream.cuda_stream)
        else:
            fg.gather(frame)
        if e2e and rank == 0:
            host_frame.copy_(frame, non_blocking=True)
            stream.synchronize()

    anim = None
    if args.config == "c5":
        anim = []
        for f in range(2):
            t = build_tris(cfg, phase=2 * math.pi * f / 60.0)
            anim.append(torch.from_numpy(t.view(np.uint8).reshape(-1, 72)).to(dev))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    align = torch.zeros(1, dtype=torch.int32, device=dev)

    def timed(e2e):
        for i in range(args.warmup):
            step(i, e2e)
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        sampler.ready.wait(timeout=3.0)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        rays = 0
        launches = 0
        t_wall0 = time.perf_counter()
        for k in range(args.steps):
            if not args.no_flush:
                flush.fill_(k & 0xFF)                       # L2 flush, outside the timed events
            barrier()
            if world > 1:
                dist.all_reduce(align)                       # untimed: lines the ranks' STREAMS up, so that ev0 is the same instant on every
            ev[k][0].record(stream)                          # device (the host-side barrier above lets them start ~0.1 ms apart)
            step(args.warmup + k, e2e)
            ev[k][1].record(stream)
            barrier()
            st = scene.frame_stats()
            rays += st["rays"]
            launches += st["launches"]
        t_wall = time.perf_counter() - t_wall0
        clocks = sampler.stop()
        ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([ms, float(rays)], dtype=torch.float64, device=dev)
        if world > 1:
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tsum = t.clone()
            dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            ms, rays = float(tmax[0]), float(tsum[1])
        return ms, rays, launches, clocks, t_wall

    ms, rays, launches, clocks, _ = timed(False)
    ms_e, rays_e, _, _, _ = timed(True)
    value = rays / (ms * 1e-3) / 1e6
    e2e_value = rays_e / (ms_e * 1e-3) / 1e6

    # ---- roofline of the dominant kernel: the render launch of THIS rank's share, on the SAME frames as the timed loop ----
    # (a) kernel time: CUDA events on the launching stream around rtb_render alone (no exchange, no copy), L2 flushed before
    #     each frame like the timed loop; (b) work: the device counters of the same frames (a slower counting variant of the
    #     kernel, never timed).  Every rank measures its own share; rank 0 reports its own and the min / median / max over ranks.
    peak, sm_mhz, peak_kind = load_peaks()
    roof = None
    cpu = None
    ref_cuda = None
    own_target = target if not (world > 1 and args.exchange == "nccl") else frame.data_ptr()
    kt, xt, variants = [], [], set()
    for k in range(args.steps):
        cam.pose = np.array(camera_pose(cfg, args.warmup + k), np.float32)
        if not args.no_flush:
            flush.fill_(k & 0xFF)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        cam.render_scene(scene, own_target, pitch, stream=C.c_void_p(stream.cuda_stream))
        e1.record(stream)
        if world > 1:                                    # the exchange / completion signal alone, after this rank's own launch
            e2 = torch.cuda.Event(enable_timing=True)
            if isinstance(fg, tiles.PeerFrame):
                fg.gather(stream.cuda_stream)
            else:
                fg.gather(frame)
            e2.record(stream)
        torch.cuda.synchronize()
        kt.append(e0.elapsed_time(e1))
        if world > 1:
            xt.append(e1.elapsed_time(e2))
        variants.add(scene.frame_stats()["kernel_variant"])
    kms = float(np.mean(kt))
    n_count = min(args.steps, 2 if cfg["mode"] == 3 else args.steps)       # path tracing: the counting variant costs seconds per frame
    scene.set_option("counters", 1)
    cnt = dict(rays=0, nodes=0, tris=0)
    for k in range(n_count):
        cam.pose = np.array(camera_pose(cfg, args.warmup + k), np.float32)
        cam.render_scene(scene, own_target, pitch, synchronize=True)
        st = scene.frame_stats()
        for key in cnt:
            cnt[key] += st[key]
    scene.set_option("counters", 0)
    barrier()
    kall = torch.tensor([kms], dtype=torch.float64, device=dev)
    if world > 1:
        gathered = [torch.zeros_like(kall) for _ in range(world)]
        dist.all_gather(gathered, kall)
        kall = torch.cat(gathered)
    kall = kall.cpu().numpy()
    if rank == 0:
        my_pixels = W * H / world
        nodes_per_ray, tris_per_ray = cnt["nodes"] / max(1, cnt["rays"]), cnt["tris"] / max(1, cnt["rays"])
        rays_f = cnt["rays"] / max(1, n_count)                               # rays per frame of this rank's share
        hops = max(0.0, rays_f - my_pixels * (cfg["spp"] if cfg["mode"] == 3 else 1)) if cfg["mode"] >= 2 else 0.0
        node_b, tri_b, hop_b = 112, 64, 64                                   # bytes per node visit / triangle test / queue hop (32 written + 32 read)
        alg_bytes = rays_f * (nodes_per_ray * node_b + tris_per_ray * tri_b) + my_pixels * 3 + hops * hop_b
        ach = alg_bytes / (kms * 1e-3) / 1e9
        flops = rays_f * flops_per_ray(nodes_per_ray, tris_per_ray)
        fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6 / 1e12                      # TFLOP/s: SMs x FP32 lanes x 2 (fma) x clock
        fp32_ach = flops / (kms * 1e-3) / 1e12
        variant = sorted(variants)[-1] if variants else -1
        kname = {2: "k_megaq (persistent path kernel with in-kernel ray queues)", 1: "k_mega (persistent path kernel, lane-owned chains)",
                 0: "k_trace_queue (staged wavefront)", -1: "k_render_direct"}.get(variant, "?")
        prof = {}
        tp = os.path.join(ROOT, "profiles", "traffic.json")                  # per-launch numbers of the committed ncu --set full capture
        if os.path.exists(tp):
            prof = json.load(open(tp)).get(args.config) or {}
        if not isinstance(prof, dict):
            prof = dict(dram_bytes=prof)
        traffic = prof.get("dram_bytes") if world == 1 else None             # the capture is of the full frame on one GPU
        roof = dict(bound=prof.get("bound", "issue"), achieved=ach, peak=peak, unit="GB/s", frac=ach / peak, traffic=traffic, peak_kind=peak_kind,
                    kernel=kname, kernel_variant=variant, kernel_ms=kms, kernel_ms_ranks=dict(min=float(kall.min()), median=float(np.median(kall)), max=float(kall.max())),
                    exchange_ms=(float(np.mean(xt)) if xt else 0.0),     # rank 0: its wait for the other ranks' bands after its own launch (includes their skew)
                    share_of_step=kms / (ms / args.steps), frames="the %d timed poses, L2 flushed, this rank's share (1/%d of the bands)" % (args.steps, world),
                    ceilings=dict(hbm=dict(achieved=ach, peak=peak, unit="GB/s", frac=ach / peak, alg_bytes=alg_bytes,
                                           dram_frac=(traffic / (kms * 1e-3) / 1e9 / peak) if traffic else None),
                                  fp32=dict(achieved=fp32_ach, peak=fp32_peak, unit="TFLOP/s", frac=fp32_ach / fp32_peak, flops=flops),
                                  issue=dict(frac=prof.get("issue_active"), source=prof.get("source"))),
                    nodes_per_ray=nodes_per_ray, tris_per_ray=tris_per_ray, rays=rays_f,
                    note="algorithmic bytes = %d B per node visit + %d B per triangle test + 3 B per pixel + %d B per queue hop; node and triangle "
                         "fetches are mostly L1/L2 hits (the scene is L2-sized), so neither the HBM nor the FP32 ceiling binds: the kernel is bound "
                         "by issue slots and the latency of dependent node fetches (ncu: profiles/, DESIGN.md 4)" % (node_b, tri_b, hop_b))

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import orc
        O = orc.Oracle()
        cores = os.cpu_count() or 1
        om = O.mesh_create(tris)
        osc = orc.OracleScene(O, [om], scene_materials(cfg), [O.build_instance(m, a, p, s) for (m, a, p, s) in scene_instances(cfg)])
        rows, y0 = 8, H // 2
        pose = camera_pose(cfg, args.warmup)
        spp = min(cfg["spp"], 2)
        t0 = time.perf_counter()
        out = osc.render(W, H, K, D, pose, mode=cfg["mode"], max_depth=cfg["depth"], spp=spp, y0=y0, y1=y0 + rows, nthreads=cores, aovs=False, stats=True)
        dt = time.perf_counter() - t0
        rows = int(max(8, min(H, rows * 12.0 / max(dt, 1e-3))))
        y0 = max(0, H // 2 - rows // 2)
        t0 = time.perf_counter()
        out = osc.render(W, H, K, D, pose, mode=cfg["mode"], max_depth=cfg["depth"], spp=spp, y0=y0, y1=y0 + rows, nthreads=cores, aovs=False, stats=True)
        dt = time.perf_counter() - t0
        cpu = dict(value=out["stats"]["rays"] / dt / 1e6, unit="Mrays/s", cores=cores, kind="port",
                   sample="%d centred rows of the %dx%d frame (%.1f s), oracle/rt_oracle.c on %d threads" % (rows, W, H, dt, cores))
        if not args.no_ref_cuda and orc.RefHost.available() and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_cuda.so")):
            try:
                sys.path.insert(0, os.path.join(ROOT, "tests"))
                import refcuda
                poses = [camera_pose(cfg, args.warmup + k) for k in range(min(args.steps, 4))]
                ref_cuda = refcuda.time_reference_kernel(tris, scene_materials(cfg), scene_instances(cfg), W, H, K, D, poses)
                if args.ref_cuda_as_shipped and os.path.exists(refcuda.PATH_ASIS):
                    ref_cuda["as_shipped"] = refcuda.time_reference_kernel(tris, scene_materials(cfg), scene_instances(cfg), W, H, K, D, poses, as_shipped=True)
            except Exception as ex:                      # the reported baseline is optional, the bench line is not
                ref_cuda = dict(error=str(ex)[:200])

    if args.config == "c5":
        info = scene.mesh_info(0)                            # build_ms of the last asynchronous rebuild
    if rank == 0:
        line = dict(metric="Mrays/s", value=value, unit="Mrays/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=ms / args.steps, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                    config=config_dict(cfg, args.config, len(tris)),
                    details=dict(parallelism="image bands of %d rows interleaved over %d GPU(s), scene replicated, bands to rank 0: %s" % (
                                     api.BAND_ROWS, world, {"direct": "kernel stores into rank 0's frame over NVLink peer memory",
                                                            "push": "one strided peer copy per rank over NVLink", "nccl": "NCCL gather"}[args.exchange] if world > 1 else "-"),
                                 signal=(args.completion if world > 1 and args.exchange != "nccl" else "-"),
                                 l2="flushed between steps with a 256 MiB write (untimed)" if not args.no_flush else "not flushed",
                                 rays_per_frame=rays / args.steps, blas=dict(nodes=info["num_nodes"], leaves=info["num_leaves"], depth=info["depth"],
                                                                          build_ms=info["build_ms"], bytes=info["device_bytes"])),
                    clocks=clocks, gpu_launches=launches,
                    e2e=dict(value=e2e_value, unit="Mrays/s", h2d_bytes_per_step=int(C.sizeof(api.RenderParams)), d2h_bytes_per_step=int(H * pitch),
                             ms_per_step=ms_e / args.steps),
                    roofline=roof, cpu_baseline=cpu)
        if ref_cuda is not None:
            line["ref_cuda"] = ref_cuda
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()                                   # nobody unmaps / frees the shared frame while a peer may still write
        if isinstance(fg, tiles.PeerFrame):
            fg.close()
        dist.destroy_process_group()
    scene.close()


if __name__ == "__main__":
    main()
