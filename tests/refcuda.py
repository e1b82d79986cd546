"""ctypes driver for oracle/_ref/libref_cuda.so: the reference's own CUDA kernel compiled for sm_100a
(TEST INFRASTRUCTURE / reported baseline only; exists where oracle/build_ref.sh ran; needs a GPU)."""
import ctypes as C
import os

import numpy as np

import orc

PATH = os.path.join(orc.ORACLE_DIR, "_ref", "libref_cuda.so")
PATH_ASIS = os.path.join(orc.ORACLE_DIR, "_ref", "libref_cuda_asis.so")


def available():
    return os.path.exists(PATH)


class RefCuda(orc.RefHost):
    """Same scene-building entry points as RefHost; render runs the reference kernel on the current device."""

    def __init__(self, path=PATH):
        self.PATH = path
        super().__init__()
        self.lib.ref_render.restype = C.c_float
        self.lib.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]

    def render(self, scene, width, height, K, D, pose, warmup=1, frames=3, aovs=True):
        import torch
        K = np.ascontiguousarray(K, np.float32).reshape(9)
        D = np.ascontiguousarray(D, np.float32)
        pose = np.ascontiguousarray(pose, np.float32)
        pitch = (width * 3 + 511) // 512 * 512
        img = torch.zeros((height, pitch), dtype=torch.uint8, device="cuda")
        prim = torch.zeros((height, width), dtype=torch.int32, device="cuda") if aovs else None
        inst = torch.zeros((height, width), dtype=torch.int32, device="cuda") if aovs else None
        t = torch.zeros((height, width), dtype=torch.float32, device="cuda") if aovs else None
        torch.cuda.synchronize()
        ms = self.lib.ref_render(scene, width, height, K.ctypes.data, D.ctypes.data, pose.ctypes.data, img.data_ptr(), pitch,
                                 prim.data_ptr() if aovs else None, inst.data_ptr() if aovs else None,
                                 t.data_ptr() if aovs else None, warmup, frames)
        out = dict(ms=ms, img=img[:, : width * 3].reshape(height, width, 3).cpu().numpy())
        if aovs:
            out.update(prim=prim.cpu().numpy(), inst=inst.cpu().numpy(), t=t.cpu().numpy())
        return out


def time_reference_kernel(tris, materials, instances, W, H, K, D, poses, as_shipped=False):
    """Reference CUDA kernel (P2 variant) on the same scene and the same camera poses as the bench's timed loop: ms/frame and
    Mrays/s (primary rays only -- the reference traces nothing else).  One untimed frame first (module load), then every pose is
    timed once by the harness's own CUDA events, L2 flushed before it like in the timed loop."""
    import time
    import torch
    R = RefCuda(PATH_ASIS if as_shipped else PATH)  # as shipped: Q_rsqrt reads a `long` (8 bytes on LP64) -- undefined behaviour, kept for the record
    t0 = time.perf_counter()
    mesh = R.mesh_create(tris)                      # the reference's host BVHTree build (single-threaded)
    t_build = time.perf_counter() - t0
    t0 = time.perf_counter()
    sc = R.scene([mesh], materials, instances)      # one cudaMalloc per leaf (BVHTree.hpp:97-114)
    t_upload = time.perf_counter() - t0
    if isinstance(poses, tuple) and not isinstance(poses[0], (tuple, list)):
        poses = [poses]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ms = []
    R.render(sc, W, H, K, D, poses[0], warmup=0, frames=1, aovs=False)
    for i, pose in enumerate(poses):
        flush.fill_(i & 0xFF)
        out = R.render(sc, W, H, K, D, pose, warmup=0, frames=1, aovs=False)
        if out["ms"] <= 0:
            raise RuntimeError("reference kernel failed")
        ms.append(out["ms"])
    m = float(np.mean(ms))
    return dict(kind="reference CUDA kernel (raycast.cu, sm_100a, %s), primary rays only" % ("exactly as shipped" if as_shipped else "32-bit Q_rsqrt"), ms_per_frame=m,
                value=W * H / (m * 1e-3) / 1e6, unit="Mrays/s", host_bvh_build_s=t_build, upload_s=t_upload, frames=len(ms),
                poses="the first %d timed poses of the bench loop, L2 flushed before each" % len(ms))
