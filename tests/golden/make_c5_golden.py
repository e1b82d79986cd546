"""Generates tests/golden/c5_primary.npz from the REFERENCE ITSELF (oracle/_ref/libref_host.so: the reference's own
BVHTree build + its own cast_ray, compiled as host C++ by oracle/build_ref.sh) for bench.py's C5 mesh -- 10,001,404
triangles, 4K, the first camera pose of the bench loop.  The reference's host BVH build of this mesh takes minutes and
/root/reference does not exist on the GPU box, so the frame is computed here once and a (8, 4)-strided sample of its
first-hit AOVs and colours is committed (tests/test_parity_gpu.py::test_c5_full_size_frame_against_reference).

    python tests/golden/make_c5_golden.py
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc  # noqa: E402
import bench  # noqa: E402

SY, SX = 8, 4


def main():
    cfg = bench.CONFIGS["c5"]
    W, H = cfg["w"], cfg["h"]
    K, D = bench.config_K(cfg)
    tris = bench.build_tris(cfg)
    pose = bench.camera_pose(cfg, 3)
    mats, inst = bench.scene_materials(cfg), bench.scene_instances(cfg)
    R = orc.RefHost()
    t0 = time.time()
    mesh = R.mesh_create(tris)
    print("reference BVHTree build of %d triangles: %.1f s" % (len(tris), time.time() - t0), flush=True)
    sc = R.scene([mesh], mats, inst)
    t0 = time.time()
    out = R.render(sc, W, H, K, D, pose)
    print("reference host render %dx%d: %.1f s, hit fraction %.3f" % (W, H, time.time() - t0, (out["prim"] >= 0).mean()), flush=True)
    np.savez_compressed(os.path.join(HERE, "c5_primary.npz"), W=W, H=H, K=K, D=D, pose=np.array(pose, np.float32), stride=np.array([SY, SX]),
                        n_tris=len(tris), prim=out["prim"][::SY, ::SX], inst=out["inst"][::SY, ::SX], t=out["t"][::SY, ::SX],
                        img=out["img"][::SY, ::SX])


if __name__ == "__main__":
    main()
