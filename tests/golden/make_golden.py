"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libref_host.so = the reference's own
sources compiled as host C++ by oracle/build_ref.sh).  Run in the build container (needs /root/reference):

    python tests/golden/make_golden.py

The fixtures pin oracle/rt_oracle.c on machines where /root/reference does not exist (the GPU box).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc  # noqa: E402
from cuda_raytracing_b200 import synth  # noqa: E402


def main():
    R = orc.RefHost()
    # known-answer values
    xs = np.array([1, 2, 0.25, 3, 100, 1e-3, 7.5, 1e6, 1e-6], np.float32)
    kat = dict(qrsqrt_x=xs, qrsqrt_bits=np.array([R.q_rsqrt(x).view(np.uint32) for x in xs], np.uint32),
               sizes=np.array([R.sizeof(i) for i in range(8)], np.int32),
               normalize_123=R.normalize([1, 2, 3]),
               euler_yaw=R.apply_euler([0.5, 0, 0], [0, 1, 0]), euler_pitch=R.apply_euler([0, 0.5, 0], [0, 1, 0]),
               euler_roll=R.apply_euler([0, 0, 0.5], [1, 0, 0]))
    rng = np.random.default_rng(7)
    poses = rng.uniform(-1, 1, (64, 6)).astype(np.float32) * np.array([15, 15, 15, 3.1, 1.0, 3.1], np.float32)
    kat["poses"] = poses
    kat["inv_poses"] = np.stack([R.invert_lre(p) for p in poses])
    np.savez_compressed(os.path.join(HERE, "kat.npz"), **kat)

    # BVH known answers: G(132,264) (C1 mesh) node/leaf statistics + a small tree in full
    v, f = synth.bumpy_sphere_mesh(132, 264)
    mh = R.mesh_create(synth.make_triangles(v, f))
    mm, ch, cnt, idx = R.mesh_nodes(mh)
    leaf = ch[:, 0] == -1
    small = synth.bumpy_sphere(12, 24, ground=True)
    ms = R.mesh_create(small)
    smm, sch, scnt, sidx = R.mesh_nodes(ms)
    np.savez_compressed(os.path.join(HERE, "bvh.npz"), c1_nodes=len(mm), c1_leaves=int(leaf.sum()), c1_max=int(cnt[leaf].max()),
                        c1_min=int(cnt[leaf].min()), c1_child_crc=np.uint64(np.bitwise_xor.reduce((ch.astype(np.int64).ravel() + 7) * np.arange(1, 2 * len(ch) + 1))),
                        small_minmax=smm, small_child=sch, small_count=scnt, small_indices=sidx)

    # rendered fixtures: two instances (one rotated/translated/scaled), one textured material, fisheye + pinhole
    tex = (rng.integers(1, 256, (16, 16, 3))).astype(np.uint8)
    mats = [dict(albedo=(0.1, 0.2, 0.9)), dict(albedo=(1, 1, 1), texture=tex)]
    soup = synth.random_soup(400, seed=3)
    msoup = R.mesh_create(soup)
    insts = [(0, 0, (0, 0, 0, 0, 0, 0), (1, 1, 1)), (1, 1, (1.5, 0.5, 0.2, 0.7, -0.3, 0.2), (1, 1, 1)),
             (1, 0, (-1.8, 0.4, 0.1, -1.1, 0.4, -0.6), (1.25, 1.25, 1.25))]
    sc = R.scene([ms, msoup], mats, insts)
    W, H = 96, 72
    K = np.array([[60, 0, 48.37], [0, 60, 36.21], [0, 0, 1]], np.float32)
    out = {}
    cams = [((0.1, -3.5, 1.2, 0.05, -0.1, 0.02), synth.ZERO_D), ((1.0, -3.0, 2.0, 0.3, -0.35, -0.1), synth.REF_D)]
    for k, (pose, D) in enumerate(cams):
        r = R.render(sc, W, H, K, D, pose, nthreads=2)
        out["img%d" % k], out["prim%d" % k], out["inst%d" % k], out["t%d" % k] = r["img"], r["prim"], r["inst"], r["t"]
        out["pose%d" % k], out["D%d" % k] = np.array(pose, np.float32), np.array(D, np.float32)
    np.savez_compressed(os.path.join(HERE, "render_small.npz"), K=K, tex=tex, soup=soup, small=small, W=W, H=H, **out)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
