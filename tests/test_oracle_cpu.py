"""CPU suite (`-m "not gpu"`): pins the oracle against fixtures generated from the reference itself
(tests/golden/make_golden.py), against the live reference host build when it exists, and checks the host side
of the product (OBJ ingestion, pose math, C-ABI exports, loud failure without a GPU, band partition over gloo)."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import orc
from cuda_raytracing_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
G = lambda name: np.load(os.path.join(HERE, "golden", name))  # noqa: E731


# ------------------------------------------------------------------ known-answer tests
def test_kat_qrsqrt_and_conventions(oracle):
    k = G("kat.npz")
    for x, bits in zip(k["qrsqrt_x"], k["qrsqrt_bits"]):
        assert oracle.q_rsqrt(x).view(np.uint32) == bits
    # survey-time values, frozen independently of the fixture
    for x, bits in ((1, 0x3F7F910F), (2, 0x3F34F95E), (0.25, 0x3FFF910F), (3, 0x3F13AC3C), (100, 0x3DCC7B79), (1e-3, 0x41FCAE36)):
        assert int(oracle.q_rsqrt(x).view(np.uint32)) == bits
    assert oracle.normalize([1, 2, 3]).tobytes() == k["normalize_123"].tobytes()
    assert np.linalg.norm(oracle.normalize([1, 2, 3])) < 1.0          # Q_rsqrt under-estimates: not unit length
    assert oracle.apply_euler([0.5, 0, 0], [0, 1, 0]).tobytes() == k["euler_yaw"].tobytes()
    assert oracle.apply_euler([0, 0.5, 0], [0, 1, 0]).tobytes() == k["euler_pitch"].tobytes()
    assert oracle.apply_euler([0, 0, 0.5], [1, 0, 0]).tobytes() == k["euler_roll"].tobytes()
    np.testing.assert_allclose(k["euler_yaw"], [-0.4794, 0.8776, 0], atol=1e-4)
    assert list(k["sizes"]) == [72, 48, 104, 48, 24, 72, 160, 24]
    assert orc.TRI.itemsize == 72 and orc.INSTANCE.itemsize == 104 and orc.MATERIAL.itemsize == 48


def test_kat_invert_lre(oracle):
    k = G("kat.npz")
    for p, ip in zip(k["poses"], k["inv_poses"]):
        assert oracle.invert_lre(p).tobytes() == ip.tobytes()


def test_bvh_known_answers(oracle):
    b = G("bvh.npz")
    v, f = synth.bumpy_sphere_mesh(132, 264)
    m = oracle.mesh_create(synth.make_triangles(v, f))
    st = m.stats()
    assert (st["nodes"], st["leaves"], st["max_tris"], st["min_tris"], st["depth"]) == (126273, 63137, 19, 1, 28)
    assert (int(b["c1_nodes"]), int(b["c1_leaves"]), int(b["c1_max"]), int(b["c1_min"])) == (126273, 63137, 19, 1)
    n = m.nodes
    ch = np.stack([n["child_a"], n["child_b"]], 1)
    crc = np.uint64(np.bitwise_xor.reduce((ch.astype(np.int64).ravel() + 7) * np.arange(1, 2 * len(ch) + 1)))
    assert crc == b["c1_child_crc"]
    # the small tree in full, bit for bit
    ms = oracle.mesh_create(synth.bumpy_sphere(12, 24, ground=True))
    n = ms.nodes
    assert np.array_equal(np.concatenate([n["min"], n["max"]], 1), b["small_minmax"])
    assert np.array_equal(np.stack([n["child_a"], n["child_b"]], 1), b["small_child"])
    assert np.array_equal(n["tri_count"], b["small_count"])
    assert np.array_equal(ms.indices, b["small_indices"])


GOLD_INSTS = [(0, 0, (0, 0, 0, 0, 0, 0), (1, 1, 1)), (1, 1, (1.5, 0.5, 0.2, 0.7, -0.3, 0.2), (1, 1, 1)),
              (1, 0, (-1.8, 0.4, 0.1, -1.1, 0.4, -0.6), (1.25, 1.25, 1.25))]


def golden_scene(oracle):
    g = G("render_small.npz")
    mats = [dict(albedo=(0.1, 0.2, 0.9)), dict(albedo=(1, 1, 1), texture=g["tex"])]
    meshes = [oracle.mesh_create(g["small"]), oracle.mesh_create(g["soup"])]
    insts = [oracle.build_instance(m, a, p, s) for (m, a, p, s) in GOLD_INSTS]
    return g, orc.OracleScene(oracle, meshes, mats, insts)


def test_render_matches_reference_fixture(oracle):
    """colour / primitive id / instance id / distance images of the reference's host path, bit for bit"""
    g, sc = golden_scene(oracle)
    for k in range(2):
        r = sc.render(int(g["W"]), int(g["H"]), g["K"], g["D%d" % k], g["pose%d" % k])
        assert np.array_equal(r["img"], g["img%d" % k])
        assert np.array_equal(r["prim"], g["prim%d" % k])
        assert np.array_equal(r["inst"], g["inst%d" % k])
        assert r["t"].tobytes() == g["t%d" % k].tobytes()
        assert (r["prim"] >= 0).mean() > 0.2


# ------------------------------------------------------------------ live cross-checks against the reference host build
def test_live_reference_bvh_and_render(oracle, refhost):
    tris = synth.bumpy_sphere(40, 80, ground=True)
    mo, mh = oracle.mesh_create(tris), refhost.mesh_create(tris)
    mm, ch, cnt, idx = refhost.mesh_nodes(mh)
    n = mo.nodes
    assert np.array_equal(np.concatenate([n["min"], n["max"]], 1), mm)
    assert np.array_equal(np.stack([n["child_a"], n["child_b"]], 1), ch)
    assert np.array_equal(n["tri_count"], cnt) and np.array_equal(mo.indices, idx)
    mats = [dict(albedo=(0.3, 0.5, 0.7))]
    pose = (0.4, -2.6, 1.5, 0.15, -0.2, 0.05)
    sc = orc.OracleScene(oracle, [mo], mats, [oracle.build_instance(0, 0, (0.2, 0.1, 0, 0.4, 0.1, -0.2), (1, 1, 1))])
    rs = refhost.scene([mh], mats, [(0, 0, (0.2, 0.1, 0, 0.4, 0.1, -0.2), (1, 1, 1))])
    a = sc.render(160, 120, synth.C1_K * np.array([[0.25], [0.25], [1]], np.float32), synth.REF_D, pose)
    b = refhost.render(rs, 160, 120, synth.C1_K * np.array([[0.25], [0.25], [1]], np.float32), synth.REF_D, pose)
    for key in ("img", "prim", "inst"):
        assert np.array_equal(a[key], b[key]), key
    assert a["t"].tobytes() == b["t"].tobytes()


def test_live_reference_instance_record(oracle, refhost):
    rng = np.random.default_rng(1)
    for _ in range(50):
        pose = rng.uniform(-1, 1, 6).astype(np.float32) * np.array([10, 10, 10, 3.1, 1.2, 3.1], np.float32)
        scale = rng.uniform(0.5, 2.0, 3).astype(np.float32)
        assert oracle.build_instance(2, 1, pose, scale).tobytes() == refhost.build_instance(2, 1, pose, scale).tobytes()


# ------------------------------------------------------------------ OBJ ingestion (oracle, reference, product)
def _write_obj(tmp_path):
    v, f = synth.bumpy_sphere_mesh(10, 20)
    p = str(tmp_path / "m.obj")
    rng = np.random.default_rng(5)
    vts = rng.uniform(0, 1, (len(v), 2))
    with open(p, "w") as fh:
        fh.write("# comment\n\nmtllib ignored.mtl\no thing\n")
        for x in v:
            fh.write("v %.6f %.6f %.6f\n" % tuple(float(c) for c in x))
        for t in vts:
            fh.write("vt %.6f %.6f\n" % tuple(t))
        fh.write("vn 0 0 1\n")
        for k, (a, b, c) in enumerate(f):
            if k % 3 == 0:
                fh.write("f %d %d %d\n" % (a + 1, b + 1, c + 1))
            elif k % 3 == 1:
                fh.write("f %d/%d %d/%d %d/%d\n" % (a + 1, a + 1, b + 1, b + 1, c + 1, c + 1))
            else:
                fh.write("f %d/%d/1 %d/%d/1 %d/%d/1\n" % (a + 1, a + 1, b + 1, b + 1, c + 1, c + 1))
        fh.write("f 1 2 3 4 5\n")               # polygon: fan triangulation
    return p


def test_obj_loader_product_vs_oracle(oracle, rtb, tmp_path):
    p = _write_obj(tmp_path)
    a = oracle.obj_load(p)
    b = rtb.OBJLoader.load(p).triangles
    assert len(a) == len(b) == 2 * 10 * 20 + 3
    assert a["v"].tobytes() == b["v"].tobytes() and a["n"].tobytes() == b["n"].tobytes()
    # faces written as plain `f a b c` carry no uv in the reference (uninitialised there, zero here and in the oracle)
    assert a["uv"].tobytes() == b["uv"].tobytes()
    with pytest.raises(Exception):
        rtb.OBJLoader.load(str(tmp_path / "missing.obj"))


def test_obj_loader_vs_reference(oracle, refhost, tmp_path):
    p = _write_obj(tmp_path)
    a = oracle.obj_load(p)
    r = refhost.mesh_triangles(refhost.obj_load(p))
    assert a["v"].tobytes() == r["v"].tobytes() and a["n"].tobytes() == r["n"].tobytes()
    has_uv = np.array([(k % 3) != 0 for k in range(400)] + [False] * 3)
    assert a["uv"][has_uv].tobytes() == r["uv"][has_uv].tobytes()


def test_obj_loader_number_spellings_and_large_file(oracle, refhost, rtb, tmp_path):
    """The product loader parses numbers with its own exact fast path and cuts the file into one chunk per host thread:
    every spelling must give strtof's bits, and a file large enough to be cut (> 256 KB per chunk) the same triangles
    as the oracle's and the reference's sequential loaders."""
    rng = np.random.default_rng(5)
    spell = ["0", "-0", "+1", "1.", ".5", "-.25", "+.125e1", "1e-3", "1E3", "123456789", "0.000001", "3.4028234e38", "1e-22", "1e22",
             "16777217", "0.1", "0.2", "0.3", "1.17549435e-38", "7.0064923e-46", "9007199254740993", "1.00000005960464477539062500",
             "1.00000017881393432617187500", "0.5000000298023223876953125", "12345678901234567890.5", "1e23", "1e-23", "4.35e-8"]
    nums = list(spell)
    for _ in range(6000):
        k = rng.integers(0, 4)
        if k == 0: nums.append("%.6f" % rng.uniform(-100, 100))
        elif k == 1: nums.append("%.9g" % rng.uniform(-1e-3, 1e-3))
        elif k == 2: nums.append("%.17g" % rng.standard_normal())
        else: nums.append("%de%d" % (rng.integers(-99999, 99999), rng.integers(-25, 25)))
    while len(nums) % 3: nums.append("1")
    p = str(tmp_path / "spell.obj")
    nv = len(nums) // 3
    with open(p, "w") as fh:
        fh.write("# numbers\r\n\n   \n")
        for i in range(nv):
            fh.write("%sv %s\t%s %s%s" % ("  " if i % 7 == 0 else "", nums[3 * i], nums[3 * i + 1], nums[3 * i + 2], "\r\n" if i % 5 == 0 else "\n"))
        for i in range(0, nv - 2, 3):
            fh.write("f %d %d %d\n" % (i + 1, i + 2, i + 3))
        fh.write("f 1 2 3 4")                                    # no newline at the end of the file
    a = oracle.obj_load(p)
    b = rtb.OBJLoader.load(p).triangles
    assert len(a) == len(b) and a["v"].tobytes() == b["v"].tobytes() and a["n"].tobytes() == b["n"].tobytes()
    # large file: > 20 chunks worth of text, mixed token forms, compared with both sequential loaders
    verts, faces = synth.bumpy_sphere_mesh(150, 300)[:2]
    vts = rng.random((len(verts), 2)).astype(np.float32)
    big = str(tmp_path / "big.obj")
    synth.write_obj(big, verts, faces, vts, faces)
    assert os.path.getsize(big) > 4 << 20
    a = oracle.obj_load(big)
    b = rtb.OBJLoader.load(big).triangles
    r = refhost.mesh_triangles(refhost.obj_load(big))
    assert len(a) == len(b) == len(r) == len(faces)
    for k in ("v", "n", "uv"):
        assert a[k].tobytes() == b[k].tobytes() == r[k].tobytes(), k
    # an index that names no vertex fails loudly (the reference reads out of bounds); `v//vn` corners carry no uv
    bad = str(tmp_path / "bad.obj")
    open(bad, "w").write("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 7\n")
    with pytest.raises(Exception):
        rtb.OBJLoader.load(bad)
    vn = str(tmp_path / "vn.obj")
    open(vn, "w").write("v 0 0 0\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\n")
    t = rtb.OBJLoader.load(vn).triangles
    assert len(t) == 1 and not t["uv"].any() and np.allclose(t["n"][0], (0, 0, 1), atol=2e-3)


# ------------------------------------------------------------------ product host math (kernel inputs) vs oracle
def test_product_host_math_bit_exact(oracle, rtb):
    rng = np.random.default_rng(11)
    for _ in range(200):
        pose = rng.uniform(-1, 1, 6).astype(np.float32) * np.array([20, 20, 20, 3.14, 1.5, 3.14], np.float32)
        assert rtb.invert_lre(pose).tobytes() == oracle.invert_lre(pose).tobytes()
        scale = rng.uniform(0.3, 3.0, 3).astype(np.float32)
        mi = rtb.MeshInstance(3, 1, pose, scale)
        assert bytes(mi.c) == oracle.build_instance(3, 1, pose, scale).tobytes()
    K = synth.REF_K
    assert rtb.invert_intrinsic(K).tobytes() == oracle.invert_intrinsic(K).tobytes()
    t = np.zeros(1, orc.TRI)
    a, b, c = (np.array(x, np.float32) for x in ([0, 0, 0], [1, 0.5, 0.2], [0.1, 1.3, -0.4]))
    rtb.lib().rtb_triangle_init(t.ctypes.data, a.ctypes.data, b.ctypes.data, c.ctypes.data)
    ref = synth.make_triangles(np.stack([a, b, c]), np.array([[0, 1, 2]]))
    assert t.tobytes() == ref.tobytes()


# ------------------------------------------------------------------ the C-ABI library
def test_cabi_exports_every_declared_symbol(rtb):
    hdr = open(os.path.join(ROOT, "include", "rtb200.h")).read()
    declared = re.findall(r"RTB_API\s+[\w\s\*]+?\b(rtb_\w+)\s*\(", hdr)
    assert len(declared) >= 24
    L = rtb.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert set(rtb.EXPORTS) <= set(declared)


def test_cabi_rejects_null_arguments_without_a_device(rtb):
    """Error behaviour of the boundary: every entry point that takes a context returns RTB_ERR_INVALID (-1) and sets
    rtb_last_error for NULL arguments -- no crash, no device needed."""
    L = rtb.lib()
    calls = {
        "rtb_set_option": (None, b"counters", 1), "rtb_mesh_create": (None, None, 0, 0, None), "rtb_mesh_rebuild": (None, 0, None, 0, 0),
        "rtb_mesh_rebuild_async": (None, 0, None, 0, 1), "rtb_mesh_commit": (None, 0),
        "rtb_mesh_info_get": (None, 0, None), "rtb_mesh_destroy": (None, 0), "rtb_mesh_read": (None, 0, 0, None, 0),
        "rtb_texture_upload": (None, None, 0, 0, None, None), "rtb_set_materials": (None, None, 0), "rtb_set_instances": (None, None, 0),
        "rtb_update_instance": (None, 0, None), "rtb_render": (None, None, None, 0, None, None), "rtb_render_host": (None, None, None, 0, None),
        "rtb_render_noalloc": (None, None, None, 0, None, None),
        "rtb_synchronize": (None,), "rtb_frame_stats_get": (None, None), "rtb_frame_alloc_shared": (None, 0, None, None),
        "rtb_frame_free_shared": (None, None), "rtb_frame_open_shared": (None, None, None), "rtb_frame_close_shared": (None, None),
        "rtb_push_bands": (None, None, None, 0, 0, 0, 1, None), "rtb_debug_math": (None, 0, None, None, 0),
        "rtb_frame_signal": (None, None, 0, 0, None), "rtb_frame_wait": (None, None, 0, 0, 0, None),
    }
    for name, args in calls.items():
        assert getattr(L, name)(*args) == -1, name
        assert name.encode() in L.rtb_last_error(), (name, L.rtb_last_error())
    L.rtb_context_destroy(None)                                   # a no-op, like free(NULL)
    assert rtb.lib().rtb_context_create(0, None) == -1


def test_cpp_dropin_builds():
    """The C++ classes with the reference's names compile and link against the C-ABI (kernel.cu-style demo)."""
    cpp = os.path.join(ROOT, "cuda_raytracing_b200", "cpp")
    subprocess.check_call(["make", "-C", cpp, "-s"])
    assert os.path.exists(os.path.join(cpp, "demo"))
    for h in ("Scene.h", "Camera.h", "MeshInstance.hpp", "Material.hpp", "MeshPrimitive.h", "OBJLoader.hpp", "TrianglePrimitive.hpp",
              "transforms.hpp", "utils.hpp", "BVHTree.hpp"):
        assert os.path.exists(os.path.join(cpp, h)), h


def test_no_gpu_fails_loudly(rtb):
    """No CPU fallback: without a usable device the context cannot be created."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ctx = C.c_void_p()
    rc = rtb.lib().rtb_context_create(0, C.byref(ctx))
    assert rc == -3 and b"no CPU fallback" in rtb.lib().rtb_last_error()
    with pytest.raises(rtb.RtbError):
        rtb.Scene(0)


def test_product_does_not_reference_oracle():
    """The oracle is test infrastructure: nothing under the package may import, link or open it."""
    pkg = os.path.join(ROOT, "cuda_raytracing_b200")
    for dp, _, files in os.walk(pkg):
        if "build" in dp.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "librt_oracle" not in txt and "import orc" not in txt and "oracle/" not in txt.replace("oracle/rt_oracle.c", ""), f


# ------------------------------------------------------------------ multi-GPU band partition over gloo (world_size 2)
GLOO_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from cuda_raytracing_b200 import tiles
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
H, pitch = 75, 96
full = (torch.arange(H * pitch, dtype=torch.int64).reshape(H, pitch) * 7 % 251).to(torch.uint8)
frame = torch.zeros((H, pitch), dtype=torch.uint8)
rows = tiles.rank_rows(H, rank, world)
frame[rows] = full[rows]                      # "render" only this rank's bands
fg = tiles.FrameGather(H, pitch, rank, world, "cpu")
for _ in range(2):
    fg.gather(frame)
if rank == 0:
    assert torch.equal(frame, full), "gathered frame differs from the single-rank frame"
    cover = torch.cat([tiles.rank_rows(H, r, world) for r in range(world)]).sort().values
    assert torch.equal(cover, torch.arange(H))
    print("GLOO_OK")
dist.destroy_process_group()
'''


def test_band_partition_gather_gloo(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(GLOO_WORKER)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                          "--master-port", "29617", str(w), ROOT], capture_output=True, text=True, timeout=240)
    assert "GLOO_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


def test_band_rows_partition():
    import torch
    from cuda_raytracing_b200 import tiles
    for H in (1, 7, 8, 9, 480, 2160):
        for world in (1, 2, 3, 4, 8):
            rows = torch.cat([tiles.rank_rows(H, r, world) for r in range(world)])
            assert sorted(rows.tolist()) == list(range(H))


def test_headless_presentation_script_overlay_png(oracle, tmp_path):
    """presentation.hpp (kernel.cu:19-139 without a window): scripted WASD steps and mouse-look reproduce the reference's pose
    arithmetic (checked with the oracle's invert_lre / apply_euler), the PNG decodes to the frame plus a green FPS overlay."""
    import struct
    import zlib
    cpp = os.path.join(ROOT, "cuda_raytracing_b200", "cpp")
    subprocess.check_call(["make", "-C", cpp, "-s"])
    script = tmp_path / "in.txt"
    script.write_text("# frame 0: drag\ndown 100 100\nmove 100 100\nmove 160 70\nup 160 70\nmove 300 300\nframe\nkey w\nkey d\nframe\nkey s\nkey a\nkey x\nkey q\nkey w\n")
    png = tmp_path / "o.png"
    W, H = 330, 48
    r = subprocess.run([os.path.join(cpp, "presentation_test"), str(script), "9", str(png), str(W), str(H), "12.5"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().split("\n")
    assert lines[-1] == "quit" and len(lines) == 9                               # frames 0..7, then 'q' ends the loop
    poses = [np.array(l.split()[1:], np.float64).astype(np.float32) for l in lines[:-1]]

    def step(p, v):                                                              # kernel.cu:54-62
        ip = oracle.invert_lre(p)
        d = np.array(v, np.float32) - np.array(ip[:3], np.float32)
        return oracle.apply_euler(np.array(ip[3:], np.float32), d)

    p = np.array([-1, -4, 2, 0, 0, 0], np.float32)
    p[3] = np.float32(np.float64(p[3]) + 60 * 0.001)                             # kernel.cu:131-132: only while the button is down
    p[4] = np.float32(np.float64(p[4]) + (-30) * -0.001)
    assert np.array_equal(poses[0], p), (poses[0], p)
    want = [p.copy()]
    for key in "wd-saxq":                                                        # one key per frame (waitKey); '-' = the bare `frame` line
        v = {"w": (0, 0.1, 0), "s": (0, -0.1, 0), "a": (-0.1, 0, 0), "d": (0.1, 0, 0)}.get(key)   # 'x' and 'q' move nothing
        if v is not None:
            p[:3] = step(p, v)
        want.append(p.copy())
    for f, (got, w) in enumerate(zip(poses, want)):
        assert np.allclose(got, w, rtol=0, atol=2e-6), (f, got, w)
    assert len(poses) == 8 == len(want) and not np.array_equal(poses[5], poses[4])

    raw = png.read_bytes()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, ihdr = 8, b"", None
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + body)
        if typ == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    assert ihdr == (W, H, 8, 2, 0, 0, 0)
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(H, 1 + 3 * W)
    assert np.all(rows[:, 0] == 0)
    rgb = rows[:, 1:].reshape(H, W, 3)
    yy, xx = np.mgrid[0:H, 0:W]
    frame = np.stack([(xx + yy) & 0xFF, yy & 0x7F, xx & 0xFF], axis=2).astype(np.uint8)    # the test image, as RGB
    overlay = np.any(rgb != frame, axis=2)
    assert overlay.any() and np.all(rgb[overlay] == (0, 255, 0))
    ys, xs = np.nonzero(overlay)
    assert ys.min() >= 9 and ys.max() <= 29 and xs.min() >= 10 and xs.max() < 10 + 18 * len("FPS: 12.500000")
    assert overlay[9:30, 10:25].all(axis=1)[0]                                   # top bar of the 'F'
