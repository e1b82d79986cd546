"""rtb_render_noalloc (include/rtb200.h): the REFLECT / PATH frame with the fused kernel's scratch in the library's static device
array, for callers that must not allocate.  Checked through the C-ABI against rtb_render on the same inputs, bit for bit (the
tolerance the variant-identity tests use: frames of the fused variants are identical), on calls in a row that share the scratch.

Collected last on purpose (test_zz_*): a failure here hides nothing after it.
"""
import ctypes as C

import numpy as np
import pytest

from cuda_raytracing_b200 import synth

pytestmark = pytest.mark.gpu

W, H, PITCH = 316, 203, 1024                      # ragged: 26 bands, the last one 3 rows high; rows padded to 1024 bytes
POSES = [(0.5, -6.0, 2.5, 0.05, -0.3, 0.02), (-1.5, -5.0, 1.2, -0.2, -0.1, 0.0), (0.0, -2.6, 1.6, 0.0, -0.22, 0.0)]
CASES = [(2, 3, 1), (3, 2, 4)]                    # (mode, max_depth, spp): reflections, path tracing


def build_scene(rtb):
    s = rtb.Scene(0)
    s.add_material(rtb.Material(albedo=(0.2, 0.5, 0.9), metallic=0.6))
    s.add_material(rtb.Material(albedo=(0.9, 0.4, 0.1), metallic=0.1))
    s.add_mesh(rtb.MeshPrimitive(synth.bumpy_sphere(24, 48, ground=True)))
    s.add_mesh(rtb.MeshPrimitive(synth.bumpy_sphere(16, 32, ground=False)))
    s.add_mesh_instance(rtb.MeshInstance(0, 0, (0, 0, 0, 0, 0, 0), (1, 1, 1)))
    s.add_mesh_instance(rtb.MeshInstance(1, 1, (2.2, 0.5, 0.8, 0.3, 0.1, 0.0), (0.7, 0.7, 0.9)))
    s.add_mesh_instance(rtb.MeshInstance(1, 0, (-2.0, 1.0, 1.1, 0.0, 0.4, 0.2), (0.5, 0.6, 0.5)))
    s.upload_to_device()
    return s


class Frame:
    """device frame + AOVs (torch is only the allocator here; the calls under test go through ctypes)"""

    def __init__(self, torch):
        self.img = torch.full((H, PITCH), 0x5a, dtype=torch.uint8, device="cuda:0")
        self.prim = torch.full((H, W), -7, dtype=torch.int32, device="cuda:0")
        self.inst = torch.full((H, W), -7, dtype=torch.int32, device="cuda:0")
        self.t = torch.full((H, W), -7.0, dtype=torch.float32, device="cuda:0")

    def aovs(self):
        return (self.prim.data_ptr(), self.inst.data_ptr(), self.t.data_ptr())

    def host(self):
        return dict(img=self.img.cpu().numpy()[:, :W * 3].copy(), prim=self.prim.cpu().numpy(), inst=self.inst.cpu().numpy(), t=self.t.cpu().numpy())


def same(a, b, what):
    for k in ("img", "prim", "inst", "t"):
        assert a[k].tobytes() == b[k].tobytes(), "%s: %s differs in %d places" % (what, k, int((a[k] != b[k]).sum()))


def camera(rtb, pose, case, tiles=(0, 1)):
    cam = rtb.Camera(W, H, synth.C1_K * 0.5, synth.REF_D)
    cam.pose = np.array(pose, np.float32)
    cam.mode, cam.max_depth, cam.spp, cam.seed = case[0], case[1], case[2], 42
    cam.tile_rank, cam.tile_count = tiles
    return cam


def test_noalloc_matches_rtb_render_on_calls_in_a_row(rtb):
    import torch
    s = build_scene(rtb)
    for case in CASES:
        want = []
        for pose in POSES:                                       # the existing entry point (also makes the scene's tables resident)
            f = Frame(torch)
            camera(rtb, pose, case).render_scene(s, f.img.data_ptr(), PITCH, synchronize=True, aovs=f.aovs())
            want.append(f.host())
        assert want[0]["img"].any() and (want[0]["prim"] >= 0).any() and want[0]["img"].tobytes() != want[1]["img"].tobytes()
        # two calls in a row on one stream, no synchronisation between them: the second starts from the first one's scratch contents
        frames = [Frame(torch) for _ in POSES]
        for pose, f in zip(POSES, frames):
            camera(rtb, pose, case).render_scene_noalloc(s, f.img.data_ptr(), PITCH, aovs=f.aovs())
        s.synchronize()
        for i, f in enumerate(frames):
            same(f.host(), want[i], "mode %d, call %d in a row" % (case[0], i))
        # and the same call repeated: identical again
        f = Frame(torch)
        for _ in range(2):
            camera(rtb, POSES[0], case).render_scene_noalloc(s, f.img.data_ptr(), PITCH, synchronize=True, aovs=f.aovs())
            same(f.host(), want[0], "mode %d, repeated" % case[0])
        st = s.frame_stats()                                     # raises on overflow
        assert st["kernel_variant"] == 1 and st["launches"] >= 3
    # a band partition (what a rank of a multi-GPU job renders): rows of the other rank stay untouched
    case = CASES[0]
    full = Frame(torch)
    camera(rtb, POSES[2], case).render_scene(s, full.img.data_ptr(), PITCH, synchronize=True, aovs=full.aovs())
    full = full.host()
    part = Frame(torch)
    camera(rtb, POSES[2], case, tiles=(1, 2)).render_scene_noalloc(s, part.img.data_ptr(), PITCH, synchronize=True, aovs=part.aovs())
    part = part.host()
    from cuda_raytracing_b200 import tiles
    mine = tiles.rank_rows(H, 1, 2).numpy()
    other = tiles.rank_rows(H, 0, 2).numpy()
    assert np.array_equal(part["img"][mine], full["img"][mine]) and np.array_equal(part["prim"][mine], full["prim"][mine])
    assert (part["img"][other] == 0x5a).all() and (part["prim"][other] == -7).all()
    s.close()


def test_noalloc_allocates_nothing_and_rejects_what_it_cannot_hold(rtb):
    import torch
    s = build_scene(rtb)
    case = CASES[0]
    f = Frame(torch)
    cam = camera(rtb, POSES[0], case)
    # the scene's device tables are not resident before the first rtb_render: refused, with a message that says what to do
    with pytest.raises(rtb.RtbError, match="rtb_render first"):
        cam.render_scene_noalloc(s, f.img.data_ptr(), PITCH)
    cam.render_scene(s, f.img.data_ptr(), PITCH, synchronize=True)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    for _ in range(3):
        cam.render_scene_noalloc(s, f.img.data_ptr(), PITCH, synchronize=True, aovs=f.aovs())
    assert torch.cuda.mem_get_info(0)[0] == free0                # the device's free memory did not move
    # beyond the static scratch: a clear error, nothing launched
    big = rtb.Camera(1920, 1080, synth.REF_K, synth.REF_D)
    big.pose = np.array(POSES[0], np.float32)
    big.mode, big.max_depth = 2, 3
    with pytest.raises(rtb.RtbError, match="RTB_NOALLOC_MAX_SLOTS"):
        big.render_scene_noalloc(s, f.img.data_ptr(), 1920 * 3)
    # modes without scratch are pointed at rtb_render; NULL arguments are RTB_ERR_INVALID
    cam.mode = 1
    with pytest.raises(rtb.RtbError, match="rtb_render"):
        cam.render_scene_noalloc(s, f.img.data_ptr(), PITCH)
    assert rtb.lib().rtb_render_noalloc(None, None, None, 0, None, None) == -1
    assert rtb.NOALLOC_MAX_SLOTS == 131072
    # a scene change makes the tables stale again until the next rtb_render
    cam.mode = 2
    s.update_mesh_instance(1, rtb.MeshInstance(1, 1, (2.0, 0.6, 0.8, 0.3, 0.1, 0.0), (0.7, 0.7, 0.9)))
    try:
        cam.render_scene_noalloc(s, f.img.data_ptr(), PITCH, synchronize=True)
        stale = False
    except rtb.RtbError:
        stale = True
    g = Frame(torch)
    cam.render_scene(s, g.img.data_ptr(), PITCH, synchronize=True, aovs=g.aovs())
    cam.render_scene_noalloc(s, f.img.data_ptr(), PITCH, synchronize=True, aovs=f.aovs())
    same(f.host(), g.host(), "after an instance update (stale tables refused first: %s)" % stale)
    s.close()
