"""ctypes binding of librtb200.so (include/rtb200.h) with the reference's class names.

Mirrors the public surface of AFIDclan/cuda-raytracing (Scene.h:10-30, Camera.h:10-30, MeshInstance.hpp:6-48,
Material.hpp:6-44, MeshPrimitive.h:26-43, OBJLoader.hpp:12-15) for harnesses written in Python; the C++ classes
in cpp/ are the drop-in for C++ callers.  Everything here forwards to the C-ABI: there is no Python or CPU
implementation of any step, and importing this module without the built CUDA library raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librtb200.so")

TRI = np.dtype([("v", "<f4", (3, 3)), ("n", "<f4", (3,)), ("uv", "<f4", (3, 2))])
MODE_REFERENCE, MODE_SHADOW, MODE_REFLECT, MODE_PATH = 0, 1, 2, 3
BAND_ROWS = 8


class Lre(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("x", "y", "z", "yaw", "pitch", "roll")]


class CInstance(C.Structure):
    _fields_ = [("mesh_index", C.c_int), ("material_index", C.c_int), ("pose", Lre), ("inv_pose", Lre),
                ("rotation", C.c_float * 3), ("inv_rotation", C.c_float * 3), ("scale", C.c_float * 3),
                ("inv_scale", C.c_float * 3)]


class CMaterial(C.Structure):
    _fields_ = [("roughness", C.c_float), ("albedo", C.c_float * 3), ("metallic", C.c_float),
                ("illumination", C.c_float), ("texture", C.c_void_p), ("texture_width", C.c_int),
                ("texture_height", C.c_int), ("texture_pitch", C.c_size_t)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("K_inv", C.c_float * 9), ("D", C.c_float * 4),
                ("pose", Lre), ("inv_pose", Lre), ("mode", C.c_int), ("max_depth", C.c_int), ("spp", C.c_int),
                ("seed", C.c_uint32), ("tile_rank", C.c_int), ("tile_count", C.c_int)]


class Aovs(C.Structure):
    _fields_ = [("prim", C.c_void_p), ("inst", C.c_void_p), ("t", C.c_void_p)]


class MeshInfo(C.Structure):
    _fields_ = [("num_triangles", C.c_int), ("num_nodes", C.c_int), ("num_leaves", C.c_int),
                ("max_leaf_size", C.c_int), ("depth", C.c_int), ("stack_need", C.c_int), ("build_ms", C.c_float),
                ("sah_cost", C.c_float), ("bounds_min", C.c_float * 3), ("bounds_max", C.c_float * 3),
                ("device_bytes", C.c_size_t)]


class FrameStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("device_ms", C.c_float), ("overflow", C.c_int), ("launches", C.c_int), ("kernel_variant", C.c_int)]


assert C.sizeof(CInstance) == 104 and C.sizeof(CMaterial) == 48 and C.sizeof(Lre) == 24

NOALLOC_MAX_SLOTS = 131072        # RTB_NOALLOC_MAX_SLOTS (include/rtb200.h)

EXPORTS = ["rtb_context_create", "rtb_context_destroy", "rtb_last_error", "rtb_version", "rtb_set_option",
           "rtb_mesh_create", "rtb_mesh_rebuild", "rtb_mesh_rebuild_async", "rtb_mesh_commit", "rtb_mesh_info_get", "rtb_mesh_read", "rtb_mesh_destroy", "rtb_texture_upload",
           "rtb_set_materials", "rtb_set_instances", "rtb_update_instance", "rtb_render", "rtb_render_host",
           "rtb_synchronize", "rtb_frame_stats_get", "rtb_debug_math", "rtb_invert_intrinsic", "rtb_invert_lre",
           "rtb_instance_build_inv", "rtb_triangle_init", "rtb_obj_load", "rtb_free",
           "rtb_frame_alloc_shared", "rtb_frame_free_shared", "rtb_frame_open_shared", "rtb_frame_close_shared", "rtb_push_bands",
           "rtb_frame_signal", "rtb_frame_wait", "rtb_render_noalloc"]

_lib = None


class RtbError(RuntimeError):
    pass


def lib():
    """Load librtb200.so; raises (never falls back) when the CUDA library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtbError("librtb200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'` "
                       "or `make -C cuda_raytracing_b200/csrc`); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    L.rtb_last_error.restype = C.c_char_p
    vp, ci = C.c_void_p, C.c_int
    L.rtb_context_create.argtypes = [ci, C.POINTER(vp)]
    L.rtb_context_destroy.argtypes = [vp]
    L.rtb_context_destroy.restype = None
    L.rtb_set_option.argtypes = [vp, C.c_char_p, ci]
    L.rtb_mesh_create.argtypes = [vp, vp, ci, ci, C.POINTER(ci)]
    L.rtb_mesh_rebuild.argtypes = [vp, ci, vp, ci, ci]
    L.rtb_mesh_rebuild_async.argtypes = [vp, ci, vp, ci, ci]
    L.rtb_mesh_commit.argtypes = [vp, ci]
    L.rtb_mesh_info_get.argtypes = [vp, ci, C.POINTER(MeshInfo)]
    L.rtb_mesh_destroy.argtypes = [vp, ci]
    L.rtb_mesh_read.argtypes = [vp, ci, ci, vp, C.c_size_t]
    L.rtb_texture_upload.argtypes = [vp, vp, ci, ci, C.POINTER(vp), C.POINTER(C.c_size_t)]
    L.rtb_set_materials.argtypes = [vp, vp, ci]
    L.rtb_set_instances.argtypes = [vp, vp, ci]
    L.rtb_update_instance.argtypes = [vp, ci, vp]
    L.rtb_render.argtypes = [vp, C.POINTER(RenderParams), vp, C.c_size_t, C.POINTER(Aovs), vp]
    L.rtb_render_noalloc.argtypes = [vp, C.POINTER(RenderParams), vp, C.c_size_t, C.POINTER(Aovs), vp]
    L.rtb_render_host.argtypes = [vp, C.POINTER(RenderParams), vp, C.c_size_t, C.POINTER(Aovs)]
    L.rtb_synchronize.argtypes = [vp]
    L.rtb_frame_stats_get.argtypes = [vp, C.POINTER(FrameStats)]
    L.rtb_debug_math.argtypes = [vp, ci, vp, vp, ci]
    L.rtb_invert_intrinsic.argtypes = [vp, vp]
    L.rtb_invert_intrinsic.restype = None
    L.rtb_invert_lre.argtypes = [C.POINTER(Lre), C.POINTER(Lre)]
    L.rtb_invert_lre.restype = None
    L.rtb_instance_build_inv.argtypes = [C.POINTER(CInstance)]
    L.rtb_instance_build_inv.restype = None
    L.rtb_triangle_init.argtypes = [vp, vp, vp, vp]
    L.rtb_triangle_init.restype = None
    L.rtb_obj_load.argtypes = [C.c_char_p, C.POINTER(vp), C.POINTER(ci)]
    L.rtb_free.argtypes = [vp]
    L.rtb_frame_alloc_shared.argtypes = [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]
    L.rtb_frame_free_shared.argtypes = [vp, vp]
    L.rtb_frame_open_shared.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.rtb_frame_close_shared.argtypes = [vp, vp]
    L.rtb_push_bands.argtypes = [vp, vp, vp, C.c_size_t, ci, ci, ci, vp]
    L.rtb_frame_signal.argtypes = [vp, vp, ci, C.c_uint, vp]
    L.rtb_frame_wait.argtypes = [vp, vp, ci, ci, C.c_uint, vp]
    L.rtb_free.restype = None
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise RtbError("rtb error %d: %s" % (rc, lib().rtb_last_error().decode()))


def invert_intrinsic(K):
    K = np.ascontiguousarray(K, np.float32).reshape(9)
    out = np.zeros(9, np.float32)
    lib().rtb_invert_intrinsic(K.ctypes.data, out.ctypes.data)
    return out.reshape(3, 3)


def invert_lre(pose):
    a, b = Lre(*[float(np.float32(v)) for v in pose]), Lre()
    lib().rtb_invert_lre(C.byref(a), C.byref(b))
    return np.array([b.x, b.y, b.z, b.yaw, b.pitch, b.roll], np.float32)


class Material:
    """Material.hpp:6-44.  upload_texture takes raw BGR texels (HxWx3 uint8) instead of a file path:
    image decoding (OpenCV) is out of scope."""

    def __init__(self, albedo=(1.0, 1.0, 1.0), roughness=0.0, metallic=0.0, illumination=0.0):
        self.albedo, self.roughness, self.metallic, self.illumination = albedo, roughness, metallic, illumination
        self.texels = None

    def upload_texture(self, texels_bgr):
        self.texels = np.ascontiguousarray(texels_bgr, np.uint8)
        assert self.texels.ndim == 3 and self.texels.shape[2] == 3


class MeshPrimitive:
    """MeshPrimitive.h:26-43: owns the host triangle array (72-byte records)."""

    def __init__(self, triangles):
        self.triangles = np.ascontiguousarray(triangles, TRI)
        self.num_triangles = len(self.triangles)


class OBJLoader:
    @staticmethod
    def load(path):
        """OBJLoader.hpp:15: parse `path` -> MeshPrimitive (parsing done by the C++ loader in librtb200)."""
        p, n = C.c_void_p(), C.c_int()
        rc = lib().rtb_obj_load(path.encode(), C.byref(p), C.byref(n))
        if rc != 0:
            raise RtbError("Could not open file %s" % path)
        # one copy out of the loader's malloc'ed array (freed right below)
        tris = np.frombuffer((C.c_ubyte * (n.value * 72)).from_address(p.value), dtype=TRI).copy() if n.value else np.zeros(0, TRI)
        lib().rtb_free(p)
        return MeshPrimitive(tris)


class MeshInstance:
    """MeshInstance.hpp:6-48."""

    def __init__(self, mesh_index, material_index, pose=(0, 0, 0, 0, 0, 0), scale=(1.0, 1.0, 1.0)):
        self.c = CInstance()
        self.c.mesh_index, self.c.material_index = mesh_index, material_index
        self.c.pose = Lre(*[float(np.float32(v)) for v in pose])
        for k in range(3):
            self.c.scale[k] = float(np.float32(scale[k]))
        self.build_inv()

    def build_inv(self):
        lib().rtb_instance_build_inv(C.byref(self.c))

    def as_array(self):
        return np.frombuffer(bytes(self.c), dtype=np.uint8).copy()


class Scene:
    """Scene.h:10-30.  add_* copy by value; nothing reaches the GPU until upload_to_device()."""

    def __init__(self, device=0):
        self._ctx = C.c_void_p()
        _check(lib().rtb_context_create(device, C.byref(self._ctx)))
        self.device = device
        self.materials, self.meshes, self.mesh_instances = [], [], []
        self.mesh_ids = []
        self.num_mesh_instances = 0

    def close(self):
        if self._ctx:
            lib().rtb_context_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, name, value):
        _check(lib().rtb_set_option(self._ctx, name.encode(), int(value)))

    def add_material(self, material):
        self.materials.append(material)

    def add_mesh(self, mesh):
        self.meshes.append(mesh)

    def add_mesh_instance(self, inst):
        self.mesh_instances.append(inst)

    def upload_to_device(self):
        """Scene.cpp:25-65: builds every BLAS on the device, uploads materials, builds the TLAS."""
        L = lib()
        for i in range(len(self.mesh_ids), len(self.meshes)):
            m = self.meshes[i]
            mid = C.c_int()
            dev_ptr = getattr(m, "device_ptr", None)
            if dev_ptr is not None:
                _check(L.rtb_mesh_create(self._ctx, dev_ptr, m.num_triangles, 1, C.byref(mid)))
            else:
                _check(L.rtb_mesh_create(self._ctx, m.triangles.ctypes.data, m.num_triangles, 0, C.byref(mid)))
            self.mesh_ids.append(mid.value)
        mats = (CMaterial * max(1, len(self.materials)))()
        for i, m in enumerate(self.materials):
            mats[i].roughness, mats[i].metallic, mats[i].illumination = m.roughness, m.metallic, m.illumination
            for k in range(3):
                mats[i].albedo[k] = float(np.float32(m.albedo[k]))
            if m.texels is not None:
                if not hasattr(m, "_dev"):
                    d, pitch = C.c_void_p(), C.c_size_t()
                    h, w = m.texels.shape[:2]
                    _check(L.rtb_texture_upload(self._ctx, m.texels.ctypes.data, w, h, C.byref(d), C.byref(pitch)))
                    m._dev = (d.value, pitch.value, w, h)
                mats[i].texture, mats[i].texture_pitch, mats[i].texture_width, mats[i].texture_height = m._dev
        _check(L.rtb_set_materials(self._ctx, mats, len(self.materials)))
        insts = (CInstance * max(1, len(self.mesh_instances)))()
        for i, mi in enumerate(self.mesh_instances):
            mi.build_inv()
            insts[i] = mi.c
            insts[i].mesh_index = self.mesh_ids[mi.c.mesh_index]
        _check(L.rtb_set_instances(self._ctx, insts, len(self.mesh_instances)))
        self.num_mesh_instances = len(self.mesh_instances)

    def update_mesh_instance(self, index, inst):
        """Scene.cpp:67-74."""
        inst.build_inv()
        self.mesh_instances[index] = inst
        c = CInstance.from_buffer_copy(bytes(inst.c))
        c.mesh_index = self.mesh_ids[inst.c.mesh_index]
        _check(lib().rtb_update_instance(self._ctx, index, C.byref(c)))

    def rebuild_mesh(self, index, triangles=None, device_ptr=None, n=None):
        """Per-frame BLAS rebuild (animated geometry, config C5)."""
        if device_ptr is not None:
            _check(lib().rtb_mesh_rebuild(self._ctx, self.mesh_ids[index], device_ptr, n, 1))
        else:
            t = np.ascontiguousarray(triangles, TRI)
            _check(lib().rtb_mesh_rebuild(self._ctx, self.mesh_ids[index], t.ctypes.data, len(t), 0))

    def rebuild_mesh_async(self, index, device_ptr, n):
        """Enqueue the rebuild on the library's build stream (next to the frame being rendered); commit_mesh() swaps it in."""
        _check(lib().rtb_mesh_rebuild_async(self._ctx, self.mesh_ids[index], device_ptr, n, 1))

    def commit_mesh(self, index):
        _check(lib().rtb_mesh_commit(self._ctx, self.mesh_ids[index]))

    def mesh_info(self, index):
        info = MeshInfo()
        _check(lib().rtb_mesh_info_get(self._ctx, self.mesh_ids[index], C.byref(info)))
        return {f: (list(getattr(info, f)) if f.startswith("bounds") else getattr(info, f)) for f, _ in MeshInfo._fields_}

    def mesh_arrays(self, index):
        """Download the wide nodes (N,7,4 float32), triangle records (T,4,4 float32) and leaf order (T uint32)."""
        info = self.mesh_info(index)
        nodes = np.zeros((info["num_nodes"], 7, 4), np.float32)
        tris = np.zeros((info["num_triangles"], 4, 4), np.float32)
        order = np.zeros(info["num_triangles"], np.uint32)
        for what, arr in ((0, nodes), (1, tris), (2, order)):
            _check(lib().rtb_mesh_read(self._ctx, self.mesh_ids[index], what, arr.ctypes.data, arr.nbytes))
        return nodes, tris, order

    def debug_math(self, what, x):
        x = np.ascontiguousarray(x, np.float32)
        out = np.zeros_like(x)
        _check(lib().rtb_debug_math(self._ctx, what, x.ctypes.data, out.ctypes.data, x.size))
        return out

    # ---- multi-GPU frame exchange over peer memory (rtb200.h: rtb_frame_*_shared, rtb_push_bands) ----
    def shared_alloc(self, nbytes):
        """-> (device pointer, 64-byte IPC handle) of a frame other processes on this node can map."""
        ptr, handle = C.c_void_p(), C.create_string_buffer(64)
        _check(lib().rtb_frame_alloc_shared(self._ctx, nbytes, C.byref(ptr), handle))
        return ptr.value, handle.raw

    def shared_free(self, ptr):
        _check(lib().rtb_frame_free_shared(self._ctx, ptr))

    def shared_open(self, handle):
        ptr = C.c_void_p()
        _check(lib().rtb_frame_open_shared(self._ctx, C.create_string_buffer(bytes(handle), 64), C.byref(ptr)))
        return ptr.value

    def shared_close(self, ptr):
        _check(lib().rtb_frame_close_shared(self._ctx, ptr))

    def push_bands(self, src_ptr, dst_ptr, pitch, height, rank, world, stream=None):
        _check(lib().rtb_push_bands(self._ctx, src_ptr, dst_ptr, pitch, height, rank, world, stream))

    def frame_signal(self, flags_ptr, slot, value, stream=None):
        """Enqueue `flags[slot] = value` (system-scope release) behind what this rank put on `stream`."""
        _check(lib().rtb_frame_signal(self._ctx, flags_ptr, slot, value & 0xffffffff, stream))

    def frame_wait(self, flags_ptr, first, count, value, stream=None):
        """Enqueue a wait on `stream` until flags[first:first+count] have all reached `value`."""
        _check(lib().rtb_frame_wait(self._ctx, flags_ptr, first, count, value & 0xffffffff, stream))

    def frame_stats(self):
        st = FrameStats()
        _check(lib().rtb_frame_stats_get(self._ctx, C.byref(st)))
        return dict(rays=st.rays, nodes=st.nodes_visited, tris=st.tris_tested, device_ms=st.device_ms, launches=st.launches, kernel_variant=st.kernel_variant)

    def synchronize(self):
        _check(lib().rtb_synchronize(self._ctx))


class Camera:
    """Camera.h:10-30 / Camera.cu:6-41."""

    def __init__(self, width, height, K, D):
        self.width, self.height = width, height
        self.K = np.ascontiguousarray(K, np.float32).reshape(3, 3)
        self.K_inv = invert_intrinsic(self.K)
        self.D = np.ascontiguousarray(D, np.float32)
        self.pose = np.zeros(6, np.float32)
        self.mode, self.max_depth, self.spp, self.seed = MODE_REFERENCE, 0, 1, 42
        self.tile_rank, self.tile_count = 0, 1

    def params(self):
        p = RenderParams()
        p.width, p.height = self.width, self.height
        for i in range(9):
            p.K_inv[i] = float(self.K_inv.reshape(9)[i])
        for i in range(4):
            p.D[i] = float(self.D[i])
        pose = np.asarray(self.pose, np.float32)
        p.pose = Lre(*[float(v) for v in pose])
        p.inv_pose = Lre(*[float(v) for v in invert_lre(pose)])          # Camera.cu:21
        p.mode, p.max_depth, p.spp, p.seed = self.mode, self.max_depth, self.spp, self.seed
        p.tile_rank, p.tile_count = self.tile_rank, self.tile_count
        return p

    def render_scene(self, scene, img_ptr, pitch, synchronize=False, aovs=None, stream=None):
        """Camera.cu:18-41: img_ptr is a DEVICE pointer (int), BGR rows `pitch` bytes apart; returns after enqueue."""
        p = self.params()
        a = Aovs(*aovs) if aovs is not None else None
        _check(lib().rtb_render(scene._ctx, C.byref(p), img_ptr, pitch, C.byref(a) if a is not None else None, stream))
        if synchronize:
            scene.synchronize()

    def render_scene_noalloc(self, scene, img_ptr, pitch, synchronize=False, aovs=None, stream=None):
        """rtb_render_noalloc: the REFLECT / PATH frame of render_scene with the kernel's scratch in the library's static device
        array -- no allocation anywhere on the call; frames above NOALLOC_MAX_SLOTS pixels are rejected (RtbError)."""
        p = self.params()
        a = Aovs(*aovs) if aovs is not None else None
        _check(lib().rtb_render_noalloc(scene._ctx, C.byref(p), img_ptr, pitch, C.byref(a) if a is not None else None, stream))
        if synchronize:
            scene.synchronize()

    def render_scene_host(self, scene, aovs=False):
        """End-to-end form: host framebuffer (and host AOVs) out; device->host copies included."""
        p = self.params()
        img = np.zeros((self.height, self.width, 3), np.uint8)
        out = dict(img=img)
        a = None
        if aovs:
            out["prim"] = np.zeros((self.height, self.width), np.int32)
            out["inst"] = np.zeros((self.height, self.width), np.int32)
            out["t"] = np.zeros((self.height, self.width), np.float32)
            a = Aovs(out["prim"].ctypes.data, out["inst"].ctypes.data, out["t"].ctypes.data)
        _check(lib().rtb_render_host(scene._ctx, C.byref(p), img.ctypes.data, self.width * 3,
                                     C.byref(a) if a is not None else None))
        return out
