"""ctypes bindings for the test oracles (TEST INFRASTRUCTURE ONLY -- never imported by the product).

* ``Oracle``  -> oracle/librt_oracle.so   : the C restatement (oracle/rt_oracle.c)
* ``RefHost`` -> oracle/_ref/libref_host.so : the reference's own sources compiled as host C++
* ``RefCuda`` -> oracle/_ref/libref_cuda.so : the reference's CUDA kernel for sm_100a (GPU box only)
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

TRI = np.dtype([("v", "<f4", (3, 3)), ("n", "<f4", (3,)), ("uv", "<f4", (3, 2))])   # 72 B, TrianglePrimitive.hpp:8-11
LRE = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("yaw", "<f4"), ("pitch", "<f4"), ("roll", "<f4")])
INSTANCE = np.dtype([("mesh_index", "<i4"), ("material_index", "<i4"), ("pose", LRE), ("inv_pose", LRE),
                     ("rotation", "<f4", (3,)), ("inv_rotation", "<f4", (3,)), ("scale", "<f4", (3,)),
                     ("inv_scale", "<f4", (3,))])                                       # 104 B, MeshInstance.hpp:6-18
MATERIAL = np.dtype([("roughness", "<f4"), ("albedo", "<f4", (3,)), ("metallic", "<f4"), ("illumination", "<f4"),
                     ("texture", "<u8"), ("texture_width", "<i4"), ("texture_height", "<i4"),
                     ("texture_pitch", "<u8")])                                         # 48 B, Material.hpp:6-19
NODE = np.dtype([("min", "<f4", (3,)), ("max", "<f4", (3,)), ("tri_offset", "<i4"), ("tri_count", "<i4"),
                 ("child_a", "<i4"), ("child_b", "<i4")])
assert TRI.itemsize == 72 and LRE.itemsize == 24 and INSTANCE.itemsize == 104 and MATERIAL.itemsize == 48

MODE_REFERENCE, MODE_SHADOW, MODE_REFLECT, MODE_PATH = 0, 1, 2, 3


class OMesh(C.Structure):
    _fields_ = [("num_triangles", C.c_int), ("triangles", C.c_void_p), ("num_nodes", C.c_int), ("nodes", C.c_void_p),
                ("num_indices", C.c_int), ("indices", C.c_void_p)]


class OScene(C.Structure):
    _fields_ = [("num_meshes", C.c_int), ("meshes", C.POINTER(C.POINTER(OMesh))), ("num_materials", C.c_int),
                ("materials", C.c_void_p), ("num_instances", C.c_int), ("instances", C.c_void_p)]


class OLre(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("x", "y", "z", "yaw", "pitch", "roll")]


class ORenderParams(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("K_inv", C.c_float * 9), ("D", C.c_float * 4),
                ("pose", OLre), ("inv_pose", OLre), ("mode", C.c_int), ("max_depth", C.c_int), ("spp", C.c_int),
                ("seed", C.c_uint32), ("y0", C.c_int), ("y1", C.c_int)]


class OStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("nodes_popped", C.c_uint64), ("boxes_tested", C.c_uint64),
                ("tris_tested", C.c_uint64)]


class OF3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class OF4(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


def build_oracle():
    """Compile oracle/librt_oracle.so if it is missing or stale (gcc only; a few seconds)."""
    so = os.path.join(ORACLE_DIR, "librt_oracle.so")
    src = os.path.join(ORACLE_DIR, "rt_oracle.c")
    if (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return so


def _vp(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """The C restatement."""

    def __init__(self):
        self.lib = L = C.CDLL(build_oracle())
        L.o_Q_rsqrt.restype = C.c_float
        L.o_Q_rsqrt.argtypes = [C.c_float]
        L.o_normalize.restype = OF3
        L.o_normalize.argtypes = [OF3]
        L.o_apply_euler.restype = OF3
        L.o_apply_euler.argtypes = [OF3, OF3]
        L.o_euler2quat.restype = OF4
        L.o_euler2quat.argtypes = [OF3]
        L.o_invert_lre.restype = OLre
        L.o_invert_lre.argtypes = [OLre]
        L.o_invert_intrinsic.argtypes = [C.c_void_p, C.c_void_p]
        L.o_instance_build_inv.argtypes = [C.c_void_p]
        L.o_mesh_create.restype = C.POINTER(OMesh)
        L.o_mesh_create.argtypes = [C.c_void_p, C.c_int]
        L.o_mesh_destroy.argtypes = [C.POINTER(OMesh)]
        L.o_mesh_stats.argtypes = [C.POINTER(OMesh), C.c_void_p]
        L.o_obj_load.restype = C.c_int
        L.o_obj_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.o_free.argtypes = [C.c_void_p]
        L.o_render.argtypes = [C.POINTER(OScene), C.POINTER(ORenderParams), C.c_void_p, C.c_size_t, C.c_void_p,
                               C.c_void_p, C.c_void_p, C.c_int, C.POINTER(OStats)]
        L.o_candidate.argtypes = [C.POINTER(OScene), OF3, OF3, C.c_int, C.c_int, C.POINTER(C.c_float)]
        L.o_cast_ray_exhaustive.argtypes = [C.POINTER(OScene), OF3, OF3, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.o_pixel_ray.argtypes = [C.POINTER(ORenderParams), C.c_int, C.c_int, C.POINTER(OF3), C.POINTER(OF3)]
        L.o_set_exhaustive.argtypes = [C.c_int]
        L.o_render_pixel.argtypes = [C.POINTER(OScene), C.POINTER(ORenderParams), C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_int32),
                                     C.POINTER(C.c_int32), C.POINTER(C.c_float)]

    def q_rsqrt(self, x):
        return np.float32(self.lib.o_Q_rsqrt(np.float32(x)))

    def normalize(self, v):
        r = self.lib.o_normalize(OF3(*[float(t) for t in v]))
        return np.array([r.x, r.y, r.z], np.float32)

    def apply_euler(self, e, v):
        r = self.lib.o_apply_euler(OF3(*[float(t) for t in e]), OF3(*[float(t) for t in v]))
        return np.array([r.x, r.y, r.z], np.float32)

    def euler2quat(self, e):
        r = self.lib.o_euler2quat(OF3(*[float(t) for t in e]))
        return np.array([r.x, r.y, r.z, r.w], np.float32)

    def invert_lre(self, p):
        r = self.lib.o_invert_lre(OLre(*[float(t) for t in p]))
        return np.array([r.x, r.y, r.z, r.yaw, r.pitch, r.roll], np.float32)

    def invert_intrinsic(self, K):
        K = np.ascontiguousarray(K, np.float32).reshape(9)
        out = np.zeros(9, np.float32)
        self.lib.o_invert_intrinsic(_vp(K), _vp(out))
        return out.reshape(3, 3)

    def build_instance(self, mesh_index, material_index, pose=(0, 0, 0, 0, 0, 0), scale=(1, 1, 1)):
        inst = np.zeros(1, INSTANCE)
        inst["mesh_index"] = mesh_index
        inst["material_index"] = material_index
        for k, n in enumerate(LRE.names):
            inst["pose"][n] = pose[k]
        inst["scale"] = scale
        self.lib.o_instance_build_inv(_vp(inst))
        return inst

    def obj_load(self, path):
        p = C.c_void_p()
        n = self.lib.o_obj_load(path.encode(), C.byref(p))
        if n < 0:
            raise FileNotFoundError(path)
        tris = np.frombuffer(C.string_at(p.value, n * 72), dtype=TRI).copy() if n else np.zeros(0, TRI)
        self.lib.o_free(p)
        return tris

    def mesh_create(self, tris):
        tris = np.ascontiguousarray(tris, TRI)
        return OracleMesh(self, self.lib.o_mesh_create(_vp(tris), len(tris)))


class OracleMesh:
    def __init__(self, orc, ptr):
        self.orc, self.ptr = orc, ptr

    @property
    def nodes(self):
        m = self.ptr.contents
        return np.frombuffer(C.string_at(m.nodes, m.num_nodes * NODE.itemsize), dtype=NODE)

    @property
    def indices(self):
        m = self.ptr.contents
        return np.frombuffer(C.string_at(m.indices, m.num_indices * 4), dtype=np.int32)

    def stats(self):
        out = np.zeros(6, np.int32)
        self.orc.lib.o_mesh_stats(self.ptr, _vp(out))
        return dict(nodes=int(out[0]), max_tris=int(out[1]), min_tris=int(out[2]), depth=int(out[3]),
                    leaves=int(out[4]), triangles=int(out[5]))

    def __del__(self):
        try:
            self.orc.lib.o_mesh_destroy(self.ptr)
        except Exception:
            pass


class OracleScene:
    """Host scene for the restatement.  materials: list of dict(albedo=, metallic=, texture=HxWx3 uint8 BGR or None)."""

    def __init__(self, orc, meshes, materials, instances):
        self.orc = orc
        self.meshes = meshes
        self._tex = []
        self.mats = np.zeros(len(materials), MATERIAL)
        for i, m in enumerate(materials):
            self.mats[i]["albedo"] = m.get("albedo", (1, 1, 1))
            self.mats[i]["roughness"] = m.get("roughness", 0.0)
            self.mats[i]["metallic"] = m.get("metallic", 0.0)
            self.mats[i]["illumination"] = m.get("illumination", 0.0)
            tex = m.get("texture")
            if tex is not None:
                tex = np.ascontiguousarray(tex, np.uint8)
                self._tex.append(tex)
                self.mats[i]["texture"] = tex.ctypes.data
                self.mats[i]["texture_width"] = tex.shape[1]
                self.mats[i]["texture_height"] = tex.shape[0]
                self.mats[i]["texture_pitch"] = tex.shape[1] * 3
        self.insts = np.concatenate(instances) if len(instances) else np.zeros(0, INSTANCE)
        self._mesh_ptrs = (C.POINTER(OMesh) * len(meshes))(*[m.ptr for m in meshes])
        self.c = OScene(len(meshes), self._mesh_ptrs, len(self.mats), self.mats.ctypes.data, len(self.insts),
                        self.insts.ctypes.data)

    def params(self, width, height, K, D, pose, mode=MODE_REFERENCE, max_depth=0, spp=1, seed=42, y0=0, y1=0):
        orc = self.orc
        p = ORenderParams()
        p.width, p.height = width, height
        Ki = orc.invert_intrinsic(K).reshape(9)
        for i in range(9):
            p.K_inv[i] = float(Ki[i])
        for i in range(4):
            p.D[i] = float(np.float32(D[i]))
        p.pose = OLre(*[float(np.float32(t)) for t in pose])
        ip = orc.invert_lre(pose)
        p.inv_pose = OLre(*[float(t) for t in ip])
        p.mode, p.max_depth, p.spp, p.seed, p.y0, p.y1 = mode, max_depth, spp, seed, y0, y1 or height
        return p

    # ---- parity-report support: why does a pixel differ?
    def pixel_ray(self, width, height, K, D, pose, x, y):
        p = self.params(width, height, K, D, pose)
        o, d = OF3(), OF3()
        self.orc.lib.o_pixel_ray(C.byref(p), int(x), int(y), C.byref(o), C.byref(d))
        return o, d

    def candidate(self, o, d, inst, prim):
        """world distance of (inst, prim) along the ray if it passes the reference's accept predicate, else None"""
        dist = C.c_float()
        ok = self.orc.lib.o_candidate(C.byref(self.c), o, d, int(inst), int(prim), C.byref(dist))
        return float(dist.value) if ok else None

    def exhaustive(self, o, d):
        """(prim, inst, distance) of the closest candidate with no BVH at all"""
        prim, inst, dist = C.c_int(), C.c_int(), C.c_float()
        self.orc.lib.o_cast_ray_exhaustive(C.byref(self.c), o, d, C.byref(prim), C.byref(inst), C.byref(dist))
        return prim.value, inst.value, float(dist.value)

    def render_pixel(self, width, height, K, D, pose, x, y, mode=MODE_REFERENCE, max_depth=0, spp=1, seed=42, exhaustive=False):
        """one pixel of render(): (bgr uint8[3], prim, inst, t).  exhaustive: with the BVH-free cast_ray (o_set_exhaustive)"""
        p = self.params(width, height, K, D, pose, mode, max_depth, spp, seed)
        bgr = np.zeros(3, np.uint8)
        prim, inst, t = C.c_int32(-1), C.c_int32(-1), C.c_float(0)
        self.orc.lib.o_set_exhaustive(1 if exhaustive else 0)
        try:
            self.orc.lib.o_render_pixel(C.byref(self.c), C.byref(p), int(x), int(y), _vp(bgr), C.byref(prim), C.byref(inst), C.byref(t))
        finally:
            self.orc.lib.o_set_exhaustive(0)
        return bgr, prim.value, inst.value, float(t.value)

    def render(self, width, height, K, D, pose, mode=MODE_REFERENCE, max_depth=0, spp=1, seed=42, y0=0, y1=0,
               nthreads=None, aovs=True, stats=False):
        orc = self.orc
        p = self.params(width, height, K, D, pose, mode, max_depth, spp, seed, y0, y1)
        img = np.zeros((height, width, 3), np.uint8)
        prim = np.full((height, width), -1, np.int32) if aovs else None
        inst = np.full((height, width), -1, np.int32) if aovs else None
        t = np.full((height, width), np.finfo(np.float32).max, np.float32) if aovs else None
        st = OStats()
        nthreads = nthreads or os.cpu_count() or 1
        orc.lib.o_render(C.byref(self.c), C.byref(p), _vp(img), width * 3, _vp(prim) if aovs else None,
                         _vp(inst) if aovs else None, _vp(t) if aovs else None, nthreads, C.byref(st) if stats else None)
        out = dict(img=img, prim=prim, inst=inst, t=t)
        if stats:
            out["stats"] = dict(rays=st.rays, nodes=st.nodes_popped, boxes=st.boxes_tested, tris=st.tris_tested)
        return out


class RefHost:
    """The reference's own sources as a host library (exists only where build_ref.sh has been run)."""

    PATH = os.path.join(ORACLE_DIR, "_ref", "libref_host.so")

    @classmethod
    def available(cls):
        return os.path.exists(cls.PATH)

    def __init__(self):
        self.lib = L = C.CDLL(self.PATH)
        L.ref_Q_rsqrt.restype = C.c_float
        L.ref_Q_rsqrt.argtypes = [C.c_float]
        for f in ("ref_mesh_create", "ref_obj_load", "ref_scene_create"):
            getattr(L, f).restype = C.c_void_p
        L.ref_mesh_create.argtypes = [C.c_void_p, C.c_int]
        L.ref_obj_load.argtypes = [C.c_char_p]
        L.ref_mesh_num_triangles.argtypes = [C.c_void_p]
        L.ref_mesh_get_triangles.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_mesh_num_nodes.argtypes = [C.c_void_p]
        L.ref_mesh_get_nodes.argtypes = [C.c_void_p] * 5
        L.ref_scene_add_material.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p,
                                             C.c_int, C.c_int]
        L.ref_scene_add_mesh.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_scene_add_instance.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.ref_scene_upload.argtypes = [C.c_void_p]
        L.ref_instance_build.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.ref_invert_lre.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_invert_intrinsic.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_apply_euler.argtypes = [C.c_void_p] * 3
        L.ref_normalize.argtypes = [C.c_void_p] * 2

    def sizeof(self, what):
        return self.lib.ref_sizeof(what)

    def q_rsqrt(self, x):
        return np.float32(self.lib.ref_Q_rsqrt(np.float32(x)))

    def normalize(self, v):
        v = np.asarray(v, np.float32)
        o = np.zeros(3, np.float32)
        self.lib.ref_normalize(_vp(v), _vp(o))
        return o

    def apply_euler(self, e, v):
        e, v = np.asarray(e, np.float32), np.asarray(v, np.float32)
        o = np.zeros(3, np.float32)
        self.lib.ref_apply_euler(_vp(e), _vp(v), _vp(o))
        return o

    def invert_lre(self, p):
        p = np.asarray(p, np.float32)
        o = np.zeros(6, np.float32)
        self.lib.ref_invert_lre(_vp(p), _vp(o))
        return o

    def build_instance(self, mesh_index, material_index, pose, scale):
        pose, scale = np.asarray(pose, np.float32), np.asarray(scale, np.float32)
        o = np.zeros(1, INSTANCE)
        self.lib.ref_instance_build(mesh_index, material_index, _vp(pose), _vp(scale), _vp(o))
        return o

    def mesh_create(self, tris):
        tris = np.ascontiguousarray(tris, TRI)
        return self.lib.ref_mesh_create(_vp(tris), len(tris))

    def obj_load(self, path):
        return self.lib.ref_obj_load(path.encode())

    def mesh_triangles(self, mesh):
        n = self.lib.ref_mesh_num_triangles(mesh)
        out = np.zeros(n, TRI)
        self.lib.ref_mesh_get_triangles(mesh, _vp(out))
        return out

    def mesh_nodes(self, mesh):
        n = self.lib.ref_mesh_num_nodes(mesh)
        mm = np.zeros((n, 6), np.float32)
        ch = np.zeros((n, 2), np.int32)
        cnt = np.zeros(n, np.int32)
        total = self.lib.ref_mesh_get_nodes(mesh, _vp(mm), _vp(ch), _vp(cnt), None)
        idx = np.zeros(total, np.int32)
        self.lib.ref_mesh_get_nodes(mesh, _vp(mm), _vp(ch), _vp(cnt), _vp(idx))
        return mm, ch, cnt, idx

    def scene(self, meshes, materials, instances):
        """meshes: handles; materials: dicts as OracleScene; instances: (mesh, mat, pose6, scale3) tuples."""
        s = self.lib.ref_scene_create()
        for m in materials:
            alb = np.asarray(m.get("albedo", (1, 1, 1)), np.float32)
            tex = m.get("texture")
            if tex is not None:
                tex = np.ascontiguousarray(tex, np.uint8)
                self.lib.ref_scene_add_material(s, _vp(alb), m.get("roughness", 0.0), m.get("metallic", 0.0),
                                                m.get("illumination", 0.0), _vp(tex), tex.shape[1], tex.shape[0])
            else:
                self.lib.ref_scene_add_material(s, _vp(alb), m.get("roughness", 0.0), m.get("metallic", 0.0),
                                                m.get("illumination", 0.0), None, 0, 0)
        for h in meshes:
            self.lib.ref_scene_add_mesh(s, h)
        for (mi, ma, pose, scale) in instances:
            pose, scale = np.asarray(pose, np.float32), np.asarray(scale, np.float32)
            self.lib.ref_scene_add_instance(s, mi, ma, _vp(pose), _vp(scale))
        self.lib.ref_scene_upload(s)
        return s

    def render(self, scene, width, height, K, D, pose, nthreads=None, y0=0, y1=0):
        K = np.ascontiguousarray(K, np.float32).reshape(9)
        D = np.ascontiguousarray(D, np.float32)
        pose = np.ascontiguousarray(pose, np.float32)
        pitch = (width * 3 + 511) // 512 * 512
        img = np.zeros((height, pitch), np.uint8)
        prim = np.full((height, width), -1, np.int32)
        inst = np.full((height, width), -1, np.int32)
        t = np.full((height, width), np.finfo(np.float32).max, np.float32)
        self.lib.ref_render(scene, width, height, _vp(K), _vp(D), _vp(pose), _vp(img), pitch, _vp(prim), _vp(inst),
                            _vp(t), nthreads or os.cpu_count() or 1, y0, y1 or height)
        return dict(img=img[:, :width * 3].reshape(height, width, 3).copy(), prim=prim, inst=inst, t=t)
