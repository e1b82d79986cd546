import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from cuda_raytracing_b200 import api as rtb, synth
import orc
K = synth.C1_K * np.array([[0.5], [0.5], [1.0]], np.float32)
one = synth.make_triangles(np.array([[-1, 0, 0], [1, 0, 0], [0, 0, 2]], np.float32), np.array([[0, 1, 2]]))
empty = np.zeros(0, orc.TRI)
def run(meshes, insts, tag):
    s = rtb.Scene(0)
    s.add_material(rtb.Material())
    for m in meshes: s.add_mesh(rtb.MeshPrimitive(m))
    for (a, b) in insts: s.add_mesh_instance(rtb.MeshInstance(a, b))
    s.upload_to_device()
    cam = rtb.Camera(64, 48, K, synth.ZERO_D); cam.pose = np.array((0.1, -3, 1, 0, 0, 0), np.float32)
    g = cam.render_scene_host(s, aovs=True)
    print(tag, "hits", int((g["prim"] >= 0).sum()), "insts", np.unique(g["inst"]), [s.mesh_info(i) for i in range(len(meshes))])
    s.close()
run([one], [(0, 0)], "one")
run([one, one], [(0, 0), (1, 0)], "one,one")
run([empty, one], [(0, 0), (1, 0)], "empty,one")
run([one, empty], [(0, 0), (1, 0)], "one,empty")
run([one], [(0, 0), (0, 0)], "one x2 inst")
