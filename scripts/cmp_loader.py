"""Build container: host OBJ loader vs the reference's mesh_loader.cpp over Assimp (oracle/_ref/frt_ref_assimp dump)."""
import sys, pathlib
R = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R)); sys.path.insert(0, str(R / "tests"))
import numpy as np
import refutil
from first_raytracer_b200 import api
refutil.REF_BIN = R / "oracle/_ref" / (sys.argv[1] if len(sys.argv) > 1 else "frt_ref_assimp")
for name, aspect in [("cornell", 1.0), ("veach", 1.5), ("teapot", 16 / 9), ("cornell_sphere", 1.0), ("cornell_glossy", 1.0)]:
    ref, info = refutil.ref_dump(name, aspect)
    scene = api.HostScene(name, aspect, 4)
    a = scene.arrays()
    tris = a["tris"]
    order = np.argsort(tris[:, 0, 3].view(np.uint32))
    is_tri = ref["kind"] == 0
    g = ref["geom"][is_tri]
    dv = max(np.abs(tris[order, k, :3] - g[:, k]).max() for k in range(3))
    dn = np.abs(a["tri_normals"][order][:, :, :3] - g[:, 3:6]).max() if a["tri_normals"].size else -1
    mats = a["materials"]; mid = tris[order, 1, 3].view(np.uint32)
    print(f"{name:16s} tris {len(g)} vs {len(order)}: max |dv| {dv:.3g}, max |dn| {dn:.3g}, material kinds equal {(mats['type'][mid] == ref['mat_kind'][is_tri]).mean():.4f}, "
          f"params {np.abs(mats['p'][mid][:, :11] - ref['mat_p'][is_tri][:, :11]).max():.3g}, nodes {scene.info['bvh_nodes']} vs {info['bvh_nodes']}")
