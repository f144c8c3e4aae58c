"""GPU box helper: closest-hit mismatches between frt_trace_rays and the float64 oracle → gpurun_out/."""
import sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / "tests"))
import oracle_c, refutil
from first_raytracer_b200 import api
name, aspect = (sys.argv[1], float(sys.argv[2])) if len(sys.argv) > 2 else ("cornell", 1.0)
scene = api.HostScene(name, aspect, 4)
info = scene.info
rays = refutil.random_rays(300_000, np.array(info["bounds_lo"]), np.array(info["bounds_hi"]), 11)
ctx = api.Context(); ctx.upload(scene)
hits, _ = ctx.trace(rays)
want = oracle_c.Oracle(scene).trace(rays)
bad = np.flatnonzero(hits["prim_id"] != want["prim"])
print(name, "mismatches", bad.size, "of", rays.shape[0])
(REPO / "gpurun_out").mkdir(exist_ok=True)
np.savez(REPO / "gpurun_out" / f"mismatch_{name}.npz", rays=rays[bad], gpu_prim=hits["prim_id"][bad], gpu_t=hits["t"][bad],
         ref_prim=want["prim"][bad], ref_t=want["t"][bad])
