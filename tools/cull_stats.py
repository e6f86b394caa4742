"""Experiment: evaluated / skipped children per warp with culling on (needs a -DRM_CULL_STATS build in RM_B200_LIB)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from ray_marching_b200 import RenderLoop, _lib
name, W, H = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
wl = bench.workload(name)
loop = RenderLoop(wl["scene"](), num_cameras=1, px_width=W, px_height=H, focal_length=bench.PX * H,
                  sensor_width=bench.PX * W, sensor_height=bench.PX * H, normals_eps=bench.EPS).cuda()
loop.cull = True
q, t = bench.ring_pose(0, 1, wl["cam_t"], torch.device("cuda"))
lib = _lib.lib()
out = (C.c_ulonglong * 8)()
lib.rm_debug_cull_stats(out, 1)
loop.render_planes(q, t, modes=wl["modes"], marching_steps=32, bounces=wl["bounces"])
lib.rm_debug_cull_stats(out, 0)
ev, bound, skipall, leaves, masked = [out[i] for i in range(5)]
n = sum(1 for _ in loop.scene.sdfs[1].sdfs) if hasattr(loop.scene, "sdfs") else 0
print(f"{name} {W}x{H}: evals/warp-level {ev}, with bound {bound} ({100*bound/ev:.1f} %), whole-union skips {skipall} "
      f"({100*skipall/ev:.1f} %), children evaluated {leaves} = {100*leaves/(ev*n):.1f} % of {n} x evals, masked {masked} "
      f"= {100*masked/(ev*n):.1f} %")
