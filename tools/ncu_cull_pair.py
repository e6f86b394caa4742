"""One full and one culled render of a workload in one process, for a single ncu invocation
(ncu -k regex:render_kernel --launch-skip 1 -c 2 python tools/ncu_cull_pair.py c5 3840 2160): launches are
full (warm-up), full, culled (warm-up), culled -> with --launch-skip 1 -c 3 the report holds full, culled, culled."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from ray_marching_b200 import RenderLoop  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c5"
wl = bench.workload(name)
W = int(sys.argv[2]) if len(sys.argv) > 3 else wl["width"]
H = int(sys.argv[3]) if len(sys.argv) > 3 else wl["height"]
dev = torch.device("cuda")
loop = RenderLoop(wl["scene"](), num_cameras=1, px_width=W, px_height=H, focal_length=bench.PX * H,
                  sensor_width=bench.PX * W, sensor_height=bench.PX * H, normals_eps=bench.EPS).to(dev)
q, t = bench.ring_pose(0, 1, wl["cam_t"], dev)
for cull in (False, True):
    loop.cull = cull
    for _ in range(2):
        loop.render_planes(q, t, modes=wl["modes"], marching_steps=bench.STEPS, bounces=wl["bounces"])
torch.cuda.synchronize()
print("done")
