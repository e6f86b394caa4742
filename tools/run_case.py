"""Render one workload a few times and print the kernel time: a short command for ncu / variant sweeps.
usage: python tools/run_case.py c3 [width height] [reps] ; RM_B200_LIB selects an alternative build."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from ray_marching_b200 import RenderLoop  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
wl = bench.workload(name)
W = int(sys.argv[2]) if len(sys.argv) > 3 else wl["width"]
H = int(sys.argv[3]) if len(sys.argv) > 3 else wl["height"]
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda")
loop = RenderLoop(wl["scene"](), num_cameras=1, px_width=W, px_height=H, focal_length=bench.PX * H,
                  sensor_width=bench.PX * W, sensor_height=bench.PX * H, normals_eps=bench.EPS).to(dev)
if os.environ.get("RM_CULL") == "1":
    loop.cull = True
if os.environ.get("RM_CULL") == "exact":
    loop.cull = "exact"
if os.environ.get("RM_SPEC") == "1":
    loop.specialise(True)
q, t = bench.ring_pose(0, 1, wl["cam_t"], dev)
shard = {}
if os.environ.get("RM_SHARD"):                       # "g/G": rank g of G, row-interleaved strips (one GPU emulating a rank)
    from ray_marching_b200 import parallel
    g_, G_ = (int(x) for x in os.environ["RM_SHARD"].split("/"))
    shard = parallel.shard_strips(W * H, g_, G_, bench.STRIP_ROWS * W)
res = None
ts = []
for i in range(reps + 1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    res = loop.render_planes(q, t, modes=wl["modes"], marching_steps=bench.STEPS, bounces=wl["bounces"], out=res, **shard)
    e1.record()
    e1.synchronize()
    ts.append(e0.elapsed_time(e1))
print(f"{name} {W}x{H} shard={os.environ.get('RM_SHARD', '-')} lib={os.environ.get('RM_B200_LIB', 'default')}: {min(ts[1:]):.3f} ms (best of {reps})")
