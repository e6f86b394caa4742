"""Time the fused loss + backward kernel of one workload (c4 / c4b): a short command for variant sweeps.
usage: python tools/run_loss.py c4b [reps] ; RM_B200_LIB selects an alternative build, RM_SPEC=1 the specialised one."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import bench  # noqa: E402
from ray_marching_b200 import RenderLoop  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4b"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
wl = bench.workload(name)
W, H = wl["width"], wl["height"]
dev = torch.device("cuda")
scene = wl["scene"]()
loop = RenderLoop(scene, num_cameras=1, px_width=W, px_height=H, focal_length=bench.PX * H,
                  sensor_width=bench.PX * W, sensor_height=bench.PX * H, normals_eps=bench.EPS).to(dev)
if os.environ.get("RM_SPEC") == "1":
    loop.specialise(True)
params = list(scene.parameters())
q0, t0 = bench.ring_pose(0, 1, wl["cam_t"], dev)
q, t = q0.clone().requires_grad_(True), t0.clone().requires_grad_(True)
gen = torch.Generator().manual_seed(1)
saved = [p.detach().clone() for p in params]
with torch.no_grad():
    for p in params:
        p.add_(0.05 * torch.randn(p.shape, generator=gen).to(dev))
    res = loop.render_planes(q0, t0, gbuffer=True, modes=(0,), marching_steps=bench.STEPS)
    lam_t = res.images[0].clone()
    depth_t = (res.pos - res.origin).norm(dim=-1, keepdim=True)
    for p, s0 in zip(params, saved):
        p.copy_(s0)
ts = []
for i in range(reps + 2):
    for p in params + [q, t]:
        p.grad = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loss = loop.inverse_loss(q, t, lam_t, depth_t, marching_steps=bench.STEPS)
    loss.backward()
    e1.record()
    e1.synchronize()
    ts.append(e0.elapsed_time(e1))
gn = torch.cat([p.grad.reshape(-1) for p in params]).norm().item()
print(f"{name} lib={os.path.basename(os.environ.get('RM_B200_LIB', 'default'))} spec={os.environ.get('RM_SPEC', '0')}: "
      f"{min(ts[2:]):.3f} ms fwd+bwd (best of {reps}), loss {loss.item():.6f}, |dpar| {gn:.6e}, dq {q.grad.cpu().numpy().round(5).tolist()}")
