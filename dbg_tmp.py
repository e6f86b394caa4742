import torch, time
from ray_marching_b200 import RenderLoop, ops, scene as S
from ray_marching_b200.tape import tape_of
dev=torch.device("cuda:0")
PX=3.45e-6
for name, build, tt in [("scene2", S.make_test_scene2, (0,0.1,1.0)), ("syn64", lambda: S.make_synthetic_scene(64,0,True), (0,0,-4.0))]:
    W=H=1024
    loop=RenderLoop(build(), num_cameras=1, px_width=W, px_height=H, focal_length=PX*H, sensor_width=PX*W, sensor_height=PX*H, normals_eps=5e-2).to(dev)
    tape=tape_of(loop.scene)
    q=torch.tensor([[1.,0,0,0]],device=dev); t=torch.tensor([list(tt)],device=dev)
    par=tape.lowered_params(dev).detach(); poses=ops.pose_records(q,t); cam=loop.camera.struct()
    g3=torch.randn(W*H,3,device=dev); code=tape.code_tensor(dev)
    ref=None
    for tol in (0.0, 1e-7, 1e-6, 1e-5, 1e-4):
        for variant in ("all", "pos-only"):
            gs = (g3,g3,g3,None,g3,None) if variant=="all" else (None,None,g3,None,None,None)
            ts=[]
            for _ in range(4):
                torch.cuda.synchronize(); t0=time.perf_counter()
                dpar,dposes=ops._render_bwd(code,par,poses,cam.n_views,cam.width,cam.height,cam.focal_length,cam.sensor_width,cam.sensor_height,32,5e-2,*gs,tape.n_nodes,tape.max_depth,tol)
                torch.cuda.synchronize(); ts.append(time.perf_counter()-t0)
            if tol==0.0: ref=ref or {}; ref[variant]=(dpar.clone(),dposes.clone())
            e=(dpar-ref[variant][0]).abs().max()/ref[variant][0].abs().max(); e2=(dposes-ref[variant][1]).abs().max()/ref[variant][1].abs().max()
            print(f"{name} tol={tol:g} {variant:8s} {min(ts)*1e3:8.3f} ms  relerr dpar {e:.2e} dpose {e2:.2e}")
