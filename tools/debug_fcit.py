import sys, numpy as np, torch, ctypes as C
sys.path.insert(0,'/root/repo')
from tygrit_b200 import B200PlannerConfig, SE2Pose, WholeBodyConfig, synth, _native as N
from tygrit_b200.planner import create_planner, _halton
from tygrit_b200.fetch_model import JOINT_LOWER, JOINT_UPPER
p2 = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 19, whole_body_algorithm="fcit_wb", timeout=20.0))
p2.update_environment(synth.room_cloud(200_000), SE2Pose(0.0, 0.0, 0.0))
print("bounds", p2._pc_bounds_xy)
x0,x1,y0,y1 = p2._pc_bounds_xy
lo = np.concatenate([[x0,y0,-np.pi], JOINT_LOWER]); hi = np.concatenate([[x1,y1,np.pi], JOINT_UPPER])
u = _halton(4096, 11, 0); q = lo + u*(hi-lo)
ok = p2.validate_configs(q[:,3:].astype(np.float32), q[:,:3].astype(np.float32))
print("numpy-halton samples valid:", ok.mean())
p2.world.set_self_collision(False)
print("without self:", p2.validate_configs(q[:,3:].astype(np.float32), q[:,:3].astype(np.float32)).mean())
p2.world.set_self_collision(True)
_, bases, _ = synth.free_base_configs(p2.world, 8, seed=5, device=torch.device("cuda", 0), base_box=(-3.2, 3.2, -2.2, 2.2))
bases = bases.cpu().numpy().astype(np.float64)
tuck = np.array([0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0])
for i in range(0, 8, 2):
    r = p2.plan_whole_body(WholeBodyConfig(tuck, SE2Pose(*bases[i])), WholeBodyConfig(tuck, SE2Pose(*bases[i + 1])))
    print(r.success, r.failure, {k: round(v, 3) for k, v in r.stats.items()}, "dist", np.linalg.norm(bases[i,:2]-bases[i+1,:2]))
