import sys, numpy as np, torch
sys.path.insert(0,'/root/repo')
from tygrit_b200 import B200PlannerConfig, SE2Pose, WholeBodyConfig, synth
from tygrit_b200.planner import create_planner
TUCK=[0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]
raised = [0.25, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]
for algo in ("fcit_wb","multilayer_rrtc"):
    p = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 19, whole_body_algorithm=algo, timeout=20.0))
    p.update_environment(synth.room_cloud(150_000), SE2Pose(0.0, 0.0, 0.0))
    arm, base, _ = synth.free_base_configs(p.world, 16, seed=3, device=torch.device("cuda", 0), base_box=(-3.2, 3.2, -2.2, 2.2))
    poses = base.cpu().numpy().astype(np.float64)
    for i in range(0, 16, 2):
        s = WholeBodyConfig(np.array(TUCK), SE2Pose(*poses[i]))
        g = WholeBodyConfig(np.array(TUCK if i % 4 else raised), SE2Pose(*poses[i + 1]))
        vs, vg = p.validate_config(s), p.validate_config(g)
        r = p.plan_whole_body(s, g)
        print(algo, i, "valid", vs, vg, "->", r.success, {k: round(v, 2) for k, v in r.stats.items()})
