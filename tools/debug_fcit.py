import sys, numpy as np, torch
sys.path.insert(0,'/root/repo')
from tygrit_b200 import B200PlannerConfig, SE2Pose, WholeBodyConfig, synth
from tygrit_b200.planner import create_planner
TUCK=[0.0, 1.32, 1.40, -0.2, 1.72, 0.0, 1.66, 0.0]
p = create_planner("fetch", "b200", config=B200PlannerConfig(max_points=1 << 12))
p.clear_pointclouds()
s = WholeBodyConfig(np.array(TUCK), SE2Pose(0.0, 0.0, 0.0)); g = WholeBodyConfig(np.array(TUCK), SE2Pose(-1.5, 0.0, 0.0))
st = {}
sv = np.concatenate([[0,0,0],TUCK]); gv = np.concatenate([[-1.5,0,0],TUCK])
path = p._multilayer(sv, gv, st)
print("multilayer path", None if path is None else np.round(path[:, :3], 3).tolist(), st)
if path is not None:
    sc = p._shortcut(path, 50.0); print("shortcut", np.round(sc[:, :3], 3).tolist())
r = p.plan_whole_body(s, g); print(r.success, {k: round(v,2) for k,v in r.stats.items()})
