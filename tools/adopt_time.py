import sys, torch, time
sys.path.insert(0,'/root/repo')
from tygrit_b200 import synth
from tygrit_b200.engine import CollisionWorld
dev=torch.device("cuda",0)
w = CollisionWorld(device=0, max_points=1 << 20, voxels_per_metre=90.0, max_voxels=192 << 20)
w.build(torch.from_numpy(synth.room_cloud(500_000)).to(dev))
n = w.used_bytes()
w2 = CollisionWorld(device=0, max_points=1 << 20, voxels_per_metre=90.0, max_voxels=192 << 20)
w2.blob_view(n).copy_(w.blob_view(n)); torch.cuda.synchronize()
for _ in range(3): w2.adopt_blob(n)
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): w2.adopt_blob(n)
e1.record(); torch.cuda.synchronize()
arm, base = synth.device_configs(300000, seed=1, device=dev)
print("adopt ms", e0.elapsed_time(e1)/20, "bytes", n, "verdicts equal", bool(torch.equal(w.check_configs(arm,base), w2.check_configs(arm,base))))
