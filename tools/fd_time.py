import sys, os, time
sys.path[:0] = ["/root/repo", "/root/repo/droid-slam_reserch_b200"]
import numpy as np, torch
import droid_backends
from droid_b200 import synth, graph
p = synth.config_c3(n_edges=2000)
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
poses, disps, intr = cu(p.poses), cu(p.disps), cu(p.intrinsics)
t = p.t1
ii, jj = torch.meshgrid(torch.arange(t), torch.arange(t), indexing="ij")
ii, jj = ii.reshape(-1).cuda(), jj.reshape(-1).cuda()
for _ in range(2): droid_backends.frame_distance(poses, disps, intr, ii, jj, 0.25)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): droid_backends.frame_distance(poses, disps, intr, ii, jj, 0.25)
e1.record(); torch.cuda.synchronize()
print("frame_distance 262144 pairs: ms", e0.elapsed_time(e1) / 5)
for _ in range(2): graph.proximity_edges(poses, disps, intr, t, 0, 0, 2, 2, 0.25, 16.0, 16 * t, False)
torch.cuda.synchronize()
e0.record()
for _ in range(5): gi, gj = graph.proximity_edges(poses, disps, intr, t, 0, 0, 2, 2, 0.25, 16.0, 16 * t, False)
e1.record(); torch.cuda.synchronize()
print("proximity_edges t=512: ms", e0.elapsed_time(e1) / 5, "edges", len(gi))
