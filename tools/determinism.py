"""Repeat the c3 (or c4) BA call and compare the results bit for bit (race detector for the solver's dataflow)."""
import os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import synth
cfg = sys.argv[1] if len(sys.argv) > 1 else "c3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
p = synth.config_c4() if cfg == "c4" else synth.config_c3()
dev = torch.device("cuda:0")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
ref = None
bad = 0
for it in range(reps):
    poses, disps = t(p.poses), t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, *args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
    cur = (dx.cpu().numpy().copy(), poses.cpu().numpy().copy())
    if ref is None:
        ref = cur
    else:
        same = np.array_equal(ref[0], cur[0]) and np.array_equal(ref[1], cur[1])
        if not same:
            bad += 1
            print(f"run {it}: differs, rel dx {np.linalg.norm(cur[0] - ref[0]) / np.linalg.norm(ref[0]):.3e}")
print(f"{cfg}: {reps} runs, {bad} differ from the first")
sys.exit(1 if bad else 0)
