"""One BA call on a mid-size graph (so that the solver runs with diagonal CTAs, staged operands and the chain) —
meant to be run under `compute-sanitizer --tool racecheck|memcheck` (small image size keeps the accumulate stage short)."""
import os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import synth
n_kf = int(sys.argv[1]) if len(sys.argv) > 1 else 120
p = synth.config_c3(n_kf=n_kf, n_edges=8 * n_kf, ht=16, wd=24)
dev = torch.device("cuda:0")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
poses, disps = t(p.poses), t(p.disps)
dx, dz = droid_backends.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj),
                           p.t0, p.t1, 1, 1e-5, 1e-2, False)
torch.cuda.synchronize()
print("race_run ok: n =", 6 * (p.t1 - p.t0), "dx norm", float(dx.norm()))
