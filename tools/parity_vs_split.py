"""How far the BA update moves with the summation grouping of the accumulate stage (DB200_BA_SPLIT) on the ill-conditioned
stereo / invalid-depth test problem, against the fp32 and the fp64 oracle (run once per split value: the library reads
the variable once per process)."""
import os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import synth
from oracle import oracle
which = sys.argv[1] if len(sys.argv) > 1 else "stereo"
if which == "stereo":
    ii, jj = synth.neighborhood_edges(0, 8, 3)
    ii = np.concatenate([ii, np.arange(8)]); jj = np.concatenate([jj, np.arange(8)])
    p = synth.make_ba_problem(ii, jj, 8, seed=77, sensor_frac=0.5, invalid_frac=0.05)
    LM, EP = 1e-4, 0.1
elif which == "c2":
    p = synth.config_c2(); LM, EP = 1e-4, 0.1
else:                                   # a c3-like backend graph the oracle finishes in seconds (full 48x64 maps)
    p = synth.config_c3(n_kf=64, n_edges=640); LM, EP = 1e-5, 1e-2
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
rel = lambda a, b: float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(np.asarray(b, np.float64)))
for iters in (1, 2):
    poses, disps = t(p.poses), t(p.disps)
    dx, dz = droid_backends.ba(poses, disps, t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj),
                               p.t0, p.t1, iters, LM, EP, False)
    a = (p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, iters, LM, EP)
    r32, r64 = oracle.ba(*a), oracle.ba(*a, dtype=np.float64)
    dx, dz = dx.cpu().numpy(), dz.cpu().numpy()
    print(f"{which} split={os.environ.get('DB200_BA_SPLIT', 'auto'):>4s} iters={iters}: dz gpu-vs-f32 {rel(dz, r32['dz']):.2e} gpu-vs-f64 {rel(dz, r64['dz']):.2e} f32-vs-f64 {rel(r32['dz'], r64['dz']):.2e} | "
          f"dx gpu-vs-f32 {rel(dx, r32['dx']):.2e} gpu-vs-f64 {rel(dx, r64['dx']):.2e} f32-vs-f64 {rel(r32['dx'], r64['dx']):.2e}")
