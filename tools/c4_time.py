"""Stage times of one BA call at BASELINE config 4 size (1536 keyframes, 20k edges, 9210^2 reduced system) on one GPU."""
import ctypes as C, os, sys
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import _capi, synth
STAGES = ["graph", "edge_prep", "accumulate", "solve", "post", "backsub"]
cfg = sys.argv[1] if len(sys.argv) > 1 else "c4"
p = {"c4": synth.config_c4, "c3": synth.config_c3, "c2": synth.config_c2, "c1": synth.config_c1}[cfg]()
dev = torch.device("cuda:0")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
lib = _capi.lib()
reps = 3 if cfg in ("c3", "c4") else 30
for it in range(reps):
    poses, disps = t(p.poses), t(p.disps)
    torch.cuda.synchronize()
    if it == 1:
        lib.db200_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    droid_backends.ba(poses, disps, *args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
    e1.record()
    torch.cuda.synchronize()
    if it < 2 or it == reps - 1:
        print(f"call {it}: {e0.elapsed_time(e1):.3f} ms for 2 iterations")
st_ms = (C.c_double * len(STAGES))(); st_n = (C.c_int * len(STAGES))()
lib.db200_profile_collect(st_ms, st_n)
lib.db200_profile_enable(0)
for k, name in enumerate(STAGES):
    if st_n[k]:
        print(f"  {name:10s}: {st_ms[k] / st_n[k]:8.3f} ms per interval ({st_n[k]} intervals)")
n = 6 * (p.t1 - p.t0)
sol = st_ms[3] / max(st_n[3], 1)
print(f"solve: n = {n}, {n ** 3 / 3 / (sol * 1e-3) / 1e12:.2f} TFLOP/s fp64")
