"""Phase times (cycles) inside chol_small_kernel on the frontend-size problem c2 (debug aid; run on the GPU box)."""
import os, sys
import ctypes as C
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import _capi, synth
p = synth.config_c2() if len(sys.argv) < 2 or sys.argv[1] == "c2" else synth.config_c1()
dev = torch.device("cuda:0")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
buf = torch.zeros(8 * 1024 + 64, dtype=torch.int64, device=dev)
for it in range(3):
    poses, disps = t(p.poses), t(p.disps)
    if it == 2:
        _capi.lib().db200_debug_chol_trace(C.c_void_p(buf.data_ptr()))
    droid_backends.ba(poses, disps, *args, p.t0, p.t1, 1, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
_capi.lib().db200_debug_chol_trace(C.c_void_p(0))
f = buf.cpu().numpy()[8 * 1024: 8 * 1024 + 24].astype(np.int64)
print("n =", 6 * (p.t1 - p.t0))
print("load:", f[1] - f[0])
for k in range(8):
    if f[2 + 2 * k] and f[3 + 2 * k]:
        prev = f[1] if k == 0 else f[3 + 2 * (k - 1)]
        print(f"step {k}: panel {f[2 + 2 * k] - prev}  update {f[3 + 2 * k] - f[2 + 2 * k]}")
print("factor total:", f[20] - f[1], " backsub:", f[21] - f[20], " all:", f[21] - f[0])
