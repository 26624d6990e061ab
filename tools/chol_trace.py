"""Timeline of the Cholesky panel chain on the c3 problem (debug aid; run on the GPU box)."""
import os, sys
import ctypes as C
import numpy as np, torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends
from droid_b200 import _capi, synth
p = synth.config_c4() if (len(sys.argv) > 1 and sys.argv[1] == 'c4') else synth.config_c3()
dev = torch.device("cuda:0")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
args = [t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
nt = (6 * (p.t1 - p.t0) + 63) // 64
buf = torch.zeros(8 * 1024 + 64, dtype=torch.int64, device=dev)
for it in range(3):
    poses, disps = t(p.poses), t(p.disps)
    if it == 2:
        _capi.lib().db200_debug_chol_trace(C.c_void_p(buf.data_ptr()))
    droid_backends.ba(poses, disps, args[0], args[1], args[2], args[3], args[4], args[5], args[6], p.t0, p.t1, 1, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
_capi.lib().db200_debug_chol_trace(C.c_void_p(0))
tr = buf.cpu().numpy().reshape(-1, 8)[: 2 * nt].astype(np.float64)
t0 = tr[1, 1]
print("tile s: factor_start(us)  factor(us) | consumer of L(s,s): pre-work done after factor start, streamed solve done after factor end | next factor start after that | period")
prev = None
for s in range(nt):
    r = tr[s]
    nxt = tr[s + 1][1] if s + 1 < nt else np.nan
    per = (r[2] - prev) if prev else 0.0
    prev = r[2]
    ntr = (nt - s) * (nt - s + 1) / 2 / 148.0          # trailing tile products per CTA in the step that follows
    print(f"{s:3d}: {(r[1]-t0)/1e3:8.1f} {(r[2]-r[1])/1e3:6.1f} | prework done {(r[6]-r[1])/1e3:6.1f} stream done {(r[5]-r[2])/1e3:6.1f} | {(nxt-r[5])/1e3:6.1f} | {per/1e3:6.1f} | {ntr:5.1f} products/CTA -> {per/1e3/max(ntr,1e-9):5.2f} us each")
xs = tr[nt:2 * nt, 0]
print("factor total us:", (tr[nt - 1, 3] - t0) / 1e3, " backsub us:", (xs[0] - xs[nt - 1]) / 1e3, " backsub per block us:", (xs[0] - xs[nt - 1]) / 1e3 / (nt - 1))

fine = buf.cpu().numpy()[8 * 1024: 8 * 1024 + 33].astype(np.float64)
print("inside the last factor_tile (cycles): per step  [cols->smem+bar, panel, bar, rank8 update]")
for P in range(8):
    a = fine[4 * P: 4 * P + 5]
    nxt = fine[4 * P + 4] if P < 7 else np.nan
    print(P, [a[1] - a[0], a[2] - a[1], a[3] - a[2], nxt - a[3]])

