"""Compares the reduced pose system [H | b] that the accumulate stage builds on the GPU (fp64, packed lower triangle,
db200_ba_system_ptr) with the CPU oracle's (dump_system, built from fp32 kernels restated in fp32 — and again in fp64 by
solving), block class by block class: which blocks differ tells a wrong term from round-off.
    gpurun -- python tools/parity_system.py [n_kf hub:deg hub:deg ...]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import droid_backends  # noqa: E402,F401
from droid_b200 import _capi, synth  # noqa: E402
from oracle import oracle  # noqa: E402

n_kf = int(sys.argv[1]) if len(sys.argv) > 1 else 80
hubs = {int(a.split(":")[0]): int(a.split(":")[1]) for a in sys.argv[2:]} or {11: 64, 0: 35}
ii, jj = synth.hub_edges(n_kf, hubs, np.random.default_rng(900 + n_kf))
p = synth.make_ba_problem(ii, jj, n_kf, ht=16, wd=24, seed=900 + n_kf)
LM, EP = 1e-5, 1e-2
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
lib = _capi.lib()
N, (B, ht, wd) = p.n_edges, p.disps.shape
M = p.eta.shape[0]
wsb = lib.db200_ba_workspace_bytes(N, B, ht, wd, p.t0, p.t1, M, 0)
ws = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
g = dict(poses=t(p.poses), disps=t(p.disps), intr=t(p.intrinsics), sens=t(p.disps_sens), tg=t(p.target), wg=t(p.weight), eta=t(p.eta),
         ii=t(p.ii), jj=t(p.jj))
P_ = lambda x: C.c_void_p(x.data_ptr())
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
rc = lib.db200_ba_accumulate(P_(g["poses"]), P_(g["disps"]), P_(g["intr"]), P_(g["sens"]), P_(g["tg"]), P_(g["wg"]), P_(g["eta"]),
                             P_(g["ii"]), P_(g["jj"]), N, B, ht, wd, p.t0, p.t1, 0, M, 1, 0, B, P_(ws), wsb, st)
torch.cuda.synchronize()
assert rc == 0 and lib.db200_ba_status(P_(ws), st) == 0
n = 6 * (p.t1 - p.t0)
ptr = lib.db200_ba_system_ptr(P_(ws), N, B, ht, wd, p.t0, p.t1, M, 0)
off = ptr - ws.data_ptr()
hb = ws[off:off + 8 * (n * (n + 1) // 2 + n)].view(torch.float64).cpu().numpy()
H = np.zeros((n, n))
H[np.tril_indices(n)] = hb[:n * (n + 1) // 2]
H = H + np.tril(H, -1).T
b = hb[n * (n + 1) // 2:]
a = (p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, LM, EP)
r32 = oracle.ba(*a, dump_system=True)
r64 = oracle.ba(*a, dtype=np.float64)
Ho, bo = r32["H"], r32["b"]
Ho = np.tril(Ho) + np.tril(Ho, -1).T
rel = lambda x, y: float(np.linalg.norm(x - y) / max(np.linalg.norm(y), 1e-300))
print(f"problem: {n_kf} keyframes, hubs {hubs}, {N} edges; system n = {n}")
print(f"whole H: rel {rel(H, Ho):.3e}   b: rel {rel(b, bo):.3e}")
Pn = p.t1 - p.t0
blk = lambda A, i, j: A[6 * i:6 * i + 6, 6 * j:6 * j + 6]
hub_rows = {h - p.t0 for h in hubs if h >= p.t0}
partners = {h: sorted(set(jj[(ii == h) & (jj != h)].tolist())) for h in hubs}
cls = {}
for i in range(Pn):
    for j in range(i + 1):
        d, o = blk(H, i, j), blk(Ho, i, j)
        if not o.any() and not d.any():
            continue
        if i == j:
            name = "diag(hub)" if i in hub_rows else "diag(other)"
        elif i in hub_rows or j in hub_rows:
            name = "hub x partner"
        else:
            name = "other x other"
            for h, ps in partners.items():
                if (i + p.t0) in ps and (j + p.t0) in ps:
                    si, sj = ps.index(i + p.t0), ps.index(j + p.t0)       # slot order = edge order of the hub's out-edges? (approx.)
                    name = f"partner x partner of hub {h}"
        e = cls.setdefault(name, [0.0, 0.0, 0])
        e[0] += float(((d - o) ** 2).sum()); e[1] += float((o ** 2).sum()); e[2] += 1
for k, (num, den, cnt) in sorted(cls.items()):
    print(f"  {k:38s} blocks {cnt:5d}  rel {np.sqrt(num / max(den, 1e-300)):.3e}")
# what each system gives as dx (numpy fp64 solve with the damping of droid_kernels.cu:1197)
def solve(Hm, bv):
    A = Hm.copy()
    dg = np.diag(A).copy()
    A[np.diag_indices(n)] = dg + EP + LM * dg
    return np.linalg.solve(A, bv)
dx_g, dx_o = solve(H, b), solve(Ho, bo)
print(f"dx from GPU system vs dx from oracle-f32 system: {rel(dx_g, dx_o):.3e};  vs oracle f64 dx: gpu {rel(dx_g, r64['dx'].reshape(-1)):.3e} "
      f"oracle-f32 system {rel(dx_o, r64['dx'].reshape(-1)):.3e}; oracle f32 own dx {rel(r32['dx'].reshape(-1), r64['dx'].reshape(-1)):.3e}")
print(f"cond(H damped) ~ {np.linalg.cond(H + np.diag(EP + LM * np.diag(H))):.3e}")
