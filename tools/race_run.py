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
if len(sys.argv) > 2 and sys.argv[2] == "extra":
    # the kernels added in round 2, on small cases: high-degree BA path (hub of degree 33 -> depth pre-pass + 6 work items),
    # altcorr forward / backward on tcgen05 (2 edges, 16x32 maps, fp32 operands -> 3 products; one tile row ragged)
    ii, jj = synth.hub_edges(40, {7: 33}, np.random.default_rng(1))
    q = synth.make_ba_problem(ii, jj, 40, ht=16, wd=24, seed=2)
    poses, disps = t(q.poses), t(q.disps)
    dx, _ = droid_backends.ba(poses, disps, t(q.intrinsics), t(q.disps_sens), t(q.target), t(q.weight), t(q.eta), t(q.ii), t(q.jj),
                              q.t0, q.t1, 1, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
    print("race_run extra: high-degree ba ok, dx norm", float(dx.norm()))
    g = torch.Generator(device=dev); g.manual_seed(3)
    B, H, W = 2, 12, 32
    f1 = torch.randn((B, H, W, 128), device=dev, generator=g); f2 = torch.randn((B, H, W, 128), device=dev, generator=g)
    yy, xx = torch.meshgrid(torch.arange(H, device=dev, dtype=torch.float32), torch.arange(W, device=dev, dtype=torch.float32), indexing="ij")
    co = (torch.stack([xx, yy], -1)[None, None].repeat(B, 1, 1, 1, 1) + 3.0 * torch.randn((B, 1, H, W, 2), device=dev, generator=g)).contiguous()
    out, = droid_backends.altcorr_forward(f1, f2, co, 3)
    g1, g2, _ = droid_backends.altcorr_backward(f1, f2, co, torch.randn((B, 1, 49, H, W), device=dev, generator=g), 3)
    torch.cuda.synchronize()
    print("race_run extra: altcorr ok", float(out.norm()), float(g1.norm()), float(g2.norm()))
