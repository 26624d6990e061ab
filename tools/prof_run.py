"""One pass over the hot kernels at benchmark sizes, for `ncu` (profiles/README.md gives the command lines).
c3 global-backend BA (one call, 2 iterations), the fused 4-level lookup (96 edges) and the CorrBlock build (48 edges)."""
import os
import sys

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(REPO, "droid-slam_reserch_b200"))
sys.path.insert(0, REPO)
import droid_backends  # noqa: E402
from droid_b200 import synth  # noqa: E402

dev = torch.device("cuda", 0)
what = sys.argv[1:] or ["ba", "lookup", "volume", "volume32", "fd"]
if "ba" in what:
    p = synth.config_c3()
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    args = [t(p.poses), t(p.disps), t(p.intrinsics), t(p.disps_sens), t(p.target), t(p.weight), t(p.eta), t(p.ii), t(p.jj)]
    droid_backends.ba(*args, p.t0, p.t1, 2, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
if "lookup" in what:
    N, ht, wd = 96, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(7)
    levels = [torch.randn((N, ht, wd, ht >> l, wd >> l), device=dev, dtype=torch.float16, generator=g) for l in range(4)]
    yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32), indexing="ij")
    coords = (torch.stack([xx, yy], 0)[None].repeat(N, 1, 1, 1) + 6.0 * torch.randn((N, 2, ht, wd), device=dev, generator=g)).contiguous()
    for _ in range(2):
        droid_backends.corr_pyramid_lookup(levels, coords, 3)
    torch.cuda.synchronize()
    del levels
if "volume" in what:
    N, C_, ht, wd = 48, 128, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(11)
    f1 = torch.randn((N, C_, ht, wd), device=dev, dtype=torch.float16, generator=g)
    f2 = torch.randn((N, C_, ht, wd), device=dev, dtype=torch.float16, generator=g)
    for _ in range(2):
        droid_backends.corr_volume_pyramid(f1, f2, 4)
    torch.cuda.synchronize()
if "volume32" in what:
    N, C_, ht, wd = 24, 128, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(12)
    f1 = torch.randn((N, C_, ht, wd), device=dev, generator=g)
    f2 = torch.randn((N, C_, ht, wd), device=dev, generator=g)
    for _ in range(2):
        droid_backends.corr_volume_pyramid(f1, f2, 4)
    torch.cuda.synchronize()
if "fd" in what:
    p = synth.config_c3(n_edges=2000)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    n = p.t1
    ii, jj = torch.meshgrid(torch.arange(n), torch.arange(n), indexing="ij")
    for _ in range(2):
        droid_backends.frame_distance(t(p.poses), t(p.disps), t(p.intrinsics), ii.reshape(-1).to(dev), jj.reshape(-1).to(dev), 0.25)
    torch.cuda.synchronize()

if "altcorr" in what:       # BASELINE config-5 shape: 88 edges, fp32 operands, all four levels through the nine-name API
    N, C_, ht, wd = 88, 128, 48, 64
    g = torch.Generator(device=dev); g.manual_seed(21)
    f1 = torch.randn((N, ht, wd, C_), device=dev, generator=g)
    pyr = [torch.randn((N, ht >> l, wd >> l, C_), device=dev, generator=g) for l in range(4)]
    yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32), indexing="ij")
    co = (torch.stack([xx, yy], -1)[None, None].repeat(N, 1, 1, 1, 1) + 4.0 * torch.randn((N, 1, ht, wd, 2), device=dev, generator=g)).contiguous()
    ga = torch.randn((N, 1, 49, ht, wd), device=dev, generator=g)
    for _ in range(2):
        for l in range(4):
            droid_backends.altcorr_forward(f1, pyr[l], (co / 2 ** l).contiguous(), 3)
        droid_backends.altcorr_backward(f1, pyr[0], co, ga, 3)
    torch.cuda.synchronize()
    print("prof_run altcorr done")
if "altcorr16" in what:     # the fused backend lookup as bench.py's `altcorr` block runs it: f16 stereo buffer, one chunk of ~94 edges
    sys.path.insert(0, os.path.join(REPO, "tests"))
    import dropin_cases as D
    from droid_b200 import corr as mirror
    n, rig, ht, wd, E = 64, 2, 48, 64, int(os.environ.get("ALT16_EDGES", "94"))
    g = torch.Generator(device=dev); g.manual_seed(33)
    fmaps = torch.randn((n, rig, 128, ht, wd), device=dev, dtype=torch.float16, generator=g)
    op = mirror.AltCorrBlock(fmaps.view(1, n * rig, 128, ht, wd))
    rng = np.random.default_rng(5)
    co = torch.from_numpy(D.smooth_coords(rng, E, ht, wd, wild=0.0)).to(dev)[None]
    ii = torch.from_numpy(rng.integers(0, n, E)).to(dev); jj = torch.from_numpy(rng.integers(0, n, E)).to(dev)
    for _ in range(3):
        out = op(co, rig * ii, rig * jj + (ii == jj).long())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = op(co, rig * ii, rig * jj + (ii == jj).long())
    e1.record(); torch.cuda.synchronize()
    print(f"prof_run altcorr16: {E} edges x 4 levels fused f16: {e0.elapsed_time(e1) / 10:.3f} ms per call")
if "altcorr16" in what and os.environ.get("DB200_ALT_DBG", "0") not in ("", "0") and int(os.environ["DB200_ALT_DBG"]) & 16:
    droid_backends.WAIT_TIMES = []
    out = op(co, rig * ii, rig * jj + (ii == jj).long())
    w = droid_backends.WAIT_TIMES[-1]
    ctas = 148
    names = ["prod: plan slot", "prod: A stage free", "prod: B stage free", "prod: TOTAL", "mma: plan", "mma: A landed", "mma: accumulator free",
             "mma: B landed", "mma: TOTAL", "epi: plan", "epi: accumulator ready", "epi: TOTAL"]
    for n_, v_ in zip(names, w):
        print(f"   {n_:24s} {v_ / ctas / 1.9e3:9.1f} us per CTA")
print("prof_run ok", what)
