"""Where a backend lookup step (bench.py altcorr_bench) spends its time: (a) driven as the reference drives it (boolean
masks per chunk: three host syncs per chunk), (b) chunk index lists precomputed (no host sync), (c) the C-ABI calls alone
on preallocated outputs.   gpurun -- python tools/altcorr_step_breakdown.py"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200"), os.path.join(REPO, "tests")]
import bench  # noqa: E402
import droid_backends  # noqa: E402
from droid_b200 import _capi, corr as mirror, synth  # noqa: E402

dev = torch.device("cuda:0")
p = bench.make_problem("c3")
n, rig = int(p.disps.shape[0]), 2
_, ht, wd = p.disps.shape
g = torch.Generator(device=dev); g.manual_seed(33)
fmaps = torch.randn((n, rig, 128, ht, wd), device=dev, dtype=torch.float16, generator=g)
co, _ = synth.reproject(p.poses.astype(np.float64), p.disps.astype(np.float64), p.intrinsics, p.ii, p.jj)
coords1 = torch.from_numpy(np.nan_to_num(co, nan=0.0, posinf=1e4, neginf=-1e4).astype(np.float32)).to(dev)[None]
ii_t, jj_t = torch.from_numpy(p.ii).to(dev), torch.from_numpy(p.jj).to(dev)
masks = [v for v in ((ii_t >= i) & (ii_t < i + 8) for i in range(0, n, 8)) if bool(v.any())]
idxs = [torch.nonzero(v).reshape(-1) for v in masks]
op = mirror.AltCorrBlock(fmaps.view(1, n * rig, 128, ht, wd))


def step_masks():
    for v in masks:
        iis, jjs = ii_t[v], jj_t[v]
        op(coords1[:, v], rig * iis, rig * jjs + (iis == jjs).long())


def step_idx():
    for ix in idxs:
        iis, jjs = ii_t[ix], jj_t[ix]
        op(coords1[:, ix], rig * iis, rig * jjs + (iis == jjs).long())


pre = []
for ix in idxs:
    iis, jjs = ii_t[ix], jj_t[ix]
    pre.append((coords1[0, ix].contiguous(), (rig * iis).contiguous(), (rig * jjs + (iis == jjs).long()).contiguous()))


def step_pre():
    for c, a, b in pre:
        droid_backends.altcorr_pyramid_forward(op._hi, a, b, c, 3)


lib = _capi.lib()
Emax = max(len(ix) for ix in idxs)
out = torch.empty((Emax, 196, ht, wd), device=dev)
wsb = lib.db200_altcorr_pyramid_workspace_bytes(Emax, 1, ht, wd)
ws = torch.empty(wsb, dtype=torch.uint8, device=dev)
arr = (C.c_void_p * 4)(*[t.data_ptr() for t in op._hi])
null = C.POINTER(C.c_void_p)()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def step_capi():
    H, W, S = ht, wd, 1
    for c, a, b in pre:
        E = c.shape[0]
        rc = lib.db200_altcorr_pyramid_forward(C.c_void_p(op._hi[0].data_ptr()), C.c_void_p(0), arr, null, C.c_void_p(0), C.c_void_p(0), null,
                                               n * rig, H, W, 128, 4, C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(c.data_ptr()),
                                               H * W * 2, 2, 2, C.c_void_p(out.data_ptr()), 196 * H * W, 1, H * W, 1, E, 1, 3,
                                               C.c_void_p(ws.data_ptr()), wsb, st)
        assert rc == 0


for name, fn in (("as the reference drives it (boolean masks)", step_masks), ("precomputed index lists", step_idx),
                 ("precomputed chunks, shim call only", step_pre), ("C-ABI calls on preallocated output", step_capi)):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(5):
        fn()
    e1.record()
    t1 = time.perf_counter()           # host time to ISSUE the work
    torch.cuda.synchronize()
    print(f"{name:48s} {e0.elapsed_time(e1) / 5:8.3f} ms/step on the device, host issue {1e3 * (t1 - t0) / 5:8.3f} ms/step")
