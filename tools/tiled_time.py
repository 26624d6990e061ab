"""Frontend lookup (96 edges x 48x64, 4 levels, f16) on the row-major pyramid and on the tiled one; CorrBlock build both ways."""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import torch
import droid_backends as db

dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(7)
N, ht, wd = 96, 48, 64
levels = [torch.randn((N, ht, wd, ht >> l, wd >> l), device=dev, dtype=torch.float16, generator=g) for l in range(4)]
yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32), indexing="ij")
coords = (torch.stack([xx, yy], 0)[None].repeat(N, 1, 1, 1) + 6.0 * torch.randn((N, 2, ht, wd), device=dev, generator=g)).contiguous()


def timed(fn, n=20, warm=3):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


import ctypes as C
from droid_b200 import _capi
lib = _capi.lib()
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for mask in (0, 1, 3):
    tl = [db.tile_level(v).contiguous() if (mask >> l) & 1 else v for l, v in enumerate(levels)]
    a = db.corr_pyramid_lookup(levels, coords, 3)
    b = db.corr_pyramid_lookup(tl, coords, 3, tiled_mask=mask)
    ms = timed(lambda: db.corr_pyramid_lookup(tl, coords, 3, tiled_mask=mask))
    arr = (C.c_void_p * 4)(*[v.data_ptr() for v in tl])
    out = torch.empty_like(b)
    raw = lambda: lib.db200_corr_pyramid_lookup_tiled(arr, C.c_void_p(coords.data_ptr()), C.c_void_p(out.data_ptr()), N, ht, wd, ht, wd, 4, 3, 1, mask, st)
    ms_raw = timed(raw, 50, 5)
    print(f"lookup mask {mask:03b}: wrapper {ms:.4f} ms, C-ABI loop {ms_raw:.4f} ms  equal={torch.equal(a, b)}  algorithmic frac {N*ht*wd*912/(ms_raw*1e-3)/1e9/6527.1:.3f}", flush=True)
    del tl
del levels
f1 = torch.randn((48, 128, ht, wd), device=dev, dtype=torch.float16, generator=g)
f2 = torch.randn((48, 128, ht, wd), device=dev, dtype=torch.float16, generator=g)
print("build row-major", timed(lambda: db.corr_volume_pyramid(f1, f2, 4), 10), "ms; tiled", timed(lambda: db.corr_volume_pyramid(f1, f2, 4, tiled=True), 10), "ms")
