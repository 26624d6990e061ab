"""corr_index_backward over the 4 levels of 88 fp32 edges (BASELINE config 5 shape) and of 96 fp16 edges"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import torch
import droid_backends as db
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(3)
ht, wd = 48, 64
for dt, N in ((torch.float32, 88), (torch.float16, 96)):
    levels = [torch.empty((N, ht, wd, ht >> l, wd >> l), device=dev, dtype=dt) for l in range(4)]
    yy, xx = torch.meshgrid(torch.arange(ht, device=dev, dtype=torch.float32), torch.arange(wd, device=dev, dtype=torch.float32), indexing="ij")
    coords = (torch.stack([xx, yy], 0)[None].repeat(N, 1, 1, 1) + 6.0 * torch.randn((N, 2, ht, wd), device=dev, generator=g)).contiguous()
    cl = [(coords / 2 ** l).contiguous() for l in range(4)]
    grads = [torch.randn((N, 7, 7, ht, wd), device=dev, generator=g).to(dt) for _ in range(4)]
    fn = lambda: [db.corr_index_backward(levels[l], cl[l], grads[l], 3) for l in range(4)]
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    nbytes = sum(v.numel() * v.element_size() for v in levels)
    print(f"{dt}: {N} edges x 4 levels backward {ms:.3f} ms, {nbytes / ms / 1e6:.0f} GB/s written")
    del levels, grads
