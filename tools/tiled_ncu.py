"""one launch of the fused f16 lookup per pyramid layout (row-major, level 0 tiled, levels 0+1 tiled) for an ncu capture"""
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
for mask in (0, 1, 3):
    tl = [db.tile_level(v).contiguous() if (mask >> l) & 1 else v for l, v in enumerate(levels)]
    torch.cuda.synchronize()
    db.corr_pyramid_lookup(tl, coords, 3, tiled_mask=mask)
    torch.cuda.synchronize()
