"""CorrBlock build (48 edges x 48x64, f16): one call with the tiled (fused-pyramid) layout and one row-major, for an ncu capture"""
import os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [REPO, os.path.join(REPO, "droid-slam_reserch_b200")]
import torch
import droid_backends as db
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(11)
f1 = torch.randn((48, 128, 48, 64), device=dev, dtype=torch.float16, generator=g)
f2 = torch.randn((48, 128, 48, 64), device=dev, dtype=torch.float16, generator=g)
for tiled in (True, False):
    db.corr_volume_pyramid(f1, f2, 4, tiled=tiled)
    torch.cuda.synchronize()
