"""BA / distance entry points of the keyframe buffer with the reference's signatures
(DepthVideo.cuda_ba / DepthVideo.distance, droid_slam/depth_video.py:160-204).  Holds only the tensors the
hot path touches; the rest of DepthVideo (images, nets, multiprocessing lock) stays with the caller."""
from __future__ import annotations

import ctypes as C

import torch

import droid_backends
from . import _capi


class BAVideo:
    def __init__(self, poses, disps, intrinsics, disps_sens=None, counter=None):
        self.poses, self.disps, self.intrinsics = poses, disps, intrinsics
        self.disps_sens = disps_sens if disps_sens is not None else torch.zeros_like(disps)
        self.counter = counter if counter is not None else poses.shape[0]

    @staticmethod
    def format_indicies(ii, jj, device):
        ii = torch.as_tensor(ii).to(device=device, dtype=torch.long).reshape(-1)
        jj = torch.as_tensor(jj).to(device=device, dtype=torch.long).reshape(-1)
        return ii, jj

    def distance(self, ii=None, jj=None, beta=0.3, bidirectional=True):
        """depth_video.py:160-190"""
        ret_matrix = ii is None
        if ret_matrix:
            N = self.counter
            ii, jj = torch.meshgrid(torch.arange(N), torch.arange(N), indexing="ij")
        ii, jj = self.format_indicies(ii, jj, self.poses.device)
        intr = self.intrinsics[0] if self.intrinsics.dim() == 2 else self.intrinsics
        if bidirectional:
            poses = self.poses[:self.counter].clone()
            d1 = droid_backends.frame_distance(poses, self.disps, intr, ii, jj, beta)
            d2 = droid_backends.frame_distance(poses, self.disps, intr, jj, ii, beta)
            d = .5 * (d1 + d2)
        else:
            d = droid_backends.frame_distance(self.poses, self.disps, intr, ii, jj, beta)
        return d.reshape(N, N) if ret_matrix else d

    def reproject(self, ii, jj, target=None):
        """depth_video.py:150-158: coords[1,N,ht,wd,2], valid[1,N,ht,wd,1] of projecting frame ii's points into jj.
        With `target` ([1,N,ht,wd,2] or [N,ht,wd,2]) the motion features of FactorGraph.update
        (factor_graph.py:203-205) come out of the same kernel: also returns motn[1,N,4,ht,wd], already clamped."""
        ii, jj = self.format_indicies(ii, jj, self.poses.device)
        B, ht, wd = self.disps.shape
        N = int(ii.shape[0])
        dev = self.poses.device
        coords = torch.empty((1, N, ht, wd, 2), dtype=torch.float32, device=dev)
        valid = torch.empty((1, N, ht, wd, 1), dtype=torch.float32, device=dev)
        motn = None
        vp = lambda x: C.c_void_p(x.data_ptr()) if x is not None and x.numel() > 0 else C.c_void_p(0)
        if target is not None:
            target = target.reshape(N, ht, wd, 2).float().contiguous()
            motn = torch.empty((1, N, 4, ht, wd), dtype=torch.float32, device=dev)
        intr = self.intrinsics.contiguous()
        per_frame = 1 if intr.dim() == 2 else 0
        _capi.check(_capi.lib().db200_reproject(vp(self.poses), vp(self.disps), vp(intr), per_frame, vp(ii), vp(jj), N,
                                                self.poses.shape[0], B, ht, wd, vp(coords), vp(valid), vp(target), vp(motn),
                                                C.c_void_p(torch.cuda.current_stream().cuda_stream)), "reproject")
        return (coords, valid) if motn is None else (coords, valid, motn)

    def cuda_ba(self, target, weight, eta, ii, jj, t0=1, t1=None, itrs=2, lm=1e-4, ep=0.1, motion_only=False):
        """depth_video.py:192-204"""
        if t1 is None:
            t1 = max(ii.max().item(), jj.max().item()) + 1
        intr = self.intrinsics[0] if self.intrinsics.dim() == 2 else self.intrinsics
        out = droid_backends.ba(self.poses, self.disps, intr, self.disps_sens, target, weight, eta, ii, jj, t0, t1, itrs,
                                lm, ep, motion_only)
        self.disps.clamp_(min=0.001)
        return out
