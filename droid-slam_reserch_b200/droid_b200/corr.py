"""CorrBlock / AltCorrBlock with the reference's public interface (droid_slam/modules/corr.py:22-139),
running on the sm_100a kernels: one GEMM launch per pyramid level writes every level once
(corr_volume_pyramid), and one launch produces all 4x49 lookup channels already concatenated
(corr_pyramid_lookup) instead of 4 launches + torch.cat."""
from __future__ import annotations

import torch
import torch.nn.functional as F

import droid_backends


class CorrSampler(torch.autograd.Function):
    """corr.py:6-20 — single-level lookup with autograd (training path)."""

    @staticmethod
    def forward(ctx, volume, coords, radius):
        ctx.save_for_backward(volume, coords)
        ctx.radius = radius
        corr, = droid_backends.corr_index_forward(volume, coords, radius)
        return corr

    @staticmethod
    def backward(ctx, grad_output):
        volume, coords = ctx.saved_tensors
        grad_volume, = droid_backends.corr_index_backward(volume, coords, grad_output.contiguous(), ctx.radius)
        return grad_volume, None, None


class _VolumePyramid(torch.autograd.Function):
    """corr.py:30-38,63-71 as ONE op: every pyramid level comes out of the fused GEMM (level l = f1^T pool_l(f2) / 16).
    backward: with G_l = d loss / d level_l [N, HW, HW_l],
        d f1 = 1/16 sum_l pool_l(f2) G_l^T,     d pool_l(f2) = 1/16 f1 G_l,   then the adjoint of the 2x2 mean chain —
    four small library GEMM pairs instead of autograd through a materialised [N,HW,HW] product and three pooling passes."""

    @staticmethod
    def forward(ctx, f1, f2, num_levels):
        ctx.save_for_backward(f1, f2)
        ctx.num_levels = num_levels
        return tuple(droid_backends.corr_volume_pyramid(f1, f2, num_levels))

    @staticmethod
    def backward(ctx, *grads):
        f1, f2 = ctx.saved_tensors
        N, C, ht, wd = f1.shape
        a = f1.float().reshape(N, C, ht * wd)
        pyr = [f2.float()]
        for _ in range(1, ctx.num_levels):
            pyr.append(F.avg_pool2d(pyr[-1], 2, stride=2))
        g1 = torch.zeros_like(a)
        gp = []
        for l, g in enumerate(grads):
            if g is None:
                gp.append(None)
                continue
            G = g.float().reshape(N, ht * wd, -1)
            g1 += torch.matmul(pyr[l].reshape(N, C, -1), G.transpose(1, 2))
            gp.append(torch.matmul(a, G).reshape(pyr[l].shape))
        g2 = None
        for l in range(ctx.num_levels - 1, -1, -1):                  # adjoint of avg_pool2d(2, stride 2), floor mode
            if g2 is not None:
                up = g2.repeat_interleave(2, dim=2).repeat_interleave(2, dim=3) * 0.25
                g2 = F.pad(up, (0, pyr[l].shape[3] - up.shape[3], 0, pyr[l].shape[2] - up.shape[2]))
                if gp[l] is not None:
                    g2 = g2 + gp[l]
            else:
                g2 = gp[l]
        g1 = (g1 / 16.0).reshape(f1.shape).to(f1.dtype)
        g2 = None if g2 is None else (g2 / 16.0).to(f2.dtype)
        return g1, g2, None


class CorrBlock:
    """corr.py:22-61.  Inference (no autograd, radius 3, fp16 128-channel maps): the volumes are PRIVATE to the block, so the
    wide levels are kept in the tiled slice layout of include/droid_b200.h ("tiled levels": 4x8-element tiles, one DRAM burst
    each), which the fused lookup reads with ~35 % fewer DRAM bytes.  `corr_pyramid` still hands out the reference's row-major
    volumes: reading it converts the block to row-major once (a copy) and the block stays that way."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=3):
        self.num_levels, self.radius = num_levels, radius
        batch, num, dim, ht, wd = fmap1.shape
        f1 = fmap1.reshape(batch * num, dim, ht, wd).contiguous()
        f2 = fmap2.reshape(batch * num, dim, ht, wd).contiguous()
        self._tiled = 0                                              # bit l: level l is stored tiled
        if torch.is_grad_enabled() and (f1.requires_grad or f2.requires_grad):
            self._pyr = list(_VolumePyramid.apply(f1, f2, num_levels))
        elif radius == 3:
            self._pyr, self._tiled = droid_backends.corr_volume_pyramid(f1, f2, num_levels, tiled=True)
        else:
            self._pyr = droid_backends.corr_volume_pyramid(f1, f2, num_levels)

    def _row_major(self):
        if self._tiled:
            self._pyr = [droid_backends.untile_level(v).contiguous() if (self._tiled >> l) & 1 else v
                         for l, v in enumerate(self._pyr)]
            self._tiled = 0

    @property
    def corr_pyramid(self):
        self._row_major()
        return self._pyr

    @corr_pyramid.setter
    def corr_pyramid(self, levels):
        self._pyr, self._tiled = list(levels), 0

    @staticmethod
    def corr(fmap1, fmap2):
        """all-pairs correlation (corr.py:63-71) -> [batch, num, ht, wd, ht, wd]"""
        batch, num, dim, ht, wd = fmap1.shape
        vol, = droid_backends.corr_volume_pyramid(fmap1.reshape(batch * num, dim, ht, wd).contiguous(),
                                                  fmap2.reshape(batch * num, dim, ht, wd).contiguous(), 1)
        return vol.view(batch, num, ht, wd, ht, wd)

    def __call__(self, coords):
        batch, num, ht, wd, _ = coords.shape
        c = coords.permute(0, 1, 4, 2, 3).contiguous().view(batch * num, 2, ht, wd)
        if torch.is_grad_enabled() and any(v.requires_grad for v in self._pyr):
            nch = (2 * self.radius + 1) ** 2
            outs = [CorrSampler.apply(self.corr_pyramid[i], c / 2 ** i, self.radius).view(batch, num, nch, ht, wd)
                    for i in range(self.num_levels)]
            return torch.cat(outs, dim=2)
        out = droid_backends.corr_pyramid_lookup(self._pyr, c, self.radius, tiled_mask=self._tiled)
        return out.view(batch, num, self.num_levels * (2 * self.radius + 1) ** 2, ht, wd)     # explicit: an empty batch has no -1

    def cat(self, other):
        if self._tiled != other._tiled:                              # e.g. one of the two was built under autograd
            self._row_major()
            other._row_major()
        for i in range(self.num_levels):                             # edges are the leading axis: the slice layout does not matter
            self._pyr[i] = torch.cat([self._pyr[i], other._pyr[i]], 0)
        return self

    def __getitem__(self, index):
        for i in range(self.num_levels):
            self._pyr[i] = self._pyr[i][index].contiguous()
        return self


class CorrLayer(torch.autograd.Function):
    """corr.py:74-88"""

    @staticmethod
    def forward(ctx, fmap1, fmap2, coords, r):
        ctx.r = r
        ctx.save_for_backward(fmap1, fmap2, coords)
        corr, = droid_backends.altcorr_forward(fmap1, fmap2, coords, r)
        return corr

    @staticmethod
    def backward(ctx, grad_corr):
        fmap1, fmap2, coords = ctx.saved_tensors
        g1, g2, gc = droid_backends.altcorr_backward(fmap1, fmap2, coords, grad_corr.contiguous(), ctx.r)
        return g1, g2, gc, None


class AltCorrBlock:
    """corr.py:91-139 with the hot call re-designed: the reference gathers `pyramid[0][:, ii]` / `pyramid[i][:, jj]`, casts
    both to fp32 and launches one kernel per level for every chunk of edges of every update step
    (factor_graph.py:261-290).  Here the channels-last pyramid is kept in the tensor-core operand format once (fp16 as it
    comes from the feature encoder; fp32 features as an fp16 hi + lo split) and `corr_fn` is ONE launch that takes the
    frame indices: no gathers, no casts, no concat (`droid_backends.altcorr_pyramid_forward`, csrc/altcorr_tc.cu).
    With autograd (fmaps that require grad) the per-level `CorrLayer` path of the reference is used."""

    def __init__(self, fmaps, num_levels=4, radius=3):
        self.num_levels, self.radius = num_levels, radius
        B, N, C, H, W = fmaps.shape
        x = fmaps.view(B * N, C, H, W) / 4.0
        self.pyramid = []
        for i in range(num_levels):
            self.pyramid.append(x.permute(0, 2, 3, 1).contiguous().view(B, N, H // 2 ** i, W // 2 ** i, C))
            x = F.avg_pool2d(x, 2, stride=2)
        self._dims = (B, N, C, H, W)
        self._fused = (C == 128 and radius == 3 and fmaps.is_cuda and fmaps.dtype in (torch.float16, torch.float32)
                       and not (torch.is_grad_enabled() and fmaps.requires_grad))
        self._lo = self._flags = self._f32 = None
        if self._fused:
            flat = [p.view((B * N,) + p.shape[2:]) for p in self.pyramid]
            if fmaps.dtype == torch.float16:
                self._hi = [droid_backends.to_channel_blocked(p) for p in flat]     # once per block: the kernels' operand layout
            else:
                parts = [droid_backends.split_hilo_f16(p) for p in flat]
                self._hi, self._lo = [q[0] for q in parts], [q[1] for q in parts]
                self._flags = parts[0][2]
                for q in parts[1:]:
                    self._flags |= q[2]
                self._f32 = flat

    def corr_fn(self, coords, ii, jj):
        B, N, H, W, S, _ = coords.shape
        if self._fused and not (torch.is_grad_enabled() and coords.requires_grad):
            nf = self._dims[1]
            ii, jj = ii.to(torch.int64), jj.to(torch.int64)
            if B > 1:
                off = (torch.arange(B, device=ii.device) * nf)[:, None]
                ii, jj = (ii[None] + off).reshape(-1), (jj[None] + off).reshape(-1)
            out = droid_backends.altcorr_pyramid_forward(self._hi, ii.contiguous(), jj.contiguous(),
                                                         coords.reshape(B * N, H, W, S, 2).contiguous(), self.radius,
                                                         lo=self._lo, flags=self._flags, f32=self._f32)
            return out.view(B, N, 49 * self.num_levels, H, W, S)
        coords = coords.permute(0, 1, 4, 2, 3, 5)
        out = []
        for i in range(self.num_levels):
            f1 = self.pyramid[0][:, ii]
            f2 = self.pyramid[i][:, jj]
            ci = (coords / 2 ** i).reshape(B * N, S, H, W, 2).contiguous()
            f1 = f1.reshape((B * N,) + f1.shape[2:]).float().contiguous()
            f2 = f2.reshape((B * N,) + f2.shape[2:]).float().contiguous()
            corr = CorrLayer.apply(f1, f2, ci, self.radius)
            out.append(corr.view(B, N, S, -1, H, W).permute(0, 1, 3, 4, 5, 2))
        return torch.cat(out, dim=2)

    def __call__(self, coords, ii, jj):
        squeeze = coords.dim() == 5
        if squeeze:
            coords = coords.unsqueeze(dim=-2)
        corr = self.corr_fn(coords, ii, jj)
        if squeeze:
            corr = corr.squeeze(dim=-1)
        return corr.contiguous()
