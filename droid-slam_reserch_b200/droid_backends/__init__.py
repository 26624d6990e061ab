"""Drop-in replacement of the reference's pybind extension `droid_backends`
(/root/reference/src/droid.cpp:237-250): same nine names, same argument order, same return types
(lists/tuples of freshly allocated tensors; `ba` mutates `poses` and `disps` in place).

Host code only allocates outputs and forwards raw device pointers + the current CUDA stream to the
C-ABI in libdroid_b200.so (include/droid_b200.h).  No CPU fallback exists: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from droid_b200 import _capi

__all__ = ["ba", "frame_distance", "projmap", "depth_filter", "iproj", "altcorr_forward", "altcorr_backward",
           "corr_index_forward", "corr_index_backward", "corr_pyramid_lookup", "corr_volume_pyramid"]

_DT = {torch.float32: _capi.F32, torch.float16: _capi.F16, torch.float64: _capi.F64}
_SYNC_STATUS = os.environ.get("DROID_B200_ASYNC", "0") != "1"


def _chk(*tensors):
    for name, t in tensors:
        if not t.is_contiguous():  # droid.cpp:84-85
            raise RuntimeError(f"{name} must be contiguous")
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor: droid_backends (B200) has no CPU path")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


# ------------------------------------------------------------------------------------------------ lookups
def corr_index_forward(volume, coords, radius):
    """droid.cpp:170-178.  volume[N,h1,w1,h2,w2], coords[N,2,h1,w1] -> [corr[N,2r+1,2r+1,h1,w1]]"""
    _chk(("volume", volume), ("coords", coords))
    N, h1, w1, h2, w2 = volume.shape
    rd = 2 * radius + 1
    corr = torch.empty((N, rd, rd, h1, w1), dtype=volume.dtype, device=volume.device)
    _capi.check(_capi.lib().db200_corr_index_forward(_ptr(volume), _ptr(coords.float()), _ptr(corr), N, h1, w1, h2, w2,
                                                     radius, _DT[volume.dtype], _stream()), "corr_index_forward")
    return [corr]


def corr_index_backward(volume, coords, corr_grad, radius):
    """droid.cpp:180-191.  -> [volume_grad like volume]"""
    _chk(("volume", volume), ("coords", coords), ("corr_grad", corr_grad))
    N, h1, w1, h2, w2 = volume.shape
    grad = torch.empty_like(volume)
    g = corr_grad.to(volume.dtype)
    _capi.check(_capi.lib().db200_corr_index_backward(_ptr(coords.float()), _ptr(g), _ptr(grad), N, h1, w1, h2, w2, radius,
                                                      _DT[volume.dtype], _stream()), "corr_index_backward")
    return [grad]


def corr_pyramid_lookup(levels, coords, radius):
    """Fused CorrBlock.__call__ (corr.py:40-50): levels = list of [N,h1,w1,h2>>l,w2>>l]; coords[N,2,h1,w1]
    at level-0 scale -> out[N, L*(2r+1)^2, h1, w1]."""
    _chk(("coords", coords), *[(f"levels[{i}]", v) for i, v in enumerate(levels)])
    N, h1, w1, h2, w2 = levels[0].shape
    L = len(levels)
    rd = 2 * radius + 1
    out = torch.empty((N, L * rd * rd, h1, w1), dtype=levels[0].dtype, device=coords.device)
    arr = (C.c_void_p * L)(*[v.data_ptr() for v in levels])
    _capi.check(_capi.lib().db200_corr_pyramid_lookup(arr, _ptr(coords.float()), _ptr(out), N, h1, w1, h2, w2, L, radius,
                                                      _DT[levels[0].dtype], _stream()), "corr_pyramid_lookup")
    return out


def corr_volume_pyramid(fmap1, fmap2, num_levels=4):
    """CorrBlock.corr + avg-pool pyramid (corr.py:30-38,63-71).  fmap1,fmap2[N,C,ht,wd] (f16/f32) ->
    list of volumes [N,ht,wd,ht>>l,wd>>l]."""
    _chk(("fmap1", fmap1), ("fmap2", fmap2))
    N, Cc, ht, wd = fmap1.shape
    dt = _DT[fmap1.dtype]
    outs = [torch.empty((N, ht, wd, ht >> l, wd >> l), dtype=fmap1.dtype, device=fmap1.device) for l in range(num_levels)]
    lib = _capi.lib()
    wsb = lib.db200_corr_volume_workspace_bytes(N, Cc, ht, wd, num_levels, dt) if N > 0 else 0
    ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device=fmap1.device)
    arr = (C.c_void_p * num_levels)(*[v.data_ptr() for v in outs])
    _capi.check(lib.db200_corr_volume_pyramid(_ptr(fmap1), _ptr(fmap2), arr, N, Cc, ht, wd, num_levels, dt, _ptr(ws), wsb,
                                              _stream()), "corr_volume_pyramid")
    return outs


def altcorr_forward(fmap1, fmap2, coords, radius):
    """droid.cpp:193-203.  fmap1[B,H1,W1,C], fmap2[B,H2,W2,C], coords[B,S,H1,W1,2] -> [corr[B,S,(2r+1)^2,H1,W1]]"""
    _chk(("fmap1", fmap1), ("fmap2", fmap2), ("coords", coords))
    dt = fmap1.dtype
    f1, f2 = fmap1.float(), fmap2.float()
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, Cc = f2.shape
    rd = 2 * radius + 1
    corr = torch.empty((B, S, rd * rd, H1, W1), dtype=torch.float32, device=f1.device)
    _capi.check(_capi.lib().db200_altcorr_forward(_ptr(f1), _ptr(f2), _ptr(coords.float()), _ptr(corr), B, S, H1, W1, H2, W2,
                                                  Cc, radius, _stream()), "altcorr_forward")
    return [corr if dt == torch.float32 else corr.to(dt)]


def altcorr_backward(fmap1, fmap2, coords, corr_grad, radius):
    """droid.cpp:205-217.  -> [fmap1_grad, fmap2_grad, coords_grad(zeros)]  (float only, like the reference)"""
    _chk(("fmap1", fmap1), ("fmap2", fmap2), ("coords", coords), ("corr_grad", corr_grad))
    f1, f2, g = fmap1.float(), fmap2.float(), corr_grad.float()
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, Cc = f2.shape
    g1, g2 = torch.empty_like(f1), torch.empty_like(f2)
    gc = torch.empty((B, S, H1, W1, 2), dtype=torch.float32, device=f1.device)
    _capi.check(_capi.lib().db200_altcorr_backward(_ptr(f1), _ptr(f2), _ptr(coords.float()), _ptr(g), _ptr(g1), _ptr(g2),
                                                   _ptr(gc), B, S, H1, W1, H2, W2, Cc, radius, _stream()), "altcorr_backward")
    return [g1, g2, gc]


# ------------------------------------------------------------------------------------------------ geometry
def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    """droid.cpp:120-136 -> dist[K] f32"""
    _chk(("poses", poses), ("disps", disps), ("intrinsics", intrinsics), ("ii", ii), ("jj", jj))
    K = ii.shape[0]
    dist = torch.empty(K, dtype=torch.float32, device=poses.device)
    _capi.check(_capi.lib().db200_frame_distance(_ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ii), _ptr(jj), K,
                                                 poses.shape[0], disps.shape[0], disps.shape[1], disps.shape[2], float(beta),
                                                 _ptr(dist), _stream()), "frame_distance")
    return dist


def iproj(poses, disps, intrinsics):
    """droid.cpp:157-167 -> points[num,ht,wd,3]"""
    _chk(("poses", poses), ("disps", disps), ("intrinsics", intrinsics))
    num, ht, wd = disps.shape
    pts = torch.empty((num, ht, wd, 3), dtype=torch.float32, device=disps.device)
    _capi.check(_capi.lib().db200_iproj(_ptr(poses), _ptr(disps), _ptr(intrinsics), num, ht, wd, _ptr(pts), _stream()), "iproj")
    return pts


def depth_filter(poses, disps, intrinsics, ix, thresh):
    """droid.cpp:220-234 -> counter[len(ix),ht,wd]"""
    _chk(("poses", poses), ("disps", disps), ("intrinsics", intrinsics), ("ix", ix), ("thresh", thresh))
    num, ht, wd = disps.shape
    cnt = torch.empty((ix.shape[0], ht, wd), dtype=torch.float32, device=disps.device)
    _capi.check(_capi.lib().db200_depth_filter(_ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ix), _ptr(thresh), ix.shape[0],
                                               num, ht, wd, _ptr(cnt), _stream()), "depth_filter")
    return cnt


def projmap(poses, disps, intrinsics, ii, jj):
    """droid.cpp:139-154 -> [coords[num,ht,wd,3], valid[num,ht,wd,1]] (no caller in the reference; kept for the surface)."""
    _chk(("poses", poses), ("disps", disps), ("intrinsics", intrinsics), ("ii", ii), ("jj", jj))
    num = ii.shape[0]
    B, ht, wd = disps.shape
    coords = torch.empty((num, ht, wd, 3), dtype=torch.float32, device=disps.device)
    valid = torch.empty((num, ht, wd, 1), dtype=torch.float32, device=disps.device)
    _capi.check(_capi.lib().db200_projmap(_ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(ii), _ptr(jj), num, poses.shape[0], B,
                                          ht, wd, _ptr(coords), _ptr(valid), _stream()), "projmap")
    return [coords, valid]


# ------------------------------------------------------------------------------------------------ bundle adjustment
_WS_CACHE = {}


def _workspace(nbytes, device):
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    ws = _WS_CACHE.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        _WS_CACHE[key] = ws
    return ws


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep, motion_only):
    """droid.cpp:88-117 / droid_kernels.cu:1314-1434.  Updates poses[B,7], disps[B,ht,wd] in place and returns
    [dx[P,6], dz[M,ht*wd]] of the last iteration (dz is None for motion_only, where the reference returns an
    undefined tensor)."""
    _chk(("targets", targets), ("weights", weights), ("poses", poses), ("disps", disps), ("intrinsics", intrinsics),
         ("disps_sens", disps_sens), ("ii", ii), ("jj", jj))
    if eta is not None:
        eta = eta.contiguous()
    B, ht, wd = disps.shape
    N = int(ii.shape[0])
    P = int(t1) - int(t0)
    M = int(eta.shape[0]) if (eta is not None and not motion_only) else 0
    lib = _capi.lib()
    wsb = lib.db200_ba_workspace_bytes(N, B, ht, wd, int(t0), int(t1), M, int(bool(motion_only)))
    if wsb == 0:
        raise RuntimeError("ba: invalid sizes")
    ws = _workspace(wsb, poses.device)
    dx = torch.empty((max(P, 0), 6), dtype=torch.float32, device=poses.device)
    dz = None if motion_only else torch.empty((M, ht * wd), dtype=torch.float32, device=poses.device)
    rc = lib.db200_ba(_ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(disps_sens), _ptr(targets), _ptr(weights), _ptr(eta),
                      _ptr(ii), _ptr(jj), N, B, ht, wd, int(t0), int(t1), int(iterations), float(lm), float(ep),
                      int(bool(motion_only)), M, _ptr(dx), _ptr(dz), _ptr(ws), ws.numel(), _stream())
    _capi.check(rc, "ba")
    if _SYNC_STATUS:
        _capi.check(lib.db200_ba_status(_ptr(ws), _stream()), "ba")
    return [dx, dz]
