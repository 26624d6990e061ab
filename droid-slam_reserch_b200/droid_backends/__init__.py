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
           "corr_index_forward", "corr_index_backward", "corr_pyramid_lookup", "corr_volume_pyramid", "altcorr_pyramid_forward",
           "corr_tiled_levels", "tile_level", "untile_level",
           "split_hilo_f16", "to_channel_blocked"]

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
    if corr.numel() == 0:                    # empty batch: the reference launches a zero-block grid (an error); nothing to do
        return [corr]
    co = coords.float()                      # kept alive in a local for the duration of the call
    _capi.check(_capi.lib().db200_corr_index_forward(_ptr(volume), _ptr(co), _ptr(corr), N, h1, w1, h2, w2,
                                                     radius, _DT[volume.dtype], _stream()), "corr_index_forward")
    return [corr]


def corr_index_backward(volume, coords, corr_grad, radius):
    """droid.cpp:180-191.  -> [volume_grad like volume]"""
    _chk(("volume", volume), ("coords", coords), ("corr_grad", corr_grad))
    N, h1, w1, h2, w2 = volume.shape
    grad = torch.empty_like(volume)
    if grad.numel() == 0:
        return [grad]
    g = corr_grad.to(volume.dtype)
    co = coords.float()
    _capi.check(_capi.lib().db200_corr_index_backward(_ptr(co), _ptr(g), _ptr(grad), N, h1, w1, h2, w2, radius,
                                                      _DT[volume.dtype], _stream()), "corr_index_backward")
    return [grad]


def corr_pyramid_lookup(levels, coords, radius, tiled_mask=0):
    """Fused CorrBlock.__call__ (corr.py:40-50): levels = list of [N,h1,w1,h2>>l,w2>>l]; coords[N,2,h1,w1]
    at level-0 scale -> out[N, L*(2r+1)^2, h1, w1].  tiled_mask: bit l set = level l is stored in the tiled slice
    layout of `corr_volume_pyramid(..., tiled=True)` (fp16, radius 3; same output bits)."""
    _chk(("coords", coords), *[(f"levels[{i}]", v) for i, v in enumerate(levels)])
    N, h1, w1, h2, w2 = levels[0].shape
    L = len(levels)
    rd = 2 * radius + 1
    out = torch.empty((N, L * rd * rd, h1, w1), dtype=levels[0].dtype, device=coords.device)
    if out.numel() == 0:
        return out
    arr = (C.c_void_p * L)(*[v.data_ptr() for v in levels])
    co = coords.float()
    if tiled_mask:
        _capi.check(_capi.lib().db200_corr_pyramid_lookup_tiled(arr, _ptr(co), _ptr(out), N, h1, w1, h2, w2, L, radius,
                                                                _DT[levels[0].dtype], int(tiled_mask), _stream()),
                    "corr_pyramid_lookup (tiled levels)")
        return out
    _capi.check(_capi.lib().db200_corr_pyramid_lookup(arr, _ptr(co), _ptr(out), N, h1, w1, h2, w2, L, radius,
                                                      _DT[levels[0].dtype], _stream()), "corr_pyramid_lookup")
    return out


def corr_tiled_levels(C_, ht, wd, num_levels, dtype):
    """bit mask of the pyramid levels `corr_volume_pyramid(..., tiled=True)` stores in the tiled slice layout (0: none)"""
    if dtype not in _DT:
        return 0
    return int(_capi.lib().db200_corr_tiled_levels(C_, ht, wd, num_levels, _DT[dtype]))


def untile_level(v):
    """tiled slices [..., h2, w2] (4x8-element tiles, include/droid_b200.h) -> the reference's row-major slices (a copy)"""
    h2, w2 = v.shape[-2:]
    lead = v.shape[:-2]
    return v.reshape(*lead, h2 // 4, w2 // 8, 4, 8).transpose(-3, -2).reshape(*lead, h2, w2)


def tile_level(v):
    """inverse of `untile_level`"""
    h2, w2 = v.shape[-2:]
    lead = v.shape[:-2]
    return v.reshape(*lead, h2 // 4, 4, w2 // 8, 8).transpose(-3, -2).reshape(*lead, h2, w2)


def corr_volume_pyramid(fmap1, fmap2, num_levels=4, tiled=False):
    """CorrBlock.corr + avg-pool pyramid (corr.py:30-38,63-71).  fmap1,fmap2[N,C,ht,wd] (f16/f32) ->
    list of volumes [N,ht,wd,ht>>l,wd>>l].  tiled=True -> (levels, tiled_mask): the levels of `tiled_mask` hold their
    [h2,w2] slices in the tiled layout that `corr_pyramid_lookup(..., tiled_mask)` reads (mask 0 when the shape has none)."""
    _chk(("fmap1", fmap1), ("fmap2", fmap2))
    N, Cc, ht, wd = fmap1.shape
    dt = _DT[fmap1.dtype]
    outs = [torch.empty((N, ht, wd, ht >> l, wd >> l), dtype=fmap1.dtype, device=fmap1.device) for l in range(num_levels)]
    lib = _capi.lib()
    mask = int(lib.db200_corr_tiled_levels(Cc, ht, wd, num_levels, dt)) if tiled else 0
    if N == 0:
        return (outs, mask) if tiled else outs
    wsb = lib.db200_corr_volume_workspace_bytes(N, Cc, ht, wd, num_levels, dt) if N > 0 else 0
    ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device=fmap1.device)
    arr = (C.c_void_p * num_levels)(*[v.data_ptr() for v in outs])
    if tiled:
        _capi.check(lib.db200_corr_volume_pyramid_tiled(_ptr(fmap1), _ptr(fmap2), arr, N, Cc, ht, wd, num_levels, dt, _ptr(ws),
                                                        wsb, mask, _stream()), "corr_volume_pyramid (tiled levels)")
        return outs, mask
    _capi.check(lib.db200_corr_volume_pyramid(_ptr(fmap1), _ptr(fmap2), arr, N, Cc, ht, wd, num_levels, dt, _ptr(ws), wsb,
                                              _stream()), "corr_volume_pyramid")
    return outs


WAIT_TIMES = None
COUNT_CELLS = None      # set to {"cells": 0, "tile_levels": 0} to count the tile x cell GEMMs of altcorr_pyramid_forward
CHECK_STATUS = os.environ.get("DROID_B200_CHECK", "0") == "1"   # tests: read the tensor-core kernels' status words back (synchronises)


def _altcorr_check(ws_status_ptr, what):
    rc = _capi.lib().db200_altcorr_status(C.c_void_p(ws_status_ptr), _stream())
    if rc < 0:
        raise RuntimeError(f"{what}: tensor-core pipeline reported a protocol error (status {rc})")
    return rc


def altcorr_forward(fmap1, fmap2, coords, radius):
    """droid.cpp:193-203.  fmap1[B,H1,W1,C], fmap2[B,H2,W2,C], coords[B,S,H1,W1,2] -> [corr[B,S,(2r+1)^2,H1,W1]]
    C = 128, radius = 3: tcgen05 kernel (csrc/altcorr_tc.cu); otherwise the SIMT kernel (csrc/altcorr.cu)."""
    _chk(("fmap1", fmap1), ("fmap2", fmap2), ("coords", coords))
    dt = fmap1.dtype
    f1, f2, co = fmap1.float(), fmap2.float(), coords.float()
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, Cc = f2.shape
    rd = 2 * radius + 1
    corr = torch.empty((B, S, rd * rd, H1, W1), dtype=torch.float32, device=f1.device)
    if corr.numel() == 0:
        return [corr.to(dt)]
    lib = _capi.lib()
    wsb = lib.db200_altcorr_workspace_bytes(B, S, H1, W1, H2, W2, Cc) if (Cc == 128 and radius == 3 and B * S > 0) else 0
    ws = torch.empty(wsb, dtype=torch.uint8, device=f1.device) if wsb else None
    _capi.check(lib.db200_altcorr_forward_ws(_ptr(f1), _ptr(f2), _ptr(co), _ptr(corr), B, S, H1, W1, H2, W2, Cc, radius,
                                             _ptr(ws), wsb, _stream()), "altcorr_forward")
    if CHECK_STATUS and ws is not None:
        _altcorr_check(ws.data_ptr() + 256, "altcorr_forward")
    return [corr if dt == torch.float32 else corr.to(dt)]


def split_hilo_f16(x, blocked=True):
    """fp32 channels-last feature maps x[F, H, W, 128] -> (hi, lo, flags): fp16 tensors with x ~= hi + lo (|x - hi - lo| <=
    2^-22 |x|) in the CHANNEL-BLOCKED operand layout [F, 2, H, W, 64] of the tensor-core altcorr kernels (`blocked=False`: same
    element order as x), and a device flag word (bit 0: some lo != 0, bit 1: some value non-finite / outside the fp16 range)."""
    _chk(("x", x))
    if x.dtype != torch.float32 or x.numel() % 4 != 0:
        raise RuntimeError("split_hilo_f16: fp32 tensor with numel % 4 == 0 expected")
    hw = 0
    shape = x.shape
    if blocked:
        if x.dim() != 4 or x.shape[-1] != 128:
            raise RuntimeError("split_hilo_f16: channels-last [F, H, W, 128] expected")
        hw = int(x.shape[1] * x.shape[2])
        shape = (x.shape[0], 2, x.shape[1], x.shape[2], 64)
    hi = torch.empty(shape, dtype=torch.float16, device=x.device)
    lo = torch.empty(shape, dtype=torch.float16, device=x.device)
    flags = torch.zeros(1, dtype=torch.int32, device=x.device)
    _capi.check(_capi.lib().db200_split_hilo_f16(_ptr(x), _ptr(hi), _ptr(lo), x.numel(), hw, _ptr(flags), _stream()), "split_hilo_f16")
    return hi, lo, flags


def to_channel_blocked(x):
    """fp16 channels-last feature maps [F, H, W, 128] -> the channel-blocked operand layout [F, 2, H, W, 64]"""
    F_, H, W, Cc = x.shape
    return x.view(F_, H, W, 2, Cc // 2).permute(0, 3, 1, 2, 4).contiguous()


def altcorr_pyramid_forward(pyramid, ii, jj, coords, radius=3, lo=None, flags=None, f32=None):
    """AltCorrBlock.corr_fn in one launch (corr.py:109-131): pyramid[l] = channel-blocked fp16 [frames, 2, H>>l, W>>l, 64]
    (`to_channel_blocked` / `split_hilo_f16`);
    ii, jj[E] int64 frame indices; coords[E,H,W,2] or [E,H,W,S,2] (level-0 scale) -> [E, 49*L, H, W(, S)] fp32.
    fp32 features: pass the fp16 split of every level (`pyramid` = hi parts, `lo`, the merged `flags` word) and the fp32
    levels themselves (`f32`)."""
    L = len(pyramid)
    _chk(("coords", coords), ("ii", ii), ("jj", jj), *[(f"pyramid[{l}]", v) for l, v in enumerate(pyramid)])
    if any(v.dtype != torch.float16 for v in pyramid) or ii.dtype != torch.int64 or jj.dtype != torch.int64:
        raise RuntimeError("altcorr_pyramid_forward: fp16 pyramid and int64 indices expected")
    if pyramid[0].dim() != 5 or pyramid[0].shape[1] != 2 or pyramid[0].shape[4] != 64:
        raise RuntimeError("altcorr_pyramid_forward: channel-blocked pyramid [frames, 2, H, W, 64] expected")
    frames, _, H, W, _ = pyramid[0].shape
    Cc = 128
    co = coords.float()
    multi = co.dim() == 5
    E = co.shape[0]
    S = co.shape[3] if multi else 1
    out = torch.empty((E, 49 * L, H, W) + ((S,) if multi else ()), dtype=torch.float32, device=co.device)
    if out.numel() == 0:
        return out
    lib = _capi.lib()
    wsb = lib.db200_altcorr_pyramid_workspace_bytes(E, S, H, W) if E > 0 else 0
    ws = torch.empty(max(wsb, 16), dtype=torch.uint8, device=co.device)
    arr = lambda ts: (C.c_void_p * L)(*[t.data_ptr() for t in ts])
    null = C.POINTER(C.c_void_p)()
    has_lo = lo is not None
    rc = lib.db200_altcorr_pyramid_forward(
        _ptr(pyramid[0]), _ptr(lo[0]) if has_lo else C.c_void_p(0), arr(pyramid), arr(lo) if has_lo else null,
        _ptr(flags) if has_lo else C.c_void_p(0), _ptr(f32[0]) if has_lo else C.c_void_p(0), arr(f32) if has_lo else null,
        frames, H, W, Cc, L, _ptr(ii), _ptr(jj), _ptr(co), H * W * S * 2, 2, 2 * S, _ptr(out), 49 * L * H * W * S, 1, H * W * S, S,
        E, S, radius, _ptr(ws), wsb, _stream())
    _capi.check(rc, "altcorr_pyramid_forward")
    if CHECK_STATUS and E > 0:
        _altcorr_check(ws.data_ptr(), "altcorr_pyramid_forward")
    if WAIT_TIMES is not None and E > 0:         # debug (DB200_ALT_DBG=16): where the roles of the forward kernel wait
        buf = (C.c_longlong * 12)()
        lib.db200_debug_altcorr_waits(C.c_void_p(ws.data_ptr()), buf, _stream())
        WAIT_TIMES.append(list(buf))
    if COUNT_CELLS is not None and E > 0:        # instrumentation (bench.py): synchronises
        COUNT_CELLS["cells"] += int(lib.db200_altcorr_cells(C.c_void_p(ws.data_ptr()), _stream()))
        COUNT_CELLS["tile_levels"] += E * S * ((H + 7) // 8) * ((W + 15) // 16) * L
    return out


def altcorr_backward(fmap1, fmap2, coords, corr_grad, radius):
    """droid.cpp:205-217.  -> [fmap1_grad, fmap2_grad, coords_grad(zeros)]  (float only, like the reference)"""
    _chk(("fmap1", fmap1), ("fmap2", fmap2), ("coords", coords), ("corr_grad", corr_grad))
    f1, f2, g, co = fmap1.float(), fmap2.float(), corr_grad.float(), coords.float()
    B, S, H1, W1, _ = coords.shape
    _, H2, W2, Cc = f2.shape
    g1, g2 = torch.empty_like(f1), torch.empty_like(f2)
    gc = torch.empty((B, S, H1, W1, 2), dtype=torch.float32, device=f1.device)
    if gc.numel() == 0:                      # no coordinate samples: zero gradients
        return [torch.zeros_like(f1), torch.zeros_like(f2), gc]
    lib = _capi.lib()
    wsb = lib.db200_altcorr_workspace_bytes(B, max(S, 1), H1, W1, H2, W2, Cc) if (Cc == 128 and radius == 3 and B > 0) else 0
    ws = torch.empty(wsb, dtype=torch.uint8, device=f1.device) if wsb else None
    _capi.check(lib.db200_altcorr_backward_ws(_ptr(f1), _ptr(f2), _ptr(co), _ptr(g), _ptr(g1), _ptr(g2), _ptr(gc), B, S, H1, W1,
                                              H2, W2, Cc, radius, _ptr(ws), wsb, _stream()), "altcorr_backward")
    if CHECK_STATUS and ws is not None and S > 0:
        _altcorr_check(ws.data_ptr() + 256, "altcorr_backward")
    return [g1, g2, gc]


# ------------------------------------------------------------------------------------------------ geometry
def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    """droid.cpp:120-136 -> dist[K] f32"""
    _chk(("poses", poses), ("disps", disps), ("intrinsics", intrinsics), ("ii", ii), ("jj", jj))
    K = ii.shape[0]
    dist = torch.empty(K, dtype=torch.float32, device=poses.device)
    if K == 0:
        return dist
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
_WS_CACHE = {}          # (device, stream) -> workspace; least recently used entries are dropped beyond _WS_CACHE_MAX
_WS_CACHE_MAX = 4


def _workspace(nbytes, device):
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    ws = _WS_CACHE.pop(key, None)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
    _WS_CACHE[key] = ws                      # re-inserted last = most recently used
    while len(_WS_CACHE) > _WS_CACHE_MAX:
        _WS_CACHE.pop(next(iter(_WS_CACHE)))
    return ws


def release_workspaces():
    """drop the cached BA workspaces (they are re-created on the next call)"""
    _WS_CACHE.clear()


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep, motion_only):
    """droid.cpp:88-117 / droid_kernels.cu:1314-1434.  Updates poses[B,7], disps[B,ht,wd] in place and returns
    [dx[P,6], dz[M,ht*wd]] of the last iteration (dz is None for motion_only, where the reference returns an
    undefined tensor)."""
    _chk(("targets", targets), ("weights", weights), ("poses", poses), ("disps", disps), ("intrinsics", intrinsics),
         ("disps_sens", disps_sens), ("ii", ii), ("jj", jj))
    if eta is not None:
        eta = eta.contiguous()
    # the reference's packed_accessor32<float / long> would throw on a wrong dtype; raw pointers would silently misread it
    for name, x in (("poses", poses), ("disps", disps), ("intrinsics", intrinsics), ("disps_sens", disps_sens), ("targets", targets),
                    ("weights", weights), ("eta", eta)):
        if x is not None and x.dtype != torch.float32:
            raise RuntimeError(f"ba: {name} must be float32, got {x.dtype}")
    for name, x in (("ii", ii), ("jj", jj)):
        if x.dtype != torch.int64:
            raise RuntimeError(f"ba: {name} must be int64, got {x.dtype}")
    B, ht, wd = disps.shape
    N = int(ii.shape[0])
    P = int(t1) - int(t0)
    M = int(eta.shape[0]) if (eta is not None and not motion_only) else 0
    if (poses.dim() != 2 or poses.shape[1] != 7 or poses.shape[0] < B or intrinsics.numel() < 4 or jj.shape != ii.shape
            or tuple(targets.shape) != (N, 2, ht, wd) or tuple(weights.shape) != (N, 2, ht, wd)
            or (not motion_only and (tuple(disps_sens.shape) != (B, ht, wd) or eta is None or tuple(eta.shape[1:]) != (ht, wd)))):
        raise RuntimeError("ba: tensor shapes do not match (poses[>=B,7], disps/disps_sens[B,ht,wd], targets/weights[N,2,ht,wd], "
                           "eta[M,ht,wd], ii/jj[N])")
    lib = _capi.lib()
    wsb = lib.db200_ba_workspace_bytes(N, B, ht, wd, int(t0), int(t1), M, int(bool(motion_only)))
    if wsb == 0:
        raise RuntimeError("ba: invalid sizes")
    ws = _workspace(wsb, poses.device)
    # the kernels write every element of dx / dz in every iteration; with no iteration the reference returns its torch::zeros
    alloc = torch.empty if int(iterations) > 0 else torch.zeros
    dx = alloc((max(P, 0), 6), dtype=torch.float32, device=poses.device)
    dz = None if motion_only else alloc((M, ht * wd), dtype=torch.float32, device=poses.device)
    rc = lib.db200_ba(_ptr(poses), _ptr(disps), _ptr(intrinsics), _ptr(disps_sens), _ptr(targets), _ptr(weights), _ptr(eta),
                      _ptr(ii), _ptr(jj), N, B, ht, wd, int(t0), int(t1), int(iterations), float(lm), float(ep),
                      int(bool(motion_only)), M, _ptr(dx), _ptr(dz), _ptr(ws), ws.numel(), _stream())
    _capi.check(rc, "ba")
    if _SYNC_STATUS:
        _capi.check(lib.db200_ba_status(_ptr(ws), _stream()), "ba")
    return [dx, dz]
