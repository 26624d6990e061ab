"""Factor-graph edge selection with the reference's semantics (droid_slam/factor_graph.py:44-55, 302-379), on the
GPU through the C-ABI (csrc/graph_edges.cu).  The edge lists that come back equal the reference's `es` element for
element; nothing here loops over edges on the host or synchronises per element (the reference does one `.item()` per
candidate)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _capi


def _vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else C.c_void_p(0)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def neighborhood_edges(t0: int, t1: int, r: int = 3, stereo: bool = False, device="cuda"):
    """factor_graph.py:302-312: every ordered pair of [t0,t1) with c < |i-j| <= r (c = 1 for stereo, else 0)."""
    ii, jj = torch.meshgrid(torch.arange(t0, t1), torch.arange(t0, t1), indexing="ij")
    ii = ii.reshape(-1).to(dtype=torch.long, device=device)
    jj = jj.reshape(-1).to(dtype=torch.long, device=device)
    c = 1 if stereo else 0
    keep = ((ii - jj).abs() > c) & ((ii - jj).abs() <= r)
    return ii[keep], jj[keep]


def proximity_edges(poses, disps, intrinsics, t, t0=0, t1=0, rad=2, nms=2, beta=0.25, thresh=16.0, max_factors=-1,
                    stereo=False, ii_old=None, jj_old=None):
    """The `es` list of FactorGraph.add_proximity_factors (factor_graph.py:315-379) as two int64 device tensors.
    `t` = video.counter.value; (ii_old, jj_old) = cat of the graph's active, bad and inactive edges."""
    lib = _capi.lib()
    dev = poses.device
    intr = intrinsics[0] if intrinsics.dim() == 2 else intrinsics
    for name, x in (("poses", poses), ("disps", disps), ("intrinsics", intr)):
        if not x.is_contiguous() or x.dtype != torch.float32 or not x.is_cuda:
            raise RuntimeError(f"{name} must be a contiguous float32 CUDA tensor")
    B, ht, wd = disps.shape
    n_old = 0 if ii_old is None else int(ii_old.shape[0])
    if n_old:
        ii_old = ii_old.to(device=dev, dtype=torch.long).contiguous()
        jj_old = jj_old.to(device=dev, dtype=torch.long).contiguous()
    cap = lib.db200_proximity_max_edges(t0, t, rad, int(bool(stereo)), max_factors)
    if cap < 0:
        raise RuntimeError("proximity_edges: edge capacity overflow")
    wsb = lib.db200_proximity_workspace_bytes(t0, t1, t)
    ws = torch.empty(max(int(wsb), 256), dtype=torch.uint8, device=dev)
    es = torch.empty((cap, 2), dtype=torch.long, device=dev)
    n = torch.zeros(1, dtype=torch.int32, device=dev)
    _capi.check(lib.db200_proximity_edges(_vp(poses), _vp(disps), _vp(intr), B, ht, wd, t0, t1, t, rad, nms, float(beta),
                                          float(thresh), max_factors, int(bool(stereo)), _vp(ii_old) if n_old else C.c_void_p(0),
                                          _vp(jj_old) if n_old else C.c_void_p(0), n_old, _vp(es), cap, _vp(n), _vp(ws),
                                          ws.numel(), _stream()), "proximity_edges")
    cnt = int(n.item())                      # the one synchronisation: the caller needs the list length
    return es[:cnt, 0].contiguous(), es[:cnt, 1].contiguous()


def filter_repeated_edges(ii, jj, ii_old, jj_old):
    """factor_graph.py:44-55: drop the new edges already present among the old (active + inactive) ones."""
    lib = _capi.lib()
    ii = ii.to(dtype=torch.long).contiguous(); jj = jj.to(dtype=torch.long).contiguous()
    n_new, n_old = int(ii.shape[0]), int(ii_old.shape[0])
    if n_new == 0:
        return ii, jj
    ii_old = ii_old.to(device=ii.device, dtype=torch.long).contiguous(); jj_old = jj_old.to(device=ii.device, dtype=torch.long).contiguous()
    keep = torch.empty(n_new, dtype=torch.uint8, device=ii.device)
    _capi.check(lib.db200_filter_repeated_edges(_vp(ii), _vp(jj), n_new, _vp(ii_old), _vp(jj_old), n_old, _vp(keep), _stream()),
                "filter_repeated_edges")
    keep = keep.bool()
    return ii[keep], jj[keep]


def edge_chunks(ii, n_frames: int, step: int = 8):
    """Index lists of the chunks `FactorGraph.update_lowmem` walks (factor_graph.py:261-290: `v = (ii >= i) & (ii < i + s)`
    for i in range(0, n_frames, s); chunks without an edge are skipped), built with ONE host synchronisation for the whole
    graph.  Indexing with an index list (`ii[idx]`, `coords1[:, idx]`) selects the same edges in the same order as the
    reference's boolean mask, but does not synchronise — with kernels that take 0.15 ms per chunk the three mask indexings
    per chunk were half of a lookup step.  Rebuild when the edge list changes; `ii` is not modified."""
    ii = torch.as_tensor(ii)
    if ii.numel() == 0:
        return []
    chunk = torch.div(ii, step, rounding_mode="floor")
    order = torch.argsort(chunk, stable=True)              # edges of a chunk stay in edge order, like the mask
    n_chunks = (int(n_frames) + step - 1) // step
    counts = torch.bincount(chunk.clamp(0, max(n_chunks - 1, 0)), minlength=n_chunks).tolist()   # the one synchronisation
    out, o = [], 0
    for c in counts:
        if c:
            out.append(order[o:o + c])
        o += c
    return out
