"""Edge-sharded multi-GPU dense BA — one process per GPU, `torch.distributed` plumbing.

The factor graph of the global backend (`droid_slam/droid_backend.py:27-41`, `factor_graph.py:246-299`) and of
this fork's multisession `Euroc_Multisession_Stereo/BackendAllMaps.py` is partitioned by SOURCE frame: rank g
owns a contiguous range of keyframes `[lo_g, hi_g)` and every edge whose `ii` lies in it, together with those
frames' `target`/`weight` rows, disparity maps and `eta` rows.  Everything the Schur complement couples (C, w,
Q, the E rows and dz of one depth map) is then rank-local, so one Gauss-Newton iteration is

    db200_ba_accumulate   (local edges)  ->  partial reduced system [H | b] (fp64)
    all_reduce(sum)        over NVLink   ->  the full reduced system on every rank          (the ONLY exchange)
    db200_ba_solve_update (replicated damped Cholesky, pose retraction; dz + disparity update of owned frames)

and after the last iteration the owned disparity rows are assembled on every rank (one more all-reduce of a
row-masked copy, 12 KB per keyframe) so that `disps` ends up identical everywhere, as the reference's
single-GPU `droid_backends.ba` leaves it.

`frame_distance` (and the lookups) are pair-/edge-parallel: `frame_distance_sharded` splits the pair list evenly and
all-gathers `dist[K]`.

`partition_frames` / `shard_edges` / `allreduce_system` / `sharded_pair_op` are pure host logic and run on CPU tensors with the
gloo backend (tests/test_dist_cpu.py); `ShardedBA` needs the CUDA library.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist


def frame_costs(ii, jj, n_frames: int) -> np.ndarray:
    """Work estimate per source frame for the fused accumulation kernel (csrc/ba_frame.cu): the per-edge pixel
    pass is linear in the number of edges d of the frame, the in-CTA Schur products grow like d(d+1)/2 pairs."""
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    deg = np.bincount(ii, minlength=n_frames).astype(np.float64)
    pose_deg = np.bincount(ii[ii != jj], minlength=n_frames).astype(np.float64)
    return 1.0 + deg + 0.125 * pose_deg * (pose_deg + 1.0)


def partition_frames(ii, jj, n_frames: int, world: int) -> np.ndarray:
    """Contiguous ownership ranges: bounds[g]..bounds[g+1] are the frames of rank g; prefix-balanced on
    `frame_costs` so that every rank's accumulation kernel runs about equally long.  Returns int64[world+1]."""
    cost = frame_costs(ii, jj, n_frames)
    csum = np.concatenate([[0.0], np.cumsum(cost)])
    targets = csum[-1] * np.arange(1, world) / world
    cuts = np.searchsorted(csum, targets, side="left")
    bounds = np.concatenate([[0], cuts, [n_frames]]).astype(np.int64)
    return np.maximum.accumulate(np.clip(bounds, 0, n_frames))


def shard_edges(ii, bounds: np.ndarray, rank: int) -> np.ndarray:
    """Indices (ascending, i.e. original edge order preserved) of the edges owned by `rank`."""
    ii = np.asarray(ii, np.int64)
    return np.nonzero((ii >= bounds[rank]) & (ii < bounds[rank + 1]))[0]


def local_depth_rows(ii_all, ii_local, t0: int, t1: int):
    """Rows of the GLOBAL depth system kx = unique(cat(arange(t0,t1), ii)) (droid_kernels.cu:1336-1344) that the
    LOCAL graph's kx contains, in order: eta_local = eta[rows]."""
    ts = np.arange(t0, t1, dtype=np.int64)
    kx_all = np.unique(np.concatenate([ts, np.asarray(ii_all, np.int64)]))
    kx_loc = np.unique(np.concatenate([ts, np.asarray(ii_local, np.int64)]))
    rows = np.searchsorted(kx_all, kx_loc)
    return kx_all, kx_loc, rows


def system_pattern(ii, jj, t0: int, t1: int) -> np.ndarray:
    """Flat indices into the packed reduced system [H (lower triangle, row by row) | b] of every entry the graph can make
    non-zero: the 6x6 blocks (a, b), a >= b, of all pose pairs that share a depth map — source frame i couples the poses
    {i} U {j_e : edges i -> j_e} (droid_kernels.cu:1244-1272, 1376-1383) — restricted to the optimised poses [t0, t1), plus b.
    Sorted ascending.  At BASELINE config 3 this is about a third of the dense triangle; the all-reduce of an iteration moves
    only these entries (everything else is zero on every rank)."""
    ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
    P = int(t1) - int(t0)
    n = 6 * P
    order = np.argsort(ii, kind="stable")
    si, sj = ii[order], jj[order]
    starts = np.flatnonzero(np.concatenate([[True], si[1:] != si[:-1]])) if len(si) else np.zeros(0, np.int64)
    ends = np.concatenate([starts[1:], [len(si)]]) if len(si) else starts
    blocks = set()
    for a, b in zip(starts.tolist(), ends.tolist()):
        fr = np.unique(np.concatenate([[si[a]], sj[a:b]])) - t0
        fr = fr[(fr >= 0) & (fr < P)]
        if fr.size:
            hi, lo = np.meshgrid(fr, fr, indexing="ij")
            m = hi >= lo
            blocks.update((hi[m] * P + lo[m]).tolist())
    blk = np.fromiter(blocks, np.int64, len(blocks)) if blocks else np.zeros(0, np.int64)
    ba, bb = blk // max(P, 1), blk % max(P, 1)
    r = (6 * ba)[:, None, None] + np.arange(6)[None, :, None]           # rows of the block
    c = (6 * bb)[:, None, None] + np.arange(6)[None, None, :]
    keep = (r >= c)                                                     # diagonal blocks: their lower triangle only
    flat = (r * (r + 1) // 2 + c)[np.broadcast_to(keep, np.broadcast_shapes(r.shape, c.shape))]
    rhs = n * (n + 1) // 2 + np.arange(n, dtype=np.int64)
    return np.sort(np.concatenate([flat.reshape(-1), rhs])).astype(np.int64)


def allreduce_system_sparse(Hb: torch.Tensor, idx: torch.Tensor, group=None) -> torch.Tensor:
    """all-reduce of the entries `idx` of the packed system only (gather -> all_reduce -> scatter back); the rest of Hb is
    zero on every rank (db200_ba_accumulate clears it) and stays untouched."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        compact = Hb.index_select(0, idx)
        dist.all_reduce(compact, op=dist.ReduceOp.SUM, group=group)
        Hb.index_copy_(0, idx, compact)
    return Hb


def allreduce_system(Hb: torch.Tensor, group=None) -> torch.Tensor:
    """Sum of the per-rank partial reduced systems, in place (fp64, lower triangle of H packed + b: n(n+1)/2 + n values).  NCCL on GPU tensors
    (NVLink 5 / NVSwitch, in-switch reduction when available), gloo on CPU tensors."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(Hb, op=dist.ReduceOp.SUM, group=group)
    return Hb


def assemble_rows(x: torch.Tensor, lo: int, hi: int, group=None) -> torch.Tensor:
    """Every rank contributes rows [lo,hi) of `x`; on return all ranks hold the assembled tensor (in place)."""
    if not (dist.is_initialized() and dist.get_world_size(group) > 1):
        return x
    if lo > 0:
        x[:lo].zero_()
    if hi < x.shape[0]:
        x[hi:].zero_()
    dist.all_reduce(x, op=dist.ReduceOp.SUM, group=group)
    return x


def pair_range(n_pairs: int, rank: int, world: int):
    """Even split of a pair list (`frame_distance` is pair-parallel): rank r gets
    [r*chunk, min((r+1)*chunk, K)) with chunk = ceil(K / world), so that one all-gather of equal chunks rebuilds the list."""
    chunk = -(-int(n_pairs) // max(int(world), 1)) if n_pairs > 0 else 0
    lo = min(rank * chunk, n_pairs)
    return lo, min(lo + chunk, n_pairs), chunk


def sharded_pair_op(fn, ii: torch.Tensor, jj: torch.Tensor, group=None) -> torch.Tensor:
    """`fn(ii_slice, jj_slice) -> f32[len]` evaluated on this rank's slice of the pair list, results all-gathered in pair
    order (K floats in total, <= 1 MB for the 512^2 meshgrid of `DepthVideo.distance`, depth_video.py:171-185).
    Every rank returns the full `dist[K]`, bit-identical to the unsharded call (pairs are independent)."""
    K = int(ii.numel())
    if not (dist.is_initialized() and dist.get_world_size(group) > 1):
        return fn(ii, jj)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi, chunk = pair_range(K, rank, world)
    mine = torch.zeros(chunk, dtype=torch.float32, device=ii.device)
    if hi > lo:
        mine[:hi - lo] = fn(ii[lo:hi].contiguous(), jj[lo:hi].contiguous())
    out = torch.empty(chunk * world, dtype=torch.float32, device=ii.device)
    dist.all_gather_into_tensor(out, mine, group=group)
    return out[:K]


def frame_distance_sharded(poses, disps, intrinsics, ii, jj, beta: float, group=None) -> torch.Tensor:
    """`droid_backends.frame_distance` (droid.cpp:120-136) with the pair list split over the ranks; no other collective."""
    import droid_backends
    return sharded_pair_op(lambda a, b: droid_backends.frame_distance(poses, disps, intrinsics, a, b, beta), ii, jj, group)


class ShardedBA:
    """`droid_backends.ba` for an edge-sharded graph.  Construct once per graph (host-side partition), call `.ba`
    with this rank's `target`/`weight` rows (`[len(local_edges),2,ht,wd]`) and the replicated small tensors."""

    def __init__(self, ii, jj, t0: int, t1: int, rank: int, world: int, device, n_frames=None, group=None):
        ii = np.asarray(ii, np.int64); jj = np.asarray(jj, np.int64)
        self.t0, self.t1, self.rank, self.world, self.device, self.group = int(t0), int(t1), rank, world, device, group
        self.n_frames = int(n_frames) if n_frames is not None else int(max(ii.max(initial=0), jj.max(initial=0), t1 - 1)) + 1
        self.bounds = partition_frames(ii, jj, self.n_frames, world)
        self.own_lo, self.own_hi = int(self.bounds[rank]), int(self.bounds[rank + 1])
        self.local_edges = shard_edges(ii, self.bounds, rank)
        self.kx_all, self.kx_loc, self.eta_rows = local_depth_rows(ii, ii[self.local_edges], self.t0, self.t1)
        self.n_local_frames = int(((self.kx_loc >= self.own_lo) & (self.kx_loc < self.own_hi)).sum())
        self.ii = torch.from_numpy(np.ascontiguousarray(ii[self.local_edges])).to(device)
        self.jj = torch.from_numpy(np.ascontiguousarray(jj[self.local_edges])).to(device)
        self._eta_rows_t = torch.from_numpy(self.eta_rows).to(device)
        self._same_rows = len(self.kx_loc) == len(self.kx_all)
        # rows of the local dz that this rank owns, and where they go in the global dz
        own = (self.kx_loc >= self.own_lo) & (self.kx_loc < self.own_hi)
        self._dz_src = torch.from_numpy(np.nonzero(own)[0]).to(device)
        self._dz_dst = torch.from_numpy(self.eta_rows[own]).to(device)
        self._ws = None
        # exchange pattern of the reduced system (host-side, once per graph): used when it is clearly smaller than the triangle
        n = 6 * (self.t1 - self.t0)
        pat = system_pattern(ii, jj, self.t0, self.t1) if world > 1 else None
        self.pattern_frac = (len(pat) / float(n * (n + 1) // 2 + n)) if pat is not None else 1.0
        self._pattern = torch.from_numpy(pat).to(device) if (pat is not None and self.pattern_frac < 0.6) else None

    def exchange(self):
        """the ONE exchange of an iteration: sum of the partial reduced systems over the ranks"""
        if self._pattern is not None:
            allreduce_system_sparse(self.system, self._pattern, self.group)
        else:
            allreduce_system(self.system, self.group)

    # ---- one call = begin -> (accumulate -> all-reduce of .system -> solve_update) x iterations -> finish
    def begin(self, poses, disps, intrinsics, disps_sens, target, weight, eta, lm, ep, motion_only=False):
        from . import _capi
        self._lib = _capi.lib()
        lib = self._lib
        B, ht, wd = disps.shape
        mo = int(bool(motion_only))
        eta_loc = None
        if not mo:
            eta_loc = eta if self._same_rows else eta.index_select(0, self._eta_rows_t)
            eta_loc = eta_loc.contiguous()
        M_loc = 0 if mo else int(eta_loc.shape[0])
        if B < self.n_frames:
            raise RuntimeError("ShardedBA: disps has fewer frames than the partition was built for")
        N = int(self.ii.shape[0])
        wsb = lib.db200_ba_workspace_bytes(N, B, ht, wd, self.t0, self.t1, M_loc, mo)
        if self._ws is None or self._ws.numel() < wsb:
            self._ws = torch.empty(int(wsb) + 1024, dtype=torch.uint8, device=self.device)
        ws = self._ws
        vp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        sys_ptr = lib.db200_ba_system_ptr(vp(ws), N, B, ht, wd, self.t0, self.t1, M_loc, mo)
        off = int(sys_ptr) - ws.data_ptr()
        n_el = int(lib.db200_ba_system_elems(self.t0, self.t1))
        self.system = ws[off:off + 8 * n_el].view(torch.float64)        # [H (packed lower triangle) | b] fp64: the all-reduced buffer
        P = self.t1 - self.t0
        self._dx = torch.empty((P, 6), dtype=torch.float32, device=self.device)
        self._dz_loc = None if mo else torch.empty((M_loc, ht * wd), dtype=torch.float32, device=self.device)
        self._call = dict(poses=poses, disps=disps, intr=intrinsics, sens=disps_sens, target=target, weight=weight, eta=eta_loc,
                          N=N, B=B, ht=ht, wd=wd, mo=mo, M=M_loc, lm=float(lm), ep=float(ep),
                          own_hi=B if self.rank == self.world - 1 else self.own_hi)

    def accumulate(self, first: bool):
        from . import _capi
        c, lib, ws = self._call, self._lib, self._ws
        vp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _capi.check(lib.db200_ba_accumulate(vp(c["poses"]), vp(c["disps"]), vp(c["intr"]), vp(c["sens"]), vp(c["target"]),
                                            vp(c["weight"]), vp(c["eta"]), vp(self.ii), vp(self.jj), c["N"], c["B"], c["ht"],
                                            c["wd"], self.t0, self.t1, c["mo"], c["M"], int(first), self.own_lo, c["own_hi"],
                                            vp(ws), ws.numel(), st), "ba_accumulate")

    def solve_update(self, last: bool):
        from . import _capi
        c, lib, ws = self._call, self._lib, self._ws
        vp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _capi.check(lib.db200_ba_solve_update(vp(c["poses"]), vp(c["disps"]), vp(c["intr"]), vp(c["target"]), vp(c["weight"]),
                                              vp(self.ii), vp(self.jj), c["N"], c["B"], c["ht"], c["wd"], self.t0, self.t1,
                                              c["lm"], c["ep"], c["mo"], c["M"], self.own_lo, c["own_hi"],
                                              vp(self._dx) if last else C.c_void_p(0),
                                              vp(self._dz_loc) if (last and not c["mo"]) else C.c_void_p(0), vp(ws), ws.numel(), st),
                    "ba_solve_update")

    def local_dz_rows(self):
        """(rows of the global dz this rank owns, their values) after the last solve_update."""
        return self._dz_dst, self._dz_loc.index_select(0, self._dz_src)

    def ba(self, poses, disps, intrinsics, disps_sens, target, weight, eta, iterations, lm, ep, motion_only=False):
        """Same contract as droid_backends.ba: poses/disps updated in place (identical on all ranks on return),
        returns [dx[P,6], dz[M,ht*wd]] of the last iteration (dz assembled from the owners)."""
        self.begin(poses, disps, intrinsics, disps_sens, target, weight, eta, lm, ep, motion_only)
        for it in range(int(iterations)):
            self.accumulate(it == 0)
            self.exchange()
            self.solve_update(it == iterations - 1)
        if motion_only:
            return [self._dx, None]
        # after the last iteration every rank owns rows of disps and of dz: ONE all-reduce of a row-masked [disps | dz] buffer
        # assembles both (12 KB per keyframe each)
        hw = self._call["ht"] * self._call["wd"]
        B, M = disps.shape[0], len(self.kx_all)
        lo, hi = self.own_lo, self._call["own_hi"]
        dst, val = self.local_dz_rows()
        if not (self.world > 1 and dist.is_initialized()):
            dz = torch.zeros((M, hw), dtype=torch.float32, device=self.device)
            dz.index_copy_(0, dst, val)
            return [self._dx, dz]
        buf = torch.zeros((B + M, hw), dtype=torch.float32, device=self.device)
        buf[lo:hi] = disps[lo:hi].reshape(hi - lo, hw)
        buf[B:].index_copy_(0, dst, val)
        dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=self.group)
        disps.copy_(buf[:B].view_as(disps))
        return [self._dx, buf[B:]]

    def status(self):
        from . import _capi
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        return _capi.lib().db200_ba_status(C.c_void_p(self._ws.data_ptr()), st)
