"""Multi-GPU host logic on CPU: world_size-2 gloo runs of the edge partition + the per-iteration all-reduce of the
reduced pose system (droid_b200/dist.py).  The per-rank partial systems come from the CPU oracle
(the checker) — the product's CUDA path is exercised by tests/test_gpu_dist.py on the B200 box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from droid_b200 import dist as bdist
from droid_b200 import synth


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_partition_covers_every_edge_once():
    p = synth.config_c3(n_kf=96, n_edges=900, ht=16, wd=24)
    for world in (1, 2, 3, 4, 8):
        b = bdist.partition_frames(p.ii, p.jj, 96, world)
        assert b[0] == 0 and b[-1] == 96 and np.all(np.diff(b) >= 0) and len(b) == world + 1
        parts = [bdist.shard_edges(p.ii, b, r) for r in range(world)]
        allidx = np.concatenate(parts)
        assert np.array_equal(np.sort(allidx), np.arange(p.n_edges))          # a partition
        for r, idx in enumerate(parts):
            assert np.all(np.diff(idx) > 0)                                      # original edge order kept
            assert np.all((p.ii[idx] >= b[r]) & (p.ii[idx] < b[r + 1]))          # source frame owned
        cost = bdist.frame_costs(p.ii, p.jj, 96)
        loads = np.array([cost[b[r]:b[r + 1]].sum() for r in range(world)])
        assert loads.max() <= loads.mean() * 1.25 + cost.max()                  # balanced up to one frame


def test_partition_edge_cases():
    # more ranks than frames with edges; empty graph
    b = bdist.partition_frames(np.array([0, 0, 1]), np.array([1, 2, 0]), 3, 8)
    assert b[0] == 0 and b[-1] == 3 and np.all(np.diff(b) >= 0)
    assert sum(len(bdist.shard_edges(np.array([0, 0, 1]), b, r)) for r in range(8)) == 3
    b = bdist.partition_frames(np.zeros(0, np.int64), np.zeros(0, np.int64), 5, 2)
    assert b[0] == 0 and b[-1] == 5


def test_local_depth_rows_with_fixed_source_frame():
    # frame 0 is fixed (t0=1) but a source of edges: only its owner has it in the local depth system
    ii = np.array([0, 0, 1, 2, 3, 3]); jj = np.array([1, 2, 0, 1, 2, 1])
    kx_all, kx_loc, rows = bdist.local_depth_rows(ii, ii[ii >= 2], 1, 4)
    assert kx_all.tolist() == [0, 1, 2, 3] and kx_loc.tolist() == [1, 2, 3] and rows.tolist() == [1, 2, 3]
    kx_all, kx_loc, rows = bdist.local_depth_rows(ii, ii[ii < 2], 1, 4)
    assert kx_loc.tolist() == [0, 1, 2, 3] and rows.tolist() == [0, 1, 2, 3]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        p = synth.config_c3(n_kf=24, n_edges=160, ht=16, wd=24, seed=77)
        n = 6 * (p.t1 - p.t0)
        bounds = bdist.partition_frames(p.ii, p.jj, 24, world)
        sel = bdist.shard_edges(p.ii, bounds, rank)
        _, _, rows = bdist.local_depth_rows(p.ii, p.ii[sel], p.t0, p.t1)
        part = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target[sel], p.weight[sel], p.eta[rows],
                         p.ii[sel], p.jj[sel], p.t0, p.t1, 1, 1e-5, 1e-2, dump_system=True)
        Hb = torch.from_numpy(np.concatenate([part["H"].reshape(-1), part["b"]]))
        # the product's packed layout (lower triangle row by row, then b), exchanged through the graph's block pattern only
        packed = torch.from_numpy(np.concatenate([part["H"][np.tril_indices(n)], part["b"]]))
        pat = bdist.system_pattern(p.ii, p.jj, p.t0, p.t1)
        assert np.isin(np.flatnonzero(packed.numpy()), pat).all()        # every partial non-zero lies inside the pattern
        sparse = bdist.allreduce_system_sparse(packed.clone(), torch.from_numpy(pat))
        dense = bdist.allreduce_system(packed.clone())
        assert torch.equal(sparse, dense) and len(pat) < 0.8 * packed.numel()
        bdist.allreduce_system(Hb)
        # owned disparity rows assembled on every rank
        d = torch.from_numpy(part["disps"].copy())
        bdist.assemble_rows(d, int(bounds[rank]), int(bounds[rank + 1]))
        q.put((rank, Hb.numpy().copy(), d.numpy().copy(), n))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world2_partial_systems_sum_to_the_full_system():
    from oracle import oracle
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    p = synth.config_c3(n_kf=24, n_edges=160, ht=16, wd=24, seed=77)
    full = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1,
                     1e-5, 1e-2, dump_system=True)
    n = res[0][3]
    want = np.concatenate([full["H"].reshape(-1), full["b"]])
    for rank, Hb, d, _ in res:
        assert Hb.shape == want.shape
        # fp32 per-edge sums regrouped per rank, fp64 across ranks: equal up to fp32 round-off of the largest entries
        assert np.abs(Hb - want).max() <= 2e-6 * np.abs(want).max()
    np.testing.assert_array_equal(res[0][1], res[1][1])                  # all-reduce: identical on all ranks
    np.testing.assert_array_equal(res[0][2], res[1][2])
    # every rank solved with its PARTIAL system above (oracle.ba is one self-contained call), so only the
    # assembly mechanics are checked here: row r of the assembled map comes from r's owner
    assert res[0][2].shape == full["disps"].shape


def _fd_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        p = synth.config_c3(n_kf=12, n_edges=60, ht=16, wd=24, seed=5)
        n = p.t1
        ii, jj = torch.meshgrid(torch.arange(n), torch.arange(n), indexing="ij")
        ii, jj = ii.reshape(-1)[:-5], jj.reshape(-1)[:-5]            # 139 pairs: not a multiple of the world size
        calls = []

        def fn(a, b):
            calls.append(len(a))
            return torch.from_numpy(oracle.frame_distance(p.poses, p.disps, p.intrinsics, a.numpy(), b.numpy(), 0.3))
        d = bdist.sharded_pair_op(fn, ii, jj)
        q.put((rank, d.numpy().copy(), calls))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_gloo_world2_frame_distance_pairs_are_split_and_gathered():
    from oracle import oracle
    assert bdist.pair_range(139, 0, 2) == (0, 70, 70) and bdist.pair_range(139, 1, 2) == (70, 139, 70)
    assert bdist.pair_range(3, 7, 8) == (3, 3, 1) and bdist.pair_range(0, 0, 2) == (0, 0, 0)
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_fd_worker, args=(r, world, port, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = sorted([q.get(timeout=240) for _ in range(world)], key=lambda t: t[0])
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    p = synth.config_c3(n_kf=12, n_edges=60, ht=16, wd=24, seed=5)
    n = p.t1
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    want = oracle.frame_distance(p.poses, p.disps, p.intrinsics, ii.reshape(-1)[:-5], jj.reshape(-1)[:-5], 0.3)
    for rank, d, calls in res:
        np.testing.assert_array_equal(d.view(np.uint32), want.view(np.uint32))      # bit-exact, in pair order, on every rank
    assert res[0][2] == [70] and res[1][2] == [69]


def test_system_pattern_is_exactly_the_nonzero_set_of_the_reduced_system():
    """droid_b200.dist.system_pattern against the oracle's dumped [H | b] (droid_kernels.cu:1244-1272, 1376-1383): with stereo
    edges, a fixed first pose range (t0 = 2) and a buffer longer than t1"""
    from oracle import oracle
    rng = np.random.default_rng(3)
    ii, jj = synth.backend_edges(30, 200, rng, rad=2, stereo_frac=0.2)
    p = synth.make_ba_problem(ii, jj, 30, ht=16, wd=24, seed=4, t0=2, buffer=34)
    r = oracle.ba(p.poses, p.disps, p.intrinsics, p.disps_sens, p.target, p.weight, p.eta, p.ii, p.jj, p.t0, p.t1, 1, 1e-4, .1,
                  dump_system=True)
    n = r["H"].shape[0]
    flat = np.concatenate([r["H"][np.tril_indices(n)], r["b"]])
    pat = bdist.system_pattern(p.ii, p.jj, p.t0, p.t1)
    assert np.all(np.diff(pat) > 0) and pat[-1] == len(flat) - 1
    assert np.array_equal(np.flatnonzero(flat), pat)
    # a backend graph at BASELINE config 3 size: a small fraction of the dense triangle
    ii, jj = synth.backend_edges(512, 6000, np.random.default_rng(1237), rad=2, stereo_frac=0.2)
    pat = bdist.system_pattern(ii, jj, 1, 512)
    assert len(pat) < 0.25 * (3066 * 3067 // 2 + 3066)
