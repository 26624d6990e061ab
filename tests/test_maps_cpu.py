"""Reconstruction folders -> fused, owner-sharded keyframe buffers (droid_b200/maps.py) against a literal restatement of
BackendAllMaps.py:43-61,101-138 (load every .npy of every folder, `[::N]`, lay the maps end to end)."""
import os

import numpy as np

from droid_b200 import maps


def _reference_concat(root, N):
    fused = {}
    for folder in maps.sorted_map_folders(root):
        for file in os.listdir(folder):
            if file.endswith(".npy"):
                fused.setdefault(file[:-4], []).append(np.load(os.path.join(folder, file))[::N])
    return {k: np.concatenate(v, 0) for k, v in fused.items()}


def _write_maps(tmp_path, sizes):
    rng = np.random.default_rng(0)
    for name, n in sizes.items():
        maps.save_reconstruction(
            str(tmp_path / name), tstamps=np.arange(n, dtype=np.float64) + 1000 * len(name), images=rng.integers(0, 255, (n, 3, 8, 8), dtype=np.uint8),
            disps=rng.random((n, 4, 6), dtype=np.float32), poses=rng.random((n, 7), dtype=np.float32),
            intrinsics=rng.random((n, 4), dtype=np.float32), fmaps=rng.random((n, 2, 5, 4, 6)).astype(np.float16),
            inps=rng.random((n, 5, 4, 6)).astype(np.float16), nets=rng.random((n, 5, 4, 6)).astype(np.float16))


def test_folder_order_follows_first_number(tmp_path):
    _write_maps(tmp_path, {"MH10": 3, "MH2": 4, "V1_03": 5, "nonumber": 2})
    names = [os.path.basename(f) for f in maps.sorted_map_folders(str(tmp_path))]
    assert names.index("nonumber") < names.index("V1_03") < names.index("MH2") < names.index("MH10")
    assert maps.sorted_map_folders(str(tmp_path / "missing")) == []


def test_fused_buffers_equal_reference_concatenation(tmp_path):
    _write_maps(tmp_path, {"MH01": 7, "MH02": 4, "MH03": 9})
    for N in (1, 2, 3):
        want = _reference_concat(str(tmp_path), N)
        got = maps.load_maps(str(tmp_path), every=N)
        assert got["counter"] == want["poses"].shape[0]
        assert got["offsets"].tolist() == np.concatenate([[0], np.cumsum([len(range(0, n, N)) for n in (7, 4, 9)])]).tolist()
        for k in maps.ARRAYS:
            np.testing.assert_array_equal(got[k].numpy(), want[k])


def test_owner_sharded_load_keeps_only_owned_heavy_rows(tmp_path):
    _write_maps(tmp_path, {"MH01": 7, "MH02": 4, "MH03": 9})
    N = 2
    want = _reference_concat(str(tmp_path), N)
    T = want["poses"].shape[0]                      # 4 + 2 + 5 = 11
    ii = np.repeat(np.arange(T), 3); jj = (ii + 1) % T
    covered = np.zeros(T, int)
    for world in (2, 3):
        covered[:] = 0
        for rank in range(world):
            lo, hi = maps.owned_range_for_rank(ii, jj, T, rank, world)
            got = maps.load_maps(str(tmp_path), every=N, own_lo=lo, own_hi=hi)
            assert got["own"] == (lo, hi)
            for k in ("fmaps", "nets", "inps", "images", "disps"):
                np.testing.assert_array_equal(got[k].numpy(), want[k][lo:hi])        # ranges straddle map boundaries
            for k in maps.REPLICATED:
                np.testing.assert_array_equal(got[k].numpy(), want[k])
            covered[lo:hi] += 1
        assert np.all(covered == 1)
    empty = maps.load_maps(str(tmp_path), every=N, own_lo=5, own_hi=5)
    assert empty["fmaps"].shape[0] == 0 and empty["poses"].shape[0] == T


def test_edge_chunks_select_what_the_reference_masks_select():
    """droid_b200.graph.edge_chunks against the mask loop of FactorGraph.update_lowmem (factor_graph.py:268-272)"""
    import torch
    from droid_b200 import graph
    g = torch.Generator().manual_seed(5)
    for n_frames, n_edges, step in ((37, 400, 8), (8, 20, 8), (64, 1, 8), (50, 300, 5)):
        ii = torch.randint(0, n_frames, (n_edges,), generator=g)
        jj = torch.randint(0, n_frames, (n_edges,), generator=g)
        want = []
        for i in range(0, n_frames, step):
            v = (ii >= i) & (ii < i + step)
            if v.any():
                want.append((ii[v], jj[v]))
        got = graph.edge_chunks(ii, n_frames, step)
        assert len(got) == len(want)
        for idx, (wi, wj) in zip(got, want):
            assert torch.equal(ii[idx], wi) and torch.equal(jj[idx], wj)
    assert graph.edge_chunks(torch.zeros(0, dtype=torch.int64), 10) == []
