"""GPU parity: correlation lookups / altcorr / corr volume vs the CPU oracle (through droid_backends -> C-ABI)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _coords(N, h1, w1, h2, w2, seed, spread=1.3):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-0.3 * w2, spread * w2, (N, 1, h1, w1))
    y = rng.uniform(-0.3 * h2, spread * h2, (N, 1, h1, w1))
    c = np.concatenate([x, y], 1).astype(np.float32)
    c[0, :, 0, 0] = [3.0, 5.0]          # exactly integer coordinates
    c[0, :, 0, 1] = [-50.0, 1e4]        # fully outside
    return c


@pytest.mark.parametrize("dtype,shape", [(np.float32, (3, 12, 16, 12, 16)), (np.float16, (2, 12, 16, 12, 16)),
                                         (np.float16, (1, 48, 64, 48, 64)), (np.float32, (2, 5, 7, 9, 11))])
def test_corr_index_forward_bit_exact(dtype, shape):
    import droid_backends
    from oracle import oracle
    N, h1, w1, h2, w2 = shape
    rng = np.random.default_rng(1)
    vol = rng.normal(0, 1, shape).astype(dtype)
    co = _coords(N, h1, w1, h2, w2, 2)
    ref = oracle.corr_index_forward(vol, co, 3)
    out, = droid_backends.corr_index_forward(torch.from_numpy(vol).cuda(), torch.from_numpy(co).cuda(), 3)
    assert out.shape == ref.shape
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint16 if dtype == np.float16 else np.uint32),
                                  ref.view(np.uint16 if dtype == np.float16 else np.uint32))


def test_corr_index_forward_f64():
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(3)
    vol = rng.normal(0, 1, (2, 6, 8, 6, 8))
    co = _coords(2, 6, 8, 6, 8, 4)
    out, = droid_backends.corr_index_forward(torch.from_numpy(vol).cuda(), torch.from_numpy(co).cuda(), 3)
    ref = oracle.corr_index_forward(vol.astype(np.float32), co, 3)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=1e-6)


def test_corr_pyramid_lookup_matches_per_level():
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(5)
    N, h, w = 3, 16, 24
    levels = [rng.normal(0, 1, (N, h, w, h >> l, w >> l)).astype(np.float16) for l in range(4)]
    co = _coords(N, h, w, h, w, 6)
    ref = np.concatenate([oracle.corr_index_forward(levels[l], (co / 2 ** l).astype(np.float32), 3).reshape(N, 49, h, w)
                          for l in range(4)], 1)
    out = droid_backends.corr_pyramid_lookup([torch.from_numpy(v).cuda() for v in levels], torch.from_numpy(co).cuda(), 3)
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint16), ref.view(np.uint16))


@pytest.mark.parametrize("dtype,N,h1,w1,h2,w2,L", [
    (np.float16, 2, 5, 40, 16, 64, 4),      # pipelined path, ragged last tile (40 = 32 + 8 source pixels per row)
    (np.float16, 1, 3, 64, 8, 64, 4),       # level 3 is a single row of 8
    (np.float16, 3, 2, 7, 24, 32, 2),       # fewer levels than the kernel's maximum, one short tile
    (np.float32, 1, 4, 33, 8, 32, 4),       # fp32 chunks of 4, levels 32/16/8/4
    (np.float32, 2, 6, 32, 12, 16, 1),
    (np.float16, 1, 4, 32, 12, 24, 2),      # level 1 width 12 is not whole 16-byte chunks -> staged (non-pipelined) kernel
])
def test_corr_pyramid_lookup_pipelined_path_bit_exact(dtype, N, h1, w1, h2, w2, L):
    """The cp.async-pipelined radius-3 kernel (csrc/corr_lookup.cu) against the oracle's per-level restatement of
    correlation_kernels.cu:20-70 — every output bit, including windows that straddle or miss the slice."""
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(11)
    levels = [rng.normal(0, 1, (N, h1, w1, h2 >> l, w2 >> l)).astype(dtype) for l in range(L)]
    co = _coords(N, h1, w1, h2, w2, 12)
    co[0, :, 1, 0] = [w2 - 4.0, h2 - 4.0]          # window ends exactly at the border
    co[0, :, 1, 1] = [-4.0 + 1e-3, 2.5]            # one column inside
    co[0, :, 1, 2] = [7.999, 0.0]                  # fractional part next to 1
    ref = np.concatenate([oracle.corr_index_forward(levels[l], (co / 2 ** l).astype(np.float32), 3).reshape(N, 49, h1, w1)
                          for l in range(L)], 1)
    out = droid_backends.corr_pyramid_lookup([torch.from_numpy(v).cuda() for v in levels], torch.from_numpy(co).cuda(), 3)
    bits = np.uint16 if dtype == np.float16 else np.uint32
    np.testing.assert_array_equal(out.cpu().numpy().view(bits), ref.view(bits))


@pytest.mark.parametrize("shape", [(2, 12, 16, 12, 16), (1, 7, 9, 5, 6)])
def test_corr_index_backward_bit_exact(shape):
    import droid_backends
    from oracle import oracle
    N, h1, w1, h2, w2 = shape
    rng = np.random.default_rng(7)
    co = _coords(N, h1, w1, h2, w2, 8)
    g = rng.normal(0, 1, (N, 7, 7, h1, w1)).astype(np.float32)
    ref = oracle.corr_index_backward(shape, co, g, 3)
    vol = torch.zeros(shape, device="cuda")
    out, = droid_backends.corr_index_backward(vol, torch.from_numpy(co).cuda(), torch.from_numpy(g).cuda(), 3)
    np.testing.assert_array_equal(out.cpu().numpy().view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("B,S,H,W,H2,W2,C", [(2, 1, 12, 16, 12, 16, 128), (1, 2, 6, 9, 3, 4, 32), (1, 1, 48, 64, 24, 32, 128)])
def test_altcorr_forward(B, S, H, W, H2, W2, C):
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(9)
    f1 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H2, W2, C)).astype(np.float32)
    co = np.stack([_coords(B, H, W, H2, W2, 10 + s) for s in range(S)], 1).transpose(0, 1, 3, 4, 2).copy()
    ref = oracle.altcorr_forward(f1, f2, co, 3)
    out, = droid_backends.altcorr_forward(torch.from_numpy(f1).cuda(), torch.from_numpy(f2).cuda(), torch.from_numpy(co).cuda(), 3)
    scale = np.sqrt(C)   # typical magnitude of a dot product of N(0,1) features
    assert np.abs(out.cpu().numpy() - ref).max() <= 1e-4 * scale   # north_star: 1e-4 relative on fp32 correlation values


def test_altcorr_backward():
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(11)
    B, S, H, W, C = 2, 1, 8, 12, 64
    f1 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    f2 = rng.normal(0, 1, (B, H, W, C)).astype(np.float32)
    co = _coords(B, H, W, H, W, 12)[:, None].transpose(0, 1, 3, 4, 2).copy()
    g = rng.normal(0, 1, (B, S, 49, H, W)).astype(np.float32)
    r1, r2, _ = oracle.altcorr_backward(f1, f2, co, g, 3)
    t = lambda a: torch.from_numpy(a).cuda()
    g1, g2, gc = droid_backends.altcorr_backward(t(f1), t(f2), t(co), t(g), 3)
    for a, b in ((g1, r1), (g2, r2)):
        assert np.abs(a.cpu().numpy() - b).max() <= 1e-4 * np.abs(b).max()
    assert float(gc.abs().max()) == 0.0 and gc.shape == (B, S, H, W, 2)


@pytest.mark.parametrize("dtype", [np.float32, np.float16])
def test_corr_volume_pyramid(dtype):
    import droid_backends
    from oracle import oracle
    rng = np.random.default_rng(13)
    N, C, ht, wd = 2, 128, 16, 24
    f1 = rng.normal(0, 1, (N, C, ht, wd)).astype(dtype)
    f2 = rng.normal(0, 1, (N, C, ht, wd)).astype(dtype)
    ref = oracle.corr_volume(f1.astype(np.float32), f2.astype(np.float32))
    outs = droid_backends.corr_volume_pyramid(torch.from_numpy(f1).cuda(), torch.from_numpy(f2).cuda(), 4)
    scale = np.sqrt(C) / 16.0
    tol = 1e-4 if dtype == np.float32 else 3e-3   # fp16 storage: ~2 ulp of the output
    for l in range(4):
        assert outs[l].shape == ref[l].shape
        assert np.abs(outs[l].float().cpu().numpy() - ref[l]).max() <= tol * scale * 4, l


def test_altcorr_equals_lookup_of_volume():
    """SURVEY §8c cross-check: altcorr == corr_index o corr on the same features."""
    import droid_backends
    rng = np.random.default_rng(15)
    N, C, ht, wd = 2, 128, 12, 16
    f1 = torch.from_numpy(rng.normal(0, 1, (N, C, ht, wd)).astype(np.float32)).cuda()
    f2 = torch.from_numpy(rng.normal(0, 1, (N, C, ht, wd)).astype(np.float32)).cuda()
    co = torch.from_numpy(_coords(N, ht, wd, ht, wd, 16)).cuda()
    vol, = droid_backends.corr_volume_pyramid(f1, f2, 1)
    a, = droid_backends.corr_index_forward(vol, co, 3)
    b, = droid_backends.altcorr_forward((f1 / 4).permute(0, 2, 3, 1).contiguous(), (f2 / 4).permute(0, 2, 3, 1).contiguous(),
                                        co.permute(0, 2, 3, 1).contiguous()[:, None], 3)
    assert float((a.reshape(N, 49, ht, wd) - b[:, 0]).abs().max()) <= 1e-4 * np.sqrt(C) / 16 * 4


def test_corr_volume_pyramid_full_size_f16():
    """48x64 maps (BASELINE configs): the tcgen05/TMA GEMM path against an fp32 torch reference of
    CorrBlock.corr + avg_pool2d (corr.py:30-38,63-71) computed on the same f16 inputs."""
    import droid_backends
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    N, C, ht, wd = 3, 128, 48, 64
    f1 = torch.randn((N, C, ht, wd), device="cuda", dtype=torch.float16, generator=g)
    f2 = torch.randn((N, C, ht, wd), device="cuda", dtype=torch.float16, generator=g)
    outs = droid_backends.corr_volume_pyramid(f1, f2, 4)
    a = (f1.float() / 4).reshape(N, C, ht * wd)
    b = f2.float() / 4
    scale = np.sqrt(C) / 16.0
    for l in range(4):
        ref = torch.matmul(a.transpose(1, 2), b.reshape(N, C, -1)).reshape(N, ht, wd, ht >> l, wd >> l)
        assert outs[l].shape == ref.shape and outs[l].dtype == torch.float16
        err = float((outs[l].float() - ref).abs().max())
        assert err <= 3e-3 * scale * 4, (l, err)
        b = torch.nn.functional.avg_pool2d(b, 2, stride=2)


def test_corr_volume_pyramid_full_size_f32():
    """fp32 volumes (training path) through the 3xTF32 tcgen05 pipeline against torch's fp32 matmul + avg_pool2d
    (corr.py:30-38,63-71): within 1e-4 relative of the value scale, i.e. fp32-faithful, not TF32-grade (1e-3)."""
    import droid_backends
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    N, C, ht, wd = 3, 128, 48, 64
    f1 = torch.randn((N, C, ht, wd), device="cuda", generator=g)
    f2 = torch.randn((N, C, ht, wd), device="cuda", generator=g)
    outs = droid_backends.corr_volume_pyramid(f1, f2, 4)
    torch.backends.cuda.matmul.allow_tf32 = False
    a = (f1.double() / 4).reshape(N, C, ht * wd)
    corr = torch.matmul(a.transpose(1, 2), (f2.double() / 4).reshape(N, C, -1)).reshape(N * ht * wd, 1, ht, wd)
    scale = np.sqrt(C) / 16.0
    for l in range(4):
        ref = corr.reshape(N, ht, wd, ht >> l, wd >> l)
        assert outs[l].shape == ref.shape and outs[l].dtype == torch.float32
        err = float((outs[l].double() - ref).abs().max())
        assert err <= 1e-4 * scale, (l, err)                 # measured ~2e-6: the dropped lo*lo terms and fp32 accumulation
        corr = torch.nn.functional.avg_pool2d(corr, 2, stride=2)


# ---- tiled levels (CorrBlock's private pyramid layout; include/droid_b200.h) ---------------------------------------------
@pytest.mark.parametrize("N,h1,w1,h2,w2,L,mask", [
    (2, 48, 64, 48, 64, 4, 0b0011),         # the frontend shape: levels 0 and 1 tiled
    (2, 5, 40, 16, 64, 4, 0b0011),          # ragged last source tile; level 1 = 8 x 32
    (1, 6, 33, 24, 32, 2, 0b0001),          # only level 0 can be tiled (level 1 is 12 x 16)
    (1, 7, 32, 48, 64, 4, 0b0010),          # a mixed pyramid: level 1 tiled, level 0 row-major
])
def test_corr_pyramid_lookup_tiled_bit_exact(N, h1, w1, h2, w2, L, mask):
    import droid_backends
    rng = np.random.default_rng(21)
    levels = [torch.from_numpy(rng.normal(0, 1, (N, h1, w1, h2 >> l, w2 >> l)).astype(np.float16)).cuda() for l in range(L)]
    co = torch.from_numpy(_coords(N, h1, w1, h2, w2, 22)).cuda()
    want = droid_backends.corr_pyramid_lookup(levels, co, 3)
    tiled = [droid_backends.tile_level(v).contiguous() if (mask >> l) & 1 else v for l, v in enumerate(levels)]
    for l in range(L):
        if (mask >> l) & 1:
            assert torch.equal(droid_backends.untile_level(tiled[l]), levels[l])
    out = droid_backends.corr_pyramid_lookup(tiled, co, 3, tiled_mask=mask)
    assert torch.equal(out.view(torch.int16), want.view(torch.int16))


def test_corr_pyramid_lookup_tiled_rejects_other_shapes():
    import droid_backends
    lv = [torch.zeros((1, 4, 8, 12, 16), dtype=torch.float16, device="cuda")]
    co = torch.zeros((1, 2, 4, 8), device="cuda")
    with pytest.raises(RuntimeError):
        droid_backends.corr_pyramid_lookup(lv, co, 3, tiled_mask=1)          # 12 x 16 is not a tiled level
    lv32 = [torch.zeros((1, 4, 8, 48, 64), dtype=torch.float32, device="cuda")]
    with pytest.raises(RuntimeError):
        droid_backends.corr_pyramid_lookup(lv32, co, 3, tiled_mask=1)        # fp16 only


def test_corr_volume_pyramid_tiled_equals_row_major():
    """Level 0 of the tiled pyramid is the row-major level 0 bit for bit (same dot products, permuted columns).  On 64-wide
    maps the pooled levels come out of the same launch as means of the fp32 accumulators rounded once: compared with the
    exact fp32 pyramid (within half an fp16 spacing at the value scale) and with the unfused row-major levels."""
    import droid_backends
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    N, C_, ht, wd = 3, 128, 48, 64
    f1 = torch.randn((N, C_, ht, wd), device="cuda", dtype=torch.float16, generator=g)
    f2 = torch.randn((N, C_, ht, wd), device="cuda", dtype=torch.float16, generator=g)
    want = droid_backends.corr_volume_pyramid(f1, f2, 4)
    got, mask = droid_backends.corr_volume_pyramid(f1, f2, 4, tiled=True)
    assert mask == 0b0011 == droid_backends.corr_tiled_levels(C_, ht, wd, 4, torch.float16)
    rows = [droid_backends.untile_level(v) if (mask >> l) & 1 else v for l, v in enumerate(got)]
    assert torch.equal(rows[0].view(torch.int16), want[0].view(torch.int16))
    exact = torch.matmul(f1.float().reshape(N, C_, -1).transpose(1, 2), f2.float().reshape(N, C_, -1)).reshape(N * ht * wd, 1, ht, wd) / 16.0
    for l in range(1, 4):
        exact = torch.nn.functional.avg_pool2d(exact, 2, stride=2)
        ref = exact.reshape(N, ht, wd, ht >> l, wd >> l)
        ulp = 2.0 ** (np.floor(np.log2(float(ref.abs().max()))) - 10)
        d_new = float((rows[l].float() - ref).abs().max()) / ulp
        d_old = float((want[l].float() - ref).abs().max()) / ulp
        print(f"level {l}: fused pyramid {d_new:.3f}, unfused {d_old:.3f} fp16 spacings from the exact mean")
        assert d_new <= 0.51 and d_old <= 2.0, (l, d_new, d_old)
    co = torch.from_numpy(_coords(N, ht, wd, ht, wd, 23)).cuda()
    a = droid_backends.corr_pyramid_lookup(rows, co, 3)
    b = droid_backends.corr_pyramid_lookup(got, co, 3, tiled_mask=mask)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    # fewer levels than four through the fused launch
    got2, mask2 = droid_backends.corr_volume_pyramid(f1, f2, 2, tiled=True)
    assert mask2 == 0b11 and torch.equal(got2[0], got[0]) and torch.equal(got2[1], got[1])
    # other widths: tiled where the shape allows it, every level from its own GEMM (bit-equal to the row-major pyramid)
    f1w, f2w = f1[:, :, :24, :32].contiguous(), f2[:, :, :24, :32].contiguous()
    got_w, mask_w = droid_backends.corr_volume_pyramid(f1w, f2w, 3, tiled=True)
    want_w = droid_backends.corr_volume_pyramid(f1w, f2w, 3)
    assert mask_w == 0b001
    for l in range(3):
        v = droid_backends.untile_level(got_w[l]) if (mask_w >> l) & 1 else got_w[l]
        assert torch.equal(v.view(torch.int16), want_w[l].view(torch.int16)), l
    # shapes without tiled levels: mask 0, plain pyramid
    f1s, f2s = f1[:, :, :12, :16].contiguous(), f2[:, :, :12, :16].contiguous()
    got_s, mask_s = droid_backends.corr_volume_pyramid(f1s, f2s, 2, tiled=True)
    assert mask_s == 0
    want_s = droid_backends.corr_volume_pyramid(f1s, f2s, 2)
    assert all(torch.equal(x, y) for x, y in zip(got_s, want_s))
