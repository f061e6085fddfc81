"""GPU parity of the trunk operators vs plain PyTorch fp32 references of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def _nchw(x):
    return x.permute(0, 3, 1, 2).contiguous()


def _ref_conv(x_nchw, w, b, s, act, slope, ups):
    y = F.conv2d(x_nchw, w, b, s, w.shape[-1] // 2)
    if act == 1:
        y = F.silu(y)
    elif act == 2:
        y = F.leaky_relu(y, slope)
    if ups == 2:
        y = F.interpolate(y, scale_factor=2, mode='nearest')
    return y


CONV_CASES = [
    # (N, H, W, Cin, Cout, k, s, act, ups)
    (2, 20, 20, 64, 64, 1, 1, 1, 1),
    (2, 20, 20, 128, 256, 3, 1, 1, 1),
    (1, 40, 40, 256, 128, 1, 1, 1, 2),      # lateral conv + fused upsample
    (2, 24, 40, 64, 128, 3, 2, 1, 1),       # stride 2
    (1, 64, 96, 16, 32, 3, 1, 1, 1),        # stem-like (Cin padded to 16, SW32)
    (1, 64, 64, 32, 64, 3, 2, 2, 1),        # Cin 32 (SW64), LeakyReLU
    (2, 20, 20, 80, 160, 3, 1, 1, 1),       # e6e widths: Cin not a multiple of 64
    (1, 20, 20, 512, 255, 1, 1, 0, 1),      # head conv (Cout 255)
    (1, 160, 160, 64, 64, 3, 1, 1, 1),      # wide map, many tiles
    (3, 12, 20, 96, 96, 3, 1, 1, 1),        # ragged tiles
    (1, 40, 40, 32, 32, 3, 1, 2, 1),        # slab kernel, Cin 32 (SW64), LeakyReLU
    (2, 40, 40, 128, 128, 3, 1, 1, 1),      # slab kernel, 40-wide map (120-pixel sub-tiles)
    (1, 80, 80, 64, 192, 3, 1, 1, 1),       # slab kernel, 16x8 sub-tiles, 3 N tiles of 64
    (2, 24, 24, 512, 64, 3, 1, 1, 1),       # slab kernel, 8 K chunks
    (2, 128, 192, 32, 64, 3, 2, 1, 1),      # pair-slab stride-2 kernel (Cin 32), several tiles per image
    (1, 36, 52, 32, 64, 3, 2, 1, 1),        # pair-slab, ragged tiles in both directions
    (1, 32, 64, 32, 32, 3, 2, 2, 1),        # pair-slab, Cout 32, LeakyReLU
    (1, 35, 52, 32, 64, 3, 2, 1, 1),        # odd height: generic kernel with the pair view
]


@pytest.mark.parametrize('case', CONV_CASES, ids=[str(c) for c in CONV_CASES])
def test_conv_fp32_reference_mode(case):
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, k, s, act, ups = case
    g = torch.Generator().manual_seed(1)
    x = torch.randn((N, Ci, H, W), generator=g)
    w = torch.randn((Co, Ci, k, k), generator=g) / (Ci * k * k) ** 0.5
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, s, act, 0.1, ups)
    y = ops.conv2d(_nhwc(x).cuda(), w, b, s, act, 0.1, ups)
    got = _nchw(y[..., :Co].float().cpu())
    assert torch.allclose(got, ref, rtol=1e-4, atol=1e-4), float((got - ref).abs().max())


@pytest.mark.parametrize('case', CONV_CASES, ids=[str(c) for c in CONV_CASES])
def test_conv_tcgen05_bf16(case):
    """bf16 operands, fp32 accumulate: compare against an fp32 conv of the bf16-rounded operands.
    Tolerance: output rounding to bf16 (2^-8 relative) + accumulation-order noise."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, k, s, act, ups = case
    g = torch.Generator().manual_seed(1)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, k, k), generator=g) / (Ci * k * k) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, s, act, 0.1, ups)
    head = Co == 255
    y = ops.conv2d(_nhwc(x).cuda().bfloat16(), w, b, s, act, 0.1, ups, out_dtype=torch.float32 if head else None)
    got = _nchw(y[..., :Co].float().cpu())
    tol = 2e-3 if head else 1.2e-2
    assert torch.allclose(got, ref, rtol=tol, atol=tol), float((got - ref).abs().max())


def test_conv_tc_writes_concat_slice():
    from yolo_series_b200 import ops
    g = torch.Generator().manual_seed(2)
    x = torch.randn((1, 64, 16, 16), generator=g).bfloat16().float()
    w = (torch.randn((64, 64, 3, 3), generator=g) / 24).bfloat16().float()
    b = torch.zeros(64)
    out = torch.full((1, 16, 16, 256), 7.0, dtype=torch.bfloat16, device='cuda')
    ops.conv2d(_nhwc(x).cuda().bfloat16(), w, b, 1, 0, 0.0, 1, out=out, out_c_off=128)
    ref = F.conv2d(x, w, b, 1, 1)
    o = out.float().cpu()
    assert torch.all(o[..., :128] == 7.0) and torch.all(o[..., 192:] == 7.0)
    assert torch.allclose(_nchw(o[..., 128:192]), ref, rtol=1.2e-2, atol=1.2e-2)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_pools_upsample_add(dtype):
    from yolo_series_b200 import ops
    g = torch.Generator().manual_seed(3)
    x = torch.randn((2, 64, 20, 28), generator=g).to(dtype)
    xc = _nhwc(x).cuda()
    xf = x.float()
    assert torch.equal(_nchw(ops.maxpool2x2(xc).float().cpu()), F.max_pool2d(xf, 2, 2))
    for k in (5, 9, 13):
        assert torch.equal(_nchw(ops.maxpool_s1(xc, k).float().cpu()), F.max_pool2d(xf, k, 1, k // 2))
    s3 = _nchw(ops.spp3(xc).float().cpu())
    for i, k in enumerate((5, 9, 13)):
        assert torch.equal(s3[:, i * 64:(i + 1) * 64], F.max_pool2d(xf, k, 1, k // 2))
    assert torch.equal(_nchw(ops.upsample2x(xc).float().cpu()), F.interpolate(xf, scale_factor=2, mode='nearest'))
    y = torch.randn((2, 64, 20, 28), generator=g).to(dtype)
    assert torch.equal(_nchw(ops.add(xc, _nhwc(y).cuda()).cpu()).float(), (x + y).float())


def test_stage_input_and_reorg():
    from yolo_series_b200 import ops
    g = torch.Generator().manual_seed(4)
    u8 = torch.randint(0, 256, (2, 3, 32, 48), generator=g, dtype=torch.uint8)
    ref = u8.float() / 255.0
    y = ops.stage_input(u8.cuda(), 4, torch.float32)
    assert torch.equal(_nchw(y.cpu())[:, :3], ref) and torch.all(y[..., 3] == 0)
    yb = ops.stage_input(ref.cuda(), 16, torch.bfloat16)
    assert torch.equal(_nchw(yb.float().cpu())[:, :3], ref.bfloat16().float())
    r = ops.stage_input(ref.cuda(), 16, torch.float32, reorg=True)
    a = ref
    want = torch.cat([a[..., ::2, ::2], a[..., 1::2, ::2], a[..., ::2, 1::2], a[..., 1::2, 1::2]], 1)
    assert torch.equal(_nchw(r.cpu())[:, :12], want) and torch.all(r[..., 12:] == 0)


@pytest.mark.parametrize('cfg', [(3, 32, 1, False, 48, 160), (3, 32, 2, False, 64, 200), (3, 64, 1, True, 64, 96),
                                 (3, 80, 1, True, 32, 288), (3, 32, 1, False, 37, 70), (3, 32, 2, False, 45, 75),
                                 (3, 32, 1, False, 256, 320), (3, 32, 2, False, 256, 640), (3, 32, 1, False, 37, 72),
                                 (3, 32, 2, False, 45, 76), (3, 64, 1, True, 256, 1280), (3, 80, 1, True, 130, 520),
                                 (3, 64, 1, True, 64, 100), (3, 80, 1, True, 2, 8)],
                         ids=['v7', 'tiny_s2', 'w6_reorg', 'e6e_reorg', 'v7_ragged', 'tiny_ragged', 'v7_ringwrap',
                              'tiny_ringwrap', 'v7_ragged4', 'tiny_ragged4', 'w6_reorg_ringwrap', 'e6e_reorg_ragged',
                              'w6_reorg_w_not_mult8_fallback', 'e6e_reorg_single_row'])
@pytest.mark.parametrize('u8', [True, False])
def test_stem_conv_fused_staging(cfg, u8):
    """Fused staging + first conv (tcgen05 kernels: k_stem_tc for RGB -> 32, k_stemr_tc for the ReOrg stems 12 -> 64 | 80;
    the mma.sync kernel remains for image widths the builders cannot load in aligned groups) vs fp32 conv of the
    bf16-rounded operands."""
    from yolo_series_b200 import ops
    C_, Co, s, reorg, H, W = cfg
    g = torch.Generator().manual_seed(6)
    img = torch.randint(0, 256, (2, C_, H, W), generator=g, dtype=torch.uint8)
    x = img.float() / 255.0
    src = img if u8 else x
    if reorg:
        x = torch.cat([x[..., ::2, ::2], x[..., 1::2, ::2], x[..., ::2, 1::2], x[..., 1::2, 1::2]], 1)
    ci = x.shape[1]
    w = (torch.randn((Co, ci, 3, 3), generator=g) / (ci * 9) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = F.silu(F.conv2d(x.bfloat16().float(), w, b, s, 1))
    y = ops.stem_conv(src.cuda(), w, b, s, 1, 0.0, reorg)
    got = _nchw(y.float().cpu())
    assert got.shape == ref.shape
    assert torch.allclose(got, ref, rtol=1.2e-2, atol=1.2e-2), float((got - ref).abs().max())


@pytest.mark.parametrize('case', [(2, 20, 20, 128, 256, 3, 1, 1, 1), (1, 80, 80, 64, 192, 3, 1, 1, 1)], ids=str)
def test_conv_tc_generic_kernel_on_3x3(case):
    """tile_n = -1 forces the generic per-tap kernel for shapes the slab kernel normally takes."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, k, s, act, ups = case
    g = torch.Generator().manual_seed(1)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, k, k), generator=g) / (Ci * k * k) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, s, act, 0.1, ups)
    y = ops.conv2d(_nhwc(x).cuda().bfloat16(), w, b, s, act, 0.1, ups, tile_n=-1)
    assert torch.allclose(_nchw(y[..., :Co].float().cpu()), ref, rtol=1.2e-2, atol=1.2e-2)


PAIR_CASES = [
    # (N, H, W, Cin, Cout, k, s, act, tile_n)   tile_n < 0 forces the generic kernel (-1: its own choice of N tile)
    (2, 20, 20, 64, 64, 1, 1, 1, -1),          # 7 M tiles: the last pair has a tile past the end
    (1, 20, 20, 512, 255, 1, 1, 0, -1),        # head-like: Cout 255, 8 K chunks
    (4, 80, 80, 256, 256, 1, 1, 1, -256),      # 200 M tiles on 74 pairs: several rounds, ring and TMEM phases wrap
    (4, 80, 80, 256, 256, 1, 1, 1, -128),      # two N tiles
    (2, 20, 20, 1024, 512, 1, 1, 1, -1),       # long K
    (2, 40, 40, 128, 64, 1, 1, 2, -64),        # N tile 64 -> 32 weight rows per CTA, LeakyReLU
    (2, 24, 40, 64, 128, 3, 2, 1, -1),         # stride-2 3x3 (element strides, zero-filled borders)
    (2, 20, 20, 128, 256, 3, 1, 1, -1),        # 3x3 stride 1 through the per-tap kernel
    (3, 12, 20, 96, 96, 3, 1, 1, -1),          # ragged spatial tiles, Cin not a multiple of 64
    (1, 8, 8, 64, 32, 1, 1, 1, -1),            # a single M tile: one CTA of the pair works on zeros
]


@pytest.mark.parametrize('case', PAIR_CASES, ids=[str(c) for c in PAIR_CASES])
def test_conv_tc_cta_pairs(case):
    """pair=1 runs the generic kernel as clusters of two CTAs (tcgen05.mma.cta_group::2, M = 256, each CTA loads half
    of the weight tile).  Same K order per output element as the single-CTA kernel -> bit-identical to it, and within
    the bf16 tolerance of the fp32 PyTorch reference."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, k, s, act, tn = case
    g = torch.Generator().manual_seed(3)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, k, k), generator=g) / (Ci * k * k) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, s, act, 0.1, 1)
    xd = _nhwc(x).cuda().bfloat16()
    y1 = ops.conv2d(xd, w, b, s, act, 0.1, 1, tile_n=tn, pair=-1)
    y2 = ops.conv2d(xd, w, b, s, act, 0.1, 1, tile_n=tn, pair=1)
    torch.cuda.synchronize()
    got = _nchw(y2[..., :Co].float().cpu())
    assert torch.allclose(got, ref, rtol=1.2e-2, atol=1.2e-2), float((got - ref).abs().max())
    assert torch.equal(y1, y2)
    # fp32 output (head logits) through the pair path
    y3 = ops.conv2d(xd, w, b, s, act, 0.1, 1, tile_n=tn if tn > -256 else -128, pair=1, out_dtype=torch.float32)
    assert torch.allclose(_nchw(y3[..., :Co].cpu()), ref, rtol=1.2e-2, atol=1.2e-2)


SLAB_PAIR_CASES = [
    # (N, H, W, Cin, Cout, act, tile_n): 3x3 stride-1 layers through the slab kernel (tile_n >= 0)
    (1, 160, 160, 64, 64, 1, 0),          # the smem-operand-bound N = 64 layer class the pairs are for; many rounds
    (2, 40, 40, 128, 128, 1, 0),          # N = 128 -> 64 weight rows per CTA
    (1, 80, 80, 64, 192, 1, 64),          # three N tiles
    (1, 40, 40, 32, 32, 2, 0),            # Cin 32 (SW64), N = 32 -> 16 weight rows per CTA, LeakyReLU
    (3, 12, 20, 96, 96, 1, 0),            # ragged tiles, odd number of M tiles: the last pair has a tile past the end
    (2, 24, 24, 512, 64, 1, 0),           # 8 K chunks
    (1, 8, 8, 64, 64, 1, 0),              # a single tile
]


@pytest.mark.parametrize('case', SLAB_PAIR_CASES, ids=[str(c) for c in SLAB_PAIR_CASES])
def test_conv3_slab_cta_pairs(case):
    """k_conv3_tc<PAIR>: the slab kernel as clusters of two CTAs (cta_group::2, M = 256, half of every weight tile per
    CTA): bit-identical to the single-CTA slab kernel, within bf16 tolerance of the fp32 PyTorch reference."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, act, tn = case
    g = torch.Generator().manual_seed(5)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, 3, 3), generator=g) / (Ci * 9) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, 1, act, 0.1, 1)
    xd = _nhwc(x).cuda().bfloat16()
    y1 = ops.conv2d(xd, w, b, 1, act, 0.1, 1, tile_n=tn, pair=-1)
    y2 = ops.conv2d(xd, w, b, 1, act, 0.1, 1, tile_n=tn, pair=1)
    torch.cuda.synchronize()
    got = _nchw(y2[..., :Co].float().cpu())
    assert torch.allclose(got, ref, rtol=1.2e-2, atol=1.2e-2), float((got - ref).abs().max())
    assert torch.equal(y1, y2)


WIDE_CASES = [
    # (N, H, W, Cin, Cout, act, tile_n): 3x3 stride-1 layers through the wide-slab kernel (YB_C3_WIDE=1 takes every shape)
    (1, 160, 160, 64, 64, 1, 0),          # N = 64, two stacked sub-tiles, many rounds
    (2, 40, 40, 128, 128, 1, 0),          # N = 128, two K chunks
    (1, 80, 80, 64, 192, 1, 64),          # three N tiles of 64
    (3, 12, 20, 96, 96, 1, 0),            # ragged tiles in both directions, Cin 96: the second K chunk is half out of bounds
    (2, 24, 24, 512, 64, 1, 0),           # 8 K chunks: slab and weight rings wrap several times per tile
    (1, 8, 8, 64, 64, 1, 0),              # a single tile, mostly halo
    (2, 33, 17, 128, 255, 2, 0),          # odd sizes, Cout 255 (head-like), LeakyReLU
    (1, 20, 20, 256, 256, 0, 0),          # the small map the rule declines, no activation
]


@pytest.mark.parametrize('case', WIDE_CASES, ids=[str(c) for c in WIDE_CASES])
def test_conv3_wide_slab_equals_slab_kernel(case, monkeypatch):
    """k_conv3w_tc (one activation slab per K chunk, filter taps as descriptor offsets into rows of 10 pixels, group stride
    1280 B) against k_conv3_tc (one slab per filter column): the MMAs accumulate in the same order, so the outputs must
    be bit-identical; both within bf16 tolerance of the fp32 PyTorch reference."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, act, tn = case
    g = torch.Generator().manual_seed(11)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, 3, 3), generator=g) / (Ci * 9) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    ref = _ref_conv(x, w, b, 1, act, 0.1, 1)
    xd = _nhwc(x).cuda().bfloat16()
    monkeypatch.setenv('YB_C3_WIDE', '0')
    y_old = ops.conv2d(xd, w, b, 1, act, 0.1, 1, tile_n=tn, pair=-1)
    monkeypatch.setenv('YB_C3_WIDE', '1')
    y_new = ops.conv2d(xd, w, b, 1, act, 0.1, 1, tile_n=tn, pair=-1)
    torch.cuda.synchronize()
    got = _nchw(y_new[..., :Co].float().cpu())
    assert torch.allclose(got, ref, rtol=1.2e-2, atol=1.2e-2), float((got - ref).abs().max())
    assert torch.equal(y_old[..., :Co], y_new[..., :Co])


POOL_CASES = [
    # (N, H, W, Cin, Cout, k, s, act, tile_n, pair)
    (2, 40, 40, 64, 64, 1, 1, 1, 0, 0),         # 1x1, one 64-channel group, several tiles per image
    (1, 80, 80, 128, 256, 1, 1, 1, 0, 0),       # N tile 256: four channel groups per tile
    (2, 20, 20, 1024, 256, 1, 1, 1, 0, 1),      # CTA pairs + pool
    (1, 22, 38, 64, 96, 1, 1, 2, 0, 0),         # ragged tiles (W, H not multiples of the tile), Cout 96, LeakyReLU
    (1, 21, 37, 64, 64, 1, 1, 1, 0, 0),         # odd H and W: floor mode drops the last row / column
    (2, 40, 40, 64, 128, 3, 1, 1, -1, 0),       # 3x3 through the generic kernel
    (1, 40, 40, 64, 64, 3, 2, 1, -1, 0),        # stride-2 conv, pooled again
]


@pytest.mark.parametrize('case', POOL_CASES, ids=[str(c) for c in POOL_CASES])
def test_conv_tc_fused_maxpool(case):
    """yb_conv_desc.pool: the conv kernel also writes MaxPool2d(2, 2) of its output (MP, models/common.py:30-36) from the
    staged tile.  The full-resolution output must equal the unfused conv bit for bit, the pooled one the separate pool
    kernel applied to it (max of the same bf16 values), and both F.max_pool2d of the reference."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, k, s, act, tn, pair = case
    g = torch.Generator().manual_seed(7)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, k, k), generator=g) / (Ci * k * k) ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    xd = _nhwc(x).cuda().bfloat16()
    y_plain = ops.conv2d(xd, w, b, s, act, 0.1, 1, tile_n=tn, pair=pair)
    ho, wo = y_plain.shape[1], y_plain.shape[2]
    pooled = torch.full((N, ho // 2, wo // 2, y_plain.shape[3]), -7.0, dtype=torch.bfloat16, device='cuda')
    y = ops.conv2d(xd, w, b, s, act, 0.1, 1, tile_n=tn, pair=pair, pool_out=pooled)
    torch.cuda.synchronize()
    assert torch.equal(y, y_plain)
    want = ops.maxpool2x2(y_plain[..., :Co].contiguous())
    assert torch.equal(pooled[..., :Co], want)
    ref = F.max_pool2d(_ref_conv(x, w, b, s, act, 0.1, 1), 2, 2)
    assert torch.allclose(_nchw(pooled[..., :Co].float().cpu()), ref, rtol=1.2e-2, atol=1.2e-2)


RES_CASES = [
    # (N, H, W, Cin, Cout, act, pair, pool)
    (2, 40, 40, 128, 160, 1, 0, False),      # e6e width 160: five 32-column chunks, flat 1x1
    (1, 20, 20, 1280, 1280, 1, 1, False),    # CTA pairs, five N tiles of 256
    (2, 40, 40, 64, 80, 1, 0, True),         # residual + fused max-pool (DownC behind a Shortcut), Cout 80: partial chunk
    (1, 22, 38, 64, 96, 2, 0, True),         # ragged 2-D tiles, LeakyReLU
    (3, 9, 13, 32, 64, 0, 0, False),         # odd sizes, no activation
]


@pytest.mark.parametrize('case', RES_CASES, ids=[str(c) for c in RES_CASES])
def test_conv_tc_fused_residual(case):
    """yb_conv_desc.res: y = act(conv(x)) + res (Shortcut, models/common.py:80-86) inside the conv epilogue: the sum is
    formed in fp32 and rounded to bf16 once (the separate add rounds the conv output first), so it is compared with the
    fp32 reference at the bf16 tolerance and with conv -> yb_add at one bf16 ulp of the larger of |conv| and |sum|; the pooled output must be the max-pool
    of the fused output bit for bit."""
    from yolo_series_b200 import ops
    N, H, W, Ci, Co, act, pair, pool = case
    g = torch.Generator().manual_seed(9)
    x = torch.randn((N, Ci, H, W), generator=g).bfloat16().float()
    w = (torch.randn((Co, Ci, 1, 1), generator=g) / Ci ** 0.5).bfloat16().float()
    b = torch.randn((Co,), generator=g)
    r = torch.randn((N, Co, H, W), generator=g).bfloat16().float()
    ref = _ref_conv(x, w, b, 1, act, 0.1, 1) + r
    xd, rd = _nhwc(x).cuda().bfloat16(), _nhwc(r).cuda().bfloat16()
    cb = (Co + 7) // 8 * 8
    if cb != Co:
        rd = torch.nn.functional.pad(rd, (0, cb - Co)).contiguous()
    pooled = torch.zeros((N, H // 2, W // 2, cb), dtype=torch.bfloat16, device='cuda') if pool else None
    y = ops.conv2d(xd, w, b, 1, act, 0.1, 1, pair=pair, res=rd, pool_out=pooled)
    torch.cuda.synchronize()
    got = _nchw(y[..., :Co].float().cpu())
    assert torch.allclose(got, ref, rtol=1.2e-2, atol=1.2e-2), float((got - ref).abs().max())
    conv = ops.conv2d(xd, w, b, 1, act, 0.1, 1, pair=pair)[..., :Co].contiguous()
    y_sep = ops.add(conv, rd[..., :Co].contiguous())
    d = (y[..., :Co].float() - y_sep.float()).abs()
    # the separate path rounds the conv output (<= half an ulp of |conv|) and then the sum (<= half an ulp of |sum|)
    scale = torch.maximum(conv.float().abs(), y_sep.float().abs()).clamp(min=1e-2)
    assert float((d / scale).max()) <= 2 ** -7
    if pool:
        assert torch.equal(pooled[..., :Co], ops.maxpool2x2(y[..., :Co].contiguous()))
