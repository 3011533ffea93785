"""tcgen05 / TMEM GEMM behind launchLinearGemm for fp16 with more than 8 token rows
(csrc/gemm_tcgen05.cu): the swapped weight-streaming mode (9..256 tokens, split-K) and the normal
prefill mode (> 256 tokens), LLaMA-7B / 70B-TP shaped and ragged sizes, against the CPU oracle's
linear (fp32 accumulate) on the same fp16-rounded inputs. Tolerance 2e-2 (fp16, north_star);
observed errors are ~1e-3 (output rounding)."""
import numpy as np
import pytest
import torch

from conftest import assert_close, rel_err
from oracle import cpu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(pkg):
    assert torch.cuda.is_available()
    return pkg.ops


def case(ops, m, n, k, seed, bias=False):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-1, 1, (m, k)).astype(np.float16)
    w = (rng.standard_normal((n, k)) * 0.05).astype(np.float16)
    b = (rng.standard_normal(n) * 0.5).astype(np.float16) if bias else None
    want = np.zeros((m, n), np.float32)
    cpu.linear(x.astype(np.float32), w.astype(np.float32), want, True, None if b is None else b.astype(np.float32))
    y = torch.full((m, n), float("nan"), dtype=torch.float16, device="cuda")
    ops.launchLinearGemm(torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda(), y, trans_b=True,
                         bias=None if b is None else torch.from_numpy(b).cuda())
    torch.cuda.synchronize()
    return y.float().cpu().numpy(), want


@pytest.mark.parametrize("m,n,k", [(32, 4096, 4096), (9, 12288, 4096), (16, 4096, 11008), (32, 22016, 4096),
                                   (64, 1024, 512), (100, 384, 1000), (128, 4096, 1024), (200, 520, 264),
                                   (256, 2048, 2048), (33, 136, 72), (17, 8, 64)])
def test_swapped_weight_streaming_mode(ops, m, n, k):
    got, want = case(ops, m, n, k, m * 7 + n)
    assert_close(got, want, "fp16", f"tcgen05 swapped {m}x{n}x{k}")
    assert rel_err(got, want) < 4e-3


@pytest.mark.parametrize("m,n,k", [(257, 256, 64), (512, 4096, 4096), (1000, 520, 264), (2048, 12288, 4096),
                                   (384, 96, 128), (300, 40, 2048), (4096, 4096, 11008),
                                   # ragged rows / columns / K across several bands of the banded tile order
                                   (4000, 2504, 328), (4100, 2500, 72), (8192, 1024, 512)])
def test_normal_prefill_mode(ops, m, n, k):
    got, want = case(ops, m, n, k, m + n * 3)
    assert_close(got, want, "fp16", f"tcgen05 normal {m}x{n}x{k}")
    assert rel_err(got, want) < 4e-3


@pytest.mark.parametrize("m,n,k", [(200, 15400, 264), (161, 15364, 72), (256, 22016, 4096)])
def test_one_row_tile_of_the_pair_kernel(ops, m, n, k):
    """161..256 token rows against >= 60 column tiles: routed to the two-CTA kernel with one (partly empty) row tile"""
    got, want = case(ops, m, n, k, m + n)
    assert_close(got, want, "fp16", f"tcgen05 pair, one row tile {m}x{n}x{k}")
    assert rel_err(got, want) < 4e-3


@pytest.mark.parametrize("m,n,k", [(2048, 12288, 4096), (4000, 2504, 328), (4100, 2500, 72), (8192, 1024, 512)])
def test_quad_cluster_multicast_variant(ops, m, n, k, monkeypatch):
    """B200_GEMM2_QUAD=1: clusters of four CTAs, the A tile multicast between the two MMA pairs (opt-in, gemm2_tcgen05.cu);
    must give the pair kernel's result bit for bit (same MMAs in the same order) and match the oracle"""
    got0, want = case(ops, m, n, k, m + n * 3)
    monkeypatch.setenv("B200_GEMM2_QUAD", "1")
    got1, _ = case(ops, m, n, k, m + n * 3)
    assert_close(got1, want, "fp16", f"tcgen05 quad {m}x{n}x{k}")
    assert np.array_equal(got0, got1)


@pytest.mark.parametrize("m,n,k", [(8192, 12288, 512), (5000, 5120, 328), (19000, 1024, 72), (2048, 12288, 4096)])
def test_wide_pair_tile_equals_the_256x256_tile(ops, m, n, k, monkeypatch):
    """B200_GEMM2_WIDE=1: the 256 x 512 pair tile (gemm2w_tcgen05_kernel, opt-in), incl. shapes whose last wave is split
    into single column tiles; same MMAs in the same K order as the default 256 x 256 tile, so bit-identical, and both
    match the oracle"""
    got_p, want = case(ops, m, n, k, m + n * 3)
    monkeypatch.setenv("B200_GEMM2_WIDE", "1")
    got_w, _ = case(ops, m, n, k, m + n * 3)
    assert_close(got_w, want, "fp16", f"tcgen05 wide {m}x{n}x{k}")
    assert np.array_equal(got_w, got_p)


@pytest.mark.parametrize("m", [24, 700])
def test_bias_epilogue(ops, m):
    got, want = case(ops, m, 1024, 512, 5, bias=True)
    assert_close(got, want, "fp16", "tcgen05 + bias")


def test_linearity_at_full_size(ops):
    # size-independent property at the cfg3 shape (8192 x 12288 x 4096): W(x1 + x2) == W x1 + W x2
    # up to fp16 output rounding; a wrong tile / swizzle / accumulate flag breaks it grossly
    m, n, k = 8192, 12288, 4096
    g = torch.Generator(device="cuda").manual_seed(1)
    x1 = (torch.rand(m, k, device="cuda", generator=g) - 0.5).half()
    x2 = (torch.rand(m, k, device="cuda", generator=g) - 0.5).half()
    w = (torch.randn(n, k, device="cuda", generator=g) * 0.02).half()
    xs = (x1.float() + x2.float()).half()
    y1, y2, ys = (torch.empty(m, n, dtype=torch.float16, device="cuda") for _ in range(3))
    ops.launchLinearGemm(x1, w, y1, trans_b=True)
    ops.launchLinearGemm(x2, w, y2, trans_b=True)
    ops.launchLinearGemm(xs, w, ys, trans_b=True)
    err = (ys.float() - (y1.float() + y2.float())).abs().max().item() / ys.float().abs().max().item()
    assert err < 1e-2, err
    # and a sampled exact check against fp32 math on 64 random rows
    rows = torch.randint(0, m, (64,), device="cuda")
    ref = x1[rows].float() @ w.float().t()
    e2 = (y1[rows].float() - ref).abs().max().item() / ref.abs().max().item()
    assert e2 < 4e-3, e2


@pytest.mark.parametrize("m,inter,k", [(257, 128, 64), (512, 1024, 512), (1000, 384, 264), (2048, 11008, 4096)])
def test_gate_up_gemm_with_swiglu_epilogue(ops, m, inter, k):
    """b200_linear_swiglu (ffn.cpp:72-80 in one launch) against the oracle's linear -> SwiGLU with the
    reference's fp16 rounding points (GEMM output, silu(gate), product)."""
    rng = np.random.default_rng(m + inter)
    f16 = lambda a: a.astype(np.float16).astype(np.float32)
    x = f16(rng.uniform(-1, 1, (m, k)))
    w = f16(rng.standard_normal((2 * inter, k)) * 0.05)
    gu = np.zeros((m, 2 * inter), np.float32)
    cpu.linear(x, w, gu, True)
    gu = f16(gu).reshape(m, 2, inter)
    g = gu[:, 0]
    want = f16(f16(g / (1.0 + np.exp(-g, dtype=np.float32))) * gu[:, 1])
    out = torch.full((m, inter), float("nan"), dtype=torch.float16, device="cuda")
    ops.linear_swiglu(torch.from_numpy(x).cuda().half(), torch.from_numpy(w).cuda().half(), out)
    torch.cuda.synchronize()
    assert_close(out.float().cpu().numpy(), want, "fp16", f"gate_up + SwiGLU {m}x{inter}x{k}")


def test_gate_up_swiglu_rejects_decode_shapes(ops):
    x = torch.zeros(8, 256, dtype=torch.float16, device="cuda")
    w = torch.zeros(512, 256, dtype=torch.float16, device="cuda")
    with pytest.raises(RuntimeError, match="unsupported"):
        ops.linear_swiglu(x, w, torch.zeros(8, 256, dtype=torch.float16, device="cuda"))


@pytest.mark.parametrize("m,n,k", [(512, 256, 64), (1000, 520, 264), (2048, 4096, 11008)])
def test_linear_with_residual_epilogue(ops, m, n, k):
    """b200_linear_residual against the oracle's linear (rounded to fp16) followed by add_residual"""
    rng = np.random.default_rng(m + n)
    f16 = lambda a: a.astype(np.float16).astype(np.float32)
    x = f16(rng.uniform(-1, 1, (m, k)))
    w = f16(rng.standard_normal((n, k)) * 0.05)
    res = f16(rng.uniform(-2, 2, (m, n)))
    y = np.zeros((m, n), np.float32)
    cpu.linear(x, w, y, True)
    want = f16(f16(y) + res)
    out = torch.full((m, n), float("nan"), dtype=torch.float16, device="cuda")
    ops.linear_residual(torch.from_numpy(x).cuda().half(), torch.from_numpy(w).cuda().half(),
                        torch.from_numpy(res).cuda().half(), out)
    torch.cuda.synchronize()
    assert_close(out.float().cpu().numpy(), want, "fp16", f"linear + residual {m}x{n}x{k}")


@pytest.mark.parametrize("knob", ["B200_GEMM2_WIDE", "B200_GEMM2_QUAD"])
def test_opt_in_tile_shapes_with_swiglu_and_residual_epilogues(ops, knob, monkeypatch):
    """the opt-in 256 x 512 pair tile / four-CTA multicast cluster also carry the SwiGLU and the residual epilogue: their
    outputs must equal the default 256 x 256 tile's bit for bit (same MMAs, same K order, same rounding points); the default
    is pinned to the oracle by the tests above"""
    g = torch.Generator(device="cpu").manual_seed(3)
    m, inter, k = 2048, 11008, 1024      # 86 GLU column tiles: 344 wide units / quad units
    x = (torch.rand(m, k, generator=g) * 2 - 1).half().cuda()
    w = (torch.randn(2 * inter, k, generator=g) * 0.05).half().cuda()
    m2, n2, k2 = 4096, 4096, 520          # 16 x 8 units, ragged K
    x2 = (torch.rand(m2, k2, generator=g) * 2 - 1).half().cuda()
    w2 = (torch.randn(n2, k2, generator=g) * 0.05).half().cuda()
    res = (torch.rand(m2, n2, generator=g) * 4 - 2).half().cuda()

    def run():
        o1 = torch.full((m, inter), float("nan"), dtype=torch.float16, device="cuda")
        o2 = torch.full((m2, n2), float("nan"), dtype=torch.float16, device="cuda")
        ops.linear_swiglu(x, w, o1)
        ops.linear_residual(x2, w2, res, o2)
        torch.cuda.synchronize()
        return o1, o2

    a1, a2 = run()
    monkeypatch.setenv(knob, "1")
    b1, b2 = run()
    assert not torch.isnan(b1).any() and not torch.isnan(b2).any()
    assert torch.equal(a1, b1), f"{knob}: SwiGLU epilogue differs from the default tile"
    assert torch.equal(a2, b2), f"{knob}: residual epilogue differs from the default tile"
