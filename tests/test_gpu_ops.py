"""Unit-level parity of the tcgen05 implicit-GEMM kernel through the C ABI (dmb_op_*),
against torch CPU fp32 on bf16-rounded operands.  Tolerance: the kernel accumulates in
fp32 and rounds the result to bf16 once, so |err| <= 2^-8 |ref| + small accumulation-order
term; stated per test."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import bf16_round, max_abs, rel_l2, report

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from dmb200 import context_for
    return context_for(torch.device("cuda", 0), "bf16")


def _check(name, got, ref, rel_tol=4e-3):
    r, m = rel_l2(got, ref), max_abs(got, ref)
    report(name, {"rel_l2": r, "max_abs": m, "ref_max": float(ref.abs().max())})
    assert r < rel_tol, (name, r, m)
    # elementwise: one bf16 rounding of the result + fp32 accumulation noise
    tol = 2.0 ** -7 * ref.abs().cpu() + 1e-3 * float(ref.abs().max())
    assert bool(((got.cpu() - ref.cpu()).abs() <= tol).all()), (name, m)


@pytest.mark.parametrize("rows,K,N,act,use_bias", [
    (300, 64, 64, 0, True), (1000, 128, 384, 0, True), (130, 256, 256, 1, True),
    (128, 64, 192, 0, False), (5, 64, 64, 1, True), (4097, 256, 768, 0, True), (257, 512, 128, 0, True),
    (40000, 64, 192, 0, True), (30000, 128, 128, 1, True)])  # the last two are large enough for CTA-pair tiles
def test_linear(ctx, rows, K, N, act, use_bias):
    g = torch.Generator().manual_seed(rows + K + N)
    x = bf16_round(torch.randn((rows, K), generator=g))
    w = bf16_round(torch.randn((N, K), generator=g) / K ** 0.5)
    b = torch.randn((N,), generator=g) if use_bias else None
    ref = F.linear(x, w, b)
    if act:
        ref = F.gelu(ref)
    got = ctx.op_linear(x.cuda(), w.cuda(), None if b is None else b.cuda(), act)
    torch.cuda.synchronize()
    _check(f"linear_{rows}x{K}x{N}", got, ref)


@pytest.mark.parametrize("B,Cin,Cout,H,W", [
    (2, 64, 64, 28, 28), (3, 128, 128, 14, 14), (2, 256, 512, 3, 3), (1, 512, 256, 7, 7),
    (2, 64, 3, 12, 12), (1, 64, 128, 1, 1), (5, 128, 64, 7, 5), (16, 256, 256, 14, 14),
    (64, 128, 128, 28, 28), (96, 64, 256, 14, 14), (40, 128, 64, 28, 28),  # large: CTA-pair (cta_group::2) tiles
    # small maps at batch >= 256: pixel-major tiles (M over samples, K over the in-image neighbours)
    (256, 128, 128, 7, 7), (300, 256, 512, 3, 3), (257, 64, 64, 7, 7), (260, 512, 256, 5, 7), (512, 64, 128, 2, 2),
    (256, 128, 64, 1, 1)])
def test_conv3x3(ctx, B, Cin, Cout, H, W):
    g = torch.Generator().manual_seed(B * 1000 + Cin + Cout + H)
    x = bf16_round(torch.randn((B, Cin, H, W), generator=g))
    w = bf16_round(torch.randn((Cout, Cin, 3, 3), generator=g) / (9 * Cin) ** 0.5)
    ref = F.conv2d(x, w, padding=1)
    got = ctx.op_conv3x3(x.cuda(), w.cuda())
    torch.cuda.synchronize()
    _check(f"conv3x3_{B}x{Cin}x{Cout}x{H}x{W}", got, ref)


@pytest.mark.parametrize("B,Cin,Cout,H,W", [(2, 64, 64, 5, 7), (1, 256, 256, 28, 28), (3, 128, 128, 1, 1)])
def test_conv_transpose(ctx, B, Cin, Cout, H, W):
    g = torch.Generator().manual_seed(B + Cin + H)
    x = bf16_round(torch.randn((B, Cin, H, W), generator=g))
    w = bf16_round(torch.randn((Cin, Cout, 4, 4), generator=g) / (4 * Cin) ** 0.5)
    b = torch.randn((Cout,), generator=g)
    ref = F.conv_transpose2d(x, w, b, stride=2, padding=1)
    got = ctx.op_conv_transpose4x4(x.cuda(), w.cuda(), b.cuda())
    torch.cuda.synchronize()
    _check(f"convT_{B}x{Cin}x{Cout}x{H}x{W}", got, ref)


def test_tf32_mode_linear_and_conv():
    from dmb200 import context_for
    c = context_for(torch.device("cuda", 0), "tf32")
    g = torch.Generator().manual_seed(5)
    x = torch.randn((333, 128), generator=g)
    w = torch.randn((256, 128), generator=g) / 128 ** 0.5
    ref = F.linear(x, w)
    got = c.op_linear(x.cuda(), w.cuda())
    r = rel_l2(got, ref)
    report("tf32_linear", {"rel_l2": r})
    assert r < 2e-3  # tf32 operands: 10-bit mantissa
    x = torch.randn((2, 64, 14, 14), generator=g)
    w = torch.randn((128, 64, 3, 3), generator=g) / (9 * 64) ** 0.5
    ref = F.conv2d(x, w, padding=1)
    got = c.op_conv3x3(x.cuda(), w.cuda())
    r = rel_l2(got, ref)
    report("tf32_conv", {"rel_l2": r})
    assert r < 2e-3


def test_bad_arguments_report_errors(ctx):
    from dmb200 import NativeError
    with pytest.raises(NativeError):
        ctx.op_linear(torch.zeros((4, 48), device="cuda"), torch.zeros((8, 48), device="cuda"))  # K not a chunk multiple
