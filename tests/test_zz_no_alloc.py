"""dmb_kl_per_sample_static: the allocation-free entry point (static __device__ scratch) against
the existing one (dmb_vae_encode's kl_sample), through the C ABI, on two calls in a row.

Collected last on purpose; passes alone in a fresh process and
after the other tests (it builds its own modules and buffers)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _call(lib, mu, logvar, out, B=None, stream=None):
    B = mu.shape[0] if B is None else B
    stream = torch.cuda.current_stream(DEV).cuda_stream if stream is None else stream
    return lib.dmb_kl_per_sample_static(0, C.c_void_p(mu.data_ptr()), C.c_void_p(logvar.data_ptr()), B,
                                        mu.numel() // mu.shape[0], C.c_float(224.0 * 224.0),
                                        C.c_void_p(out.data_ptr()), C.c_void_p(stream))


def test_static_scratch_kl_matches_vae_encode():
    from dmb200 import context_for, load_library
    from dmb200.native import KL_STATIC_MAX_B
    from models.vae import VAE
    torch.manual_seed(0)
    vae = VAE().eval().to(DEV)
    ctx = context_for(DEV)
    ctx.ensure_vae_enc(vae)
    lib = load_library()
    g = torch.Generator().manual_seed(5)
    batches = []
    for B in (5, 3):                                   # the second call is SMALLER: stale partials of
        img = torch.rand((B, 3, 224, 224), generator=g).to(DEV)   # the first must not leak into it
        eps = torch.randn((B, 4, 28, 28), generator=g).to(DEV)
        _, kl, mu, logvar = ctx.vae_encode(img, eps, want_moments=True)
        batches.append((kl, mu, logvar, torch.full((B,), float("nan"), device=DEV)))
    torch.cuda.synchronize()
    free0 = None
    for i, (kl, mu, logvar, out) in enumerate(batches):   # two calls in a row, same stream
        if i == 1:
            free0 = torch.cuda.mem_get_info(DEV)[0]
        rc = _call(lib, mu, logvar, out)
        assert rc == 0, lib.dmb_last_error_string()
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(DEV)[0] >= free0      # the warmed-up call took no device memory
    for kl, mu, logvar, out in batches:
        assert torch.isfinite(out).all()
        # float64 sums of the same terms in another order, rounded once to fp32
        torch.testing.assert_close(out, kl, rtol=1e-5, atol=1e-7)
    # and against the definition, in float64 on the host
    kl, mu, logvar, out = batches[0]
    m, lv = mu.double().cpu(), logvar.double().cpu()
    want = 0.5 * (torch.exp(lv.float()).double() + m * m - 1.0 - lv).flatten(1).sum(1) / (224.0 * 224.0)
    torch.testing.assert_close(out.double().cpu(), want, rtol=1e-5, atol=1e-7)
    # the wrapper, on a side stream
    from dmb200.native import kl_per_sample_static
    s = torch.cuda.Stream(DEV)
    out2 = torch.empty_like(out)
    s.wait_stream(torch.cuda.current_stream(DEV))
    with torch.cuda.stream(s):
        kl_per_sample_static(mu, logvar, out2)
    s.synchronize()
    assert torch.equal(out2, out)                        # fixed summation order: same bits
    # beyond the static array: refused, nothing launched, clear message
    big_mu = torch.zeros((KL_STATIC_MAX_B + 1, 4), device=DEV)
    big_out = torch.zeros((KL_STATIC_MAX_B + 1,), device=DEV)
    rc = _call(lib, big_mu, big_mu, big_out)
    assert rc != 0
    msg = lib.dmb_last_error_string().decode()
    assert "DMB_KL_STATIC_MAX_B" in msg and str(KL_STATIC_MAX_B) in msg, msg
    rc = _call(lib, big_mu, big_mu, big_out, B=KL_STATIC_MAX_B)   # exactly the maximum is accepted
    assert rc == 0, lib.dmb_last_error_string()
    torch.cuda.synchronize()
    assert float(big_out[:KL_STATIC_MAX_B].abs().max()) == 0.0    # exp(0) + 0 - 1 - 0
