"""tcgen05/TMA GEMM tile (csrc/tc_gemm.cuh) against a plain PyTorch fp32 reference of the same op
on the same fp16 operands.  Tolerances: fp32 accumulation of fp16 products -> 1e-3 relative to the
row scale for fp32 outputs, one fp16 ulp (2^-10 relative) more for fp16 outputs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _gemm(a, b, mode, alpha=1.0, bias=None, relu=0, want_t=False, n_atoms=51, z0=-25.0, dz=1.0,
          q_scale=1.0):
    import torch
    from hanabi_agents_b200 import _lib
    M, K = a.shape
    N = b.shape[0]
    dev = a.device
    out_t = q = None
    if mode == 0:
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device=dev)
        if want_t:
            out_t = torch.full((N, M + 8 - M % 8 if M % 8 else M), float("nan"), dtype=torch.float16, device=dev)
    else:
        out = torch.full((M, N), float("nan"), dtype=torch.float32, device=dev)
    if mode == 2:
        q = torch.full((M, N // 64), float("nan"), dtype=torch.float32, device=dev)
    _lib.check(_lib.lib().hb_gemm_f16(
        a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), M, N, K, mode, alpha,
        None if bias is None else bias.data_ptr(), relu, out.data_ptr(), out.stride(0),
        None if out_t is None else out_t.data_ptr(), 0 if out_t is None else out_t.stride(0),
        None if q is None else q.data_ptr(), n_atoms, z0, dz, q_scale, _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    return out, out_t, q


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (128, 128, 256), (256, 512, 768), (32, 64, 128),
                                   (300, 200, 72), (1000, 1280, 512), (65, 8, 8)])
def test_gemm_f32_out(M, N, K):
    import torch
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    a = torch.randn(M, K, device="cuda", generator=g).half()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).half()
    bias = torch.randn(N, device="cuda", generator=g)
    out, _, _ = _gemm(a, b, 1, alpha=0.5, bias=bias)
    ref = 0.5 * (a.float() @ b.float().t()) + bias
    assert torch.isfinite(out).all()
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("M,N,K", [(256, 512, 768), (100, 136, 200)])
def test_gemm_f16_relu_and_transpose(M, N, K):
    import torch
    g = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randn(M, K, device="cuda", generator=g).half()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).half()
    bias = torch.randn(N, device="cuda", generator=g)
    out, out_t, _ = _gemm(a, b, 0, bias=bias, relu=1, want_t=True)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    torch.testing.assert_close(out.float(), ref, rtol=2e-3, atol=2e-3)
    assert torch.equal(out_t[:, :M], out.t())


def test_gemm_bits_exact():
    """0/1 operands (observation bits) and small integer weights: every product and sum is exact."""
    import torch
    g = torch.Generator(device="cuda").manual_seed(5)
    a = (torch.rand(384, 704, device="cuda", generator=g) < 0.15).half()
    b = torch.randint(-3, 4, (256, 704), device="cuda", generator=g).half()
    out, _, _ = _gemm(a, b, 1)
    assert torch.equal(out, a.float() @ b.float().t())


def test_gemm_categorical_head():
    import torch
    M, A, K = 200, 20, 512
    g = torch.Generator(device="cuda").manual_seed(9)
    a = torch.randn(M, K, device="cuda", generator=g).half()
    b = (torch.randn(A * 64, K, device="cuda", generator=g) / K ** 0.5).half()
    bias = torch.randn(A * 64, device="cuda", generator=g)
    logits, _, q = _gemm(a, b, 2, bias=bias, n_atoms=51, z0=-25.0, dz=1.0, q_scale=1.0 / 51)
    ref = (a.float() @ b.float().t() + bias).view(M, A, 64)
    torch.testing.assert_close(logits.view(M, A, 64), ref, rtol=1e-3, atol=1e-3)
    z = torch.linspace(-25, 25, 51, device="cuda")
    qref = (torch.softmax(ref[:, :, :51], -1) * z).mean(-1)
    torch.testing.assert_close(q, qref, rtol=1e-3, atol=1e-3)
