"""tcgen05 grouped-GEMM engine vs a plain PyTorch fp32 reference of the same op (GPU only).

Tolerances: operands are bf16 (exactly representable inputs are generated), accumulation is fp32, so
plain products match to fp32 round-off (rtol 1e-4 on K<=1024 sums); bf16 outputs to 1 bf16 ulp (2^-8 rel);
GELU uses tanh.approx (abs err ~5e-4 on tanh) -> atol 3e-3 on activations.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _bf(x):
    return x.to(torch.bfloat16)


def _gelu_tanh(z):
    return torch.nn.functional.gelu(z, approximate="tanh")


def _gelu_tanh_grad(z):
    k0, k1 = 0.7978845608028654, 0.044715
    u = k0 * (z + k1 * z ** 3)
    t = torch.tanh(u)
    return 0.5 * (1 + t) + 0.5 * z * (1 - t * t) * k0 * (1 + 3 * k1 * z * z)


def _operands(P, M, N, K, a_mn, b_mn, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = _bf(torch.randn(P, M, K, device="cuda", generator=g) * 0.5)
    B = _bf(torch.randn(P, N, K, device="cuda", generator=g) * 0.5)
    ref = torch.einsum("pmk,pnk->pmn", A.float(), B.float())
    A_dev = A.transpose(1, 2).contiguous() if a_mn else A.contiguous()
    B_dev = B.transpose(1, 2).contiguous() if b_mn else B.contiguous()
    return A_dev, B_dev, ref


def _common(P, M, N, K, a_mn, b_mn, A, B, n_sub=1):
    return dict(M=M, N=N, K=K, P=P, a_mn=int(a_mn), b_mn=int(b_mn), n_sub=n_sub,
                A=A, a_ld=A.shape[2], a_ms=A.shape[1] * A.shape[2],
                B=B, b_ld=B.shape[2], b_ms=B.shape[1] * B.shape[2], ln_eps=1e-6)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(256, 512, 512), (128, 128, 64), (384, 256, 1024)])
def test_store_f32_all_majors(a_mn, b_mn, M, N, K):
    from fqlpop import _lib
    P = 2
    A, B, ref = _operands(P, M, N, K, a_mn, b_mn)
    out = torch.full((P, M, N), float("nan"), device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, a_mn, b_mn, A, B), epi_kind=_lib.EPI_STORE_F32, out_f32=out)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)


def test_store_f32_nsub4_and_n64():
    from fqlpop import _lib
    P, M, N, K = 3, 256, 512, 256
    A, B, ref = _operands(P, M, N, K, 0, 0, seed=1)
    out = torch.full((P, M, N), float("nan"), device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 0, 0, A, B, n_sub=4), epi_kind=_lib.EPI_STORE_F32, out_f32=out)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)
    # N = 64 with an MN-major B (layer-0 input-gradient shape)
    P, M, N, K = 2, 256, 64, 512
    A, B, ref = _operands(P, M, N, K, 0, 1, seed=2)
    out = torch.full((P, M, N), float("nan"), device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 0, 1, A, B), epi_kind=_lib.EPI_STORE_F32, out_f32=out)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)
    # wgrad shape: both MN-major, N = 64
    P, M, N, K = 2, 512, 64, 256
    A, B, ref = _operands(P, M, N, K, 1, 1, seed=3)
    out = torch.full((P, M, N), float("nan"), device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 1, 1, A, B), epi_kind=_lib.EPI_STORE_F32, out_f32=out)
    torch.cuda.synchronize()
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-3)


def test_member_list_and_active_count():
    from fqlpop import _lib
    P, M, N, K = 4, 128, 128, 128
    A, B, ref = _operands(P, M, N, K, 0, 0, seed=4)
    out = torch.zeros((P, M, N), device="cuda")
    n_active = torch.tensor([2], dtype=torch.int32, device="cuda")
    members = torch.tensor([3, 1, 0, 0], dtype=torch.int32, device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 0, 0, A, B), epi_kind=_lib.EPI_STORE_F32, out_f32=out,
                   n_active=n_active, member_list=members)
    torch.cuda.synchronize()
    torch.testing.assert_close(out[3], ref[3], rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(out[1], ref[1], rtol=1e-4, atol=1e-3)
    assert out[0].abs().max().item() == 0 and out[2].abs().max().item() == 0


@pytest.mark.parametrize("act", [1, 2])
def test_fwd_bias_act(act):
    from fqlpop import _lib
    P, M, N, K = 2, 256, 512, 512
    A, B, ref = _operands(P, M, N, K, 0, 0, seed=5)
    ref = ref / math.sqrt(K) * 4
    B = _bf(B.float() / math.sqrt(K) * 4)
    ref = torch.einsum("pmk,pnk->pmn", A.float(), B.float())
    bias = torch.randn(P, N, device="cuda") * 0.1
    z = ref + bias[:, None, :]
    h_ref = _gelu_tanh(z) if act == 1 else torch.relu(z)
    gp_ref = _gelu_tanh_grad(z) if act == 1 else (z > 0).float()
    out_h = torch.zeros(P, M, N, device="cuda", dtype=torch.bfloat16)
    out_gp = torch.zeros_like(out_h)
    _lib.gemm_test(**_common(P, M, N, K, 0, 0, A, B), epi_kind=_lib.EPI_FWD, act=act, bias=bias,
                   out_h=out_h, out_gp=out_gp)
    torch.cuda.synchronize()
    torch.testing.assert_close(out_h.float(), h_ref, rtol=1e-2, atol=4e-3)
    if act == 1:
        torch.testing.assert_close(out_gp.float(), gp_ref, rtol=1e-2, atol=4e-3)


def test_fwd_layernorm():
    from fqlpop import _lib
    P, M, N, K = 2, 256, 512, 512
    A, B, _ = _operands(P, M, N, K, 0, 0, seed=6)
    B = _bf(B.float() / math.sqrt(K) * 4)
    acc = torch.einsum("pmk,pnk->pmn", A.float(), B.float())
    bias = torch.randn(P, N, device="cuda") * 0.1
    gamma = 1 + 0.1 * torch.randn(P, N, device="cuda")
    beta = 0.1 * torch.randn(P, N, device="cuda")
    g = _gelu_tanh(acc + bias[:, None, :])
    mean = g.mean(-1, keepdim=True)
    var = (g * g).mean(-1, keepdim=True) - mean * mean
    rstd = torch.rsqrt(var + 1e-6)
    xhat = (g - mean) * rstd
    h_ref = xhat * gamma[:, None, :] + beta[:, None, :]
    out_h = torch.zeros(P, M, N, device="cuda", dtype=torch.bfloat16)
    out_gp = torch.zeros_like(out_h)
    out_xhat = torch.zeros_like(out_h)
    out_rstd = torch.zeros(P, M, device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 0, 0, A, B, n_sub=4), epi_kind=_lib.EPI_FWD_LN, act=1, bias=bias,
                   gamma=gamma, beta=beta, out_h=out_h, out_gp=out_gp, out_xhat=out_xhat, out_rstd=out_rstd)
    torch.cuda.synchronize()
    torch.testing.assert_close(out_rstd, rstd.squeeze(-1), rtol=5e-3, atol=1e-3)
    torch.testing.assert_close(out_xhat.float(), xhat, rtol=1e-2, atol=2e-2)
    torch.testing.assert_close(out_h.float(), h_ref, rtol=1e-2, atol=2e-2)


def test_dgrad_with_bias_partials():
    from fqlpop import _lib
    P, M, N, K = 2, 256, 512, 512
    A, B, acc = _operands(P, M, N, K, 0, 1, seed=7)
    gp = _bf(torch.rand(P, M, N, device="cuda"))
    dz_ref = acc * gp.float()
    out_dz = torch.zeros(P, M, N, device="cuda", dtype=torch.bfloat16)
    dbp = torch.zeros(P, M // 128, N, device="cuda")
    _lib.gemm_test(**_common(P, M, N, K, 0, 1, A, B), epi_kind=_lib.EPI_DGRAD, gp=gp, out_h=out_dz, dbias_part=dbp)
    torch.cuda.synchronize()
    torch.testing.assert_close(out_dz.float(), dz_ref, rtol=1e-2, atol=1e-2)
    db_ref = out_dz.float().sum(1)
    torch.testing.assert_close(dbp.sum(1), db_ref, rtol=1e-4, atol=1e-2)


def test_dgrad_layernorm():
    from fqlpop import _lib
    P, M, N, K = 2, 256, 512, 512
    A, B, dh = _operands(P, M, N, K, 0, 1, seed=8)
    dh = dh / math.sqrt(K)
    B = _bf(B.float() / math.sqrt(K))
    dh = torch.einsum("pmk,pkn->pmn", A.float(), B.float()) if False else torch.einsum(
        "pmk,pnk->pmn", A.float(), B.transpose(1, 2).float())
    gamma = 1 + 0.1 * torch.randn(P, N, device="cuda")
    xhat = _bf(torch.randn(P, M, N, device="cuda"))
    gp = _bf(torch.rand(P, M, N, device="cuda"))
    rstd = 0.5 + torch.rand(P, M, device="cuda")
    dxh = dh * gamma[:, None, :]
    c1 = dxh.mean(-1, keepdim=True)
    c2 = (dxh * xhat.float()).mean(-1, keepdim=True)
    dg = rstd[..., None] * (dxh - c1 - xhat.float() * c2)
    dz_ref = dg * gp.float()
    out_dz = torch.zeros(P, M, N, device="cuda", dtype=torch.bfloat16)
    dbp = torch.zeros(P, M // 128, N, device="cuda")
    dgp = torch.zeros_like(dbp)
    dbetap = torch.zeros_like(dbp)
    _lib.gemm_test(**_common(P, M, N, K, 0, 1, A, B, n_sub=4), epi_kind=_lib.EPI_DGRAD_LN, gamma=gamma, gp=gp,
                   xhat=xhat, rstd=rstd, out_h=out_dz, dbias_part=dbp, dgamma_part=dgp, dbeta_part=dbetap)
    torch.cuda.synchronize()
    torch.testing.assert_close(out_dz.float(), dz_ref, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(dbetap.sum(1), dh.sum(1), rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dgp.sum(1), (dh * xhat.float()).sum(1), rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(dbp.sum(1), out_dz.float().sum(1), rtol=1e-4, atol=1e-2)
