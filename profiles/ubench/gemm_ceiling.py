"""Ceiling of the cta_group::1 128x128 tcgen05 tile configuration for the operand layouts the step uses."""
import sys, torch
sys.path.insert(0, "flow-q-learning_b200")
from fqlpop import _lib
dev = "cuda"
for (M, N, K, P, n_sub, a_mn, b_mn) in [(8192, 8192, 8192, 1, 1, 0, 0), (8192, 8192, 8192, 1, 1, 0, 1), (8192, 8192, 8192, 1, 1, 1, 1),
                                         (8192, 512, 8192, 1, 4, 0, 0), (8192, 512, 8192, 1, 4, 0, 1), (18944, 512, 512, 1, 4, 0, 1), (18944, 512, 512, 1, 4, 0, 0)]:
    A = (torch.randn(P, M, K, device=dev) * 0.1).to(torch.bfloat16)   # [M][K] (k-major) or read as [K][M] when a_mn
    B = (torch.randn(P, N, K, device=dev) * 0.1).to(torch.bfloat16)
    out = torch.empty(P, M, N, device=dev, dtype=torch.bfloat16)
    kw = dict(M=M, N=N, K=K, P=P, a_mn=a_mn, b_mn=b_mn, n_sub=n_sub, A=A, a_ld=(M if a_mn else K), a_ms=M * K, B=B,
              b_ld=(N if b_mn else K), b_ms=N * K, ln_eps=1e-6, epi_kind=_lib.EPI_STORE_BF16, act=0, out_h=out)
    for _ in range(3): _lib.gemm_test(**kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    it = 10
    e0.record()
    for _ in range(it): _lib.gemm_test(**kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / it
    print(f"M={M} N={N} K={K} n_sub={n_sub} a_mn={a_mn} b_mn={b_mn}: {ms*1e3:.1f} us, {2.0*P*M*N*K/ms/1e9:.1f} TFLOP/s")
