// C-ABI of libfqlpop.so (declared in include/fqlpop.h).
#include "../../include/fqlpop.h"
#include "gemm.h"

#include <cstdio>
#include <cstring>
#include <string>

namespace {
thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return 1;
}
}  // namespace

namespace fqlpop {
void set_last_error(const std::string& m) { g_err = m; }
}

extern "C" const char* fqlpop_last_error(void) { return g_err.c_str(); }
extern "C" int fqlpop_version(void) { return 1; }

extern "C" int fqlpop_gemm_test(const fqlpop_gemm_test_t* d, void* stream) {
  using namespace fqlpop;
  if (!d) return fail("null desc");
  if (d->N % 64 || d->K % 64) return fail("N and K must be multiples of 64");
  if (d->n_sub != 1 && d->n_sub != 4) return fail("n_sub must be 1 or 4");
  if (cudaError_t e = gemm_init(); e != cudaSuccess) return fail(std::string("gemm_init: ") + cudaGetErrorString(e));
  static GemmLaunch L;  // ~8 KB: keep off the stack
  memset(&L, 0, sizeof(L));
  L.num_groups = 1;
  L.P = d->P;
  L.n_active = d->n_active;
  L.member_list = d->member_list;
  GemmGroup& g = L.g[0];
  g.M = d->M;
  g.N = d->N;
  g.k_blocks = d->K / 64;
  g.a_mn = d->a_mn;
  g.b_mn = d->b_mn;
  int rc;
  if (!d->a_mn) rc = encode_tmap_bf16(&g.mapA, d->A, d->K, d->M, d->P, d->a_ld, d->a_ms, 128);
  else rc = encode_tmap_bf16(&g.mapA, d->A, d->M, d->K, d->P, d->a_ld, d->a_ms, 64);
  if (rc) return fail("encode mapA failed rc=" + std::to_string(rc));
  if (!d->b_mn) rc = encode_tmap_bf16(&g.mapB, d->B, d->K, d->N, d->P, d->b_ld, d->b_ms, 128);
  else rc = encode_tmap_bf16(&g.mapB, d->B, d->N, d->K, d->P, d->b_ld, d->b_ms, 64);
  if (rc) return fail("encode mapB failed rc=" + std::to_string(rc));
  EpiParams& E = g.epi;
  E.kind = d->epi_kind;
  E.act = d->act;
  E.ln_eps = d->ln_eps;
  const long long MN = (long long)d->M * d->N;
  const int mt = (d->M + 127) / 128;
  E.bias = {d->bias, d->N, d->N, 0};
  E.gamma = {d->gamma, d->N, d->N, 0};
  E.beta = {d->beta, d->N, d->N, 0};
  E.out_h = {(__nv_bfloat16*)d->out_h, MN, d->N, 0};
  E.out_gp = {(__nv_bfloat16*)d->out_gp, MN, d->N, 0};
  E.out_xhat = {(__nv_bfloat16*)d->out_xhat, MN, d->N, 0};
  E.out_rstd = {d->out_rstd, d->M, 1, 0};
  E.gp = {(const __nv_bfloat16*)d->gp, MN, d->N, 0};
  E.xhat = {(const __nv_bfloat16*)d->xhat, MN, d->N, 0};
  E.rstd = {d->rstd, d->M, 1, 0};
  E.out_f32 = {d->out_f32, MN, d->N, 0};
  E.out_dbias_part = {d->dbias_part, (long long)mt * d->N, d->N, 0};
  E.out_dgamma_part = {d->dgamma_part, (long long)mt * d->N, d->N, 0};
  E.out_dbeta_part = {d->dbeta_part, (long long)mt * d->N, d->N, 0};
  const int grid = finalize_launch(&L, d->n_sub);
  cudaError_t e = launch_gemm(L, d->n_sub, grid, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(std::string("launch: ") + cudaGetErrorString(e));
  return 0;
}
