// Population engine: arenas, activation buffers, the dependency-ordered multi-stream schedule of one lockstep
// FQL update (captured once into a CUDA graph and replayed), forward-only loss evaluation, parameter I/O.
//
// Schedule of one step (SURVEY.md Appendix A maps each pass to the upstream loss term):
//   s0: prep -> Euler flow chain (F x [4 GEMM + head]) ------------------------> pi_dout -> PI backward -> Adam(pi)
//   s1: PI forward (3B rows) -> TQ forward -> QC forward -> QC backward -> [QA done] -> Adam(critic)
//   s2: BC forward -> BC backward -> QA forward -> QA input-gradient chain -> [Euler done] -> Adam(bc)
#include "pop.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "chain.h"
#include "philox.cuh"

namespace fqlpop {

#define CK(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) {                                                                      \
      if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
      return 1;                                                                                   \
    }                                                                                             \
  } while (0)

// One member group: a window [off, off + n) of the active list that steps on its own three streams, independently
// of the other groups (members share nothing but read-only data), so one group's long dependent chain -- the F
// Euler steps -- overlaps the other groups' wide passes instead of leaving the SMs idle behind a step barrier.
struct GroupPlan {
  int off = 0, n = 0;
  // s[0..4]: compute branches of the step (high priority); s[5..7]: the three optimizer branches (low priority: their
  // grids flood every free SM and would otherwise delay the small head kernels of the other groups)
  cudaStream_t s[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // low-priority twins of s[0..4]: the layer-chain launches (SM-filling, long) run here, so that whenever an SM frees
  // up the short kernels between them (heads, weight gradients, optimizer) are served first
  cudaStream_t c[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  // g_step: the update whose gradients stay readable (the last step of a call: grad stats, fqlpop_leaf which = grad);
  // g_fast: the same update without the gradient stores (every earlier step of a multi-step call)
  cudaGraph_t g_step = nullptr, g_fast = nullptr, g_loss = nullptr;
  cudaGraphExec_t x_step = nullptr, x_fast = nullptr, x_loss = nullptr;
  cudaEvent_t done = nullptr;
};
struct Plan {
  std::vector<GroupPlan> groups;
  // A second, finer split (groups of 2) for single-step training calls -- the pattern of a loop that feeds every step a
  // fresh host batch: such a call joins all groups at its end, and the smaller the groups, the less of the GPU idles
  // around that join (16 members, bench.py e2e: 17.2k -> 18.0k updates/s; K-step calls stay on `groups`: 18.6k vs 18.4k)
  std::vector<GroupPlan> groups1;
  std::vector<cudaEvent_t> ev;
  int launches_per_step = 0;  // kernels of ours inside one replay of every group's step graph
};

namespace {

int round_up(int x, int m) { return (x + m - 1) / m * m; }

template <class T>
int dev_alloc(Population* P, T** out, long long count, std::string* err) {
  void* p = nullptr;
  const size_t bytes = (size_t)(count > 0 ? count : 1) * sizeof(T);
  CK(cudaMalloc(&p, bytes));
  CK(cudaMemset(p, 0, bytes));
  P->allocs.push_back(p);
  *out = reinterpret_cast<T*>(p);
  return 0;
}

NetLayout make_layout(int nb, int in_dim, int H, int n_hidden, int out_dim, int ln) {
  NetLayout L;
  L.nb = nb; L.in_dim = in_dim; L.in_pad = round_up(in_dim, 64); L.H = H; L.n_hidden = n_hidden;
  L.out_dim = out_dim; L.ln = ln;
  long long o = 0;
  for (int l = 0; l <= n_hidden; ++l) {
    L.offW[l] = o; o += (long long)L.K(l) * L.N(l);
    L.offb[l] = o; o += L.N(l);
    if (ln && l < n_hidden) {
      L.offg[l] = o; o += H;
      L.offbeta[l] = o; o += H;
    }
  }
  L.S = o;  // every leaf size is a multiple of 64 floats, so S and all offsets are too
  return L;
}

int alloc_arena(Population* P, Arena& ar, const NetLayout& L, bool with_target, std::string* err) {
  ar.L = L;
  ar.V = P->c.P * L.nb;
  const long long n = (long long)ar.V * L.S;
  if (dev_alloc(P, &ar.theta, n, err) || dev_alloc(P, &ar.mu, n, err) || dev_alloc(P, &ar.nu, n, err) ||
      dev_alloc(P, &ar.grad, n, err) || dev_alloc(P, &ar.shadow, n, err))
    return 1;
  if (with_target && (dev_alloc(P, &ar.target, n, err) || dev_alloc(P, &ar.tshadow, n, err))) return 1;
  return 0;
}

int alloc_pass(Population* P, PassBuf& pb, const NetLayout& L, int V, int M, bool save, std::string* err) {
  pb.V = V; pb.M = M; pb.in_pad = L.in_pad; pb.H = L.H; pb.n_hidden = L.n_hidden; pb.save = save; pb.ln = L.ln;
  const long long act = (long long)V * M * L.H;
  if (dev_alloc(P, &pb.X0, (long long)V * M * L.in_pad, err)) return 1;
  const int nh = save ? L.n_hidden : 2;
  for (int l = 0; l < nh; ++l)
    if (dev_alloc(P, &pb.h[l], act, err)) return 1;
  if (save)
    for (int l = 0; l < L.n_hidden; ++l) {
      if (dev_alloc(P, &pb.gp[l], act, err)) return 1;
      if (L.ln && (dev_alloc(P, &pb.xhat[l], act, err) || dev_alloc(P, &pb.rstd[l], (long long)V * M, err))) return 1;
    }
  return 0;
}

int alloc_bwd(Population* P, BwdBuf& bw, const NetLayout& L, int V, int M, bool partials, std::string* err) {
  bw.V = V; bw.M = M; bw.H = L.H; bw.m_tiles = (M + kBlockM - 1) / kBlockM;
  if (dev_alloc(P, &bw.dout, (long long)V * M * kHeadPad, err)) return 1;
  for (int l = 0; l < L.n_hidden; ++l) {
    if (dev_alloc(P, &bw.dz[l], (long long)V * M * L.H, err)) return 1;
    if (partials) {
      const long long n = (long long)V * bw.m_tiles * 4 * L.H;  // chain kernels write one row per TMEM lane quarter
      if (dev_alloc(P, &bw.dbias_part[l], n, err)) return 1;
      if (L.ln && (dev_alloc(P, &bw.dgamma_part[l], n, err) || dev_alloc(P, &bw.dbeta_part[l], n, err))) return 1;
    }
  }
  return 0;
}

// Glorot-uniform kernels (true fan-in / fan-out), zero biases, LayerNorm scale 1; padding stays zero.
__global__ void init_arena_kernel(float* theta, NetLayout L, int V, const unsigned long long* seeds, int net_id) {
  const int v = blockIdx.y;
  const int p = v / L.nb, e = v - p * L.nb;
  float* th = theta + (long long)v * L.S;
  for (int l = 0; l <= L.n_hidden; ++l) {
    const int K = L.K(l), N = L.N(l), Kt = L.K_true(l), Nt = L.N_true(l);
    const float lim = sqrtf(6.0f / (float)(Kt + Nt));
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K * N; i += gridDim.x * blockDim.x) {
      const int k = i / N, n = i - k * N;
      float w = 0.f;
      if (k < Kt && n < Nt) {
        const unsigned long long stream = 0xF00D000000000000ull + (unsigned long long)(net_id * 64 + l * 4 + e);
        w = (2.f * philox_uniform(seeds[p], stream, (uint32_t)k, (uint32_t)n) - 1.f) * lim;
      }
      th[L.offW[l] + i] = w;
    }
    if (L.ln && l < L.n_hidden)
      for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < L.H; i += gridDim.x * blockDim.x) th[L.offg[l] + i] = 1.f;
  }
}

// ---------------------------------------------------------------------------------------------------
// GEMM group builders
// ---------------------------------------------------------------------------------------------------
struct Builder {
  Population* P;
  std::string* err;
  bool ok = true;
  GemmLaunch* L = nullptr;  // scratch launch descriptor (heap)
  int count = 0;            // kernels launched through this builder

  void begin(int nb) {
    memset(L, 0, sizeof(GemmLaunch));
    L->P = P->cur_n() * nb;
    L->n_active = nullptr;
    L->member_list = nb == 1 ? P->cur_list() : P->cur_list_e();
  }
  GemmGroup& add() { return L->g[L->num_groups++]; }
  void tmap(CUtensorMap* m, const void* base, long long inner, long long rows, long long batch, long long ld,
            long long ms, int box_rows) {
    const int rc = encode_tmap_bf16(m, base, inner, rows, batch, ld, ms, box_rows);
    if (rc != 0 && ok) {
      ok = false;
      if (err) *err = "cuTensorMapEncodeTiled failed rc=" + std::to_string(rc);
    }
  }
  void launch(int n_sub, cudaStream_t s) {
    const int grid = finalize_launch(L, n_sub);
    cudaError_t e = launch_gemm(*L, n_sub, grid, s);
    if (e != cudaSuccess && ok) {
      ok = false;
      if (err) *err = std::string("launch_gemm: ") + cudaGetErrorString(e);
    }
    ++count;
  }

  // forward layer l of `pass` on rows [row0, row0 + M): h_l = act(x W_l + b_l) (+ LayerNorm)
  void fwd_group(PassBuf& pass, const Arena& ar, const float* params, const __nv_bfloat16* wsh, int l, int row0, int M,
                 bool save) {
    const NetLayout& N = ar.L;
    GemmGroup& g = add();
    const int K = N.K(l);
    const __nv_bfloat16* X = l == 0 ? pass.X0 : pass.hl(l - 1);
    g.M = M; g.N = N.H; g.k_blocks = K / 64; g.a_mn = 0; g.b_mn = 1;
    g.a_row0 = row0;
    tmap(&g.mapA, X, K, pass.M, pass.V, K, (long long)pass.M * K, 128);
    tmap(&g.mapB, wsh + N.offW[l], N.H, K, ar.V, N.H, N.S, 64);
    EpiParams& E = g.epi;
    E.kind = N.ln ? EPI_FWD_LN : EPI_FWD;
    E.act = ACT_GELU;
    E.ln_eps = P->c.ln_eps;
    const long long ams = (long long)pass.M * N.H, roff = (long long)row0 * N.H;
    E.bias = {params + N.offb[l], N.S, N.H, 0};
    if (N.ln) {
      E.gamma = {params + N.offg[l], N.S, N.H, 0};
      E.beta = {params + N.offbeta[l], N.S, N.H, 0};
    }
    E.out_h = {pass.hl(l) + roff, ams, N.H, 0};
    if (save) {
      E.out_gp = {pass.gp[l] + roff, ams, N.H, 0};
      if (N.ln) {
        E.out_xhat = {pass.xhat[l] + roff, ams, N.H, 0};
        E.out_rstd = {pass.rstd[l] + row0, (long long)pass.M, 1, 0};
      }
    }
  }

  // backward through layer l (l >= 1): dz_{l-1} = act'(z_{l-1}) * LNbwd(dz_l W_l^T); dz_{n_hidden} = dout
  void dgrad_group(PassBuf& pass, BwdBuf& bw, const Arena& ar, int l, int row0, bool partials) {
    const NetLayout& N = ar.L;
    GemmGroup& g = add();
    const int Kd = N.N(l);  // reduction = out features of layer l (64 for the head)
    const __nv_bfloat16* dZ = l == N.n_hidden ? bw.dout : bw.dz[l];
    g.M = bw.M; g.N = N.H; g.k_blocks = Kd / 64; g.a_mn = 0; g.b_mn = 0;
    tmap(&g.mapA, dZ, Kd, bw.M, bw.V, Kd, (long long)bw.M * Kd, 128);
    tmap(&g.mapB, ar.shadow + N.offW[l], Kd, N.K(l), ar.V, Kd, N.S, 128);
    EpiParams& E = g.epi;
    E.kind = N.ln ? EPI_DGRAD_LN : EPI_DGRAD;
    const long long ams = (long long)pass.M * N.H, roff = (long long)row0 * N.H;
    E.gp = {pass.gp[l - 1] + roff, ams, N.H, 0};
    if (N.ln) {
      E.gamma = {ar.theta + N.offg[l - 1], N.S, N.H, 0};
      E.xhat = {pass.xhat[l - 1] + roff, ams, N.H, 0};
      E.rstd = {pass.rstd[l - 1] + row0, (long long)pass.M, 1, 0};
    }
    E.out_h = {bw.dz[l - 1], (long long)bw.M * N.H, N.H, 0};
    if (partials) {
      const long long pms = (long long)bw.m_tiles * N.H;
      E.out_dbias_part = {bw.dbias_part[l - 1], pms, N.H, 0};
      if (N.ln) {
        E.out_dgamma_part = {bw.dgamma_part[l - 1], pms, N.H, 0};
        E.out_dbeta_part = {bw.dbeta_part[l - 1], pms, N.H, 0};
      }
    }
  }

  // gradient with respect to the (padded) network input: dx0 = dz_0 W_0^T, fp32
  void dinput_group(BwdBuf& bw, const Arena& ar, float* out) {
    const NetLayout& N = ar.L;
    GemmGroup& g = add();
    g.M = bw.M; g.N = N.in_pad; g.k_blocks = N.H / 64; g.a_mn = 0; g.b_mn = 0;
    tmap(&g.mapA, bw.dz[0], N.H, bw.M, bw.V, N.H, (long long)bw.M * N.H, 128);
    tmap(&g.mapB, ar.shadow + N.offW[0], N.H, N.in_pad, ar.V, N.H, N.S, 128);
    g.epi.kind = EPI_STORE_F32;
    g.epi.out_f32 = {out, (long long)bw.M * N.in_pad, N.in_pad, 0};
  }

  // weight gradient of layer l: dW_l = X_l^T dz_l  (fp32, written straight into the grad arena)
  // keep_grad < 0: plain gradient store (the optimizer runs as its own kernel); 0 / 1: Adam + Polyak + shadow refresh
  // fused into the GEMM's epilogue, the gradient itself stored only when 1
  int fuse_adam = -1;
  void wgrad_group(PassBuf& pass, BwdBuf& bw, const Arena& ar, int l, int row0) {
    const NetLayout& N = ar.L;
    GemmGroup& g = add();
    const int Kin = N.K(l), Nout = N.N(l);
    const __nv_bfloat16* X = l == 0 ? pass.X0 : pass.hl(l - 1);
    const __nv_bfloat16* dZ = l == N.n_hidden ? bw.dout : bw.dz[l];
    g.M = Kin; g.N = Nout; g.k_blocks = bw.M / 64; g.a_mn = 1; g.b_mn = 1;
    g.a_row0 = row0;
    tmap(&g.mapA, X, Kin, pass.M, pass.V, Kin, (long long)pass.M * Kin, 64);
    tmap(&g.mapB, dZ, Nout, bw.M, bw.V, Nout, (long long)bw.M * Nout, 64);
    g.epi.kind = EPI_STORE_F32;
    g.epi.out_f32 = {ar.grad + N.offW[l], N.S, Nout, 0};
    if (fuse_adam >= 0) {
      g.epi.kind = EPI_ADAM;
      AdamTile& T = g.epi.adam;
      T.theta = ar.theta + N.offW[l]; T.mu = ar.mu + N.offW[l]; T.nu = ar.nu + N.offW[l];
      T.shadow = ar.shadow + N.offW[l];
      T.target = ar.target ? ar.target + N.offW[l] : nullptr;
      T.tshadow = ar.tshadow ? ar.tshadow + N.offW[l] : nullptr;
      T.adam_bc = P->adam_bc; T.nb = N.nb; T.keep_grad = fuse_adam;
      T.lr = P->c.lr; T.b1 = P->c.adam_b1; T.b2 = P->c.adam_b2; T.eps = P->c.adam_eps; T.tau = P->c.tau;
    }
  }
};

// ---------------------------------------------------------------------------------------------------
// layer-chain job builders (chain.cu): one launch = a whole forward pass / Euler flow / input-gradient chain
// ---------------------------------------------------------------------------------------------------
void record(Plan* pl, cudaEvent_t* out, cudaStream_t s);

struct ChainBuilder {
  Population* P;
  std::string* err;
  GroupPlan* gp = nullptr;  // chain launches hop to the low-priority twin of their stream (FQLPOP_SCHED bit 2)
  Plan* pl = nullptr;
  bool ok = true;
  ChainLaunch* L = nullptr;
  int count = 0;

  void begin(int nb) {
    memset(L, 0, sizeof(ChainLaunch));
    L->n_virtual = P->cur_n() * nb;
    L->member_list = nb == 1 ? P->cur_list() : P->cur_list_e();
  }
  ChainJob& add() { return L->job[L->num_jobs++]; }
  void tmap(CUtensorMap* m, const void* base, long long inner, long long rows, long long batch, long long ld,
            long long ms, int box_rows) {
    const int rc = encode_tmap_bf16(m, base, inner, rows, batch, ld, ms, box_rows);
    if (rc != 0 && ok) {
      ok = false;
      if (err) *err = "cuTensorMapEncodeTiled failed rc=" + std::to_string(rc);
    }
  }
  void tmap64(CUtensorMap* m, const void* base, long long inner, long long rows, long long batch, long long ld,
              long long ms, int box_rows) {
    const int rc = encode_tmap_bf16_sw64(m, base, inner, rows, batch, ld, ms, box_rows);
    if (rc != 0 && ok) {
      ok = false;
      if (err) *err = "cuTensorMapEncodeTiled (64B swizzle) failed rc=" + std::to_string(rc);
    }
  }
  // backward chains on chain16.cu: dz leaves through TMA stores too (box 64 columns x 32 rows of the [V][M][H] buffer)
  void encode_backward_maps() {
    const ChainJob& j0 = L->job[0];
    L->population_active = (int)P->h_members.size();
    const bool c16 = chain16_eligible(*L) &&
                     chain16_selected_shape(j0.H, 64, j0.ln != 0, j0.iters, (int)P->h_members.size());
    for (int i = 0; i < L->num_jobs; ++i) {
      ChainJob& j = L->job[i];
      for (int l = 0; l < j.n_layers; ++l) {
        j.layer[l].h_tma = 0;
        if (!c16 || j.layer[l].kind != CL_HIDDEN || !j.layer[l].out_dz) continue;
        const HArgs& h = hargs[i][l];
        if (encode_tmap_bf16(&j.mapH[l], h.base, j.H, j.row0 + j.M, h.V, j.H, h.pass_rows * j.H, 32) == 0) j.layer[l].h_tma = 1;
      }
    }
  }
  void launch(int mode, cudaStream_t s) {
    if (mode == 0) encode_forward_maps();
    else encode_backward_maps();
    finalize_chain(L);
    L->timeline = P->timeline;
    L->tag = (long long)count | ((long long)P->g_off << 8);
#ifdef FQLPOP_DEBUG_SWITCHES
    if (const char* e = getenv("FQLPOP_CHAIN_DEBUG")) L->debug_flags = atoi(e);  // (make TRACE=1 only) bit 0: no weight stream
#endif
    L->population_active = (int)P->h_members.size();
    {  // (debug) FQLPOP_CHAIN_TRACE=1: clock stamps of CTA 0 of chain launch FQLPOP_TRACE_LAUNCH (default 0 = Euler)
      const char* tl = getenv("FQLPOP_TRACE_LAUNCH");
      L->trace = (P->chain_trace && count == (tl ? atoi(tl) : 0) && P->g_off == 0) ? P->chain_trace : nullptr;
    }
    cudaStream_t run = s;
    if (gp)
      for (int i = 0; i < 5; ++i)
        if (gp->s[i] == s) run = gp->c[i];
    cudaEvent_t e_in = nullptr, e_out = nullptr;
    if (run != s) {
      record(pl, &e_in, s);
      cudaStreamWaitEvent(run, e_in, 0);
    }
    cudaError_t e = launch_chain(*L, mode, run);
    if (e != cudaSuccess && ok) {
      ok = false;
      if (err) *err = std::string("launch_chain: ") + cudaGetErrorString(e);
    }
    if (run != s) {
      record(pl, &e_out, run);
      cudaStreamWaitEvent(s, e_out, 0);
    }
    ++count;
  }
  ChainJob& common(PassBuf& pass, const Arena& ar, const float* params, int row0, int M) {
    const NetLayout& N = ar.L;
    ChainJob& j = add();
    j.M = M; j.row0 = row0; j.iters = 1; j.ln = N.ln; j.H = N.H; j.ln_eps = P->c.ln_eps;
    j.params = params; j.S = N.S;
    j.pass_ms = (long long)pass.M * N.H; j.pass_rs = pass.M;
    return j;
  }

  // forward pass over rows [row0, row0 + M): every hidden layer; the last hidden activations always go to HBM
  // (the head kernels read them), the others only when the backward pass needs them
  // Weight maps of a forward chain are encoded when the launch is complete (launch()): the second-generation kernel
  // (chain16.cu) fetches a weight stage with one TMA instruction whose box spans the CTA's 64-column blocks -- 4 blocks x
  // 32 k for a whole 256-column accumulator tile, 2 blocks x 64 k for one CTA's half in pair mode, 1 block for the narrow
  // layers (encode_tmap_bf16_wide) -- and whether it runs, paired or not, depends on every job of the launch.
  struct WArgs { const void* base; long long inner, rows, batch, ms; };
  WArgs wargs[kChainMaxJobs][kChainMaxLayers];
  struct HArgs { const void* base; long long pass_rows, V; };   // the layer's h buffer [V][pass rows][H] (out_h set)
  HArgs hargs[kChainMaxJobs][kChainMaxLayers];
  void wmap_later(int layer, const void* base, long long inner, long long rows, long long batch, long long ms) {
    wargs[L->num_jobs - 1][layer] = WArgs{base, inner, rows, batch, ms};
  }
  void encode_forward_maps() {
    const ChainJob& j0 = L->job[0];
    const bool c16 = chain16_eligible(*L) &&
                     chain16_selected_shape(j0.H, 64, j0.ln != 0, j0.iters, (int)P->h_members.size());
    bool pair = c16 && chain16_pair_mode();
    for (int i = 0; i < L->num_jobs; ++i) pair = pair && chain16_pairs_rows(L->job[i].M);
    L->population_active = (int)P->h_members.size();
    for (int i = 0; i < L->num_jobs; ++i)      // (chain2_will_run looks at the layer shapes only)
      for (int l = 0; l < L->job[i].n_layers; ++l) L->job[i].layer[l].b_box_rows = 0;
    const bool c2 = !c16 && chain2_will_run(*L);
    for (int i = 0; i < L->num_jobs; ++i) {
      ChainJob& j = L->job[i];
      for (int l = 0; l < j.n_layers; ++l) {
        const WArgs& a = wargs[i][l];
        int rc;
        j.layer[l].h_tma = 0;
        if (c16 && j.layer[l].out_h && l < kChainMaxLayers && j.layer[l].kind == CL_HIDDEN) {
          // chain16.cu writes h with TMA stores: box 64 columns x 32 rows (one epilogue warp's rows of one ACT block),
          // rows clipped at the job's last row
          const HArgs& h = hargs[i][l];
          const int hrc = encode_tmap_bf16(&j.mapH[l], h.base, j.H, j.row0 + j.M, h.V, j.H, h.pass_rows * j.H, 32);
          if (hrc == 0) j.layer[l].h_tma = 1;
        }
        if (c16) {
          const int boxes = j.layer[l].n_pad >= 256 ? (pair ? 2 : 4) : 1;
          rc = encode_tmap_bf16_wide(&j.mapW[l], a.base, a.inner, a.rows, a.batch, a.inner, a.ms, pair ? 64 : 32, boxes);
          j.layer[l].b_box_rows = boxes;
        } else if (c2) {   // N-split cluster kernel: one CTA stages 256 columns of a hidden layer, all of a narrow one
          const int boxes = j.layer[l].kind == CL_HIDDEN ? 4 : j.layer[l].n_pad / 64;
          rc = encode_tmap_bf16_wide(&j.mapW[l], a.base, a.inner, a.rows, a.batch, a.inner, a.ms, 32, boxes);
          j.layer[l].b_box_rows = boxes;
        } else {
          rc = encode_tmap_bf16(&j.mapW[l], a.base, a.inner, a.rows, a.batch, a.inner, a.ms, 32);
          j.layer[l].b_box_rows = 0;
        }
        if (rc != 0 && ok) {
          ok = false;
          if (err) *err = "cuTensorMapEncodeTiled (forward weights) failed rc=" + std::to_string(rc);
        }
      }
    }
  }
  ChainJob& fwd_job(PassBuf& pass, const Arena& ar, const float* params, const __nv_bfloat16* wsh, int row0, int M,
                    bool save) {
    const NetLayout& N = ar.L;
    ChainJob& j = common(pass, ar, params, row0, M);
    j.n_layers = N.n_hidden;
    tmap(&j.mapA0, pass.X0, N.in_pad, pass.M, pass.V, N.in_pad, (long long)pass.M * N.in_pad, 128);
    const long long roff = (long long)row0 * N.H;
    for (int l = 0; l < N.n_hidden; ++l) {
      wmap_later(l, wsh + N.offW[l], N.H, N.K(l), ar.V, N.S);
      ChainLayer& y = j.layer[l];
      y.kind = CL_HIDDEN; y.k_blocks = N.K(l) / 64; y.n_pad = N.H; y.a_from_x0 = l == 0;
      y.off_b = N.offb[l]; y.off_g = N.offg[l]; y.off_beta = N.offbeta[l];
      if (save || l == N.n_hidden - 1) y.out_h = pass.hl(l) + roff;
      hargs[L->num_jobs - 1][l] = HArgs{pass.hl(l), pass.M, pass.V};
      if (save) {
        y.out_gp = pass.gp[l] + roff;
        if (N.ln) {
          y.out_xhat = pass.xhat[l] + roff;
          y.out_rstd = pass.rstd[l] + row0;
        }
      }
    }
    return j;
  }

  // the F Euler steps of compute_flow_actions: hidden layers + the head as a 64-column GEMM, iterated on chip
  void euler_job() {
    const Arena& ar = P->bc;
    const NetLayout& N = ar.L;
    ChainJob& j = fwd_job(P->EU, ar, ar.theta, ar.shadow, 0, P->c.B, false);
    j.layer[N.n_hidden - 1].out_h = nullptr;
    const int l = N.n_hidden;
    wmap_later(l, ar.shadow + N.offW[l], kHeadPad, N.H, ar.V, N.S);
    ChainLayer& y = j.layer[l];
    y.kind = CL_EULER_HEAD; y.k_blocks = N.H / 64; y.n_pad = kHeadPad; y.a_from_x0 = 0;
    y.off_b = N.offb[l];
    j.n_layers = N.n_hidden + 1;
    j.iters = P->c.F;
    j.O = P->c.O; j.A = P->c.A; j.F = P->c.F; j.B = P->c.B;
    j.xe = P->xe; j.target_a = P->target_a;
  }

  // input-gradient chain: dz_{l-1} = act'(z_{l-1}) * LNbwd(dz_l W_l^T) for l = n_hidden .. 1 (dz_{n_hidden} = dout),
  // optionally followed by the gradient with respect to the network input (fp32)
  void bwd_job(PassBuf& pass, BwdBuf& bw, const Arena& ar, int row0, bool partials, float* dinput_out) {
    const NetLayout& N = ar.L;
    ChainJob& j = common(pass, ar, ar.theta, 0, bw.M);
    j.dz_ms = (long long)bw.M * N.H;
    j.part_ms = (long long)bw.part_rows * N.H;
    tmap(&j.mapA0, bw.dout, kHeadPad, bw.M, bw.V, kHeadPad, (long long)bw.M * kHeadPad, 128);
    const long long roff = (long long)row0 * N.H;
    // rows of a weight box = output columns one CTA stages per k step: 256, or 128 when CTAs run as cta_group::2 pairs
    const int cap = (chain16_selected_shape(N.H, N.in_pad, N.ln != 0, 1, (int)P->h_members.size()) && chain16_pair_mode() && chain16_pairs_rows(bw.M)) ? 128 : 256;
    int s = 0;
    for (int l = N.n_hidden; l >= 1; --l, ++s) {
      const int Kd = N.N(l);
      ChainLayer& y = j.layer[s];
      y.b_box_rows = N.K(l) < cap ? N.K(l) : cap;
      tmap64(&j.mapW[s], ar.shadow + N.offW[l], Kd, N.K(l), ar.V, Kd, N.S, y.b_box_rows);
      y.kind = CL_HIDDEN; y.k_blocks = Kd / 64; y.n_pad = N.H; y.a_from_x0 = s == 0;
      y.off_g = N.offg[l - 1];
      y.gp = pass.gp[l - 1] + roff;
      if (N.ln) {
        y.xhat = pass.xhat[l - 1] + roff;
        y.rstd = pass.rstd[l - 1] + row0;
      }
      y.out_dz = bw.dz[l - 1];
      hargs[L->num_jobs - 1][s] = HArgs{bw.dz[l - 1], bw.M, bw.V};
      if (partials) {
        y.dbias_part = bw.dbias_part[l - 1];
        if (N.ln) {
          y.dgamma_part = bw.dgamma_part[l - 1];
          y.dbeta_part = bw.dbeta_part[l - 1];
        }
      }
    }
    if (dinput_out) {
      ChainLayer& y = j.layer[s];
      y.b_box_rows = N.in_pad < cap ? N.in_pad : cap;
      tmap64(&j.mapW[s], ar.shadow + N.offW[0], N.H, N.in_pad, ar.V, N.H, N.S, y.b_box_rows);
      y.kind = CL_DINPUT; y.k_blocks = N.H / 64; y.n_pad = N.in_pad; y.a_from_x0 = 0;
      y.out_f32 = dinput_out; y.out_ld = N.in_pad;
      j.f32_ms = (long long)bw.M * N.in_pad;
      ++s;
    }
    j.n_layers = s;
  }
};

void record(Plan* pl, cudaEvent_t* out, cudaStream_t s) {
  cudaEvent_t e;
  cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
  pl->ev.push_back(e);
  cudaEventRecord(e, s);
  *out = e;
}

// Issues one step (train = update; !train = losses only) of the members in P's current window on the group's
// streams. Called under stream capture; s[0] is the origin stream.
int issue_step(Population* P, bool train, GroupPlan& gp, std::string* err, bool keep_grads = true) {
  Plan* pl = P->plan;
  // schedule knobs (measured on B200): FQLPOP_SCHED bit 0 = critic forward / actor-critic pass
  // on their own streams (s3, s4) instead of queued behind s1 / s2; bit 1 = optimizer kernels on separate streams;
  // FQLPOP_ADAM_PRIO = priority of those streams (0 low, 1 same as compute)
  // Switches that change RESULTS (skip Adam, run without the weight stream, ...) exist only in the trace build
  // (make TRACE=1 -> lib/libfqlpop_trace.so); the product library cannot be talked into skipping work by an env var.
#ifdef FQLPOP_DEBUG_SWITCHES
  static const int dbg_skip = [] { const char* e = getenv("FQLPOP_DEBUG_SKIP"); return e ? atoi(e) : 0; }();
#else
  constexpr int dbg_skip = 0;
#endif
  static const int sched = [] { const char* e = getenv("FQLPOP_SCHED"); return e ? atoi(e) : 7; }();
  cudaStream_t s0 = gp.s[0], s1 = gp.s[1], s2 = gp.s[2];
  cudaStream_t s3 = (sched & 1) ? gp.s[3] : s1, s4 = (sched & 1) ? gp.s[4] : s2;
  const PopParams& c = P->c;
  Builder b{P, err};
  b.L = new GemmLaunch;
  // FQLPOP_FUSE_ADAM=1: the optimizer inside the weight-gradient GEMM's epilogue (gemm.cu EPI_ADAM) instead of separate
  // kernels after plain weight-gradient GEMMs. Parity-tested, and it removes the fp32 gradient round trip (34 -> 26 bytes
  // per parameter), but MEASURED SLOWER: 17.1k vs 18.4k updates/s at 16 members (16.3k before the epilogue walked its tile
  // row by row in 512-byte segments). 2 CTAs x 128 epilogue threads per SM keep at most 64 KB of loads in flight per SM, and
  // only half of the time (a batch of rows is loaded, waited for, updated, stored), against the 160 KB of the stand-alone
  // adam_kernel; double-buffering the batches needs ~64 more registers than the 168 two resident CTAs leave.
  // Off by default; read at graph-capture time.
  const int fuse = [] { const char* e = getenv("FQLPOP_FUSE_ADAM"); return e ? atoi(e) : 0; }();
  b.fuse_adam = fuse ? (keep_grads ? 1 : 0) : -1;
  int pw = 0;  // pointwise launches
  const int B = c.B;
  const int nsub_q = P->critic.L.ln ? 4 : 1, nsub_a = P->pi.L.ln ? 4 : 1;
  const int nh_q = P->critic.L.n_hidden, nh_a = P->pi.L.n_hidden;

  cudaEvent_t e_prep, e_pi, e_qa, e_euler, e_s1, e_s2, e_qcf;
  launch_prep(P, s0, train ? 1 : 0); ++pw;
  record(pl, &e_prep, s0);
  cudaStreamWaitEvent(s1, e_prep, 0);
  cudaStreamWaitEvent(s2, e_prep, 0);
  if (s3 != s1) cudaStreamWaitEvent(s3, e_prep, 0);

  ChainBuilder cb{P, err};
  if (sched & 4) { cb.gp = &gp; cb.pl = pl; }
  cb.L = new ChainLaunch;
  const bool chain = P->use_chain;
  // forward pass of one network over [row0, row0 + M) (+ a second, unsaved row range)
  auto fwd_pass = [&](PassBuf& pass, const Arena& ar, const float* params, const __nv_bfloat16* wsh, int nsub, int M0,
                      bool save0, int M1, cudaStream_t s) {
    const int nh = ar.L.n_hidden;
    if (chain) {
      cb.begin(ar.L.nb);
      cb.fwd_job(pass, ar, params, wsh, 0, M0, save0);
      if (M1 > 0) cb.fwd_job(pass, ar, params, wsh, M0, M1, false);
      cb.launch(0, s);
    } else {
      for (int l = 0; l < nh; ++l) {
        b.begin(ar.L.nb);
        b.fwd_group(pass, ar, params, wsh, l, 0, M0, save0);
        if (M1 > 0) b.fwd_group(pass, ar, params, wsh, l, M0, M1, false);
        b.launch(nsub, s);
      }
    }
  };
  auto bwd_pass = [&](PassBuf& pass, BwdBuf& bw, const Arena& ar, int nsub, bool partials, float* dinput,
                      cudaStream_t s) {
    const int nh = ar.L.n_hidden;
    if (chain) {
      cb.begin(ar.L.nb);
      cb.bwd_job(pass, bw, ar, 0, partials, dinput);
      cb.launch(1, s);
    } else {
      for (int l = nh; l >= 1; --l) {
        b.begin(ar.L.nb);
        b.dgrad_group(pass, bw, ar, l, 0, partials);
        b.launch(nsub, s);
      }
      if (dinput) {
        b.begin(ar.L.nb);
        b.dinput_group(bw, ar, dinput);
        b.launch(1, s);
      }
    }
  };

  // ---- s0: Euler flow chain --------------------------------------------------------------------
  if (chain) {
    cb.begin(1);
    cb.euler_job();
    cb.launch(0, s0);
  } else {
    for (int i = 0; i < c.F; ++i) {
      for (int l = 0; l < nh_a; ++l) {
        b.begin(1);
        b.fwd_group(P->EU, P->bc, P->bc.theta, P->bc.shadow, l, 0, B, false);
        b.launch(nsub_a, s0);
      }
      launch_euler_head(P, i, s0); ++pw;
    }
  }
  record(pl, &e_euler, s0);

  // ---- s1: one-step policy forward (rows [0,B) kept), TD target, critic forward/backward --------
  fwd_pass(P->PI, P->pi, P->pi.theta, P->pi.shadow, nsub_a, B, train, 2 * B, s1);
  launch_pi_head(P, s1); ++pw;
  record(pl, &e_pi, s1);
  fwd_pass(P->TQ, P->critic, P->critic.target, P->critic.tshadow, nsub_q, B, false, 0, s1);
  launch_tq_head(P, s1); ++pw;
  // (the critic's own forward pass needs nothing from the policy / target passes: it runs beside them on s3 and
  // joins s1 at the loss head)
  fwd_pass(P->QC, P->critic, P->critic.theta, P->critic.shadow, nsub_q, B, train, 0, s3);
  record(pl, &e_qcf, s3);
  cudaStreamWaitEvent(s1, e_qcf, 0);
  launch_qc_head(P, s1); ++pw;
  if (train) {
    bwd_pass(P->QC, P->bQC, P->critic, nsub_q, true, nullptr, s1);
    if (!fuse) {
      b.begin(c.E);
      for (int l = 0; l <= nh_q; ++l) b.wgrad_group(P->QC, P->bQC, P->critic, l, 0);
      if (!(dbg_skip & 2)) b.launch(1, s1);
    }
  }

  // ---- s2: BC flow forward/backward, then the critic pass on (s, clip(a_pi)) -----------------------
  fwd_pass(P->BC, P->bc, P->bc.theta, P->bc.shadow, nsub_a, B, train, 0, s2);
  launch_bc_head(P, s2); ++pw;
  if (train) {
    bwd_pass(P->BC, P->bBC, P->bc, nsub_a, true, nullptr, s2);
    if (!fuse) {
      b.begin(1);
      for (int l = 0; l <= nh_a; ++l) b.wgrad_group(P->BC, P->bBC, P->bc, l, 0);
      if (!(dbg_skip & 2)) b.launch(1, s2);
    }
  }
  // critic pass on (s, clip(a_pi)): its own stream, as soon as the policy head is out
  cudaStreamWaitEvent(s4, e_pi, 0);
  fwd_pass(P->QA, P->critic, P->critic.theta, P->critic.shadow, nsub_q, B, train, 0, s4);
  launch_qa_head(P, s4); ++pw;
  if (train) bwd_pass(P->QA, P->bQA, P->critic, nsub_q, false, P->dx0_qa, s4);
  record(pl, &e_qa, s4);

  // ---- s0: distillation + policy backward -------------------------------------------------------
  cudaStreamWaitEvent(s0, e_qa, 0);
  launch_pi_dout(P, s0); ++pw;
  if (train) {
    bwd_pass(P->PI, P->bPI, P->pi, nsub_a, true, nullptr, s0);
    // the optimizer branches run on the low-priority streams
    cudaStream_t a0 = (sched & 2) ? gp.s[5] : s1, a1 = (sched & 2) ? gp.s[6] : s2, a2 = (sched & 2) ? gp.s[7] : s0;
    cudaEvent_t e_w0, e_w1, e_w2, e_a0, e_a1, e_a2;
    if (!fuse) {
      b.begin(1);
      for (int l = 0; l <= nh_a; ++l) b.wgrad_group(P->PI, P->bPI, P->pi, l, 0);
      if (!(dbg_skip & 2)) b.launch(1, s0);
    }
    // Fused mode: the weight-gradient GEMM of a network IS its optimizer step (gemm.cu EPI_ADAM), so it may only start once
    // every reader of that network's weights is done; the biases / LayerNorm leaves follow in a small kernel.
    auto optimizer = [&](int which, PassBuf& pass, BwdBuf& bw, Arena& ar, int nb, int nh, cudaStream_t st) {
      if (fuse) {
        b.begin(nb);
        for (int l = 0; l <= nh; ++l) b.wgrad_group(pass, bw, ar, l, 0);
        if (!(dbg_skip & 2)) b.launch(1, st);
      }
      if (!(dbg_skip & 4)) { launch_small_grads(P, which, st); ++pw; }
      if (!(dbg_skip & 1)) {
        if (fuse) launch_adam_leaves(P, which, st);
        else launch_adam(P, which, st);
        ++pw;
      }
    };
    record(pl, &e_w2, s0);
    cudaStreamWaitEvent(a2, e_w2, 0);
    optimizer(2, P->PI, P->bPI, P->pi, 1, nh_a, a2);
    record(pl, &e_a2, a2);
    // critic optimizer: after every reader of the critic weights (QA chain on s4) is done
    record(pl, &e_w0, s1);
    cudaStreamWaitEvent(a0, e_w0, 0);
    cudaStreamWaitEvent(a0, e_qa, 0);
    optimizer(0, P->QC, P->bQC, P->critic, c.E, nh_q, a0);
    record(pl, &e_a0, a0);
    // BC flow optimizer: after the Euler chain (s0) has stopped reading the BC weights
    record(pl, &e_w1, s2);
    cudaStreamWaitEvent(a1, e_w1, 0);
    cudaStreamWaitEvent(a1, e_euler, 0);
    optimizer(1, P->BC, P->bBC, P->bc, 1, nh_a, a1);
    record(pl, &e_a1, a1);
    cudaStreamWaitEvent(s0, e_a0, 0);
    cudaStreamWaitEvent(s0, e_a1, 0);
    cudaStreamWaitEvent(s0, e_a2, 0);
  }
  record(pl, &e_s1, s1);
  record(pl, &e_s2, s2);
  cudaStreamWaitEvent(s0, e_s1, 0);
  cudaStreamWaitEvent(s0, e_s2, 0);
  launch_finish_info(P, train ? 1 : 0, s0); ++pw;
  delete b.L;
  delete cb.L;
  if (!b.ok || !cb.ok) return 1;
  CK(cudaGetLastError());
  if (train) pl->launches_per_step += b.count + cb.count + pw;
  return 0;
}

void destroy_graphs(Plan* pl) {
  for (std::vector<GroupPlan>* set : {&pl->groups, &pl->groups1})
  for (GroupPlan& g : *set) {
    if (g.x_step) cudaGraphExecDestroy(g.x_step);
    if (g.x_fast) cudaGraphExecDestroy(g.x_fast);
    if (g.x_loss) cudaGraphExecDestroy(g.x_loss);
    if (g.g_step) cudaGraphDestroy(g.g_step);
    if (g.g_fast) cudaGraphDestroy(g.g_fast);
    if (g.g_loss) cudaGraphDestroy(g.g_loss);
    g.x_step = g.x_fast = g.x_loss = nullptr;
    g.g_step = g.g_fast = g.g_loss = nullptr;
  }
  for (cudaEvent_t e : pl->ev) cudaEventDestroy(e);
  pl->ev.clear();
  pl->launches_per_step = 0;
}

void destroy_groups(Plan* pl) {
  destroy_graphs(pl);
  for (std::vector<GroupPlan>* set : {&pl->groups, &pl->groups1}) {
    for (GroupPlan& g : *set) {
      for (int i = 0; i < 8; ++i)
        if (g.s[i]) cudaStreamDestroy(g.s[i]);
      for (int i = 0; i < 5; ++i)
        if (g.c[i]) cudaStreamDestroy(g.c[i]);
      if (g.done) cudaEventDestroy(g.done);
    }
    set->clear();
  }
}

// Splits the active list into member groups of FQLPOP_GROUP_SIZE members (0 = one lockstep group). Default: 4, or 2 while
// at most 10 members are active (small populations are latency-bound: more, smaller graphs in flight; measured 4 members
// 9.8k -> 10.5k updates/s, 8 members 15.3k -> 15.6k; from 12 members on the group size is within +-3 %).
static int make_group_set(std::vector<GroupPlan>& set, int n, int gs, std::string* err) {
  if (gs <= 0 || gs > n) gs = n;
  const int G = (n + gs - 1) / gs;
  set.resize(G);
  for (int g = 0; g < G; ++g) {
    GroupPlan& gp = set[g];
    // spread the remainder so group sizes differ by at most one
    gp.off = (int)((long long)n * g / G);
    gp.n = (int)((long long)n * (g + 1) / G) - gp.off;
    int prio_lo = 0, prio_hi = 0;  // numerically: lowest priority, highest priority
    const char* ape = getenv("FQLPOP_ADAM_PRIO");
    const int adam_prio = ape ? atoi(ape) : 1;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int i = 0; i < 8; ++i)
      CK(cudaStreamCreateWithPriority(&gp.s[i], cudaStreamNonBlocking, (i < 5 || adam_prio) ? prio_hi : prio_lo));
    for (int i = 0; i < 5; ++i) CK(cudaStreamCreateWithPriority(&gp.c[i], cudaStreamNonBlocking, prio_lo));
    CK(cudaEventCreateWithFlags(&gp.done, cudaEventDisableTiming));
  }
  return 0;
}
int make_groups(Population* P, std::string* err) {
  Plan* pl = P->plan;
  destroy_groups(pl);
  const int n = (int)P->h_members.size();
  int gs = n <= 10 ? 2 : 4;
  const char* e = getenv("FQLPOP_GROUP_SIZE");
  if (e) gs = atoi(e);
  if (make_group_set(pl->groups, n, gs, err)) return 1;
  // single-step calls: groups of 2 (FQLPOP_GROUP_SIZE fixes one split for every call)
  if (!e && gs > 2 && n > 2 && make_group_set(pl->groups1, n, 2, err)) return 1;
  return 0;
}

struct Window {  // RAII: narrows the population's launch window to one group
  Population* P;
  int off, n;
  Window(Population* p, const GroupPlan& g) : P(p), off(p->g_off), n(p->g_n) { P->g_off = g.off; P->g_n = g.n; }
  ~Window() { P->g_off = off; P->g_n = n; }
};

int build_graph(Population* P, bool train, GroupPlan& gp, std::string* err, bool keep_grads = true) {
  Window w(P, gp);
  CK(cudaStreamBeginCapture(gp.s[0], cudaStreamCaptureModeThreadLocal));
  const int rc = issue_step(P, train, gp, err, keep_grads);
  cudaGraph_t g = nullptr;
  cudaError_t e = cudaStreamEndCapture(gp.s[0], &g);
  if (rc) {
    if (g) cudaGraphDestroy(g);
    return 1;
  }
  CK(e);
  cudaGraphExec_t x = nullptr;
  CK(cudaGraphInstantiate(&x, g, 0));
  if (train && keep_grads) { gp.g_step = g; gp.x_step = x; }
  else if (train) { gp.g_fast = g; gp.x_fast = x; }
  else { gp.g_loss = g; gp.x_loss = x; }
  return 0;
}

}  // namespace

// =======================================================================================================
// public (pop.h) entry points
// =======================================================================================================
static int upload_members(Population* P, std::string* err) {
  const int n = (int)P->h_members.size();
  std::vector<int> ve((size_t)n * P->c.E);
  for (int i = 0; i < n; ++i)
    for (int e = 0; e < P->c.E; ++e) ve[(size_t)i * P->c.E + e] = P->h_members[i] * P->c.E + e;
  CK(cudaMemcpy(P->member_list, P->h_members.data(), sizeof(int) * n, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(P->member_list_e, ve.data(), sizeof(int) * ve.size(), cudaMemcpyHostToDevice));
  const int ne = n * P->c.E;
  CK(cudaMemcpy(P->n_active, &n, sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(P->n_active_e, &ne, sizeof(int), cudaMemcpyHostToDevice));
  P->g_off = 0;
  P->g_n = n;
  if (P->plan) return make_groups(P, err);
  return 0;
}

int pop_create_impl(const PopParams& c, const float* alphas, const unsigned long long* seeds, Population** out,
                    std::string* err) {
  auto bad = [&](const char* m) {
    if (err) *err = m;
    return 1;
  };
  if (c.P <= 0 || c.B <= 0 || c.O <= 0 || c.A <= 0) return bad("P, B, O, A must be positive");
  if (c.B % 64) return bad("batch_size must be a multiple of 64 (the weight-gradient GEMM reduces over 64-row blocks)");
  if (c.A > kMaxHeadOut) return bad("action_dim > 8 is not supported by the head kernels");
  if (c.E < 1 || c.E > 4) return bad("num_ensembles must be in [1, 4]");
  if (c.H_actor % 64 || c.H_value % 64 || c.H_actor <= 0 || c.H_value <= 0 || c.H_actor > 512 || c.H_value > 512)
    return bad("hidden widths must be multiples of 64, at most 512");
  if ((c.layer_norm && c.H_value != 512) || (c.actor_layer_norm && c.H_actor != 512))
    return bad("LayerNorm networks need hidden width 512 (one CTA owns the full 512-wide row)");
  if (c.n_hidden_actor < 1 || c.n_hidden_value < 1 || c.n_hidden_actor >= kMaxLayers || c.n_hidden_value >= kMaxLayers)
    return bad("between 1 and 7 hidden layers are supported");
  if (c.F < 1) return bad("flow_steps must be >= 1");
  if (cudaError_t e = gemm_init(); e != cudaSuccess) {
    if (err) *err = std::string("gemm_init: ") + cudaGetErrorString(e);
    return 1;
  }
  if (cudaError_t e = chain_init(); e != cudaSuccess) {
    if (err) *err = std::string("chain_init: ") + cudaGetErrorString(e);
    return 1;
  }
  if (int rc = pointwise_init()) {
    if (err) *err = std::string("pointwise_init: ") + cudaGetErrorString((cudaError_t)rc);
    return 1;
  }
  Population* P = new Population;
  P->c = c;
  const int V = c.P * c.E;
  const NetLayout Lq = make_layout(c.E, c.O + c.A, c.H_value, c.n_hidden_value, 1, c.layer_norm);
  const NetLayout Lb = make_layout(1, c.O + c.A + 1, c.H_actor, c.n_hidden_actor, c.A, c.actor_layer_norm);
  const NetLayout Lp = make_layout(1, c.O + c.A, c.H_actor, c.n_hidden_actor, c.A, c.actor_layer_norm);
  int rc = 0;
  rc |= alloc_arena(P, P->critic, Lq, true, err);
  rc |= alloc_arena(P, P->bc, Lb, false, err);
  rc |= alloc_arena(P, P->pi, Lp, false, err);
  rc = rc || alloc_pass(P, P->PI, Lp, c.P, 3 * c.B, true, err);
  rc = rc || alloc_pass(P, P->TQ, Lq, V, c.B, false, err);
  rc = rc || alloc_pass(P, P->QC, Lq, V, c.B, true, err);
  rc = rc || alloc_pass(P, P->QA, Lq, V, c.B, true, err);
  rc = rc || alloc_pass(P, P->BC, Lb, c.P, c.B, true, err);
  rc = rc || alloc_pass(P, P->EU, Lb, c.P, c.B, false, err);
  rc = rc || alloc_bwd(P, P->bQC, Lq, V, c.B, true, err);
  rc = rc || alloc_bwd(P, P->bQA, Lq, V, c.B, false, err);
  rc = rc || alloc_bwd(P, P->bBC, Lb, c.P, c.B, true, err);
  rc = rc || alloc_bwd(P, P->bPI, Lp, c.P, c.B, true, err);
  // The layer-chain kernels need every network input to fit one 64-wide block (true for all OGBench shapes the
  // reference uses) and at most kChainMaxLayers GEMMs per chain; otherwise the per-layer kernels are used.
  P->use_chain = Lq.in_pad == 64 && Lb.in_pad == 64 && Lp.in_pad == 64 && c.n_hidden_actor + 1 <= kChainMaxLayers &&
                 c.n_hidden_value + 1 <= kChainMaxLayers;
#ifdef FQLPOP_DEBUG_SWITCHES
  if (getenv("FQLPOP_NO_CHAIN")) P->use_chain = false;
#endif
  if (getenv("FQLPOP_CHAIN_TRACE")) rc = rc || dev_alloc(P, &P->chain_trace, 8192, err);
  if (getenv("FQLPOP_TIMELINE")) rc = rc || dev_alloc(P, &P->timeline, 8 + 4 * 65536, err);
  for (BwdBuf* bw : {&P->bQC, &P->bQA, &P->bBC, &P->bPI}) bw->part_rows = bw->m_tiles * (P->use_chain ? 4 : 1);
  const long long PB = (long long)c.P * c.B;
  rc = rc || dev_alloc(P, &P->act_f, PB * c.A, err) || dev_alloc(P, &P->vel, PB * c.A, err) ||
       dev_alloc(P, &P->xe, PB * c.A, err) || dev_alloc(P, &P->target_a, PB * c.A, err) ||
       dev_alloc(P, &P->api, PB * c.A, err) || dev_alloc(P, &P->rew, PB, err) || dev_alloc(P, &P->mask, PB, err) ||
       dev_alloc(P, &P->y, PB, err) || dev_alloc(P, &P->dx0_qa, (long long)V * c.B * Lq.in_pad, err) ||
       dev_alloc(P, &P->info, (long long)c.P * kInfoSlots, err) ||
       dev_alloc(P, &P->leaf_stats, (long long)c.P * 64 * 4, err) || dev_alloc(P, &P->adam_bc, 2LL * c.P, err) ||
       dev_alloc(P, &P->adam_step, c.P, err) || dev_alloc(P, &P->alphas, c.P, err) ||
       dev_alloc(P, &P->seeds, c.P, err) || dev_alloc(P, &P->step_counter, c.P, err) ||
       dev_alloc(P, &P->member_list, c.P, err) || dev_alloc(P, &P->member_list_e, V, err) ||
       dev_alloc(P, &P->n_active, 1, err) || dev_alloc(P, &P->n_active_e, 1, err);
  if (rc) {
    pop_destroy_impl(P);
    return 1;
  }
  cudaMemcpy(P->alphas, alphas, sizeof(float) * c.P, cudaMemcpyHostToDevice);
  cudaMemcpy(P->seeds, seeds, sizeof(unsigned long long) * c.P, cudaMemcpyHostToDevice);
  P->h_members.resize(c.P);
  for (int i = 0; i < c.P; ++i) P->h_members[i] = i;
  if (upload_members(P, err)) {
    pop_destroy_impl(P);
    return 1;
  }
  // parameters: Glorot-uniform kernels, zero biases, LayerNorm scale 1, target = critic
  int net = 0;
  for (Arena* ar : {&P->critic, &P->bc, &P->pi}) {
    init_arena_kernel<<<dim3(64, ar->V), 256>>>(ar->theta, ar->L, ar->V, P->seeds, net++);
  }
  cudaMemcpy(P->critic.target, P->critic.theta, sizeof(float) * (size_t)P->critic.V * P->critic.L.S,
             cudaMemcpyDeviceToDevice);
  launch_refresh_shadow(P, 0);
  launch_fill_qa_dout(P, 0);
  P->plan = new Plan;
  if (make_groups(P, err)) {
    pop_destroy_impl(P);
    return 1;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    if (err) *err = std::string("pop_create: ") + cudaGetErrorString(e);
    pop_destroy_impl(P);
    return 1;
  }
  *out = P;
  return 0;
}

void pop_destroy_impl(Population* P) {
  if (!P) return;
  cudaDeviceSynchronize();
  if (P->plan) {
    destroy_groups(P->plan);
    delete P->plan;
  }
  for (void* p : P->allocs) cudaFree(p);
  delete P;
}

void pop_invalidate_plan(Population* P) {
  if (P->plan) {
    cudaDeviceSynchronize();
    destroy_graphs(P->plan);
  }
}

int pop_set_active_impl(Population* P, const int* members, int n, std::string* err) {
  if (n < 1 || n > P->c.P) {
    if (err) *err = "active member count out of range";
    return 1;
  }
  std::vector<char> seen(P->c.P, 0);
  for (int i = 0; i < n; ++i) {
    if (members[i] < 0 || members[i] >= P->c.P || seen[members[i]]) {
      if (err) *err = "invalid or duplicate member id";
      return 1;
    }
    seen[members[i]] = 1;
  }
  pop_invalidate_plan(P);
  P->h_members.assign(members, members + n);
  return upload_members(P, err);
}

static int run_plan(Population* P, bool train, int n_steps, cudaStream_t stream, std::string* err) {
  Plan* pl = P->plan;
  if (!P->ds.obs) {
    if (err) *err = "no dataset / batch bound (call fqlpop_set_dataset first)";
    return 1;
  }
#ifdef FQLPOP_DEBUG_SWITCHES
  const bool eager = getenv("FQLPOP_EAGER") != nullptr;
#else
  constexpr bool eager = false;
#endif
  cudaEvent_t e_in;
  CK(cudaEventCreateWithFlags(&e_in, cudaEventDisableTiming));
  CK(cudaEventRecord(e_in, stream));
  int rc = 0;
  const bool fine = train && n_steps == 1 && !eager && !pl->groups1.empty();
  std::vector<GroupPlan>& groups = fine ? pl->groups1 : pl->groups;
  if (!eager && !(train ? groups[0].x_step : groups[0].x_loss)) {
    // (launches_per_step describes the split of the K-step calls; if only single-step calls are ever made, theirs)
    const int keep = pl->launches_per_step;
    if (train) pl->launches_per_step = 0;
    for (GroupPlan& gp : groups)
      if (!rc) rc = build_graph(P, train, gp, err);
    if (train && fine && keep > 0) pl->launches_per_step = keep;
  }
  // multi-step calls: every step but the last runs the variant that does not store the weight gradients
  if (!eager && train && n_steps > 1 && !groups[0].x_fast) {
    const int keep = pl->launches_per_step;
    for (GroupPlan& gp : groups)
      if (!rc) rc = build_graph(P, true, gp, err, false);
    pl->launches_per_step = keep;
  }
  // every group runs its n_steps back to back on its own streams; the groups only meet again on `stream`. The launches
  // are submitted ROUND-ROBIN over the groups: the stream launch queue is finite (~1k pending launches), so submitting
  // all n_steps of group 0 first would block the host inside group 0's loop and serialise the groups for large n_steps
  // (the Trainer steps in chunks of log_interval = 5000).
  for (GroupPlan& gp : groups) cudaStreamWaitEvent(gp.s[0], e_in, 0);
  if (eager) {
    for (GroupPlan& gp : groups) {
      if (rc) break;
      Window w(P, gp);
      if (train) pl->launches_per_step = 0;
      for (int i = 0; i < n_steps && !rc; ++i) rc = issue_step(P, train, gp, err);
    }
  } else {
    for (int i = 0; i < n_steps && !rc; ++i) {
      for (GroupPlan& gp : groups) {
        cudaGraphExec_t x = !train ? gp.x_loss : (i + 1 < n_steps ? gp.x_fast : gp.x_step);
        cudaError_t e = cudaGraphLaunch(x, gp.s[0]);
        if (e != cudaSuccess) {
          if (err) *err = std::string("cudaGraphLaunch: ") + cudaGetErrorString(e);
          rc = 1;
          break;
        }
      }
    }
  }
  if (!rc) {
    for (GroupPlan& gp : groups) {
      cudaEventRecord(gp.done, gp.s[0]);
      cudaStreamWaitEvent(stream, gp.done, 0);
    }
  }
  cudaEventDestroy(e_in);
  return rc;
}

int pop_step_impl(Population* P, int n_steps, cudaStream_t stream, std::string* err) {
  return run_plan(P, true, n_steps, stream, err);
}
int pop_total_loss_impl(Population* P, cudaStream_t stream, std::string* err) {
  return run_plan(P, false, 1, stream, err);
}
int pop_refresh_shadow_impl(Population* P, cudaStream_t stream, std::string* err) {
  launch_refresh_shadow(P, stream);
  CK(cudaGetLastError());
  return 0;
}
int pop_launches_per_step(const Population* P) { return P->plan ? P->plan->launches_per_step : 0; }

// (measurement) `iters` back-to-back launches, on `stream`, of one layer-chain launch of the step covering every
// active member: which = 0 the Euler flow chain, 1 the one-step policy forward pass (3B rows, first B saved), 2 the critic
// forward pass (LayerNorm + saves), 3 the critic input-gradient chain, 4 the target-critic forward pass.
// The pass buffers must hold a step's inputs (call after fqlpop_step). Leaves the population state untouched except
// for those scratch buffers.
int pop_debug_chain_launch(Population* P, int which, int iters, cudaStream_t stream, std::string* err) {
  if (!P->use_chain) {
    if (err) *err = "layer-chain kernels are not in use for this shape";
    return 1;
  }
  ChainBuilder cb{P, err};
  cb.L = new ChainLaunch;
  const int B = P->c.B;
  for (int i = 0; i < iters && cb.ok; ++i) {
    int mode = 0;
    if (which == 0) {
      cb.begin(1);
      cb.euler_job();
    } else if (which == 1) {
      cb.begin(P->pi.L.nb);
      cb.fwd_job(P->PI, P->pi, P->pi.theta, P->pi.shadow, 0, B, true);
      cb.fwd_job(P->PI, P->pi, P->pi.theta, P->pi.shadow, B, 2 * B, false);
    } else if (which == 2) {  // critic forward pass (LayerNorm, tensors saved for the backward pass)
      cb.begin(P->critic.L.nb);
      cb.fwd_job(P->QC, P->critic, P->critic.theta, P->critic.shadow, 0, B, true);
    } else if (which == 3) {  // critic input-gradient chain (LayerNorm backward, bias / LayerNorm gradient partials)
      cb.begin(P->critic.L.nb);
      cb.bwd_job(P->QC, P->bQC, P->critic, 0, true, nullptr);
      mode = 1;
    } else {                  // target-critic forward pass (LayerNorm, nothing saved)
      cb.begin(P->critic.L.nb);
      cb.fwd_job(P->TQ, P->critic, P->critic.target, P->critic.tshadow, 0, B, false);
    }
    cb.launch(mode, stream);
  }
  delete cb.L;
  if (!cb.ok) return 1;
  CK(cudaGetLastError());
  return 0;
}

// ---- member migration (survivors of a successive-halving rung are spread over the GPUs again) ------------
// One member's whole training state as ONE contiguous device blob of 32-bit words, so that the host side can hand it to
// a single NCCL send / recv: an 8-word header {adam_step, alpha (fp32 bits), seed lo/hi, Philox step counter lo/hi,
// blob words, layout check} followed by critic theta / mu / nu / target (E blocks each), BC-flow theta / mu / nu and
// one-step-policy theta / mu / nu, each exactly as it lies in its arena (padded kernels included: padding is zero on
// both sides). Gradients, bf16 shadows and activations are not state: the shadows are rebuilt on import.
constexpr int kMemberHeaderWords = 8;
static unsigned member_layout_check(const Population* P) {
  // two populations exchange members only if their arenas have the same shape
  unsigned long long h = 1469598103934665603ull;
  for (const Arena* ar : {&P->critic, &P->bc, &P->pi})
    for (long long v : {(long long)ar->L.nb, ar->L.S, (long long)ar->L.in_pad, (long long)ar->L.H, (long long)ar->L.n_hidden,
                        (long long)ar->L.out_dim, (long long)ar->L.ln})
      h = (h ^ (unsigned long long)v) * 1099511628211ull;
  return (unsigned)(h ^ (h >> 32));
}
long long pop_member_state_words(const Population* P) {
  long long n = kMemberHeaderWords;
  for (const Arena* ar : {&P->critic, &P->bc, &P->pi}) n += (long long)ar->L.nb * ar->L.S * (ar->target ? 4 : 3);
  return n;
}
__global__ void member_header_kernel(unsigned* blob, int* adam_step, float* alphas, unsigned long long* seeds,
                                     unsigned long long* step_counter, int member, unsigned words, unsigned check,
                                     int import) {
  if (threadIdx.x || blockIdx.x) return;
  if (!import) {
    blob[0] = (unsigned)adam_step[member];
    blob[1] = __float_as_uint(alphas[member]);
    blob[2] = (unsigned)seeds[member]; blob[3] = (unsigned)(seeds[member] >> 32);
    blob[4] = (unsigned)step_counter[member]; blob[5] = (unsigned)(step_counter[member] >> 32);
    blob[6] = words; blob[7] = check;
  } else {
    adam_step[member] = (int)blob[0];
    alphas[member] = __uint_as_float(blob[1]);
    seeds[member] = (unsigned long long)blob[2] | ((unsigned long long)blob[3] << 32);
    step_counter[member] = (unsigned long long)blob[4] | ((unsigned long long)blob[5] << 32);
  }
}
int pop_member_io(Population* P, int member, float* blob, int import, cudaStream_t stream, std::string* err) {
  if (member < 0 || member >= P->c.P || !blob) {
    if (err) *err = "member_io: bad member or null buffer";
    return 1;
  }
  const unsigned words = (unsigned)pop_member_state_words(P), check = member_layout_check(P);
  if (import) {   // the header is checked on the host before a single word of the arenas is touched
    unsigned head[kMemberHeaderWords];
    CK(cudaMemcpyAsync(head, blob, sizeof(head), cudaMemcpyDeviceToHost, stream));
    CK(cudaStreamSynchronize(stream));
    if (head[6] != words || head[7] != check) {
      if (err) *err = "member_import: the blob was not exported by a population of the same network shapes";
      return 1;
    }
  }
  member_header_kernel<<<1, 32, 0, stream>>>(reinterpret_cast<unsigned*>(blob), P->adam_step, P->alphas, P->seeds,
                                             P->step_counter, member, words, check, import);
  float* w = blob + kMemberHeaderWords;
  for (Arena* ar : {&P->critic, &P->bc, &P->pi}) {
    const long long n = (long long)ar->L.nb * ar->L.S, off = (long long)member * n;
    for (float* base : {ar->theta, ar->mu, ar->nu, ar->target}) {
      if (!base) continue;
      CK(cudaMemcpyAsync(import ? base + off : w, import ? w : base + off, sizeof(float) * (size_t)n,
                         cudaMemcpyDeviceToDevice, stream));
      w += n;
    }
    if (import) {
      launch_to_bf16(ar->theta + off, ar->shadow + off, n / 4, stream);
      if (ar->target) launch_to_bf16(ar->target + off, ar->tshadow + off, n / 4, stream);
    }
  }
  CK(cudaGetLastError());
  return 0;
}

// leaf I/O: kind 0 kernel [nb][K_true][N_true], 1 bias [nb][N_true], 2 LayerNorm scale, 3 LayerNorm bias [nb][H];
// which 0 theta, 1 mu, 2 nu, 3 target, 4 grad. Host buffers are unpadded fp32.
int pop_leaf_io(Population* P, int net, int layer, int kind, int which, int member, float* host, int to_host,
                std::string* err) {
  Arena* ar = net == 0 ? &P->critic : (net == 1 ? &P->bc : (net == 2 ? &P->pi : nullptr));
  if (!ar || layer < 0 || layer > ar->L.n_hidden || member < 0 || member >= P->c.P) {
    if (err) *err = "leaf_io: bad net/layer/member";
    return 1;
  }
  float* base = which == 0 ? ar->theta : which == 1 ? ar->mu : which == 2 ? ar->nu : which == 3 ? ar->target
                                                                                 : which == 4 ? ar->grad : nullptr;
  if (!base) {
    if (err) *err = "leaf_io: this arena has no such array";
    return 1;
  }
  const NetLayout& L = ar->L;
  long long off;
  int rows, cols, ld;
  if (kind == 0) { off = L.offW[layer]; rows = L.K_true(layer); cols = L.N_true(layer); ld = L.N(layer); }
  else if (kind == 1) { off = L.offb[layer]; rows = 1; cols = L.N_true(layer); ld = L.N(layer); }
  else if ((kind == 2 || kind == 3) && L.ln && layer < L.n_hidden) {
    off = kind == 2 ? L.offg[layer] : L.offbeta[layer]; rows = 1; cols = L.H; ld = L.H;
  } else {
    if (err) *err = "leaf_io: no such leaf";
    return 1;
  }
  for (int e = 0; e < L.nb; ++e) {
    float* d = base + ((long long)member * L.nb + e) * L.S + off;
    float* h = host + (long long)e * rows * cols;
    if (to_host) CK(cudaMemcpy2D(h, sizeof(float) * cols, d, sizeof(float) * ld, sizeof(float) * cols, rows,
                                 cudaMemcpyDeviceToHost));
    else CK(cudaMemcpy2D(d, sizeof(float) * ld, h, sizeof(float) * cols, sizeof(float) * cols, rows,
                         cudaMemcpyHostToDevice));
  }
  return 0;
}

}  // namespace fqlpop
