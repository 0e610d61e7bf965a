// Layer-chain kernels (see chain.h for what they replace).
//
// CTA = 10 warps: warp 0 = TMA producer (streams every layer's bf16 weight tiles, 16 KB stages), warp 1 = TMEM
// allocator + tcgen05.mma issuer, warps 2..9 = epilogue. Two epilogue warps share each TMEM lane quarter and split
// every 128-column accumulator sub-tile in halves, so a thread owns one row x 64 columns of each sub-tile.
// Shared memory: ACT (128 rows x 512 bf16 = the A operand of every hidden layer, K-major, 128-byte swizzle),
// X0 (128 x 64, the first layer's A operand, loaded by TMA), kStages weight stages, mbarriers.
// TMEM: 512 fp32 columns = one full 128 x 512 layer output, one mbarrier per 128-column sub-tile.
//
// Overlap inside a layer: the MMA warp issues the sub-tiles in order and commits each to its own barrier; the
// epilogue turns sub-tile s into bf16 *registers* while the tensor core works on sub-tile s+1 (ACT is still being
// read as the A operand, so it cannot be overwritten yet); once the last sub-tile has landed every MMA of the layer
// has retired and the kept registers are flushed into ACT for the next layer. LayerNorm layers park the activated
// values back in TMEM during the overlapped pass and normalise in a second pass once the row statistics are known.
#include "chain.h"

#include "gemm.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kThreads = 320;
constexpr int kStages = 5;
constexpr int kActBlockBytes = 128 * 64 * 2;   // one 64-wide k-block of the A operand = 16 KB
constexpr int kActBytes = 8 * kActBlockBytes;  // 128 KB
constexpr int kStageBytes = 16384;
constexpr int kBoxBytes = 8192;                // forward weight box: 64 n x 64 k bf16
constexpr int kBarBytes = 256;
constexpr int kSmemBytes = kActBytes + kActBlockBytes + kStages * kStageBytes + kBarBytes + 1024;

__device__ __forceinline__ float clip1(float x) { return fminf(fmaxf(x, -1.f), 1.f); }
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// lane l ends up with the sum over the warp's 32 rows of column l of a 32x32 block held one row per lane.
__device__ __forceinline__ float warp_colsum32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int j = 0; j < s; ++j) {
      const float send = up ? v[j] : v[j + s];
      const float keep = up ? v[j + s] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// 16 packed words (32 bf16) -> 64 contiguous bytes of global memory
__device__ __forceinline__ void st_global_packed32(__nv_bfloat16* dst, const uint32_t* w) {
  uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) d4[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
}
__device__ __forceinline__ void ld_global_bf16x32(const __nv_bfloat16* src, float (&v)[32]) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 u = __ldg(s4 + i);
    v[8 * i + 0] = bf16_lo(u.x); v[8 * i + 1] = bf16_hi(u.x);
    v[8 * i + 2] = bf16_lo(u.y); v[8 * i + 3] = bf16_hi(u.y);
    v[8 * i + 4] = bf16_lo(u.z); v[8 * i + 5] = bf16_hi(u.z);
    v[8 * i + 6] = bf16_lo(u.w); v[8 * i + 7] = bf16_hi(u.w);
  }
}
__device__ __forceinline__ void ld_global_f32x32(const float* src, float (&v)[32]) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 u = __ldg(s4 + i);
    v[4 * i + 0] = u.x; v[4 * i + 1] = u.y; v[4 * i + 2] = u.z; v[4 * i + 3] = u.w;
  }
}
// 4 of the 8 16-byte chunks of row r of a 128-byte-swizzled K-major block (chunks c0 .. c0 + 3) from 16 packed words
__device__ __forceinline__ void st_act_half_row(uint8_t* block, int r, int c0, const uint32_t* w) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(block + r * 128 + (((c0 + i) ^ (r & 7)) << 4)) =
        make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
}

// ---------------------------------------------------------------------------------------------------------------
// Forward hidden layer epilogue (MODE 0): h = LayerNorm?(gelu(acc + b)).
// Pass 1 runs sub-tile by sub-tile behind the tensor core (packed FFMA2 math, double-buffered tcgen05.ld): the bf16
// activations cannot go to ACT yet -- the MMAs of the later sub-tiles still read it -- so they are PARKED in the TMEM
// columns they were just drained from (16 packed columns per 32 accumulator columns). Pass 2, once every MMA of the
// layer has retired (and, for LayerNorm, the row statistics are complete), moves them TMEM -> registers -> ACT / HBM.
// Without LayerNorm the last sub-tile skips the detour.
// ---------------------------------------------------------------------------------------------------------------
struct EpiCtx {
  uint32_t tmem_row;
  uint64_t* tmem_full;
  uint32_t fullph;
  uint8_t* act;
  const float* params;
  int r, half, q, lane, H;
  bool row_ok;
  long long prow;    // element offset of this row in the pass buffers
  long long rs_off;  // offset of this row in the rstd buffers
  float ln_eps;
  long long* trace;  // non-null on the one traced thread
  int tr;
};

template <bool LN, bool SAVE>
__device__ __forceinline__ void fwd_hidden_layer(EpiCtx& C, const ChainLayer& Ly) {
  const int n_pad = Ly.n_pad;
  const int n_sub = (n_pad + 127) >> 7;
  const int my0 = C.half * 64;  // first column of this thread inside every 128-column sub-tile
  f32x2 s1 = pk2(0.f, 0.f), s2 = pk2(0.f, 0.f);
  const float* bias = C.params + Ly.off_b;

#pragma unroll 1
  for (int sub = 0; sub < n_sub; ++sub) {
    mbar_wait(&C.tmem_full[sub], (C.fullph >> sub) & 1u);
    C.fullph ^= 1u << sub;
    tc_fence_after();
    if (C.trace && C.tr < 4000) C.trace[4096 + C.tr++] = clock64();
    if (sub * 128 + my0 >= n_pad) continue;
    const bool direct = !LN && sub == n_sub - 1;  // every MMA of the layer has retired: ACT is free
    const int col0 = sub * 128 + my0;
    uint32_t rr[2][32];
    tmem_ld32(C.tmem_row + col0, rr[0]);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int col = col0 + c * 32;
      float4 b4[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) b4[i] = __ldg(reinterpret_cast<const float4*>(bias + col) + i);
      tmem_ld_wait();
      if (c == 0) tmem_ld32(C.tmem_row + col0 + 32, rr[1]);
      uint32_t pk[16], gpk[16];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f32x2 g0, g1, p0, p1;
        gelu2<SAVE>(add2(pk2u(rr[c][4 * i + 0], rr[c][4 * i + 1]), pk2(b4[i].x, b4[i].y)), g0, p0);
        gelu2<SAVE>(add2(pk2u(rr[c][4 * i + 2], rr[c][4 * i + 3]), pk2(b4[i].z, b4[i].w)), g1, p1);
        if (LN) {
          s1 = add2(s1, add2(g0, g1));
          s2 = fma2(g0, g0, fma2(g1, g1, s2));
        }
        pk[2 * i] = pack_bf16x2_v(g0);
        pk[2 * i + 1] = pack_bf16x2_v(g1);
        if (SAVE) {
          gpk[2 * i] = pack_bf16x2_v(p0);
          gpk[2 * i + 1] = pack_bf16x2_v(p1);
        }
      }
      if (SAVE && C.row_ok) st_global_packed32(Ly.out_gp + C.prow + col, gpk);
      if (direct) {
        st_act_half_row(C.act + (sub * 2 + C.half) * kActBlockBytes, C.r, c * 4, pk);
        if (C.row_ok && Ly.out_h) st_global_packed32(Ly.out_h + C.prow + col, pk);
      } else {
        tmem_st16(C.tmem_row + col0 + c * 16, pk);
      }
    }
  }
  tmem_st_wait();
  if (C.trace && C.tr < 4000) C.trace[4096 + C.tr++] = clock64();

  float mean = 0.f, rstd = 1.f;
  if (LN) {
    // exchange the half-row statistics through the (now free) first bytes of ACT
    float a0, a1, q0, q1;
    upk2(s1, a0, a1);
    upk2(s2, q0, q1);
    float t1 = a0 + a1, t2 = q0 + q1;
    float2* ex = reinterpret_cast<float2*>(C.act);
    ex[C.half * 128 + C.r] = make_float2(t1, t2);
    epi_bar_sync();
    const float2 o = ex[(C.half ^ 1) * 128 + C.r];
    epi_bar_sync();
    t1 += o.x;
    t2 += o.y;
    const float inv_n = 1.0f / (float)C.H;
    mean = t1 * inv_n;
    const float var = fmaxf(t2 * inv_n - mean * mean, 0.f);  // flax LayerNorm: E[x^2] - E[x]^2
    rstd = rsqrtf(var + C.ln_eps);
    if (C.row_ok && C.half == 0 && Ly.out_rstd) Ly.out_rstd[C.rs_off] = rstd;
  }
  const int n_park = LN ? n_sub : n_sub - 1;
  const f32x2 rstd2 = pk2(rstd, rstd), shift2 = pk2(-mean * rstd, -mean * rstd);
#pragma unroll 1
  for (int sub = 0; sub < n_park; ++sub) {
    if (sub * 128 + my0 >= n_pad) continue;
    const int col0 = sub * 128 + my0;
    uint32_t pr[2][16];
    tmem_ld16(C.tmem_row + col0, pr[0]);
    tmem_ld16(C.tmem_row + col0 + 16, pr[1]);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int col = col0 + c * 32;
      if (LN) {
        const float4* g4 = reinterpret_cast<const float4*>(C.params + Ly.off_g + col);
        const float4* e4 = reinterpret_cast<const float4*>(C.params + Ly.off_beta + col);
        uint32_t hk[16], xk[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 ga = __ldg(g4 + i), be = __ldg(e4 + i);
          const uint32_t w0 = pr[c][2 * i], w1 = pr[c][2 * i + 1];
          const f32x2 x0 = fma2(pk2u(w0 << 16, w0 & 0xFFFF0000u), rstd2, shift2);
          const f32x2 x1 = fma2(pk2u(w1 << 16, w1 & 0xFFFF0000u), rstd2, shift2);
          xk[2 * i] = pack_bf16x2_v(x0);
          xk[2 * i + 1] = pack_bf16x2_v(x1);
          hk[2 * i] = pack_bf16x2_v(fma2(x0, pk2(ga.x, ga.y), pk2(be.x, be.y)));
          hk[2 * i + 1] = pack_bf16x2_v(fma2(x1, pk2(ga.z, ga.w), pk2(be.z, be.w)));
        }
        st_act_half_row(C.act + (sub * 2 + C.half) * kActBlockBytes, C.r, c * 4, hk);
        if (C.row_ok) {
          if (Ly.out_h) st_global_packed32(Ly.out_h + C.prow + col, hk);
          if (SAVE && Ly.out_xhat) st_global_packed32(Ly.out_xhat + C.prow + col, xk);
        }
      } else {
        st_act_half_row(C.act + (sub * 2 + C.half) * kActBlockBytes, C.r, c * 4, pr[c]);
        if (C.row_ok && Ly.out_h) st_global_packed32(Ly.out_h + C.prow + col, pr[c]);
      }
    }
  }
}

template <int MODE>  // 0 = forward / Euler flow, 1 = backward (input-gradient chain)
__global__ void __launch_bounds__(kThreads, 1) fql_chain_kernel(const __grid_constant__ ChainLaunch L) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* x0s = smem + kActBytes;
  uint8_t* stage_base = x0s + kActBlockBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_base + kStages * kStageBytes);
  uint64_t* full_bar = bars;                  // [kStages]
  uint64_t* empty_bar = bars + kStages;       // [kStages]
  uint64_t* tmem_full = bars + 2 * kStages;   // [4] one per 128-column accumulator sub-tile
  uint64_t* act_ready = bars + 2 * kStages + 4;
  uint64_t* a0_full = bars + 2 * kStages + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 6);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---- tile decode ---------------------------------------------------------------------------------
  int ji = 0;
#pragma unroll 1
  for (int i = 1; i < L.num_jobs; ++i)
    if ((int)blockIdx.x >= L.job[i].tile_begin) ji = i;
  const ChainJob& J = L.job[ji];
  const int rel = (int)blockIdx.x - J.tile_begin;
  const int bidx = rel / J.m_tiles;
  const int mt = rel - bidx * J.m_tiles;
  const int v = L.member_list ? __ldg(L.member_list + bidx) : bidx;
  const int n_layers = J.n_layers;
  const int iters = J.iters;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&J.mapA0);
    for (int i = 0; i < n_layers; ++i) tma_prefetch_desc(&J.mapW[i]);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 4; ++s) mbar_init(&tmem_full[s], 1);
    mbar_init(act_ready, 256);
    mbar_init(a0_full, 1);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ========================================= TMA producer =========================================
    if (lane == 0) {
      mbar_arrive_expect_tx(a0_full, kActBlockBytes);
      tma_load_3d(x0s, &J.mapA0, a0_full, 0, J.row0 + mt * 128, v);
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < iters; ++it) {
        for (int li = 0; li < n_layers; ++li) {
          const ChainLayer& Ly = J.layer[li];
          for (int n0 = 0; n0 < Ly.n_pad; n0 += 128) {
            const int ncols = min(128, Ly.n_pad - n0);
            for (int kb = 0; kb < Ly.k_blocks; ++kb) {
              mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sb = stage_base + stage * kStageBytes;
              if (MODE == 0) {
                // W[k][n], n contiguous: boxes of 64 n x 64 k
                mbar_arrive_expect_tx(&full_bar[stage], (ncols / 64) * kBoxBytes);
                tma_load_3d(sb, &J.mapW[li], &full_bar[stage], n0, kb * 64, v);
                if (ncols > 64) tma_load_3d(sb + kBoxBytes, &J.mapW[li], &full_bar[stage], n0 + 64, kb * 64, v);
              } else {
                // the same array read as B^T: rows = output columns of this GEMM, 64 reduction elements per row
                mbar_arrive_expect_tx(&full_bar[stage], ncols * 128);
                tma_load_3d(sb, &J.mapW[li], &full_bar[stage], kb * 64, n0, v);
              }
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ========================================= MMA issuer ===========================================
    // The whole warp walks the loop (waits are warp-uniform); one elected lane issues the MMAs and commits.
    {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t itc = 0;
      int tr = 0;
      const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
      const uint64_t bdesc0 = MODE == 0 ? make_smem_desc_sw128(smem_u32(stage_base), kBoxBytes, 1024)
                                        : make_smem_desc_sw128(smem_u32(stage_base), 16, 1024);
      constexpr uint32_t kBStep = (MODE == 0 ? 2048u : 32u) >> 4;  // descriptor units (16 B) per K = 16 step
      for (int it = 0; it < iters; ++it) {
        for (int li = 0; li < n_layers; ++li, ++itc) {
          const ChainLayer& Ly = J.layer[li];
          if (itc == 0) mbar_wait(a0_full, 0);
          else mbar_wait(act_ready, (itc - 1) & 1);
          tc_fence_after();
          if (L.trace && blockIdx.x == 0 && lane == 0 && tr < 4000) L.trace[tr++] = clock64();
          const uint64_t adesc0 = make_smem_desc_sw128(Ly.a_from_x0 ? x0_addr : act_addr, 16, 1024);
          const int k_blocks = Ly.k_blocks;
          int sub = 0;
          for (int n0 = 0; n0 < Ly.n_pad; n0 += 128, ++sub) {
            const int ncols = min(128, Ly.n_pad - n0);
            const uint32_t idesc = make_idesc_bf16(128, (uint32_t)ncols, 0, MODE == 0 ? 1 : 0);
            const uint32_t tmem_d = tmem_base + n0;
            for (int kb = 0; kb < k_blocks; ++kb) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              if (elect_one()) {
                const uint64_t ad = adesc0 + (uint64_t)(kb * (kActBlockBytes >> 4));
                const uint64_t bd = bdesc0 + (uint64_t)(stage * (kStageBytes >> 4));
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  umma_bf16(tmem_d, ad + 2 * k, bd + kBStep * k, idesc, (kb | k) != 0 ? 1u : 0u);
                umma_commit(&empty_bar[stage]);
                if (kb == k_blocks - 1) umma_commit(&tmem_full[sub]);
              }
              __syncwarp();
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
            if (L.trace && blockIdx.x == 0 && lane == 0 && tr < 4000) L.trace[tr++] = clock64();
          }
        }
      }
    }
  } else {
    // ========================================= epilogue =============================================
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;    // which 64 columns of every 128-column sub-tile
    const int r = q * 32 + lane;         // row of the tile == TMEM lane
    const int grow = mt * 128 + r;       // row within the job
    const bool row_ok = grow < J.M;
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int H = J.H;
    const float* params = J.params + (long long)v * J.S;
    const long long prow = (long long)v * J.pass_ms + (long long)grow * H;  // element offset in pass buffers
    uint32_t fullph = 0;  // bit s: parity to wait for on tmem_full[s]
    auto wait_sub = [&](int sub) {
      mbar_wait(&tmem_full[sub], (fullph >> sub) & 1u);
      fullph ^= 1u << sub;
      tc_fence_after();
    };

    int etr = 0;
    float xs[8];  // Euler flow: the action iterate of this row (fp32)
    if (MODE == 0 && J.xe) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        xs[j] = (row_ok && half == 0 && j < J.A) ? J.xe[((long long)v * J.B + grow) * J.A + j] : 0.f;
    }

    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li) {
        const ChainLayer& Ly = J.layer[li];
        const int n_pad = Ly.n_pad;
        if (MODE == 0 && Ly.kind == CL_HIDDEN) {
          EpiCtx C{tmem_row, tmem_full, fullph, act, params, r, half, q, lane, H, row_ok, prow,
                   (long long)v * J.pass_rs + grow, J.ln_eps,
                   (L.trace && blockIdx.x == 0 && threadIdx.x == 64) ? L.trace : nullptr, etr};
          const bool save = Ly.out_gp != nullptr;
          if (J.ln) {
            if (save) fwd_hidden_layer<true, true>(C, Ly);
            else fwd_hidden_layer<true, false>(C, Ly);
          } else {
            if (save) fwd_hidden_layer<false, true>(C, Ly);
            else fwd_hidden_layer<false, false>(C, Ly);
          }
          fullph = C.fullph;
          etr = C.tr;
        } else if (MODE == 1 && Ly.kind == CL_HIDDEN && !J.ln) {
          // ------------------------------------------------------------------------------------------
          // bias + GELU (forward) / act'(z) product (backward); sub-tiles 0..2 are kept in registers while the
          // tensor core still reads ACT
          // ------------------------------------------------------------------------------------------
          uint32_t keep[3][32];
          float csum[4][2];
#pragma unroll
          for (int sub = 0; sub < 4; ++sub) {
            if (sub * 128 < n_pad) {
              wait_sub(sub);
              if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                  const int col = sub * 128 + half * 64 + c * 32;
                  uint32_t rr[32];
                  tmem_ld32(tmem_row + col, rr);
                  uint32_t pk[16];
                  if (MODE == 0) {
                    uint32_t gpk[16];
                    const float4* b4 = reinterpret_cast<const float4*>(params + Ly.off_b + col);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                      const float4 b = __ldg(b4 + i);
                      float g0, g1, g2, g3, p0, p1, p2, p3;
                      gelu_fwd(__uint_as_float(rr[4 * i + 0]) + b.x, g0, p0);
                      gelu_fwd(__uint_as_float(rr[4 * i + 1]) + b.y, g1, p1);
                      gelu_fwd(__uint_as_float(rr[4 * i + 2]) + b.z, g2, p2);
                      gelu_fwd(__uint_as_float(rr[4 * i + 3]) + b.w, g3, p3);
                      pk[2 * i] = pack_bf16x2(g0, g1);
                      pk[2 * i + 1] = pack_bf16x2(g2, g3);
                      gpk[2 * i] = pack_bf16x2(p0, p1);
                      gpk[2 * i + 1] = pack_bf16x2(p2, p3);
                    }
                    if (row_ok) {
                      if (Ly.out_h) st_global_packed32(Ly.out_h + prow + col, pk);
                      if (Ly.out_gp) st_global_packed32(Ly.out_gp + prow + col, gpk);
                    }
                  } else {
                    float gp[32];
                    if (Ly.gp && row_ok) ld_global_bf16x32(Ly.gp + prow + col, gp);
                    else {
#pragma unroll
                      for (int j = 0; j < 32; ++j) gp[j] = row_ok ? 1.f : 0.f;
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                      pk[i] = pack_bf16x2(__uint_as_float(rr[2 * i]) * gp[2 * i],
                                          __uint_as_float(rr[2 * i + 1]) * gp[2 * i + 1]);
                    if (row_ok) st_global_packed32(Ly.out_dz + (long long)v * J.dz_ms + (long long)grow * H + col, pk);
                    if (Ly.dbias_part) {
                      // column sums use the bf16-rounded dz, the same values the weight-gradient GEMM consumes
#pragma unroll
                      for (int i = 0; i < 16; ++i) { gp[2 * i] = bf16_lo(pk[i]); gp[2 * i + 1] = bf16_hi(pk[i]); }
                      csum[sub][c] = warp_colsum32(gp, lane);
                    }
                  }
                  if (sub < 3) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) keep[sub < 3 ? sub : 0][c * 16 + i] = pk[i];
                  } else {
                    st_act_half_row(act + (sub * 2 + half) * kActBlockBytes, r, c * 4, pk);
                  }
                }
              }
            }
          }
          // every MMA of this layer has retired: ACT may be overwritten
#pragma unroll
          for (int sub = 0; sub < 3; ++sub) {
            if (sub * 128 + half * 64 < n_pad) {
              if (sub * 128 + 128 >= n_pad) {
                // this was the last sub-tile of a narrow layer: its wait above already covered every MMA
              }
              st_act_half_row(act + (sub * 2 + half) * kActBlockBytes, r, 0, &keep[sub][0]);
              st_act_half_row(act + (sub * 2 + half) * kActBlockBytes, r, 4, &keep[sub][16]);
            }
          }
          if (MODE == 1 && Ly.dbias_part) {
            float* dst = Ly.dbias_part + (long long)v * J.part_ms + (long long)(mt * 4 + q) * H;
#pragma unroll
            for (int sub = 0; sub < 4; ++sub)
              if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                for (int c = 0; c < 2; ++c) dst[sub * 128 + half * 64 + c * 32 + lane] = csum[sub][c];
              }
          }
        } else if (MODE == 1 && Ly.kind == CL_HIDDEN) {
          // ------------------------------------------------------------------------------------------
          // LayerNorm layers: pass 1 (overlapped with the MMAs of later sub-tiles) parks per-element values in
          // TMEM and accumulates the row statistics; pass 2 normalises into ACT.
          // ------------------------------------------------------------------------------------------
          float s1 = 0.f, s2 = 0.f;
          float cs_a[4][2], cs_b[4][2], cs_c[4][2];
#pragma unroll
          for (int sub = 0; sub < 4; ++sub) {
            if (sub * 128 < n_pad) {
              wait_sub(sub);
              if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                  const int col = sub * 128 + half * 64 + c * 32;
                  uint32_t rr[32];
                  tmem_ld32(tmem_row + col, rr);
                  if (MODE == 0) {
                    uint32_t gpk[16];
                    const float4* b4 = reinterpret_cast<const float4*>(params + Ly.off_b + col);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                      const float4 b = __ldg(b4 + i);
                      float g0, g1, g2, g3, p0, p1, p2, p3;
                      gelu_fwd(__uint_as_float(rr[4 * i + 0]) + b.x, g0, p0);
                      gelu_fwd(__uint_as_float(rr[4 * i + 1]) + b.y, g1, p1);
                      gelu_fwd(__uint_as_float(rr[4 * i + 2]) + b.z, g2, p2);
                      gelu_fwd(__uint_as_float(rr[4 * i + 3]) + b.w, g3, p3);
                      s1 += (g0 + g1) + (g2 + g3);
                      s2 = fmaf(g0, g0, fmaf(g1, g1, fmaf(g2, g2, fmaf(g3, g3, s2))));
                      rr[4 * i + 0] = __float_as_uint(g0); rr[4 * i + 1] = __float_as_uint(g1);
                      rr[4 * i + 2] = __float_as_uint(g2); rr[4 * i + 3] = __float_as_uint(g3);
                      gpk[2 * i] = pack_bf16x2(p0, p1);
                      gpk[2 * i + 1] = pack_bf16x2(p2, p3);
                    }
                    tmem_st32(tmem_row + col, rr);  // park the activations in place of the accumulator
                    if (row_ok && Ly.out_gp) st_global_packed32(Ly.out_gp + prow + col, gpk);
                  } else {
                    // acc = dL/dh (h = xhat * gamma + beta): dxhat = acc * gamma
                    float xh[32], ga[32], t0[32];
                    if (row_ok) ld_global_bf16x32(Ly.xhat + prow + col, xh);
                    else {
#pragma unroll
                      for (int j = 0; j < 32; ++j) xh[j] = 0.f;
                    }
                    ld_global_f32x32(params + Ly.off_g + col, ga);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                      const float dh = row_ok ? __uint_as_float(rr[j]) : 0.f;
                      const float dxh = dh * ga[j];
                      s1 += dxh;
                      s2 = fmaf(dxh, xh[j], s2);
                      t0[j] = dh;
                      xh[j] = dh * xh[j];
                      rr[j] = __float_as_uint(dxh);
                    }
                    tmem_st32(tmem_row + col, rr);
                    if (Ly.dbeta_part) {
                      cs_b[sub][c] = warp_colsum32(t0, lane);   // dbeta = colsum(dh)
                      cs_c[sub][c] = warp_colsum32(xh, lane);   // dgamma = colsum(dh * xhat)
                    }
                  }
                }
              }
            }
          }
          tmem_st_wait();
          // exchange the half-row statistics through the (now free) first bytes of ACT
          {
            float2* ex = reinterpret_cast<float2*>(act);
            ex[half * 128 + r] = make_float2(s1, s2);
            epi_bar_sync();
            const float2 o = ex[(half ^ 1) * 128 + r];
            s1 += o.x;
            s2 += o.y;
            epi_bar_sync();
          }
          const float inv_n = 1.0f / (float)H;
          if (MODE == 0) {
            const float mean = s1 * inv_n;
            const float var = fmaxf(s2 * inv_n - mean * mean, 0.f);  // flax LayerNorm: E[x^2] - E[x]^2
            const float rstd = rsqrtf(var + J.ln_eps);
            if (row_ok && half == 0 && Ly.out_rstd) Ly.out_rstd[(long long)v * J.pass_rs + grow] = rstd;
#pragma unroll
            for (int sub = 0; sub < 4; ++sub) {
              if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                  const int col = sub * 128 + half * 64 + c * 32;
                  uint32_t rr[32];
                  tmem_ld32(tmem_row + col, rr);
                  const float4* g4 = reinterpret_cast<const float4*>(params + Ly.off_g + col);
                  const float4* e4 = reinterpret_cast<const float4*>(params + Ly.off_beta + col);
                  uint32_t pk[16], xpk[16];
                  tmem_ld_wait();
#pragma unroll
                  for (int i = 0; i < 8; ++i) {
                    const float4 ga = __ldg(g4 + i), be = __ldg(e4 + i);
                    const float x0 = (__uint_as_float(rr[4 * i + 0]) - mean) * rstd;
                    const float x1 = (__uint_as_float(rr[4 * i + 1]) - mean) * rstd;
                    const float x2 = (__uint_as_float(rr[4 * i + 2]) - mean) * rstd;
                    const float x3 = (__uint_as_float(rr[4 * i + 3]) - mean) * rstd;
                    xpk[2 * i] = pack_bf16x2(x0, x1);
                    xpk[2 * i + 1] = pack_bf16x2(x2, x3);
                    pk[2 * i] = pack_bf16x2(fmaf(x0, ga.x, be.x), fmaf(x1, ga.y, be.y));
                    pk[2 * i + 1] = pack_bf16x2(fmaf(x2, ga.z, be.z), fmaf(x3, ga.w, be.w));
                  }
                  st_act_half_row(act + (sub * 2 + half) * kActBlockBytes, r, c * 4, pk);
                  if (row_ok) {
                    if (Ly.out_h) st_global_packed32(Ly.out_h + prow + col, pk);
                    if (Ly.out_xhat) st_global_packed32(Ly.out_xhat + prow + col, xpk);
                  }
                }
              }
            }
          } else {
            // dg = rstd * (dxhat - mean(dxhat) - xhat * mean(dxhat * xhat)); dz = act'(z) * dg
            const float c1 = s1 * inv_n, c2 = s2 * inv_n;
            const float rstd = row_ok ? __ldg(Ly.rstd + (long long)v * J.pass_rs + grow) : 0.f;
#pragma unroll
            for (int sub = 0; sub < 4; ++sub) {
              if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                  const int col = sub * 128 + half * 64 + c * 32;
                  uint32_t rr[32];
                  tmem_ld32(tmem_row + col, rr);
                  float xh[32], gp[32];
                  if (row_ok) {
                    ld_global_bf16x32(Ly.xhat + prow + col, xh);
                    if (Ly.gp) ld_global_bf16x32(Ly.gp + prow + col, gp);
                    else {
#pragma unroll
                      for (int j = 0; j < 32; ++j) gp[j] = 1.f;
                    }
                  } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) { xh[j] = 0.f; gp[j] = 0.f; }
                  }
                  uint32_t pk[16];
                  tmem_ld_wait();
#pragma unroll
                  for (int i = 0; i < 16; ++i) {
                    const float d0 = gp[2 * i] * (rstd * (__uint_as_float(rr[2 * i]) - c1 - xh[2 * i] * c2));
                    const float d1 = gp[2 * i + 1] * (rstd * (__uint_as_float(rr[2 * i + 1]) - c1 - xh[2 * i + 1] * c2));
                    pk[i] = pack_bf16x2(d0, d1);
                  }
                  st_act_half_row(act + (sub * 2 + half) * kActBlockBytes, r, c * 4, pk);
                  if (row_ok) st_global_packed32(Ly.out_dz + (long long)v * J.dz_ms + (long long)grow * H + col, pk);
                  if (Ly.dbias_part) {
#pragma unroll
                    for (int i = 0; i < 16; ++i) { gp[2 * i] = bf16_lo(pk[i]); gp[2 * i + 1] = bf16_hi(pk[i]); }
                    cs_a[sub][c] = warp_colsum32(gp, lane);
                  }
                }
              }
            }
            if (Ly.dbias_part) {
              const long long o = (long long)v * J.part_ms + (long long)(mt * 4 + q) * H;
#pragma unroll
              for (int sub = 0; sub < 4; ++sub)
                if (sub * 128 + half * 64 < n_pad) {
#pragma unroll
                  for (int c = 0; c < 2; ++c) {
                    const int col = sub * 128 + half * 64 + c * 32 + lane;
                    Ly.dbias_part[o + col] = cs_a[sub][c];
                    if (Ly.dbeta_part) {
                      Ly.dbeta_part[o + col] = cs_b[sub][c];
                      Ly.dgamma_part[o + col] = cs_c[sub][c];
                    }
                  }
                }
            }
          }
        } else if (Ly.kind == CL_EULER_HEAD) {
          // ------------------------------------------------------------------------------------------
          // BC-flow head inside the Euler rollout: x <- x + v / F; the next step's input differs from this
          // step's only in the action columns and in t, which are patched in place in the X0 block.
          // ------------------------------------------------------------------------------------------
          wait_sub(0);
          if (half == 0) {
            uint32_t rr[32];
            tmem_ld32(tmem_row, rr);
            tmem_ld_wait();
            const float invF = 1.f / (float)J.F;
            const bool last = it == iters - 1;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j < J.A) {
                xs[j] = fmaf(__uint_as_float(rr[j]) + __ldg(params + Ly.off_b + j), invF, xs[j]);
                if (last) {
                  if (row_ok) J.target_a[((long long)v * J.B + grow) * J.A + j] = clip1(xs[j]);
                } else {
                  const int col = J.O + j;
                  *reinterpret_cast<__nv_bfloat16*>(x0s + r * 128 + (((col >> 3) ^ (r & 7)) << 4) + (col & 7) * 2) =
                      __float2bfloat16_rn(xs[j]);
                }
              }
            }
            if (!last) {
              const int col = J.O + J.A;
              *reinterpret_cast<__nv_bfloat16*>(x0s + r * 128 + (((col >> 3) ^ (r & 7)) << 4) + (col & 7) * 2) =
                  __float2bfloat16_rn((float)(it + 1) / (float)J.F);
            }
          }
        } else {  // CL_DINPUT
          wait_sub(0);
          if (half * 64 < n_pad && half == 0) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (c * 32 < n_pad) {
                uint32_t rr[32];
                tmem_ld32(tmem_row + c * 32, rr);
                tmem_ld_wait();
                if (row_ok) {
                  float4* d4 = reinterpret_cast<float4*>(Ly.out_f32 + (long long)v * J.f32_ms +
                                                         (long long)grow * Ly.out_ld + c * 32);
#pragma unroll
                  for (int i = 0; i < 8; ++i)
                    d4[i] = make_float4(__uint_as_float(rr[4 * i]), __uint_as_float(rr[4 * i + 1]),
                                        __uint_as_float(rr[4 * i + 2]), __uint_as_float(rr[4 * i + 3]));
                }
              }
            }
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        mbar_arrive(act_ready);
        if (L.trace && blockIdx.x == 0 && threadIdx.x == 64 && etr < 4000) L.trace[4096 + etr++] = clock64();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

int finalize_chain(ChainLaunch* L) {
  int t = 0;
  for (int i = 0; i < L->num_jobs; ++i) {
    ChainJob& j = L->job[i];
    j.m_tiles = (j.M + 127) / 128;
    j.tile_begin = t;
    t += j.m_tiles * L->n_virtual;
  }
  return t;
}

cudaError_t chain_init() {
  cudaError_t e = cudaFuncSetAttribute(fql_chain_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(fql_chain_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
}

cudaError_t launch_chain(const ChainLaunch& L, int mode, cudaStream_t stream) {
  int grid = 0;
  for (int i = 0; i < L.num_jobs; ++i) grid += L.job[i].m_tiles * L.n_virtual;
  if (grid <= 0) return cudaSuccess;
  if (mode == 0) fql_chain_kernel<0><<<grid, kThreads, kSmemBytes, stream>>>(L);
  else fql_chain_kernel<1><<<grid, kThreads, kSmemBytes, stream>>>(L);
  return cudaGetLastError();
}

}  // namespace fqlpop
