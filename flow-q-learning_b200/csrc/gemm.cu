// Population-batched grouped GEMM for sm_100a: TMA-staged bf16 operand tiles (128-byte swizzle),
// tcgen05.mma with fp32 accumulators in TMEM, fused epilogues read back with tcgen05.ld.
//
// CTA = 6 warps: warp 0 = TMA producer (one elected lane), warp 1 = TMEM allocator + MMA issuer (one
// elected lane), warps 2..5 = epilogue (one thread per accumulator row; TMEM lane quarter = warp_id % 4).
// One output tile (128 rows x 128*N_SUB columns) per CTA. N_SUB = 1: two CTAs are co-resident per SM, so
// one CTA's epilogue overlaps the other's main loop. N_SUB = 4: the CTA owns full 512-wide rows (all 512
// TMEM columns) so LayerNorm statistics are thread-local; the statistics pass of sub-tile s overlaps the
// MMAs of sub-tile s+1.
//
// Replaces: the Dense/GELU/LayerNorm chain XLA emits for FQLAgent.update
// (/root/reference/trainer/experiment.py:109) and its autodiff transposes.
#include "gemm.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kThreads = 192;
constexpr int kStageBytesA = kBlockM * kBlockK * 2;  // 16 KB
constexpr int kStageBytesB = kBlockN * kBlockK * 2;  // 16 KB
constexpr int kStageBytes = kStageBytesA + kStageBytesB;
constexpr int kBoxBytesMN = 64 * kBlockK * 2;  // one (64 mn x 64 k) box of an MN-major operand = 8 KB

template <int N_SUB>
struct Cfg {
  static constexpr int kStages = (N_SUB == 1) ? 3 : 5;
  static constexpr int kTmemCols = 128 * N_SUB;
  // epilogue scratch: column-sum partials [3 quantities][4 warps][128*N_SUB] floats
  static constexpr int kEpiBytes = 3 * 4 * 128 * N_SUB * 4;
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + kBarBytes + 1024 /*align slack*/;
};

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

__device__ __forceinline__ void store_bf16x32(__nv_bfloat16* dst, const float (&v)[32]) {
  uint4* d4 = reinterpret_cast<uint4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 u;
    u.x = pack_bf16x2(v[8 * i + 0], v[8 * i + 1]);
    u.y = pack_bf16x2(v[8 * i + 2], v[8 * i + 3]);
    u.z = pack_bf16x2(v[8 * i + 4], v[8 * i + 5]);
    u.w = pack_bf16x2(v[8 * i + 6], v[8 * i + 7]);
    d4[i] = u;
  }
}
__device__ __forceinline__ void load_bf16x32(const __nv_bfloat16* src, float (&v)[32]) {
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
__device__ __forceinline__ void load_f32x32(const float* src, float (&v)[32]) {
  const float4* s4 = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 u = __ldg(s4 + i);
    v[4 * i + 0] = u.x; v[4 * i + 1] = u.y; v[4 * i + 2] = u.z; v[4 * i + 3] = u.w;
  }
}
__device__ __forceinline__ void store_f32x32(float* dst, const float (&v)[32]) {
  float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
  for (int i = 0; i < 8; ++i) d4[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <int N_SUB>
__global__ void __launch_bounds__(kThreads, (N_SUB == 1) ? 2 : 1)
fql_gemm_kernel(const __grid_constant__ GemmLaunch L) {
  using C = Cfg<N_SUB>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  float* epi_scratch = reinterpret_cast<float*>(smem + C::kStages * kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kStages * kStageBytes + C::kEpiBytes);
  uint64_t* full_bar = bars;                     // [kStages]
  uint64_t* empty_bar = bars + C::kStages;       // [kStages]
  uint64_t* tmem_full_bar = bars + 2 * C::kStages;  // [N_SUB]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::kStages + N_SUB);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // ---- tile decode -------------------------------------------------------------------------
  const int tile = blockIdx.x;
  int gi = 0;
#pragma unroll 1
  for (int i = 1; i < L.num_groups; ++i)
    if (tile >= L.g[i].tile_begin) gi = i;
  const GemmGroup& G = L.g[gi];
  const int tiles_per_member = G.m_tiles * G.n_tiles;
  const int rel = tile - G.tile_begin;
  const int bidx = rel / tiles_per_member;
  const int rem = rel - bidx * tiles_per_member;
  const int mt = rem / G.n_tiles;
  const int nt = rem - mt * G.n_tiles;
  const int n_active = L.n_active ? __ldg(L.n_active) : L.P;
  if (bidx >= n_active) return;
  const int member = L.member_list ? __ldg(L.member_list + bidx) : bidx;

  const int k_blocks = G.k_blocks;
  const int n_base = nt * (kBlockN * N_SUB);  // first column of this CTA
  // valid columns of each sub-tile (N is a multiple of 64; a sub-tile is 64 or 128 wide)
  auto sub_cols = [&](int sub) {
    const int left = G.N - (n_base + sub * kBlockN);
    return left >= kBlockN ? kBlockN : (left > 0 ? left : 0);
  };

  // ---- one-time setup ------------------------------------------------------------------------
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&G.mapA);
    tma_prefetch_desc(&G.mapB);
    for (int s = 0; s < C::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < N_SUB; ++s) mbar_init(&tmem_full_bar[s], 1);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int sub = 0; sub < N_SUB; ++sub) {
        const int ncols = sub_cols(sub);
        if (ncols == 0) break;
        const int n0 = n_base + sub * kBlockN;
        const uint32_t bytes_b = G.b_mn ? (ncols / 64) * kBoxBytesMN : kStageBytesB;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = stage_base + stage * kStageBytes;
          uint8_t* sb = sa + kStageBytesA;
          mbar_arrive_expect_tx(&full_bar[stage], kStageBytesA + bytes_b);
          if (!G.a_mn) {
            tma_load_3d(sa, &G.mapA, &full_bar[stage], G.a_col0 + kb * kBlockK, G.a_row0 + mt * kBlockM, member);
          } else {
            tma_load_3d(sa, &G.mapA, &full_bar[stage], G.a_col0 + mt * kBlockM, G.a_row0 + kb * kBlockK, member);
            tma_load_3d(sa + kBoxBytesMN, &G.mapA, &full_bar[stage], G.a_col0 + mt * kBlockM + 64,
                        G.a_row0 + kb * kBlockK, member);
          }
          if (!G.b_mn) {
            tma_load_3d(sb, &G.mapB, &full_bar[stage], G.b_col0 + kb * kBlockK, G.b_row0 + n0, member);
          } else {
            tma_load_3d(sb, &G.mapB, &full_bar[stage], G.b_col0 + n0, G.b_row0 + kb * kBlockK, member);
            if (ncols > 64)
              tma_load_3d(sb + kBoxBytesMN, &G.mapB, &full_bar[stage], G.b_col0 + n0 + 64,
                          G.b_row0 + kb * kBlockK, member);
          }
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer =======================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      // per-UMMA (K = 16) advance of the descriptor start address, in bytes
      const uint32_t a_kstep = G.a_mn ? 2048u : 32u;
      const uint32_t b_kstep = G.b_mn ? 2048u : 32u;
      const uint32_t a_lbo = G.a_mn ? (uint32_t)kBoxBytesMN : 16u;
      const uint32_t b_lbo = G.b_mn ? (uint32_t)kBoxBytesMN : 16u;
      for (int sub = 0; sub < N_SUB; ++sub) {
        const int ncols = sub_cols(sub);
        if (ncols == 0) break;
        const uint32_t idesc = make_idesc_bf16(kBlockM, (uint32_t)ncols, G.a_mn, G.b_mn);
        const uint32_t tmem_d = tmem_base + sub * kBlockN;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(stage_base + stage * kStageBytes);
          const uint32_t sb = sa + kStageBytesA;
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(sa + k * a_kstep, a_lbo, 1024);
            const uint64_t bdesc = make_smem_desc_sw128(sb + k * b_kstep, b_lbo, 1024);
            umma_bf16(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // frees this smem stage once the MMAs above retire
          if (kb == k_blocks - 1) umma_commit(&tmem_full_bar[sub]);
          if (++stage == C::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================================== epilogue =========================================
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int et = q * 32 + lane;        // row within the tile == epilogue thread id 0..127
    const int grow = mt * kBlockM + et;  // row within the group
    const bool row_ok = grow < G.M;
    const EpiParams& E = G.epi;
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int kind = E.kind;
    const int ncols_total = kBlockN * N_SUB;
    float* part0 = epi_scratch;                            // [4][ncols_total]
    float* part1 = epi_scratch + 4 * ncols_total;          // [4][ncols_total]
    float* part2 = epi_scratch + 8 * ncols_total;          // [4][ncols_total]

    if (kind == EPI_STORE_F32 || kind == EPI_STORE_BF16 || kind == EPI_FWD || kind == EPI_DGRAD || kind == EPI_ADAM) {
      // ---------------- single-pass epilogues ----------------
      for (int sub = 0; sub < N_SUB; ++sub) {
        const int ncols = sub_cols(sub);
        if (ncols == 0) break;
        mbar_wait(&tmem_full_bar[sub], 0);
        tc_fence_after();
        for (int c = 0; c < ncols / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(tmem_row + sub * kBlockN + c * 32, r);
          tmem_ld_wait();
          const int lcol = sub * kBlockN + c * 32;  // column within the CTA tile
          const int col = n_base + lcol;            // column within the group
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          if (kind == EPI_ADAM) {
            // Weight-gradient tile -> optimizer update in place (optax.adam + Polyak + bf16 shadow refresh),
            // without the gradient ever travelling to HBM and back. Same transpose as the plain store below, then every
            // lane owns 16 contiguous bytes of 8 rows: theta / mu / nu (/ target) are read and written with coalesced
            // 16-byte accesses, 4 rows' worth of loads in flight per lane.
            if (N_SUB == 1) {
              // The warp's whole 32 x ncols block is staged first (16 KB of the idle operand stages per warp, 16-byte chunks
              // XOR-swizzled by row), then walked ROW by row: every load / store instruction of the update covers one
              // contiguous row segment of up to 512 bytes (lane = 16-byte chunk) instead of four 128-byte pieces of four
              // rows -- DRAM pages are touched once, not once per 32-column chunk.
              const AdamTile& T = E.adam;
              uint8_t* stg = stage_base + q * 16384;
#pragma unroll
              for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(stg + lane * 512 + (((c * 8 + i) ^ lane) & 31) * 16) =
                    make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
              if (c == ncols / 32 - 1) {
                __syncwarp();
                const int nch = ncols >> 2;   // 16-byte chunks per row of this tile
                const int rows_valid = G.M - (mt * kBlockM + q * 32);
                const long long base = member * E.out_f32.ms + (long long)(mt * kBlockM + q * 32) * E.out_f32.ld + n_base +
                                       sub * kBlockN + lane * 4;
                const int pm = member / T.nb;
                const float ibc1 = 1.f / __ldg(T.adam_bc + 2 * pm), ibc2 = 1.f / __ldg(T.adam_bc + 2 * pm + 1);
                const bool has_t = T.target != nullptr;
                if (lane < nch) {
#pragma unroll 1
                  for (int r0 = 0; r0 < 32; r0 += 4) {
                    float4 th[4], m[4], vv[4], tg[4];
                    long long o[4];
                    bool ok[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                      const int r = r0 + j;
                      ok[j] = r < rows_valid;
                      o[j] = base + (long long)(ok[j] ? r : 0) * E.out_f32.ld;
                      th[j] = __ldcs(reinterpret_cast<const float4*>(T.theta + o[j]));
                      m[j] = __ldcs(reinterpret_cast<const float4*>(T.mu + o[j]));
                      vv[j] = __ldcs(reinterpret_cast<const float4*>(T.nu + o[j]));
                      tg[j] = has_t ? __ldcs(reinterpret_cast<const float4*>(T.target + o[j])) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                      const int r = r0 + j;
                      const float4 g = *reinterpret_cast<const float4*>(stg + r * 512 + ((lane ^ r) & 31) * 16);
                      if (!ok[j]) continue;
#define ADAM1(c)                                                                \
  m[j].c = T.b1 * m[j].c + (1.f - T.b1) * g.c;                                  \
  vv[j].c = T.b2 * vv[j].c + (1.f - T.b2) * g.c * g.c;                          \
  th[j].c = th[j].c - T.lr * (m[j].c * ibc1) / (sqrtf(vv[j].c * ibc2) + T.eps);
                      ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
                      __stcs(reinterpret_cast<float4*>(T.theta + o[j]), th[j]);
                      __stcs(reinterpret_cast<float4*>(T.mu + o[j]), m[j]);
                      __stcs(reinterpret_cast<float4*>(T.nu + o[j]), vv[j]);
                      *reinterpret_cast<uint2*>(T.shadow + o[j]) =
                          make_uint2(pack_bf16x2(th[j].x, th[j].y), pack_bf16x2(th[j].z, th[j].w));
                      if (has_t) {
                        tg[j].x = T.tau * th[j].x + (1.f - T.tau) * tg[j].x;
                        tg[j].y = T.tau * th[j].y + (1.f - T.tau) * tg[j].y;
                        tg[j].z = T.tau * th[j].z + (1.f - T.tau) * tg[j].z;
                        tg[j].w = T.tau * th[j].w + (1.f - T.tau) * tg[j].w;
                        __stcs(reinterpret_cast<float4*>(T.target + o[j]), tg[j]);
                        *reinterpret_cast<uint2*>(T.tshadow + o[j]) =
                            make_uint2(pack_bf16x2(tg[j].x, tg[j].y), pack_bf16x2(tg[j].z, tg[j].w));
                      }
                      if (T.keep_grad) *reinterpret_cast<float4*>(E.out_f32.p + o[j]) = g;
                    }
                  }
                }
                __syncwarp();
              }
            }
          } else if (kind == EPI_STORE_F32) {
            if (N_SUB == 1) {
              // One tile per CTA: once the accumulator is complete the operand stages are idle. A thread owns a ROW, so
              // storing its 128 bytes directly costs 32 partial-line requests per warp instruction; transposing the
              // 32 x 32 block through a warp-private 4 KB slice of stage memory (16-byte chunks XOR-swizzled by row)
              // makes every store instruction cover 4 rows x 128 contiguous bytes.
              uint8_t* stg = stage_base + q * 4096;
#pragma unroll
              for (int i = 0; i < 8; ++i)
                *reinterpret_cast<float4*>(stg + lane * 128 + ((i ^ (lane & 7)) << 4)) =
                    make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
              __syncwarp();
              const int rows_valid = G.M - (mt * kBlockM + q * 32);
              float* dst = E.out_f32.p + member * E.out_f32.ms + (long long)(mt * kBlockM + q * 32) * E.out_f32.ld + col;
              const int ch = lane & 7, rsub = lane >> 3;
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int r = i * 4 + rsub;
                const float4 u = *reinterpret_cast<const float4*>(stg + r * 128 + ((ch ^ (r & 7)) << 4));
                if (r < rows_valid) *reinterpret_cast<float4*>(dst + (long long)r * E.out_f32.ld + ch * 4) = u;
              }
              __syncwarp();
            } else {
              if (row_ok) store_f32x32(E.out_f32.p + member * E.out_f32.ms + (long long)grow * E.out_f32.ld + col, v);
            }
          } else if (kind == EPI_STORE_BF16) {
            if (row_ok) store_bf16x32(E.out_h.p + member * E.out_h.ms + (long long)grow * E.out_h.ld + col, v);
          } else if (kind == EPI_FWD) {
            float b[32];
            if (E.bias.p) load_f32x32(E.bias.p + member * E.bias.ms + col, b);
            else {
#pragma unroll
              for (int j = 0; j < 32; ++j) b[j] = 0.f;
            }
            float gp[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float z = v[j] + b[j];
              if (E.act == ACT_GELU) {
                gelu_fwd(z, v[j], gp[j]);
              } else if (E.act == ACT_RELU) {
                v[j] = fmaxf(z, 0.f);
                gp[j] = z > 0.f ? 1.f : 0.f;
              } else {
                v[j] = z;
                gp[j] = 1.f;
              }
            }
            if (row_ok) {
              store_bf16x32(E.out_h.p + member * E.out_h.ms + (long long)grow * E.out_h.ld + col, v);
              if (E.out_gp.p) store_bf16x32(E.out_gp.p + member * E.out_gp.ms + (long long)grow * E.out_gp.ld + col, gp);
            }
          } else {  // EPI_DGRAD
            if (E.gp.p) {
              float gp[32];
              if (row_ok) load_bf16x32(E.gp.p + member * E.gp.ms + (long long)grow * E.gp.ld + col, gp);
              else {
#pragma unroll
                for (int j = 0; j < 32; ++j) gp[j] = 0.f;
              }
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] *= gp[j];
            }
            if (row_ok) store_bf16x32(E.out_h.p + member * E.out_h.ms + (long long)grow * E.out_h.ld + col, v);
            if (E.out_dbias_part.p) {
              // column sums use the bf16-rounded dz, the same values the weight-gradient GEMM consumes
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = row_ok ? __bfloat162float(__float2bfloat16_rn(v[j])) : 0.f;
              const float cs = warp_colsum32(v, lane);
              part0[q * ncols_total + lcol + lane] = cs;
            }
          }
        }
      }
      if (kind == EPI_DGRAD && E.out_dbias_part.p) {
        epi_bar_sync();
        for (int cc = et; cc < ncols_total; cc += 128) {
          if (n_base + cc < G.N) {
            const float s = part0[cc] + part0[ncols_total + cc] + part0[2 * ncols_total + cc] + part0[3 * ncols_total + cc];
            E.out_dbias_part.p[member * E.out_dbias_part.ms + (long long)mt * G.N + n_base + cc] = s;
          }
        }
      }
    } else if (kind == EPI_FWD_LN) {
      // ---------------- forward with LayerNorm over the full row (N_SUB == 4, N == 512) ------------
      float sum = 0.f, sumsq = 0.f;
      for (int sub = 0; sub < N_SUB; ++sub) {
        mbar_wait(&tmem_full_bar[sub], 0);
        tc_fence_after();
        for (int c = 0; c < kBlockN / 32; ++c) {
          uint32_t r[32];
          const uint32_t taddr = tmem_row + sub * kBlockN + c * 32;
          tmem_ld32(taddr, r);
          tmem_ld_wait();
          const int col = n_base + sub * kBlockN + c * 32;
          float b[32], gp[32];
          load_f32x32(E.bias.p + member * E.bias.ms + col, b);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float z = __uint_as_float(r[j]) + b[j];
            float g;
            if (E.act == ACT_GELU) {
              gelu_fwd(z, g, gp[j]);
            } else if (E.act == ACT_RELU) {
              g = fmaxf(z, 0.f);
              gp[j] = z > 0.f ? 1.f : 0.f;
            } else {
              g = z;
              gp[j] = 1.f;
            }
            sum += g;
            sumsq = fmaf(g, g, sumsq);
            r[j] = __float_as_uint(g);
          }
          tmem_st32(taddr, r);  // park the activations in place of the accumulator
          if (row_ok && E.out_gp.p)
            store_bf16x32(E.out_gp.p + member * E.out_gp.ms + (long long)grow * E.out_gp.ld + col, gp);
        }
      }
      tmem_st_wait();
      const float inv_n = 1.0f / (float)G.N;
      const float mean = sum * inv_n;
      const float var = fmaxf(sumsq * inv_n - mean * mean, 0.f);  // flax LayerNorm: E[x^2] - E[x]^2
      const float rstd = rsqrtf(var + E.ln_eps);
      if (row_ok && E.out_rstd.p) E.out_rstd.p[member * E.out_rstd.ms + grow] = rstd;
      for (int sub = 0; sub < N_SUB; ++sub) {
        for (int c = 0; c < kBlockN / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(tmem_row + sub * kBlockN + c * 32, r);
          tmem_ld_wait();
          const int col = n_base + sub * kBlockN + c * 32;
          float ga[32], be[32], xh[32], h[32];
          load_f32x32(E.gamma.p + member * E.gamma.ms + col, ga);
          load_f32x32(E.beta.p + member * E.beta.ms + col, be);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            xh[j] = (__uint_as_float(r[j]) - mean) * rstd;
            h[j] = fmaf(xh[j], ga[j], be[j]);
          }
          if (row_ok) {
            store_bf16x32(E.out_h.p + member * E.out_h.ms + (long long)grow * E.out_h.ld + col, h);
            if (E.out_xhat.p)
              store_bf16x32(E.out_xhat.p + member * E.out_xhat.ms + (long long)grow * E.out_xhat.ld + col, xh);
          }
        }
      }
    } else if (kind == EPI_DGRAD_LN) {
      // ---------------- LayerNorm backward + activation backward (N_SUB == 4, N == 512) -------------
      // acc = dL/dh (h = xhat*gamma + beta). dxhat = acc*gamma; dg = rstd*(dxhat - mean(dxhat) - xhat*mean(dxhat*xhat));
      // dz = dg * act'(z).  Partials: dbeta = colsum(acc), dgamma = colsum(acc*xhat), dbias = colsum(dz).
      float s1 = 0.f, s2 = 0.f;
      for (int sub = 0; sub < N_SUB; ++sub) {
        mbar_wait(&tmem_full_bar[sub], 0);
        tc_fence_after();
        for (int c = 0; c < kBlockN / 32; ++c) {
          uint32_t r[32];
          const uint32_t taddr = tmem_row + sub * kBlockN + c * 32;
          tmem_ld32(taddr, r);
          tmem_ld_wait();
          const int lcol = sub * kBlockN + c * 32;
          const int col = n_base + lcol;
          float xh[32], ga[32], t0[32], t1[32];
          if (row_ok) load_bf16x32(E.xhat.p + member * E.xhat.ms + (long long)grow * E.xhat.ld + col, xh);
          else {
#pragma unroll
            for (int j = 0; j < 32; ++j) xh[j] = 0.f;
          }
          load_f32x32(E.gamma.p + member * E.gamma.ms + col, ga);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float dh = row_ok ? __uint_as_float(r[j]) : 0.f;
            const float dxh = dh * ga[j];
            s1 += dxh;
            s2 = fmaf(dxh, xh[j], s2);
            t0[j] = dh;
            t1[j] = dh * xh[j];
            r[j] = __float_as_uint(dxh);
          }
          tmem_st32(taddr, r);
          if (E.out_dbeta_part.p) {
            const float cs0 = warp_colsum32(t0, lane);
            const float cs1 = warp_colsum32(t1, lane);
            part1[q * ncols_total + lcol + lane] = cs0;
            part2[q * ncols_total + lcol + lane] = cs1;
          }
        }
      }
      tmem_st_wait();
      const float inv_n = 1.0f / (float)G.N;
      const float c1 = s1 * inv_n, c2 = s2 * inv_n;
      const float rstd = row_ok ? __ldg(E.rstd.p + member * E.rstd.ms + grow) : 0.f;
      for (int sub = 0; sub < N_SUB; ++sub) {
        for (int c = 0; c < kBlockN / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(tmem_row + sub * kBlockN + c * 32, r);
          tmem_ld_wait();
          const int lcol = sub * kBlockN + c * 32;
          const int col = n_base + lcol;
          float xh[32], gp[32], dz[32];
          if (row_ok) {
            load_bf16x32(E.xhat.p + member * E.xhat.ms + (long long)grow * E.xhat.ld + col, xh);
            if (E.gp.p) load_bf16x32(E.gp.p + member * E.gp.ms + (long long)grow * E.gp.ld + col, gp);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) { xh[j] = 0.f; gp[j] = 0.f; }
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float dg = rstd * (__uint_as_float(r[j]) - c1 - xh[j] * c2);
            dz[j] = (E.gp.p ? gp[j] : 1.f) * dg;
          }
          if (row_ok) store_bf16x32(E.out_h.p + member * E.out_h.ms + (long long)grow * E.out_h.ld + col, dz);
          if (E.out_dbias_part.p) {
#pragma unroll
            for (int j = 0; j < 32; ++j) dz[j] = row_ok ? __bfloat162float(__float2bfloat16_rn(dz[j])) : 0.f;
            const float cs = warp_colsum32(dz, lane);
            part0[q * ncols_total + lcol + lane] = cs;
          }
        }
      }
      epi_bar_sync();
      for (int cc = et; cc < ncols_total; cc += 128) {
        const long long o = (long long)mt * G.N + n_base + cc;
        if (E.out_dbias_part.p)
          E.out_dbias_part.p[member * E.out_dbias_part.ms + o] =
              part0[cc] + part0[ncols_total + cc] + part0[2 * ncols_total + cc] + part0[3 * ncols_total + cc];
        if (E.out_dbeta_part.p) {
          E.out_dbeta_part.p[member * E.out_dbeta_part.ms + o] =
              part1[cc] + part1[ncols_total + cc] + part1[2 * ncols_total + cc] + part1[3 * ncols_total + cc];
          E.out_dgamma_part.p[member * E.out_dgamma_part.ms + o] =
              part2[cc] + part2[ncols_total + cc] + part2[2 * ncols_total + cc] + part2[3 * ncols_total + cc];
        }
      }
    }
  }

  // ---- teardown --------------------------------------------------------------------------------
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C::kTmemCols);
  }
}

}  // namespace

// ================================== host side ===================================================
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

int encode_tmap_bf16(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                     long long ld, long long batch_stride, int box_rows) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)batch_stride * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (batch <= 1) strides[1] = (cuuint64_t)ld * 2 * (cuuint64_t)rows;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

// One box = n_boxes adjacent 64-column blocks x box_rows rows, landing in shared memory as n_boxes consecutive
// [box_rows][64] 128-byte-swizzled tiles -- the layout n_boxes separate encode_tmap_bf16 boxes would give, for ONE
// cp.async.bulk.tensor instruction. Coordinates: (0, row, column block, batch). `inner` must be a multiple of 64.
int encode_tmap_bf16_wide(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                          long long ld, long long batch_stride, int box_rows, int n_boxes) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn || inner % 64 != 0) return -1;
  cuuint64_t dims[4] = {64, (cuuint64_t)rows, (cuuint64_t)(inner / 64), (cuuint64_t)(batch < 1 ? 1 : batch)};
  cuuint64_t strides[3] = {(cuuint64_t)ld * 2, 128, (cuuint64_t)batch_stride * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_rows, (cuuint32_t)n_boxes, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  if (batch <= 1) strides[2] = (cuuint64_t)ld * 2 * (cuuint64_t)rows;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

// Weights re-tiled as [K / ks][N / 64][ks k][64 n] (every 64 n x ks k box a contiguous run, the boxes of one k-slab
// adjacent; ks = 32 or 64): a box of n_boxes * ks rows is one CTA's weight stage, fetched from consecutive addresses.
// Coordinates: (0, (k_slab * (N / 64) + column block) * ks, batch).
int encode_tmap_bf16_tiled(CUtensorMap* out, const void* base, long long K, long long N, long long batch,
                           long long batch_stride, int n_boxes, int ks) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn || N % 64 != 0 || K % ks != 0 || ks * n_boxes > 256) return -1;
  cuuint64_t dims[3] = {64, (cuuint64_t)(K * (N / 64)), (cuuint64_t)(batch < 1 ? 1 : batch)};
  cuuint64_t strides[2] = {128, (cuuint64_t)batch_stride * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)(ks * n_boxes), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (batch <= 1) strides[1] = (cuuint64_t)K * N * 2;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

int encode_tmap_bf16_sw64(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                          long long ld, long long batch_stride, int box_rows) {
  PFN_encodeTiled fn = get_encode_fn();
  if (!fn) return -1;
  cuuint64_t dims[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)batch};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)batch_stride * 2};
  cuuint32_t box[3] = {32, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (batch <= 1) strides[1] = (cuuint64_t)ld * 2 * (cuuint64_t)rows;
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : (int)r;
}

int finalize_launch(GemmLaunch* L, int n_sub) {
  int t = 0;
  for (int i = 0; i < L->num_groups; ++i) {
    GemmGroup& g = L->g[i];
    g.m_tiles = (g.M + kBlockM - 1) / kBlockM;
    g.n_tiles = (g.N + kBlockN * n_sub - 1) / (kBlockN * n_sub);
    g.tile_begin = t;
    t += g.m_tiles * g.n_tiles * L->P;
  }
  L->tiles_per_member_total = L->P > 0 ? t / L->P : 0;
  return t;
}

cudaError_t gemm_init() {
  cudaError_t e = cudaFuncSetAttribute(fql_gemm_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<1>::kSmemBytes);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(fql_gemm_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<4>::kSmemBytes);
}

cudaError_t launch_gemm(const GemmLaunch& L, int n_sub, int grid, cudaStream_t stream) {
  if (grid <= 0) return cudaSuccess;
  if (n_sub == 1)
    fql_gemm_kernel<1><<<grid, kThreads, Cfg<1>::kSmemBytes, stream>>>(L);
  else
    fql_gemm_kernel<4><<<grid, kThreads, Cfg<4>::kSmemBytes, stream>>>(L);
  return cudaGetLastError();
}

}  // namespace fqlpop
