// Fused persistent world-model rollout kernel, second generation (see rollout.h for the reference path it replaces:
// evaluate_actor_fn /root/reference/evaluator/evaluation.py:84-114 driving Task.step
// /root/reference/task/offline_task_simulated.py:85-107, policy = FQLAgent.sample_actions).
//
// Same program as fql_rollout_kernel (rollout.cu) -- per lockstep step: one-step policy 4 x 512 (GELU) + head, state
// predictor (input LayerNorm, ReLU MLP, residual), termination predictor (ReLU MLP), episode bookkeeping -- with the
// epilogue organisation of the second-generation layer-chain kernel (chain16.cu):
//   * 20 warps: TMA weight producer, MMA issuer, 2 idle, 16 epilogue warps = 4 per TMEM lane quarter, one per COLUMN GROUP
//     cg. A thread owns one episode (row) and columns [64 cg, 64 cg + 64) of every 256-column accumulator tile.
//   * 512-wide layers: tile 0 is processed while tile 1's MMAs still read ACT and held in registers; ACT is handed to the
//     tensor core per 64-column block (blk_full[8]); `drained` barriers guard TMEM re-use. Narrower layers are one tile.
//   * The episode's observation (fp32) is DISTRIBUTED over the row's four column-group threads: thread cg keeps columns
//     [16 cg, 16 cg + 16) of the 64-wide observation in registers. The three 64-column "head" layers (policy head, state
//     output, termination logit) are read the same way -- 16 accumulator columns per thread -- so all 16 warps share that
//     work too; the only row-wide quantities (the input LayerNorm's two sums) cross the four threads through shared memory
//     and a 128-thread named barrier. Episode bookkeeping (done / success / length) lives in the cg = 0 thread.
//   * PAIR: clusters of two CTAs (adjacent episode tiles of a member), tcgen05.mma.cta_group::2, half of every weight stage
//     per CTA (as chain16.cu); the pair leaves the step loop together.
// Covers RO_MODE_ROLLOUT with a LayerNorm-free policy and layer widths of 64 / 128 / 256 / 512; rollout.cu keeps the
// single-step modes (sample_actions, env_step) and everything else.
#include <climits>

#include "philox.cuh"
#include "pop.h"
#include "ptx_sm100.cuh"
#include "rollout.h"

namespace fqlpop {

namespace {

constexpr int kEpiWarps = 16;
constexpr int kFirstEpiWarp = 4;
constexpr int kThreadsR = (kFirstEpiWarp + kEpiWarps) * 32;  // 640
constexpr int kStages = 5;
constexpr int kMaxSlots = 2 * kStages;
constexpr int kActBlockBytes = 128 * 64 * 2;
constexpr int kActBytes = 8 * kActBlockBytes;
constexpr int kStageBytes = 32 * 256 * 2;
constexpr int kBoxBytes = 64 * 32 * 2;
constexpr int kBarBytes = 1024;   // barriers and flags
constexpr int kSmemBytesR = kActBytes + kActBlockBytes + kStages * kStageBytes + kBarBytes + 1024;

#ifdef FQLPOP_EPI_TRACE
// (make TRACE=1) CTA (0,0): [0,2048) MMA issuer, 8 words per layer: start, cycles waited for drained / A operand / weight
// stages, end; [2048,4096) epilogue lane 0 of the cg = 0 and cg = 3 warps of quarter 0: enter, tile 0 ready, layer done
__device__ long long g_ro16_trace[4096];
// weight stages of lockstep step 1, policy layers 1-2 (64 stages): producer saw the slot empty / issued the TMA, MMA
// issuer saw the stage full / issued its MMAs
__device__ long long g_ro16_stage[4 * 64];
__device__ long long g_ro16_stage_peer[4 * 64];   // the same two producer stamps for CTA (1,0): the peer of a pair
#define RT_T0() const long long rt_t0 = clock64()
#define RT_ACC(x) x += clock64() - rt_t0
#else
#define RT_T0()
#define RT_ACC(x)
#endif

__device__ __forceinline__ float clip1(float x) { return fminf(fmaxf(x, -1.f), 1.f); }
__device__ __forceinline__ void sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
// 16 columns starting at `col` (multiple of 16) of row r -> ACT (two 16-byte chunks of the 128-byte-swizzled row)
__device__ __forceinline__ void st_act16(uint32_t act, int r, int col, const uint32_t* w) {
  const uint32_t row = act + (uint32_t)((col >> 6) * kActBlockBytes + r * 128);
  const int c0 = (col & 63) >> 3;
  sts128(row + ((c0 ^ (r & 7)) << 4), w[0], w[1], w[2], w[3]);
  sts128(row + (((c0 + 1) ^ (r & 7)) << 4), w[4], w[5], w[6], w[7]);
}
// this thread's 16 columns [16 cg, 16 cg + 16) of row r -> the 64-wide X0 block
__device__ __forceinline__ void st_x0_16(uint32_t x0, int r, int cg, const float (&x)[16]) {
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = pack_bf16x2(x[2 * i], x[2 * i + 1]);
  const uint32_t row = x0 + (uint32_t)(r * 128);
  sts128(row + (((2 * cg) ^ (r & 7)) << 4), w[0], w[1], w[2], w[3]);
  sts128(row + (((2 * cg + 1) ^ (r & 7)) << 4), w[4], w[5], w[6], w[7]);
}
__device__ __forceinline__ int ld_volatile_s32(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_s32(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void ld_vec16(const float* p, float4 (&v)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __ldg(reinterpret_cast<const float4*>(p) + i);
}

struct REpi {
  uint32_t tmem_row;
  uint64_t* tmem_full;
  uint32_t blk_full_a, drained_a, x0_ready_a;  // leader's barriers (shared::cluster address when this CTA is the peer)
  bool remote;
  uint32_t act, x0;
  int r, q, cg, lane;
};
__device__ __forceinline__ void arrive_on(const REpi& C, uint32_t addr) {
  if (C.remote) mbar_arrive_cluster(addr);
  else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
// this warp's rows of an A-operand block are written, its accumulator columns read (see chain16.cu publish)
__device__ __forceinline__ void publish(const REpi& C, uint32_t bar_addr) {
  tc_fence_before();
  fence_proxy_async_smem();
  __syncwarp();
  if (C.lane == 0) arrive_on(C, bar_addr);
}
__device__ __forceinline__ void arrive_drained(const REpi& C, int t) {
  tc_fence_before();
  __syncwarp();
  if (C.lane == 0) arrive_on(C, C.drained_a + 8u * (uint32_t)t);
}
__device__ __forceinline__ void wait_full(const REpi& C, int t, uint32_t parity) {
  mbar_wait(&C.tmem_full[t], parity);
  tc_fence_after();
}

// bias + activation of one 16-column chunk -> 8 packed bf16 pairs
template <int ACT>
__device__ __forceinline__ void act_chunk(const uint32_t (&rr)[16], const float4 (&b4)[4], uint32_t* pk) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float z0 = __uint_as_float(rr[4 * i + 0]) + b4[i].x, z1 = __uint_as_float(rr[4 * i + 1]) + b4[i].y;
    const float z2 = __uint_as_float(rr[4 * i + 2]) + b4[i].z, z3 = __uint_as_float(rr[4 * i + 3]) + b4[i].w;
    float g0, g1, g2, g3, u;
    if (ACT == ACT_GELU) {
      gelu1<false>(z0, g0, u); gelu1<false>(z1, g1, u); gelu1<false>(z2, g2, u); gelu1<false>(z3, g3, u);
    } else {
      g0 = fmaxf(z0, 0.f); g1 = fmaxf(z1, 0.f); g2 = fmaxf(z2, 0.f); g3 = fmaxf(z3, 0.f);
    }
    pk[2 * i] = pack_bf16x2(g0, g1);
    pk[2 * i + 1] = pack_bf16x2(g2, g3);
  }
}
// this thread's 64 columns of accumulator tile `tile`: DEST 0 keep in pk[32], 1 write to ACT
// (measured: handing every block of a tile over as soon as all 16 warps have written their 16 columns of it -- blocks
// completing one after the other -- costs more in fences than the earlier MMA start gains: 324 vs 328 M env-steps/s)
template <int ACT, int DEST>
__device__ __forceinline__ void hidden_pass(const REpi& C, const float* bias, int tile, uint32_t (&pk)[32]) {
  const int colw = tile * 256 + C.cg * 64;
  const uint32_t tbase = C.tmem_row + colw;
  uint32_t ra[16], rb[16];
  tmem_ld16(tbase, ra);
#pragma unroll
  for (int c = 0; c < 4; c += 2) {
    float4 b4[4];
    ld_vec16(bias + colw + c * 16, b4);
    tmem_ld_wait();
    tmem_ld16(tbase + (c + 1) * 16, rb);
    act_chunk<ACT>(ra, b4, pk + 8 * c);
    if (DEST == 1) st_act16(C.act, C.r, colw + c * 16, pk + 8 * c);
    ld_vec16(bias + colw + (c + 1) * 16, b4);
    tmem_ld_wait();
    if (c + 2 < 4) tmem_ld16(tbase + (c + 2) * 16, ra);
    act_chunk<ACT>(rb, b4, pk + 8 * (c + 1));
    if (DEST == 1) st_act16(C.act, C.r, colw + (c + 1) * 16, pk + 8 * (c + 1));
  }
}
// hidden layer h = act(acc + b) -> next layer's A operand. n_pad in {128, 256, 512}.
template <int ACT>
__device__ __forceinline__ void hidden_layer(const REpi& C, const float* bias, int n_pad, uint32_t par0, uint32_t par1) {
  uint32_t pk0[32];
  wait_full(C, 0, par0);
  if (n_pad > 256) {
    hidden_pass<ACT, 0>(C, bias, 0, pk0);
    arrive_drained(C, 0);
    wait_full(C, 1, par1);  // every MMA of the layer has retired: ACT may be overwritten
#pragma unroll
    for (int c = 0; c < 4; ++c) st_act16(C.act, C.r, C.cg * 64 + c * 16, pk0 + 8 * c);
    publish(C, C.blk_full_a + 8u * (uint32_t)C.cg);
    uint32_t pk1[32];
    hidden_pass<ACT, 1>(C, bias, 1, pk1);
    publish(C, C.blk_full_a + 8u * (uint32_t)(4 + C.cg));
    arrive_drained(C, 1);
  } else {
    // one accumulator tile: all of the layer's MMAs have retired, write straight to ACT
    if (C.cg * 64 < n_pad) {
      hidden_pass<ACT, 1>(C, bias, 0, pk0);
      publish(C, C.blk_full_a + 8u * (uint32_t)C.cg);
    }
    arrive_drained(C, 0);
  }
}
template <bool PAIR>
__global__ void __launch_bounds__(kThreadsR, 1) fql_rollout16_kernel(const __grid_constant__ RoParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* x0s = smem + kActBytes;
  uint8_t* stage_base = x0s + kActBlockBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_base + kStages * kStageBytes);
  uint64_t* full_bar = bars;                 // [kMaxSlots]
  uint64_t* empty_bar = bars + kMaxSlots;    // [kMaxSlots]
  uint64_t* tmem_full = bars + 2 * kMaxSlots;  // [2]
  uint64_t* blk_full = tmem_full + 2;        // [8]
  uint64_t* drained = blk_full + 8;          // [2]
  uint64_t* x0_ready = drained + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x0_ready + 1);
  int* first_done = reinterpret_cast<int*>(tmem_slot + 1);  // first step at which every episode of the CTA was done
  int* steps_done = first_done + 1;                         // lockstep steps fully book-kept
  int* peer_first_done = steps_done + 1;                    // (pair) the other CTA's two flags, written by it
  int* peer_steps_done = peer_first_done + 1;
  int* warp_all = peer_steps_done + 1;                      // [4]
  // a weight stage: 32 k x 256 n; in pair mode 64 k x this CTA's 128 n (half as many waits and commits per MMA for the one
  // thread that feeds the M = 256 products)
  constexpr int kSlotBytes = kStageBytes;
  constexpr int kSlots = kStages;
  constexpr int kStageK = PAIR ? 64 : 32;
  constexpr int kBoxB = kStageK * 128;          // one 64-column box of a stage
  constexpr uint32_t kCtas = PAIR ? 2 : 1;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int tile = PAIR ? 2 * ((int)blockIdx.x >> 1) + (int)rank : (int)blockIdx.x;
  const int member = P.member_list ? __ldg(P.member_list + blockIdx.y) : (int)blockIdx.y;
  const int n_layers = P.n_layers;
  const int n_steps = P.n_steps;

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < n_layers; ++i) tma_prefetch_desc(&P.maps[i]);
    for (int s = 0; s < kMaxSlots; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_full[1], 1);
    for (int i = 0; i < 8; ++i) mbar_init(&blk_full[i], 4 * kCtas);
    mbar_init(&drained[0], kEpiWarps * kCtas);
    mbar_init(&drained[1], kEpiWarps * kCtas);
    mbar_init(x0_ready, kEpiWarps * kCtas);
    *first_done = INT_MAX;
    *steps_done = 0;
    *peer_first_done = PAIR ? INT_MAX : 0;
    *peer_steps_done = PAIR ? 0 : INT_MAX;
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
    else { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t full_bar_l = (PAIR && !leader) ? mapa_u32(smem_u32(full_bar), 0) : smem_u32(full_bar);
  auto pair_first_done = [&]() { return max(ld_volatile_s32(first_done), ld_volatile_s32(peer_first_done)); };

  if (warp == 0) {
    // ======================================= TMA producer ========================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int step = 0; step < n_steps; ++step) {
        if (step >= 2) {   // at most one lockstep step ahead of the bookkeeping
          const long long t0 = clock64();
          while (min(ld_volatile_s32(steps_done), ld_volatile_s32(peer_steps_done)) < step - 1) {
            __nanosleep(64);
            if (clock64() - t0 > 8000000000LL) __trap();
          }
          if (pair_first_done() <= step - 2) break;
        }
        for (int li = 0; li < n_layers; ++li) {
          const RoLayer& Ly = P.layer[li];
          const int batch = Ly.per_member ? member : 0;
          for (int n0 = 0; n0 < Ly.n_pad; n0 += 256) {
            const int ncols = min(256, Ly.n_pad - n0);
            const int mycols = PAIR ? max(ncols / 2, 64) : ncols;
            const int c0 = n0 + (int)rank * mycols;
            for (int ks = 0; ks < Ly.k_blocks * (64 / kStageK); ++ks) {  // one stage = kStageK k x mycols n
              mbar_wait(&empty_bar[stage], phase ^ 1);
#ifdef FQLPOP_EPI_TRACE
              const bool st_tr = blockIdx.x <= 1 && blockIdx.y == 0 && step == 1 && (li == 1 || li == 2);
              const int st_idx = (li - 1) * 32 + (n0 / 256) * 16 + ks;
              long long* st_log = blockIdx.x == 0 ? g_ro16_stage : g_ro16_stage_peer;
              if (st_tr) st_log[4 * st_idx] = clock64();
#endif
              uint8_t* sb = stage_base + stage * kSlotBytes;
#ifdef FQLPOP_DEBUG_SWITCHES
              if (P.debug_flags & 2) {
                if (leader) mbar_arrive(&full_bar[stage]);
                if (++stage == kSlots) { stage = 0; phase ^= 1; }
                continue;
              }
#endif
              if (leader) mbar_arrive_expect_tx(&full_bar[stage], (mycols / 64) * kBoxB * kCtas);
              // one instruction per stage: the weights are tiled [K / kStageK][N / 64][kStageK][64] and the map's box is this CTA's
              // mycols / 64 adjacent boxes (Program::widen) -- one contiguous run of global memory
              const int row = (ks * (Ly.n_pad >> 6) + (c0 >> 6)) * kStageK;
              if (PAIR) tma_load_3d_2sm(sb, &P.maps[li], full_bar_l + 8u * (uint32_t)stage, 0, row, batch);
              else tma_load_3d(sb, &P.maps[li], &full_bar[stage], 0, row, batch);
#ifdef FQLPOP_EPI_TRACE
              if (st_tr) st_log[4 * st_idx + 1] = clock64();
#endif
              if (++stage == kSlots) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer ==========================================
    if (leader) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t lc = 0, tc = 0, xc = 0;   // layers / two-tile layers / X0-fed layers issued so far
      uint32_t blk_par = 0;              // bit kb: parity to wait for on blk_full[kb]
      const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
      for (int step = 0; step < n_steps; ++step) {
        if (step >= 2 && pair_first_done() <= step - 2) break;
        for (int li = 0; li < n_layers; ++li, ++lc) {
          const RoLayer& Ly = P.layer[li];
          const bool from_act = !Ly.a_from_x0;
#ifdef FQLPOP_EPI_TRACE
          long long w_dr = 0, w_a = 0, w_wt = 0;
          const long long t_begin = clock64();
#endif
          if (!from_act) {
            RT_T0();
            mbar_wait(x0_ready, xc & 1);
            ++xc;
            tc_fence_after();
            RT_ACC(w_a);
          }
          const uint32_t a_base = from_act ? act_addr : x0_addr;
          const bool two = Ly.n_pad > 256;
          int t = 0;
          for (int n0 = 0; n0 < Ly.n_pad; n0 += 256, ++t) {
            const int ncols = min(256, Ly.n_pad - n0);
            const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, PAIR ? (uint32_t)max(ncols, 128) : (uint32_t)ncols, 0, 1);
            {
              RT_T0();
              if (t == 0) {
                if (lc > 0) mbar_wait(&drained[0], (lc - 1) & 1);
              } else if (tc > 0) {
                mbar_wait(&drained[1], (tc - 1) & 1);
              }
              tc_fence_after();
              RT_ACC(w_dr);
            }
            for (int kb = 0; kb < Ly.k_blocks; ++kb) {
              if (from_act && t == 0) {
                RT_T0();
                mbar_wait(&blk_full[kb], (blk_par >> kb) & 1u);
                blk_par ^= 1u << kb;
                tc_fence_after();
                RT_ACC(w_a);
              }
              for (int h = 0; h < 64 / kStageK; ++h) {
                {
                  RT_T0();
                  mbar_wait(&full_bar[stage], phase);
                  tc_fence_after();
                  RT_ACC(w_wt);
                }
#ifdef FQLPOP_EPI_TRACE
                const bool st_tr = blockIdx.x == 0 && blockIdx.y == 0 && step == 1 && (li == 1 || li == 2) && lane == 0;
                const int st_idx = (li - 1) * 32 + t * 16 + kb * 2 + h;
                if (st_tr) g_ro16_stage[4 * st_idx + 2] = clock64();
#endif
#ifdef FQLPOP_DEBUG_SWITCHES
                if (P.debug_flags & 1) {   // free the slot (both CTAs) without reading it
                  if (lane == 0) {
                    mbar_arrive(&empty_bar[stage]);
                    if (PAIR) mbar_arrive_cluster(mapa_u32(smem_u32(&empty_bar[stage]), 1));
                    if (kb == Ly.k_blocks - 1 && h == 64 / kStageK - 1) {
                      if (PAIR) umma_commit_2sm(&tmem_full[t], 3);
                      else umma_commit(&tmem_full[t]);
                    }
                  }
                } else
#endif
                if (elect_one()) {
                  const uint32_t sa = a_base + kb * kActBlockBytes + h * 64;
                  const uint32_t sb = smem_u32(stage_base + stage * kSlotBytes);
                  const bool last = kb == Ly.k_blocks - 1 && h == 64 / kStageK - 1;
#pragma unroll
                  for (int k = 0; k < kStageK / 16; ++k) {
                    const uint64_t adesc = make_smem_desc_sw128(sa + k * 32, 16, 1024);
                    const uint64_t bdesc = make_smem_desc_sw128(sb + k * 2048, kBoxB, 1024);
                    const uint32_t acc = (kb | h | k) != 0 ? 1u : 0u;
                    if (PAIR) umma_bf16_2sm(tmem_base + n0, adesc, bdesc, idesc, acc);
                    else umma_bf16(tmem_base + n0, adesc, bdesc, idesc, acc);
                  }
                  if (PAIR) {
                    umma_commit_2sm(&empty_bar[stage], 3);
                    if (last) umma_commit_2sm(&tmem_full[t], 3);
                  } else {
                    umma_commit(&empty_bar[stage]);
                    if (last) umma_commit(&tmem_full[t]);
                  }
                }
                __syncwarp();
#ifdef FQLPOP_EPI_TRACE
                if (st_tr) g_ro16_stage[4 * st_idx + 3] = clock64();
#endif
                if (++stage == kSlots) { stage = 0; phase ^= 1; }
              }
            }
          }
          if (two) ++tc;
#ifdef FQLPOP_EPI_TRACE
          if (blockIdx.x == 0 && blockIdx.y == 0 && lane == 0 && lc < 256) {
            long long* o = g_ro16_trace + 8 * lc;
            o[0] = t_begin; o[1] = w_dr; o[2] = w_a; o[3] = w_wt; o[4] = clock64();
          }
#endif
        }
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ======================================= epilogue ============================================
    const int q = warp & 3;
    const int cg = (warp - kFirstEpiWarp) >> 2;
    const int r = q * 32 + lane;
    const int e = tile * 128 + r;   // episode id within the member
    const bool row_ok = e < P.n;
    REpi C;
    C.tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    C.tmem_full = tmem_full;
    C.remote = PAIR && !leader;
    C.blk_full_a = C.remote ? mapa_u32(smem_u32(blk_full), 0) : smem_u32(blk_full);
    C.drained_a = C.remote ? mapa_u32(smem_u32(drained), 0) : smem_u32(drained);
    C.x0_ready_a = C.remote ? mapa_u32(smem_u32(x0_ready), 0) : smem_u32(x0_ready);
    C.act = smem_u32(act); C.x0 = smem_u32(x0s);
    C.r = r; C.q = q; C.cg = cg; C.lane = lane;
    const int O = P.O, A = P.A, OA = P.O + P.A;
    const long long me = (long long)member * P.n + e;
    const unsigned long long seed = P.seeds ? P.seeds[member] : 0ull;
    const int k0 = 16 * cg;   // this thread's 16 columns of every 64-wide row quantity
    float2* ex = reinterpret_cast<float2*>(act);   // [4][128]: ACT is idle while the policy head's epilogue runs

    float o[16];  // this thread's share of the episode's observation (fp32, columns >= O are zero)
    {
      const float* src = P.init_obs + (long long)member * P.init_obs_ms + (long long)e * O;
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i] = (row_ok && k0 + i < O) ? __ldg(src + k0 + i) : 0.f;
    }
    bool done_once = false, success = false;   // (cg == 0 thread) episode bookkeeping
    int length = 0, steps_executed = 0;

    // Philox policy noise is drawn one lockstep step ahead, while this warp would otherwise wait for the tensor core, and
    // parked in global scratch (L2): thread (row, cg) draws action slots cg, cg + 4, ... of its row, one share per policy
    // layer. Step 0's is drawn inline.
    int pi_head = 0;
    while (pi_head < n_layers - 1 && P.layer[pi_head].kind != RO_PI_HEAD) ++pi_head;
    float* nz_next = P.noise_next + ((long long)member * 2 * P.n + e) * A;   // + (step & 1) * n * A
    auto draw_ahead = [&](int step, int li) {   // called on entry to policy layer li of lockstep step `step`
      if (P.noise || !row_ok || li < 1 || li > pi_head || step + 1 >= n_steps) return;
      float* dst = nz_next + (long long)((step + 1) & 1) * P.n * A;
      for (int j = cg + 4 * (li - 1); j < A; j += 4 * pi_head)
        __stcg(dst + j, philox_normal(seed, P.eval_counter + (unsigned long long)(step + 1), (uint32_t)e, (uint32_t)j));
    };
    // policy input [o, z_step] -> X0 (this thread's 16 columns)
    auto build_policy_input = [&](int step) {
      float x[16];
      const float* nz = P.noise ? P.noise + (long long)member * P.noise_ms + ((long long)step * P.n + e) * A : nullptr;
      const float* nzs = nz_next + (long long)(step & 1) * P.n * A;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int k = k0 + i;
        float v = o[i];
        if (k >= O && k < OA) {
          if (!row_ok) v = 0.f;
          else if (nz) v = __ldg(nz + (k - O));
          else if (step > 0) v = __ldcg(nzs + (k - O));
          else v = philox_normal(seed, P.eval_counter, (uint32_t)e, (uint32_t)(k - O));
        }
        x[i] = v;
      }
      st_x0_16(C.x0, r, cg, x);
      if (P.traj_obs && row_ok) {
        float* dst = P.traj_obs + (((long long)member * n_steps + step) * P.n + e) * O;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (k0 + i < O) dst[k0 + i] = o[i];
      }
    };
    build_policy_input(0);
    publish(C, C.x0_ready_a);

    // The CTA's done / progress flags of lockstep step s are published by ONE thread, off the critical path: the other
    // three bookkeeping warps only arrive on named barrier 9 after the termination layer; the first one completes the
    // barrier and stores the flags on entry to layer 1 of step s + 1 -- before its next mbarrier arrival, which every role's
    // check of the flags (two steps after they were set) is ordered after.
    int pending_step = -1;
    auto publish_flags = [&]() {
      asm volatile("bar.sync 9, 128;" ::: "memory");
      if (r == 0) {
        const int s1 = pending_step + 1;
        int fd = ld_volatile_s32(first_done);
        if (warp_all[0] && warp_all[1] && warp_all[2] && warp_all[3] && fd == INT_MAX) {
          fd = pending_step;
          st_volatile_s32(first_done, fd);
        }
        if (PAIR) {  // mirror both flags into the peer, first_done before steps_done
          const uint32_t peer = rank ^ 1u;
          asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(mapa_u32(smem_u32(peer_first_done), peer)), "r"(fd) : "memory");
          asm volatile("fence.acq_rel.cluster;" ::: "memory");
          asm volatile("st.shared::cluster.s32 [%0], %1;" ::"r"(mapa_u32(smem_u32(peer_steps_done), peer)), "r"(s1) : "memory");
          asm volatile("fence.acq_rel.cluster;" ::: "memory");
        }
        __threadfence_block();
        st_volatile_s32(steps_done, s1);
      }
      __syncwarp();
      pending_step = -1;
    };

    uint32_t lc = 0, tc = 0;
    for (int step = 0; step < n_steps; ++step) {
      if (step >= 2 && pair_first_done() <= step - 2) break;
      steps_executed = step + 1;
      for (int li = 0; li < n_layers; ++li, ++lc) {
        const RoLayer& Ly = P.layer[li];
        const uint32_t par0 = lc & 1, par1 = tc & 1;
        const float* bias = Ly.bias + (Ly.per_member ? (long long)member * Ly.bias_ms : 0);
#ifdef FQLPOP_EPI_TRACE
        long long* tr = (blockIdx.x == 0 && blockIdx.y == 0 && q == 0 && lane == 0 && (cg == 0 || cg == 3) && lc < 256)
                            ? g_ro16_trace + 2048 + 8 * lc + (cg ? 4 : 0) : nullptr;
        if (tr) tr[0] = clock64();
#endif
        if (li == 1 && pending_step >= 0) publish_flags();
        draw_ahead(step, li);
        if (Ly.kind == RO_HIDDEN) {
          if (Ly.act == ACT_GELU) hidden_layer<ACT_GELU>(C, bias, Ly.n_pad, par0, par1);
          else hidden_layer<ACT_RELU>(C, bias, Ly.n_pad, par0, par1);
          if (Ly.n_pad > 256) ++tc;
#ifdef FQLPOP_EPI_TRACE
          if (tr) tr[2] = clock64();
#endif
          continue;
        }
        // head layers: 64 accumulator columns, 16 per thread
        wait_full(C, 0, par0);
#ifdef FQLPOP_EPI_TRACE
        if (tr) tr[1] = clock64();
#endif
        float x[16];
        {
          uint32_t rr[16];
          tmem_ld16(C.tmem_row + k0, rr);
          float4 b4[4];
          ld_vec16(bias + k0, b4);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            x[4 * i + 0] = __uint_as_float(rr[4 * i + 0]) + b4[i].x;
            x[4 * i + 1] = __uint_as_float(rr[4 * i + 1]) + b4[i].y;
            x[4 * i + 2] = __uint_as_float(rr[4 * i + 2]) + b4[i].z;
            x[4 * i + 3] = __uint_as_float(rr[4 * i + 3]) + b4[i].w;
          }
        }
        arrive_drained(C, 0);
        if (Ly.kind == RO_PI_HEAD) {
          // the head image places action j in column O + j, i.e. where concat(o, a) wants it; then the state predictor's
          // input LayerNorm over the O + A valid columns (row sums cross the four column-group threads)
          float s = 0.f, ss = 0.f;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int k = k0 + i;
            x[i] = (k < O) ? o[i] : ((k < OA && row_ok) ? clip1(x[i]) : 0.f);
            if (k < OA) { s += x[i]; ss = fmaf(x[i], x[i], ss); }
          }
          if (P.traj_act && row_ok) {
            float* dst = P.traj_act + (((long long)member * n_steps + step) * P.n + e) * A;
#pragma unroll
            for (int i = 0; i < 16; ++i)
              if (k0 + i >= O && k0 + i < OA) dst[k0 + i - O] = x[i];
          }
          ex[cg * 128 + r] = make_float2(s, ss);
          asm volatile("bar.sync %0, 128;" ::"r"(q + 1) : "memory");
          float t1 = 0.f, t2 = 0.f;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 v = ex[j * 128 + r];
            t1 += v.x;
            t2 += v.y;
          }
          const float inv = 1.f / (float)OA;
          const float mean = t1 * inv;
          const float rstd = rsqrtf(fmaxf(t2 * inv - mean * mean, 0.f) + P.ln_eps_env);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int k = k0 + i;
            x[i] = (k < OA) ? fmaf((x[i] - mean) * rstd, __ldg(P.env_ln_scale + k), __ldg(P.env_ln_bias + k)) : 0.f;
          }
          st_x0_16(C.x0, r, cg, x);
          publish(C, C.x0_ready_a);
        } else if (Ly.kind == RO_STATE_OUT) {
#pragma unroll
          for (int i = 0; i < 16; ++i) o[i] = (k0 + i < O && row_ok) ? o[i] + x[i] : 0.f;
          st_x0_16(C.x0, r, cg, o);
          publish(C, C.x0_ready_a);
        } else {  // RO_TERM_OUT
          const int s1 = step + 1;
          if (cg == 0) {
            const float logit = x[0];
            if (P.traj_logit && row_ok) P.traj_logit[((long long)member * n_steps + step) * P.n + e] = logit;
            const bool term = logit > 0.f;
            const bool trunc = s1 >= P.max_episode_steps;
            if ((term || trunc) && !done_once && row_ok) {
              success = term;
              length = s1;
            }
            done_once = done_once || term || trunc;
            const int all = __all_sync(0xffffffffu, done_once || !row_ok);
            if (lane == 0) warp_all[q] = all;
            if (q != 0) {
              __threadfence_block();
              __syncwarp();
              asm volatile("bar.arrive 9, 128;" ::: "memory");
            } else {
              pending_step = step;   // publish_flags() on entry to layer 1 of the next step
            }
          }
#ifdef FQLPOP_EPI_TRACE
          if (tr) tr[3] = clock64();
#endif
          if (s1 < n_steps) {
            build_policy_input(s1);
            publish(C, C.x0_ready_a);
          }
        }
#ifdef FQLPOP_EPI_TRACE
        if (tr) tr[2] = clock64();
#endif
      }
    }

    if (row_ok) {
      float* dst = P.obs_work + me * O;
#pragma unroll
      for (int i = 0; i < 16; ++i)
        if (k0 + i < O) dst[k0 + i] = o[i];
    }
    if (cg == 0) {
      if (row_ok) {
        P.success[me] = success ? 1 : 0;
        P.length[me] = length;
      }
      float fs = (row_ok && success) ? 1.f : 0.f;
      float fr = row_ok ? (-(float)(length > 0 ? length : steps_executed) + (success ? 1.f : 0.f)) : 0.f;
#pragma unroll
      for (int s = 16; s >= 1; s >>= 1) {
        fs += __shfl_xor_sync(0xffffffffu, fs, s);
        fr += __shfl_xor_sync(0xffffffffu, fr, s);
      }
      if (lane == 0) {
        atomicAdd(P.success_sum + member, fs);
        atomicAdd(P.return_sum + member, fr);
        if (q == 0) atomicMax(P.steps_run + member, steps_executed);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_2sm(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

cudaError_t rollout16_init() {
  cudaError_t e = cudaFuncSetAttribute(fql_rollout16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesR);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(fql_rollout16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytesR);
}

// RO_MODE_ROLLOUT, LayerNorm-free policy, every layer 64 / 128 / 256 / 512 wide with hidden layers >= 128
bool rollout16_eligible(const RoParams& R) {
  if (R.mode != RO_MODE_ROLLOUT) return false;
  for (int i = 0; i < R.n_layers; ++i) {
    const RoLayer& L = R.layer[i];
    if (L.ln) return false;
    if (L.kind == RO_HIDDEN) {
      if (L.n_pad != 128 && L.n_pad != 256 && L.n_pad != 512) return false;
      if (L.act != ACT_GELU && L.act != ACT_RELU) return false;
    } else if (L.n_pad != 64) {
      return false;
    }
    if (L.a_from_x0 ? L.k_blocks != 1 : (L.k_blocks != 2 && L.k_blocks != 4 && L.k_blocks != 8)) return false;
  }
  return true;
}

cudaError_t launch_rollout16(const RoParams& R_in, int n_members, bool pair, cudaStream_t stream) {
  RoParams R = R_in;
#ifdef FQLPOP_DEBUG_SWITCHES
  if (const char* e = getenv("FQLPOP_RO_DEBUG")) R.debug_flags = atoi(e);
#endif
  if (!pair || R.tiles_per_member < 2) {
    fql_rollout16_kernel<false><<<dim3(R.tiles_per_member, n_members), kThreadsR, kSmemBytesR, stream>>>(R);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * ((R.tiles_per_member + 1) / 2), n_members);   // an odd last tile gets an all-invalid partner
  cfg.blockDim = dim3(kThreadsR);
  cfg.dynamicSmemBytes = kSmemBytesR;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, fql_rollout16_kernel<true>, R);
}

}  // namespace fqlpop

#ifdef FQLPOP_EPI_TRACE
extern "C" int fqlpop_debug_rollout_trace(long long* host_out, int n) {
  cudaDeviceSynchronize();
  return cudaMemcpyFromSymbol(host_out, fqlpop::g_ro16_trace, sizeof(long long) * (n < 4096 ? n : 4096)) == cudaSuccess ? 0 : -1;
}
extern "C" int fqlpop_debug_rollout_stage_trace(long long* host_out) {
  cudaDeviceSynchronize();
  return cudaMemcpyFromSymbol(host_out, fqlpop::g_ro16_stage, sizeof(long long) * 256) == cudaSuccess &&
         cudaMemcpyFromSymbol(host_out + 256, fqlpop::g_ro16_stage_peer, sizeof(long long) * 256) == cudaSuccess ? 0 : -1;
}
#endif
