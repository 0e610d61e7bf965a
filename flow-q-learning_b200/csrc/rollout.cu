// Fused persistent world-model rollout kernel (see rollout.h for the reference path it replaces).
//
// CTA = 6 warps: warp 0 = TMA producer (streams every layer's bf16 weights from L2, 16 KB stages),
// warp 1 = TMEM allocator + tcgen05.mma issuer, warps 2..5 = epilogue (one thread per episode; the thread keeps
// its episode's observation, done / success / length state in registers for the whole evaluation).
// Shared memory: ACT (128 rows x 512 bf16, the A operand of every hidden layer, K-major, 128-B swizzle),
// X0 (128 x 64, the A operand of each network's first layer), kStages weight stages, mbarriers.
// TMEM: 512 fp32 columns = one full 128 x 512 layer output.
//
// Layer pipeline: MMA warp waits `act_ready`, issues the layer's MMAs as weight stages land, commits
// `layer_done`; epilogue waits `layer_done`, turns the accumulator into the next A operand (bias, GELU / ReLU,
// optional LayerNorm, bf16) or into actions / next observations / termination logits, fences and arrives on
// `act_ready`. The producer runs up to one lockstep step ahead of the bookkeeping.
#include "rollout.h"

#include <climits>
#include <cstring>

#include "gemm.h"
#include "philox.cuh"
#include "pop.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kThreads = 320;  // TMA, MMA, 4 episode warps (own the rows' state), 4 helper warps (share the wide hidden epilogues)
constexpr int kStages = 5;
constexpr int kActBlockBytes = 128 * 64 * 2;  // one 64-wide k-block of the A operand = 16 KB
constexpr int kActBytes = 8 * kActBlockBytes;  // 128 KB
constexpr int kStageBytes = 32 * 256 * 2;      // (32 k) x (256 n) bf16 = 16 KB = four TMA boxes
constexpr int kBoxBytes = 64 * 32 * 2;         // 64 n x 32 k
constexpr int kBarBytes = 512;
constexpr int kSmemBytes = kActBytes + kActBlockBytes + kStages * kStageBytes + kBarBytes + 1024;
constexpr int kMaxSlots = 2 * kStages;  // pair mode: 8 KB weight slots (each CTA stages half of every tile's columns)

__device__ __forceinline__ float clip1(float x) { return fminf(fmaxf(x, -1.f), 1.f); }

// writes 8 consecutive bf16 (one 16-byte chunk `chunk` of row r) into a 128-B-swizzled K-major block
__device__ __forceinline__ void st_chunk(uint8_t* block, int r, int chunk, const float* v8) {
  uint4 u;
  u.x = pack_bf16x2(v8[0], v8[1]);
  u.y = pack_bf16x2(v8[2], v8[3]);
  u.z = pack_bf16x2(v8[4], v8[5]);
  u.w = pack_bf16x2(v8[6], v8[7]);
  *reinterpret_cast<uint4*>(block + r * 128 + ((chunk ^ (r & 7)) << 4)) = u;
}
// 64-wide row -> X0
__device__ __forceinline__ void st_row64(uint8_t* x0, int r, const float (&x)[64]) {
#pragma unroll
  for (int c = 0; c < 8; ++c) st_chunk(x0, r, c, &x[8 * c]);
}
// 32 columns starting at col0 (multiple of 32) -> ACT
__device__ __forceinline__ void st_act32(uint8_t* act, int r, int col0, const float (&v)[32]) {
  uint8_t* blk = act + (col0 >> 6) * kActBlockBytes;
  const int cbase = (col0 & 63) >> 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) st_chunk(blk, r, cbase + i, &v[8 * i]);
}

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ int ld_volatile_s32(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_s32(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

// PAIR: CTAs run as clusters of two (adjacent 128-episode tiles of the same member) and every product is ONE
// tcgen05.mma.cta_group::2 of M = 256 issued by the leader (cluster rank 0): each CTA stages only half of the N columns of
// every weight stage, which halves the bytes a CTA pulls through shared memory per lockstep step (1.93 MB -> 0.97 MB; the
// weight stream's stage round trip of ~3000 cycles is what bounds the single-CTA kernel). The pair leaves
// the step loop together: "every episode is done" is tracked per CTA and mirrored into the peer.
template <bool PAIR>
__global__ void __launch_bounds__(kThreads, 1) fql_rollout_kernel(const __grid_constant__ RoParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* x0s = smem + kActBytes;
  uint8_t* stage_base = x0s + kActBlockBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_base + kStages * kStageBytes);
  uint64_t* full_bar = bars;                // [kMaxSlots]
  uint64_t* empty_bar = bars + kMaxSlots;   // [kMaxSlots]
  uint64_t* act_ready = bars + 2 * kMaxSlots;
  uint64_t* layer_done = bars + 2 * kMaxSlots + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kMaxSlots + 2);
  int* first_done = reinterpret_cast<int*>(tmem_slot + 1);  // first step at which every episode of the CTA was done
  int* steps_done = first_done + 1;                         // lockstep steps fully book-kept
  int* peer_first_done = steps_done + 1;                    // (pair) the other CTA's two flags, written by it
  int* peer_steps_done = peer_first_done + 1;
  int* warp_all = peer_steps_done + 1;                      // [4]
  constexpr int kSlotBytes = PAIR ? kStageBytes / 2 : kStageBytes;
  constexpr int kSlots = PAIR ? kMaxSlots : kStages;
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
    mbar_init(act_ready, 8 * kCtas);   // one arrival per epilogue warp (4 episode + 4 helper warps per CTA)
    mbar_init(layer_done, 1);
    *first_done = INT_MAX;
    *steps_done = 0;
    *peer_first_done = PAIR ? INT_MAX : 0;   // unpaired: the "peer" is always done
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
  // the pair leaves at the first step both CTAs were done at; both flags of the peer are settled once its bookkeeping of
  // that step is (the producer checks that explicitly, everybody else is ordered behind it by the MMA barriers)
  auto pair_first_done = [&]() { return max(ld_volatile_s32(first_done), ld_volatile_s32(peer_first_done)); };
  // barriers both CTAs signal live in the leader
  const uint32_t act_ready_l = (PAIR && !leader) ? mapa_u32(smem_u32(act_ready), 0) : smem_u32(act_ready);
  const uint32_t full_bar_l = (PAIR && !leader) ? mapa_u32(smem_u32(full_bar), 0) : smem_u32(full_bar);
  // this warp's rows of the next A operand are written (generic proxy -> async proxy fence per lane), its accumulator
  // columns read: one arrival per warp
  auto arrive_ready = [&]() {
    tc_fence_before();
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      if (PAIR && !leader) mbar_arrive_cluster(act_ready_l);
      else mbar_arrive(act_ready);
    }
  };

  if (warp == 0) {
    // ======================================= TMA producer ========================================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int step = 0; step < n_steps; ++step) {
        if (step >= 2) {
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
            // columns of this MMA tile staged by THIS CTA (a 64-column layer is issued as N = 128 in pair mode: the
            // second CTA's half lies outside the weight tensor and is zero-filled)
            const int mycols = PAIR ? max(ncols / 2, 64) : ncols;
            const int c0 = n0 + (int)rank * mycols;
            for (int ks = 0; ks < Ly.k_blocks * 2; ++ks) {  // one stage = 32 k x ncols n
              mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sb = stage_base + stage * kSlotBytes;
              if (leader) mbar_arrive_expect_tx(&full_bar[stage], (mycols / 64) * kBoxBytes * kCtas);
              for (int b = 0; b < mycols / 64; ++b) {
                if (PAIR) tma_load_3d_2sm(sb + b * kBoxBytes, &P.maps[li], full_bar_l + 8u * (uint32_t)stage, c0 + 64 * b, ks * 32, batch);
                else tma_load_3d(sb + b * kBoxBytes, &P.maps[li], &full_bar[stage], c0 + 64 * b, ks * 32, batch);
              }
              if (++stage == kSlots) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer ==========================================
    if (lane == 0 && leader) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t it = 0;
      const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
      for (int step = 0; step < n_steps; ++step) {
        if (step >= 2 && pair_first_done() <= step - 2) break;
        for (int li = 0; li < n_layers; ++li, ++it) {
          const RoLayer& Ly = P.layer[li];
          mbar_wait(act_ready, it & 1);
          tc_fence_after();
          const uint32_t a_base = Ly.a_from_x0 ? x0_addr : act_addr;
          // 128 x 256 x 16 MMAs (256 x 256 x 16 across the pair): a stage holds 32 k x up to 256 n as 64n x 32k boxes
          for (int n0 = 0; n0 < Ly.n_pad; n0 += 256) {
            const int ncols = min(256, Ly.n_pad - n0);
            const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, PAIR ? (uint32_t)max(ncols, 128) : (uint32_t)ncols, 0, 1);
            for (int ks = 0; ks < Ly.k_blocks * 2; ++ks) {
              mbar_wait(&full_bar[stage], phase);
              tc_fence_after();
              const uint32_t sa = a_base + (ks >> 1) * kActBlockBytes + (ks & 1) * 64;
              const uint32_t sb = smem_u32(stage_base + stage * kSlotBytes);
#pragma unroll
              for (int k = 0; k < 2; ++k) {
                const uint64_t adesc = make_smem_desc_sw128(sa + k * 32, 16, 1024);
                const uint64_t bdesc = make_smem_desc_sw128(sb + k * 2048, kBoxBytes, 1024);
                if (PAIR) umma_bf16_2sm(tmem_base + n0, adesc, bdesc, idesc, (ks | k) != 0 ? 1u : 0u);
                else umma_bf16(tmem_base + n0, adesc, bdesc, idesc, (ks | k) != 0 ? 1u : 0u);
              }
              if (PAIR) umma_commit_2sm(&empty_bar[stage], 3);
              else umma_commit(&empty_bar[stage]);
              if (++stage == kSlots) { stage = 0; phase ^= 1; }
            }
          }
          if (PAIR) umma_commit_2sm(layer_done, 3);
          else umma_commit(layer_done);
        }
      }
    }
  } else if (warp >= 6) {
    // ======================================= helper epilogue =====================================
    // Same TMEM lane quarter as episode warp (warp - 4); takes the odd 32-column chunks of the bias + activation
    // epilogue of the plain hidden layers, so a row's 512 columns are split over two threads. Holds no episode state.
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t it = 0;
    arrive_ready();
    for (int step = 0; step < n_steps; ++step) {
      if (step >= 2 && pair_first_done() <= step - 2) break;
      for (int li = 0; li < n_layers; ++li, ++it) {
        const RoLayer& Ly = P.layer[li];
        mbar_wait(layer_done, it & 1);
        tc_fence_after();
        if (Ly.kind == RO_HIDDEN && !Ly.ln) {
          const float* bias = Ly.bias + (Ly.per_member ? (long long)member * Ly.bias_ms : 0);
          const int nchunks = Ly.n_pad / 32;
          for (int c = 1; c < nchunks; c += 2) {
            uint32_t rr[32];
            tmem_ld32(tmem_row + c * 32, rr);
            tmem_ld_wait();
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float z = __uint_as_float(rr[j]) + __ldg(bias + c * 32 + j);
              if (Ly.act == ACT_GELU) {
                float gp;
                gelu_fwd(z, v[j], gp);
              } else {
                v[j] = fmaxf(z, 0.f);
              }
            }
            st_act32(act, r, c * 32, v);
          }
        }
        arrive_ready();
      }
    }
  } else if (warp >= 2) {
    // ======================================= epilogue ============================================
    const int q = warp & 3;
    const int r = q * 32 + lane;  // row of the tile == TMEM lane
    const int e = tile * 128 + r;  // episode id within the member
    const bool row_ok = e < P.n;
    const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    const int O = P.O, A = P.A, OA = P.O + P.A;
    const long long me = (long long)member * P.n + e;
    const unsigned long long seed = P.seeds ? P.seeds[member] : 0ull;

    float o[64];  // this episode's observation (fp32, cols >= O are zero)
    {
      const float* src = P.init_obs + (long long)member * P.init_obs_ms + (long long)e * O;
#pragma unroll
      for (int k = 0; k < 64; ++k) o[k] = (row_ok && k < O) ? __ldg(src + k) : 0.f;
    }
    bool done_once = false, success = false;
    int length = 0;
    uint32_t it = 0;
    int steps_executed = 0;

    // policy input [o, z_step] -> X0
    auto build_policy_input = [&](int step) {
      float x[64];
      const float* nz = P.noise ? P.noise + (long long)member * P.noise_ms + ((long long)step * P.n + e) * A : nullptr;
#pragma unroll
      for (int k = 0; k < 64; ++k) {
        float v = o[k];
        if (k >= O && k < OA) {
          if (!row_ok) v = 0.f;
          else if (nz) v = __ldg(nz + (k - O));
          else v = philox_normal(seed, P.eval_counter + (unsigned long long)step, (uint32_t)e, (uint32_t)(k - O));
        }
        x[k] = v;
      }
      st_row64(x0s, r, x);
      if (P.traj_obs && row_ok) {
        float* dst = P.traj_obs + (((long long)member * n_steps + step) * P.n + e) * O;
#pragma unroll
        for (int k = 0; k < 64; ++k)
          if (k < O) dst[k] = o[k];
      }
    };
    // state-predictor input LN(concat(o, a)) -> X0; a occupies x[O .. O+A)
    auto build_state_input = [&](float (&x)[64]) {
      float s = 0.f, ss = 0.f;
#pragma unroll
      for (int k = 0; k < 64; ++k)
        if (k < OA) { s += x[k]; ss = fmaf(x[k], x[k], ss); }
      const float inv = 1.f / (float)OA;
      const float mean = s * inv;
      const float var = fmaxf(ss * inv - mean * mean, 0.f);
      const float rstd = rsqrtf(var + P.ln_eps_env);
#pragma unroll
      for (int k = 0; k < 64; ++k)
        x[k] = (k < OA) ? fmaf((x[k] - mean) * rstd, __ldg(P.env_ln_scale + k), __ldg(P.env_ln_bias + k)) : 0.f;
      st_row64(x0s, r, x);
    };

    if (P.mode == RO_MODE_ENV) {
      float x[64];
#pragma unroll
      for (int k = 0; k < 64; ++k) {
        float v = o[k];
        if (k >= O && k < OA) v = row_ok ? clip1(__ldg(P.actions_in + (long long)e * A + (k - O))) : 0.f;
        x[k] = v;
      }
      build_state_input(x);
    } else {
      build_policy_input(0);
    }
    arrive_ready();

    for (int step = 0; step < n_steps; ++step) {
      if (step >= 2 && pair_first_done() <= step - 2) break;
      steps_executed = step + 1;
      for (int li = 0; li < n_layers; ++li, ++it) {
        const RoLayer& Ly = P.layer[li];
        mbar_wait(layer_done, it & 1);
        tc_fence_after();
        const float* bias = Ly.bias + (Ly.per_member ? (long long)member * Ly.bias_ms : 0);
        if (Ly.kind == RO_HIDDEN) {
          const int nchunks = Ly.n_pad / 32;
          if (!Ly.ln) {
            for (int c = 0; c < nchunks; c += 2) {  // odd chunks: helper warps
              uint32_t rr[32];
              tmem_ld32(tmem_row + c * 32, rr);
              tmem_ld_wait();
              float v[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float z = __uint_as_float(rr[j]) + __ldg(bias + c * 32 + j);
                if (Ly.act == ACT_GELU) {
                  float gp;
                  gelu_fwd(z, v[j], gp);
                } else {
                  v[j] = fmaxf(z, 0.f);
                }
              }
              st_act32(act, r, c * 32, v);
            }
          } else {
            // activation, then LayerNorm over the n_true-wide row (statistics are thread-local)
            const float* gamma = Ly.gamma + (Ly.per_member ? (long long)member * Ly.bias_ms : 0);
            const float* beta = Ly.beta + (Ly.per_member ? (long long)member * Ly.bias_ms : 0);
            float s = 0.f, ss = 0.f;
            for (int c = 0; c < nchunks; ++c) {
              uint32_t rr[32];
              tmem_ld32(tmem_row + c * 32, rr);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float z = __uint_as_float(rr[j]) + __ldg(bias + c * 32 + j);
                float g, gp;
                if (Ly.act == ACT_GELU) gelu_fwd(z, g, gp);
                else g = fmaxf(z, 0.f);
                s += g;
                ss = fmaf(g, g, ss);
                rr[j] = __float_as_uint(g);
              }
              tmem_st32(tmem_row + c * 32, rr);
            }
            tmem_st_wait();
            const float inv = 1.f / (float)Ly.n_true;
            const float mean = s * inv;
            const float rstd = rsqrtf(fmaxf(ss * inv - mean * mean, 0.f) + P.ln_eps_pi);
            for (int c = 0; c < nchunks; ++c) {
              uint32_t rr[32];
              tmem_ld32(tmem_row + c * 32, rr);
              tmem_ld_wait();
              float v[32];
#pragma unroll
              for (int j = 0; j < 32; ++j)
                v[j] = fmaf((__uint_as_float(rr[j]) - mean) * rstd, __ldg(gamma + c * 32 + j), __ldg(beta + c * 32 + j));
              st_act32(act, r, c * 32, v);
            }
          }
        } else {
          // head layers: 64 accumulator columns
          float x[64];
          {
            uint32_t r0[32], r1[32];
            tmem_ld32(tmem_row, r0);
            tmem_ld32(tmem_row + 32, r1);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              x[j] = __uint_as_float(r0[j]) + __ldg(bias + j);
              x[32 + j] = __uint_as_float(r1[j]) + __ldg(bias + 32 + j);
            }
          }
          if (Ly.kind == RO_PI_HEAD) {
            // the head image places action j in column O + j, i.e. where concat(o, a) wants it
#pragma unroll
            for (int k = 0; k < 64; ++k) x[k] = (k < O) ? o[k] : ((k < OA && row_ok) ? clip1(x[k]) : 0.f);
            if (P.mode == RO_MODE_POLICY) {
              if (row_ok) {
#pragma unroll
                for (int k = 0; k < 64; ++k)
                  if (k >= O && k < OA) P.actions_out[(long long)e * A + (k - O)] = x[k];
              }
            } else {
              if (P.traj_act && row_ok) {
                float* dst = P.traj_act + (((long long)member * n_steps + step) * P.n + e) * A;
#pragma unroll
                for (int k = 0; k < 64; ++k)
                  if (k >= O && k < OA) dst[k - O] = x[k];
              }
              build_state_input(x);
            }
          } else if (Ly.kind == RO_STATE_OUT) {
#pragma unroll
            for (int k = 0; k < 64; ++k) o[k] = (k < O && row_ok) ? o[k] + x[k] : 0.f;
            st_row64(x0s, r, o);
            if (P.mode == RO_MODE_ENV && row_ok) {
#pragma unroll
              for (int k = 0; k < 64; ++k)
                if (k < O) P.next_obs_out[(long long)e * O + k] = o[k];
            }
          } else {  // RO_TERM_OUT
            const float logit = x[0];
            if (P.mode == RO_MODE_ENV) {
              if (row_ok) P.logits_out[e] = logit;
            } else {
              if (P.traj_logit && row_ok) P.traj_logit[((long long)member * n_steps + step) * P.n + e] = logit;
              const bool term = logit > 0.f;
              const int s1 = step + 1;
              const bool trunc = s1 >= P.max_episode_steps;
              if ((term || trunc) && !done_once && row_ok) {
                success = term;
                length = s1;
              }
              done_once = done_once || term || trunc;
              const int all = __all_sync(0xffffffffu, done_once || !row_ok);
              if (lane == 0) warp_all[q] = all;
              epi_bar_sync();
              if (r == 0) {
                int fd = ld_volatile_s32(first_done);
                if (warp_all[0] && warp_all[1] && warp_all[2] && warp_all[3] && fd == INT_MAX) {
                  fd = step;
                  st_volatile_s32(first_done, step);
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
              epi_bar_sync();  // the flags are published before any warp of this CTA arrives for the next step
              if (s1 < n_steps) build_policy_input(s1);
            }
          }
        }
        arrive_ready();
      }
    }

    if (P.mode == RO_MODE_ROLLOUT) {
      if (row_ok) {
        P.success[me] = success ? 1 : 0;
        P.length[me] = length;
        float* dst = P.obs_work + me * O;
#pragma unroll
        for (int k = 0; k < 64; ++k)
          if (k < O) dst[k] = o[k];
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

// shifted one-step-policy head image: column O + j of the image = column j of the flax kernel, so the head's
// accumulator lines up with concat(obs, action). w: [P][H][64] bf16, b: [P][64] fp32.
__global__ void build_head_image_kernel(const float* __restrict__ theta, long long S, long long offW, long long offb,
                                        int H, int O, int A, __nv_bfloat16* __restrict__ w, float* __restrict__ b) {
  const int p = blockIdx.y;
  const float* th = theta + (long long)p * S;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < H * 64; i += gridDim.x * blockDim.x) {
    const int k = i >> 6, c = i & 63;
    const int j = c - O;
    w[((long long)p * H + k) * 64 + c] = __float2bfloat16_rn((j >= 0 && j < A) ? th[offW + (long long)k * kHeadPad + j] : 0.f);
  }
  if (blockIdx.x == 0 && threadIdx.x < 64) {
    const int j = (int)threadIdx.x - O;
    b[p * 64 + threadIdx.x] = (j >= 0 && j < A) ? th[offb + j] : 0.f;
  }
}

// hidden-layer kernel [K][H] (bf16 shadow) -> tiled [K / ks][H / 64][ks][64], ks = 32 << ks_shift; grid.y = member
__global__ void tile_policy_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long S,
                                   long long offW, int K, int H, const int* member_list, int ks_shift) {
  const int p = member_list ? member_list[blockIdx.y] : (int)blockIdx.y;
  const uint4* s = reinterpret_cast<const uint4*>(src + (long long)p * S + offW);
  uint4* d = reinterpret_cast<uint4*>(dst + (long long)p * S + offW);
  const int NB = H / 64, per_row = H / 8;   // 16-byte chunks
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K * per_row; i += gridDim.x * blockDim.x) {
    const int k = i / per_row, c = i % per_row;   // chunk c = columns [8 c, 8 c + 8)
    const int sh = 5 + ks_shift, ks = 1 << sh;
    d[((long long)((k >> sh) * NB + (c >> 3)) * ks + (k & (ks - 1))) * 8 + (c & 7)] = s[i];
  }
}

__global__ void finish_rollout_kernel(const int* member_list, int n_active, int n, const float* success_sum,
                                      const float* return_sum, float* success_rate, float* return_mean) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_active) return;
  const int p = member_list ? member_list[i] : i;
  if (success_rate) success_rate[p] = success_sum[p] / (float)n;
  if (return_mean) return_mean[p] = return_sum[p] / (float)n;
}

}  // namespace

cudaError_t rollout_init() {
  cudaError_t e = cudaFuncSetAttribute(fql_rollout_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fql_rollout_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  return rollout16_init();
}

// FQLPOP_ROLLOUT_PAIR: clusters of two CTAs on adjacent episode tiles (tcgen05.mma.cta_group::2). Unset: on for the
// second-generation kernel (348 vs 327 M env-steps/s at 16 x 4096 x 200), off for this file's (165 vs 170).
static bool rollout_pair_mode(bool gen2) {
  static const int v = [] { const char* e = getenv("FQLPOP_ROLLOUT_PAIR"); return e ? (atoi(e) != 0 ? 1 : 0) : -1; }();
  return v < 0 ? gen2 : v != 0;
}
// FQLPOP_ROLLOUT_GEN=1 keeps whole rollouts on this file's kernel; default: rollout16.cu where it applies
static int rollout_gen() {
  static const int gen = [] { const char* e = getenv("FQLPOP_ROLLOUT_GEN"); return e ? atoi(e) : 2; }();
  return gen;
}
static cudaError_t launch_rollout(const RoParams& R, int n_members, cudaStream_t stream) {
  const bool pair_on = rollout_pair_mode(false);
  bool pair_ok = pair_on && R.tiles_per_member >= 2;
  for (int i = 0; i < R.n_layers && pair_ok; ++i) {   // every MMA tile must split into two halves of whole 64-column boxes
    const int rem = R.layer[i].n_pad % 256;
    pair_ok = rem == 0 || rem == 64 || rem == 128;
  }
  if (!pair_ok) {
    fql_rollout_kernel<false><<<dim3(R.tiles_per_member, n_members), kThreads, kSmemBytes, stream>>>(R);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * ((R.tiles_per_member + 1) / 2), n_members);   // an odd last tile gets an all-invalid partner
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = kSmemBytes;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, fql_rollout_kernel<true>, R);
}

#define CK(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess) {                                                                      \
      if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
      return 1;                                                                                   \
    }                                                                                             \
  } while (0)

static int pad64(int x) { return (x + 63) / 64 * 64; }

// ------------------------------------------------------------------------------------------------------
// world model
// ------------------------------------------------------------------------------------------------------
static int upload_layer(EnvModel* m, EnvModel::Layer& L, int K, int N, const float* kernel, const float* bias,
                        std::string* err) {
  L.K = K; L.N = N; L.K_pad = pad64(K); L.N_pad = pad64(N);
  std::vector<__nv_bfloat16> w((size_t)L.K_pad * L.N_pad, __float2bfloat16(0.f));
  std::vector<float> b((size_t)L.N_pad, 0.f);
  for (int k = 0; k < K; ++k)
    for (int n = 0; n < N; ++n) w[(size_t)k * L.N_pad + n] = __float2bfloat16(kernel[(size_t)k * N + n]);
  for (int n = 0; n < N; ++n) b[n] = bias[n];
  std::vector<__nv_bfloat16> wt(w.size()), wt64(w.size());
  const int NB = L.N_pad / 64;
  for (int k = 0; k < L.K_pad; ++k)
    for (int n = 0; n < L.N_pad; ++n) {
      wt[((size_t)((k / 32) * NB + n / 64) * 32 + k % 32) * 64 + n % 64] = w[(size_t)k * L.N_pad + n];
      wt64[((size_t)((k / 64) * NB + n / 64) * 64 + k % 64) * 64 + n % 64] = w[(size_t)k * L.N_pad + n];
    }
  void *dw = nullptr, *db = nullptr, *dwt = nullptr, *dwt64 = nullptr;
  CK(cudaMalloc(&dw, w.size() * sizeof(__nv_bfloat16)));
  m->allocs.push_back(dw);
  CK(cudaMalloc(&dwt, wt.size() * sizeof(__nv_bfloat16)));
  m->allocs.push_back(dwt);
  CK(cudaMemcpy(dwt, wt.data(), wt.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  L.wt = (__nv_bfloat16*)dwt;
  CK(cudaMalloc(&dwt64, wt64.size() * sizeof(__nv_bfloat16)));
  m->allocs.push_back(dwt64);
  CK(cudaMemcpy(dwt64, wt64.data(), wt64.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  L.wt64 = (__nv_bfloat16*)dwt64;
  CK(cudaMalloc(&db, b.size() * sizeof(float)));
  m->allocs.push_back(db);
  CK(cudaMemcpy(dw, w.data(), w.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, b.data(), b.size() * sizeof(float), cudaMemcpyHostToDevice));
  L.w = (__nv_bfloat16*)dw;
  L.b = (float*)db;
  return 0;
}

int envmodel_create_impl(int O, int A, int n_state_hidden, const int* state_hidden, int n_term_hidden,
                         const int* term_hidden, float ln_eps, const float* const* state_kernels,
                         const float* const* state_biases, const float* ln_scale, const float* ln_bias,
                         const float* const* term_kernels, const float* const* term_biases, EnvModel** out,
                         std::string* err) {
  auto bad = [&](const char* msg) {
    if (err) *err = msg;
    return 1;
  };
  if (O < 1 || A < 1 || O + A > 64) return bad("world model: need ob_dim + action_dim <= 64");
  if (n_state_hidden < 1 || n_state_hidden > kRoMaxHidden || n_term_hidden < 1 || n_term_hidden > kRoMaxHidden)
    return bad("world model: between 1 and 6 hidden layers per network");
  for (int i = 0; i < n_state_hidden; ++i)
    if (state_hidden[i] < 1 || state_hidden[i] > 512) return bad("world model: hidden widths must be in [1, 512]");
  for (int i = 0; i < n_term_hidden; ++i)
    if (term_hidden[i] < 1 || term_hidden[i] > 512) return bad("world model: hidden widths must be in [1, 512]");
  if (cudaError_t e = rollout_init(); e != cudaSuccess) return bad(cudaGetErrorString(e));
  EnvModel* m = new EnvModel;
  m->O = O; m->A = A; m->n_state_hidden = n_state_hidden; m->n_term_hidden = n_term_hidden; m->ln_eps = ln_eps;
  int rc = 0;
  int k = O + A;
  m->state.resize(n_state_hidden + 1);
  for (int i = 0; i <= n_state_hidden && !rc; ++i) {
    const int n = i < n_state_hidden ? state_hidden[i] : O;
    if (i < n_state_hidden) m->state_hidden[i] = n;
    rc = upload_layer(m, m->state[i], k, n, state_kernels[i], state_biases[i], err);
    k = n;
  }
  k = O;
  m->term.resize(n_term_hidden + 1);
  for (int i = 0; i <= n_term_hidden && !rc; ++i) {
    const int n = i < n_term_hidden ? term_hidden[i] : 1;
    if (i < n_term_hidden) m->term_hidden[i] = n;
    rc = upload_layer(m, m->term[i], k, n, term_kernels[i], term_biases[i], err);
    k = n;
  }
  if (!rc) {
    std::vector<float> sc(64, 0.f), bi(64, 0.f);
    for (int i = 0; i < O + A; ++i) { sc[i] = ln_scale[i]; bi[i] = ln_bias[i]; }
    void *ds = nullptr, *db = nullptr;
    if (cudaMalloc(&ds, 64 * sizeof(float)) != cudaSuccess || cudaMalloc(&db, 64 * sizeof(float)) != cudaSuccess) {
      rc = bad("cudaMalloc failed");
    } else {
      m->allocs.push_back(ds);
      m->allocs.push_back(db);
      cudaMemcpy(ds, sc.data(), 64 * sizeof(float), cudaMemcpyHostToDevice);
      cudaMemcpy(db, bi.data(), 64 * sizeof(float), cudaMemcpyHostToDevice);
      m->ln_scale = (float*)ds;
      m->ln_bias = (float*)db;
    }
  }
  if (rc) {
    envmodel_destroy_impl(m);
    return 1;
  }
  *out = m;
  return 0;
}

void envmodel_destroy_impl(EnvModel* m) {
  if (!m) return;
  cudaDeviceSynchronize();
  for (void* p : m->allocs) cudaFree(p);
  delete m;
}

// ------------------------------------------------------------------------------------------------------
// layer programs
// ------------------------------------------------------------------------------------------------------
namespace {

struct Program {
  RoParams* R;
  std::string* err;
  bool ok = true;
  struct MapArgs { const void* w; const void* wt; const void* wt64; int K_pad, N_pad, batch; long long ms; };
  MapArgs margs[kRoMaxLayers];
  // rollout16.cu fetches a whole weight stage (the 64-column boxes of one CTA's share of an accumulator tile x 32 k) with
  // one TMA instruction from the TILED copy of the weights, where that stage is one contiguous run: re-encode every map
  bool widen(bool pair) {
    for (int i = 0; i < R->n_layers; ++i) {
      const MapArgs& a = margs[i];
      const int ncols = a.N_pad < 256 ? a.N_pad : 256;
      const int mycols = pair ? (ncols / 2 > 64 ? ncols / 2 : 64) : ncols;
      const void* img = pair ? a.wt64 : a.wt;   // pair mode: 64-k weight stages
      if (!img || encode_tmap_bf16_tiled(&R->maps[i], img, a.K_pad, a.N_pad, a.batch, a.ms, mycols / 64, pair ? 64 : 32) != 0) {
        if (err) *err = "rollout: cuTensorMapEncodeTiled (tiled weights) failed";
        return ok = false;
      }
    }
    return true;
  }
  int add(const void* w, int K_pad, int N_pad, int n_true, int batch, long long ms, int kind, int act, int ln,
          int a_from_x0, const float* bias, long long bias_ms, int per_member, const void* wt = nullptr,
          const void* wt64 = nullptr) {
    if (R->n_layers >= kRoMaxLayers) {
      ok = false;
      if (err) *err = "rollout: too many layers";
      return -1;
    }
    const int i = R->n_layers++;
    margs[i] = MapArgs{w, wt, wt64 ? wt64 : wt, K_pad, N_pad, batch, ms};
    if (encode_tmap_bf16(&R->maps[i], w, N_pad, K_pad, batch, N_pad, ms, 32) != 0) {
      ok = false;
      if (err) *err = "rollout: cuTensorMapEncodeTiled failed";
    }
    RoLayer& L = R->layer[i];
    L.kind = kind; L.act = act; L.ln = ln; L.a_from_x0 = a_from_x0;
    L.k_blocks = K_pad / 64; L.n_pad = N_pad; L.n_true = n_true; L.per_member = per_member;
    L.bias = bias; L.bias_ms = bias_ms; L.gamma = nullptr; L.beta = nullptr;
    return i;
  }
};

int ensure_scratch(Population* P, int n, std::string* err) {
  RolloutScratch& S = P->ro;
  const int Pn = P->c.P;
  if (!S.head_w) {
    void* p = nullptr;
    CK(cudaMalloc(&p, (size_t)Pn * P->pi.L.H * 64 * sizeof(__nv_bfloat16))); P->allocs.push_back(p); S.head_w = (__nv_bfloat16*)p;
    CK(cudaMalloc(&p, (size_t)Pn * 64 * sizeof(float))); P->allocs.push_back(p); S.head_b = (float*)p;
    CK(cudaMalloc(&p, (size_t)Pn * P->pi.L.S * sizeof(__nv_bfloat16))); P->allocs.push_back(p); S.pi_tiled = (__nv_bfloat16*)p;
    CK(cudaMalloc(&p, (size_t)Pn * sizeof(float))); P->allocs.push_back(p); S.success_sum = (float*)p;
    CK(cudaMalloc(&p, (size_t)Pn * sizeof(float))); P->allocs.push_back(p); S.return_sum = (float*)p;
    CK(cudaMalloc(&p, (size_t)Pn * sizeof(int))); P->allocs.push_back(p); S.steps_run = (int*)p;
    CK(cudaMalloc(&p, (size_t)Pn * sizeof(int))); P->allocs.push_back(p); S.identity = (int*)p;
    std::vector<int> id(Pn);
    for (int i = 0; i < Pn; ++i) id[i] = i;
    CK(cudaMemcpy(S.identity, id.data(), sizeof(int) * Pn, cudaMemcpyHostToDevice));
  }
  if (S.ep_cap < n) {
    void* p = nullptr;
    CK(cudaMalloc(&p, (size_t)Pn * n * P->c.O * sizeof(float))); P->allocs.push_back(p); S.obs_work = (float*)p;
    CK(cudaMalloc(&p, (size_t)Pn * n)); P->allocs.push_back(p); S.success = (unsigned char*)p;
    CK(cudaMalloc(&p, (size_t)Pn * n * sizeof(int))); P->allocs.push_back(p); S.length = (int*)p;
    CK(cudaMalloc(&p, (size_t)Pn * 2 * n * P->c.A * sizeof(float))); P->allocs.push_back(p); S.noise_next = (float*)p;
    S.ep_cap = n;
  }
  return 0;
}

void add_policy(Program& pg, Population* P) {
  const Arena& ar = P->pi;
  const NetLayout& N = ar.L;
  for (int l = 0; l < N.n_hidden; ++l) {
    const int i = pg.add(ar.shadow + N.offW[l], N.K(l), N.H, N.H, ar.V, N.S, RO_HIDDEN, ACT_GELU, N.ln, l == 0,
                         ar.theta + N.offb[l], N.S, 1, P->ro.pi_tiled ? P->ro.pi_tiled + N.offW[l] : nullptr);
    if (i >= 0 && N.ln) {
      pg.R->layer[i].gamma = ar.theta + N.offg[l];
      pg.R->layer[i].beta = ar.theta + N.offbeta[l];
    }
  }
  // (a 64-column kernel is its own tiled image)
  pg.add(P->ro.head_w, N.H, 64, N.out_dim, ar.V, (long long)N.H * 64, RO_PI_HEAD, ACT_NONE, 0, 0, P->ro.head_b, 64, 1,
         P->ro.head_w, P->ro.head_w);
}
void add_env(Program& pg, const EnvModel* m) {
  for (size_t i = 0; i < m->state.size(); ++i) {
    const EnvModel::Layer& L = m->state[i];
    const bool last = i + 1 == m->state.size();
    pg.add(L.w, L.K_pad, L.N_pad, L.N, 1, 0, last ? RO_STATE_OUT : RO_HIDDEN, ACT_RELU, 0, i == 0, L.b, 0, 0, L.wt, L.wt64);
  }
  for (size_t i = 0; i < m->term.size(); ++i) {
    const EnvModel::Layer& L = m->term[i];
    const bool last = i + 1 == m->term.size();
    pg.add(L.w, L.K_pad, L.N_pad, L.N, 1, 0, last ? RO_TERM_OUT : RO_HIDDEN, ACT_RELU, 0, i == 0, L.b, 0, 0, L.wt, L.wt64);
  }
}
void launch_head_image(Population* P, cudaStream_t s) {
  const NetLayout& N = P->pi.L;
  build_head_image_kernel<<<dim3(32, P->c.P), 256, 0, s>>>(P->pi.theta, N.S, N.offW[N.n_hidden], N.offb[N.n_hidden], N.H,
                                                          P->c.O, P->c.A, P->ro.head_w, P->ro.head_b);
}

}  // namespace

int rollout_impl(Population* P, EnvModel* m, const RolloutArgs& a, cudaStream_t stream, std::string* err) {
  auto bad = [&](const char* msg) {
    if (err) *err = msg;
    return 1;
  };
  if (!P || !m) return bad("rollout: null handle");
  if (m->O != P->c.O || m->A != P->c.A) return bad("rollout: world model and population disagree on ob_dim / action_dim");
  if (a.n_episodes < 1 || a.max_episode_steps < 1 || !a.init_obs) return bad("rollout: bad arguments");
  if (cudaError_t e = rollout_init(); e != cudaSuccess) return bad(cudaGetErrorString(e));
  if (ensure_scratch(P, a.n_episodes, err)) return 1;
  static RoParams R;  // ~4 KB
  memset(&R, 0, sizeof(R));
  Program pg{&R, err};
  add_policy(pg, P);
  add_env(pg, m);
  if (!pg.ok) return 1;
  const int n_active = (int)P->h_members.size();
  R.mode = RO_MODE_ROLLOUT;
  R.O = P->c.O; R.A = P->c.A; R.n = a.n_episodes;
  R.tiles_per_member = (a.n_episodes + 127) / 128;
  R.max_episode_steps = a.max_episode_steps;
  R.n_steps = (a.horizon > 0 && a.horizon < a.max_episode_steps) ? a.horizon : a.max_episode_steps;
  R.ln_eps_env = m->ln_eps; R.ln_eps_pi = P->c.ln_eps;
  R.env_ln_scale = m->ln_scale; R.env_ln_bias = m->ln_bias;
  R.member_list = P->member_list;
  R.seeds = P->seeds;
  R.eval_counter = 0x8000000000000000ull | ((a.eval_seed & 0x7FFFFFFFull) << 32);
  R.init_obs = a.init_obs; R.init_obs_ms = a.init_obs_ms;
  R.noise = a.noise; R.noise_ms = a.noise_ms;
  R.obs_work = P->ro.obs_work;
  R.noise_next = P->ro.noise_next;
  R.success = a.success ? a.success : P->ro.success;
  R.length = a.length ? a.length : P->ro.length;
  R.success_sum = P->ro.success_sum; R.return_sum = P->ro.return_sum;
  R.steps_run = a.steps_run ? a.steps_run : P->ro.steps_run;
  R.traj_obs = a.traj_obs; R.traj_act = a.traj_act; R.traj_logit = a.traj_logit;
  launch_head_image(P, stream);
  CK(cudaMemsetAsync(P->ro.success_sum, 0, sizeof(float) * P->c.P, stream));
  CK(cudaMemsetAsync(P->ro.return_sum, 0, sizeof(float) * P->c.P, stream));
  CK(cudaMemsetAsync(R.steps_run, 0, sizeof(int) * P->c.P, stream));
  if (rollout_gen() >= 2 && rollout16_eligible(R)) {
    const bool pair = rollout_pair_mode(true) && R.tiles_per_member >= 2;
    if (!pg.widen(pair)) return 1;
    const NetLayout& N = P->pi.L;
    for (int l = 0; l < N.n_hidden; ++l)
      tile_policy_kernel<<<dim3(16, n_active), 256, 0, stream>>>(P->pi.shadow, P->ro.pi_tiled, N.S, N.offW[l], N.K(l), N.H,
                                                               P->member_list, pair ? 1 : 0);
    P->ro.launches += N.n_hidden;
    CK(launch_rollout16(R, n_active, pair, stream));
  } else {
    CK(launch_rollout(R, n_active, stream));
  }
  CK(cudaGetLastError());
  finish_rollout_kernel<<<(n_active + 127) / 128, 128, 0, stream>>>(P->member_list, n_active, a.n_episodes,
                                                                    P->ro.success_sum, P->ro.return_sum,
                                                                    a.success_rate, a.return_mean);
  CK(cudaGetLastError());
  P->ro.launches += 3;
  return 0;
}

int sample_actions_impl(Population* P, int member, const float* obs, int n, const float* noise,
                        unsigned long long seed, float* out, cudaStream_t stream, std::string* err) {
  auto bad = [&](const char* msg) {
    if (err) *err = msg;
    return 1;
  };
  if (!P || !obs || !out || n < 1 || member < 0 || member >= P->c.P) return bad("sample_actions: bad arguments");
  if (P->c.O + P->c.A > 64) return bad("sample_actions: need ob_dim + action_dim <= 64");
  if (cudaError_t e = rollout_init(); e != cudaSuccess) return bad(cudaGetErrorString(e));
  if (ensure_scratch(P, 1, err)) return 1;
  static RoParams R;
  memset(&R, 0, sizeof(R));
  Program pg{&R, err};
  add_policy(pg, P);
  if (!pg.ok) return 1;
  R.mode = RO_MODE_POLICY;
  R.O = P->c.O; R.A = P->c.A; R.n = n;
  R.tiles_per_member = (n + 127) / 128;
  R.max_episode_steps = 1; R.n_steps = 1;
  R.ln_eps_pi = P->c.ln_eps; R.ln_eps_env = 1e-6f;
  // one CTA row of the grid = the requested member
  R.member_list = P->ro.identity + member;
  R.seeds = P->seeds;
  R.eval_counter = 0xC000000000000000ull | ((seed & 0x3FFFFFFFull) << 32);
  // init_obs / noise are indexed by member id with stride 0: the caller's buffers are for this member only
  R.init_obs = obs; R.init_obs_ms = 0;
  R.noise = noise; R.noise_ms = 0;
  R.actions_out = out;
  launch_head_image(P, stream);
  CK(launch_rollout(R, 1, stream));
  CK(cudaGetLastError());
  P->ro.launches += 2;
  return 0;
}

int env_step_impl(EnvModel* m, const float* obs, const float* act, int n, float* next_obs, float* logits,
                  cudaStream_t stream, std::string* err) {
  auto bad = [&](const char* msg) {
    if (err) *err = msg;
    return 1;
  };
  if (!m || !obs || !act || !next_obs || !logits || n < 1) return bad("env_step: bad arguments");
  if (cudaError_t e = rollout_init(); e != cudaSuccess) return bad(cudaGetErrorString(e));
  static RoParams R;
  memset(&R, 0, sizeof(R));
  Program pg{&R, err};
  add_env(pg, m);
  if (!pg.ok) return 1;
  R.mode = RO_MODE_ENV;
  R.O = m->O; R.A = m->A; R.n = n;
  R.tiles_per_member = (n + 127) / 128;
  R.max_episode_steps = 1; R.n_steps = 1;
  R.ln_eps_env = m->ln_eps; R.ln_eps_pi = 1e-6f;
  R.env_ln_scale = m->ln_scale; R.env_ln_bias = m->ln_bias;
  R.init_obs = obs; R.actions_in = act;
  R.next_obs_out = next_obs; R.logits_out = logits;
  CK(launch_rollout(R, 1, stream));
  CK(cudaGetLastError());
  return 0;
}

}  // namespace fqlpop
