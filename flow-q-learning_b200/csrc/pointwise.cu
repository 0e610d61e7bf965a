// HBM-bound kernels of the population FQL step: batch gather + network-input staging, the CUDA-core head
// layers fused with their loss / Euler / TD-target arithmetic, small-gradient reductions, gradient statistics
// and the fused Adam + Polyak + bf16-shadow refresh.
//
// Reference semantics (call site /root/reference/trainer/experiment.py:109):
//   critic_loss, actor_loss (bc_flow + alpha * distill + q), compute_flow_actions (F Euler steps),
//   optax.adam, target_update. Gather: Dataset.sample (/root/reference/task/offline_task_simulated.py:53-58).
#include <cfloat>

#include "philox.cuh"
#include <string>

#include "pop.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kHeadThreads = 256;
constexpr int kRowsPerWarp = 2;  // short serial chains: the heads sit on the step's critical path between two layer-chain launches
constexpr int kRowsPerBlock = (kHeadThreads / 32) * kRowsPerWarp;  // 16

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, s));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, s));
  return v;
}
__device__ __forceinline__ void atomic_max_f(float* a, float v) {
  if (v >= 0.f) atomicMax(reinterpret_cast<int*>(a), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_min_f(float* a, float v) {
  if (v >= 0.f) atomicMin(reinterpret_cast<int*>(a), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned int*>(a), __float_as_uint(v));
}
__device__ __forceinline__ float clip1(float x) { return fminf(fmaxf(x, -1.f), 1.f); }
__device__ __forceinline__ __nv_bfloat16 to_bf(float x) { return __float2bfloat16_rn(x); }

// wT[(e*out + j)*H + k] = theta[(v0+e)*S + offW + k*kHeadPad + j]   (transposed head kernel in shared memory)
__device__ __forceinline__ void stage_head(float* wT, float* bs, const float* theta, long long S, long long offW,
                                           long long offb, int v0, int nb, int out, int H) {
  const int total = nb * out * H;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int e = i / (out * H);
    const int r = i - e * out * H;
    const int k = r / out;
    const int j = r - k * out;
    wT[(e * out + j) * H + k] = theta[(long long)(v0 + e) * S + offW + (long long)k * kHeadPad + j];
  }
  for (int i = threadIdx.x; i < nb * out; i += blockDim.x) {
    const int e = i / out, j = i - e * out;
    bs[i] = theta[(long long)(v0 + e) * S + offb + j];
  }
  __syncthreads();
}

// acc[j] = sum_k h[k] * wT[j*H + k] for j < out; result replicated on every lane.
template <int MAXO>
__device__ __forceinline__ void head_dot(const __nv_bfloat16* hrow, const float* wT, int H, int out, int lane,
                                         float (&acc)[MAXO]) {
#pragma unroll
  for (int j = 0; j < MAXO; ++j) acc[j] = 0.f;
  for (int k0 = lane * 8; k0 < H; k0 += 256) {
    const uint4 u = *reinterpret_cast<const uint4*>(hrow + k0);
    float x[8];
    x[0] = bf16_lo(u.x); x[1] = bf16_hi(u.x); x[2] = bf16_lo(u.y); x[3] = bf16_hi(u.y);
    x[4] = bf16_lo(u.z); x[5] = bf16_hi(u.z); x[6] = bf16_lo(u.w); x[7] = bf16_hi(u.w);
#pragma unroll
    for (int j = 0; j < MAXO; ++j) {
      if (j < out) {
        const float4* w = reinterpret_cast<const float4*>(wT + j * H + k0);
        const float4 w0 = w[0], w1 = w[1];
        float a = acc[j];
        a = fmaf(x[0], w0.x, a); a = fmaf(x[1], w0.y, a); a = fmaf(x[2], w0.z, a); a = fmaf(x[3], w0.w, a);
        a = fmaf(x[4], w1.x, a); a = fmaf(x[5], w1.y, a); a = fmaf(x[6], w1.z, a); a = fmaf(x[7], w1.w, a);
        acc[j] = a;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < MAXO; ++j)
    if (j < out) acc[j] = warp_sum(acc[j]);
}

// ------------------------------------------------------------------------------------------------------
// prep: gather the batch rows, draw (or read) the noise, stage every network input in bf16
// ------------------------------------------------------------------------------------------------------
struct PrepArgs {
  const float *obs, *act, *nobs, *rew, *mask;
  long long n;
  const long long* indices;
  const float *z_critic, *x0, *t, *z_actor, *z_mse;
  const unsigned long long* seeds;
  const unsigned long long* step_counter;
  const int* member_list;
  int O, A, B, E, pad_q, pad_bc;
  float b1, b2;
  __nv_bfloat16 *Xpi, *Xqc, *Xbc, *Xe, *Xtq, *Xqa;
  float *act_f, *vel, *xe, *rew_o, *mask_o, *info, *adam_bc;
  int* adam_step;
  int bump_adam;
};

__global__ void __launch_bounds__(256) prep_kernel(const PrepArgs a) {
  const int p = a.member_list[blockIdx.y];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.x * 8 + warp;
  if (b >= a.B) return;
  const int O = a.O, A = a.A, B = a.B, E = a.E;
  const unsigned long long step = a.step_counter[p];
  const unsigned long long seed = a.seeds[p];
  const long long pb = (long long)p * B + b;

  if (b == 0) {  // per-member, per-step bookkeeping
    if (lane < kInfoSlots) {
      float init = 0.f;
      if (lane == I_Q_MAX) init = -FLT_MAX;
      if (lane == I_Q_MIN) init = FLT_MAX;
      a.info[p * kInfoSlots + lane] = init;
    }
    if (lane == 0 && a.bump_adam) {
      const int t = a.adam_step[p] + 1;
      a.adam_step[p] = t;
      a.adam_bc[2 * p + 0] = (float)(1.0 - pow((double)a.b1, (double)t));
      a.adam_bc[2 * p + 1] = (float)(1.0 - pow((double)a.b2, (double)t));
    }
  }

  long long idx;
  if (a.indices) idx = a.indices[pb];
  else idx = (long long)philox_index(seed, step, (uint32_t)b, (uint32_t)(4 * A + 1), (uint32_t)a.n);
  const float tt = a.t ? a.t[pb] : philox_uniform(seed, step, (uint32_t)b, (uint32_t)(4 * A));
  if (lane == 0) {
    a.rew_o[pb] = a.rew[idx];
    a.mask_o[pb] = a.mask[idx];
  }
  const int pad = a.pad_bc > a.pad_q ? a.pad_bc : a.pad_q;
  for (int c = lane; c < pad; c += 32) {
    float o = 0.f, no = 0.f, ac = 0.f, zc = 0.f, x0 = 0.f, za = 0.f, zm = 0.f;
    const bool is_o = c < O, is_a = (c >= O) && (c < O + A);
    if (is_o) {
      o = a.obs[idx * O + c];
      no = a.nobs[idx * O + c];
    } else if (is_a) {
      const int j = c - O;
      ac = a.act[idx * A + j];
      zc = a.z_critic ? a.z_critic[pb * A + j] : philox_normal(seed, step, (uint32_t)b, (uint32_t)j);
      x0 = a.x0 ? a.x0[pb * A + j] : philox_normal(seed, step, (uint32_t)b, (uint32_t)(A + j));
      za = a.z_actor ? a.z_actor[pb * A + j] : philox_normal(seed, step, (uint32_t)b, (uint32_t)(2 * A + j));
      zm = a.z_mse ? a.z_mse[pb * A + j] : philox_normal(seed, step, (uint32_t)b, (uint32_t)(3 * A + j));
      a.act_f[pb * A + j] = ac;
      a.vel[pb * A + j] = ac - x0;
      a.xe[pb * A + j] = za;
    }
    if (c < a.pad_q) {
      const long long rpi = ((long long)p * 3 * B) * a.pad_q;
      // PI rows: [0,B) = (obs, z_actor) kept for backward; [B,2B) = (next_obs, z_critic); [2B,3B) = (obs, z_mse)
      a.Xpi[rpi + (long long)b * a.pad_q + c] = to_bf(is_o ? o : za);
      a.Xpi[rpi + (long long)(B + b) * a.pad_q + c] = to_bf(is_o ? no : zc);
      a.Xpi[rpi + (long long)(2 * B + b) * a.pad_q + c] = to_bf(is_o ? o : zm);
      for (int e = 0; e < E; ++e) {
        const long long rq = (((long long)p * E + e) * B + b) * a.pad_q + c;
        a.Xqc[rq] = to_bf(is_o ? o : ac);
        if (!is_a) {  // the action columns of these two are written by pi_head
          a.Xtq[rq] = to_bf(no);
          a.Xqa[rq] = to_bf(o);
        }
      }
    }
    if (c < a.pad_bc) {
      const long long rb = (pb)*a.pad_bc + c;
      const float xt = (1.f - tt) * x0 + tt * ac;
      a.Xbc[rb] = to_bf(is_o ? o : (is_a ? xt : (c == O + A ? tt : 0.f)));
      a.Xe[rb] = to_bf(is_o ? o : za);  // t = 0 for the first Euler step
    }
  }
}

// ------------------------------------------------------------------------------------------------------
// head layers (H -> out <= 8 on CUDA cores, fp32 weights, bf16 activations) fused with what follows them
// ------------------------------------------------------------------------------------------------------
// (debug) one timeline record: start ns, end ns, SM id, tag | blockIdx.x << 32
__device__ __forceinline__ void tl_record(long long* timeline, long long t0, long long tag) {
  long long t1;
  unsigned smid;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
  const unsigned long long slot = atomicAdd(reinterpret_cast<unsigned long long*>(timeline), 1ull);
  if (slot < 65536ull) {
    long long* rec = timeline + 8 + 4 * slot;
    rec[0] = t0; rec[1] = t1; rec[2] = smid; rec[3] = tag | ((long long)blockIdx.x << 32);
  }
}

struct HeadArgs {
  long long* timeline;     // (debug) per-CTA records, see fqlpop_debug_timeline
  long long tag;
  const __nv_bfloat16* h;  // [V][M][H] last hidden activations of the pass
  const float* theta;      // arena the head weights come from (theta, or target for the TD target)
  long long S, offW, offb;
  int H, out, nb, M, B, O, A, E, pad_in, F, euler_i, last, q_agg_min, normalize_q;
  float discount;
  const int* member_list;
  const float *act_f, *vel, *rew, *mask, *alphas;
  float *api, *y, *xe, *target_a, *info;
  const float* y_in;
  __nv_bfloat16 *Xtq, *Xqa, *Xe, *dout;
};

#define HEAD_PROLOGUE()                                                                  \
  long long tl_t0 = 0;                                                                   \
  if (a.timeline && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0)); \
  extern __shared__ float sm_head[];                                                     \
  float* wT = sm_head;                                                                   \
  float* bs = sm_head + a.nb * a.out * a.H;                                              \
  const int p = a.member_list[blockIdx.y];                                               \
  stage_head(wT, bs, a.theta, a.S, a.offW, a.offb, p * a.nb, a.nb, a.out, a.H);          \
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;                            \
  const int row_begin = blockIdx.x * kRowsPerBlock + warp * kRowsPerWarp;

#define HEAD_EPILOGUE()                                                                  \
  if (a.timeline && threadIdx.x == 0) tl_record(a.timeline, tl_t0, a.tag);

template <int MAXO>
__device__ __forceinline__ float pick_lane(const float (&acc)[MAXO], const float* bs, int lane) {
  float v = 0.f;
#pragma unroll
  for (int j = 0; j < MAXO; ++j)
    if (j == lane) v = acc[j] + bs[j];
  return v;
}

// one-step policy head on the 3B rows: a_pi (kept), clip(a_pi) -> critic input of the actor pass,
// clip(next action) -> target-critic input, and the mse metric.
__global__ void __launch_bounds__(kHeadThreads) pi_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  float mse_acc = 0.f;
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int r = row_begin + i;
    if (r >= a.M) break;
    float acc[kMaxHeadOut];
    head_dot<kMaxHeadOut>(a.h + ((long long)p * a.M + r) * a.H, wT, a.H, a.out, lane, acc);
    const float v = pick_lane<kMaxHeadOut>(acc, bs, lane);
    const int seg = r / a.B, b = r - seg * a.B;
    if (lane < a.out) {
      const long long pb = (long long)p * a.B + b;
      if (seg == 0) {
        a.api[pb * a.A + lane] = v;
        for (int e = 0; e < a.E; ++e)
          a.Xqa[(((long long)p * a.E + e) * a.B + b) * a.pad_in + a.O + lane] = to_bf(clip1(v));
      } else if (seg == 1) {
        for (int e = 0; e < a.E; ++e)
          a.Xtq[(((long long)p * a.E + e) * a.B + b) * a.pad_in + a.O + lane] = to_bf(clip1(v));
      } else {
        const float d = clip1(v) - a.act_f[pb * a.A + lane];
        mse_acc = fmaf(d, d, mse_acc);
      }
    }
  }
  mse_acc = warp_sum(mse_acc);
  if (lane == 0 && mse_acc != 0.f) atomicAdd(&a.info[p * kInfoSlots + I_MSE], mse_acc / (float)(a.B * a.A));
  HEAD_EPILOGUE();
}

// target critic head: y = r + discount * mask * agg_e Qbar_e(s', a')
__global__ void __launch_bounds__(kHeadThreads) tq_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int b = row_begin + i;
    if (b >= a.B) break;
    float qsum = 0.f, qmin = FLT_MAX;
    for (int e = 0; e < a.E; ++e) {
      float acc[1];
      head_dot<1>(a.h + (((long long)p * a.E + e) * a.B + b) * a.H, wT + e * a.H, a.H, 1, lane, acc);
      const float q = acc[0] + bs[e];
      qsum += q;
      qmin = fminf(qmin, q);
    }
    if (lane == 0) {
      const long long pb = (long long)p * a.B + b;
      const float nq = a.q_agg_min ? qmin : qsum / (float)a.E;
      a.y[pb] = a.rew[pb] + a.discount * a.mask[pb] * nq;
    }
  }
  HEAD_EPILOGUE();
}

// critic head on (s, a): TD loss, its gradient (bf16, column 0 of dout) and the q statistics
__global__ void __launch_bounds__(kHeadThreads) qc_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  float loss = 0.f, qs = 0.f, qmx = -FLT_MAX, qmn = FLT_MAX;
  const float gscale = 2.f / (float)(a.E * a.B);
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int b = row_begin + i;
    if (b >= a.B) break;
    const float y = a.y_in[(long long)p * a.B + b];
    for (int e = 0; e < a.E; ++e) {
      float acc[1];
      const long long vb = ((long long)p * a.E + e) * a.B + b;
      head_dot<1>(a.h + vb * a.H, wT + e * a.H, a.H, 1, lane, acc);
      const float q = acc[0] + bs[e];
      const float d = q - y;
      loss = fmaf(d, d, loss);
      qs += q;
      qmx = fmaxf(qmx, q);
      qmn = fminf(qmn, q);
      if (lane == 0) a.dout[vb * kHeadPad] = to_bf(d * gscale);
    }
  }
  if (lane == 0 && row_begin < a.B) {
    const float inv = 1.f / (float)(a.E * a.B);
    atomicAdd(&a.info[p * kInfoSlots + I_CRITIC_LOSS], loss * inv);
    atomicAdd(&a.info[p * kInfoSlots + I_Q_MEAN], qs * inv);
    atomic_max_f(&a.info[p * kInfoSlots + I_Q_MAX], qmx);
    atomic_min_f(&a.info[p * kInfoSlots + I_Q_MIN], qmn);
  }
  HEAD_EPILOGUE();
}

// critic head on (s, clip(a_pi)): mean_e Q_e -> sums for q_loss / q / the normalisation constant
__global__ void __launch_bounds__(kHeadThreads) qa_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  float qsum = 0.f, qabs = 0.f;
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int b = row_begin + i;
    if (b >= a.B) break;
    float qa = 0.f;
    for (int e = 0; e < a.E; ++e) {
      float acc[1];
      head_dot<1>(a.h + (((long long)p * a.E + e) * a.B + b) * a.H, wT + e * a.H, a.H, 1, lane, acc);
      qa += acc[0] + bs[e];
    }
    qa /= (float)a.E;
    qsum += qa;
    qabs += fabsf(qa);
  }
  if (lane == 0 && row_begin < a.B) {
    atomicAdd(&a.info[p * kInfoSlots + I_Q], qsum / (float)a.B);
    atomicAdd(&a.info[p * kInfoSlots + I_QABS], qabs / (float)a.B);
  }
  HEAD_EPILOGUE();
}

// BC flow head on (s, x_t, t): flow-matching loss and its gradient
__global__ void __launch_bounds__(kHeadThreads) bc_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  float loss = 0.f;
  const float inv = 1.f / (float)(a.B * a.A);
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int b = row_begin + i;
    if (b >= a.B) break;
    float acc[kMaxHeadOut];
    const long long pb = (long long)p * a.B + b;
    head_dot<kMaxHeadOut>(a.h + pb * a.H, wT, a.H, a.out, lane, acc);
    const float v = pick_lane<kMaxHeadOut>(acc, bs, lane);
    if (lane < a.out) {
      const float d = v - a.vel[pb * a.A + lane];
      loss = fmaf(d, d, loss);
      a.dout[pb * kHeadPad + lane] = to_bf(2.f * d * inv);
    }
  }
  loss = warp_sum(loss);
  if (lane == 0 && row_begin < a.B) atomicAdd(&a.info[p * kInfoSlots + I_BC_LOSS], loss * inv);
  HEAD_EPILOGUE();
}

// BC flow head inside the Euler rollout: x <- x + v / F, restage the next step's input (bf16) with t = (i+1)/F
__global__ void __launch_bounds__(kHeadThreads) euler_head_kernel(const HeadArgs a) {
  HEAD_PROLOGUE();
  const float invF = 1.f / (float)a.F;
  const float tnext = (float)(a.euler_i + 1) / (float)a.F;
  for (int i = 0; i < kRowsPerWarp; ++i) {
    const int b = row_begin + i;
    if (b >= a.B) break;
    float acc[kMaxHeadOut];
    const long long pb = (long long)p * a.B + b;
    head_dot<kMaxHeadOut>(a.h + pb * a.H, wT, a.H, a.out, lane, acc);
    const float v = pick_lane<kMaxHeadOut>(acc, bs, lane);
    if (lane < a.out) {
      const float x = a.xe[pb * a.A + lane] + v * invF;
      a.xe[pb * a.A + lane] = x;
      if (a.last) a.target_a[pb * a.A + lane] = clip1(x);
      else a.Xe[pb * a.pad_in + a.O + lane] = to_bf(x);
    } else if (lane == a.out && !a.last) {
      a.Xe[pb * a.pad_in + a.O + a.A] = to_bf(tnext);
    }
  }
}

// d(actor loss)/d a_pi = alpha * 2 (a_pi - a_flow) / (B A) + 1[|a_pi| <= 1] * s * sum_e dQ_e/da  (bf16 into dout)
struct PiDoutArgs {
  const float *api, *target_a, *dx0, *alphas;
  float* info;
  __nv_bfloat16* dout;
  const int* member_list;
  int B, A, E, O, pad_in, normalize_q;
};
__global__ void __launch_bounds__(256) pi_dout_kernel(const PiDoutArgs a) {
  const int p = a.member_list[blockIdx.y];
  const int i = blockIdx.x * 256 + threadIdx.x;
  float d2 = 0.f;
  if (i < a.B * a.A) {
    const int b = i / a.A, j = i - b * a.A;
    const long long pb = (long long)p * a.B + b;
    const float ap = a.api[pb * a.A + j];
    const float d = ap - a.target_a[pb * a.A + j];
    d2 = d * d;
    float dq = 0.f;
    for (int e = 0; e < a.E; ++e) dq += a.dx0[(((long long)p * a.E + e) * a.B + b) * a.pad_in + a.O + j];
    const float s = a.normalize_q ? 1.f / a.info[p * kInfoSlots + I_QABS] : 1.f;
    const float g = a.alphas[p] * 2.f * d / (float)(a.B * a.A) + (fabsf(ap) <= 1.f ? s * dq : 0.f);
    a.dout[pb * kHeadPad + j] = to_bf(g);
  }
  d2 = warp_sum(d2);
  if ((threadIdx.x & 31) == 0 && d2 != 0.f)
    atomicAdd(&a.info[p * kInfoSlots + I_DISTILL_LOSS], d2 / (float)(a.B * a.A));
}

// ------------------------------------------------------------------------------------------------------
// bias / LayerNorm gradients: sum the per-m-tile partials the dgrad epilogues wrote; head bias from dout
// ------------------------------------------------------------------------------------------------------
struct SmallGradJob {
  float* dst;          // arena grad base + leaf offset
  long long dst_ms;    // S
  const float* part;   // [V][m_tiles][N] or nullptr
  long long part_ms;
  const __nv_bfloat16* dout;  // [V][M][64] (head bias job)
  long long dout_ms;
  int m_tiles, N, M, nb;
};
constexpr int kMaxSmallJobs = 48;
struct SmallGradArgs {
  SmallGradJob j[kMaxSmallJobs];
  const int* member_list;
};
__global__ void __launch_bounds__(256) small_grads_kernel(const __grid_constant__ SmallGradArgs a) {
  const SmallGradJob& J = a.j[blockIdx.x];
  const int p = a.member_list[blockIdx.y];
  __shared__ float red[4][kHeadPad];
  for (int e = 0; e < J.nb; ++e) {
    const long long v = (long long)p * J.nb + e;
    if (J.part) {
      for (int n = threadIdx.x; n < J.N; n += 256) {
        float s = 0.f;
        for (int mt = 0; mt < J.m_tiles; ++mt) s += J.part[v * J.part_ms + (long long)mt * J.N + n];
        J.dst[v * J.dst_ms + n] = s;
      }
    } else {
      const int col = threadIdx.x & 63, rg = threadIdx.x >> 6;
      float s = 0.f;
      for (int r = rg; r < J.M; r += 4) s += __bfloat162float(J.dout[v * J.dout_ms + (long long)r * kHeadPad + col]);
      red[rg][col] = s;
      __syncthreads();
      if (rg == 0) J.dst[v * J.dst_ms + col] = red[0][col] + red[1][col] + red[2][col] + red[3][col];
      __syncthreads();
    }
  }
}

// ------------------------------------------------------------------------------------------------------
// gradient statistics (upstream TrainState.apply_loss_fn: grad/max, grad/min, grad/norm = sum of leaf norms)
// ------------------------------------------------------------------------------------------------------
struct LeafRef {
  const float* g;  // arena grad base
  long long ms;    // S
  long long off;
  int count, nb;
};
constexpr int kMaxLeaves = 64;
struct StatsArgs {
  LeafRef l[kMaxLeaves];
  const int* member_list;
  float* leaf_stats;  // [P][kMaxLeaves][4]: max, min, sumsq
};
__global__ void __launch_bounds__(256) grad_stats_kernel(const __grid_constant__ StatsArgs a) {
  const LeafRef& L = a.l[blockIdx.x];
  const int p = a.member_list[blockIdx.y];
  float mx = 0.f, mn = 0.f, ss = 0.f;  // zero is always among the leaves upstream (target params have zero grads)
  for (int e = 0; e < L.nb; ++e) {
    const float4* g = reinterpret_cast<const float4*>(L.g + ((long long)p * L.nb + e) * L.ms + L.off);
    for (int i = threadIdx.x; i < L.count / 4; i += 256) {
      const float4 x = g[i];
      mx = fmaxf(mx, fmaxf(fmaxf(x.x, x.y), fmaxf(x.z, x.w)));
      mn = fminf(mn, fminf(fminf(x.x, x.y), fminf(x.z, x.w)));
      ss = fmaf(x.x, x.x, fmaf(x.y, x.y, fmaf(x.z, x.z, fmaf(x.w, x.w, ss))));
    }
  }
  __shared__ float r[3][8];
  mx = warp_max(mx); mn = warp_min(mn); ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) {
    r[0][threadIdx.x >> 5] = mx; r[1][threadIdx.x >> 5] = mn; r[2][threadIdx.x >> 5] = ss;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) {
      mx = fmaxf(mx, r[0][w]); mn = fminf(mn, r[1][w]); ss += r[2][w];
    }
    float* o = a.leaf_stats + ((long long)p * kMaxLeaves + blockIdx.x) * 4;
    o[0] = mx; o[1] = mn; o[2] = ss;
  }
}
__global__ void grad_stats_final_kernel(const float* leaf_stats, int n_leaves, const int* member_list, float* info) {
  const int p = member_list[blockIdx.x];
  if (threadIdx.x != 0) return;
  float mx = 0.f, mn = 0.f, nrm = 0.f;
  for (int l = 0; l < n_leaves; ++l) {
    const float* o = leaf_stats + ((long long)p * kMaxLeaves + l) * 4;
    mx = fmaxf(mx, o[0]); mn = fminf(mn, o[1]); nrm += sqrtf(o[2]);
  }
  info[p * kInfoSlots + I_GRAD_MAX] = mx;
  info[p * kInfoSlots + I_GRAD_MIN] = mn;
  info[p * kInfoSlots + I_GRAD_NORM] = nrm;
}

// ------------------------------------------------------------------------------------------------------
// fused Adam (optax.adam: b1, b2, eps, bias-corrected) + Polyak target + bf16 shadow refresh. HBM-bound:
// per trainable parameter 16 B read (g, p, mu, nu) + 14 B written (p, mu, nu, bf16 shadow), critic + 4 B read
// + 6 B written for the target and its shadow.
// ------------------------------------------------------------------------------------------------------
struct AdamArgs {
  long long* timeline;
  long long tag;
  float *theta, *mu, *nu;
  const float* grad;
  __nv_bfloat16* shadow;
  float* target;
  __nv_bfloat16* tshadow;
  long long S;
  int nb;
  int n_virtual;  // active members x nb
  const int* member_list;
  const float* adam_bc;
  float lr, b1, b2, eps, tau;
};
// Persistent, deliberately NARROW grid (a few dozen CTAs of 1024 threads, grid-stride over every active member): the
// optimizer is HBM-bound and needs bytes in flight, not SMs -- while a grid of thousands of small CTAs keeps every SM
// partially occupied until it drains, which starves the layer-chain kernels of the other member groups (their CTAs
// need a whole SM). With ~32 CTAs per launch the three optimizer launches of a group occupy well under 148 SMs and
// overlap the chains. theta / mu / nu / target are touched once per step: streaming loads and stores (evict-first)
// keep them from pushing the bf16 shadow weights -- which the next chains stream from L2 -- out of the cache.
constexpr int kAdamThreads = 1024;
constexpr int kAdamUnroll = 2;  // float4 quads per thread per trip: all their loads are issued before the first use
                                // (2 x 4-5 x 16 B in flight per thread, ~160 KB per SM at 64 registers per thread)
__device__ __forceinline__ void adam_quad(const AdamArgs& a, long long o, float ibc1, float ibc2, float4 g, float4 th,
                                          float4 m, float4 v, float4 tg) {
#define ADAM1(c)                                                        \
  m.c = a.b1 * m.c + (1.f - a.b1) * g.c;                                \
  v.c = a.b2 * v.c + (1.f - a.b2) * g.c * g.c;                          \
  th.c = th.c - a.lr * (m.c * ibc1) / (sqrtf(v.c * ibc2) + a.eps);
  ADAM1(x) ADAM1(y) ADAM1(z) ADAM1(w)
#undef ADAM1
  __stcs(reinterpret_cast<float4*>(a.theta + o), th);
  __stcs(reinterpret_cast<float4*>(a.mu + o), m);
  __stcs(reinterpret_cast<float4*>(a.nu + o), v);
  uint2 s;
  s.x = pack_bf16x2(th.x, th.y);
  s.y = pack_bf16x2(th.z, th.w);
  *reinterpret_cast<uint2*>(a.shadow + o) = s;
  if (a.target) {
    tg.x = a.tau * th.x + (1.f - a.tau) * tg.x;
    tg.y = a.tau * th.y + (1.f - a.tau) * tg.y;
    tg.z = a.tau * th.z + (1.f - a.tau) * tg.z;
    tg.w = a.tau * th.w + (1.f - a.tau) * tg.w;
    __stcs(reinterpret_cast<float4*>(a.target + o), tg);
    uint2 ts;
    ts.x = pack_bf16x2(tg.x, tg.y);
    ts.y = pack_bf16x2(tg.z, tg.w);
    *reinterpret_cast<uint2*>(a.tshadow + o) = ts;
  }
}
__global__ void __launch_bounds__(kAdamThreads, 1) adam_kernel(const AdamArgs a) {
  long long tl_t0 = 0;
  const bool tl_on = a.timeline && threadIdx.x == 0;
  if (tl_on) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0));
  // virtual member by virtual member: the element index needs no 64-bit division (a grid-stride walk over
  // n_virtual x S / 4 quads spent most of its issue slots -- 63-67 % busy, ncu -- on `idx / n4`), and the member's
  // pointers and bias corrections are computed once per member instead of once per quad
  const int n4 = (int)(a.S >> 2);
  const int stride = (int)gridDim.x * kAdamThreads;
  for (int av = 0; av < a.n_virtual; ++av) {
    const int am = av / a.nb, e = av - am * a.nb;
    const int p = a.member_list[am];
    const long long base = ((long long)p * a.nb + e) * a.S;
    const float ibc1 = 1.f / a.adam_bc[2 * p], ibc2 = 1.f / a.adam_bc[2 * p + 1];
    for (int i0 = (int)blockIdx.x * kAdamThreads + (int)threadIdx.x; i0 < n4; i0 += kAdamUnroll * stride) {
      long long o[kAdamUnroll];
      float4 g[kAdamUnroll], th[kAdamUnroll], m[kAdamUnroll], v[kAdamUnroll], tg[kAdamUnroll];
      bool ok[kAdamUnroll];
#pragma unroll
      for (int u = 0; u < kAdamUnroll; ++u) {
        const int i = i0 + u * stride;
        ok[u] = i < n4;
        o[u] = base + 4ll * (ok[u] ? i : i0);
      }
#pragma unroll
      for (int u = 0; u < kAdamUnroll; ++u) {
        g[u] = __ldcs(reinterpret_cast<const float4*>(a.grad + o[u]));
        th[u] = __ldcs(reinterpret_cast<const float4*>(a.theta + o[u]));
        m[u] = __ldcs(reinterpret_cast<const float4*>(a.mu + o[u]));
        v[u] = __ldcs(reinterpret_cast<const float4*>(a.nu + o[u]));
        tg[u] = a.target ? __ldcs(reinterpret_cast<const float4*>(a.target + o[u])) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < kAdamUnroll; ++u)
        if (ok[u]) adam_quad(a, o[u], ibc1, ibc2, g[u], th[u], m[u], v[u], tg[u]);
    }
  }
  if (tl_on) tl_record(a.timeline, tl_t0, a.tag);
}

// The same update for the leaves no weight-gradient GEMM produces (biases, LayerNorm scale / bias): the kernels are
// updated inside the GEMM epilogue (gemm.cu, EPI_ADAM), these few thousand parameters per virtual member here.
constexpr int kMaxAdamSegs = 32;
struct AdamLeafArgs {
  AdamArgs a;
  long long seg_off[kMaxAdamSegs];
  int seg_n[kMaxAdamSegs];
};
__global__ void __launch_bounds__(128) adam_leaves_kernel(const __grid_constant__ AdamLeafArgs L) {
  const AdamArgs& a = L.a;
  const int av = blockIdx.y;
  const int am = av / a.nb, e = av - am * a.nb;
  const int p = a.member_list[am];
  const float ibc1 = 1.f / a.adam_bc[2 * p], ibc2 = 1.f / a.adam_bc[2 * p + 1];
  const long long base = ((long long)p * a.nb + e) * a.S + L.seg_off[blockIdx.x];
  for (int i = threadIdx.x; i < L.seg_n[blockIdx.x] / 4; i += 128) {
    const long long o = base + 4ll * i;
    const float4 g = *reinterpret_cast<const float4*>(a.grad + o);
    const float4 th = *reinterpret_cast<const float4*>(a.theta + o);
    const float4 m = *reinterpret_cast<const float4*>(a.mu + o);
    const float4 v = *reinterpret_cast<const float4*>(a.nu + o);
    const float4 tg = a.target ? *reinterpret_cast<const float4*>(a.target + o) : make_float4(0.f, 0.f, 0.f, 0.f);
    adam_quad(a, o, ibc1, ibc2, g, th, m, v, tg);
  }
}

// theta -> bf16 shadow (after importing parameters)
__global__ void __launch_bounds__(256) to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                      long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 x = reinterpret_cast<const float4*>(src)[i];
    uint2 s;
    s.x = pack_bf16x2(x.x, x.y);
    s.y = pack_bf16x2(x.z, x.w);
    reinterpret_cast<uint2*>(dst)[i] = s;
  }
}

__global__ void fill_col0_kernel(__nv_bfloat16* dout, long long rows, float value) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < rows) dout[i * kHeadPad] = __float2bfloat16_rn(value);
}

struct FinishArgs {
  float* info;
  const float* alphas;
  const int* member_list;
  unsigned long long* step_counter;
  int n_active, normalize_q, bump;
};
__global__ void finish_info_kernel(const FinishArgs a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_active) return;
  const int p = a.member_list[i];
  if (a.bump) a.step_counter[p] += 1ull;
  float* f = a.info + p * kInfoSlots;
  const float s = a.normalize_q ? 1.f / f[I_QABS] : 1.f;
  f[I_Q_LOSS] = -f[I_Q] * s;
  f[I_ACTOR_LOSS] = f[I_BC_LOSS] + a.alphas[p] * f[I_DISTILL_LOSS] + f[I_Q_LOSS];
}

HeadArgs head_args(const Population* P, const PassBuf& pass, const Arena& ar, const float* theta) {
  HeadArgs a{};
  const NetLayout& L = ar.L;
  a.h = pass.hl(L.n_hidden - 1);
  a.theta = theta;
  a.S = L.S;
  a.offW = L.offW[L.n_hidden];
  a.offb = L.offb[L.n_hidden];
  a.H = L.H;
  a.out = L.out_dim;
  a.nb = L.nb;
  a.M = pass.M;
  a.B = P->c.B;
  a.O = P->c.O;
  a.A = P->c.A;
  a.E = P->c.E;
  a.pad_in = pass.in_pad;
  a.F = P->c.F;
  a.q_agg_min = P->c.q_agg_min;
  a.normalize_q = P->c.normalize_q_loss;
  a.discount = P->c.discount;
  a.member_list = P->cur_list();
  a.act_f = P->act_f; a.vel = P->vel; a.rew = P->rew; a.mask = P->mask; a.alphas = P->alphas;
  a.api = P->api; a.y = P->y; a.xe = P->xe; a.target_a = P->target_a; a.info = P->info;
  a.y_in = P->y;
  a.Xtq = P->TQ.X0; a.Xqa = P->QA.X0; a.Xe = P->EU.X0;
  return a;
}
size_t head_smem(const HeadArgs& a) { return (size_t)(a.nb * a.out * a.H + a.nb * a.out) * sizeof(float); }
dim3 head_grid(const Population* P, int rows) {
  return dim3((rows + kRowsPerBlock - 1) / kRowsPerBlock, (unsigned)P->cur_n());
}

}  // namespace

// ====================================== launchers ======================================================
void launch_prep(Population* P, cudaStream_t s, int bump_adam) {
  PrepArgs a{};
  a.obs = P->ds.obs; a.act = P->ds.act; a.nobs = P->ds.next_obs; a.rew = P->ds.rew; a.mask = P->ds.mask;
  a.n = P->ds.n;
  a.indices = P->noise.indices;
  a.z_critic = P->noise.z_critic; a.x0 = P->noise.x0; a.t = P->noise.t; a.z_actor = P->noise.z_actor;
  a.z_mse = P->noise.z_mse;
  a.seeds = P->seeds;
  a.step_counter = P->step_counter;
  a.member_list = P->cur_list();
  a.O = P->c.O; a.A = P->c.A; a.B = P->c.B; a.E = P->c.E;
  a.pad_q = P->PI.in_pad; a.pad_bc = P->BC.in_pad;
  a.b1 = P->c.adam_b1; a.b2 = P->c.adam_b2;
  a.Xpi = P->PI.X0; a.Xqc = P->QC.X0; a.Xbc = P->BC.X0; a.Xe = P->EU.X0; a.Xtq = P->TQ.X0; a.Xqa = P->QA.X0;
  a.act_f = P->act_f; a.vel = P->vel; a.xe = P->xe; a.rew_o = P->rew; a.mask_o = P->mask; a.info = P->info;
  a.adam_bc = P->adam_bc;
  a.adam_step = P->adam_step;
  a.bump_adam = bump_adam;
  prep_kernel<<<dim3((P->c.B + 7) / 8, (unsigned)P->cur_n()), 256, 0, s>>>(a);
}

void launch_pi_head(Population* P, cudaStream_t s) {
  HeadArgs a = head_args(P, P->PI, P->pi, P->pi.theta);
  a.timeline = P->timeline; a.tag = 20 | ((long long)P->g_off << 8);
  pi_head_kernel<<<head_grid(P, a.M), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_tq_head(Population* P, cudaStream_t s) {
  HeadArgs a = head_args(P, P->TQ, P->critic, P->critic.target);
  a.timeline = P->timeline; a.tag = 21 | ((long long)P->g_off << 8);
  tq_head_kernel<<<head_grid(P, a.B), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_qc_head(Population* P, cudaStream_t s) {
  HeadArgs a = head_args(P, P->QC, P->critic, P->critic.theta);
  a.dout = P->bQC.dout;
  a.timeline = P->timeline; a.tag = 22 | ((long long)P->g_off << 8);
  qc_head_kernel<<<head_grid(P, a.B), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_qa_head(Population* P, cudaStream_t s) {
  HeadArgs a = head_args(P, P->QA, P->critic, P->critic.theta);
  a.timeline = P->timeline; a.tag = 23 | ((long long)P->g_off << 8);
  qa_head_kernel<<<head_grid(P, a.B), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_bc_head(Population* P, cudaStream_t s) {
  HeadArgs a = head_args(P, P->BC, P->bc, P->bc.theta);
  a.dout = P->bBC.dout;
  a.timeline = P->timeline; a.tag = 24 | ((long long)P->g_off << 8);
  bc_head_kernel<<<head_grid(P, a.B), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_euler_head(Population* P, int i, cudaStream_t s) {
  HeadArgs a = head_args(P, P->EU, P->bc, P->bc.theta);
  a.euler_i = i;
  a.last = (i == P->c.F - 1) ? 1 : 0;
  euler_head_kernel<<<head_grid(P, a.B), kHeadThreads, head_smem(a), s>>>(a);
}
void launch_pi_dout(Population* P, cudaStream_t s) {
  PiDoutArgs a{};
  a.api = P->api; a.target_a = P->target_a; a.dx0 = P->dx0_qa; a.alphas = P->alphas; a.info = P->info;
  a.dout = P->bPI.dout;
  a.member_list = P->cur_list();
  a.B = P->c.B; a.A = P->c.A; a.E = P->c.E; a.O = P->c.O; a.pad_in = P->QA.in_pad;
  a.normalize_q = P->c.normalize_q_loss;
  pi_dout_kernel<<<dim3((P->c.B * P->c.A + 255) / 256, (unsigned)P->cur_n()), 256, 0, s>>>(a);
}

static int add_small_jobs(SmallGradArgs& a, int n, const Arena& ar, const BwdBuf& bw) {
  const NetLayout& L = ar.L;
  for (int l = 0; l < L.n_hidden; ++l) {
    SmallGradJob j{};
    j.dst_ms = L.S; j.part_ms = (long long)bw.part_rows * L.H; j.m_tiles = bw.part_rows; j.N = L.H; j.nb = L.nb;
    j.dst = ar.grad + L.offb[l]; j.part = bw.dbias_part[l];
    a.j[n++] = j;
    if (L.ln) {
      j.dst = ar.grad + L.offg[l]; j.part = bw.dgamma_part[l];
      a.j[n++] = j;
      j.dst = ar.grad + L.offbeta[l]; j.part = bw.dbeta_part[l];
      a.j[n++] = j;
    }
  }
  SmallGradJob j{};
  j.dst = ar.grad + L.offb[L.n_hidden]; j.dst_ms = L.S; j.dout = bw.dout; j.dout_ms = (long long)bw.M * kHeadPad;
  j.M = bw.M; j.nb = L.nb; j.N = kHeadPad;
  a.j[n++] = j;
  return n;
}
void launch_small_grads(Population* P, int which, cudaStream_t s) {
  static SmallGradArgs a;  // ~3 KB
  int n = 0;
  if (which == 0) n = add_small_jobs(a, n, P->critic, P->bQC);
  else if (which == 1) n = add_small_jobs(a, n, P->bc, P->bBC);
  else n = add_small_jobs(a, n, P->pi, P->bPI);
  a.member_list = P->cur_list();
  small_grads_kernel<<<dim3(n, (unsigned)P->cur_n()), 256, 0, s>>>(a);
}

static int add_leaves(StatsArgs& a, int n, const Arena& ar) {
  const NetLayout& L = ar.L;
  auto add = [&](long long off, int count) {
    a.l[n].g = ar.grad; a.l[n].ms = L.S; a.l[n].off = off; a.l[n].count = count; a.l[n].nb = L.nb;
    ++n;
  };
  for (int l = 0; l <= L.n_hidden; ++l) {
    add(L.offW[l], L.K(l) * L.N(l));
    add(L.offb[l], L.N(l));
    if (L.ln && l < L.n_hidden) {
      add(L.offg[l], L.H);
      add(L.offbeta[l], L.H);
    }
  }
  return n;
}
void launch_grad_stats(Population* P, cudaStream_t s) {
  static StatsArgs a;
  int n = 0;
  n = add_leaves(a, n, P->critic);
  n = add_leaves(a, n, P->bc);
  n = add_leaves(a, n, P->pi);
  a.member_list = P->cur_list();
  a.leaf_stats = P->leaf_stats;
  grad_stats_kernel<<<dim3(n, (unsigned)P->cur_n()), 256, 0, s>>>(a);
  grad_stats_final_kernel<<<(unsigned)P->cur_n(), 32, 0, s>>>(P->leaf_stats, n, P->cur_list(), P->info);
}

static void launch_adam_one(Population* P, Arena& ar, cudaStream_t s) {
  AdamArgs a{};
  a.theta = ar.theta; a.mu = ar.mu; a.nu = ar.nu; a.grad = ar.grad; a.shadow = ar.shadow;
  a.target = ar.target; a.tshadow = ar.tshadow;
  a.S = ar.L.S; a.nb = ar.L.nb;
  a.member_list = P->cur_list();
  a.adam_bc = P->adam_bc;
  a.lr = P->c.lr; a.b1 = P->c.adam_b1; a.b2 = P->c.adam_b2; a.eps = P->c.adam_eps; a.tau = P->c.tau;
  a.timeline = P->timeline;
  a.tag = (&ar == &P->critic ? 30 : (&ar == &P->bc ? 31 : 32)) | ((long long)P->g_off << 8);
  a.n_virtual = P->cur_n() * ar.L.nb;
  // Tried and dropped in round 2 (profiles/r02_adam_experiments.txt): a bulk-copy (cp.async.bulk) streaming variant with a
  // small SM footprint -- it moves fewer bytes per SM and second than this register-path kernel (49 vs 68-85 GB/s per
  // SM at 8-12 CTAs per launch), and the step time is insensitive to the CTA count / group phasing anyway: the optimizer's
  // ~310 us per step is its HBM time (1.8 GB at ~5.8 TB/s) and adds to the chain time whatever the schedule.
  static const int ctas = [] { const char* e = getenv("FQLPOP_ADAM_CTAS"); return e ? atoi(e) : 32; }();
  adam_kernel<<<ctas > 0 ? ctas : 32, kAdamThreads, 0, s>>>(a);
}
void launch_adam(Population* P, int which, cudaStream_t s) {
  launch_adam_one(P, which == 0 ? P->critic : (which == 1 ? P->bc : P->pi), s);
}

// Adam on the bias / LayerNorm leaves only (the kernels are updated by the weight-gradient GEMM's epilogue).
void launch_adam_leaves(Population* P, int which, cudaStream_t s) {
  Arena& ar = which == 0 ? P->critic : (which == 1 ? P->bc : P->pi);
  static AdamLeafArgs L;
  AdamArgs& a = L.a;
  a = AdamArgs{};
  a.theta = ar.theta; a.mu = ar.mu; a.nu = ar.nu; a.grad = ar.grad; a.shadow = ar.shadow;
  a.target = ar.target; a.tshadow = ar.tshadow;
  a.S = ar.L.S; a.nb = ar.L.nb;
  a.member_list = P->cur_list();
  a.adam_bc = P->adam_bc;
  a.lr = P->c.lr; a.b1 = P->c.adam_b1; a.b2 = P->c.adam_b2; a.eps = P->c.adam_eps; a.tau = P->c.tau;
  a.n_virtual = P->cur_n() * ar.L.nb;
  const NetLayout& N = ar.L;
  int n = 0;
  for (int l = 0; l <= N.n_hidden; ++l) {
    L.seg_off[n] = N.offb[l]; L.seg_n[n] = N.N(l); ++n;
    if (N.ln && l < N.n_hidden) {
      L.seg_off[n] = N.offg[l]; L.seg_n[n] = N.H; ++n;
      L.seg_off[n] = N.offbeta[l]; L.seg_n[n] = N.H; ++n;
    }
  }
  adam_leaves_kernel<<<dim3(n, (unsigned)a.n_virtual), 128, 0, s>>>(L);
}

void launch_finish_info(Population* P, int bump, cudaStream_t s) {
  FinishArgs a{};
  a.info = P->info; a.alphas = P->alphas; a.member_list = P->cur_list(); a.step_counter = P->step_counter;
  a.n_active = P->cur_n(); a.normalize_q = P->c.normalize_q_loss; a.bump = bump;
  finish_info_kernel<<<(a.n_active + 127) / 128, 128, 0, s>>>(a);
}

void launch_refresh_shadow(Population* P, cudaStream_t s) {
  for (Arena* ar : {&P->critic, &P->bc, &P->pi}) {
    const long long n4 = (long long)ar->V * ar->L.S / 4;
    to_bf16_kernel<<<1184, 256, 0, s>>>(ar->theta, ar->shadow, n4);
    if (ar->target) to_bf16_kernel<<<1184, 256, 0, s>>>(ar->target, ar->tshadow, n4);
  }
}

void launch_to_bf16(const float* src, __nv_bfloat16* dst, long long n4, cudaStream_t s) {
  to_bf16_kernel<<<(unsigned)min(1184LL, (n4 + 255) / 256), 256, 0, s>>>(src, dst, n4);
}

void launch_fill_qa_dout(Population* P, cudaStream_t s) {
  const long long rows = (long long)P->bQA.V * P->bQA.M;
  const float v = -1.f / (float)(P->c.E * P->c.B);
  fill_col0_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(P->bQA.dout, rows, v);
}

// device Philox stream exposed for the bit-exact parity test against oracle/philox_oracle.py
__global__ void philox_fill_kernel(unsigned long long seed, unsigned long long step, int rows, int slots, int kind,
                                   unsigned int n, float* out_f, unsigned int* out_u) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * slots) return;
  const uint32_t row = (uint32_t)(i / slots), slot = (uint32_t)(i % slots);
  if (kind == 0) {
    const Philox4 r = philox4x32_10(seed, step, row, slot);
    out_u[4 * i + 0] = r.x; out_u[4 * i + 1] = r.y; out_u[4 * i + 2] = r.z; out_u[4 * i + 3] = r.w;
  } else if (kind == 1) {
    out_f[i] = philox_uniform(seed, step, row, slot);
  } else if (kind == 2) {
    out_u[i] = philox_index(seed, step, row, slot, n);
  } else {
    out_f[i] = philox_normal(seed, step, row, slot);
  }
}
int philox_fill(unsigned long long seed, unsigned long long step, int rows, int slots, int kind, unsigned int n,
                void* out_host) {
  const size_t count = (size_t)rows * slots * (kind == 0 ? 4 : 1);
  void* d = nullptr;
  if (cudaMalloc(&d, count * 4) != cudaSuccess) return 1;
  philox_fill_kernel<<<(rows * slots + 255) / 256, 256>>>(seed, step, rows, slots, kind, n, (float*)d, (unsigned int*)d);
  cudaError_t e = cudaMemcpy(out_host, d, count * 4, cudaMemcpyDeviceToHost);
  cudaFree(d);
  return e == cudaSuccess ? 0 : 1;
}

// ---- fqlpop_philox_fill_noalloc: the same kernel, staging through a fixed file-scope device array -------------------
// (no counterpart in the reference; added for callers that must not allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.) Every GPU that loads the
// library holds kPhiloxNoAllocMaxWords * 4 bytes = 1 MiB until exit. Every call shares the array and it keeps the last
// call's contents: the kernel writes each of the `count` words a call copies out, so nothing has to start at zero, but
// two calls that overlap on one GPU (different host threads / streams) race on it.
static __device__ __align__(256) unsigned int g_philox_scratch[kPhiloxNoAllocMaxWords];

int philox_fill_noalloc(unsigned long long seed, unsigned long long step, int rows, int slots, int kind, unsigned int n,
                        void* out_host, std::string* err) {
  auto bad = [&](const std::string& m) {
    if (err) *err = m;
    return 1;
  };
  const unsigned long long count = (unsigned long long)rows * (unsigned long long)slots * (kind == 0 ? 4ull : 1ull);
  if (count > (unsigned long long)kPhiloxNoAllocMaxWords)
    return bad("philox_fill_noalloc: " + std::to_string(count) + " output words exceed the fixed scratch of " +
               std::to_string(kPhiloxNoAllocMaxWords) + " words (use fqlpop_philox_fill)");
  // the array's name is a host-side stand-in in host code: ask the runtime for its device address, once per device,
  // with that device current
  constexpr int kMaxDevices = 64;
  static unsigned int* dev_ptr[kMaxDevices] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return bad(std::string("philox_fill_noalloc: cudaGetDevice: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return bad("philox_fill_noalloc: device ordinal out of range");
  if (!dev_ptr[dev]) {
    unsigned int* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_philox_scratch);
    if (e != cudaSuccess || !ptr)
      return bad(std::string("philox_fill_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    dev_ptr[dev] = ptr;
  }
  unsigned int* d = dev_ptr[dev];
  philox_fill_kernel<<<(rows * slots + 255) / 256, 256>>>(seed, step, rows, slots, kind, n, (float*)d, d);
  e = cudaGetLastError();
  if (e != cudaSuccess) return bad(std::string("philox_fill_noalloc: launch: ") + cudaGetErrorString(e));
  e = cudaMemcpy(out_host, d, (size_t)count * 4, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return bad(std::string("philox_fill_noalloc: copy: ") + cudaGetErrorString(e));
  return 0;
}

int pointwise_init() {
  // head kernels stage up to nb*out*H floats; raise the dynamic shared-memory limit once
  const int bytes = 64 * 1024;
  cudaError_t e = cudaSuccess;
  for (auto f : {(const void*)pi_head_kernel, (const void*)tq_head_kernel, (const void*)qc_head_kernel,
                 (const void*)qa_head_kernel, (const void*)bc_head_kernel, (const void*)euler_head_kernel}) {
    e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

}  // namespace fqlpop
