// World-model training kernels (see wm.h). fp32 throughout; every reduction runs in a fixed order, so a step is
// deterministic and reproducible across launches.
#include "wm.h"

#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "philox.cuh"

namespace fqlpop {

// ---------------------------------------------------------------------------------------------------------------------
// C[M][N] (+)= epi(op(A) op(B)), op(A)(m,k) = TA ? A[k*lda+m] : A[m*lda+k], op(B)(k,n) = TB ? B[n*ldb+k] : B[k*ldb+n].
// 32 x 32 output tiles, 128-deep k tiles (K <= 256 here: at most two rounds of global loads, all of a round's loads in
// flight together), 256 threads, 4 rows x 1 column per thread; the k loop runs in index order (deterministic).
// One kernel serves the forward Dense layers (bias / relu / residual), the input-gradient products (TB, relu mask taken
// from the saved activation) and the weight-gradient products (TA, reduction over the batch).
// ---------------------------------------------------------------------------------------------------------------------
enum WmEpi : int { WE_NONE = 0, WE_BIAS = 1, WE_BIAS_RELU = 2, WE_BIAS_RESID = 3, WE_MASK = 4 };

constexpr int kWmTile = 32;
constexpr int kWmBK = 128;

template <bool TA, bool TB>
__global__ void __launch_bounds__(256) wm_gemm_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                      int ldb, float* __restrict__ C, int ldc, int M, int N, int K,
                                                      int epi, const float* __restrict__ bias,
                                                      const float* __restrict__ aux, int ldaux, int accumulate,
                                                      int k_chunk, long long c_split_stride) {
  __shared__ float As[kWmBK][kWmTile + 1];  // [k][m]
  __shared__ float Bs[kWmBK][kWmTile + 1];  // [k][n]
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int m0 = blockIdx.y * kWmTile, n0 = blockIdx.x * kWmTile;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  // split-K (weight gradients over a long batch axis): slice blockIdx.z reduces k in [z * k_chunk, ...) into its own
  // partial C; wm_reduce_kernel adds the slices in index order
  const int kbeg = blockIdx.z * k_chunk;
  K = min(K, kbeg + k_chunk);
  C += blockIdx.z * c_split_stride;
  float ra[16], rb[16];
  auto load_round = [&](int k0) {  // every global load of a round is issued before anything consumes one
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int e = tid + i * 256;
      const int am = TA ? (e & 31) : (e >> 7), ak = TA ? (e >> 5) : (e & 127);
      const int bk = TB ? (e & 127) : (e >> 5), bn = TB ? (e >> 7) : (e & 31);
      const int gm = m0 + am, gka = k0 + ak, gkb = k0 + bk, gn = n0 + bn;
      ra[i] = (gm < M && gka < K) ? (TA ? A[(long long)gka * lda + gm] : A[(long long)gm * lda + gka]) : 0.f;
      rb[i] = (gkb < K && gn < N) ? (TB ? B[(long long)gn * ldb + gkb] : B[(long long)gkb * ldb + gn]) : 0.f;
    }
  };
  if (kbeg < K) load_round(kbeg);
  for (int k0 = kbeg; k0 < K; k0 += kWmBK) {
    const int kt = min(kWmBK, K - k0);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int e = tid + i * 256;
      As[TA ? (e >> 5) : (e & 127)][TA ? (e & 31) : (e >> 7)] = ra[i];
      Bs[TB ? (e & 127) : (e >> 5)][TB ? (e >> 7) : (e & 31)] = rb[i];
    }
    __syncthreads();
    if (k0 + kWmBK < K) load_round(k0 + kWmBK);  // next round's loads fly while this round is multiplied
#pragma unroll 8
    for (int k = 0; k < kt; ++k) {
      const float b = Bs[k][tx];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(As[k][ty * 4 + i], b, acc[i]);
    }
    __syncthreads();
  }
  const int gn = n0 + tx;
  if (gn >= N) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
    float r = acc[i];
    if (epi == WE_BIAS || epi == WE_BIAS_RELU || epi == WE_BIAS_RESID) r += bias[gn];
    if (epi == WE_BIAS_RELU) r = fmaxf(r, 0.f);
    if (epi == WE_BIAS_RESID) r += aux[(long long)gm * ldaux + gn];
    if (epi == WE_MASK) r = aux[(long long)gm * ldaux + gn] > 0.f ? r : 0.f;
    float* c = C + (long long)gm * ldc + gn;
    *c = accumulate ? *c + r : r;
  }
}

// The same product for the LARGE problems (the multistep trainer's products over all T*B rows: the frozen termination
// predictor forward / backward and the split-K weight gradients): 64 x 64 tiles, 4 x 4 outputs per thread, 16-deep k slabs
// with the next slab's global loads in flight while the current one is multiplied -- 16 FMAs per two 16-byte shared-memory
// loads instead of 4 per two scalar ones. Same arguments and epilogues as wm_gemm_kernel.
constexpr int kWmTile64 = 64;
constexpr int kWmBK64 = 16;
template <bool TA, bool TB>
__global__ void __launch_bounds__(256) wm_gemm64_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                        int ldb, float* __restrict__ C, int ldc, int M, int N, int K,
                                                        int epi, const float* __restrict__ bias,
                                                        const float* __restrict__ aux, int ldaux, int accumulate,
                                                        int k_chunk, long long c_split_stride) {
  __shared__ __align__(16) float As[kWmBK64][kWmTile64 + 4];  // [k][m]
  __shared__ __align__(16) float Bs[kWmBK64][kWmTile64 + 4];  // [k][n]
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * kWmTile64, n0 = blockIdx.x * kWmTile64;
  const int kbeg = blockIdx.z * k_chunk;
  K = min(K, kbeg + k_chunk);
  C += blockIdx.z * c_split_stride;
  float acc[4][4] = {};
  float ra[4], rb[4];
  auto load_round = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      const int am = TA ? (e & 63) : (e >> 4), ak = TA ? (e >> 6) : (e & 15);
      const int bk = TB ? (e & 15) : (e >> 6), bn = TB ? (e >> 4) : (e & 63);
      const int gm = m0 + am, gka = k0 + ak, gkb = k0 + bk, gn = n0 + bn;
      ra[i] = (gm < M && gka < K) ? (TA ? A[(long long)gka * lda + gm] : A[(long long)gm * lda + gka]) : 0.f;
      rb[i] = (gkb < K && gn < N) ? (TB ? B[(long long)gn * ldb + gkb] : B[(long long)gkb * ldb + gn]) : 0.f;
    }
  };
  if (kbeg < K) load_round(kbeg);
  for (int k0 = kbeg; k0 < K; k0 += kWmBK64) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * 256;
      As[TA ? (e >> 6) : (e & 15)][TA ? (e & 63) : (e >> 4)] = ra[i];
      Bs[TB ? (e & 15) : (e >> 6)][TB ? (e >> 4) : (e & 63)] = rb[i];
    }
    __syncthreads();
    if (k0 + kWmBK64 < K) load_round(k0 + kWmBK64);
#pragma unroll
    for (int k = 0; k < kWmBK64; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int gn = n0 + tx * 4 + j;
      if (gn >= N) continue;
      float r = acc[i][j];
      if (epi == WE_BIAS || epi == WE_BIAS_RELU || epi == WE_BIAS_RESID) r += bias[gn];
      if (epi == WE_BIAS_RELU) r = fmaxf(r, 0.f);
      if (epi == WE_BIAS_RESID) r += aux[(long long)gm * ldaux + gn];
      if (epi == WE_MASK) r = aux[(long long)gm * ldaux + gn] > 0.f ? r : 0.f;
      float* c = C + (long long)gm * ldc + gn;
      *c = accumulate ? *c + r : r;
    }
  }
}

// out[n] = sum_m X[m][n] (* Y[m][n]): bias gradients and the LayerNorm scale / bias gradients. One thread per column.
// (both at once for the LayerNorm: out = sum X, out_xy = sum X * Y). blockIdx.y = row slice (rows_per_slice rows each,
// its sums go to out + y * N: partials for wm_reduce_kernel when gridDim.y > 1).
__global__ void __launch_bounds__(256) wm_colsum_kernel(const float* __restrict__ X, const float* __restrict__ Y, int ld,
                                                        int M, int N, float* __restrict__ out,
                                                        float* __restrict__ out_xy, int rows_per_slice) {
  // 32 columns x 8 row lanes per CTA: row lane ty sums rows m0 + ty, m0 + ty + 8, ... (loads of a warp are 128 contiguous
  // bytes, 8 independent accumulation chains per column), then the 8 partials are added in lane order.
  __shared__ float sm[2][8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + tx;
  const int m0 = blockIdx.y * rows_per_slice, m1 = min(M, m0 + rows_per_slice);
  float s = 0.f, sxy = 0.f;
  if (n < N) {
    for (int mb = m0 + ty; mb < m1; mb += 256) {  // 32 rows per round, all loads of a round independent
      float x[32], y[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int m = mb + 8 * i;
        x[i] = m < m1 ? X[(long long)m * ld + n] : 0.f;
        y[i] = (Y && m < m1) ? Y[(long long)m * ld + n] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        s += x[i];
        sxy += x[i] * y[i];
      }
    }
  }
  sm[0][ty][tx] = s;
  sm[1][ty][tx] = sxy;
  __syncthreads();
  if (ty == 0 && n < N) {
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a += sm[0][k][tx];
      b += sm[1][k][tx];
    }
    out[(long long)blockIdx.y * N + n] = a;
    if (Y) out_xy[(long long)blockIdx.y * N + n] = b;
  }
}

// out[i] = sum_s part[s][i], slices added in index order
__global__ void wm_reduce_kernel(const float* __restrict__ part, int S, long long n, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float s = 0.f;
  for (int k = 0; k < S; ++k) s += part[(long long)k * n + i];
  out[i] = s;
}

// x0 = LayerNorm(concat(obs, act)) * scale + bias (envmodel/baseline.py:31-32; flax: var = E[x^2] - E[x]^2, eps inside
// the rsqrt). One warp per row; keeps xhat for the scale gradient.
// act_stride: elements between the action rows of consecutive batch entries (A, or T * A inside a [B][T][A] window).
__global__ void wm_ln_concat_kernel(const float* __restrict__ obs, const float* __restrict__ act, int act_stride, int O,
                                    int A, int rows, const float* __restrict__ scale, const float* __restrict__ bias,
                                    float eps, float* __restrict__ x0, float* __restrict__ xhat,
                                    float* __restrict__ rstd_out) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int D = O + A;
  float s = 0.f, s2 = 0.f;
  for (int j = lane; j < D; j += 32) {
    const float v = j < O ? obs[(long long)row * O + j] : act[(long long)row * act_stride + (j - O)];
    s += v;
    s2 += v * v;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  const float mean = s / D, var = fmaxf(s2 / D - mean * mean, 0.f), rstd = rsqrtf(var + eps);
  for (int j = lane; j < D; j += 32) {
    const float v = j < O ? obs[(long long)row * O + j] : act[(long long)row * act_stride + (j - O)];
    const float h = (v - mean) * rstd;
    xhat[(long long)row * D + j] = h;
    x0[(long long)row * D + j] = h * scale[j] + bias[j];
  }
  if (rstd_out && lane == 0) rstd_out[row] = rstd;
}

// Back-propagation through time, one step: dx = LayerNorm backward of dx0 (d loss / d LN output) to concat(obs, act);
// the observation part plus this step's output gradient (residual connection o' = f(o, a) + o) is what the PREVIOUS
// step's prediction receives: g_prev[b][j] += g_cur[b][j] + dx[b][j], j < O. One warp per row.
__global__ void wm_ln_bwd_kernel(const float* __restrict__ dx0, const float* __restrict__ xhat,
                                 const float* __restrict__ rstd, const float* __restrict__ scale, int O, int D, int rows,
                                 const float* __restrict__ g_cur, float* __restrict__ g_prev) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  float s1 = 0.f, s2 = 0.f;
  for (int j = lane; j < D; j += 32) {
    const float d = dx0[(long long)row * D + j] * scale[j];
    s1 += d;
    s2 += d * xhat[(long long)row * D + j];
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  const float m1 = s1 / D, m2 = s2 / D, r = rstd[row];
  for (int j = lane; j < O; j += 32) {
    const float d = dx0[(long long)row * D + j] * scale[j];
    const float dx = r * (d - m1 - xhat[(long long)row * D + j] * m2);
    g_prev[(long long)row * O + j] += g_cur[(long long)row * O + j] + dx;
  }
}

// o_0 = observations[:, 0, :] of a [B][T][O] window (envmodel/multistep.py:50)
__global__ void wm_take_first_kernel(const float* __restrict__ obs, int B, int T, int O, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * O) out[i] = obs[(long long)(i / O) * T * O + (i % O)];
}

// termination predictor input: next_obs (* dropout multiplier)
__global__ void wm_mask_copy_kernel(const float* __restrict__ x, const float* __restrict__ mask, long long n,
                                    float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = mask ? x[i] * mask[i] : x[i];
}

// StepLoader.sample on the device: uniform rows with replacement, Philox(seed, step, row, 0)
__global__ void wm_gather_kernel(const float* __restrict__ obs, const float* __restrict__ act,
                                 const float* __restrict__ next_obs, const float* __restrict__ rewards, long long n_rows,
                                 int O, int A, int batch, unsigned long long seed, const int* __restrict__ step,
                                 float* __restrict__ b_obs, float* __restrict__ b_act, float* __restrict__ b_next,
                                 float* __restrict__ b_rew) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= batch) return;
  const long long idx = philox_index(seed, (unsigned long long)*step, (uint32_t)row, 0u, (uint32_t)n_rows);
  for (int j = lane; j < O; j += 32) {
    b_obs[(long long)row * O + j] = obs[idx * O + j];
    b_next[(long long)row * O + j] = next_obs[idx * O + j];
  }
  for (int j = lane; j < A; j += 32) b_act[(long long)row * A + j] = act[idx * A + j];
  if (lane == 0) b_rew[row] = rewards[idx];
}

// MultistepLoader.sample on the device (utils/data_loader.py:25-47): batch row b draws an episode (Philox slot 0) and a
// window start inside it (slot 1, uniform in [0, episode_len - T)), then copies the T transitions of the window.
// One warp per (b, t).
__global__ void wm_gather_windows_kernel(const float* __restrict__ obs, const float* __restrict__ act,
                                         const float* __restrict__ next_obs, const float* __restrict__ rewards,
                                         long long n_rows, int episode_len, int O, int A, int batch, int T,
                                         unsigned long long seed, const int* __restrict__ step, float* __restrict__ b_obs,
                                         float* __restrict__ b_act, float* __restrict__ b_next, float* __restrict__ b_rew) {
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= (long long)batch * T) return;
  const int b = (int)(w / T), t = (int)(w % T);
  const unsigned long long st = (unsigned long long)*step;
  const long long episode = philox_index(seed, st, (uint32_t)b, 0u, (uint32_t)(n_rows / episode_len));
  const long long start = philox_index(seed, st, (uint32_t)b, 1u, (uint32_t)(episode_len - T));
  const long long src = episode * episode_len + start + t;
  for (int j = lane; j < O; j += 32) {
    b_obs[w * O + j] = obs[src * O + j];
    b_next[w * O + j] = next_obs[src * O + j];
  }
  for (int j = lane; j < A; j += 32) b_act[w * A + j] = act[src * A + j];
  if (lane == 0) b_rew[w] = rewards[src];
}

// fixed-order block reduction of up to 8 values per thread (256 threads); result valid in thread 0
template <int NV>
__device__ __forceinline__ void wm_block_reduce(float (&v)[NV], float (*sm)[NV]) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int o = 16; o; o >>= 1) v[i] += __shfl_xor_sync(0xffffffffu, v[i], o);
  if (lane == 0)
#pragma unroll
    for (int i = 0; i < NV; ++i) sm[w][i] = v[i];
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      float s = 0.f;
      for (int k = 0; k < 8; ++k) s += sm[k][i];
      v[i] = s;
    }
  }
}

// Training steps only (step != nullptr): hand the optimiser its count (step[1]) and advance the counter the sampler of
// the NEXT step reads (step[0]). Runs in the single-CTA loss kernel: after this step's gather, before its Adam launch.
__device__ __forceinline__ void wm_tick(int* step, float* logs) {
  if (!step) return;
  const int t = step[0];
  step[1] = t;
  step[0] = t + 1;
  logs[WL_STEP] = (float)(t + 1);
}

__device__ __forceinline__ float wm_sigmoid(float x) { return 1.f / (1.f + __expf(-x)); }
// optax.sigmoid_binary_cross_entropy(logits, labels), stable form
__device__ __forceinline__ float wm_bce(float x, float t) { return fmaxf(x, 0.f) - x * t + log1pf(expf(-fabsf(x))); }

// state_prediction_loss (envmodel/loss.py:65-111) with weighted_binary_cross_entropy (:12-28): logs, d loss / d pred of
// the squared-error term and d loss / d logit of the termination term; `terminated` = rewards == 0.
// Rows of pred / logit / dpred / dlogit are TIME-major (row = t * B + b) while next_obs / rewards arrive batch-major
// ([B][T][...]); T == 1 makes the two the same. Any grid: every CTA leaves its partial sums in `part`, the LAST CTA to
// finish (ticket counter) adds them in CTA order and writes the logs, so the result does not depend on scheduling.
__global__ void __launch_bounds__(256) wm_state_loss_kernel(const float* __restrict__ pred,
                                                            const float* __restrict__ next_obs,
                                                            const float* __restrict__ logit,
                                                            const float* __restrict__ rewards, int rows, int B, int T,
                                                            int O, float ttw, float tw, float rw,
                                                            float* __restrict__ dpred, float* __restrict__ dlogit,
                                                            float* __restrict__ logs, int* __restrict__ step,
                                                            float* __restrict__ part, unsigned int* __restrict__ ticket) {
  __shared__ float sm[8][5];
  __shared__ bool last;
  const float denom = 1.f + tw + rw;
  const bool term = tw > 0.f;
  float v[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // squared error, true ce, false ce, n true, n false
  for (int r = blockIdx.x * 256 + threadIdx.x; r < rows; r += gridDim.x * 256) {
    const long long src = T > 1 ? (long long)(r % B) * T + r / B : r;
    float se = 0.f;
#pragma unroll 7
    for (int j = 0; j < O; ++j) {
      const float d = pred[(long long)r * O + j] - next_obs[src * O + j];
      se += d * d;
      if (dpred) dpred[(long long)r * O + j] = 2.f * d / ((float)rows * O * denom);
    }
    v[0] += se;
    if (term) {
      const float x = logit[r], t = rewards[src] == 0.f ? 1.f : 0.f, ce = wm_bce(x, t);
      if (t == 1.f) { v[1] += ce; v[3] += 1.f; } else { v[2] += ce; v[4] += 1.f; }
      if (dlogit) dlogit[r] = (tw / denom) * (t == 1.f ? ttw : 1.f) * (wm_sigmoid(x) - t) / ((ttw + 1.f) * rows);
    }
  }
  wm_block_reduce<5>(v, sm);
  if (threadIdx.x == 0) {
    for (int i = 0; i < 5; ++i) part[blockIdx.x * 5 + i] = v[i];
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    if (last) {
      __threadfence();
      *ticket = 0u;
      for (int i = 0; i < 5; ++i) v[i] = 0.f;
      for (unsigned int c = 0; c < gridDim.x; ++c)
        for (int i = 0; i < 5; ++i) v[i] += ((volatile float*)part)[c * 5 + i];
      const float mse = v[0] / ((float)rows * O);
      const float tl = term ? (ttw * v[1] + v[2]) / ((ttw + 1.f) * rows) : 0.f;
      logs[WL_NEXT_OBS_LOSS] = mse;
      logs[WL_TERM_LOSS] = tl;
      logs[WL_TRUE_LOSS] = term ? v[1] / v[3] : 0.f;   // 0 / 0 = NaN without positives, as in the reference
      logs[WL_FALSE_LOSS] = term ? v[2] / v[4] : 0.f;
      logs[WL_RECON_LOSS] = 0.f;
      logs[WL_LOSS] = (mse + tw * tl) / denom;
      wm_tick(step, logs);
    }
  }
}

// focal_loss (envmodel/loss.py:31-62) + the eval metrics of termination_predictor_trainer.py:88-126.
__global__ void __launch_bounds__(256) wm_focal_loss_kernel(const float* __restrict__ logit,
                                                            const float* __restrict__ rewards, int rows, float alpha,
                                                            float gamma, float* __restrict__ dlogit,
                                                            float* __restrict__ logs, int* __restrict__ step) {
  __shared__ float sm[8][6];
  float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // loss sum, n true, correct, true positives, predicted positives, -
  for (int b = threadIdx.x; b < rows; b += 256) {
    const float x = logit[b], t = rewards[b] == 0.f ? 1.f : 0.f, p = wm_sigmoid(x), ce = wm_bce(x, t);
    float l, g;
    if (t == 1.f) {
      const float mod = powf(1.f - p, gamma);
      l = alpha * mod * ce;
      g = alpha * mod * (-gamma * p * ce - (1.f - p));
    } else {
      const float mod = powf(p, gamma);
      l = (1.f - alpha) * mod * ce;
      g = (1.f - alpha) * mod * (gamma * (1.f - p) * ce + p);
    }
    if (dlogit) dlogit[b] = g / rows;
    const float pr = x > 0.f ? 1.f : 0.f;
    v[0] += l; v[1] += t; v[2] += (pr == t) ? 1.f : 0.f; v[3] += pr * t; v[4] += pr;
  }
  wm_block_reduce<6>(v, sm);
  if (threadIdx.x == 0) {
    const float loss = v[0] / rows, nt = v[1], nf = rows - v[1];
    logs[WL_LOSS] = loss;
    logs[WL_TRUE_LOSS] = loss * nt / (nt + 1e-8f);   // the reference scales the MEAN loss by the class count (loss.py:57-60)
    logs[WL_FALSE_LOSS] = loss * nf / (nf + 1e-8f);
    const float prec = v[4] > 0.f ? v[3] / v[4] : 1.f, rec = nt > 0.f ? v[3] / nt : 1.f;
    logs[WL_ACCURACY] = v[2] / rows;
    logs[WL_PRECISION] = prec;
    logs[WL_RECALL] = rec;
    logs[WL_F1] = (prec + rec) > 0.f ? 2.f * prec * rec / (prec + rec) : 0.f;
    wm_tick(step, logs);
  }
}

// optax.adam(cosine_decay_schedule(lr0, decay_steps)): count starts at 0, the schedule is read BEFORE the increment,
// bias correction uses count + 1; eps = 1e-8 outside the square root.
__global__ void wm_adam_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                               const float* __restrict__ g, long long n, const int* __restrict__ step, float lr0,
                               int decay_steps, float* __restrict__ logs) {
  const int t = step[1];
  const float frac = (float)min(t, decay_steps) / (float)decay_steps;
  const float lr = lr0 * 0.5f * (1.f + cospif(frac));
  const float bc1 = 1.f - powf(0.9f, (float)(t + 1)), bc2 = 1.f - powf(0.999f, (float)(t + 1));
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) logs[WL_LR] = lr;
  if (i >= n) return;
  const float gi = g[i];
  const float mi = 0.9f * m[i] + 0.1f * gi, vi = 0.999f * v[i] + 0.001f * gi * gi;
  m[i] = mi;
  v[i] = vi;
  p[i] -= lr * (mi / bc1) / (sqrtf(vi / bc2) + 1e-8f);
}



// ---------------------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------------------
WmNet wm_layout(int in_dim, int n_hidden, const int* hidden, int out_dim, bool input_ln) {
  WmNet n;
  memset(&n, 0, sizeof(n));
  n.n_layers = n_hidden + 1;
  n.dims[0] = in_dim;
  for (int i = 0; i < n_hidden; ++i) n.dims[i + 1] = hidden[i];
  n.dims[n.n_layers] = out_dim;
  long long off = 0;
  n.off_ln_scale = n.off_ln_bias = -1;
  if (input_ln) {
    n.off_ln_scale = off; off += in_dim;
    n.off_ln_bias = off; off += in_dim;
  }
  for (int l = 0; l < n.n_layers; ++l) {
    n.offW[l] = off; off += (long long)n.dims[l] * n.dims[l + 1];
    n.offB[l] = off; off += n.dims[l + 1];
  }
  n.n_params = off;
  return n;
}

struct WmTrainer {
  WmConfig cfg;
  WmNet net, term;
  int device = 0;
  float *params = nullptr, *m = nullptr, *v = nullptr, *grads = nullptr, *term_params = nullptr;
  float* x[kWmMaxLayers + 1] = {};    // x[0] = network input, x[l] = relu output of layer l-1, x[n_layers] = raw output
  float* tx[kWmMaxLayers + 1] = {};   // frozen termination predictor on the prediction (tx[0] aliases the prediction)
  float *xhat = nullptr, *pred = nullptr, *dbuf[2] = {nullptr, nullptr}, *dlogit = nullptr;
  float* gx[kWmMaxLayers + 1] = {};   // gx[l] = d loss / d x[l], one buffer per layer (side streams read them late)
  float *logs = nullptr, *loss_part = nullptr, *rstd = nullptr, *ocur = nullptr, *carry = nullptr;
  float* wparts[kWmMaxLayers] = {};      // multistep: [kWmSplits][din][dout] partial weight gradients, one buffer per layer
  float* cparts[kWmMaxLayers + 1] = {};  // [kWmSplits][dout] partial bias sums; [nl]: LayerNorm bias / scale, 2 x [kWmSplits][D0]
  unsigned int* ticket = nullptr;
  int* step = nullptr;
  int rows = 0;  // batch * sequence length
  // multistep: persistent scan kernels (wm_scan.cuh)
  float* wt = nullptr;        // transposed kernels for the backward scan
  bool scan_ok = false;
  unsigned scan_mask = 0;     // layers whose kernel the scan kernels keep in shared memory
  int scan_smem = 0, scan_maxdim = 0;
  // resident dataset + sampled batch
  const float *ds_obs = nullptr, *ds_act = nullptr, *ds_next = nullptr, *ds_rew = nullptr;
  long long ds_rows = 0;
  float *b_obs = nullptr, *b_act = nullptr, *b_next = nullptr, *b_rew = nullptr;
  // one captured graph per (mode, pointer set)
  struct Key { int train; const void* p[5]; bool operator==(const Key& o) const { return train == o.train && !memcmp(p, o.p, sizeof(p)); } };
  struct Entry { Key key; cudaGraphExec_t exec; int launches; };
  std::vector<Entry> graphs;
  long long launches = 0;
  std::vector<void*> owned;
};

namespace {
#define WM_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      if (err) *err = std::string(#call) + ": " + cudaGetErrorString(e_);                          \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

int wm_alloc(WmTrainer* t, float** p, long long n, std::string* err) {
  WM_CUDA(cudaMalloc((void**)p, sizeof(float) * (size_t)(n > 0 ? n : 1)));
  WM_CUDA(cudaMemset(*p, 0, sizeof(float) * (size_t)(n > 0 ? n : 1)));
  t->owned.push_back(*p);
  return 0;
}

// Kernels go to `cur`. fork(): the side streams (weight gradients / column sums) pick up after what `s` holds so far;
// join(): `s` waits for both. Under stream capture the events become graph edges, so the weight-gradient products and
// the bias sums leave the critical path (only the input-gradient chain is sequential).
struct Launcher {
  cudaStream_t s, w, c, cur;
  int n = 0;
  std::vector<cudaEvent_t> evs;
  cudaEvent_t ev() {
    cudaEvent_t e;
    cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    evs.push_back(e);
    return e;
  }
  void fork() {
    cudaEvent_t e = ev();
    cudaEventRecord(e, s);
    cudaStreamWaitEvent(w, e, 0);
    cudaStreamWaitEvent(c, e, 0);
  }
  void join() {
    cudaEvent_t a = ev(), b = ev();
    cudaEventRecord(a, w);
    cudaEventRecord(b, c);
    cudaStreamWaitEvent(s, a, 0);
    cudaStreamWaitEvent(s, b, 0);
  }
  // splits > 1: slice z of the k range goes to C + z * M * N (ldc must be N); wm_reduce adds the slices
  void gemm(bool ta, bool tb, const float* A, int lda, const float* B, int ldb, float* C, int ldc, int M, int N, int K,
            int epi, const float* bias, const float* aux, int ldaux, int accumulate, int splits = 1,
            int min_tiles64 = 148) {
    const int kc = splits > 1 ? ((K + splits - 1) / splits + kWmBK - 1) / kWmBK * kWmBK : K;
    const long long cs = splits > 1 ? (long long)M * N : 0;
    // enough 64 x 64 tiles for every SM: the register-tiled kernel (the small per-step products keep the 32 x 32 one,
    // whose finer tiles give a latency-bound step more CTAs)
    const long long tiles64 = (long long)((N + kWmTile64 - 1) / kWmTile64) * ((M + kWmTile64 - 1) / kWmTile64) * splits;
    static const bool big_on = [] { const char* e = getenv("FQLPOP_WM_GEMM64"); return e ? atoi(e) != 0 : true; }();
    if (big_on && tiles64 >= min_tiles64) {
      dim3 g64((N + kWmTile64 - 1) / kWmTile64, (M + kWmTile64 - 1) / kWmTile64, splits);
      if (ta) wm_gemm64_kernel<true, false><<<g64, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
      else if (tb) wm_gemm64_kernel<false, true><<<g64, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
      else wm_gemm64_kernel<false, false><<<g64, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
      ++n;
      return;
    }
    dim3 grid((N + kWmTile - 1) / kWmTile, (M + kWmTile - 1) / kWmTile, splits);
    if (ta) wm_gemm_kernel<true, false><<<grid, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
    else if (tb) wm_gemm_kernel<false, true><<<grid, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
    else wm_gemm_kernel<false, false><<<grid, 256, 0, cur>>>(A, lda, B, ldb, C, ldc, M, N, K, epi, bias, aux, ldaux, accumulate, kc, cs);
    ++n;
  }
  // slices > 1: out / out_xy receive [slices][N] partials
  void colsum(const float* X, const float* Y, int ld, int M, int N, float* out, float* out_xy, int slices = 1) {
    dim3 grid((N + 31) / 32, slices);
    wm_colsum_kernel<<<grid, 256, 0, cur>>>(X, Y, ld, M, N, out, out_xy, (M + slices - 1) / slices);
    ++n;
  }
  void reduce(const float* part, int S, long long cnt, float* out) {
    wm_reduce_kernel<<<(int)((cnt + 255) / 256), 256, 0, cur>>>(part, S, cnt, out);
    ++n;
  }
};

// forward Dense stack: act[0] -> act[n_layers]; relu on hidden layers; last layer adds `resid` when given
void wm_forward(Launcher& L, const WmNet& net, const float* params, float* const* act, int rows, const float* resid,
                int ld_resid) {
  for (int l = 0; l < net.n_layers; ++l) {
    const bool last = l == net.n_layers - 1;
    const int epi = last ? (resid ? WE_BIAS_RESID : WE_BIAS) : WE_BIAS_RELU;
    L.gemm(false, false, act[l], net.dims[l], params + net.offW[l], net.dims[l + 1], act[l + 1], net.dims[l + 1], rows,
           net.dims[l + 1], net.dims[l], epi, params + net.offB[l], last ? resid : nullptr, ld_resid, 0);
  }
}
}  // namespace

long long wm_num_params(const WmTrainer* t) { return t->net.n_params; }
long long wm_launches(const WmTrainer* t) { return t->launches; }

}  // namespace fqlpop

namespace fqlpop {
namespace {

// Enqueues one step on L.s. gather != 0: the batch is first drawn from the resident dataset into the b_* buffers.
void wm_issue(WmTrainer* t, Launcher& L, int train, int gather, const float* obs, const float* act, const float* next_obs,
              const float* rew, const float* mask) {
  const WmConfig& c = t->cfg;
  const WmNet& net = t->net;
  const int B = c.batch, O = c.O, A = c.A, nl = net.n_layers;
  cudaStream_t s = L.s;
  if (gather) {
    wm_gather_kernel<<<(B * 32 + 255) / 256, 256, 0, s>>>(t->ds_obs, t->ds_act, t->ds_next, t->ds_rew, t->ds_rows, O, A, B,
                                                         c.seed, t->step, t->b_obs, t->b_act, t->b_next, t->b_rew);
    ++L.n;
    obs = t->b_obs; act = t->b_act; next_obs = t->b_next; rew = t->b_rew;
  }
  const bool state = c.kind == WM_STATE;
  const bool with_term = state && c.termination_weight > 0.f;
  float* g = nullptr;  // gradient wrt the output of the layer being walked
  if (state) {
    wm_ln_concat_kernel<<<(B * 32 + 255) / 256, 256, 0, s>>>(obs, act, A, O, A, B, t->params + net.off_ln_scale,
                                                            t->params + net.off_ln_bias, c.ln_eps, t->x[0], t->xhat,
                                                            nullptr);
    ++L.n;
    wm_forward(L, net, t->params, t->x, B, obs, O);  // x[nl] = Dense_O(x) + obs
    if (with_term) wm_forward(L, t->term, t->term_params, t->tx, B, nullptr, 0);  // tx[0] aliases x[nl]
    wm_state_loss_kernel<<<1, 256, 0, s>>>(t->x[nl], next_obs, with_term ? t->tx[t->term.n_layers] : nullptr, rew, B, B, 1,
                                           O, c.true_termination_weight, c.termination_weight, c.reconstruction_weight,
                                           train ? t->pred : nullptr, train ? t->dlogit : nullptr, t->logs,
                                           train ? t->step : nullptr, t->loss_part, t->ticket);
    ++L.n;
    if (!train) return;
    if (with_term) {  // d loss / d prediction through the frozen termination predictor, added onto the squared-error part
      const WmNet& tn = t->term;
      const float* gt = t->dlogit;
      for (int l = tn.n_layers - 1; l >= 0; --l) {
        float* out = l == 0 ? t->pred : t->dbuf[l & 1];
        L.gemm(false, true, gt, tn.dims[l + 1], t->term_params + tn.offW[l], tn.dims[l + 1], out, tn.dims[l], B, tn.dims[l],
               tn.dims[l + 1], l == 0 ? WE_NONE : WE_MASK, nullptr, l == 0 ? nullptr : t->tx[l], tn.dims[l], l == 0);
        gt = out;
      }
    }
    g = t->pred;
  } else {
    wm_mask_copy_kernel<<<(int)(((long long)B * O + 255) / 256), 256, 0, s>>>(next_obs, train ? mask : nullptr,
                                                                             (long long)B * O, t->x[0]);
    ++L.n;
    wm_forward(L, net, t->params, t->x, B, nullptr, 0);
    wm_focal_loss_kernel<<<1, 256, 0, s>>>(t->x[nl], rew, B, c.focal_alpha, c.focal_gamma, train ? t->dlogit : nullptr,
                                           t->logs, train ? t->step : nullptr);
    ++L.n;
    if (!train) return;
    g = t->dlogit;
  }
  // backward through the trained network: the input-gradient chain stays on the main stream, weight gradients and
  // bias sums of every layer branch off as soon as that layer's output gradient exists
  for (int l = nl - 1; l >= 0; --l) {
    const int din = net.dims[l], dout = net.dims[l + 1];
    L.fork();
    L.cur = L.w;
    L.gemm(true, false, t->x[l], din, g, dout, t->grads + net.offW[l], dout, din, dout, B, WE_NONE, nullptr, nullptr, 0, 0);
    L.cur = L.c;
    L.colsum(g, nullptr, dout, B, dout, t->grads + net.offB[l], nullptr);
    L.cur = L.s;
    if (l > 0 || state) {
      float* out = t->gx[l];
      L.gemm(false, true, g, dout, t->params + net.offW[l], dout, out, din, B, din, dout, l > 0 ? WE_MASK : WE_NONE, nullptr,
             l > 0 ? t->x[l] : nullptr, din, 0);
      g = out;
    }
  }
  if (state)  // input LayerNorm: d scale = sum_b dx0 * xhat, d bias = sum_b dx0
    L.colsum(g, t->xhat, net.dims[0], B, net.dims[0], t->grads + net.off_ln_bias, t->grads + net.off_ln_scale);
  L.join();
  wm_adam_kernel<<<(int)((net.n_params + 255) / 256), 256, 0, s>>>(t->params, t->m, t->v, t->grads, net.n_params, t->step,
                                                                  c.lr, c.decay_steps, t->logs);
  ++L.n;
}


#include "wm_scan.cuh"

constexpr int kWmLossCtas = 128, kWmSplits = 64;
constexpr int kWmEpisodeLen = 1000;  // OGBench episodes (stride 1000 at utils/data_loader.py:16,30)

// One step of the multistep predictor (envmodel/multistep.py:34-54: nn.scan of the baseline cell over the window, each
// prediction fed back as the next observation; loss over all T predictions, state_predictor_trainer.py:79-97).
// Activations are kept TIME-major ([T][B][width] slabs) so that every time step is a plain [B][width] matrix, the frozen
// termination predictor and the loss run ONCE over all T*B rows, and after back-propagation through time each weight
// gradient is ONE product over the T*B rows (split-K, slices added in order) instead of T small ones.
void wm_issue_multistep(WmTrainer* t, Launcher& L, int train, int gather, const float* obs, const float* act,
                        const float* next_obs, const float* rew) {
  const WmConfig& c = t->cfg;
  const WmNet& net = t->net;
  const int B = c.batch, T = c.seq_len, O = c.O, A = c.A, nl = net.n_layers, D0 = net.dims[0], R = t->rows;
  cudaStream_t s = L.s;
  const bool with_term = c.termination_weight > 0.f;
  const float* scale = t->params + net.off_ln_scale;
  if (gather) {
    wm_gather_windows_kernel<<<(int)(((long long)R * 32 + 255) / 256), 256, 0, s>>>(
        t->ds_obs, t->ds_act, t->ds_next, t->ds_rew, t->ds_rows, kWmEpisodeLen, O, A, B, T, c.seed, t->step, t->b_obs,
        t->b_act, t->b_next, t->b_rew);
    ++L.n;
    obs = t->b_obs; act = t->b_act; next_obs = t->b_next; rew = t->b_rew;
  }
  wm_take_first_kernel<<<(B * O + 255) / 256, 256, 0, s>>>(obs, B, T, O, t->ocur);
  ++L.n;
  // FQLPOP_WM_SCAN=0 keeps the per-step launches (2 x T x 5 dependent small kernels) for A/B runs
  static const bool scan_on = [] { const char* e = getenv("FQLPOP_WM_SCAN"); return e ? atoi(e) != 0 : true; }();
  const bool scan = scan_on && t->scan_ok;
  // Time chunks. The scans occupy (B / 4) SMs (64 of 148 at batch 256) and are latency-bound; everything else of the step
  // is products over ALL T*B rows. Rows are time-major, so a chunk of time steps is a contiguous slab of rows: the
  // window is cut into C chunks, the scans run chunk after chunk on the main stream, and the work over a chunk's rows
  // runs beside the NEXT chunk's scan on the idle SMs -- forward: the frozen termination predictor; backward: its
  // input-gradient chain (one chunk ahead) and the split-K weight-gradient / bias / LayerNorm sums (one chunk behind,
  // every chunk filling its own slices of the partial-sum buffers, added in slice order at the end: deterministic).
  // FQLPOP_WM_CHUNKS = 1 / 2 / 4 / 8 (default 4; windows shorter than 16 steps are not cut).
  static const int chunks_env = [] { const char* e = getenv("FQLPOP_WM_CHUNKS"); const int v = e ? atoi(e) : 4; return v >= 8 ? 8 : (v >= 4 ? 4 : (v >= 2 ? 2 : 1)); }();
  const int C = (scan && T >= 16) ? chunks_env : 1;
  const int Tc = (T + C - 1) / C, Sc = kWmSplits / C;
  auto t_lo = [&](int ch) { return std::min(T, ch * Tc); };
  auto t_hi = [&](int ch) { return std::min(T, (ch + 1) * Tc); };
  auto after = [&](cudaStream_t waiter, cudaStream_t done) {
    cudaEvent_t e = L.ev();
    cudaEventRecord(e, done);
    cudaStreamWaitEvent(waiter, e, 0);
  };
  ScanArgs sa{};
  if (scan) {
    sa.net = net; sa.params = t->params; sa.wt = t->wt;
    for (int l = 0; l <= nl; ++l) { sa.x[l] = t->x[l]; sa.gx[l] = t->gx[l]; }
    sa.ocur = t->ocur; sa.xhat = t->xhat; sa.rstd = t->rstd; sa.act = act;
    sa.B = B; sa.T = T; sa.O = O; sa.A = A; sa.maxdim = t->scan_maxdim; sa.eps = c.ln_eps; sa.cache_mask = t->scan_mask;
    sa.carry_g = t->carry;
  }
  // frozen termination predictor over the rows of chunk ch (tx[0] aliases the predictions)
  auto term_forward = [&](int ch) {
    const long long r0 = (long long)t_lo(ch) * B;
    const int rows = (t_hi(ch) - t_lo(ch)) * B;
    if (rows <= 0) return;
    float* tx[kWmMaxLayers + 1];
    for (int l = 0; l <= t->term.n_layers; ++l) tx[l] = t->tx[l] + r0 * t->term.dims[l];
    wm_forward(L, t->term, t->term_params, tx, rows, nullptr, 0);
  };
  if (scan) {
    for (int ch = 0; ch < C; ++ch) {
      sa.t0 = t_lo(ch); sa.t1 = t_hi(ch);
      if (sa.t1 > sa.t0) {
        wm_scan_fwd_kernel<<<(B + kScanRB - 1) / kScanRB, kScanThreads, t->scan_smem, s>>>(sa);
        ++L.n;
      }
      if (with_term) {
        if (C > 1) { after(L.w, s); L.cur = L.w; }
        term_forward(ch);
        L.cur = s;
      }
    }
    if (with_term && C > 1) after(s, L.w);
  } else {
    for (int ti = 0; ti < T; ++ti) {
      float* act_l[kWmMaxLayers + 1];
      for (int l = 0; l < nl; ++l) act_l[l] = t->x[l] + (long long)ti * B * net.dims[l];
      float* o_t = t->ocur + (long long)ti * B * O;
      act_l[nl] = o_t + (long long)B * O;  // o_{t+1}: prediction of this step, input of the next
      wm_ln_concat_kernel<<<(B * 32 + 255) / 256, 256, 0, s>>>(o_t, act + (long long)ti * A, T * A, O, A, B, scale,
                                                              t->params + net.off_ln_bias, c.ln_eps, act_l[0],
                                                              t->xhat + (long long)ti * B * D0, t->rstd + ti * B);
      ++L.n;
      wm_forward(L, net, t->params, act_l, B, o_t, O);
    }
    if (with_term) term_forward(0);   // (C == 1: all rows)
  }
  float* pred = t->ocur + (long long)B * O;  // [T][B][O]
  wm_state_loss_kernel<<<kWmLossCtas, 256, 0, s>>>(pred, next_obs, with_term ? t->tx[t->term.n_layers] : nullptr, rew, R, B,
                                                  T, O, c.true_termination_weight, c.termination_weight,
                                                  c.reconstruction_weight, train ? t->gx[nl] : nullptr,
                                                  train ? t->dlogit : nullptr, t->logs, train ? t->step : nullptr,
                                                  t->loss_part, t->ticket);
  ++L.n;
  if (!train) return;
  // d loss / d prediction through the frozen termination predictor, rows of chunk ch, added onto the squared-error part
  auto term_backward = [&](int ch) {
    const long long r0 = (long long)t_lo(ch) * B;
    const int rows = (t_hi(ch) - t_lo(ch)) * B;
    if (rows <= 0) return;
    const WmNet& tn = t->term;
    const float* gt = t->dlogit + r0;
    for (int l = tn.n_layers - 1; l >= 0; --l) {
      float* out = (l == 0 ? t->gx[nl] : t->dbuf[l & 1]) + r0 * tn.dims[l];
      L.gemm(false, true, gt, tn.dims[l + 1], t->term_params + tn.offW[l], tn.dims[l + 1], out, tn.dims[l], rows, tn.dims[l],
             tn.dims[l + 1], l == 0 ? WE_NONE : WE_MASK, nullptr, l == 0 ? nullptr : t->tx[l] + r0 * tn.dims[l], tn.dims[l],
             l == 0);
      gt = out;
    }
  };
  // parameter-gradient partial sums over the rows of chunk ch: slices [ch * Sc, (ch + 1) * Sc) of every layer's buffers
  auto weight_grads = [&](int ch) {
    const long long r0 = (long long)t_lo(ch) * B;
    const int rows = (t_hi(ch) - t_lo(ch)) * B;
    for (int l = nl - 1; l >= 0; --l) {
      const int din = net.dims[l], dout = net.dims[l + 1];
      float* wp = t->wparts[l] + (long long)ch * Sc * din * dout;
      float* cp = t->cparts[l] + (long long)ch * Sc * dout;
      if (rows <= 0) {   // (a window shorter than the chunk grid: this chunk's slices are zero)
        cudaMemsetAsync(wp, 0, sizeof(float) * (size_t)Sc * din * dout, L.cur);
        cudaMemsetAsync(cp, 0, sizeof(float) * (size_t)Sc * dout, L.cur);
        continue;
      }
      // (a chunk's products run beside a scan on the ~84 SMs it leaves idle: fewer 64 x 64 tiles are enough)
      L.gemm(true, false, t->x[l] + r0 * din, din, t->gx[l + 1] + r0 * dout, dout, wp, dout, din, dout, rows, WE_NONE, nullptr,
             nullptr, 0, 0, Sc, C > 1 ? 84 : 148);
      L.colsum(t->gx[l + 1] + r0 * dout, nullptr, dout, rows, dout, cp, nullptr, Sc);
    }
    float* lp = t->cparts[nl] + (long long)ch * Sc * D0;
    if (rows <= 0) {
      cudaMemsetAsync(lp, 0, sizeof(float) * (size_t)Sc * D0, L.cur);
      cudaMemsetAsync(lp + (long long)kWmSplits * D0, 0, sizeof(float) * (size_t)Sc * D0, L.cur);
    } else {
      L.colsum(t->gx[0] + r0 * D0, t->xhat + r0 * D0, D0, rows, D0, lp, lp + (long long)kWmSplits * D0, Sc);
    }
  };
  if (scan) {
    for (int l = 0; l < nl; ++l) {
      const int n = net.dims[l] * net.dims[l + 1];
      wm_transpose_kernel<<<(n + 255) / 256, 256, 0, s>>>(t->params + net.offW[l], t->wt + net.offW[l], net.dims[l],
                                                         net.dims[l + 1]);
      ++L.n;
    }
    cudaEvent_t tb_done[8] = {};
    if (with_term) {
      term_backward(C - 1);              // the top chunk's rows first, on the main stream: the backward scan starts there
      if (C > 1) {
        after(L.w, s);                   // (the side stream has seen the loss kernel through this)
        L.cur = L.w;
        for (int ch = C - 2; ch >= 0; --ch) {
          term_backward(ch);
          tb_done[ch] = L.ev();
          cudaEventRecord(tb_done[ch], L.w);
        }
        L.cur = s;
      }
    }
    for (int ch = C - 1; ch >= 0; --ch) {
      if (with_term && ch < C - 1) cudaStreamWaitEvent(s, tb_done[ch], 0);
      sa.t0 = t_lo(ch); sa.t1 = t_hi(ch);
      if (sa.t1 > sa.t0) {
        wm_scan_bwd_kernel<<<(B + kScanRB - 1) / kScanRB, kScanThreads, t->scan_smem, s>>>(sa);
        ++L.n;
      }
      if (C > 1) { after(L.c, s); L.cur = L.c; }
      weight_grads(ch);
      L.cur = s;
    }
    if (C > 1) L.cur = L.c;
  } else {
    if (with_term) term_backward(0);
    // back-propagation through time: input-gradient chain of step ti, then LayerNorm backward + residual into step ti-1
    for (int ti = T - 1; ti >= 0; --ti) {
      const float* g = t->gx[nl] + (long long)ti * B * O;
      for (int l = nl - 1; l >= 0; --l) {
        const int din = net.dims[l], dout = net.dims[l + 1];
        float* out = t->gx[l] + (long long)ti * B * din;
        L.gemm(false, true, g, dout, t->params + net.offW[l], dout, out, din, B, din, dout, l > 0 ? WE_MASK : WE_NONE, nullptr,
               l > 0 ? t->x[l] + (long long)ti * B * din : nullptr, din, 0);
        g = out;
      }
      if (ti > 0) {
        wm_ln_bwd_kernel<<<(B * 32 + 255) / 256, 256, 0, s>>>(g, t->xhat + (long long)ti * B * D0, t->rstd + ti * B, scale, O,
                                                             D0, B, t->gx[nl] + (long long)ti * B * O,
                                                             t->gx[nl] + (long long)(ti - 1) * B * O);
        ++L.n;
      }
    }
    weight_grads(0);
  }
  // the slices of every buffer, added in slice order
  for (int l = nl - 1; l >= 0; --l) {
    const int din = net.dims[l], dout = net.dims[l + 1];
    L.reduce(t->wparts[l], kWmSplits, (long long)din * dout, t->grads + net.offW[l]);
    L.reduce(t->cparts[l], kWmSplits, dout, t->grads + net.offB[l]);
  }
  L.reduce(t->cparts[nl], kWmSplits, D0, t->grads + net.off_ln_bias);
  L.reduce(t->cparts[nl] + (long long)kWmSplits * D0, kWmSplits, D0, t->grads + net.off_ln_scale);
  L.cur = s;
  if (scan && C > 1) {
    after(s, L.c);
    if (with_term) after(s, L.w);   // (already ordered through tb_done; joins the side stream into the capture's end)
  }
  wm_adam_kernel<<<(int)((net.n_params + 255) / 256), 256, 0, s>>>(t->params, t->m, t->v, t->grads, net.n_params, t->step,
                                                                  c.lr, c.decay_steps, t->logs);
  ++L.n;
}

int wm_run(WmTrainer* t, int train, int gather, int repeats, const float* obs, const float* act, const float* next_obs,
           const float* rew, const float* mask, cudaStream_t stream, std::string* err) {
  WM_CUDA(cudaSetDevice(t->device));
  WmTrainer::Key key{train + 2 * gather, {obs, act, next_obs, rew, mask}};
  WmTrainer::Entry* e = nullptr;
  for (auto& g : t->graphs)
    if (g.key == key) e = &g;
  if (!e) {
    if (t->graphs.size() >= 8) {
      cudaGraphExecDestroy(t->graphs.front().exec);
      t->graphs.erase(t->graphs.begin());
    }
    cudaStream_t cs;
    WM_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    WM_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    cudaStream_t sw, sc;
    WM_CUDA(cudaStreamCreateWithFlags(&sw, cudaStreamNonBlocking));
    WM_CUDA(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking));
    Launcher L{cs, sw, sc, cs};
    if (t->cfg.kind == WM_MULTISTEP) wm_issue_multistep(t, L, train, gather, obs, act, next_obs, rew);
    else wm_issue(t, L, train, gather, obs, act, next_obs, rew, mask);
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(cs, &graph);
    cudaError_t le = cudaGetLastError();
    for (cudaEvent_t e : L.evs) cudaEventDestroy(e);
    cudaStreamDestroy(sw);
    cudaStreamDestroy(sc);
    if (ce != cudaSuccess || le != cudaSuccess) {
      if (err) *err = std::string("world-model step capture: ") + cudaGetErrorString(ce != cudaSuccess ? ce : le);
      cudaStreamDestroy(cs);
      return 1;
    }
    cudaGraphExec_t exec;
    WM_CUDA(cudaGraphInstantiate(&exec, graph, 0));
    cudaGraphDestroy(graph);
    cudaStreamDestroy(cs);
    t->graphs.push_back({key, exec, L.n});
    e = &t->graphs.back();
  }
  for (int i = 0; i < repeats; ++i) WM_CUDA(cudaGraphLaunch(e->exec, stream));
  t->launches += (long long)repeats * e->launches;
  return 0;
}
}  // namespace

int wm_create_impl(const WmConfig& cfg, const float* params_host, const float* term_params_host, WmTrainer** out,
                   std::string* err) {
  auto bad = [&](const char* m) { if (err) *err = m; return 1; };
  if (cfg.kind != WM_STATE && cfg.kind != WM_TERMINATION && cfg.kind != WM_MULTISTEP)
    return bad("world model: kind must be 0 (state), 1 (termination) or 2 (multistep)");
  if (cfg.O < 1 || cfg.A < 0 || cfg.batch < 1) return bad("world model: bad dimensions");
  if (cfg.n_hidden < 1 || cfg.n_hidden >= kWmMaxLayers) return bad("world model: between 1 and 7 hidden layers");
  if (cfg.decay_steps < 1) return bad("world model: decay_steps must be positive");
  if (!params_host) return bad("world model: initial parameters are required");
  const bool multi = cfg.kind == WM_MULTISTEP;
  if (multi && (cfg.seq_len < 1 || (long long)cfg.seq_len * cfg.batch > (1 << 24)))
    return bad("world model: multistep needs 1 <= seq_len and batch * seq_len <= 2^24");
  const bool state = cfg.kind != WM_TERMINATION;
  const bool with_term = state && cfg.termination_weight > 0.f;
  if (with_term && (!term_params_host || cfg.term_n_hidden < 1 || cfg.term_n_hidden >= kWmMaxLayers))
    return bad("world model: termination_weight > 0 needs the trained termination predictor (parameters + hidden dims)");
  WmTrainer* t = new WmTrainer();
  t->cfg = cfg;
  if (!multi) t->cfg.seq_len = 1;
  cudaGetDevice(&t->device);
  t->net = state ? wm_layout(cfg.O + cfg.A, cfg.n_hidden, cfg.hidden, cfg.O, true)
                 : wm_layout(cfg.O, cfg.n_hidden, cfg.hidden, 1, false);
  const long long n = t->net.n_params;
  const int B = cfg.batch, nl = t->net.n_layers;
  const long long R = (long long)B * t->cfg.seq_len;  // rows every per-row buffer holds (time-major slabs for multistep)
  t->rows = (int)R;
  int maxdim = 1;
  for (int l = 0; l <= nl; ++l) maxdim = std::max(maxdim, t->net.dims[l]);
  int rc = 0;
  rc |= wm_alloc(t, &t->params, n, err) || wm_alloc(t, &t->m, n, err) || wm_alloc(t, &t->v, n, err) ||
        wm_alloc(t, &t->grads, n, err);
  for (int l = 0; l <= nl && !rc; ++l) {
    if (!(multi && l == nl)) rc |= wm_alloc(t, &t->x[l], R * t->net.dims[l], err);  // multistep: predictions live in ocur
    rc |= wm_alloc(t, &t->gx[l], R * t->net.dims[l], err);
  }
  if (!rc && state) rc |= wm_alloc(t, &t->xhat, R * t->net.dims[0], err) || wm_alloc(t, &t->rstd, R, err);
  if (!rc && multi) {
    rc |= wm_alloc(t, &t->ocur, (R + B) * cfg.O, err) || wm_alloc(t, &t->carry, (long long)B * cfg.O, err);
    for (int l = 0; l < nl && !rc; ++l)
      rc |= wm_alloc(t, &t->wparts[l], (long long)kWmSplits * t->net.dims[l] * t->net.dims[l + 1], err) ||
            wm_alloc(t, &t->cparts[l], (long long)kWmSplits * t->net.dims[l + 1], err);
    if (!rc) rc |= wm_alloc(t, &t->cparts[nl], 2LL * kWmSplits * t->net.dims[0], err);
    t->x[nl] = t->ocur ? t->ocur + (long long)B * cfg.O : nullptr;
    // persistent scan kernels: widths up to kScanMaxDim; the largest kernels that fit stay in shared memory
    if (!rc && maxdim <= kScanMaxDim && t->net.dims[0] <= kScanMaxIn) {
      rc |= wm_alloc(t, &t->wt, n, err);
      t->scan_maxdim = scan_round32(maxdim);
      const int base = scan_base_floats(t->scan_maxdim, cfg.O);
      int budget = (220 * 1024) / 4 - base;   // floats left of the opt-in shared memory of one SM
      // a cached kernel is padded to a multiple of 4 columns in BOTH orientations (the backward kernel holds its transpose)
      auto padded = [&](int l) {
        const int K = t->net.dims[l], N = t->net.dims[l + 1];
        return std::max(K * scan_round4(N), N * scan_round4(K));
      };
      bool used[kWmMaxLayers] = {};
      for (int pick = 0; pick < nl; ++pick) {
        int best = -1;
        for (int l = 0; l < nl; ++l)
          if (!used[l] && padded(l) <= budget && (best < 0 || padded(l) > padded(best))) best = l;
        if (best < 0) break;
        used[best] = true;
        t->scan_mask |= 1u << best;
        budget -= padded(best);
      }
      t->scan_smem = ((220 * 1024) / 4 - budget) * 4;
      if (!rc && (cudaFuncSetAttribute(wm_scan_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t->scan_smem) != cudaSuccess ||
                  cudaFuncSetAttribute(wm_scan_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, t->scan_smem) != cudaSuccess))
        rc = bad("world model: raising the scan kernels' shared-memory limit failed");
      t->scan_ok = !rc;
    }
  }
  if (!rc && with_term) {
    t->term = wm_layout(cfg.O, cfg.term_n_hidden, cfg.term_hidden, 1, false);
    for (int l = 0; l <= t->term.n_layers; ++l) maxdim = std::max(maxdim, t->term.dims[l]);
    rc |= wm_alloc(t, &t->term_params, t->term.n_params, err);
    t->tx[0] = t->x[nl];
    for (int l = 1; l <= t->term.n_layers && !rc; ++l) rc |= wm_alloc(t, &t->tx[l], R * t->term.dims[l], err);
    if (!rc && cudaMemcpy(t->term_params, term_params_host, sizeof(float) * t->term.n_params, cudaMemcpyHostToDevice) != cudaSuccess)
      rc = bad("world model: copying the termination predictor failed");
  }
  if (!rc)
    rc |= wm_alloc(t, &t->pred, (long long)B * cfg.O, err) || wm_alloc(t, &t->dbuf[0], R * maxdim, err) ||
          wm_alloc(t, &t->dbuf[1], R * maxdim, err) || wm_alloc(t, &t->dlogit, R, err) ||
          wm_alloc(t, &t->logs, kWmLogs, err) || wm_alloc(t, (float**)&t->step, 2, err) ||
          wm_alloc(t, &t->loss_part, kWmLossCtas * 5, err) || wm_alloc(t, (float**)&t->ticket, 1, err) ||
          wm_alloc(t, &t->b_obs, R * cfg.O, err) || wm_alloc(t, &t->b_act, R * std::max(cfg.A, 1), err) ||
          wm_alloc(t, &t->b_next, R * cfg.O, err) || wm_alloc(t, &t->b_rew, R, err);
  if (!rc && cudaMemcpy(t->params, params_host, sizeof(float) * n, cudaMemcpyHostToDevice) != cudaSuccess)
    rc = bad("world model: copying the initial parameters failed");
  if (rc) {
    wm_destroy_impl(t);
    return 1;
  }
  *out = t;
  return 0;
}

void wm_destroy_impl(WmTrainer* t) {
  if (!t) return;
  cudaSetDevice(t->device);
  cudaDeviceSynchronize();
  for (auto& g : t->graphs) cudaGraphExecDestroy(g.exec);
  for (void* p : t->owned) cudaFree(p);
  delete t;
}

int wm_step_impl(WmTrainer* t, int train, const float* obs, const float* act, const float* next_obs, const float* rewards,
                 const float* dropout_mask, cudaStream_t stream, std::string* err) {
  if (!next_obs || !rewards || (t->cfg.kind != WM_TERMINATION && (!obs || !act))) {
    if (err) *err = "world-model step: null batch pointer";
    return 1;
  }
  return wm_run(t, train ? 1 : 0, 0, 1, obs, act, next_obs, rewards, train ? dropout_mask : nullptr, stream, err);
}

int wm_set_dataset_impl(WmTrainer* t, const float* obs, const float* act, const float* next_obs, const float* rewards,
                        long long n_rows, std::string* err) {
  if (!obs || !act || !next_obs || !rewards || n_rows < 1 || n_rows > 0xffffffffLL) {
    if (err) *err = "world-model dataset: null pointer or row count outside [1, 2^32)";
    return 1;
  }
  t->ds_obs = obs; t->ds_act = act; t->ds_next = next_obs; t->ds_rew = rewards; t->ds_rows = n_rows;
  for (auto& g : t->graphs) cudaGraphExecDestroy(g.exec);  // captured gathers hold the old pointers
  t->graphs.clear();
  return 0;
}

int wm_train_impl(WmTrainer* t, int n_steps, const float* dropout_mask, cudaStream_t stream, std::string* err) {
  if (t->cfg.kind == WM_MULTISTEP && t->ds_rows &&
      (t->ds_rows % kWmEpisodeLen != 0 || t->cfg.seq_len >= kWmEpisodeLen)) {
    if (err) *err = "world-model train: multistep windows need a dataset of whole 1000-step episodes and seq_len < 1000";
    return 1;
  }
  if (!t->ds_rows) {
    if (err) *err = "world-model train: no resident dataset (fqlpop_wm_set_dataset)";
    return 1;
  }
  if (n_steps < 0) n_steps = 0;
  return wm_run(t, 1, 1, n_steps, t->ds_obs, t->ds_act, t->ds_next, t->ds_rew, dropout_mask, stream, err);
}

int wm_logs_impl(WmTrainer* t, float* out_host, cudaStream_t stream, std::string* err) {
  WM_CUDA(cudaSetDevice(t->device));
  WM_CUDA(cudaMemcpyAsync(out_host, t->logs, sizeof(float) * kWmLogs, cudaMemcpyDeviceToHost, stream));
  WM_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

int wm_export_impl(WmTrainer* t, int which, float* out_host, long long n, cudaStream_t stream, std::string* err) {
  if (which < 0 || which > 3 || n != t->net.n_params) {
    if (err) *err = "world-model export: which must be 0..3 and n the parameter count";
    return 1;
  }
  const float* src = which == 0 ? t->params : which == 1 ? t->m : which == 2 ? t->v : t->grads;
  WM_CUDA(cudaSetDevice(t->device));
  WM_CUDA(cudaMemcpyAsync(out_host, src, sizeof(float) * n, cudaMemcpyDeviceToHost, stream));
  WM_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

}  // namespace fqlpop

// ---------------------------------------------------------------------------------------------------------------------
// initial-observation generator (VAE decoder)
// ---------------------------------------------------------------------------------------------------------------------
namespace fqlpop {

__global__ void wm_normal_kernel(unsigned long long seed, int rows, int cols, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (long long)rows * cols) out[i] = philox_normal(seed, 0ull, (uint32_t)(i / cols), (uint32_t)(i % cols));
}

struct WmDecoder {
  WmNet net;
  int device = 0, cap = 0;
  float* params = nullptr;
  float* act[kWmMaxLayers + 1] = {};  // act[0] = z ... act[n_layers - 1]; the last layer writes straight into `out`
};

int wm_decoder_create_impl(int latent_dim, int n_hidden, const int* hidden, int out_dim, const float* params_host,
                           WmDecoder** out, std::string* err) {
  if (latent_dim < 1 || out_dim < 1 || n_hidden < 1 || n_hidden >= kWmMaxLayers || !params_host) {
    if (err) *err = "decoder: bad dimensions, layer count outside [1, 7] or null parameters";
    return 1;
  }
  WmDecoder* d = new WmDecoder();
  d->net = wm_layout(latent_dim, n_hidden, hidden, out_dim, false);
  cudaGetDevice(&d->device);
  if (cudaMalloc((void**)&d->params, sizeof(float) * d->net.n_params) != cudaSuccess ||
      cudaMemcpy(d->params, params_host, sizeof(float) * d->net.n_params, cudaMemcpyHostToDevice) != cudaSuccess) {
    if (err) *err = "decoder: allocating / copying the parameters failed";
    wm_decoder_destroy_impl(d);
    return 1;
  }
  *out = d;
  return 0;
}

void wm_decoder_destroy_impl(WmDecoder* d) {
  if (!d) return;
  cudaSetDevice(d->device);
  cudaDeviceSynchronize();
  cudaFree(d->params);
  for (float* a : d->act) cudaFree(a);
  delete d;
}

int wm_decoder_sample_impl(WmDecoder* d, int n, const float* noise, unsigned long long seed, float* out,
                           cudaStream_t stream, std::string* err) {
  if (n < 1 || !out) {
    if (err) *err = "decoder sample: n must be positive and out non-null";
    return 1;
  }
  WM_CUDA(cudaSetDevice(d->device));
  const WmNet& net = d->net;
  if (n > d->cap) {  // grow the activation buffers (stream-ordered with respect to earlier samples: sync first)
    WM_CUDA(cudaDeviceSynchronize());
    for (int l = 0; l < net.n_layers; ++l) {
      cudaFree(d->act[l]);
      d->act[l] = nullptr;
      WM_CUDA(cudaMalloc((void**)&d->act[l], sizeof(float) * (size_t)n * net.dims[l]));
    }
    d->cap = n;
  }
  float* act[kWmMaxLayers + 1];
  for (int l = 0; l < net.n_layers; ++l) act[l] = d->act[l];
  act[net.n_layers] = out;
  if (noise) {
    WM_CUDA(cudaMemcpyAsync(act[0], noise, sizeof(float) * (size_t)n * net.dims[0], cudaMemcpyDeviceToDevice, stream));
  } else {
    const long long tot = (long long)n * net.dims[0];
    wm_normal_kernel<<<(int)((tot + 255) / 256), 256, 0, stream>>>(seed, n, net.dims[0], act[0]);
  }
  Launcher L{stream, stream, stream, stream};
  wm_forward(L, net, d->params, act, n, nullptr, 0);
  WM_CUDA(cudaGetLastError());
  return 0;
}

#include "wm_vae.cuh"

}  // namespace fqlpop
