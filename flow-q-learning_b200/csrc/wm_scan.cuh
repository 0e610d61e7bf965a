// Persistent scan kernels of the multistep world-model trainer (included by wm.cu).
//
// Reference: envmodel/multistep.py:31-54 (nn.scan of the baseline cell over the window, each prediction fed back as the
// next observation) and its reverse-mode derivative. The batch rows of a window are independent, only time is
// sequential: one CTA owns kScanRB batch rows and walks all T steps of them -- forward: LayerNorm(concat(o_t, a_t)) ->
// Dense / relu stack -> o_{t+1} = f + o_t; backward: the input-gradient chain of step t, the LayerNorm backward and the
// residual carry into step t - 1 -- without returning to the launch queue (the per-step version is 2 x T x 5 dependent
// small launches, launch-latency-bound). Everything the rest of the step needs is written where the per-step kernels
// wrote it (time-major slabs x[l], xhat, rstd, predictions, gx[l]), so the loss, the frozen termination predictor, the
// split-K weight gradients and Adam are unchanged. fp32 SIMT like the reference.
//
// A layer of the stack for kScanRB = 4 rows is a skinny product. Thread = 4 rows x 4 adjacent output columns (16
// accumulators) over one slice of the reduction: per reduction index one 16-byte load of the rows' inputs (shared memory,
// [k][row], broadcast) and one of the kernel ([k][column]) feed 16 FMAs. The slices' partial sums meet in shared memory
// and the epilogue (bias, relu / residual, stores) adds them. Kernel matrices are read from shared memory for the layers
// that fit (chosen by the host, largest first; rows padded to a multiple of 4 columns), otherwise from L2 through the
// read-only path (the whole network is 310 KB of fp32: L2-resident). The backward kernel reads TRANSPOSED kernels
// ([out][in], rebuilt once per optimiser step by wm_transpose_kernel) so that its products have the same access pattern.
//
// Measured (T = 256, B = 256, antsoccer shape, 64 CTAs): 6.9 ms per optimiser step against 14.6 ms for the per-step
// launches (5.5 ms with termination_weight = 0, the reference scripts' setting); 6.0 / 5.3 ms with wm_gemm64_kernel;
// 5.6 / 5.3 ms with the window cut into 4 time chunks (wm.cu: the products over a chunk's rows overlap the next chunk's
// scan); 5.4 / 5.1 ms with the LayerNorm scale / bias of a lane's inputs read once instead of on every step and the
// slices' partial sums fetched four loads at a time. The per-step critical path is a chain of
// 9 short phases whose latencies (barrier, shared-memory round trips, L2 for the one kernel that does not fit) are paid in
// full at 2 warps per scheduler: issue slots 24 % busy. Tried and dropped: 512 / 1024 threads with finer reduction
// slices (7.4 / 9.0 ms: more partial sums to add, same latencies); clusters of 2 or 4 CTAs that split every layer's
// columns, keep the whole network shared-memory resident and broadcast their outputs with st.shared::cluster + one
// barrier.cluster per layer (products 2x faster, but the cluster barrier and the remote stores cost ~1.5 k cycles per
// layer: 8.0 ms); rotating the epilogues' four [column][row] activation stores per lane pair so that a warp's store
// spreads over 8 banks instead of 2 (the bank-conflict replays ncu shows): the selects cost more than the replays (6.1 ms).
#pragma once

constexpr int kScanRB = 4;         // batch rows per CTA
constexpr int kScanThreads = 256;
constexpr int kScanMaxDim = 512;   // widest layer the scan kernels take (wider networks keep the per-step path)
constexpr int kScanMaxSlices = 8;  // reduction slices per product
constexpr int kScanMaxIn = 128;    // widest network input (ob_dim + action_dim): 4 LayerNorm slots per lane
constexpr int kScanLnPer = kScanMaxIn / 32;
constexpr int kScanPartFloats = kScanThreads * 16;  // every thread parks its 4 x 4 partial sums

struct ScanArgs {
  WmNet net;
  const float* params;   // flat parameter vector (LayerNorm, kernels [in][out], biases)
  const float* wt;       // backward: transposed kernels [out][in] at the same offW offsets
  float* x[kWmMaxLayers + 1];    // [T][B][dims[l]]; x[n_layers] = predictions o_1 .. o_T
  float* gx[kWmMaxLayers + 1];   // d loss / d x[l], same shapes; gx[n_layers] arrives holding the loss gradient
  float* ocur;           // [T + 1][B][O]: slab 0 = first observations of the windows
  float* xhat;           // [T][B][O + A]
  float* rstd;           // [T][B]
  const float* act;      // [B][T][A]
  int B, T, O, A;
  int maxdim;            // widest layer, rounded up to 32
  float eps;
  unsigned cache_mask;   // bit l: kernel l is copied into shared memory at kernel start
  // One launch walks the time steps [t0, t1) (forward: upwards from o_{t0} = ocur slab t0; backward: downwards, the
  // residual carry of step t1 arriving in / the one for step t0 - 1 leaving through carry_g [B][O]): the host cuts the
  // window into chunks so that the work over a chunk's rows (termination predictor, weight gradients) overlaps the
  // next chunk's scan on the SMs the scan leaves idle.
  int t0, t1;
  float* carry_g;
};

__host__ __device__ inline int scan_round4(int n) { return (n + 3) & ~3; }
__host__ __device__ inline int scan_round32(int n) { return (n + 31) & ~31; }
// reduction slices of a product with N <= kScanMaxDim output columns: as many as the CTA has threads for, at most
// kScanMaxSlices (slices x column groups <= kScanThreads, so slices x RB x round4(N) <= kScanPartFloats)
__host__ __device__ inline int scan_slices(int N) {
  const int g = kScanThreads / (scan_round4(N) / 4);
  return g > kScanMaxSlices ? kScanMaxSlices : (g < 1 ? 1 : g);
}
// shared-memory carve-up (floats): two activation buffers [maxdim][RB], the partial sums, one [RB][O] row state
__host__ __device__ inline int scan_base_floats(int maxdim, int O) {
  return 2 * maxdim * kScanRB + kScanPartFloats + scan_round4(kScanRB * O);
}

__device__ __forceinline__ float4 scan_lds4(uint32_t a) {
  float4 v;
  // volatile + memory clobber: shared memory changes under these loads (other threads' stores, barriers)
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ float4 scan_ldg4(const float* p) {
  float4 v;
  asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t scan_saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// i / d for 0 <= i < 2^15, 1 <= d <= 2^10: exact through a float reciprocal (a non-constant integer division is a
// ~100-cycle dependent chain, and these kernels pay every such chain in full: two warps per scheduler)
__device__ __forceinline__ int scan_div(int i, int d) { return __float2int_rd(((float)i + 0.5f) * __frcp_rn((float)d)); }

#define SCAN_FMA16(acc, xv, wv)                                                                     \
  acc[0][0] = fmaf(xv.x, wv.x, acc[0][0]); acc[0][1] = fmaf(xv.x, wv.y, acc[0][1]);                 \
  acc[0][2] = fmaf(xv.x, wv.z, acc[0][2]); acc[0][3] = fmaf(xv.x, wv.w, acc[0][3]);                 \
  acc[1][0] = fmaf(xv.y, wv.x, acc[1][0]); acc[1][1] = fmaf(xv.y, wv.y, acc[1][1]);                 \
  acc[1][2] = fmaf(xv.y, wv.z, acc[1][2]); acc[1][3] = fmaf(xv.y, wv.w, acc[1][3]);                 \
  acc[2][0] = fmaf(xv.z, wv.x, acc[2][0]); acc[2][1] = fmaf(xv.z, wv.y, acc[2][1]);                 \
  acc[2][2] = fmaf(xv.z, wv.z, acc[2][2]); acc[2][3] = fmaf(xv.z, wv.w, acc[2][3]);                 \
  acc[3][0] = fmaf(xv.w, wv.x, acc[3][0]); acc[3][1] = fmaf(xv.w, wv.y, acc[3][1]);                 \
  acc[3][2] = fmaf(xv.w, wv.z, acc[3][2]); acc[3][3] = fmaf(xv.w, wv.w, acc[3][3]);

#ifdef FQLPOP_EPI_TRACE
__device__ long long g_scan_prod[4];   // thread 0 of CTA 0: index math, reduction loop, partial stores, barrier
#define SCAN_PT(i) do { if (blockIdx.x == 0 && tid == 0) { const long long n_ = clock64(); g_scan_prod[i] += n_ - pt_; pt_ = n_; } } while (0)
#else
#define SCAN_PT(i)
#endif
// One layer's partial products. W: [K][ldw] (ldw = N rounded up to 4 in shared memory, N in global memory), xs: [K][RB].
// part[(g * RB + r) * Np + c] (Np = round4(N)) = slice g of sum_k xs[k][r] * W[k][c]. Ends with __syncthreads().
__device__ __forceinline__ void scan_product(const float* __restrict__ W, bool cached, int K, int N, const float* xs,
                                             float* part, int tid) {
  static_assert(kScanRB == 4, "the rows' inputs are read as one float4 per reduction index");
  const int Np = scan_round4(N), CG = Np >> 2, G = scan_slices(N);
  const uint32_t xa = scan_saddr(xs), pa = scan_saddr(part);
#ifdef FQLPOP_EPI_TRACE
  long long pt_ = clock64();
#endif
  if (tid < CG * G) {
    const int g = scan_div(tid, CG), cg = tid - g * CG;
    const int Kc = scan_div(K + G - 1, G), k0 = g * Kc, k1 = min(K, k0 + Kc);
    float acc[4][4] = {};
    SCAN_PT(0);
    // Software-pipelined reduction loop: the operands of the NEXT batch of 4 (8 for a kernel in global memory) reduction
    // indices are requested before the FMAs of the current one are issued (the loads are volatile asm, so they stay where
    // they are written). With two warps per scheduler nothing else hides the load latency.
    const int nk = max(k1 - k0, 0);   // the last slices of a short reduction are empty
    uint32_t xk = xa + (uint32_t)k0 * 16u;
    if (cached) {
      uint32_t wa = scan_saddr(W) + (uint32_t)(k0 * Np + 4 * cg) * 4u;
      const uint32_t wstep = (uint32_t)Np * 4u;
      float4 wA[4], xA[4], wB[4], xB[4];
#define SCAN_LD4(wv, xv)                                                                           \
  _Pragma("unroll") for (int i = 0; i < 4; ++i) {                                                  \
    wv[i] = scan_lds4(wa + (uint32_t)i * wstep);                                                   \
    xv[i] = scan_lds4(xk + (uint32_t)i * 16u);                                                     \
  }                                                                                                \
  wa += 4u * wstep;                                                                                \
  xk += 64u;
#define SCAN_FM4(wv, xv) _Pragma("unroll") for (int i = 0; i < 4; ++i) { SCAN_FMA16(acc, xv[i], wv[i]) }
      const int nb = nk >> 2;
      if (nb > 0) { SCAN_LD4(wA, xA) }
      for (int b = 0; b < nb; b += 2) {
        if (b + 1 < nb) { SCAN_LD4(wB, xB) }
        SCAN_FM4(wA, xA)
        if (b + 2 < nb) { SCAN_LD4(wA, xA) }
        if (b + 1 < nb) { SCAN_FM4(wB, xB) }
      }
      for (int k = nb << 2; k < nk; ++k) {
        const float4 wv = scan_lds4(wa), xv = scan_lds4(xk);
        wa += wstep;
        xk += 16u;
        SCAN_FMA16(acc, xv, wv)
      }
#undef SCAN_LD4
#undef SCAN_FM4
    } else if ((N & 3) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0) {
      const float* wp = W + (long long)k0 * N + 4 * cg;
      const long long wstep = N;
      float4 wA[8], wB[8];
#define SCAN_LDG8(wv) _Pragma("unroll") for (int i = 0; i < 8; ++i) wv[i] = scan_ldg4(wp + (long long)i * wstep); wp += 8LL * wstep;
#define SCAN_FM8(wv)                                                                               \
  _Pragma("unroll") for (int i = 0; i < 8; ++i) {                                                  \
    const float4 xv = scan_lds4(xk + (uint32_t)i * 16u);                                           \
    SCAN_FMA16(acc, xv, wv[i])                                                                     \
  }                                                                                                \
  xk += 128u;
      const int nb = nk >> 3;
      if (nb > 0) { SCAN_LDG8(wA) }
      for (int b = 0; b < nb; b += 2) {
        if (b + 1 < nb) { SCAN_LDG8(wB) }
        SCAN_FM8(wA)
        if (b + 2 < nb) { SCAN_LDG8(wA) }
        if (b + 1 < nb) { SCAN_FM8(wB) }
      }
      for (int k = nb << 3; k < nk; ++k) {
        const float4 wv = scan_ldg4(wp), xv = scan_lds4(xk);
        wp += wstep;
        xk += 16u;
        SCAN_FMA16(acc, xv, wv)
      }
#undef SCAN_LDG8
#undef SCAN_FM8
    } else {  // kernel in global memory whose rows are not 16-byte aligned
      const float* wp = W + (long long)k0 * N + 4 * cg;
      const int c0 = 4 * cg;
      for (int k = k0; k < k1; ++k) {
        float4 wv;
        wv.x = __ldg(wp);
        wv.y = c0 + 1 < N ? __ldg(wp + 1) : 0.f;
        wv.z = c0 + 2 < N ? __ldg(wp + 2) : 0.f;
        wv.w = c0 + 3 < N ? __ldg(wp + 3) : 0.f;
        const float4 xv = scan_lds4(xk);
        wp += N;
        xk += 16u;
        SCAN_FMA16(acc, xv, wv)
      }
    }
    SCAN_PT(1);
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const uint32_t a = pa + (uint32_t)((g * kScanRB + r) * Np + 4 * cg) * 4u;
      asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(acc[r][0]), "f"(acc[r][1]), "f"(acc[r][2]),
                   "f"(acc[r][3])
                   : "memory");
    }
    SCAN_PT(2);
  }
  __syncthreads();
  SCAN_PT(3);
}
// sum over the slices of 4 adjacent outputs of row r
__device__ __forceinline__ float4 scan_gather4(const float* part, int N, int r, int cg) {
  const int Np = scan_round4(N), G = scan_slices(N);
  const uint32_t a = scan_saddr(part) + (uint32_t)(r * Np + 4 * cg) * 4u;
  // the slices' loads are issued back to back (four at a time), then added in slice order
  float4 v = scan_lds4(a);
  const uint32_t step = (uint32_t)(kScanRB * Np) * 4u;
  int g = 1;
  for (; g + 3 < G; g += 4) {
    const float4 p0 = scan_lds4(a + (uint32_t)g * step), p1 = scan_lds4(a + (uint32_t)(g + 1) * step);
    const float4 p2 = scan_lds4(a + (uint32_t)(g + 2) * step), p3 = scan_lds4(a + (uint32_t)(g + 3) * step);
    v.x += p0.x; v.y += p0.y; v.z += p0.z; v.w += p0.w;
    v.x += p1.x; v.y += p1.y; v.z += p1.z; v.w += p1.w;
    v.x += p2.x; v.y += p2.y; v.z += p2.z; v.w += p2.w;
    v.x += p3.x; v.y += p3.y; v.z += p3.z; v.w += p3.w;
  }
  for (; g < G; ++g) {
    const float4 p = scan_lds4(a + (uint32_t)g * step);
    v.x += p.x; v.y += p.y; v.z += p.z; v.w += p.w;
  }
  return v;
}

#ifdef FQLPOP_EPI_TRACE
// (make TRACE=1) thread 0 of CTA 0 accumulates the cycles of every phase over the T steps and prints them at the end
#define SCAN_TR_DECL long long tr_ph[2 + 2 * kWmMaxLayers] = {}; long long tr_t = clock64()
#define SCAN_TR(i) do { const long long n_ = clock64(); tr_ph[i] += n_ - tr_t; tr_t = n_; } while (0)
#define SCAN_TR_PRINT(name, nl) do { if (blockIdx.x == 0 && threadIdx.x == 0) { printf("%s cycles per step: head %lld tail %lld |", name, tr_ph[0] / T, tr_ph[1] / T); \
    for (int l_ = 0; l_ < nl; ++l_) printf(" L%d product %lld epilogue %lld |", l_, tr_ph[2 + 2 * l_] / T, tr_ph[3 + 2 * l_] / T); printf("\n"); } } while (0)
#else
#define SCAN_TR_DECL
#define SCAN_TR(i)
#define SCAN_TR_PRINT(name, nl)
#endif

struct ScanSmem {
  float *xa, *xb, *part, *rowst, *cache;
};
__device__ __forceinline__ ScanSmem scan_carve(float* sm, int maxdim, int O) {
  ScanSmem s;
  s.xa = sm;
  s.xb = s.xa + maxdim * kScanRB;
  s.part = s.xb + maxdim * kScanRB;
  s.rowst = s.part + kScanPartFloats;
  s.cache = s.rowst + scan_round4(kScanRB * O);
  return s;
}
// kernels of the layers in cache_mask -> shared memory, rows padded with zeros to a multiple of 4 columns
// (src holds [K][N] with K = kdim(l), N = ndim(l)); w[l] = where layer l's kernel is read from
template <bool TRANSPOSED>
__device__ __forceinline__ void scan_load_weights(const ScanArgs& a, const float* src, float* cache, const float* (&w)[kWmMaxLayers]) {
  float* cp = cache;
  for (int l = 0; l < a.net.n_layers; ++l) {
    const float* g = src + a.net.offW[l];
    const int K = TRANSPOSED ? a.net.dims[l + 1] : a.net.dims[l], N = TRANSPOSED ? a.net.dims[l] : a.net.dims[l + 1];
    const int Np = scan_round4(N);
    if (a.cache_mask >> l & 1u) {
      for (int i = threadIdx.x; i < K * Np; i += kScanThreads) {
        const int k = i / Np, c = i - k * Np;
        cp[i] = c < N ? __ldg(g + (long long)k * N + c) : 0.f;
      }
      w[l] = cp;
      cp += K * Np;
    } else {
      w[l] = g;
    }
  }
}

__global__ void __launch_bounds__(kScanThreads) wm_scan_fwd_kernel(const __grid_constant__ ScanArgs a) {
  extern __shared__ __align__(16) float scan_sm[];
  const ScanSmem S = scan_carve(scan_sm, a.maxdim, a.O);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int B = a.B, T = a.T, O = a.O, A = a.A, D = O + A, nl = a.net.n_layers;
  const int row0 = blockIdx.x * kScanRB;
  const float* w[kWmMaxLayers];
  scan_load_weights<false>(a, a.params, S.cache, w);
  float* ob = S.rowst;  // [RB][O] current observation of the rows
  for (int i = tid; i < kScanRB * O; i += kScanThreads) {
    const int r = scan_div(i, O), row = row0 + r;
    ob[i] = row < B ? a.ocur[((long long)a.t0 * B + row) * O + (i - r * O)] : 0.f;
  }
  __syncthreads();
  const float* ln_scale = a.params + a.net.off_ln_scale;
  const float* ln_bias = a.params + a.net.off_ln_bias;
  // LayerNorm warps (one per row): lane j + 32 i holds input j + 32 i; the action part is fetched one step ahead
  const int ln_row = row0 + warp;
  const bool ln_ok = warp < kScanRB && ln_row < B;
  float a_pre[kScanLnPer];
  auto fetch_actions = [&](int ti) {
#pragma unroll
    for (int i = 0; i < kScanLnPer; ++i) {
      const int j = lane + 32 * i;
      a_pre[i] = (ln_ok && ti < T && j >= O && j < D) ? __ldg(a.act + ((long long)ln_row * T + ti) * A + (j - O)) : 0.f;
    }
  };
  fetch_actions(a.t0);
  // this lane's LayerNorm scale / bias: read once, not on every step's critical path
  float ln_s[kScanLnPer], ln_b[kScanLnPer];
#pragma unroll
  for (int i = 0; i < kScanLnPer; ++i) {
    const int j = lane + 32 * i;
    ln_s[i] = j < D ? __ldg(ln_scale + j) : 0.f;
    ln_b[i] = j < D ? __ldg(ln_bias + j) : 0.f;
  }
  SCAN_TR_DECL;
  for (int ti = a.t0; ti < a.t1; ++ti) {
    // ---- x0 = LayerNorm(concat(o_t, a_t)) * scale + bias: one warp per row (as wm_ln_concat_kernel) ----
    if (warp < kScanRB) {
      const int row = ln_row;
      const bool ok = ln_ok;
      float vin[kScanLnPer];
      float s = 0.f, s2 = 0.f;
#pragma unroll
      for (int i = 0; i < kScanLnPer; ++i) {
        const int j = lane + 32 * i;
        const float v = (ok && j < O) ? ob[warp * O + j] : a_pre[i];
        vin[i] = v;
        s += v;
        s2 += v * v;
      }
      fetch_actions(ti + 1);   // in flight during this step's products
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      }
      const float mean = s / D, var = fmaxf(s2 / D - mean * mean, 0.f), rstd = rsqrtf(var + a.eps);
      const long long base = ((long long)ti * B + row) * D;
#pragma unroll
      for (int i = 0; i < kScanLnPer; ++i) {
        const int j = lane + 32 * i;
        if (j < D) {
          const float h = (vin[i] - mean) * rstd;
          const float x0 = ok ? h * ln_s[i] + ln_b[i] : 0.f;
          if (ok) {
            a.xhat[base + j] = h;
            a.x[0][base + j] = x0;
          }
          S.xa[j * kScanRB + warp] = x0;
        }
      }
      if (ok && lane == 0) a.rstd[(long long)ti * B + row] = rstd;
    }
    __syncthreads();
    SCAN_TR(0);
    float *cur = S.xa, *nxt = S.xb;
    for (int l = 0; l < nl; ++l) {
      const int K = a.net.dims[l], N = a.net.dims[l + 1], CG = scan_round4(N) >> 2;
      const bool last = l == nl - 1;
      // this thread's epilogue items (row r, 4 adjacent columns): the biases are fetched before the product
      constexpr int kEpiPer = (kScanRB * (kScanMaxDim / 4) + kScanThreads - 1) / kScanThreads;
      const float* bias = a.params + a.net.offB[l];
      const bool vec = (N & 3) == 0;
      float4 bv[kEpiPer];
#pragma unroll
      for (int e = 0; e < kEpiPer; ++e) {
        const int i = tid + e * kScanThreads;
        const int r = scan_div(i, CG), c0 = 4 * (i - r * CG);
        bv[e] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < kScanRB * CG) {
          bv[e].x = __ldg(bias + c0);
          if (c0 + 1 < N) bv[e].y = __ldg(bias + c0 + 1);
          if (c0 + 2 < N) bv[e].z = __ldg(bias + c0 + 2);
          if (c0 + 3 < N) bv[e].w = __ldg(bias + c0 + 3);
        }
      }
      scan_product(w[l], (a.cache_mask >> l & 1u) != 0, K, N, cur, S.part, tid);
      SCAN_TR(2 + 2 * l);
      float* out = a.x[l + 1] + (long long)ti * B * N;
#pragma unroll
      for (int e = 0; e < kEpiPer; ++e) {
        const int i = tid + e * kScanThreads;
        if (i >= kScanRB * CG) break;
        const int r = scan_div(i, CG), cg = i - r * CG, row = row0 + r, c0 = 4 * cg;
        const float4 p = scan_gather4(S.part, N, r, cg);
        float y[4] = {p.x + bv[e].x, p.y + bv[e].y, p.z + bv[e].z, p.w + bv[e].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (!last) y[q] = fmaxf(y[q], 0.f);
          else if (c0 + q < N) y[q] += ob[r * O + c0 + q];  // residual: o_{t+1} = f(o_t, a_t) + o_t (envmodel/baseline.py:37)
          if (row >= B) y[q] = 0.f;
        }
        if (row < B) {
          float* po = out + (long long)row * N + c0;
          if (vec) {
            *reinterpret_cast<float4*>(po) = make_float4(y[0], y[1], y[2], y[3]);
          } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              if (c0 + q < N) po[q] = y[q];
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (c0 + q >= N) break;
          if (last) ob[r * O + c0 + q] = y[q];
          else nxt[(c0 + q) * kScanRB + r] = y[q];
        }
      }
      __syncthreads();
      SCAN_TR(3 + 2 * l);
      float* t = cur; cur = nxt; nxt = t;
    }
  }
  SCAN_TR_PRINT("wm_scan_fwd", nl);
#ifdef FQLPOP_EPI_TRACE
  if (blockIdx.x == 0 && tid == 0)
    printf("products of thread 0, cycles per step (fwd + earlier kernels): index %lld loop %lld stores %lld barrier %lld\n",
           g_scan_prod[0] / T, g_scan_prod[1] / T, g_scan_prod[2] / T, g_scan_prod[3] / T);
#endif
}

__global__ void __launch_bounds__(kScanThreads) wm_scan_bwd_kernel(const __grid_constant__ ScanArgs a) {
  extern __shared__ __align__(16) float scan_sm[];
  const ScanSmem S = scan_carve(scan_sm, a.maxdim, a.O);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int B = a.B, T = a.T, O = a.O, D = a.O + a.A, nl = a.net.n_layers;
  const int row0 = blockIdx.x * kScanRB;
  const float* w[kWmMaxLayers];
  scan_load_weights<true>(a, a.wt, S.cache, w);
  float* carry = S.rowst;  // [RB][O]: what step t hands to the prediction of step t - 1
  for (int i = tid; i < kScanRB * O; i += kScanThreads) {
    const int r = scan_div(i, O), row = row0 + r;
    carry[i] = (a.t1 < T && row < B) ? a.carry_g[(long long)row * O + (i - r * O)] : 0.f;
  }
  __syncthreads();
  const float* ln_scale = a.params + a.net.off_ln_scale;
  // global operands are fetched ahead of the phase that needs them: the loss gradient of step t - 1 during step t, the
  // LayerNorm operands of step t at its start, the relu masks of a layer before its product
  constexpr int kHeadPer = (kScanRB * kScanMaxIn + kScanThreads - 1) / kScanThreads;
  constexpr int kMaskPer = (kScanRB * (kScanMaxDim / 4) + kScanThreads - 1) / kScanThreads;
  float g_pre[kHeadPer];
  auto fetch_loss_grad = [&](int ti) {
#pragma unroll
    for (int e = 0; e < kHeadPer; ++e) {
      const int i = tid + e * kScanThreads;
      const int r = scan_div(i, O), c = i - r * O, row = row0 + r;
      g_pre[e] = (ti >= a.t0 && i < kScanRB * O && row < B) ? a.gx[nl][((long long)ti * B + row) * O + c] : 0.f;
    }
  };
  fetch_loss_grad(a.t1 - 1);
  const int ln_row = row0 + warp;
  const bool ln_ok = warp < kScanRB && ln_row < B;
  float ln_s[kScanLnPer];   // this lane's LayerNorm scales, read once
#pragma unroll
  for (int i = 0; i < kScanLnPer; ++i) ln_s[i] = (lane + 32 * i) < D ? __ldg(ln_scale + lane + 32 * i) : 0.f;
  SCAN_TR_DECL;
  for (int ti = a.t1 - 1; ti >= a.t0; --ti) {
    // d loss / d o_{t+1}: the loss kernel's gradient plus the carry of step t + 1; the total goes back to gx[nl] (the
    // last layer's weight gradient reads it) and stays in `carry` for the residual path
    float *cur = S.xa, *nxt = S.xb;
    {
      float* g = a.gx[nl] + (long long)ti * B * O;
#pragma unroll
      for (int e = 0; e < kHeadPer; ++e) {
        const int i = tid + e * kScanThreads;
        if (i >= kScanRB * O) break;
        const int r = scan_div(i, O), c = i - r * O, row = row0 + r;
        float v = 0.f;
        if (row < B) {
          v = g_pre[e] + carry[i];
          g[(long long)row * O + c] = v;
        }
        carry[i] = v;
        cur[c * kScanRB + r] = v;
      }
    }
    fetch_loss_grad(ti - 1);
    float xh_pre[kScanLnPer], rs_pre = 0.f;
#pragma unroll
    for (int i = 0; i < kScanLnPer; ++i) {
      const int j = lane + 32 * i;
      xh_pre[i] = (ln_ok && ti > 0 && j < D) ? a.xhat[((long long)ti * B + ln_row) * D + j] : 0.f;
    }
    if (ln_ok && ti > 0) rs_pre = a.rstd[(long long)ti * B + ln_row];
    __syncthreads();
    SCAN_TR(0);
    for (int l = nl - 1; l >= 0; --l) {
      const int din = a.net.dims[l], dout = a.net.dims[l + 1], CG = scan_round4(din) >> 2;
      const bool vec = (din & 3) == 0;
      const float* xl = a.x[l] + (long long)ti * B * din;
      float4 xm[kMaskPer];
#pragma unroll
      for (int e = 0; e < kMaskPer; ++e) {
        const int i = tid + e * kScanThreads;
        const int r = scan_div(i, CG), cg = i - r * CG, row = row0 + r, c0 = 4 * cg;
        xm[e] = make_float4(1.f, 1.f, 1.f, 1.f);
        if (l > 0 && i < kScanRB * CG && row < B) {
          const float* p = xl + (long long)row * din + c0;
          if (vec) {
            xm[e] = *reinterpret_cast<const float4*>(p);
          } else {
            xm[e].x = p[0];
            if (c0 + 1 < din) xm[e].y = p[1];
            if (c0 + 2 < din) xm[e].z = p[2];
            if (c0 + 3 < din) xm[e].w = p[3];
          }
        }
      }
      scan_product(w[l], (a.cache_mask >> l & 1u) != 0, dout, din, cur, S.part, tid);
      SCAN_TR(2 + 2 * l);
      float* out = a.gx[l] + (long long)ti * B * din;
#pragma unroll
      for (int e = 0; e < kMaskPer; ++e) {
        const int i = tid + e * kScanThreads;
        if (i >= kScanRB * CG) break;
        const int r = scan_div(i, CG), cg = i - r * CG, row = row0 + r, c0 = 4 * cg;
        float4 v = scan_gather4(S.part, din, r, cg);
        // relu mask of the saved activation
        if (!(xm[e].x > 0.f)) v.x = 0.f;
        if (!(xm[e].y > 0.f)) v.y = 0.f;
        if (!(xm[e].z > 0.f)) v.z = 0.f;
        if (!(xm[e].w > 0.f)) v.w = 0.f;
        if (row >= B) v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < B) {
          float* p = out + (long long)row * din + c0;
          if (vec) {
            *reinterpret_cast<float4*>(p) = v;
          } else {
            p[0] = v.x;
            if (c0 + 1 < din) p[1] = v.y;
            if (c0 + 2 < din) p[2] = v.z;
            if (c0 + 3 < din) p[3] = v.w;
          }
        }
        nxt[(c0 + 0) * kScanRB + r] = v.x;
        if (c0 + 1 < din) nxt[(c0 + 1) * kScanRB + r] = v.y;
        if (c0 + 2 < din) nxt[(c0 + 2) * kScanRB + r] = v.z;
        if (c0 + 3 < din) nxt[(c0 + 3) * kScanRB + r] = v.w;
      }
      __syncthreads();
      SCAN_TR(3 + 2 * l);
      float* t = cur; cur = nxt; nxt = t;
    }
    // LayerNorm backward of d loss / d x0 to the observation part of concat(o_t, a_t) (as wm_ln_bwd_kernel)
    if (ti > 0 && ln_ok) {
      float d[kScanLnPer];
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int i = 0; i < kScanLnPer; ++i) {
        const int j = lane + 32 * i;
        d[i] = j < D ? cur[j * kScanRB + warp] * ln_s[i] : 0.f;
        s1 += d[i];
        s2 += d[i] * xh_pre[i];
      }
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        s2 += __shfl_xor_sync(0xffffffffu, s2, o);
      }
      const float m1 = s1 / D, m2 = s2 / D;
#pragma unroll
      for (int i = 0; i < kScanLnPer; ++i) {
        const int j = lane + 32 * i;
        if (j < O) carry[warp * O + j] += rs_pre * (d[i] - m1 - xh_pre[i] * m2);
      }
    }
    __syncthreads();
    SCAN_TR(1);
  }
  if (a.t0 > 0)
    for (int i = tid; i < kScanRB * O; i += kScanThreads) {
      const int r = scan_div(i, O), row = row0 + r;
      if (row < B) a.carry_g[(long long)row * O + (i - r * O)] = carry[i];
    }
  SCAN_TR_PRINT("wm_scan_bwd", nl);
}

// wt[n * K + k] = p[k * N + n]
__global__ void wm_transpose_kernel(const float* __restrict__ p, float* __restrict__ wt, int K, int N) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= K * N) return;
  const int n = i / K, k = i - n * K;
  wt[i] = p[(long long)k * N + n];
}
