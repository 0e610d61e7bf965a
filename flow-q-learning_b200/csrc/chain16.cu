// Layer-chain kernels, second generation (see chain.h for what a chain launch is and what it replaces).
//
// Same contract as fql_chain_kernel (chain.cu) -- one CTA carries a 128-row tile of one virtual member through every
// layer of a forward pass / the F Euler steps / an input-gradient chain, activations never leave the SM -- but the
// epilogue is laid out for throughput and the hand-off to the tensor core is per 64-column block:
//
//   * 20 warps: warp 0 = TMA weight producer, warp 1 = TMEM allocator + tcgen05.mma issuer (warps 2, 3 idle: the first
//     warpgroup hands its registers to the others via setmaxnreg), warps 4..19 = epilogue.
//     Four epilogue warps per TMEM lane quarter (= per scheduler), one per COLUMN GROUP cg: a thread owns one row and
//     columns [64 cg, 64 cg + 64) of each 256-column accumulator tile, i.e. exactly its row of ACT k-blocks cg and
//     4 + cg. 64 values per tile and thread fit in registers, so:
//   * tile 0 of a layer is processed while tile 1's MMAs still read ACT and its bf16 results are HELD IN REGISTERS
//     (no parking in TMEM, no flush pass); when tile 1 has landed they go to ACT block cg, the block is published
//     (blk_full[cg], 128 arrivals) and the next layer's MMAs start on it while tile 1's epilogue fills blocks 4..7 one
//     by one. The tensor core therefore only idles for the register -> ACT copy of a layer (a few hundred cycles).
//   * TMEM re-use is guarded by two `drained` barriers (every epilogue thread arrives once it has read its part of an
//     accumulator tile) instead of the act_lo / act_hi hand-shake: the next layer's tile-0 MMAs wait for drained[0],
//     its tile-1 MMAs for drained[1].
//   * TMEM loads are double-buffered in registers (the load of chunk c + 1 is in flight while chunk c is computed).
//   * No CTA-wide barriers in the epilogue: LayerNorm row statistics are exchanged between the four warps of a lane
//     quarter through a named barrier of 128 threads; bias / LayerNorm vectors are read through L1 (prefetched while
//     the warp waits for its accumulator).
//   * Saved tensors (act'(z), xhat) and the backward pass's inputs go through a 2 KB per-warp staging block (overlaying
//     weight stages 3 and 4) so that the global side is coalesced; h / dz are copied out of ACT after the hand-off.
//
// Covers the production shape (every hidden layer 512 wide, reductions of 64 or 512); launch_chain (chain.cu) falls
// back to the first-generation kernels otherwise.
#include "chain.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kEpiWarps = 16;
constexpr int kFirstEpiWarp = 4;                              // warps 0..3 = one warpgroup: producer, MMA issuer, two idle
constexpr int kThreads16 = (kFirstEpiWarp + kEpiWarps) * 32;  // 640
// Register budget: a scheduler hosts 5 warps (16384 registers): the kernel launches at 96 registers per thread, then the
// first warpgroup gives its registers back (24 each) and the four epilogue warpgroups grow to 120.
constexpr int kRegsLight = 64, kRegsEpi = 112;
constexpr bool kRebalanceRegs = false;  // ptxas 12.9 spills MORE with the setmaxnreg split than at a flat 96 (measured: 2-3x the spill bytes)
constexpr int kStages = 5;
constexpr int kStagesStaged = 3;
constexpr int kStgBytes = 2048;                 // per epilogue warp (16 x 2 KB = weight stages 3 and 4)
constexpr int kActBlockBytes = 128 * 64 * 2;    // one 64-wide k-block of the A operand = 16 KB
constexpr int kActBytes = 8 * kActBlockBytes;   // 128 KB
constexpr int kStageBytes = 16384;              // 32 k x 256 n bf16
constexpr int kStageK = 32;
constexpr int kTileN = 256;
constexpr int kFwdBoxBytes = 64 * kStageK * 2;
constexpr int kBarBytes = 512;
constexpr int kMaxSlots = 2 * kStages;         // pair mode: 8 KB weight slots
constexpr int kSmemBytes16 = kActBytes + kActBlockBytes + kStages * kStageBytes + kBarBytes + 1024;

__device__ __forceinline__ float clip1(float x) { return fminf(fmaxf(x, -1.f), 1.f); }
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void quarter_bar_sync(int q) { asm volatile("bar.sync %0, 128;" ::"r"(q + 1) : "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// lanes 0..15 end up with the sum over the warp's 32 rows of column (lane & 15) of a 32 x 16 block held one row per lane
__device__ __forceinline__ float warp_colsum16(float (&v)[16], int lane) {
#pragma unroll
  for (int s = 8; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int j = 0; j < s; ++j) {
      const float send = up ? v[j] : v[j + s];
      const float keep = up ? v[j + s] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 16);
}

// Shared memory is addressed explicitly (32-bit shared-space addresses, st.shared / ld.shared): through generic pointers the
// compiler emits generic ST.E / LD.E with 64-bit addresses for these (it cannot prove the address space across the Epi struct).
__device__ __forceinline__ void sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}
// 16 columns starting at `col` (multiple of 16) of row r -> ACT (two 16-byte chunks of the 128-byte-swizzled row)
__device__ __forceinline__ void st_act16(uint32_t act, int r, int col, const uint32_t* w) {
  const uint32_t row = act + (uint32_t)((col >> 6) * kActBlockBytes + r * 128);
  const int c0 = (col & 63) >> 3;
  sts128(row + ((c0 ^ (r & 7)) << 4), w[0], w[1], w[2], w[3]);
  sts128(row + (((c0 + 1) ^ (r & 7)) << 4), w[4], w[5], w[6], w[7]);
}

// ---------------------------------------------------------------------------------------------------------------
// Staging blocks of the BACKWARD chains (per warp, 2 KB = two quarter blocks; the forward passes store what they save
// straight from registers and stage nothing).
//   quarter block: 32 rows x 16 bf16 (32 B per row), 1 KB; chunk c (0..1) of row r at r * 32 + ((c ^ ((r >> 2) & 1)) << 4)
// Row-wise accesses (lane = row) and the coalesced sweeps are bank-conflict free.
// ---------------------------------------------------------------------------------------------------------------
// global rows -> quarter block, asynchronously (rows that do not exist are zero-filled). One commit group.
__device__ __forceinline__ void stage_in_quarter(uint32_t blk, const __nv_bfloat16* g, int ld, int lane, int rows_valid) {
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int p = lane + 32 * i;
    const int r = p >> 1, c = p & 1;
    const uint32_t dst = blk + r * 32 + ((c ^ ((r >> 2) & 1)) << 4);
    if (r < rows_valid)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(g + (long long)r * ld + c * 8) : "memory");
    else sts128(dst, 0u, 0u, 0u, 0u);
  }
}
__device__ __forceinline__ void ld_quarter16(uint32_t blk, int lane, uint32_t* w) {
  const uint32_t row = blk + lane * 32;
  const int sw = (lane >> 2) & 1;
  const uint4 a = lds128(row + (sw << 4));
  const uint4 b = lds128(row + ((sw ^ 1) << 4));
  w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}
// 64-column block of ACT (this warp's 32 rows: 32 x 128 bytes) -> global; 4 rows x 128 contiguous bytes per warp instruction
__device__ __forceinline__ void copy_out_block(uint32_t blk, __nv_bfloat16* g, int ld, int lane, int rows_valid) {
  const int ch = lane & 7, rsub = lane >> 3;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i * 4 + rsub;
    const uint4 v = lds128(blk + r * 128 + ((ch ^ (r & 7)) << 4));
    if (r < rows_valid) *reinterpret_cast<uint4*>(g + (long long)r * ld + ch * 8) = v;
  }
}
// the same sweep, also summing every column over the warp's 32 rows (rows that do not exist hold zeros)
__device__ __forceinline__ void copy_out_block_colsum(uint32_t blk, __nv_bfloat16* g, int ld, int lane,
                                                      int rows_valid, float* part) {
  const int ch = lane & 7, rsub = lane >> 3;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i * 4 + rsub;
    const uint4 v = lds128(blk + r * 128 + ((ch ^ (r & 7)) << 4));
    if (g != nullptr && r < rows_valid) *reinterpret_cast<uint4*>(g + (long long)r * ld + ch * 8) = v;  // (nullptr: TMA stores it)
    acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x);
    acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
    acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z);
    acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 8);
    acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], 16);
  }
  if (lane < 8) {
    float4* p4 = reinterpret_cast<float4*>(part + ch * 8);
    p4[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    p4[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

// Per-thread view of the epilogue.
struct Epi {
  uint32_t tmem_row;       // TMEM address of this thread's lane, column 0
  uint64_t* tmem_full;     // this CTA's accumulator-ready barriers (the MMA commit is multicast to both CTAs of a pair)
  // shared::cluster addresses of the barriers the MMA issuer waits on: blk_full[0], drained[0], x0_ready of the pair's
  // LEADER CTA (of this CTA itself when CTAs are not paired). One arrival per WARP.
  uint32_t blk_full_a, drained_a, x0_ready_a;
  bool remote;             // those barriers live in the other CTA of the pair
  const float* pb;         // shared-memory copy of the current layer's bias | LayerNorm scale | LayerNorm bias restricted to
                           // this column group's 128 columns ([3][128] floats), or nullptr -> read them through L1 / L2
  float* pb_base;          // this column group's two parameter buffers (layer parity), nullptr when the X0 block is in use
  uint32_t act;            // shared-space address of ACT
  const float* params;     // this virtual member's fp32 parameter block
  float2* ex;              // LayerNorm row-sum exchange [4 column groups][128 rows] (the X0 block, idle by then)
  uint32_t stg;            // shared-space address of this warp's staging block (launches that stage tensors only)
  int r, q, cg, lane, H;
  bool row_ok;
  int rows_valid;          // how many of the warp's 32 rows exist
  long long wrow;          // element offset of the warp's first row in the pass buffers ([rows][H])
  long long wdz;           // ... in the dz buffers
  long long rs_off;        // offset of this thread's row in the rstd buffers
  long long part_off;      // offset of this (row tile, lane quarter) row in the partial-sum buffers
  float ln_eps;
  const CUtensorMap* map_h;  // this layer's h store map (ChainLayer::h_tma)
  int h_row, h_v;          // first row of this warp in the pass buffers, virtual member (TMA store coordinates)
  bool h_pending;          // this warp has TMA stores in flight that read its rows of ACT
  long long* tr;           // (make TRACE=1) stamp log of the traced epilogue thread, 8 slots per layer
};
__device__ __forceinline__ void stamp(const Epi& C, int slot) {
#ifdef FQLPOP_EPI_TRACE
  if (C.tr) C.tr[slot] = clock64();
#endif
}

// this warp's rows of ACT block `blk` are written and its accumulator columns are read: let the tensor core have them.
// Every lane orders its own generic-proxy writes before the async proxy (the fence also waits for them to be performed),
// the warp converges, one lane arrives. The barrier may live in the other CTA of the pair; the arrive is deliberately NOT
// .release.cluster (measured: that form costs ~5000 cycles per layer -- a cluster-scope fence per warp that also drops
// the L1 lines the bias loads hit): what it publishes is this CTA's own shared memory, read by this CTA's own tensor
// core, already fenced for the async proxy above.
__device__ __forceinline__ void arrive_on(const Epi& C, uint32_t addr) {
  if (C.remote) mbar_arrive_cluster(addr);
  else asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void publish(const Epi& C, int blk) {
  tc_fence_before();
  fence_proxy_async_smem();
  __syncwarp();
  if (C.lane == 0) arrive_on(C, C.blk_full_a + 8u * (uint32_t)blk);
}
__device__ __forceinline__ void arrive_drained(const Epi& C, int t) {
  tc_fence_before();
  __syncwarp();
  if (C.lane == 0) arrive_on(C, C.drained_a + 8u * (uint32_t)t);
}
__device__ __forceinline__ void wait_full(const Epi& C, int t, uint32_t parity) {
  mbar_wait(&C.tmem_full[t], parity);
  tc_fence_after();
}
__device__ __forceinline__ void prefetch_params(const Epi& C, const ChainLayer& Ly, bool fwd, bool ln) {
  // the 128 columns this thread touches per layer = 4 lines per vector; lanes 0..3 / 4..7 / 8..11 take one vector each
  const int line = C.lane & 3, which = C.lane >> 2;
  const int col = (line >> 1) * kTileN + C.cg * 64 + (line & 1) * 32;
  if (fwd) {
    if (which == 0) prefetch_l1(C.params + Ly.off_b + col);
    else if (ln && which == 1) prefetch_l1(C.params + Ly.off_g + col);
    else if (ln && which == 2) prefetch_l1(C.params + Ly.off_beta + col);
  } else if (ln && which == 0) {
    prefetch_l1(C.params + Ly.off_g + col);
  }
}
__device__ __forceinline__ void ld_vec16(const float* p, float4 (&v)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = __ldg(reinterpret_cast<const float4*>(p) + i);
}
// 16 parameters of array `arr` (0 bias, 1 LayerNorm scale, 2 LayerNorm bias; element offset `off` in the parameter block)
// at absolute column `col` of this thread's column group: from the shared-memory copy when there is one
__device__ __forceinline__ void ld_param16(const Epi& C, int arr, long long off, int col, float4 (&v)[4]) {
  if (C.pb) {
    const int k = (col >= kTileN ? 64 : 0) + (col & 63);
    const float4* p = reinterpret_cast<const float4*>(C.pb + arr * 128 + k);
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = p[i];
  } else {
    ld_vec16(C.params + off + col, v);
  }
}
// Every bias / LayerNorm vector element is used once per CTA and layer, so each read through L1 is an L2 round trip
// (~600 cycles, a dozen of them per thread and layer in the LayerNorm passes: measured as the epilogue's main
// long-scoreboard stall). One warp of each column group copies the group's 3 x 128 values into shared memory while
// the group waits for its accumulator (the X0 block is idle after the first layer; two buffers alternate by layer, so
// nobody can still be reading the one being written), then the group's four warps meet on a named barrier.
__device__ __forceinline__ void load_layer_params(Epi& C, const ChainLayer& Ly, int layer_parity, bool need_bias, bool need_ln,
                                                  bool need_gamma_only) {
  if (C.pb_base == nullptr || Ly.a_from_x0) { C.pb = nullptr; return; }
  float* buf = C.pb_base + layer_parity * (3 * 128);
  if (C.q == 0) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const bool need = j == 0 ? need_bias : (j == 1 ? (need_ln || need_gamma_only) : need_ln);
      if (!need) continue;
      const long long off = j == 0 ? Ly.off_b : (j == 1 ? Ly.off_g : Ly.off_beta);
      const int k = C.lane;                                    // float4 index within the group's 128 columns
      const int col = (k < 16 ? 0 : kTileN) + C.cg * 64 + (k & 15) * 4;
      reinterpret_cast<float4*>(buf + j * 128)[k] = __ldg(reinterpret_cast<const float4*>(C.params + off + col));
    }
  }
  asm volatile("bar.sync %0, 128;" ::"r"(5 + C.cg) : "memory");
  C.pb = buf;
}

// ---------------------------------------------------------------------------------------------------------------
// forward hidden layer: h = LayerNorm?(gelu(acc + b))
// ---------------------------------------------------------------------------------------------------------------
// bias + GELU (+ act'(z)) of one 16-column chunk; LN accumulates the row sums of g and g^2
template <bool LN, bool SAVE>
__device__ __forceinline__ void fwd_chunk(const uint32_t (&rr)[16], const float4 (&b4)[4], uint32_t* pk, uint32_t* gpk,
                                          float& s1, float& s2) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float g0, g1, g2, g3, p0, p1, p2, p3;
    gelu1<SAVE>(__uint_as_float(rr[4 * i + 0]) + b4[i].x, g0, p0);
    gelu1<SAVE>(__uint_as_float(rr[4 * i + 1]) + b4[i].y, g1, p1);
    gelu1<SAVE>(__uint_as_float(rr[4 * i + 2]) + b4[i].z, g2, p2);
    gelu1<SAVE>(__uint_as_float(rr[4 * i + 3]) + b4[i].w, g3, p3);
    if (LN) {
      s1 += (g0 + g1) + (g2 + g3);
      s2 = fmaf(g0, g0, fmaf(g1, g1, fmaf(g2, g2, fmaf(g3, g3, s2))));
    }
    pk[2 * i] = pack_bf16x2(g0, g1);
    pk[2 * i + 1] = pack_bf16x2(g2, g3);
    if (SAVE) {
      gpk[2 * i] = pack_bf16x2(p0, p1);
      gpk[2 * i + 1] = pack_bf16x2(p2, p3);
    }
  }
}

// 16 bf16 values (one 32-byte sector) of this thread's row straight to global memory: one 256-bit store. Used for the
// tensors a forward pass saves for the backward pass (act'(z), the LayerNorm's xhat): forward launches then need no staging
// area and keep all five weight slots (with three, the policy pass spent 60 % of its MMA issuer's time waiting for weights).
__device__ __forceinline__ void st_row32(__nv_bfloat16* dst, const uint32_t (&w)[8], bool ok) {
  if (ok)
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}

// DEST: 0 = keep the 64 bf16 results in pk[32]; 1 = write them to ACT; 2 = park them in the drained TMEM columns
template <bool LN, bool SAVE, int DEST>
__device__ __forceinline__ void fwd_pass1(const Epi& C, const ChainLayer& Ly, int tile, uint32_t (&pk)[32], float& s1,
                                          float& s2) {
  const int colw = tile * kTileN + C.cg * 64;
  const uint32_t tbase = C.tmem_row + colw;
  uint32_t ra[16], rb[16];
  tmem_ld16(tbase, ra);
#pragma unroll
  for (int c = 0; c < 4; c += 2) {
    uint32_t gpk[8];
    float4 b4[4];
    // ---- even chunk (ra) ----
    ld_param16(C, 0, Ly.off_b, colw + c * 16, b4);
    tmem_ld_wait();
    tmem_ld16(tbase + (c + 1) * 16, rb);
    fwd_chunk<LN, SAVE>(ra, b4, pk + 8 * c, gpk, s1, s2);
    if (SAVE) st_row32(Ly.out_gp + C.wrow + (long long)C.lane * C.H + colw + c * 16, gpk, C.row_ok);
    if (DEST == 1) st_act16(C.act, C.r, colw + c * 16, pk + 8 * c);
    if (DEST == 2) tmem_st8(tbase + c * 8, *reinterpret_cast<const uint32_t(*)[8]>(pk + 8 * c));
    // ---- odd chunk (rb) ----
    ld_param16(C, 0, Ly.off_b, colw + (c + 1) * 16, b4);
    tmem_ld_wait();
    if (c + 2 < 4) tmem_ld16(tbase + (c + 2) * 16, ra);
    fwd_chunk<LN, SAVE>(rb, b4, pk + 8 * (c + 1), gpk, s1, s2);
    if (SAVE) st_row32(Ly.out_gp + C.wrow + (long long)C.lane * C.H + colw + (c + 1) * 16, gpk, C.row_ok);
    if (DEST == 1) st_act16(C.act, C.r, colw + (c + 1) * 16, pk + 8 * (c + 1));
    if (DEST == 2) tmem_st8(tbase + (c + 1) * 8, *reinterpret_cast<const uint32_t(*)[8]>(pk + 8 * (c + 1)));
  }
}

// LayerNorm of 16 held / parked bf16 values: x = (g - mean) rstd, h = x gamma + beta -> ACT (+ xhat -> global)
template <bool SAVE>
__device__ __forceinline__ void ln_chunk(const Epi& C, const ChainLayer& Ly, int col, const uint32_t* pr, float rstd,
                                         float shift, int cw, bool save_x) {
  float4 ga[4], be[4];
  ld_param16(C, 1, Ly.off_g, col, ga);
  ld_param16(C, 2, Ly.off_beta, col, be);
  uint32_t hk[8], xk[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t w0 = pr[2 * i], w1 = pr[2 * i + 1];
    const float x0 = fmaf(bf16_lo(w0), rstd, shift), x1 = fmaf(bf16_hi(w0), rstd, shift);
    const float x2 = fmaf(bf16_lo(w1), rstd, shift), x3 = fmaf(bf16_hi(w1), rstd, shift);
    if (SAVE) {
      xk[2 * i] = pack_bf16x2(x0, x1);
      xk[2 * i + 1] = pack_bf16x2(x2, x3);
    }
    hk[2 * i] = pack_bf16x2(fmaf(x0, ga[i].x, be[i].x), fmaf(x1, ga[i].y, be[i].y));
    hk[2 * i + 1] = pack_bf16x2(fmaf(x2, ga[i].z, be[i].z), fmaf(x3, ga[i].w, be[i].w));
  }
  st_act16(C.act, C.r, col, hk);
  if (SAVE && save_x) st_row32(Ly.out_xhat + C.wrow + (long long)C.lane * C.H + col, xk, C.row_ok);
}

// parked tile (bf16 pairs in the drained accumulator columns) -> LayerNorm -> ACT (+ xhat -> global)
template <bool SAVE>
__device__ __forceinline__ void ln_flush_tile(const Epi& C, const ChainLayer& Ly, int tile, float rstd, float shift, bool save_x) {
  const int colw = tile * kTileN + C.cg * 64;
  uint32_t pa[8], pb[8];
  tmem_ld8(C.tmem_row + colw, pa);
#pragma unroll
  for (int c = 0; c < 4; c += 2) {
    tmem_ld_wait();
    tmem_ld8(C.tmem_row + colw + (c + 1) * 8, pb);
    ln_chunk<SAVE>(C, Ly, colw + c * 16, pa, rstd, shift, 0, save_x);
    tmem_ld_wait();
    if (c + 2 < 4) tmem_ld8(C.tmem_row + colw + (c + 2) * 8, pa);
    ln_chunk<SAVE>(C, Ly, colw + (c + 1) * 16, pb, rstd, shift, 16, save_x);
  }
}

template <bool LN, bool SAVE>
__device__ __forceinline__ void fwd_hidden_layer(Epi& C, const ChainLayer& Ly, uint32_t par0, uint32_t par1, bool pub,
                                                 int layer_parity) {
  // the previous layer's h stores (TMA, below) must have read this warp's rows of ACT before they are rewritten; they
  // were issued a whole layer of MMAs ago, so this wait does not wait
  if (C.h_pending) {
    if (C.lane == 0) bulk_wait_group_read0();
    __syncwarp();
    C.h_pending = false;
  }
  load_layer_params(C, Ly, layer_parity, true, LN, false);
  if (!C.pb) prefetch_params(C, Ly, true, LN);
  float s1 = 0.f, s2 = 0.f;
  if (!LN) {
    uint32_t pk0[32];
    wait_full(C, 0, par0);
    stamp(C, 1);
    fwd_pass1<false, SAVE, 0>(C, Ly, 0, pk0, s1, s2);
    arrive_drained(C, 0);
    stamp(C, 2);
    wait_full(C, 1, par1);  // every MMA of the layer has retired: ACT may be overwritten
    stamp(C, 3);
#pragma unroll
    for (int c = 0; c < 4; ++c) st_act16(C.act, C.r, C.cg * 64 + c * 16, pk0 + 8 * c);
    if (pub) publish(C, C.cg);
    stamp(C, 4);
    uint32_t pk1[32];
    fwd_pass1<false, SAVE, 1>(C, Ly, 1, pk1, s1, s2);
    if (pub) publish(C, 4 + C.cg);
    arrive_drained(C, 1);
    stamp(C, 5);
  } else {
    // LayerNorm needs the whole row before anything can be normalised: both tiles are parked as bf16 in the accumulator
    // columns they were drained from (holding tile 0 in registers instead costs the LayerNorm instantiation spills in
    // its inner loops), then flushed tile by tile once the row statistics are known.
    uint32_t pk[32];
    wait_full(C, 0, par0);
    stamp(C, 1);
    fwd_pass1<true, SAVE, 2>(C, Ly, 0, pk, s1, s2);
    stamp(C, 2);
    wait_full(C, 1, par1);
    stamp(C, 3);
    fwd_pass1<true, SAVE, 2>(C, Ly, 1, pk, s1, s2);
    tmem_st_wait();
    // row statistics over all 512 columns: the four column-group warps of this lane quarter
    C.ex[C.cg * 128 + C.r] = make_float2(s1, s2);
    quarter_bar_sync(C.q);
    float t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 o = C.ex[j * 128 + C.r];
      t1 += o.x;
      t2 += o.y;
    }
    const float inv_n = 1.0f / (float)C.H;
    const float mean = t1 * inv_n;
    const float var = fmaxf(t2 * inv_n - mean * mean, 0.f);  // flax LayerNorm: E[x^2] - E[x]^2
    const float rstd = rsqrtf(var + C.ln_eps);
    const float shift = -mean * rstd;
    if (C.row_ok && C.cg == 0 && Ly.out_rstd) Ly.out_rstd[C.rs_off] = rstd;
    const bool save_x = SAVE && Ly.out_xhat != nullptr;
    ln_flush_tile<SAVE>(C, Ly, 0, rstd, shift, save_x);
    if (pub) publish(C, C.cg);
    arrive_drained(C, 0);
    stamp(C, 4);
    ln_flush_tile<SAVE>(C, Ly, 1, rstd, shift, save_x);
    if (pub) publish(C, 4 + C.cg);
    arrive_drained(C, 1);
    stamp(C, 5);
  }
  // h -> HBM out of ACT, off the hand-off path: these rows of the two blocks are next written by this very warp
  // (measured: storing h from the registers like act'(z), one 32-byte sector per thread and chunk, makes the policy pass
  // slower -- 59.7 vs 54.1 us; the copy-out writes whole 128-byte row segments)
  if (Ly.out_h) {
    const uint32_t b0 = C.act + C.cg * kActBlockBytes + C.q * (32 * 128);
    if (Ly.h_tma) {
      // the warp's 32 rows of its two 64-column blocks are two 4 KB boxes in the 128-byte-swizzled layout TMA knows: one
      // lane hands them to the copy engine (rows beyond the job's last row are clipped by the map) -- no epilogue
      // instructions for the copy-out (it was ~5 k cycles per layer of a saving pass)
      fence_proxy_async_smem();
      __syncwarp();
      if (C.lane == 0 && C.rows_valid > 0) {
        tma_store_3d(C.map_h, b0, C.cg * 64, C.h_row, C.h_v);
        tma_store_3d(C.map_h, b0 + 4 * kActBlockBytes, kTileN + C.cg * 64, C.h_row, C.h_v);
        bulk_commit_group();
      }
      C.h_pending = true;
    } else {
      __syncwarp();
      copy_out_block(b0, Ly.out_h + C.wrow + C.cg * 64, C.H, C.lane, C.rows_valid);
      copy_out_block(b0 + 4 * kActBlockBytes, Ly.out_h + C.wrow + kTileN + C.cg * 64, C.H, C.lane, C.rows_valid);
    }
  }
  stamp(C, 6);
}

// ---------------------------------------------------------------------------------------------------------------
// backward hidden layer: dz = act'(z) * LNbwd?(acc), acc = dz_next W^T
// ---------------------------------------------------------------------------------------------------------------
// chunk cc (0..7 over both tiles) of this warp's columns: absolute column
__device__ __forceinline__ int chunk_col(const Epi& C, int cc) { return (cc >> 2) * kTileN + C.cg * 64 + (cc & 3) * 16; }

// no LayerNorm. The saved act'(z) streams through the two quarter-block slots, two chunks ahead.
__device__ __forceinline__ void bwd_hidden_plain(const Epi& C, const ChainLayer& Ly, uint32_t par0, uint32_t par1, bool pub) {
  const bool has_gp = Ly.gp != nullptr;
  const __nv_bfloat16* gsrc = has_gp ? Ly.gp + C.wrow : nullptr;
  auto prefetch = [&](int cc) {
    if (has_gp && cc < 8) stage_in_quarter(C.stg + (cc & 1) * 1024, gsrc + chunk_col(C, cc), C.H, C.lane, C.rows_valid);
    cp_async_commit();
  };
  prefetch(0);
  prefetch(1);
  uint32_t pk0[32];
  uint32_t ra[16], rb[16];
  auto gp_chunk = [&](int cc, uint32_t (&gw)[8]) {
    if (has_gp) {
      cp_async_wait_group<1>();
      __syncwarp();
      ld_quarter16(C.stg + (cc & 1) * 1024, C.lane, gw);
      __syncwarp();
      prefetch(cc + 2);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) gw[i] = C.row_ok ? 0x3F803F80u : 0u;  // bf16 (1, 1) / (0, 0)
    }
  };
  auto mul_chunk = [&](const uint32_t (&rr)[16], const uint32_t (&gw)[8], uint32_t* pk) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      pk[i] = pack_bf16x2(__uint_as_float(rr[2 * i]) * bf16_lo(gw[i]), __uint_as_float(rr[2 * i + 1]) * bf16_hi(gw[i]));
  };
  // ---- tile 0 -> registers ----
  wait_full(C, 0, par0);
  {
    const uint32_t tbase = C.tmem_row + C.cg * 64;
    tmem_ld16(tbase, ra);
#pragma unroll
    for (int c = 0; c < 4; c += 2) {
      uint32_t gw[8];
      gp_chunk(c, gw);
      tmem_ld_wait();
      tmem_ld16(tbase + (c + 1) * 16, rb);
      mul_chunk(ra, gw, pk0 + 8 * c);
      gp_chunk(c + 1, gw);
      tmem_ld_wait();
      if (c + 2 < 4) tmem_ld16(tbase + (c + 2) * 16, ra);
      mul_chunk(rb, gw, pk0 + 8 * (c + 1));
    }
  }
  arrive_drained(C, 0);
  wait_full(C, 1, par1);
#pragma unroll
  for (int c = 0; c < 4; ++c) st_act16(C.act, C.r, C.cg * 64 + c * 16, pk0 + 8 * c);
  if (pub) publish(C, C.cg);
  // ---- tile 1 -> ACT ----
  {
    const int colw = kTileN + C.cg * 64;
    const uint32_t tbase = C.tmem_row + colw;
    tmem_ld16(tbase, ra);
#pragma unroll
    for (int c = 0; c < 4; c += 2) {
      uint32_t gw[8], pk[8];
      gp_chunk(4 + c, gw);
      tmem_ld_wait();
      tmem_ld16(tbase + (c + 1) * 16, rb);
      mul_chunk(ra, gw, pk);
      st_act16(C.act, C.r, colw + c * 16, pk);
      gp_chunk(4 + c + 1, gw);
      tmem_ld_wait();
      if (c + 2 < 4) tmem_ld16(tbase + (c + 2) * 16, ra);
      mul_chunk(rb, gw, pk);
      st_act16(C.act, C.r, colw + (c + 1) * 16, pk);
    }
  }
  if (pub) publish(C, 4 + C.cg);
  arrive_drained(C, 1);
  cp_async_wait_group<0>();
}

// LayerNorm backward. Statistics pass (the accumulator stays in TMEM untouched): h = xhat gamma + beta, acc = dL/dh ->
// dxhat = acc gamma; s1 = sum dxhat, s2 = sum dxhat xhat; partial column sums dbeta = colsum(acc), dgamma = colsum(acc xhat).
// Output pass: dg = rstd (dxhat - mean(dxhat) - xhat mean(dxhat xhat)); dz = act'(z) dg -> ACT.
__device__ __forceinline__ void bwd_hidden_ln(const Epi& C, const ChainLayer& Ly, uint32_t par0, uint32_t par1, bool pub) {
  const __nv_bfloat16* xsrc = Ly.xhat + C.wrow;
  const bool has_gp = Ly.gp != nullptr;
  const __nv_bfloat16* gsrc = has_gp ? Ly.gp + C.wrow : nullptr;
  auto prefetch_x = [&](int cc) {
    if (cc < 8) stage_in_quarter(C.stg + (cc & 1) * 1024, xsrc + chunk_col(C, cc), C.H, C.lane, C.rows_valid);
    cp_async_commit();
  };
  prefetch_x(0);
  prefetch_x(1);
  float s1 = 0.f, s2 = 0.f;
  uint32_t rr[16];
#pragma unroll 1
  for (int cc = 0; cc < 8; ++cc) {
    if (cc == 0) wait_full(C, 0, par0);
    if (cc == 4) wait_full(C, 1, par1);
    const int col = chunk_col(C, cc);
    tmem_ld16(C.tmem_row + col, rr);
    uint32_t xw[8];
    cp_async_wait_group<1>();
    __syncwarp();
    ld_quarter16(C.stg + (cc & 1) * 1024, C.lane, xw);
    __syncwarp();
    prefetch_x(cc + 2);
    float4 ga[4];
    ld_param16(C, 1, Ly.off_g, col, ga);
    tmem_ld_wait();
    float t0[16], t1[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float gam[4] = {ga[i].x, ga[i].y, ga[i].z, ga[i].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float dh = C.row_ok ? __uint_as_float(rr[4 * i + e]) : 0.f;
        const uint32_t w = xw[2 * i + (e >> 1)];
        const float xh = (e & 1) ? bf16_hi(w) : bf16_lo(w);
        const float dxh = dh * gam[e];
        s1 += dxh;
        s2 = fmaf(dxh, xh, s2);
        t0[4 * i + e] = dh;
        t1[4 * i + e] = dh * xh;
      }
    }
    if (Ly.dbeta_part) {
      const float cb = warp_colsum16(t0, C.lane);
      const float cgm = warp_colsum16(t1, C.lane);
      if (C.lane < 16) {
        Ly.dbeta_part[C.part_off + col + C.lane] = cb;
        Ly.dgamma_part[C.part_off + col + C.lane] = cgm;
      }
    }
  }
  cp_async_wait_group<0>();
  C.ex[C.cg * 128 + C.r] = make_float2(s1, s2);
  quarter_bar_sync(C.q);
  float t1s = 0.f, t2s = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 o = C.ex[j * 128 + C.r];
    t1s += o.x;
    t2s += o.y;
  }
  const float inv_n = 1.0f / (float)C.H;
  const float rstd = C.row_ok ? __ldg(Ly.rstd + C.rs_off) : 0.f;
  const float k1 = -(t1s * inv_n) * rstd, k2 = -(t2s * inv_n) * rstd;
  // ---- output pass: slot 0 = xhat, slot 1 = act'(z) of the current chunk ----
  auto prefetch_xg = [&](int cc) {
    if (cc < 8) {
      stage_in_quarter(C.stg, xsrc + chunk_col(C, cc), C.H, C.lane, C.rows_valid);
      if (has_gp) stage_in_quarter(C.stg + 1024, gsrc + chunk_col(C, cc), C.H, C.lane, C.rows_valid);
    }
    cp_async_commit();
  };
  prefetch_xg(0);
#pragma unroll 1
  for (int cc = 0; cc < 8; ++cc) {
    const int col = chunk_col(C, cc);
    tmem_ld16(C.tmem_row + col, rr);
    uint32_t xw[8], gw[8];
    cp_async_wait_group<0>();
    __syncwarp();
    ld_quarter16(C.stg, C.lane, xw);
    if (has_gp) ld_quarter16(C.stg + 1024, C.lane, gw);
    else {
#pragma unroll
      for (int i = 0; i < 8; ++i) gw[i] = C.row_ok ? 0x3F803F80u : 0u;
    }
    __syncwarp();
    prefetch_xg(cc + 1);
    float4 ga[4];
    ld_param16(C, 1, Ly.off_g, col, ga);
    tmem_ld_wait();
    uint32_t pk[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float gam[4] = {ga[i].x, ga[i].y, ga[i].z, ga[i].w};
      float dz[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint32_t wx = xw[2 * i + (e >> 1)], wg = gw[2 * i + (e >> 1)];
        const float xh = (e & 1) ? bf16_hi(wx) : bf16_lo(wx);
        const float gpv = (e & 1) ? bf16_hi(wg) : bf16_lo(wg);
        const float dxh = __uint_as_float(rr[4 * i + e]) * gam[e];
        dz[e] = gpv * fmaf(xh, k2, fmaf(dxh, rstd, k1));
      }
      pk[2 * i] = pack_bf16x2(dz[0], dz[1]);
      pk[2 * i + 1] = pack_bf16x2(dz[2], dz[3]);
    }
    st_act16(C.act, C.r, col, pk);
    if (cc == 3) {
      if (pub) publish(C, C.cg);
      arrive_drained(C, 0);
    }
  }
  cp_async_wait_group<0>();
  if (pub) publish(C, 4 + C.cg);
  arrive_drained(C, 1);
}

template <bool LN>
__device__ __forceinline__ void bwd_hidden_layer(Epi& C, const ChainLayer& Ly, uint32_t par0, uint32_t par1, bool pub,
                                                 int layer_parity) {
  if (C.h_pending) {   // the previous layer's dz stores (TMA, below) have read this warp's rows of ACT
    if (C.lane == 0) bulk_wait_group_read0();
    __syncwarp();
    C.h_pending = false;
  }
  if (LN) load_layer_params(C, Ly, layer_parity, false, false, true);
  else C.pb = nullptr;
  if (!C.pb) prefetch_params(C, Ly, false, LN);
  if (LN) bwd_hidden_ln(C, Ly, par0, par1, pub);
  else bwd_hidden_plain(C, Ly, par0, par1, pub);
  stamp(C, 5);
  // dz -> HBM (+ bias-gradient partial sums) out of ACT, off the hand-off path. With a store map (ChainLayer::h_tma) the
  // copy engine moves the warp's two 4 KB boxes (as the forward passes' h) and only the column sums are left to the warp.
  const uint32_t b0 = C.act + C.cg * kActBlockBytes + C.q * (32 * 128);
  if (Ly.h_tma) {
    fence_proxy_async_smem();
    __syncwarp();
    if (C.lane == 0 && C.rows_valid > 0) {
      tma_store_3d(C.map_h, b0, C.cg * 64, C.h_row, C.h_v);
      tma_store_3d(C.map_h, b0 + 4 * kActBlockBytes, kTileN + C.cg * 64, C.h_row, C.h_v);
      bulk_commit_group();
    }
    C.h_pending = true;
  } else {
    __syncwarp();
  }
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int col = t * kTileN + C.cg * 64;
    __nv_bfloat16* dst = Ly.h_tma ? nullptr : Ly.out_dz + C.wdz + col;
    if (Ly.dbias_part)
      copy_out_block_colsum(b0 + t * 4 * kActBlockBytes, dst, C.H, C.lane, C.rows_valid, Ly.dbias_part + C.part_off + col);
    else if (dst) copy_out_block(b0 + t * 4 * kActBlockBytes, dst, C.H, C.lane, C.rows_valid);
  }
  stamp(C, 6);
}

// ===============================================================================================================
// PAIR: CTAs run as clusters of two (the two CTAs carry adjacent 128-row tiles of the same virtual member) and every
// product is ONE tcgen05.mma.cta_group::2 of M = 256 issued by the leader (cluster rank 0): each CTA stages only HALF of
// the N columns of every weight stage (8 KB instead of 16 KB per 32-deep k step), so the bytes a CTA has to pull through
// its shared memory per layer halve -- the weight stream (stage round trip ~3000 cycles, measured) is what bounds the
// single-CTA kernel. Everything the leader's MMA issuer waits on (ACT blocks, drained accumulators, X0, weight stages)
// collects arrivals from both CTAs; its commits are multicast to both.
template <int MODE, bool LN, bool PAIR>  // MODE 0 = forward / Euler flow, 1 = backward (input-gradient chain)
__global__ void __launch_bounds__(kThreads16, 1) fql_chain16_kernel(const __grid_constant__ ChainLaunch L) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* x0s = smem + kActBytes;
  uint8_t* stage_base = x0s + kActBlockBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_base + kStages * kStageBytes);
  uint64_t* full_bar = bars;                    // [kMaxSlots]
  uint64_t* empty_bar = bars + kMaxSlots;       // [kMaxSlots]
  uint64_t* tmem_full = bars + 2 * kMaxSlots;   // [2] one per 256-column accumulator tile
  uint64_t* blk_full = tmem_full + 2;           // [8] ACT k-block j holds the next layer's A operand
  uint64_t* drained = blk_full + 8;             // [2] every epilogue warp has read its part of accumulator tile t
  uint64_t* a0_full = drained + 2;
  uint64_t* x0_ready = a0_full + 1;             // Euler: the X0 block holds the next step's input
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x0_ready + 1);
  // weight slot = one stage of THIS CTA: 32 k x 256 n; pair mode, backward: 32 k x this CTA's 128 rows (8 KB slots, twice
  // as many); pair mode, forward: 64 k x this CTA's 128 n -- four MMAs per wait / commit of the one issuing thread, whose
  // loop overhead is what bounds a pair (measured in rollout16.cu: 310 -> 348 M env-steps/s)
  constexpr bool kDeep = PAIR && MODE == 0;
  constexpr int kSlotBytes = (PAIR && MODE == 1) ? kStageBytes / 2 : kStageBytes;
  constexpr int kStgK = kDeep ? 64 : kStageK;       // reduction elements per weight stage
  constexpr int kFwdBox = 64 * kStgK * 2;           // forward weight box: 64 n x kStgK k
  constexpr uint32_t kCtas = PAIR ? 2 : 1;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  long long tl_t0 = 0;
  if (L.timeline && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0));

  // ---- tile decode ---------------------------------------------------------------------------------
  // work unit = one CTA (a 128-row tile) or one PAIR of CTAs (two adjacent row tiles; the second one may lie beyond
  // the job's rows: it then computes on zero / foreign rows and stores nothing)
  const int unit = PAIR ? (int)blockIdx.x >> 1 : (int)blockIdx.x;
  int ji = 0, ub = 0;
#pragma unroll 1
  for (int i = 0; i < L.num_jobs; ++i) {
    const int per = PAIR ? (L.job[i].m_tiles + 1) >> 1 : L.job[i].m_tiles;
    const int n_units = per * L.n_virtual;
    if (unit < ub + n_units || i == L.num_jobs - 1) { ji = i; break; }
    ub += n_units;
  }
  const ChainJob& J = L.job[ji];
  const int per_member = PAIR ? (J.m_tiles + 1) >> 1 : J.m_tiles;
  const int rel = unit - ub;
  const int bidx = rel / per_member;
  const int mt = PAIR ? 2 * (rel - bidx * per_member) + (int)rank : rel - bidx * per_member;
  const int v = L.member_list ? __ldg(L.member_list + bidx) : bidx;
  const int n_layers = J.n_layers;
  const int iters = J.iters;
  const int n_stages = L.stages;  // weight slots in use

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&J.mapA0);
    for (int i = 0; i < n_layers; ++i) tma_prefetch_desc(&J.mapW[i]);
    for (int s = 0; s < kMaxSlots; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_full[1], 1);
    for (int i = 0; i < 8; ++i) mbar_init(&blk_full[i], 4 * kCtas);      // one arrival per producing warp
    mbar_init(&drained[0], kEpiWarps * kCtas);
    mbar_init(&drained[1], kEpiWarps * kCtas);
    mbar_init(a0_full, 1);
    mbar_init(x0_ready, 4 * kCtas);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc_2sm(tmem_slot, 512); tmem_relinquish_2sm(); }
    else { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled to them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (kRebalanceRegs) {
    if (warp < kFirstEpiWarp) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsLight));
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsEpi));
  }
  // shared::cluster addresses of the leader's copies of the barriers both CTAs signal
  const uint32_t full_bar_l = PAIR ? mapa_u32(smem_u32(full_bar), 0) : smem_u32(full_bar);
  const uint32_t a0_full_l = PAIR ? mapa_u32(smem_u32(a0_full), 0) : smem_u32(a0_full);

  if (warp == 0) {
    // ========================================= TMA producer =========================================
    // (both CTAs of a pair: each streams its own half of every weight stage and its own X0 block; the bytes are counted on
    // the leader's barriers, which the leader arms with the total)
    if (lane == 0) {
      if (leader) mbar_arrive_expect_tx(a0_full, kActBlockBytes * kCtas);
      if (PAIR) tma_load_3d_2sm(x0s, &J.mapA0, a0_full_l, 0, J.row0 + mt * 128, v);
      else tma_load_3d(x0s, &J.mapA0, a0_full, 0, J.row0 + mt * 128, v);
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < iters; ++it) {
        for (int li = 0; li < n_layers; ++li) {
          const ChainLayer& Ly = J.layer[li];
          const int k_steps = Ly.k_blocks * (64 / kStgK);
          for (int n0 = 0; n0 < Ly.n_pad; n0 += kTileN) {
            const int ncols = min(kTileN, Ly.n_pad - n0);
            // columns of this MMA tile staged by THIS CTA: [c0, c0 + mycols). A narrow layer (64 columns) is issued as an
            // N = 128 product in pair mode: the second CTA's half lies outside the weight tensor and is zero-filled.
            const int mycols = PAIR ? max(ncols / 2, 64) : ncols;
            const int c0 = n0 + (int)rank * mycols;
            const uint32_t my_bytes = MODE == 0 ? (uint32_t)mycols * kStgK * 2
                                                : (uint32_t)(PAIR ? mycols : Ly.b_box_rows) * kStageK * 2;
            for (int ks = 0; ks < k_steps; ++ks) {
              mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sb = stage_base + stage * kSlotBytes;
              const uint32_t fb = full_bar_l + 8u * (uint32_t)stage;
#ifdef FQLPOP_DEBUG_SWITCHES
              if (L.debug_flags & 1) {  // (make TRACE=1 only) no weight stream: the MMAs run on whatever the stage holds
                if (leader) mbar_arrive(&full_bar[stage]);
              } else
#endif
              {
                if (leader) mbar_arrive_expect_tx(&full_bar[stage], my_bytes * kCtas);
                if (MODE == 0) {
                  // W[k][n], n contiguous: boxes of 64 n x 32 k
                  const int wb = Ly.b_box_rows;   // 64-column blocks per TMA instruction
                  for (int b = 0; b < mycols / 64; b += wb) {
                    if (PAIR) tma_load_4d_2sm(sb + b * kFwdBox, &J.mapW[li], fb, 0, ks * kStgK, (c0 >> 6) + b, v);
                    else tma_load_4d(sb + b * kFwdBox, &J.mapW[li], &full_bar[stage], 0, ks * kStgK, (c0 >> 6) + b, v);
                  }
                } else {
                  // the same array read as B^T: rows = output columns of this GEMM, 32 reduction elements per 64-byte row
                  if (PAIR) tma_load_3d_2sm(sb, &J.mapW[li], fb, ks * kStageK, c0, v);
                  else tma_load_3d(sb, &J.mapW[li], &full_bar[stage], ks * kStageK, c0, v);
                }
              }
              if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ========================================= MMA issuer ===========================================
    // The whole warp walks the loop (waits are warp-uniform); one elected lane issues the MMAs and commits. In a pair only
    // the leader CTA issues.
    if (leader) {
    int stage = 0;
    uint32_t phase = 0;
    int lc = 0;  // layers issued so far (every layer uses accumulator tile 0)
    int tc = 0;  // two-tile layers issued so far
    int pc = 0;  // ACT productions consumed so far
    const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
    const uint64_t bdesc0 = MODE == 0 ? make_smem_desc(smem_u32(stage_base), kFwdBox, 1024, 2)
                                      : make_smem_desc(smem_u32(stage_base), 16, 512, 4);
    constexpr uint32_t kBStep = (MODE == 0 ? 2048u : 32u) >> 4;  // descriptor units (16 B) per K = 16 step
#ifdef FQLPOP_EPI_TRACE
    // (make TRACE=1) per layer of CTA 0: start, cycles waited for drained / ACT blocks / weight stages, end
    const bool tr_on = L.trace && blockIdx.x == 0;
    long long w_dr = 0, w_blk = 0, w_wt = 0, t_begin = 0;
#define TR_T0() const long long tr_t0 = clock64()
#define TR_ACC(x) x += clock64() - tr_t0
#else
#define TR_T0()
#define TR_ACC(x)
#endif
    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li, ++lc) {
        const ChainLayer& Ly = J.layer[li];
        const bool from_act = !Ly.a_from_x0;
#ifdef FQLPOP_EPI_TRACE
        w_dr = w_blk = w_wt = 0;
        t_begin = clock64();
#endif
        if (!from_act) {
          TR_T0();
          if (lc == 0) mbar_wait(a0_full, 0);
          else mbar_wait(x0_ready, (it - 1) & 1);
          tc_fence_after();
          TR_ACC(w_blk);
        }
        const uint64_t adesc0 = make_smem_desc(from_act ? act_addr : x0_addr, 16, 1024, 2);
        const int nkb = Ly.k_blocks;
        const bool two = Ly.n_pad > kTileN;
        int tile = 0;
        for (int n0 = 0; n0 < Ly.n_pad; n0 += kTileN, ++tile) {
          const int ncols = min(kTileN, Ly.n_pad - n0);
          const uint32_t n_inst = PAIR ? (uint32_t)max(ncols, 128) : (uint32_t)ncols;
          const uint32_t idesc = make_idesc_bf16(PAIR ? 256 : 128, n_inst, 0, MODE == 0 ? 1 : 0);
          const uint32_t tmem_d = tmem_base + n0;
          // the accumulator tile must have been read by every epilogue warp (of both CTAs)
          {
            TR_T0();
            if (tile == 0) {
              if (lc > 0) mbar_wait(&drained[0], (lc - 1) & 1);
            } else if (tc > 0) {
              mbar_wait(&drained[1], (tc - 1) & 1);
            }
            tc_fence_after();
            TR_ACC(w_dr);
          }
          for (int kb = 0; kb < nkb; ++kb) {
            if (from_act && tile == 0) {
              TR_T0();
              mbar_wait(&blk_full[kb], pc & 1);
              tc_fence_after();
              TR_ACC(w_blk);
            }
            for (int h = 0; h < 64 / kStgK; ++h) {
              {
                TR_T0();
                mbar_wait(&full_bar[stage], phase);
                TR_ACC(w_wt);
              }
              tc_fence_after();
              if (elect_one()) {
                const uint64_t ad = adesc0 + (uint64_t)(kb * (kActBlockBytes >> 4) + h * 4);
                const uint64_t bd = bdesc0 + (uint64_t)(stage * (kSlotBytes >> 4));
                const uint32_t acc = (kb != 0 || h != 0) ? 1u : 0u;
                const bool last = kb == nkb - 1 && h == 64 / kStgK - 1;
                if (PAIR) {
                  umma_bf16_2sm(tmem_d, ad, bd, idesc, acc);
                  umma_bf16_2sm(tmem_d, ad + 2, bd + kBStep, idesc, 1u);
                  if (kDeep) {
                    umma_bf16_2sm(tmem_d, ad + 4, bd + 2 * kBStep, idesc, 1u);
                    umma_bf16_2sm(tmem_d, ad + 6, bd + 3 * kBStep, idesc, 1u);
                  }
                  umma_commit_2sm(&empty_bar[stage], 3);
                  if (last) umma_commit_2sm(&tmem_full[tile], 3);
                } else {
                  umma_bf16(tmem_d, ad, bd, idesc, acc);
                  umma_bf16(tmem_d, ad + 2, bd + kBStep, idesc, 1u);
                  umma_commit(&empty_bar[stage]);
                  if (last) umma_commit(&tmem_full[tile]);
                }
              }
              __syncwarp();
              if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
          }
        }
        if (from_act) ++pc;
        if (two) ++tc;
#ifdef FQLPOP_EPI_TRACE
        if (tr_on && lane == 0 && lc < 400) {
          long long* o = L.trace + 8 * lc;
          o[0] = t_begin; o[1] = w_dr; o[2] = w_blk; o[3] = w_wt; o[4] = clock64();
        }
#endif
      }
    }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ========================================= epilogue =============================================
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int cg = (warp - kFirstEpiWarp) >> 2;  // column group: columns [64 cg, 64 cg + 64) of every accumulator tile
    const int r = q * 32 + lane;         // row of the tile == TMEM lane
    const int grow = mt * 128 + r;       // row within the job
    const bool row_ok = grow < J.M;
    Epi C;
    C.tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    C.tmem_full = tmem_full;
    C.remote = PAIR && !leader;
    C.blk_full_a = C.remote ? mapa_u32(smem_u32(blk_full), 0) : smem_u32(blk_full);
    C.drained_a = C.remote ? mapa_u32(smem_u32(drained), 0) : smem_u32(drained);
    C.x0_ready_a = C.remote ? mapa_u32(smem_u32(x0_ready), 0) : smem_u32(x0_ready);
    C.act = smem_u32(act);
    C.params = J.params + (long long)v * J.S;
    C.ex = reinterpret_cast<float2*>(x0s);
    C.r = r; C.q = q; C.cg = cg; C.lane = lane; C.H = J.H;
    C.row_ok = row_ok;
    C.rs_off = (long long)v * J.pass_rs + grow;
    C.part_off = (long long)v * J.part_ms + (long long)(mt * 4 + q) * J.H;
    C.ln_eps = J.ln_eps;
    C.stg = L.staged ? smem_u32(stage_base + kStagesStaged * kStageBytes + (warp - kFirstEpiWarp) * kStgBytes) : 0u;
    const int grow0 = mt * 128 + q * 32;
    C.wrow = (long long)v * J.pass_ms + (long long)grow0 * J.H;
    C.wdz = (long long)v * J.dz_ms + (long long)grow0 * J.H;
    C.rows_valid = min(32, max(0, J.M - grow0));
    C.pb = nullptr;
    // the X0 block after the first layer: [0, 4 KB) LayerNorm row-sum exchange, then per column group two parameter
    // buffers of 3 x 128 floats (the Euler flow keeps re-using X0 as its first layer's operand: no buffers there)
    C.pb_base = iters == 1 ? reinterpret_cast<float*>(x0s + 4096) + cg * (2 * 3 * 128) : nullptr;
    C.tr = nullptr;
    C.map_h = nullptr; C.h_row = J.row0 + grow0; C.h_v = v; C.h_pending = false;
#ifdef FQLPOP_EPI_TRACE
    const bool ep_traced = L.trace && blockIdx.x == 0 && warp == kFirstEpiWarp && lane == 0;
#endif

    float xs[8];  // Euler flow: the action iterate of this row (fp32)
    if (MODE == 0 && J.xe) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        xs[j] = (row_ok && cg == 0 && j < J.A) ? J.xe[((long long)v * J.B + grow) * J.A + j] : 0.f;
    }

    int lc = 0, tc = 0;
    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li, ++lc) {
        const ChainLayer& Ly = J.layer[li];
        const uint32_t par0 = lc & 1, par1 = tc & 1;
#ifdef FQLPOP_EPI_TRACE
        C.tr = (ep_traced && lc < 400) ? L.trace + 4096 + 8 * lc : nullptr;
#endif
        stamp(C, 0);
        C.map_h = &J.mapH[li];
        if (Ly.kind == CL_HIDDEN) {
          const bool pub = !(it == iters - 1 && li == n_layers - 1);  // the last layer's output is consumed by nobody on chip
          if (MODE == 0) {
            if (Ly.out_gp != nullptr) fwd_hidden_layer<LN, true>(C, Ly, par0, par1, pub, lc & 1);
            else fwd_hidden_layer<LN, false>(C, Ly, par0, par1, pub, lc & 1);
          } else {
            bwd_hidden_layer<LN>(C, Ly, par0, par1, pub, lc & 1);
          }
          ++tc;
        } else if (Ly.kind == CL_EULER_HEAD) {
          // BC-flow head inside the Euler rollout: x <- x + v / F; the next step's input differs from this step's
          // only in the action columns and in t, which are patched in place in the X0 block.
          wait_full(C, 0, par0);  // every thread follows every phase of the barrier (parity waits must not skip one)
          if (cg == 0) {
            uint32_t rr[16];
            tmem_ld16(C.tmem_row, rr);
            tmem_ld_wait();
            const float invF = 1.f / (float)J.F;
            const bool last = it == iters - 1;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j < J.A) {
                xs[j] = fmaf(__uint_as_float(rr[j]) + __ldg(C.params + Ly.off_b + j), invF, xs[j]);
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
            tc_fence_before();
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) arrive_on(C, C.x0_ready_a);
          }
          arrive_drained(C, 0);
        } else {  // CL_DINPUT: gradient with respect to the (padded, <= 64 wide) network input, fp32
          wait_full(C, 0, par0);
          if (cg == 0) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (c * 32 < Ly.n_pad) {
                uint32_t rr[32];
                tmem_ld32(C.tmem_row + c * 32, rr);
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
          arrive_drained(C, 0);
        }
      }
    }
    // the h stores of the last layer complete before the CTA (and its shared memory) goes away
    if (C.h_pending && lane == 0) bulk_wait_group0();
  }

  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // no CTA leaves (or frees TMEM) while its peer may still signal it or run MMAs into it
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_2sm(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
  if (L.timeline && threadIdx.x == 0) {
    long long t1;
    unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    const unsigned long long slot = atomicAdd(reinterpret_cast<unsigned long long*>(L.timeline), 1ull);
    if (slot < 65536ull) {
      long long* rec = L.timeline + 8 + 4 * slot;
      rec[0] = tl_t0; rec[1] = t1; rec[2] = smid; rec[3] = L.tag | ((long long)blockIdx.x << 32);
    }
  }
}

}  // namespace

template <int MODE, bool LN, bool PAIR>
static cudaError_t set_smem() {
  return cudaFuncSetAttribute(fql_chain16_kernel<MODE, LN, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes16);
}
cudaError_t chain16_init() {
  cudaError_t e;
  if ((e = set_smem<0, false, false>()) != cudaSuccess) return e;
  if ((e = set_smem<0, true, false>()) != cudaSuccess) return e;
  if ((e = set_smem<1, false, false>()) != cudaSuccess) return e;
  if ((e = set_smem<1, true, false>()) != cudaSuccess) return e;
  if ((e = set_smem<0, false, true>()) != cudaSuccess) return e;
  if ((e = set_smem<0, true, true>()) != cudaSuccess) return e;
  if ((e = set_smem<1, false, true>()) != cudaSuccess) return e;
  return set_smem<1, true, true>();
}

// Pair mode (cta_group::2) is the default; FQLPOP_CHAIN_PAIR=0 runs the same kernel one CTA per row tile. Read once: the
// weight tensor maps of a backward chain are encoded for one mode or the other (pop.cu asks chain16_pair_mode()).
bool chain16_pair_mode() {
  static const bool on = [] { const char* e = getenv("FQLPOP_CHAIN_PAIR"); return e ? atoi(e) != 0 : true; }();
  return on;
}
// a job's row tiles are paired only when there is an even number of them (no half-empty pairs)
bool chain16_pairs_rows(int M) { return (((M + 127) / 128) & 1) == 0; }

// The 16-warp kernel covers the production shape: every hidden layer 512 wide, reductions of 64 (first layer) or 512,
// narrow layers (Euler head, input gradient) of one 64-column block.
bool chain16_eligible(const ChainLaunch& L) {
  const int ln = L.job[0].ln;
  for (int i = 0; i < L.num_jobs; ++i) {
    const ChainJob& j = L.job[i];
    if (j.H != 512 || j.ln != ln) return false;
    if (j.ln && j.iters > 1) return false;  // the LayerNorm row-sum exchange lives in the X0 block, which the Euler flow re-uses
    for (int l = 0; l < j.n_layers; ++l) {
      const ChainLayer& y = j.layer[l];
      if (y.kind == CL_HIDDEN && y.n_pad != 512) return false;
      if (y.kind != CL_HIDDEN && y.n_pad != 64) return false;
      if (y.a_from_x0 ? y.k_blocks != 1 : y.k_blocks != 8) return false;
      if (y.a_from_x0 && l != 0) return false;
    }
  }
  return true;
}
// ... whose forward weight maps were encoded with wide boxes (pop.cu wide_boxes)
static bool chain16_maps_ok(const ChainLaunch& L, int mode, bool pair) {
  if (mode != 0) return true;
  for (int i = 0; i < L.num_jobs; ++i) {
    const ChainJob& j = L.job[i];
    for (int l = 0; l < j.n_layers; ++l) {
      const int wb = j.layer[l].b_box_rows;
      const int per_cta = j.layer[l].n_pad >= 256 ? (pair ? 2 : 4) : 1;
      if (wb != per_cta) return false;   // (the box is also 64 k deep in pair mode, 32 otherwise: ChainBuilder::launch)
    }
  }
  return true;
}

template <int MODE, bool LN>
static cudaError_t launch_variant(ChainLaunch& L, bool pair, int units, cudaStream_t stream) {
  if (!pair) {
    fql_chain16_kernel<MODE, LN, false><<<units, kThreads16, kSmemBytes16, stream>>>(L);
    return cudaGetLastError();
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(2 * units);
  cfg.blockDim = dim3(kThreads16);
  cfg.dynamicSmemBytes = kSmemBytes16;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, fql_chain16_kernel<MODE, LN, true>, L);
}

cudaError_t launch_chain16(ChainLaunch& L, int mode, int grid, cudaStream_t stream) {
  (void)grid;
  const bool ln = L.job[0].ln != 0;
  bool pair = chain16_pair_mode();
  for (int i = 0; i < L.num_jobs; ++i) pair = pair && chain16_pairs_rows(L.job[i].M);
  int units = 0;
  for (int i = 0; i < L.num_jobs; ++i)
    units += (pair ? (L.job[i].m_tiles + 1) / 2 : L.job[i].m_tiles) * L.n_virtual;
  // weight slots: 16 KB each (8 KB in pair mode); launches that stage tensors through shared memory give up the last 32 KB
  if (!chain16_maps_ok(L, mode, pair)) return cudaErrorInvalidValue;
  if (mode == 0) L.stages = kStages;   // forward passes store what they save from registers: nothing is staged
  L.staged = L.stages != kStages;
  L.stages = (L.staged ? kStagesStaged : kStages) * ((pair && mode == 1) ? 2 : 1);
  if (mode == 0) return ln ? launch_variant<0, true>(L, pair, units, stream) : launch_variant<0, false>(L, pair, units, stream);
  return ln ? launch_variant<1, true>(L, pair, units, stream) : launch_variant<1, false>(L, pair, units, stream);
}

}  // namespace fqlpop
