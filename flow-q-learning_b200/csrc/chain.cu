// Layer-chain kernels (see chain.h for what they replace).
//
// CTA = 10 warps: warp 0 = TMA producer (streams every layer's bf16 weights, 16 KB stages = 32 k x 256 n), warp 1 =
// TMEM allocator + tcgen05.mma issuer (whole warp walks the loop, one elected lane issues), warps 2..9 = epilogue
// (two warps per TMEM lane quarter; a thread owns one row and 128 of the 256 columns of an accumulator tile).
// Shared memory: ACT (128 rows x 512 bf16 = the A operand of every hidden layer, K-major, 128-byte swizzle, eight
// 64-column blocks), X0 (128 x 64, the first layer's A operand, loaded by TMA), kStages weight stages, mbarriers.
// TMEM: 512 fp32 columns = one full 128 x 512 layer output = two 128 x 256 accumulator tiles.
//
// MMA shape: M = 128, N = 256, K = 16 (measured on B200: the 128 x 128 shape sustains 57% of the tensor peak from
// shared-memory operands, the 128 x 256 shape 96% -- profiles/ubench/chain_trace.py).
//
// Pipeline of one layer (tiles T0 = columns 0..255, T1 = 256..511):
//   tensor core :  T0 ........ T1 ........ | next layer: T0'(k 0..255) T0'(k 256..511) T1' ...
//   epilogue    :             pass 1 of T0 | flush T0 -> ACT[k 0..255]  pass 1 of T1 -> ACT[k 256..511]
//                             (parked)     |  `act_lo` ^                              `act_hi` ^
// Pass 1 of T0 overlaps the MMAs of T1; its bf16 results cannot go to ACT yet (T1's MMAs still read it), so they are
// parked in the TMEM columns they were drained from. Once T1 has landed, T0 is flushed to the low half of ACT and
// `act_lo` lets the next layer's MMAs start on it while T1's epilogue fills the high half. LayerNorm layers need the
// whole row before normalising: both tiles are parked, then flushed low half first.
//
// Global traffic of the epilogue (saved h / act'(z) / xhat of a forward pass, dz and the saved tensors of a backward
// pass): a thread owns a ROW, so a direct 16-byte access per lane touches 32 different 128-byte lines per warp
// instruction. Everything therefore goes through shared memory: h / dz are copied out of ACT (they are there anyway,
// as the next layer's A operand), the other tensors through a warp-private 4 KB staging block (32 rows x 128 bytes,
// same XOR swizzle as ACT, so no cross-warp synchronisation is needed), and the global side is always 4 rows x 128
// contiguous bytes per warp instruction (stores: ld.shared + st.global; loads: cp.async, issued a block ahead). The
// per-column sums of dz (bias gradients) fall out of the copy-out loop. Launches that need the staging area run on 3
// weight stages instead of 5 (the Euler chain keeps 5).
#include "chain.h"

#include "gemm.h"
#include "ptx_sm100.cuh"

namespace fqlpop {

namespace {

constexpr int kThreads = 320;
constexpr int kThreads2 = 352;  // cluster kernel: + one warp that forwards finished ACT blocks to the peer CTA
constexpr int kStages = 5;          // weight stages of a launch without saved tensors (Euler chain, target critic)
constexpr int kStagesStaged = 3;    // ... of a launch whose epilogue stages tensors through shared memory
constexpr int kStagingBytes = 4096; // per epilogue warp: 32 rows x 128 bytes (overlays weight stages 3 and 4)
constexpr int kActBlockBytes = 128 * 64 * 2;   // one 64-wide k-block of the A operand = 16 KB
constexpr int kActBytes = 8 * kActBlockBytes;  // 128 KB
constexpr int kStageBytes = 16384;             // 32 k x 256 n bf16
constexpr int kStageK = 32;
constexpr int kTileN = 256;
constexpr int kFwdBoxBytes = 64 * kStageK * 2;  // forward weight box: 64 n x 32 k (MN-major, 128-byte swizzle)
constexpr int kBarBytes = 256 + 1024;  // mbarriers + the cluster kernel's cross-CTA row-sum buffer (128 x float2)
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
__device__ __forceinline__ void ld_global_packed32(const __nv_bfloat16* src, uint32_t* w) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 u = __ldg(s4 + i);
    w[4 * i] = u.x; w[4 * i + 1] = u.y; w[4 * i + 2] = u.z; w[4 * i + 3] = u.w;
  }
}
// 4 of the 8 16-byte chunks of row r of a 128-byte-swizzled K-major block (chunks c0 .. c0 + 3) from 16 packed words
__device__ __forceinline__ void st_act_half_row(uint8_t* block, int r, int c0, const uint32_t* w) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(block + r * 128 + (((c0 + i) ^ (r & 7)) << 4)) =
        make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
}
// 32 columns starting at `col` (multiple of 32) of row r -> ACT
__device__ __forceinline__ void st_act32(uint8_t* act, int r, int col, const uint32_t* w) {
  st_act_half_row(act + (col >> 6) * kActBlockBytes, r, (col & 63) >> 3, w);
}
__device__ __forceinline__ f32x2 bf2_to_f2(uint32_t w) { return pk2u(w << 16, w & 0xFFFF0000u); }

// Per-thread view of the epilogue.
struct EpiCtx {
  uint32_t tmem_row;     // TMEM address of this thread's lane, column 0
  uint64_t* tmem_full;   // [2]
  uint64_t *act_lo, *act_hi;
  uint8_t* act;
  const float* params;   // this virtual member's fp32 parameter block
  int r, half, q, lane, H;
  bool row_ok;
  long long prow;        // element offset of this row in the pass buffers ([rows][H])
  long long rs_off;      // offset of this row in the rstd buffers
  long long dzrow;       // element offset of this row in the dz buffers
  long long part_off;    // offset of this (tile, lane quarter) row in the partial-sum buffers
  float ln_eps;
  // warp-level view for the staged global traffic
  uint8_t* stg;          // this warp's 4 KB staging block (nullptr in launches that keep all 5 weight stages)
  long long wrow;        // element offset of the warp's first row in the pass buffers
  long long wdz;         // ... in the dz buffers
  int rows_valid;        // how many of the warp's 32 rows exist (0 .. 32)
  // 2-CTA cluster variant (chain2): this CTA owns output columns [g_off, g_off + 256) of every hidden layer and
  // alternates between two 256-column accumulators (tm_off = 0 / 256); the single-CTA kernel keeps both at 0
  int tm_off, g_off;
  bool cl;               // cluster variant: publish every finished 64-column block of ACT (local barrier + copy to the peer)
  bool publish;          // ... unless nobody consumes this layer's output on chip
  uint64_t* blk_full;    // [8] local barriers: k-block j of ACT holds the next layer's A operand
  uint32_t peer_act, peer_blk_full;  // shared::cluster addresses of the peer's ACT / blk_full[0]
  float* pbuf;           // smem copy of the current layer's bias [512] / LayerNorm scale [512] / LayerNorm bias [512] (the X0
                         // block, idle after the first layer), or nullptr -> read them through L1
  mutable bool puse;     // the current layer reads its parameters from pbuf
  long long* tr;         // (debug) stamp log of the traced epilogue thread: tr[0] = count, stamps (clock, id) from tr[2]
};

// (debug) clock stamp with a site id; only the traced thread has C.tr != nullptr
// Compiled in only with -DFQLPOP_EPI_TRACE (make TRACE=1): even the dormant branch costs registers in the epilogue.
__device__ __forceinline__ void stamp(const EpiCtx& C, int id) {
#ifdef FQLPOP_EPI_TRACE
  if (C.tr) {
    const long long i = C.tr[0];
    if (i < 1900) {
      C.tr[2 + 2 * i] = clock64();
      C.tr[3 + 2 * i] = id;
      C.tr[0] = i + 1;
    }
  }
#endif
}

// CL = cluster kernel. The single-CTA kernel instantiates everything with CL = false: offsets fold to 0 and the
// publish hooks vanish at compile time.
template <bool CL> __device__ __forceinline__ int goff(const EpiCtx& C) { return CL ? C.g_off : 0; }
template <bool CL> __device__ __forceinline__ uint32_t tmoff(const EpiCtx& C) { return CL ? (uint32_t)C.tm_off : 0u; }

// Every bias / LayerNorm line of a layer is touched exactly once per CTA, i.e. each first use is a compulsory L1 miss
// (measured: 600-1800 cycles under load, 8 of them per thread and LayerNorm tile pass). The epilogue threads therefore
// pull the layer's parameter lines into L1 while they wait for the accumulator: thread t of the 256 takes line t % 16 of
// array t / 16 (0 bias, 1 LayerNorm scale, 2 LayerNorm bias). ncols = 512 (single CTA) or this CTA's 256 columns.
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
// four consecutive parameters: from the shared-memory copy (array arr: 0 bias, 1 LayerNorm scale, 2 LayerNorm bias; absolute
// column gcol) or through L1
__device__ __forceinline__ float4 ldp4(const EpiCtx& C, const float* g, int arr, int gcol, int i) {
  return C.puse ? *reinterpret_cast<const float4*>(C.pbuf + arr * 512 + gcol + 4 * i)
                : __ldg(reinterpret_cast<const float4*>(g + gcol) + i);
}
// Cooperative copy of the layer's parameter vectors into pbuf (all 256 epilogue threads; barrier before -- nobody still
// reads the previous layer's copy -- and after).
__device__ __forceinline__ void load_layer_params(const EpiCtx& C, const ChainLayer& Ly, bool fwd, bool ln) {
  const int t = (C.half * 4 + C.q) * 32 + C.lane;
  epi_bar_sync();
  float4* dst = reinterpret_cast<float4*>(C.pbuf);
#pragma unroll 1
  for (int i = t; i < 384; i += 256) {
    const int arr = i >> 7, k = i & 127;
    const long long off = arr == 0 ? Ly.off_b : (arr == 1 ? Ly.off_g : Ly.off_beta);
    const bool need = fwd ? (arr == 0 || ln) : (ln && arr == 1);
    if (need) dst[i] = __ldg(reinterpret_cast<const float4*>(C.params + off) + k);
  }
  epi_bar_sync();
}
template <bool CL>
__device__ __forceinline__ void prefetch_layer_params(const EpiCtx& C, const ChainLayer& Ly, bool fwd, bool ln) {
  const int t = (C.half * 4 + C.q) * 32 + C.lane;
  const int line = t & 15, which = t >> 4;
  const int ncols = CL ? 256 : Ly.n_pad;
  if (line * 32 >= ncols) return;
  const float* base = C.params + goff<CL>(C) + line * 32;
  if (fwd) {
    if (which == 0) prefetch_l1(base + Ly.off_b);
    else if (ln && which == 1) prefetch_l1(base + Ly.off_g);
    else if (ln && which == 2) prefetch_l1(base + Ly.off_beta);
  } else if (ln && which == 0) {
    prefetch_l1(base + Ly.off_g);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Staged global traffic. A "block" is 32 rows x 64 bf16 columns = 32 x 128 bytes in shared memory, 16-byte chunk c
// of row r at r * 128 + ((c ^ (r & 7)) << 4) -- the layout of a warp's 32 rows inside a 64-column block of ACT, and
// of the warp-private staging block. Row-wise accesses (lane = row) and the coalesced sweep below (a quarter-warp per
// row) are both bank-conflict free.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_stage16(uint8_t* stg, int lane, int cw, const uint32_t* w) {
  uint8_t* row = stg + lane * 128;
  const int c0 = cw >> 3;
  *reinterpret_cast<uint4*>(row + ((c0 ^ (lane & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
  *reinterpret_cast<uint4*>(row + (((c0 + 1) ^ (lane & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
}
__device__ __forceinline__ void ld_stage16(const uint8_t* stg, int lane, int cw, uint32_t* w) {
  const uint8_t* row = stg + lane * 128;
  const int c0 = cw >> 3;
  const uint4 a = *reinterpret_cast<const uint4*>(row + ((c0 ^ (lane & 7)) << 4));
  const uint4 b = *reinterpret_cast<const uint4*>(row + (((c0 + 1) ^ (lane & 7)) << 4));
  w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}
// block -> global rows (row stride ld elements); 4 rows x 128 contiguous bytes per warp instruction
__device__ __forceinline__ void copy_out_block(const uint8_t* blk, __nv_bfloat16* g, int ld, int lane, int rows_valid) {
  const int ch = lane & 7, rsub = lane >> 3;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i * 4 + rsub;
    const uint4 v = *reinterpret_cast<const uint4*>(blk + r * 128 + ((ch ^ (r & 7)) << 4));
    if (r < rows_valid) *reinterpret_cast<uint4*>(g + (long long)r * ld + ch * 8) = v;
  }
}
// the same sweep, also summing every column over the warp's 32 rows (rows that do not exist hold zeros):
// lanes 0..7 write 8 column sums each
__device__ __forceinline__ void copy_out_block_colsum(const uint8_t* blk, __nv_bfloat16* g, int ld, int lane,
                                                      int rows_valid, float* part) {
  const int ch = lane & 7, rsub = lane >> 3;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i * 4 + rsub;
    const uint4 v = *reinterpret_cast<const uint4*>(blk + r * 128 + ((ch ^ (r & 7)) << 4));
    if (r < rows_valid) *reinterpret_cast<uint4*>(g + (long long)r * ld + ch * 8) = v;
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
// global rows -> block, asynchronously (cp.async; rows that do not exist are zero-filled). One commit group.
__device__ __forceinline__ void stage_in_block(uint8_t* blk, const __nv_bfloat16* g, int ld, int lane, int rows_valid) {
  const int ch = lane & 7, rsub = lane >> 3;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = i * 4 + rsub;
    uint8_t* dst = blk + r * 128 + ((ch ^ (r & 7)) << 4);
    if (r < rows_valid) cp_async16(dst, g + (long long)r * ld + ch * 8);
    else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
  }
  cp_async_commit();
}
// Half blocks (32 rows x 32 columns = 32 x 64 bytes, 2 KB): chunk c (0..3) of row r at r * 64 + ((c ^ ((r >> 1) & 3))
// << 4). Used where two tensors have to be resident at once (LayerNorm backward).
__device__ __forceinline__ void stage_in_half(uint8_t* blk, const __nv_bfloat16* g, int ld, int lane, int rows_valid) {
  const int ch = lane & 3, rsub = lane >> 2;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = i * 8 + rsub;
    uint8_t* dst = blk + r * 64 + ((ch ^ ((r >> 1) & 3)) << 4);
    if (r < rows_valid) cp_async16(dst, g + (long long)r * ld + ch * 8);
    else *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
  }
  cp_async_commit();
}
__device__ __forceinline__ void ld_half16(const uint8_t* blk, int lane, int cw, uint32_t* w) {
  const uint8_t* row = blk + lane * 64;
  const int c0 = cw >> 3, sw = (lane >> 1) & 3;
  const uint4 a = *reinterpret_cast<const uint4*>(row + ((c0 ^ sw) << 4));
  const uint4 b = *reinterpret_cast<const uint4*>(row + (((c0 + 1) ^ sw) << 4));
  w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}

__device__ __forceinline__ void wait_tile(const EpiCtx& C, uint32_t& fullph, int t) {
  mbar_wait(&C.tmem_full[t], (fullph >> t) & 1u);  // bit t of fullph: parity to wait for on tmem_full[t]
  fullph ^= 1u << t;
  tc_fence_after();
}
// make the generic-proxy writes to ACT / the TMEM reads visible to the MMA warp, then signal
__device__ __forceinline__ void release(uint64_t* bar) {
  tc_fence_before();
  fence_proxy_async_smem();
  mbar_arrive(bar);
}
// number of 32-column chunks of this thread inside accumulator tile `tile` (0 .. 4)
__device__ __forceinline__ int tile_chunks(const EpiCtx& C, int tile, int n_pad) {
  const int left = n_pad - (tile * kTileN + C.half * 128);
  return left <= 0 ? 0 : (left >= 128 ? 4 : left >> 5);
}

// ---------------------------------------------------------------------------------------------------------------
// 16-column chunk helpers. A thread walks its 128 columns of an accumulator tile in chunks of 16 (register budget:
// 168 per thread at 10 warps), two chunks per loop trip so the TMEM load of the next chunk is always in flight.
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int tile_chunks16(const EpiCtx& C, int tile, int n_pad) {
  const int left = n_pad - (tile * kTileN + C.half * 128);
  return left <= 0 ? 0 : (left >= 128 ? 8 : left >> 4);  // n_pad is a multiple of 64 -> a multiple of 4
}
__device__ __forceinline__ void st_global_packed16(__nv_bfloat16* dst, const uint32_t* w) {
  uint4* d4 = reinterpret_cast<uint4*>(dst);
  d4[0] = make_uint4(w[0], w[1], w[2], w[3]);
  d4[1] = make_uint4(w[4], w[5], w[6], w[7]);
}
__device__ __forceinline__ void ld_global_packed16(const __nv_bfloat16* src, uint32_t* w) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  const uint4 a = __ldg(s4), b = __ldg(s4 + 1);
  w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
  w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
}
// 16 columns starting at `col` (multiple of 16) of row r -> ACT (two 16-byte chunks of the swizzled row)
__device__ __forceinline__ void st_act16(uint8_t* act, int r, int col, const uint32_t* w) {
  uint8_t* row = act + (col >> 6) * kActBlockBytes + r * 128;
  const int c0 = (col & 63) >> 3;
  *reinterpret_cast<uint4*>(row + ((c0 ^ (r & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
  *reinterpret_cast<uint4*>(row + (((c0 + 1) ^ (r & 7)) << 4)) = make_uint4(w[4], w[5], w[6], w[7]);
}
// lanes 0..15 end up with the sum over the warp's 32 rows of column (lane & 15) of a 32 x 16 block
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
// column sums of the bf16-rounded values (the numbers the weight-gradient GEMM consumes) -> per-quarter partials
__device__ __forceinline__ void store_colsum16(float* part, const EpiCtx& C, int col, const uint32_t* pk) {
  float v[16];
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[2 * i] = bf16_lo(pk[i]); v[2 * i + 1] = bf16_hi(pk[i]); }
  const float cs = warp_colsum16(v, C.lane);
  if (C.lane < 16) part[C.part_off + col + C.lane] = cs;
}

__device__ __forceinline__ void publish_block(const EpiCtx& C, int gblk);

// This warp's part of accumulator tile `tile` (its 32 rows x its 128 columns = two 64-column blocks of ACT) -> global
// rows of `dst` ([rows][H] bf16, already offset to the job's row 0 of virtual member 0), optionally with the per-column
// sums over the warp's rows -> part[part_off + column].
template <bool CL>
__device__ __forceinline__ void copy_out_act(const EpiCtx& C, int tile, int n_pad, __nv_bfloat16* dst, long long wofs,
                                             float* part) {
  const int nblk = tile_chunks16(C, tile, n_pad) >> 2;
  const int col0 = tile * kTileN + C.half * 128;
  for (int b = 0; b < nblk; ++b) {
    const int col = goff<CL>(C) + col0 + b * 64;
    const uint8_t* blk = C.act + (col >> 6) * kActBlockBytes + C.q * (32 * 128);
    if (part) copy_out_block_colsum(blk, dst + wofs + col, C.H, C.lane, C.rows_valid, part + C.part_off + col);
    else copy_out_block(blk, dst + wofs + col, C.H, C.lane, C.rows_valid);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// forward hidden layer: h = LayerNorm?(gelu(acc + b))
// ---------------------------------------------------------------------------------------------------------------
// Pass 1 over this thread's chunks of one accumulator tile: bias + GELU (+ act'(z) for the backward pass, staged and
// copied out per 64-column block); the bf16 result is parked in the drained TMEM columns or -- `direct` -- written
// to ACT (the caller copies h out of ACT once the layer has been handed to the tensor core).
__device__ __forceinline__ void ld_bias16(const EpiCtx& C, const float* bias, int gcol, float4 (&b4)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) b4[i] = ldp4(C, bias, 0, gcol, i);
}
template <bool LN, bool SAVE, bool CL>
__device__ __forceinline__ void fwd_chunk(const EpiCtx& C, const float4 (&b4)[4], const uint32_t (&rr)[16], int col0, int c,
                                          bool direct, float& s1, float& s2) {
  const int col = col0 + c * 16;
  uint32_t pk[8], gpk[8];
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
  if (SAVE) st_stage16(C.stg, C.lane, (c & 3) * 16, gpk);
  if (direct) st_act16(C.act, C.r, goff<CL>(C) + col, pk);
  else tmem_st8(C.tmem_row + tmoff<CL>(C) + col0 + c * 8, pk);
}
// DB: the accumulator chunks are double-buffered in registers -- the TMEM load of chunk c + 1 is in flight while chunk c
// is being computed (tcgen05.wait::ld waits for every outstanding load, so the next one is issued right after the
// wait). Pays off in the cluster kernel (one pass per layer, nothing else to overlap with); the single-CTA kernel,
// whose passes already overlap the other tile's MMAs, is faster with the plain loop (fewer live registers).
template <bool LN, bool SAVE, bool DB, bool CL = DB>
__device__ __forceinline__ void fwd_pass1(const EpiCtx& C, const ChainLayer& Ly, int tile, bool direct, float& s1, float& s2) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);  // a multiple of 4
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
  const float* bias = C.params + Ly.off_b;
  const int g0 = goff<CL>(C) + col0;  // absolute column of this thread's first chunk
  const uint32_t tbase = C.tmem_row + tmoff<CL>(C) + col0;
  uint32_t ra[16], rb[16];
  if (DB) tmem_ld16(tbase, ra);
#pragma unroll 1
  for (int cb = 0; cb < nch; cb += 4) {
    if (DB) {
#pragma unroll 1
      for (int c = cb; c < cb + 4; c += 2) {
        float4 b4[4];
        ld_bias16(C, bias, g0 + c * 16, b4);
        tmem_ld_wait();
        tmem_ld16(tbase + (c + 1) * 16, rb);
        fwd_chunk<LN, SAVE, CL>(C, b4, ra, col0, c, direct, s1, s2);
        ld_bias16(C, bias, g0 + (c + 1) * 16, b4);
        tmem_ld_wait();
        if (c + 2 < nch) tmem_ld16(tbase + (c + 2) * 16, ra);
        fwd_chunk<LN, SAVE, CL>(C, b4, rb, col0, c + 1, direct, s1, s2);
      }
    } else {
#pragma unroll 1
      for (int c = cb; c < cb + 4; ++c) {
        float4 b4[4];
        ld_bias16(C, bias, g0 + c * 16, b4);
        tmem_ld16(tbase + c * 16, ra);
        tmem_ld_wait();
        fwd_chunk<LN, SAVE, CL>(C, b4, ra, col0, c, direct, s1, s2);
      }
    }
    if (SAVE) {
      __syncwarp();
      copy_out_block(C.stg, Ly.out_gp + C.wrow + goff<CL>(C) + col0 + cb * 16, C.H, C.lane, C.rows_valid);
      __syncwarp();
    }
    if (CL && direct) publish_block(C, (goff<CL>(C) + col0 + cb * 16) >> 6);
  }
}

// Parked tile -> ACT; LayerNorm layers normalise on the way: x = (g - mean) rstd, h = x gamma + beta (xhat staged and
// copied out per 64-column block).
template <bool LN, bool SAVE, bool CL>
__device__ __forceinline__ void fwd_flush(const EpiCtx& C, const ChainLayer& Ly, int tile, float mean, float rstd) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
  const float shift = -mean * rstd;
  const bool save_x = LN && SAVE && Ly.out_xhat != nullptr;
  uint32_t pr[8];
#pragma unroll 1
  for (int cb = 0; cb < nch; cb += 4) {
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      const int c = cb + c4;
      const int col = col0 + c * 16;
      tmem_ld8(C.tmem_row + tmoff<CL>(C) + col0 + c * 8, pr);
      tmem_ld_wait();
      if (LN) {
        const int gcol = goff<CL>(C) + col;
        uint32_t hk[8], xk[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 ga = ldp4(C, C.params + Ly.off_g, 1, gcol, i), be = ldp4(C, C.params + Ly.off_beta, 2, gcol, i);
          const uint32_t w0 = pr[2 * i], w1 = pr[2 * i + 1];
          const float x0 = fmaf(bf16_lo(w0), rstd, shift), x1 = fmaf(bf16_hi(w0), rstd, shift);
          const float x2 = fmaf(bf16_lo(w1), rstd, shift), x3 = fmaf(bf16_hi(w1), rstd, shift);
          xk[2 * i] = pack_bf16x2(x0, x1);
          xk[2 * i + 1] = pack_bf16x2(x2, x3);
          hk[2 * i] = pack_bf16x2(fmaf(x0, ga.x, be.x), fmaf(x1, ga.y, be.y));
          hk[2 * i + 1] = pack_bf16x2(fmaf(x2, ga.z, be.z), fmaf(x3, ga.w, be.w));
        }
        st_act16(C.act, C.r, goff<CL>(C) + col, hk);
        if (save_x) st_stage16(C.stg, C.lane, c4 * 16, xk);
      } else {
        st_act16(C.act, C.r, goff<CL>(C) + col, pr);
      }
    }
    if (save_x) {
      __syncwarp();
      copy_out_block(C.stg, Ly.out_xhat + C.wrow + goff<CL>(C) + col0 + cb * 16, C.H, C.lane, C.rows_valid);
      __syncwarp();
    }
    if (CL) publish_block(C, (goff<CL>(C) + col0 + cb * 16) >> 6);
  }
}

// The two halves of a row are owned by two threads (column halves of every tile): exchange partial sums through the
// first bytes of ACT, which is free once every MMA of the layer has retired.
__device__ __forceinline__ void exchange_row_sums(const EpiCtx& C, float& t1, float& t2) {
  float2* ex = reinterpret_cast<float2*>(C.act);
  ex[C.half * 128 + C.r] = make_float2(t1, t2);
  epi_bar_sync();
  const float2 o = ex[(C.half ^ 1) * 128 + C.r];
  epi_bar_sync();
  t1 += o.x;
  t2 += o.y;
}

template <bool LN, bool SAVE>
__device__ __forceinline__ void fwd_hidden_layer(const EpiCtx& C, uint32_t& fullph, const ChainLayer& Ly) {
  constexpr bool CL = false;
  const int n_tiles = (Ly.n_pad + kTileN - 1) / kTileN;
  float s1 = 0.f, s2 = 0.f;
  C.puse = C.pbuf != nullptr && !Ly.a_from_x0;
  if (C.puse) load_layer_params(C, Ly, true, LN);
  else prefetch_layer_params<CL>(C, Ly, true, LN);
  if (!LN) {
    wait_tile(C, fullph, 0);
    stamp(C, 10);
    if (n_tiles == 1) {
      fwd_pass1<false, SAVE, false>(C, Ly, 0, true, s1, s2);
      stamp(C, 11);
      release(C.act_lo);
      stamp(C, 12);
    } else {
      fwd_pass1<false, SAVE, false>(C, Ly, 0, false, s1, s2);
      stamp(C, 13);
      tmem_st_wait();
      stamp(C, 14);
      wait_tile(C, fullph, 1);  // every MMA of the layer has retired: ACT may be overwritten
      fwd_flush<false, SAVE, CL>(C, Ly, 0, 0.f, 1.f);
      stamp(C, 15);
      release(C.act_lo);
      stamp(C, 16);
      fwd_pass1<false, SAVE, false>(C, Ly, 1, true, s1, s2);
      stamp(C, 17);
    }
    release(C.act_hi);
    stamp(C, 18);
    // h -> HBM out of ACT, off the hand-off path: these rows of ACT are next written by this very warp
    if (Ly.out_h) {
      __syncwarp();
      for (int t = 0; t < n_tiles; ++t) copy_out_act<CL>(C, t, Ly.n_pad, Ly.out_h, C.wrow, nullptr);
    }
  } else {
    wait_tile(C, fullph, 0);
    stamp(C, 19);
    fwd_pass1<true, SAVE, false>(C, Ly, 0, false, s1, s2);
    stamp(C, 20);
    if (n_tiles > 1) {
      wait_tile(C, fullph, 1);
      stamp(C, 21);
      fwd_pass1<true, SAVE, false>(C, Ly, 1, false, s1, s2);
      stamp(C, 22);
    }
    tmem_st_wait();
    stamp(C, 23);
    float t1 = s1, t2 = s2;
    exchange_row_sums(C, t1, t2);
    stamp(C, 24);
    const float inv_n = 1.0f / (float)C.H;
    const float mean = t1 * inv_n;
    const float var = fmaxf(t2 * inv_n - mean * mean, 0.f);  // flax LayerNorm: E[x^2] - E[x]^2
    const float rstd = rsqrtf(var + C.ln_eps);
    if (C.row_ok && C.half == 0 && Ly.out_rstd) Ly.out_rstd[C.rs_off] = rstd;
    // LayerNorm layers copy h out BEFORE the hand-off: the next layer's row-sum exchange reuses the head of ACT and
    // must not overtake a slower warp's copy
    fwd_flush<true, SAVE, CL>(C, Ly, 0, mean, rstd);
    stamp(C, 25);
    if (Ly.out_h) { __syncwarp(); copy_out_act<CL>(C, 0, Ly.n_pad, Ly.out_h, C.wrow, nullptr); }
    release(C.act_lo);
    stamp(C, 26);
    if (n_tiles > 1) {
      fwd_flush<true, SAVE, CL>(C, Ly, 1, mean, rstd);
      stamp(C, 27);
      if (Ly.out_h) { __syncwarp(); copy_out_act<CL>(C, 1, Ly.n_pad, Ly.out_h, C.wrow, nullptr); }
    }
    release(C.act_hi);
    stamp(C, 28);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// backward hidden layer: dz = act'(z) * LNbwd?(acc), acc = dz_next W^T
// ---------------------------------------------------------------------------------------------------------------
// act'(z) of block `blk` (64 columns) of this warp's part of tile `tile` -> staging (asynchronous)
template <bool CL>
__device__ __forceinline__ void bwd_prefetch_gp(const EpiCtx& C, const ChainLayer& Ly, int tile, int blk) {
  if (!Ly.gp) return;
  if (blk * 4 >= tile_chunks16(C, tile, Ly.n_pad)) return;
  stage_in_block(C.stg, Ly.gp + C.wrow + goff<CL>(C) + tile * kTileN + C.half * 128 + blk * 64, C.H, C.lane, C.rows_valid);
}

// no LayerNorm: one pass, parked or direct exactly like the forward layer. The first block of act'(z) has been
// prefetched by the caller; `next_tile` >= 0 prefetches that tile's first block on the way out.
template <bool CL>
__device__ __forceinline__ void bwd_pass1(const EpiCtx& C, const ChainLayer& Ly, int tile, bool direct, int next_tile) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
  const bool has_gp = Ly.gp != nullptr;
  uint32_t rr[16];
#pragma unroll 1
  for (int cb = 0; cb < nch; cb += 4) {
    if (has_gp) {
      cp_async_wait_all();
      __syncwarp();
    }
#pragma unroll 1
    for (int c4 = 0; c4 < 4; ++c4) {
      const int c = cb + c4;
      const int col = col0 + c * 16;
      uint32_t gw[8];
      if (has_gp) ld_stage16(C.stg, C.lane, c4 * 16, gw);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) gw[i] = C.row_ok ? 0x3F803F80u : 0u;  // bf16 (1, 1) / (0, 0)
      }
      tmem_ld16(C.tmem_row + tmoff<CL>(C) + col, rr);
      tmem_ld_wait();
      uint32_t pk[8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        pk[i] = pack_bf16x2(__uint_as_float(rr[2 * i]) * bf16_lo(gw[i]),
                            __uint_as_float(rr[2 * i + 1]) * bf16_hi(gw[i]));
      if (direct) st_act16(C.act, C.r, goff<CL>(C) + col, pk);
      else tmem_st8(C.tmem_row + tmoff<CL>(C) + col0 + c * 8, pk);
    }
    if (has_gp) {
      __syncwarp();
      if (cb + 4 < nch) bwd_prefetch_gp<CL>(C, Ly, tile, (cb >> 2) + 1);
      else if (next_tile >= 0) bwd_prefetch_gp<CL>(C, Ly, next_tile, 0);
    }
    if (CL && direct) publish_block(C, (goff<CL>(C) + col0 + cb * 16) >> 6);
  }
}
template <bool CL>
__device__ __forceinline__ void bwd_flush(const EpiCtx& C, const ChainLayer& Ly, int tile) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
#pragma unroll 1
  for (int cp = 0; cp < nch; cp += 4) {
    uint32_t pr[4][8];
#pragma unroll
    for (int j = 0; j < 4; ++j) tmem_ld8(C.tmem_row + tmoff<CL>(C) + col0 + (cp + j) * 8, pr[j]);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 4; ++j) st_act16(C.act, C.r, goff<CL>(C) + col0 + (cp + j) * 16, pr[j]);
  }
}

// LayerNorm backward works on half blocks (32 columns): the staging block holds two of them.
// sub-block `sb` (0 .. 3 within the tile's 128 columns of this warp) of a saved tensor -> staging slot `slot`
template <bool CL>
__device__ __forceinline__ void ln_prefetch(const EpiCtx& C, const __nv_bfloat16* src, int n_pad, int tile, int sb, int slot) {
  if (sb * 2 >= tile_chunks16(C, tile, n_pad)) return;
  stage_in_half(C.stg + slot * 2048, src + C.wrow + goff<CL>(C) + tile * kTileN + C.half * 128 + sb * 32, C.H, C.lane, C.rows_valid);
}

// LayerNorm, statistics pass (overlaps the MMAs of the other tile; the accumulator stays in TMEM untouched):
// h = xhat gamma + beta, acc = dL/dh  ->  dxhat = acc gamma; s1 = sum dxhat, s2 = sum dxhat xhat;
// partial column sums dbeta = colsum(acc), dgamma = colsum(acc xhat). xhat is double-buffered through the two
// staging slots; sub-block 0 has been prefetched into slot 0 by the caller.
template <bool CL>
__device__ __forceinline__ void bwd_ln_stats(const EpiCtx& C, const ChainLayer& Ly, int tile, int next_tile, float& s1, float& s2) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
  const int nsb = nch >> 1;
  uint32_t rr[16];
#pragma unroll 1
  for (int sb = 0; sb < nsb; ++sb) {
    cp_async_wait_all();
    __syncwarp();
    if (sb + 1 < nsb) ln_prefetch<CL>(C, Ly.xhat, Ly.n_pad, tile, sb + 1, (sb + 1) & 1);
    else if (next_tile >= 0) ln_prefetch<CL>(C, Ly.xhat, Ly.n_pad, next_tile, 0, (sb + 1) & 1);
    const uint8_t* slot = C.stg + (sb & 1) * 2048;
#pragma unroll 1
    for (int ch = 0; ch < 2; ++ch) {
      const int col = col0 + sb * 32 + ch * 16;
      uint32_t xw[8];
      ld_half16(slot, C.lane, ch * 16, xw);
      const int gcol = goff<CL>(C) + col;
      tmem_ld16(C.tmem_row + tmoff<CL>(C) + col, rr);
      tmem_ld_wait();
      float t0[16], t1[16];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 ga = ldp4(C, C.params + Ly.off_g, 1, gcol, i);
        const float gam[4] = {ga.x, ga.y, ga.z, ga.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float dh = C.row_ok ? __uint_as_float(rr[4 * i + e]) : 0.f;
          const uint32_t w = xw[2 * i + (e >> 1)];
          const float xh = (e & 1) ? bf16_hi(w) : bf16_lo(w);
          const float dxh = dh * gam[e];
          s1 += dxh;
          s2 = fmaf(dxh, xh, s2);
          if (Ly.dbeta_part) {
            t0[4 * i + e] = dh;
            t1[4 * i + e] = dh * xh;
          }
        }
      }
      if (Ly.dbeta_part) {
        const float cb = warp_colsum16(t0, C.lane);
        const float cg = warp_colsum16(t1, C.lane);
        if (C.lane < 16) {
          Ly.dbeta_part[C.part_off + goff<CL>(C) + col + C.lane] = cb;
          Ly.dgamma_part[C.part_off + goff<CL>(C) + col + C.lane] = cg;
        }
      }
    }
  }
}
// LayerNorm, output pass: dg = rstd (dxhat - mean(dxhat) - xhat mean(dxhat xhat)); dz = act'(z) dg -> ACT.
// Slot 0 = xhat, slot 1 = act'(z) of the current 32-column sub-block.
template <bool CL>
__device__ __forceinline__ void bwd_ln_out(const EpiCtx& C, const ChainLayer& Ly, int tile, float c1, float c2, float rstd) {
  const int nch = tile_chunks16(C, tile, Ly.n_pad);
  if (nch == 0) return;
  const int col0 = tile * kTileN + C.half * 128;
  const float k1 = -c1 * rstd, k2 = -c2 * rstd;
  const int nsb = nch >> 1;
  const bool has_gp = Ly.gp != nullptr;
  uint32_t rr[16];
#pragma unroll 1
  for (int sb = 0; sb < nsb; ++sb) {
    __syncwarp();  // the previous sub-block has been read by every lane
    ln_prefetch<CL>(C, Ly.xhat, Ly.n_pad, tile, sb, 0);
    if (has_gp) ln_prefetch<CL>(C, Ly.gp, Ly.n_pad, tile, sb, 1);
    cp_async_wait_all();
    __syncwarp();
#pragma unroll 1
    for (int ch = 0; ch < 2; ++ch) {
      const int col = col0 + sb * 32 + ch * 16;
      uint32_t xw[8], gw[8];
      ld_half16(C.stg, C.lane, ch * 16, xw);
      if (has_gp) ld_half16(C.stg + 2048, C.lane, ch * 16, gw);
      else {
#pragma unroll
        for (int i = 0; i < 8; ++i) gw[i] = C.row_ok ? 0x3F803F80u : 0u;
      }
      const int gcol = goff<CL>(C) + col;
      tmem_ld16(C.tmem_row + tmoff<CL>(C) + col, rr);
      tmem_ld_wait();
      uint32_t pk[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 ga = ldp4(C, C.params + Ly.off_g, 1, gcol, i);
        const float gam[4] = {ga.x, ga.y, ga.z, ga.w};
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
      st_act16(C.act, C.r, goff<CL>(C) + col, pk);
    }
    if (CL && (sb & 1)) publish_block(C, (goff<CL>(C) + col0 + sb * 32) >> 6);
  }
}

template <bool LN>
__device__ __forceinline__ void bwd_hidden_layer(const EpiCtx& C, uint32_t& fullph, const ChainLayer& Ly) {
  constexpr bool CL = false;
  const int n_tiles = (Ly.n_pad + kTileN - 1) / kTileN;
  C.puse = LN && C.pbuf != nullptr && !Ly.a_from_x0;
  if (C.puse) load_layer_params(C, Ly, false, LN);
  else prefetch_layer_params<CL>(C, Ly, false, LN);
  if (!LN) {
    bwd_prefetch_gp<CL>(C, Ly, 0, 0);
    wait_tile(C, fullph, 0);
    stamp(C, 10);
    if (n_tiles == 1) {
      bwd_pass1<CL>(C, Ly, 0, true, -1);
      stamp(C, 11);
      release(C.act_lo);
      stamp(C, 12);
    } else {
      bwd_pass1<CL>(C, Ly, 0, false, 1);
      stamp(C, 13);
      tmem_st_wait();
      stamp(C, 14);
      wait_tile(C, fullph, 1);
      stamp(C, 15);
      bwd_flush<CL>(C, Ly, 0);
      stamp(C, 16);
      release(C.act_lo);
      stamp(C, 17);
      bwd_pass1<CL>(C, Ly, 1, true, -1);
      stamp(C, 18);
    }
    release(C.act_hi);
    stamp(C, 19);
    // dz -> HBM (+ bias-gradient partial sums) out of ACT, off the hand-off path
    __syncwarp();
    for (int t = 0; t < n_tiles; ++t) copy_out_act<CL>(C, t, Ly.n_pad, Ly.out_dz, C.wdz, Ly.dbias_part);
  } else {
    float s1 = 0.f, s2 = 0.f;
    ln_prefetch<CL>(C, Ly.xhat, Ly.n_pad, 0, 0, 0);
    wait_tile(C, fullph, 0);
    stamp(C, 20);
    bwd_ln_stats<CL>(C, Ly, 0, n_tiles > 1 ? 1 : -1, s1, s2);
    stamp(C, 21);
    if (n_tiles > 1) {
      wait_tile(C, fullph, 1);
      stamp(C, 22);
      bwd_ln_stats<CL>(C, Ly, 1, -1, s1, s2);
      stamp(C, 23);
    }
    float t1 = s1, t2 = s2;
    exchange_row_sums(C, t1, t2);
    stamp(C, 24);
    const float inv_n = 1.0f / (float)C.H;
    const float rstd = C.row_ok ? __ldg(Ly.rstd + C.rs_off) : 0.f;
    // copies happen before the hand-off (see fwd_hidden_layer)
    bwd_ln_out<CL>(C, Ly, 0, t1 * inv_n, t2 * inv_n, rstd);
    stamp(C, 25);
    __syncwarp();
    copy_out_act<CL>(C, 0, Ly.n_pad, Ly.out_dz, C.wdz, Ly.dbias_part);
    release(C.act_lo);
    stamp(C, 26);
    if (n_tiles > 1) {
      bwd_ln_out<CL>(C, Ly, 1, t1 * inv_n, t2 * inv_n, rstd);
      stamp(C, 27);
      __syncwarp();
      copy_out_act<CL>(C, 1, Ly.n_pad, Ly.out_dz, C.wdz, Ly.dbias_part);
    }
    release(C.act_hi);
    stamp(C, 28);
  }
}

// ===============================================================================================================
// chain2: the same layer chains on a CLUSTER OF TWO CTAs per 128-row tile. Every hidden layer's 512 output columns are
// split: CTA `rank` computes columns [256 rank, 256 rank + 256) -- one 128 x 256 accumulator, double-buffered in TMEM
// (layer parity), so the next layer's MMAs never wait for the epilogue to drain it. The epilogue writes finished
// 64-column blocks straight into ACT (no parking, no flush pass), publishes each block to the local tensor core
// (mbarrier) and copies it into the peer's ACT (cp.async.bulk shared::cta -> shared::cluster, completion counted on
// the PEER's block barrier), and the next layer's MMAs start on whichever K blocks have landed. Per-layer latency:
// max(MMA of 128 x 256 x 512, epilogue of 128 columns per thread) instead of their sum over 512 columns; two SMs'
// fp32 pipes work on every row tile. Layers narrower than 257 columns (Euler head, input gradient) are computed by
// both CTAs redundantly (each patches its own copy of the layer-0 operand; rank 0 stores).
//
// Cross-CTA ordering: peer_done = "the peer's MMAs of this layer have retired" (remote arrive once per layer): nobody
// writes into the other CTA's ACT -- nor rewrites a block of its own ACT that an outgoing copy may still be reading
// -- before that. LayerNorm row statistics cross the cluster through st.shared::cluster + a remote mbarrier.
// ===============================================================================================================
struct Cl2 {
  uint64_t *tmem_full, *peer_done, *layer_ready, *ln_x;
  uint32_t peer_peer_done, peer_ln_x, peer_ex2;  // shared::cluster addresses in the peer CTA
  const float2* ex2;                             // [128] row sums written by the peer
  bool ln;
};

__device__ __forceinline__ void publish_block(const EpiCtx& C, int gblk) {
  if (!C.publish) return;
  // this thread has written its row of the block (generic proxy) and is done with the accumulator columns behind
  // it; the block barrier completes when all 128 threads of the half-group have arrived. The local tensor core and
  // the copy warp (which forwards the block to the peer) both wait on it.
  tc_fence_before();
  fence_proxy_async_smem();
  mbar_arrive(&C.blk_full[gblk]);
}

// wait for this layer's accumulator, tell the peer our ACT is free, wait until the peer's is
__device__ __forceinline__ void c2_begin_layer(EpiCtx& C, const Cl2& X, int lc) {
  const int p = lc & 1;
  mbar_wait(&X.tmem_full[p], (lc >> 1) & 1);
  tc_fence_after();
  // LayerNorm chains reuse the head of ACT for the row-sum exchange: every warp must be done copying the previous
  // layer out of ACT before the peer may write there
  if (X.ln) epi_bar_sync();
  if (threadIdx.x == 64) mbar_arrive_cluster(X.peer_peer_done);
  mbar_wait_cluster(X.peer_done, lc & 1);
  C.tm_off = p * 256;
}

// row sums over all 512 columns: the other half-group through the head of ACT, the other CTA through its ex2 area
__device__ __forceinline__ void exchange_row_sums2(const EpiCtx& C, const Cl2& X, int lnc, float& t1, float& t2) {
  float2* ex = reinterpret_cast<float2*>(C.act);
  const float2* ex2 = X.ex2;
  ex[C.half * 128 + C.r] = make_float2(t1, t2);
  epi_bar_sync();
  const float2 o = ex[(C.half ^ 1) * 128 + C.r];
  t1 += o.x;
  t2 += o.y;
  if (C.half == 0) {
    st_cluster_f32x2(X.peer_ex2 + C.r * 8, t1, t2);
    mbar_arrive_cluster_release(X.peer_ln_x);
  }
  mbar_wait_cluster(X.ln_x, lnc & 1);
  const float2 w = ex2[C.r];
  t1 += w.x;
  t2 += w.y;
  epi_bar_sync();  // the exchange area has been read: ACT may be rewritten
}

template <bool LN, bool SAVE>
__device__ __forceinline__ void fwd2_hidden_layer(EpiCtx& C, const Cl2& X, const ChainLayer& Ly, int lc, int lnc) {
  C.puse = C.pbuf != nullptr && !Ly.a_from_x0;
  if (C.puse) load_layer_params(C, Ly, true, LN);
  else prefetch_layer_params<true>(C, Ly, true, LN);
  c2_begin_layer(C, X, lc);
  stamp(C, 40);
  float s1 = 0.f, s2 = 0.f;
  if (!LN) {
    fwd_pass1<false, SAVE, true>(C, Ly, 0, true, s1, s2);
  } else {
    fwd_pass1<true, SAVE, true>(C, Ly, 0, false, s1, s2);
    tmem_st_wait();
    float t1 = s1, t2 = s2;
    exchange_row_sums2(C, X, lnc, t1, t2);
    const float inv_n = 1.0f / (float)C.H;
    const float mean = t1 * inv_n;
    const float var = fmaxf(t2 * inv_n - mean * mean, 0.f);
    const float rstd = rsqrtf(var + C.ln_eps);
    if (C.row_ok && C.half == 0 && C.g_off == 0 && Ly.out_rstd) Ly.out_rstd[C.rs_off] = rstd;
    fwd_flush<true, SAVE, true>(C, Ly, 0, mean, rstd);
  }
  stamp(C, 41);
  if (Ly.out_h) {
    __syncwarp();
    copy_out_act<true>(C, 0, Ly.n_pad, Ly.out_h, C.wrow, nullptr);
  }
  stamp(C, 42);
}

template <bool LN>
__device__ __forceinline__ void bwd2_hidden_layer(EpiCtx& C, const Cl2& X, const ChainLayer& Ly, int lc, int lnc) {
  C.puse = LN && C.pbuf != nullptr && !Ly.a_from_x0;
  if (C.puse) load_layer_params(C, Ly, false, LN);
  else prefetch_layer_params<true>(C, Ly, false, LN);
  if (!LN) bwd_prefetch_gp<true>(C, Ly, 0, 0);
  else ln_prefetch<true>(C, Ly.xhat, Ly.n_pad, 0, 0, 0);
  c2_begin_layer(C, X, lc);
  stamp(C, 50);
  if (!LN) {
    bwd_pass1<true>(C, Ly, 0, true, -1);
  } else {
    float s1 = 0.f, s2 = 0.f;
    bwd_ln_stats<true>(C, Ly, 0, -1, s1, s2);
    float t1 = s1, t2 = s2;
    exchange_row_sums2(C, X, lnc, t1, t2);
    const float inv_n = 1.0f / (float)C.H;
    const float rstd = C.row_ok ? __ldg(Ly.rstd + C.rs_off) : 0.f;
    bwd_ln_out<true>(C, Ly, 0, t1 * inv_n, t2 * inv_n, rstd);
  }
  stamp(C, 51);
  __syncwarp();
  copy_out_act<true>(C, 0, Ly.n_pad, Ly.out_dz, C.wdz, Ly.dbias_part);
  stamp(C, 52);
}

// K-block order of a 512-deep layer in CTA `r`: first the blocks each half-group (here and in the peer) finishes
// first, own blocks ahead of the peer's
__device__ __forceinline__ int c2_kb_order(int i, int r) {
  const int j = i & 3;
  const int cta = (j < 2) ? r : (r ^ 1);
  return cta * 4 + (j & 1) * 2 + (i >> 2);
}

template <int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1) fql_chain2_kernel(const __grid_constant__ ChainLaunch L) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* x0s = smem + kActBytes;
  uint8_t* stage_base = x0s + kActBlockBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage_base + kStages * kStageBytes);
  uint64_t* full_bar = bars;                // [kStages]
  uint64_t* empty_bar = bars + kStages;     // [kStages]
  uint64_t* tmem_full = bars + 2 * kStages; // [2] one per accumulator buffer
  uint64_t* blk_full = tmem_full + 2;       // [8] ACT k-blocks
  uint64_t* peer_done = blk_full + 8;
  uint64_t* layer_ready = peer_done + 1;    // after a narrow (unsplit) layer: 256 arrivals
  uint64_t* a0_full = layer_ready + 1;
  uint64_t* ln_x = a0_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ln_x + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();
  long long tl_t0 = 0;
  if (L.timeline && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0));

  const int pair = (int)blockIdx.x >> 1;
  int ji = 0;
#pragma unroll 1
  for (int i = 1; i < L.num_jobs; ++i)
    if (pair >= L.job[i].tile_begin) ji = i;
  const ChainJob& J = L.job[ji];
  const int rel = pair - J.tile_begin;
  const int bidx = rel / J.m_tiles;
  const int mt = rel - bidx * J.m_tiles;
  const int v = L.member_list ? __ldg(L.member_list + bidx) : bidx;
  const int n_layers = J.n_layers;
  const int iters = J.iters;
  const int n_stages = L.stages;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&J.mapA0);
    for (int i = 0; i < n_layers; ++i) tma_prefetch_desc(&J.mapW[i]);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_full[1], 1);
    // own blocks: one arrival per producing thread; the peer's blocks: its copy warp's expect_tx + the bytes
    for (int i = 0; i < 8; ++i) mbar_init(&blk_full[i], (i >> 2) == rank ? 128 : 1);
    mbar_init(peer_done, 1);
    mbar_init(layer_ready, 256);
    mbar_init(a0_full, 1);
    mbar_init(ln_x, 128);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // the peer's barriers are initialised before anything is sent to them
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
          const bool split = Ly.kind == CL_HIDDEN;
          const int n0 = split ? rank * 256 : 0;
          const int ncols = split ? 256 : Ly.n_pad;
          const int nkb = Ly.k_blocks;
          for (int i = 0; i < nkb; ++i) {
            const int kb = (nkb == 8 && !Ly.a_from_x0) ? c2_kb_order(i, rank) : i;
            for (int h = 0; h < 2; ++h) {
              const int ks = kb * 2 + h;
              mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sb = stage_base + stage * kStageBytes;
              if (MODE == 0) {
                mbar_arrive_expect_tx(&full_bar[stage], ncols * kStageK * 2);
                if (Ly.b_box_rows > 0) {
                  // wide-box weight map (pop.cu encode_forward_maps): the whole stage is ONE TMA instruction -- four
                  // separate 4 KB box requests deliver their first bytes thousands of cycles later (measured in
                  // rollout16.cu / chain16.cu: weight waits per layer 9 k -> 2 k cycles)
                  tma_load_4d(sb, &J.mapW[li], &full_bar[stage], 0, ks * kStageK, n0 >> 6, v);
                } else {
                  for (int b = 0; b < ncols / 64; ++b)
                    tma_load_3d(sb + b * kFwdBoxBytes, &J.mapW[li], &full_bar[stage], n0 + 64 * b, ks * kStageK, v);
                }
              } else {
                mbar_arrive_expect_tx(&full_bar[stage], (split ? 256 : Ly.b_box_rows) * kStageK * 2);
                tma_load_3d(sb, &J.mapW[li], &full_bar[stage], ks * kStageK, n0, v);
              }
              if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ========================================= MMA issuer ===========================================
    int stage = 0;
    uint32_t phase = 0;
    int lc = 0, sc = 0, hc = 0;  // layers, split layers, narrow layers issued so far
    int tr = 0;
    const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
    const uint64_t bdesc0 = MODE == 0 ? make_smem_desc(smem_u32(stage_base), kFwdBoxBytes, 1024, 2)
                                      : make_smem_desc(smem_u32(stage_base), 16, 512, 4);
    constexpr uint32_t kBStep = (MODE == 0 ? 2048u : 32u) >> 4;
    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li, ++lc) {
        const ChainLayer& Ly = J.layer[li];
        const bool split = Ly.kind == CL_HIDDEN;
        const int ncols = split ? 256 : Ly.n_pad;
        const int nkb = Ly.k_blocks;
        const bool from_act = !Ly.a_from_x0;
        if (!from_act) {
          if (lc == 0) mbar_wait(a0_full, 0);
          else mbar_wait(layer_ready, (hc - 1) & 1);
          tc_fence_after();
        }
        const uint64_t adesc0 = make_smem_desc(from_act ? act_addr : x0_addr, 16, 1024, 2);
        const uint32_t idesc = make_idesc_bf16(128, (uint32_t)ncols, 0, MODE == 0 ? 1 : 0);
        const uint32_t tmem_d = tmem_base + (lc & 1) * 256;
        for (int i = 0; i < nkb; ++i) {
          const int kb = (nkb == 8 && from_act) ? c2_kb_order(i, rank) : i;
          if (from_act) {
            mbar_wait(&blk_full[kb], (sc - 1) & 1);
            tc_fence_after();
          }
          if (L.trace && blockIdx.x == 0 && lane == 0 && tr < 4000) L.trace[tr++] = clock64();
          for (int h = 0; h < 2; ++h) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
              const uint64_t ad = adesc0 + (uint64_t)(kb * (kActBlockBytes >> 4) + h * 4);
              const uint64_t bd = bdesc0 + (uint64_t)(stage * (kStageBytes >> 4));
              umma_bf16(tmem_d, ad, bd, idesc, (i != 0 || h != 0) ? 1u : 0u);
              umma_bf16(tmem_d, ad + 2, bd + kBStep, idesc, 1u);
              umma_commit(&empty_bar[stage]);
              if (i == nkb - 1 && h == 1) umma_commit(&tmem_full[lc & 1]);
            }
            __syncwarp();
            if (++stage == n_stages) { stage = 0; phase ^= 1; }
          }
        }
        if (split) ++sc; else ++hc;
      }
    }
  } else if (warp == 10) {
    // ========================================= block forwarder ======================================
    // finished blocks of this CTA's column range -> the peer's ACT (asynchronous bulk copy; completion is counted on
    // the peer's block barrier, armed from here)
    if (lane == 0) {
      const uint32_t peer = (uint32_t)(rank ^ 1);
      const uint32_t peer_act = mapa_u32(smem_u32(act), peer);
      const uint32_t peer_blk = mapa_u32(smem_u32(blk_full), peer);
      int lc = 0, sc = 0;
      for (int it = 0; it < iters; ++it) {
        for (int li = 0; li < n_layers; ++li, ++lc) {
          if (J.layer[li].kind != CL_HIDDEN) continue;
          if (it == iters - 1 && li == n_layers - 1) continue;  // nobody consumes the last layer on chip
          mbar_wait_cluster(peer_done, lc & 1);  // the peer's tensor core has stopped reading its ACT
          for (int j = 0; j < 4; ++j) {
            const int gblk = rank * 4 + (j & 1) * 2 + (j >> 1);  // order of completion: block 0 of either half-group first
            mbar_wait(&blk_full[gblk], sc & 1);
            const uint32_t pb = peer_blk + gblk * 8;
            mbar_arrive_expect_tx_cluster(pb, kActBlockBytes);
            dsmem_bulk_copy(peer_act + gblk * kActBlockBytes, act + gblk * kActBlockBytes, kActBlockBytes, pb);
          }
          ++sc;
        }
      }
    }
  } else {
    // ========================================= epilogue =============================================
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const int grow = mt * 128 + r;
    const bool row_ok = grow < J.M;
    EpiCtx C;
    C.tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    C.tmem_full = tmem_full; C.act_lo = nullptr; C.act_hi = nullptr;
    C.act = act;
    C.params = J.params + (long long)v * J.S;
    C.r = r; C.half = half; C.q = q; C.lane = lane; C.H = J.H;
    C.row_ok = row_ok;
    C.prow = (long long)v * J.pass_ms + (long long)grow * J.H;
    C.rs_off = (long long)v * J.pass_rs + grow;
    C.dzrow = (long long)v * J.dz_ms + (long long)grow * J.H;
    C.part_off = (long long)v * J.part_ms + (long long)(mt * 4 + q) * J.H;
    C.ln_eps = J.ln_eps;
    C.stg = n_stages == kStages ? nullptr : stage_base + kStagesStaged * kStageBytes + (warp - 2) * kStagingBytes;
    const int grow0 = mt * 128 + q * 32;
    C.wrow = (long long)v * J.pass_ms + (long long)grow0 * J.H;
    C.wdz = (long long)v * J.dz_ms + (long long)grow0 * J.H;
    C.rows_valid = min(32, max(0, J.M - grow0));
    C.pbuf = iters == 1 ? reinterpret_cast<float*>(x0s) : nullptr;  // the Euler chain re-stages X0 every iteration
    C.puse = false;
    C.tm_off = 0; C.g_off = rank * 256; C.cl = true; C.publish = true;
    C.blk_full = blk_full;
    const uint32_t peer = (uint32_t)(rank ^ 1);
    C.peer_act = mapa_u32(smem_u32(act), peer);
    C.peer_blk_full = mapa_u32(smem_u32(blk_full), peer);
    Cl2 X;
    X.tmem_full = tmem_full; X.peer_done = peer_done; X.layer_ready = layer_ready; X.ln_x = ln_x;
    X.peer_peer_done = mapa_u32(smem_u32(peer_done), peer);
    X.peer_ln_x = mapa_u32(smem_u32(ln_x), peer);
    X.ex2 = reinterpret_cast<const float2*>(reinterpret_cast<uint8_t*>(bars) + 256);
    X.peer_ex2 = mapa_u32(smem_u32(X.ex2), peer);
    X.ln = J.ln != 0;
    const bool traced = L.trace && blockIdx.x == 0 && threadIdx.x == 64;
    C.tr = traced ? L.trace + 4096 : nullptr;
    if (traced) L.trace[4096] = 0;

    float xs[8];
    if (MODE == 0 && J.xe) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        xs[j] = (row_ok && half == 0 && j < J.A) ? J.xe[((long long)v * J.B + grow) * J.A + j] : 0.f;
    }

    int lc = 0, lnc = 0;
    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li, ++lc) {
        const ChainLayer& Ly = J.layer[li];
        if (Ly.kind == CL_HIDDEN) {
          // the last layer's output is consumed by nobody on chip
          C.publish = !(it == iters - 1 && li == n_layers - 1);
          if (MODE == 0) {
            const bool save = Ly.out_gp != nullptr;
            if (J.ln) {
              if (save) fwd2_hidden_layer<true, true>(C, X, Ly, lc, lnc);
              else fwd2_hidden_layer<true, false>(C, X, Ly, lc, lnc);
              ++lnc;
            } else {
              if (save) fwd2_hidden_layer<false, true>(C, X, Ly, lc, lnc);
              else fwd2_hidden_layer<false, false>(C, X, Ly, lc, lnc);
            }
          } else {
            if (J.ln) { bwd2_hidden_layer<true>(C, X, Ly, lc, lnc); ++lnc; }
            else bwd2_hidden_layer<false>(C, X, Ly, lc, lnc);
          }
        } else if (Ly.kind == CL_EULER_HEAD) {
          c2_begin_layer(C, X, lc);
          if (half == 0) {
            uint32_t rr[16];
            tmem_ld16(C.tmem_row + C.tm_off, rr);
            tmem_ld_wait();
            const float invF = 1.f / (float)J.F;
            const bool last = it == iters - 1;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (j < J.A) {
                xs[j] = fmaf(__uint_as_float(rr[j]) + __ldg(C.params + Ly.off_b + j), invF, xs[j]);
                if (last) {
                  if (row_ok && rank == 0) J.target_a[((long long)v * J.B + grow) * J.A + j] = clip1(xs[j]);
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
          release(layer_ready);
        } else {  // CL_DINPUT (both CTAs hold the full result; rank 0 stores)
          c2_begin_layer(C, X, lc);
          if (half == 0 && rank == 0) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (c * 32 < Ly.n_pad) {
                uint32_t rr[32];
                tmem_ld32(C.tmem_row + C.tm_off + c * 32, rr);
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
          release(layer_ready);
        }
        stamp(C, 99);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();  // no CTA leaves while its peer may still signal it or copy into it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
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
  uint64_t* tmem_full = bars + 2 * kStages;   // [2] one per 256-column accumulator tile
  uint64_t* act_lo = bars + 2 * kStages + 2;  // ACT columns 0..255 (and X0) hold the next layer's A operand
  uint64_t* act_hi = bars + 2 * kStages + 3;  // ... columns 256..511 too, and all of TMEM has been drained
  uint64_t* a0_full = bars + 2 * kStages + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 5);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  long long tl_t0 = 0;
  if (L.timeline && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tl_t0));

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
  const int n_stages = L.stages;  // 5, or 3 when stages 3 and 4 serve as the epilogue's staging area

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&J.mapA0);
    for (int i = 0; i < n_layers; ++i) tma_prefetch_desc(&J.mapW[i]);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full[0], 1);
    mbar_init(&tmem_full[1], 1);
    mbar_init(act_lo, 256);
    mbar_init(act_hi, 256);
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
          const int k_steps = Ly.k_blocks * 2;
          for (int n0 = 0; n0 < Ly.n_pad; n0 += kTileN) {
            const int ncols = min(kTileN, Ly.n_pad - n0);
            for (int ks = 0; ks < k_steps; ++ks) {
              mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sb = stage_base + stage * kStageBytes;
#ifdef FQLPOP_DEBUG_SWITCHES
              if (L.debug_flags & 1) {  // (make TRACE=1 only) no weight stream: the MMAs run on whatever the stage holds
                mbar_arrive(&full_bar[stage]);
              } else
#endif
              if (MODE == 0) {
                // W[k][n], n contiguous: boxes of 64 n x 32 k
                mbar_arrive_expect_tx(&full_bar[stage], ncols * kStageK * 2);
                for (int b = 0; b < ncols / 64; ++b)
                  tma_load_3d(sb + b * kFwdBoxBytes, &J.mapW[li], &full_bar[stage], n0 + 64 * b, ks * kStageK, v);
              } else {
                // the same array read as B^T: rows = output columns of this GEMM, 32 reduction elements per 64-byte row
                mbar_arrive_expect_tx(&full_bar[stage], Ly.b_box_rows * kStageK * 2);
                tma_load_3d(sb, &J.mapW[li], &full_bar[stage], ks * kStageK, n0, v);
              }
              if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ========================================= MMA issuer ===========================================
    // The whole warp walks the loop (waits are warp-uniform); one elected lane issues the MMAs and commits.
    int stage = 0;
    uint32_t phase = 0;
    uint32_t itc = 0;
    int tr = 0;
    const uint32_t act_addr = smem_u32(act), x0_addr = smem_u32(x0s);
    const uint64_t bdesc0 = MODE == 0 ? make_smem_desc(smem_u32(stage_base), kFwdBoxBytes, 1024, 2)
                                      : make_smem_desc(smem_u32(stage_base), 16, 512, 4);
    constexpr uint32_t kBStep = (MODE == 0 ? 2048u : 32u) >> 4;  // descriptor units (16 B) per K = 16 step
    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li, ++itc) {
        const ChainLayer& Ly = J.layer[li];
        bool hi_pending = false;
        if (itc == 0) mbar_wait(a0_full, 0);
        else {
          mbar_wait(act_lo, (itc - 1) & 1);
          hi_pending = true;
        }
        tc_fence_after();
        if (L.trace && blockIdx.x == 0 && lane == 0 && tr < 4000) L.trace[tr++] = clock64();
        const uint64_t adesc0 = make_smem_desc(Ly.a_from_x0 ? x0_addr : act_addr, 16, 1024, 2);
        const int k_steps = Ly.k_blocks * 2;
        int tile = 0;
        for (int n0 = 0; n0 < Ly.n_pad; n0 += kTileN, ++tile) {
          const int ncols = min(kTileN, Ly.n_pad - n0);
          const uint32_t idesc = make_idesc_bf16(128, (uint32_t)ncols, 0, MODE == 0 ? 1 : 0);
          const uint32_t tmem_d = tmem_base + n0;
          for (int ks = 0; ks < k_steps; ++ks) {
            if (hi_pending && (tile > 0 || ks * kStageK >= 256)) {
              mbar_wait(act_hi, (itc - 1) & 1);
              tc_fence_after();
              hi_pending = false;
            }
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
              const uint64_t ad = adesc0 + (uint64_t)((ks >> 1) * (kActBlockBytes >> 4) + (ks & 1) * 4);
              const uint64_t bd = bdesc0 + (uint64_t)(stage * (kStageBytes >> 4));
              umma_bf16(tmem_d, ad, bd, idesc, ks != 0 ? 1u : 0u);
              umma_bf16(tmem_d, ad + 2, bd + kBStep, idesc, 1u);
              umma_commit(&empty_bar[stage]);
              if (ks == k_steps - 1) umma_commit(&tmem_full[tile]);
            }
            __syncwarp();
            if (++stage == n_stages) { stage = 0; phase ^= 1; }
          }
          if (L.trace && blockIdx.x == 0 && lane == 0 && tr < 4000) L.trace[tr++] = clock64();
        }
        if (hi_pending) {
          mbar_wait(act_hi, (itc - 1) & 1);
          tc_fence_after();
        }
      }
    }
  } else {
    // ========================================= epilogue =============================================
    const int q = warp & 3;              // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;    // which 128 columns of every 256-column accumulator tile
    const int r = q * 32 + lane;         // row of the tile == TMEM lane
    const int grow = mt * 128 + r;       // row within the job
    const bool row_ok = grow < J.M;
    EpiCtx C;
    C.tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    C.tmem_full = tmem_full; C.act_lo = act_lo; C.act_hi = act_hi;
    C.act = act;
    C.params = J.params + (long long)v * J.S;
    C.r = r; C.half = half; C.q = q; C.lane = lane; C.H = J.H;
    C.row_ok = row_ok;
    C.prow = (long long)v * J.pass_ms + (long long)grow * J.H;
    C.rs_off = (long long)v * J.pass_rs + grow;
    C.dzrow = (long long)v * J.dz_ms + (long long)grow * J.H;
    C.part_off = (long long)v * J.part_ms + (long long)(mt * 4 + q) * J.H;
    C.ln_eps = J.ln_eps;
    C.stg = n_stages == kStages ? nullptr : stage_base + kStagesStaged * kStageBytes + (warp - 2) * kStagingBytes;
    const int grow0 = mt * 128 + q * 32;
    C.wrow = (long long)v * J.pass_ms + (long long)grow0 * J.H;
    C.wdz = (long long)v * J.dz_ms + (long long)grow0 * J.H;
    C.rows_valid = min(32, max(0, J.M - grow0));
    C.pbuf = iters == 1 ? reinterpret_cast<float*>(x0s) : nullptr;  // the Euler chain re-stages X0 every iteration
    C.puse = false;
    C.tm_off = 0; C.g_off = 0; C.cl = false; C.publish = false; C.blk_full = nullptr; C.peer_act = 0; C.peer_blk_full = 0;
    const bool traced = L.trace && blockIdx.x == 0 && threadIdx.x == 64;
    C.tr = traced ? L.trace + 4096 : nullptr;
    if (traced) L.trace[4096] = 0;
    uint32_t fullph = 0;

    float xs[8];  // Euler flow: the action iterate of this row (fp32)
    if (MODE == 0 && J.xe) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        xs[j] = (row_ok && half == 0 && j < J.A) ? J.xe[((long long)v * J.B + grow) * J.A + j] : 0.f;
    }

    for (int it = 0; it < iters; ++it) {
      for (int li = 0; li < n_layers; ++li) {
        const ChainLayer& Ly = J.layer[li];
        if (Ly.kind == CL_HIDDEN) {
          if (MODE == 0) {
            const bool save = Ly.out_gp != nullptr;
            if (J.ln) {
              if (save) fwd_hidden_layer<true, true>(C, fullph, Ly);
              else fwd_hidden_layer<true, false>(C, fullph, Ly);
            } else {
              if (save) fwd_hidden_layer<false, true>(C, fullph, Ly);
              else fwd_hidden_layer<false, false>(C, fullph, Ly);
            }
          } else {
            if (J.ln) bwd_hidden_layer<true>(C, fullph, Ly);
            else bwd_hidden_layer<false>(C, fullph, Ly);
          }
        } else if (Ly.kind == CL_EULER_HEAD) {
          // BC-flow head inside the Euler rollout: x <- x + v / F; the next step's input differs from this step's
          // only in the action columns and in t, which are patched in place in the X0 block.
          wait_tile(C, fullph, 0);
          if (half == 0) {
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
          }
          release(act_lo);
          release(act_hi);
        } else {  // CL_DINPUT: gradient with respect to the (padded, <= 64 wide) network input, fp32
          wait_tile(C, fullph, 0);
          if (half == 0) {
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
          release(act_lo);
          release(act_hi);
        }
        stamp(C, 99);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
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
  e = cudaFuncSetAttribute(fql_chain_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fql_chain2_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(fql_chain2_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
  if (e != cudaSuccess) return e;
  return chain16_init();
}

constexpr int kClusterMaxMembers = 8;

// The cluster kernel covers the production shape: every hidden layer 512 wide, reductions of 64 (first layer) or 512.
static bool chain2_eligible(const ChainLaunch& L);
// (pop.cu) true when launch_chain will run this forward launch on the N-split cluster kernel: its weight maps may then be
// wide-box maps (4 column blocks per hidden-layer stage, 1 for the narrow layers)
bool chain2_will_run(const ChainLaunch& L) {
  if (chain16_eligible(L) && chain16_selected_shape(L.job[0].H, 64, L.job[0].ln != 0, L.job[0].iters, L.population_active))
    return false;
  return chain2_eligible(L);
}
static bool chain2_eligible(const ChainLaunch& L) {
  // FQLPOP_CHAIN2 = 0: single-CTA kernel; 1: cluster kernel for every chain; 2: cluster kernel for the Euler chain only;
  // unset: by population size. The cluster kernel cuts a chain's latency 1.4-1.7x but costs ~20 % more SM time: with few
  // active members the step is latency-bound and it wins (4 members 6.5k -> 9.8k updates/s, 8 members 11.8k -> 15.3k), from
  // ~12 members on the GPU is saturated and SM time decides (16 members 17.2k vs 16.2k). Successive halving walks through
  // exactly these sizes. Read at every launch (graph capture time; pruning the population re-captures the graphs).
  const char* e = getenv("FQLPOP_CHAIN2");
  const int mode = e ? atoi(e) : ((L.population_active > 0 && L.population_active <= kClusterMaxMembers) ? 1 : 0);
  if (!mode) return false;
  if (mode == 2 && L.job[0].iters <= 1) return false;  // 2: only the Euler chain (the step's longest serial path)
  for (int i = 0; i < L.num_jobs; ++i) {
    const ChainJob& j = L.job[i];
    if (j.H != 512) return false;
    for (int l = 0; l < j.n_layers; ++l) {
      const ChainLayer& y = j.layer[l];
      if (y.kind == CL_HIDDEN && y.n_pad != 512) return false;
      if (y.kind != CL_HIDDEN && y.n_pad > 256) return false;
      if (y.a_from_x0 ? y.k_blocks != 1 : y.k_blocks != 8) return false;
    }
  }
  return true;
}

// Which kernel generation runs a chain of this shape: the second-generation kernel (chain16.cu) unless
// FQLPOP_CHAIN_GEN=1 (first-generation kernels: A/B measurements, parity cross-checks), FQLPOP_CHAIN2=1/2 forces the
// N-split cluster kernel, or few members are active (below). The builders in pop.cu ask the same question: the weight boxes of a backward chain are sized
// for the kernel that will read them.
// Population size decides between the two cluster designs (measured on B200, profiles/r02_small_populations.txt): up to
// kClusterMaxMembers active members the step is latency-bound and the N-split cluster kernel of the first generation
// (two SMs work on every layer of a row tile: shortest chain latency) wins -- 1 member 2.9k vs 2.4k updates/s, 4 members
// 10.5k vs 8.8k, 8 members 15.5k vs 14.9k; above, SM time decides and the second-generation pair kernel wins.
bool chain16_selected_shape(int H, int in_pad, bool ln, int iters, int population_active) {
  static const int gen = [] { const char* e = getenv("FQLPOP_CHAIN_GEN"); return e ? atoi(e) : 2; }();
  if (gen < 2 || H != 512 || in_pad != 64 || (ln && iters > 1)) return false;
  const char* e = getenv("FQLPOP_CHAIN2");
  const int m = e ? atoi(e) : ((population_active > 0 && population_active <= kClusterMaxMembers) ? 1 : 0);
  return !(m == 1 || (m == 2 && iters > 1));
}

cudaError_t launch_chain(ChainLaunch& L, int mode, cudaStream_t stream) {
  int grid = 0;
  for (int i = 0; i < L.num_jobs; ++i) grid += L.job[i].m_tiles * L.n_virtual;
  if (grid <= 0) return cudaSuccess;
  // the staging area (tensors saved for / read by the backward pass) takes the place of weight stages 3 and 4
  bool staged = mode == 1;
  for (int i = 0; i < L.num_jobs; ++i)
    for (int l = 0; l < L.job[i].n_layers; ++l)
      if (L.job[i].layer[l].out_gp || L.job[i].layer[l].out_xhat) staged = true;
  L.stages = staged ? kStagesStaged : kStages;
  if (chain16_eligible(L) && chain16_selected_shape(L.job[0].H, 64, L.job[0].ln != 0, L.job[0].iters, L.population_active))
    return launch_chain16(L, mode, grid, stream);
  const bool c2 = chain2_eligible(L);
  if (mode == 0)  // wide-box weight maps are read by chain16.cu and the cluster kernel only, each with its own box shape
    for (int i = 0; i < L.num_jobs; ++i)
      for (int l = 0; l < L.job[i].n_layers; ++l) {
        const int wb = L.job[i].layer[l].b_box_rows;
        if (wb != 0 && !(c2 && wb == (L.job[i].layer[l].kind == CL_HIDDEN ? 4 : L.job[i].layer[l].n_pad / 64)))
          return cudaErrorInvalidValue;
      }
  if (c2) {
    if (mode == 0) fql_chain2_kernel<0><<<2 * grid, kThreads2, kSmemBytes, stream>>>(L);
    else fql_chain2_kernel<1><<<2 * grid, kThreads2, kSmemBytes, stream>>>(L);
    return cudaGetLastError();
  }
  if (mode == 0) fql_chain_kernel<0><<<grid, kThreads, kSmemBytes, stream>>>(L);
  else fql_chain_kernel<1><<<grid, kThreads, kSmemBytes, stream>>>(L);
  return cudaGetLastError();
}

}  // namespace fqlpop
