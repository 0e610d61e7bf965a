// Grouped / population-batched GEMM on tcgen05 (sm_100a): launch descriptors shared by host and device.
//
// One launch = up to kMaxGroups "groups". A group is one logical GEMM  D[M,N] = A[M,K] * B[N,K]^T  that is
// repeated for every active population member (the batch coordinate of the 3-D TMA maps). Operands are bf16,
// accumulation is fp32 in TMEM, and each group picks a fused epilogue (bias+GELU(+LayerNorm), the
// matching backward epilogues, or a plain store).
//
// Reference path this replaces: every `nn.Dense` inside FQLAgent.update (called at
// /root/reference/trainer/experiment.py:109) -- one XLA executable per agent per step in the reference,
// one tile of this kernel per (member, network, 128-row block) here.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace fqlpop {

constexpr int kMaxGroups = 12;
constexpr int kBlockM = 128;
constexpr int kBlockN = 128;  // columns per accumulator sub-tile
constexpr int kBlockK = 64;   // 64 bf16 = 128 B = one swizzle row

enum EpiKind : int {
  EPI_STORE_F32 = 0,   // out_f32[row, col] = acc
  EPI_STORE_BF16 = 1,  // out_h[row, col] = bf16(acc)
  EPI_FWD = 2,         // z = acc + bias; h = act(z); store h (bf16) and optionally act'(z) (bf16)
  EPI_FWD_LN = 3,      // as EPI_FWD followed by LayerNorm over the full 512-wide row (needs N_SUB = 4)
  EPI_DGRAD = 4,       // dz = acc * gp; store dz (bf16); column sums of dz -> bias-gradient partials
  EPI_DGRAD_LN = 5,    // LayerNorm backward on acc, then as EPI_DGRAD; also gamma/beta gradient partials
  EPI_ADAM = 6,        // acc = dL/dW: Adam + Polyak + bf16 shadow refresh of that weight tile, straight out of the accumulator
                       // (the gradient itself is stored only when adam.keep_grad is set)
};

// Optimizer state of the weight tile a weight-gradient GEMM produces (EPI_ADAM). All arrays share the gradient's layout
// (out_f32: per-virtual-member stride ms, leading dimension ld) and are already offset to the leaf.
struct AdamTile {
  float *theta, *mu, *nu, *target;        // target == nullptr: no Polyak copy (actor networks)
  __nv_bfloat16 *shadow, *tshadow;
  const float* adam_bc;                   // [population][2]: 1 - b1^t, 1 - b2^t of the current step
  int nb;                                 // virtual members per population member (adam_bc is indexed by member / nb)
  int keep_grad;                          // also write the gradient to out_f32 (steps whose gradients are read back)
  float lr, b1, b2, eps, tau;
  int _pad;
};
enum ActKind : int { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };

// pointer + per-member stride (elements) + leading dimension (elements)
template <class T>
struct MPtr {
  T* p;
  long long ms;
  int ld;
  int _pad;
};

struct EpiParams {
  int kind;
  int act;
  float ln_eps;
  int _pad;
  MPtr<const float> bias;    // [P][N]
  MPtr<const float> gamma;   // [P][N]   (LayerNorm scale of THIS layer for FWD_LN; of the layer whose
  MPtr<const float> beta;    //           output gradient is being produced for DGRAD_LN)
  MPtr<__nv_bfloat16> out_h;     // FWD*: activations; STORE_BF16: plain output; DGRAD*: dz
  MPtr<__nv_bfloat16> out_gp;    // FWD*: act'(z) saved for backward (optional)
  MPtr<__nv_bfloat16> out_xhat;  // FWD_LN: normalised activations saved for backward (optional)
  MPtr<float> out_rstd;          // FWD_LN: 1/sqrt(var+eps) per row (optional); ld unused
  MPtr<const __nv_bfloat16> gp;    // DGRAD*: saved act'(z) (nullptr -> 1)
  MPtr<const __nv_bfloat16> xhat;  // DGRAD_LN
  MPtr<const float> rstd;          // DGRAD_LN
  MPtr<float> out_f32;             // STORE_F32
  MPtr<float> out_dbias_part;      // DGRAD*: [P][m_tiles][N] column sums of dz (optional)
  MPtr<float> out_dgamma_part;     // DGRAD_LN: [P][m_tiles][N]
  MPtr<float> out_dbeta_part;      // DGRAD_LN: [P][m_tiles][N]
  AdamTile adam;                   // EPI_ADAM
};

struct GemmGroup {
  CUtensorMap mapA;  // K-major: dims (K, rows, P) box (64,128,1); MN-major: dims (MN, K, P) box (64,64,1)
  CUtensorMap mapB;
  int tile_begin;  // first linear tile of this group
  int m_tiles;     // ceil(M / 128)
  int n_tiles;     // ceil(N / (128 * N_SUB))
  int M;           // valid rows
  int N;           // valid columns (multiple of 64)
  int k_blocks;    // K / 64
  int a_mn;        // 1 = A is MN-major in global memory ([k][m], m contiguous)
  int b_mn;        // 1 = B is MN-major in global memory ([k][n], n contiguous)
  int a_row0;      // K-major: first row (m) ; MN-major: first k row
  int a_col0;      // K-major: first k column ; MN-major: first m column
  int b_row0;
  int b_col0;
  EpiParams epi;
};

struct GemmLaunch {
  int num_groups;
  int tiles_per_member_total;  // unused by device; bookkeeping
  int P;                       // population size (grid covers P members)
  int _pad;
  const int* n_active;         // device scalar: number of active members (nullptr -> P)
  const int* member_list;      // device [P]: active member ids (nullptr -> identity)
  GemmGroup g[kMaxGroups];
};

// Host API ---------------------------------------------------------------------------------------
// Encodes a 3-D bf16 tensor map. `inner` is the contiguous extent; `rows` the second extent with pitch
// `ld` (elements); `batch` the third with pitch `batch_stride` (elements). box = (64, box_rows, 1),
// 128-byte swizzle. Returns 0 on success.
int encode_tmap_bf16(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                     long long ld, long long batch_stride, int box_rows);

// Same tensor, 64-byte swizzle and a (32, box_rows, 1) box: K-major operand stages of 32 reduction elements per row.
int encode_tmap_bf16_wide(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                          long long ld, long long batch_stride, int box_rows, int n_boxes);
int encode_tmap_bf16_tiled(CUtensorMap* out, const void* base, long long K, long long N, long long batch,
                           long long batch_stride, int n_boxes, int ks);
int encode_tmap_bf16_sw64(CUtensorMap* out, const void* base, long long inner, long long rows, long long batch,
                          long long ld, long long batch_stride, int box_rows);

// Fills tile_begin of every group (tiles are ordered group-major, then member, m-tile, n-tile) and
// returns the grid size. n_sub is 1 or 4 (4 = full 512-wide rows per CTA, needed by the *_LN epilogues).
int finalize_launch(GemmLaunch* L, int n_sub);

// Launches the kernel variant for n_sub in {1,4}. All groups of a launch must use the same n_sub.
cudaError_t launch_gemm(const GemmLaunch& L, int n_sub, int grid, cudaStream_t stream);
// One-time: raises the dynamic shared memory limit of both variants.
cudaError_t gemm_init();

}  // namespace fqlpop
