// Layer-chain kernels of the population FQL update (sm_100a): one CTA carries a 128-row tile of one (virtual)
// member through EVERY Dense layer of a forward pass -- or through the whole input-gradient chain of a backward
// pass, or through all F Euler steps of the BC flow -- without the activations leaving the SM. Only the bf16
// weights stream (TMA, from L2), the products run on tcgen05 with the accumulator in TMEM, and the activations
// live in shared memory as the next layer's A operand.
//
// Reference path this replaces: the Dense -> GELU -> (LayerNorm) chains XLA emits for FQLAgent.update,
// compute_flow_actions and their autodiff transposes (call site /root/reference/trainer/experiment.py:109;
// math in SURVEY.md Appendix A). One launch here = one whole MLP evaluation (or backward sweep) of every active
// member, instead of one launch per layer (gemm.cu keeps the per-layer kernels for the weight gradients and
// for hidden widths this kernel does not cover).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>

namespace fqlpop {

constexpr int kChainMaxLayers = 8;  // hidden layers + head / input-gradient layer
constexpr int kChainMaxJobs = 3;

enum ChainLayerKind : int {
  CL_HIDDEN = 0,      // forward: h = act(acc + b) (+ LayerNorm); backward: dz = act'(z) * LNbwd(acc)
  CL_EULER_HEAD = 1,  // forward: v = acc + b; x <- x + v / F; restage (obs, x, t) as the next step's input
  CL_DINPUT = 2,      // backward: dL/d(network input) = acc, stored as fp32
};

struct ChainLayer {
  int kind;
  int k_blocks;    // reduction length / 64
  int n_pad;       // output columns (multiple of 64, <= 512)
  int a_from_x0;   // A operand = the 128 x 64 input block instead of the activation buffer
  int b_box_rows;  // backward: rows of the weight TMA box (min(256, output columns)); forward: 0 = weight map with 64 n x
                   // 32 k boxes, else a wide map whose box spans this many 64-column blocks (chain16.cu only)
  int h_tma;       // (chain16.cu, forward) out_h is written by TMA stores through ChainJob::mapH of this layer
  long long off_b, off_g, off_beta;  // element offsets of bias / LayerNorm scale / LayerNorm bias in `params`
  // forward outputs (nullptr = not stored). Row-major [V][pass rows][H] bf16, already offset to the job's row 0.
  __nv_bfloat16* out_h;
  __nv_bfloat16* out_gp;
  __nv_bfloat16* out_xhat;
  float* out_rstd;  // [V][pass rows]
  // backward inputs saved by the forward pass, and outputs
  const __nv_bfloat16* gp;
  const __nv_bfloat16* xhat;
  const float* rstd;
  __nv_bfloat16* out_dz;  // [V][M][H]
  float* dbias_part;      // [V][m_tiles * 4][H] column sums per TMEM lane quarter (nullptr = skip)
  float* dgamma_part;
  float* dbeta_part;
  float* out_f32;         // CL_DINPUT: [V][M][out_ld]
  int out_ld;
  int _pad;
};

struct ChainJob {
  CUtensorMap mapA0;                   // first A operand: dims (64, pass rows, V), box 64 x 128
  CUtensorMap mapW[kChainMaxLayers];   // forward: [K][N] n-contiguous, box 64n x 32k; backward: same array read
                                       // as B^T, 64-byte swizzle, box 32 k x min(256, n) rows
  CUtensorMap mapH[kChainMaxLayers];   // (chain16.cu, forward) the layer's h buffer: dims (H, row0 + M, V), box 64 x 32
  ChainLayer layer[kChainMaxLayers];
  int tile_begin, m_tiles, M, row0;    // rows [row0, row0 + M) of the pass buffers
  int n_layers, iters, ln, H;
  float ln_eps;
  int O, A, F, B;                      // Euler flow only
  int _pad;
  const float* params;                 // fp32 parameter arena (biases, LayerNorm), stride S per virtual member
  long long S;
  long long pass_ms, pass_rs;          // per-virtual-member stride of the pass buffers ([rows][H], [rows])
  long long dz_ms, part_ms;            // ... of the backward dz buffers and of the partial-sum buffers
  long long f32_ms;
  float* xe;                           // Euler: [P][B][A] fp32 start point (noise) ...
  float* target_a;                     // ... and clip(x_F) out
};

struct ChainLaunch {
  int num_jobs;
  int n_virtual;            // active virtual members (grid = sum_jobs m_tiles * n_virtual)
  const int* member_list;   // (dev) [n_virtual] virtual member ids
  int debug_flags;          // (debug) bit 0: no weight TMA / no stage waits (timing of the bare MMA stream; results are garbage)
  int population_active;    // active members of the whole population on this GPU (picks the kernel: see chain2_eligible)
  int stages;               // weight stages (set by launch_chain: 5, or 3 when the epilogue stages tensors through smem)
  int staged;               // (chain16) the epilogue stages tensors through shared memory
  int _pad1;
  long long* trace;         // (dev, debug) clock64 stamps of CTA 0: [0,4096) MMA thread, [4096,8192) epilogue thread 0
  long long* timeline;      // (dev, debug) [0] = record counter; records of 4 words from [8]: start ns, end ns, smid, tag
  long long tag;            // (debug) launch index within the step | first member of the group << 8 | blockIdx << 32
  ChainJob job[kChainMaxJobs];
};

// mode 0: forward / Euler, 1: backward. Returns the launch status.
cudaError_t launch_chain(ChainLaunch& L, int mode, cudaStream_t stream);
cudaError_t chain_init();
// Fills tile_begin of every job; returns the grid size.
int finalize_chain(ChainLaunch* L);

// Second-generation kernel (chain16.cu): 16 epilogue warps, per-k-block hand-off. launch_chain picks it whenever the
// launch has the production shape (chain16_eligible).
cudaError_t chain16_init();
bool chain16_eligible(const ChainLaunch& L);
bool chain2_will_run(const ChainLaunch& L);  // the N-split cluster kernel will take this forward launch
bool chain16_selected_shape(int H, int in_pad, bool ln, int iters, int population_active);
bool chain16_pair_mode();
bool chain16_pairs_rows(int M);  // ... and only for jobs with an even number of 128-row tiles  // clusters of two CTAs, tcgen05.mma.cta_group::2 (default; FQLPOP_CHAIN_PAIR=0 disables)
cudaError_t launch_chain16(ChainLaunch& L, int mode, int grid, cudaStream_t stream);

}  // namespace fqlpop
