// Device-resident population of Flow Q-Learning agents (one per BC coefficient alpha) stepped in lockstep.
//
// Replaces, for the whole population at once, what the reference does one agent at a time:
//   task.sample("train", B)      /root/reference/trainer/experiment.py:108, task/offline_task_simulated.py:53-58
//   agent.update(batch)          /root/reference/trainer/experiment.py:109   (FQLAgent.update)
//   agent.total_loss(val_batch)  /root/reference/trainer/experiment.py:115
//   agent.sample_actions(...)    /root/reference/evaluator/evaluation.py:58-63
//
// Memory layout (all in HBM, owned by the Population):
//   * three parameter arenas (critic with E "virtual members" per member, BC flow field, one-step policy);
//     per virtual member one contiguous block of S floats holding every leaf, flax layout kernel[in][out]
//     with in padded to a multiple of 64 and the head's out padded to 64 (padding stays exactly zero);
//     the same block shape is used for theta / Adam mu / Adam nu / grad (fp32) and the bf16 shadow the GEMMs
//     read; the critic additionally has target (fp32) + target shadow (bf16).
//   * per forward pass activation buffers in bf16: X0 [V][M][in_pad], h_l / act'(z_l) / xhat_l [V][M][H],
//     rstd_l [V][M] (fp32).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "gemm.h"

namespace fqlpop {

constexpr int kMaxLayers = 8;   // hidden layers + head
constexpr int kMaxHeadOut = 8;  // action_dim limit of the CUDA-core head kernels
constexpr int kHeadPad = 64;    // head out dimension padded to one swizzle row
constexpr int kInfoSlots = 16;

// info slots (per member); the first 13 follow oracle.INFO_KEYS / upstream FQL info dict
enum InfoSlot : int {
  I_CRITIC_LOSS = 0, I_Q_MEAN, I_Q_MAX, I_Q_MIN, I_ACTOR_LOSS, I_BC_LOSS, I_DISTILL_LOSS, I_Q_LOSS, I_Q, I_MSE,
  I_GRAD_MAX, I_GRAD_MIN, I_GRAD_NORM,
  I_QABS,  // scratch: mean |q(s, a_pi)| for normalize_q_loss
};

struct NetLayout {
  int nb = 1;  // virtual members per population member (ensemble size for the critic)
  int in_dim = 0, in_pad = 0, H = 0, n_hidden = 0, out_dim = 0, ln = 0;
  long long offW[kMaxLayers] = {0}, offb[kMaxLayers] = {0}, offg[kMaxLayers] = {0}, offbeta[kMaxLayers] = {0};
  long long S = 0;  // floats per virtual member
  __host__ __device__ int n_layers() const { return n_hidden + 1; }
  __host__ __device__ int K(int l) const { return l == 0 ? in_pad : H; }                // rows of kernel l (padded)
  __host__ __device__ int N(int l) const { return l == n_hidden ? kHeadPad : H; }       // cols of kernel l (padded)
  __host__ __device__ int K_true(int l) const { return l == 0 ? in_dim : H; }
  __host__ __device__ int N_true(int l) const { return l == n_hidden ? out_dim : H; }
};

struct Arena {
  NetLayout L;
  int V = 0;  // virtual members = P * nb
  float *theta = nullptr, *mu = nullptr, *nu = nullptr, *grad = nullptr;
  __nv_bfloat16* shadow = nullptr;
  float* target = nullptr;            // critic only
  __nv_bfloat16* tshadow = nullptr;   // critic only
};

// Activation storage of one forward pass (optionally kept for the backward pass).
struct PassBuf {
  int V = 0, M = 0, in_pad = 0, H = 0, n_hidden = 0;
  bool save = false, ln = false;
  __nv_bfloat16* X0 = nullptr;                  // [V][M][in_pad]
  __nv_bfloat16* h[kMaxLayers] = {nullptr};     // [V][M][H]  (save: one per layer; else two, ping-pong)
  __nv_bfloat16* gp[kMaxLayers] = {nullptr};    // save only: act'(z_l)
  __nv_bfloat16* xhat[kMaxLayers] = {nullptr};  // save && ln
  float* rstd[kMaxLayers] = {nullptr};          // save && ln: [V][M]
  const __nv_bfloat16* hl(int l) const { return save ? h[l] : h[l & 1]; }
  __nv_bfloat16* hl(int l) { return save ? h[l] : h[l & 1]; }
};

// Backward scratch of one trainable pass.
struct BwdBuf {
  int V = 0, M = 0, H = 0, m_tiles = 0;
  int part_rows = 0;  // rows of the partial-sum buffers actually written (m_tiles, or 4 * m_tiles by the chain kernels)
  __nv_bfloat16* dout = nullptr;   // [V][M][64]   head output gradient (bf16, zero padded)
  __nv_bfloat16* dz[kMaxLayers] = {nullptr};   // [V][M][H] per hidden layer (wgrad reads them later)
  float* dbias_part[kMaxLayers] = {nullptr};   // [V][m_tiles][H]
  float* dgamma_part[kMaxLayers] = {nullptr};
  float* dbeta_part[kMaxLayers] = {nullptr};
};

struct PopParams {
  int O, A, H_actor, H_value, n_hidden_actor, n_hidden_value, E, F, B, P;
  float lr, discount, tau, adam_b1, adam_b2, adam_eps, ln_eps;
  int q_agg_min, normalize_q_loss, layer_norm, actor_layer_norm;
};

struct Dataset {
  const float *obs = nullptr, *act = nullptr, *next_obs = nullptr, *rew = nullptr, *mask = nullptr;
  long long n = 0;
};

// explicit noise for parity mode; all (dev) [P][B][...] or nullptr -> counter-based Philox on device
struct NoisePtrs {
  const float *z_critic = nullptr, *x0 = nullptr, *t = nullptr, *z_actor = nullptr, *z_mse = nullptr;
  const long long* indices = nullptr;  // (dev) [P][B] rows of the dataset, nullptr -> Philox
};

struct Plan;  // captured step (streams, events, CUDA graph)

// Device scratch of the fused rollout / sample_actions kernels (rollout.cu), allocated on first use.
struct RolloutScratch {
  __nv_bfloat16* head_w = nullptr;  // [P][H][64] one-step-policy head, action j in column O + j
  float* head_b = nullptr;          // [P][64]
  __nv_bfloat16* pi_tiled = nullptr;  // [P][S] hidden-layer kernels of the one-step policy, tiled [K/32][H/64][32][64]
  float *obs_work = nullptr, *success_sum = nullptr, *return_sum = nullptr;
  int* steps_run = nullptr;
  unsigned char* success = nullptr;
  int* length = nullptr;
  float* noise_next = nullptr;  // [P][2][ep_cap][A] policy noise of the next lockstep step (rollout16.cu draws it early)
  int* identity = nullptr;  // [P] = 0..P-1 (a one-entry member list for single-member launches)
  long long ep_cap = 0;
  long long launches = 0;   // kernels launched by rollout / sample_actions calls so far
};

struct Population {
  PopParams c{};
  Arena critic, bc, pi;
  PassBuf PI, TQ, QC, QA, BC, EU;
  BwdBuf bQC, bBC, bPI, bQA;
  float *act_f = nullptr, *vel = nullptr, *xe = nullptr, *target_a = nullptr, *api = nullptr, *rew = nullptr,
        *mask = nullptr, *y = nullptr;
  float* dx0_qa = nullptr;  // [P*E][B][in_pad] f32: dL_q / d input of the critic in the actor pass
  float* info = nullptr;    // [P][kInfoSlots]
  float* leaf_stats = nullptr;  // [P][64][4] per-leaf max / min / sum of squares of the gradients
  float* adam_bc = nullptr;     // [P][2] Adam bias corrections 1-b1^t, 1-b2^t of the current step
  int* adam_step = nullptr;  // [P]
  float* alphas = nullptr;   // [P]
  unsigned long long* seeds = nullptr;  // [P]
  unsigned long long* step_counter = nullptr;  // (dev) [P] per-member step index (Philox counter)
  int *member_list = nullptr, *member_list_e = nullptr, *n_active = nullptr, *n_active_e = nullptr;
  std::vector<int> h_members;  // host copy of the active member ids
  // Window of the active list the launchers cover: the whole list, or one member group while a group's step is
  // being issued (groups step independently of each other, see pop.cu).
  int g_off = 0, g_n = 0;
  int cur_n() const { return g_n; }
  const int* cur_list() const { return member_list + g_off; }
  const int* cur_list_e() const { return member_list_e + (long long)g_off * c.E; }
  Dataset ds;
  NoisePtrs noise;
  Plan* plan = nullptr;
  long long* timeline = nullptr;     // (dev, debug) FQLPOP_TIMELINE=1: per-CTA start/end stamps of every chain launch
  long long* chain_trace = nullptr;  // (dev, debug) FQLPOP_CHAIN_TRACE=1: clock stamps of the Euler chain's CTA 0
  bool use_chain = false;  // layer-chain kernels (chain.cu) instead of one launch per layer
  RolloutScratch ro;
  std::vector<void*> allocs;
};

// ---- pop.cu ------------------------------------------------------------------------------------------
int pop_create_impl(const PopParams& c, const float* alphas, const unsigned long long* seeds, Population** out,
                    std::string* err);
void pop_destroy_impl(Population* p);
int pop_set_active_impl(Population* p, const int* members, int n, std::string* err);
int pop_step_impl(Population* p, int n_steps, cudaStream_t stream, std::string* err);
int pop_total_loss_impl(Population* p, cudaStream_t stream, std::string* err);
int pop_refresh_shadow_impl(Population* p, cudaStream_t stream, std::string* err);
void pop_invalidate_plan(Population* p);
int pop_launches_per_step(const Population* p);
int pop_debug_chain_launch(Population* p, int which, int iters, cudaStream_t stream, std::string* err);
int pop_leaf_io(Population* p, int net, int layer, int kind, int which, int member, float* host, int to_host,
                std::string* err);
// one member's training state as one contiguous device blob (survivor migration between GPUs)
long long pop_member_state_words(const Population* p);
int pop_member_io(Population* p, int member, float* blob_dev, int import, cudaStream_t stream, std::string* err);

// ---- pointwise.cu launchers (all take the population and a stream; grid covers the active members) ----
void launch_prep(Population* p, cudaStream_t s, int bump_adam);
void launch_pi_head(Population* p, cudaStream_t s);
void launch_tq_head(Population* p, cudaStream_t s);
void launch_qc_head(Population* p, cudaStream_t s);
void launch_qa_head(Population* p, cudaStream_t s);
void launch_bc_head(Population* p, cudaStream_t s);
void launch_euler_head(Population* p, int i, cudaStream_t s);
void launch_pi_dout(Population* p, cudaStream_t s);
void launch_small_grads(Population* p, int which, cudaStream_t s);
void launch_grad_stats(Population* p, cudaStream_t s);
void launch_adam(Population* p, int which, cudaStream_t s);  // 0 critic, 1 bc flow, 2 one-step policy
void launch_adam_leaves(Population* p, int which, cudaStream_t s);  // biases / LayerNorm leaves only (kernels: gemm.cu EPI_ADAM)
void launch_finish_info(Population* p, int bump_step, cudaStream_t s);
void launch_refresh_shadow(Population* p, cudaStream_t s);
void launch_to_bf16(const float* src, __nv_bfloat16* dst, long long n4, cudaStream_t s);  // n4 = float4 groups
void launch_fill_qa_dout(Population* p, cudaStream_t s);
int pointwise_init();
int philox_fill(unsigned long long seed, unsigned long long step, int rows, int slots, int kind, unsigned int n,
                void* out_host);
// the same through a fixed file-scope device array instead of cudaMalloc / cudaFree (pointwise.cu)
constexpr int kPhiloxNoAllocMaxWords = 262144;  // 1 MiB of uint32 / fp32 outputs per call
int philox_fill_noalloc(unsigned long long seed, unsigned long long step, int rows, int slots, int kind, unsigned int n,
                        void* out_host, std::string* err);

}  // namespace fqlpop
