// Fused persistent world-model rollout (sm_100a): one CTA advances 128 simulated episodes of one population
// member through every lockstep step of an evaluation without leaving the SM -- one-step policy
// (obs, z) -> 512^4 -> action, state predictor LN(concat(o, a)) -> relu MLP -> + o, termination predictor
// relu MLP -> logit > 0, done / invalid / success bookkeeping -- with the weights streamed from L2 by TMA,
// the matrix products on tcgen05 (fp32 accumulators in TMEM) and the activations resident in shared memory.
//
// Reference path this replaces (one Python iteration per step, per-element device syncs):
//   evaluate_actor_fn                              /root/reference/evaluator/evaluation.py:67-114
//   OfflineTaskWithSimulatedEvaluations.step       /root/reference/task/offline_task_simulated.py:85-107
//   BaselineStatePredictor.__call__                /root/reference/envmodel/baseline.py:25-37
//   TerminationPredictor.__call__                  /root/reference/envmodel/termination_predictor.py:14-21
//   FQLAgent.sample_actions                        /root/reference/evaluator/evaluation.py:58-63,94
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

namespace fqlpop {

struct Population;

constexpr int kRoMaxLayers = 20;
constexpr int kRoMaxHidden = 6;  // hidden layers per world-model network

enum RoKind : int {
  RO_HIDDEN = 0,     // h = act(acc + b) (+ LayerNorm) -> next layer's A operand in shared memory
  RO_PI_HEAD = 1,    // a = clip(acc + b, -1, 1); then LN(concat(o, a)) -> state-predictor input
  RO_STATE_OUT = 2,  // o' = acc + b + o; -> termination-predictor input
  RO_TERM_OUT = 3,   // logit = acc + b; terminated = logit > 0; episode bookkeeping; next policy input
};
enum RoMode : int {
  RO_MODE_ROLLOUT = 0,  // policy + state + termination, many steps
  RO_MODE_POLICY = 1,   // FQLAgent.sample_actions only (one step, actions out)
  RO_MODE_ENV = 2,      // Task.step only (one step, actions in; next_obs + logits out)
};

struct RoLayer {
  int kind, act, ln, a_from_x0;
  int k_blocks;   // reduction length / 64 (padded)
  int n_pad;      // output columns, padded to a multiple of 64 (<= 512)
  int n_true;     // valid output columns
  int per_member; // 1: weights / biases are indexed by the population member
  const float* bias;   long long bias_ms;   // fp32 [n_pad] (+ member * bias_ms)
  const float* gamma;  const float* beta;   // LayerNorm after the activation (policy with actor_layer_norm)
};

struct RoParams {
  CUtensorMap maps[kRoMaxLayers];  // bf16 weights [K_pad][N_pad] (N contiguous), box 64 x 64, 128-B swizzle
  RoLayer layer[kRoMaxLayers];
  int n_layers, mode;
  int O, A, n;          // observation / action dims, episodes per member
  int tiles_per_member; // ceil(n / 128)
  int n_steps;          // lockstep steps to run at most (<= max_episode_steps)
  int max_episode_steps;
  float ln_eps_env;     // eps of the state predictor's input LayerNorm (flax default 1e-6)
  float ln_eps_pi;
  const float* env_ln_scale; const float* env_ln_bias;  // [O + A]
  const int* member_list;  // (dev) active member ids, grid.y-major order; nullptr -> identity
  const unsigned long long* seeds;  // (dev) [P] member seeds (Philox keys)
  unsigned long long eval_counter;  // Philox counter base of this evaluation
  // inputs
  const float* init_obs;  long long init_obs_ms;  // [n][O] (+ member * ms; ms = 0 -> shared)
  const float* noise;     long long noise_ms;     // [T][n][A] explicit policy noise or nullptr -> Philox
  const float* actions_in;                        // RO_MODE_ENV: [n][A]
  // state + outputs (all indexed by member id)
  float* obs_work;         // [P][n][O] fp32 current observations (scratch, also the final observations)
  float* noise_next;       // [P][2][n][A] scratch: Philox policy noise of step s at [s & 1], drawn during step s - 1
  unsigned char* success;  // [P][n]
  int* length;             // [P][n] step of the first termination / truncation (0 = never within n_steps)
  float* success_sum;      // [P] sum of success flags   (host divides by n)
  float* return_sum;       // [P] sum of returns -(length) + success
  int* steps_run;          // [P] max over CTAs of lockstep steps executed
  float* traj_obs;         // optional [P][T][n][O]: observation fed to the policy at step t
  float* traj_act;         // optional [P][T][n][A]: clipped action
  float* traj_logit;       // optional [P][T][n]
  float* actions_out;      // RO_MODE_POLICY: [n][A]
  float* next_obs_out;     // RO_MODE_ENV: [n][O]
  float* logits_out;       // RO_MODE_ENV: [n]
  int debug_flags;         // (make TRACE=1, rollout16.cu) 1: no MMAs, 2: no weight TMA -- timing only, results are garbage
};

// Device-resident world model (shared by every member): bf16 padded kernels for the tensor cores, fp32 biases.
struct EnvModel {
  int O = 0, A = 0;
  int n_state_hidden = 0, n_term_hidden = 0;
  int state_hidden[kRoMaxHidden] = {0}, term_hidden[kRoMaxHidden] = {0};
  float ln_eps = 1e-6f;
  // per layer (hidden..., output): padded dims and device buffers
  struct Layer {
    int K = 0, N = 0, K_pad = 0, N_pad = 0;
    __nv_bfloat16* w = nullptr;  // [K_pad][N_pad]
    __nv_bfloat16* wt = nullptr;   // the same, tiled [K_pad / 32][N_pad / 64][32][64] (rollout16.cu)
    __nv_bfloat16* wt64 = nullptr; // ... and tiled [K_pad / 64][N_pad / 64][64][64] (rollout16.cu, pair mode)
    float* b = nullptr;          // [N_pad]
  };
  std::vector<Layer> state, term;
  float *ln_scale = nullptr, *ln_bias = nullptr;  // [O + A]
  std::vector<void*> allocs;
};

int envmodel_create_impl(int O, int A, int n_state_hidden, const int* state_hidden, int n_term_hidden,
                         const int* term_hidden, float ln_eps, const float* const* state_kernels,
                         const float* const* state_biases, const float* ln_scale, const float* ln_bias,
                         const float* const* term_kernels, const float* const* term_biases, EnvModel** out,
                         std::string* err);
void envmodel_destroy_impl(EnvModel* m);

struct RolloutArgs {
  int n_episodes = 0, max_episode_steps = 0, horizon = 0;
  const float* init_obs = nullptr; long long init_obs_ms = 0;
  const float* noise = nullptr; long long noise_ms = 0;
  unsigned long long eval_seed = 0;
  unsigned char* success = nullptr; int* length = nullptr;
  float* success_rate = nullptr; float* return_mean = nullptr; int* steps_run = nullptr;
  float *traj_obs = nullptr, *traj_act = nullptr, *traj_logit = nullptr;
};
int rollout_impl(Population* P, EnvModel* m, const RolloutArgs& a, cudaStream_t stream, std::string* err);
int sample_actions_impl(Population* P, int member, const float* obs, int n, const float* noise,
                        unsigned long long seed, float* out, cudaStream_t stream, std::string* err);
int env_step_impl(EnvModel* m, const float* obs, const float* act, int n, float* next_obs, float* logits,
                  cudaStream_t stream, std::string* err);
cudaError_t rollout_init();
// rollout16.cu: second-generation whole-rollout kernel (16 epilogue warps)
cudaError_t rollout16_init();
bool rollout16_eligible(const RoParams& R);
cudaError_t launch_rollout16(const RoParams& R, int n_members, bool pair, cudaStream_t stream);

}  // namespace fqlpop
