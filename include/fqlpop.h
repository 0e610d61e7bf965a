/* fqlpop.h -- C-ABI of the B200-native population FQL hot path (libfqlpop.so).
 *
 * The reference (Mouafak-Khaled/Flow-Q-Learning) has no FFI of its own: its hot path is reached through
 * Python duck-typing (FQLAgent.update at trainer/experiment.py:109, FQLAgent.sample_actions at
 * evaluator/evaluation.py:58-63, OfflineTaskWithSimulatedEvaluations.step at
 * task/offline_task_simulated.py:85-107, Dataset.sample at task/offline_task_simulated.py:53-58).
 * This header is the boundary a maintainer binds with ctypes to replace those calls (INTEGRATION.md shows
 * the stub). Plain pointers and sizes only; every function returns 0 on success and a non-zero status
 * otherwise, with a message available from fqlpop_last_error(). No exceptions cross the ABI.
 * Device pointers are marked (dev); all others are host pointers. `stream` is a cudaStream_t (may be NULL).
 */
#ifndef FQLPOP_H_
#define FQLPOP_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

const char* fqlpop_last_error(void);
int fqlpop_version(void);

/* ---- kernel-level entry: one grouped GEMM with a fused epilogue (unit tests / micro-benchmarks) ------
 * Replaces a single nn.Dense (+GELU(+LayerNorm)) or its transpose inside FQLAgent.update.            */
typedef struct {
  int32_t M, N, K;        /* rows, cols (multiple of 64), reduction (multiple of 64)                   */
  int32_t P;              /* batch = population members                                                */
  int32_t a_mn, b_mn;     /* 0: operand is K-major ([rows][K]); 1: MN-major ([K][rows])                */
  int32_t n_sub;          /* 1: 128x128 tiles; 4: 128x512 tiles (required by the LayerNorm epilogues)  */
  int32_t epi_kind, act;  /* EpiKind / ActKind (see csrc/gemm.h)                                       */
  float ln_eps;
  const void* A; int64_t a_ld, a_ms;   /* (dev) bf16 */
  const void* B; int64_t b_ld, b_ms;   /* (dev) bf16 */
  const float* bias; const float* gamma; const float* beta;          /* (dev) [P][N]                   */
  void* out_h; void* out_gp; void* out_xhat; float* out_rstd;        /* (dev) bf16 [P][M][N]; [P][M]   */
  const void* gp; const void* xhat; const float* rstd;               /* (dev)                          */
  float* out_f32;                                                    /* (dev) [P][M][N]                */
  float* dbias_part; float* dgamma_part; float* dbeta_part;          /* (dev) [P][ceil(M/128)][N]      */
  const int32_t* n_active; const int32_t* member_list;               /* (dev) optional                 */
} fqlpop_gemm_test_t;
int fqlpop_gemm_test(const fqlpop_gemm_test_t* desc, void* stream);

/* ---- population of FQL agents --------------------------------------------------------------------------
 * One handle = P agents (one per BC coefficient alpha) resident on the current CUDA device, stepped in
 * lockstep. Mirrors AgentConfig (/root/reference/trainer/config.py:6-23); ob_dim/action_dim are the two
 * fields FQLAgent.create fills from the example batch (/root/reference/trainer/experiment.py:44-49).      */
typedef struct {
  int32_t ob_dim, action_dim;
  int32_t hidden_actor, hidden_value;       /* uniform hidden width (multiple of 64, <= 512)              */
  int32_t n_hidden_actor, n_hidden_value;   /* number of hidden layers (reference: 4)                     */
  int32_t num_ensembles;                    /* critic ensemble size (upstream FQL: 2)                     */
  int32_t flow_steps, batch_size, population;
  float lr, discount, tau, adam_b1, adam_b2, adam_eps, ln_eps;
  int32_t q_agg_min;                        /* 0: mean over the ensemble ("mean"), 1: "min"               */
  int32_t normalize_q_loss, layer_norm, actor_layer_norm;
} fqlpop_config_t;

/* Replaces FQLAgent.create for every member (trainer/experiment.py:44-49): Glorot-uniform kernels, zero
 * biases, LayerNorm scale 1, target = critic. alphas/seeds: host [population].                            */
int fqlpop_create(const fqlpop_config_t* cfg, const float* alphas, const uint64_t* seeds, void** out_handle);
int fqlpop_destroy(void* handle);

/* Binds the offline dataset the on-device uniform gather reads (replaces Dataset.sample reached through
 * Task.sample("train", B), task/offline_task_simulated.py:53-58). (dev) row-major fp32, borrowed.         */
int fqlpop_set_dataset(void* handle, const float* obs, const float* act, const float* next_obs,
                       const float* rew, const float* mask, int64_t n_rows);
/* Parity mode: explicit row indices [P][B] and noise tensors [P][B][A] / t [P][B] (dev, borrowed).
 * Any NULL pointer falls back to the counter-based Philox stream for that tensor.                         */
int fqlpop_set_noise(void* handle, const int64_t* indices, const float* z_critic, const float* x0,
                     const float* t, const float* z_actor, const float* z_mse);
/* Members that keep training (successive-halving survivors, hpo/successive_halving.py:64-100).           */
int fqlpop_set_active(void* handle, const int32_t* members, int32_t n);

/* n lockstep updates of every active member: gather + FQLAgent.update (trainer/experiment.py:108-109).   */
int fqlpop_step(void* handle, int32_t n_steps, void* stream);
/* FQLAgent.total_loss(batch, grad_params=None) on the bound batch, no update (trainer/experiment.py:115). */
int fqlpop_total_loss(void* handle, void* stream);
/* Info dict of the last step / loss evaluation: host float [population][16], slots as oracle.INFO_KEYS
 * (13 used). with_grad_stats != 0 also reduces grad/max, grad/min, grad/norm. Synchronises.               */
int fqlpop_info(void* handle, float* out_host, int32_t with_grad_stats, void* stream);
/* Same block without the synchronisation: the copy is enqueued on `stream` into PINNED host memory and the caller
 * waits on its own event -- lets a training loop read step i's info while step i+1 runs.                  */
int fqlpop_info_async(void* handle, float* out_pinned, void* stream);

/* Parameter / optimizer-state I/O for one member (utils/agent.py:9-41 load_agent, experiment.py:128-141
 * save_agent). net: 0 critic, 1 actor_bc_flow, 2 actor_onestep_flow; kind: 0 kernel [E][in][out], 1 bias,
 * 2 LayerNorm scale, 3 LayerNorm bias; which: 0 params, 1 Adam mu, 2 Adam nu, 3 target params, 4 grads.
 * host: unpadded fp32 (critic leaves carry the leading ensemble axis). Synchronous.                       */
int fqlpop_leaf(void* handle, int32_t net, int32_t layer, int32_t kind, int32_t which, int32_t member,
                float* host, int32_t to_host);
/* Re-derives the bf16 GEMM copies after parameters were written with fqlpop_leaf.                         */
int fqlpop_refresh_shadow(void* handle, void* stream);
/* Adam step counts [population] and the per-member step counters [population] (position of each member's Philox
 * stream: batch indices and noise of its next update); set != 0 writes. NULL skips an array.                   */
int fqlpop_counters(void* handle, int32_t* adam_steps, uint64_t* step_counter, int32_t set);
/* Survivor migration (the reference is single-process and has no counterpart: after a successive-halving
 * rung, /root/reference/hpo/successive_halving.py:64-100, the members that go on may sit on a few GPUs only).
 * A member's whole training state -- parameters, Adam moments, target critic, Adam step count, alpha, seed and the
 * position of its Philox stream -- as ONE device buffer of fqlpop_member_state_words() 32-bit words, so that the host
 * side moves it with one NCCL send / recv (or cudaMemcpyPeer). export reads slot `member`; import overwrites slot
 * `member` of a population with the same network shapes (checked: a foreign blob fails and writes nothing) and
 * rebuilds its bf16 GEMM copies. Both are stream-ordered on `stream`; a member continues on the importing GPU with
 * exactly the batches and noise it would have drawn where it was.                                          */
int64_t fqlpop_member_state_words(void* handle);
int fqlpop_member_export(void* handle, int32_t member, float* blob_dev, void* stream);
int fqlpop_member_import(void* handle, int32_t member, const float* blob_dev, void* stream);
/* Rewards / masks [population][batch] (host) of the batch the last step gathered: lets a test recover the row
 * indices the on-device uniform gather drew (Dataset.sample, task/offline_task_simulated.py:53-58).         */
int fqlpop_last_batch(void* handle, float* rewards_host, float* masks_host);
/* The device's counter-based Philox4x32-10 stream (every random draw of the path), for bit-exact checks
 * against oracle/philox_oracle.py. kind 0: 4 uint32 words per (row, slot); 1: uniform fp32; 2: index in
 * [0, n) as uint32; 3: normal fp32. out_host: [rows][slots] (x4 words for kind 0).                           */
int fqlpop_philox_fill(uint64_t seed, uint64_t step, int32_t rows, int32_t slots, int32_t kind, uint32_t n,
                       void* out_host);
/* fqlpop_philox_fill without any allocation -- no counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it). Same arguments, same output; the device staging buffer is a fixed 1 MiB file-scope device array
 * (fqlpop_philox_fill_noalloc_max_words() = 262144 output words: rows * slots, x4 for kind 0), which every GPU that
 * loads the library holds until exit. A larger request fails with a clear fqlpop_last_error(); nothing is truncated. The
 * entry point calls no cudaMalloc / cudaFree or allocating library routine. All calls share the array (it keeps the last
 * call's contents; each call overwrites every word it returns): calls that overlap on one GPU race -- serialise them. */
int fqlpop_philox_fill_noalloc(uint64_t seed, uint64_t step, int32_t rows, int32_t slots, int32_t kind, uint32_t n,
                               void* out_host);
int32_t fqlpop_philox_fill_noalloc_max_words(void);
/* Debug: clock64 stamps of CTA 0 of the Euler layer-chain kernel (needs FQLPOP_CHAIN_TRACE=1 at create time):
 * [0,4096) MMA thread (layer start, after each sub-tile's last MMA issue), [4096,8192) first epilogue thread.  */
int fqlpop_debug_chain_trace(void* handle, long long* host_out, int32_t n);
/* Debug: per-CTA records of every layer-chain launch since the last call (needs FQLPOP_TIMELINE=1 at create time):
 * 4 words each -- start ns, end ns (globaltimer), SM id, tag (launch index in the step | first member of the group
 * << 8 | blockIdx << 32). *n_out = number of records copied; the log is reset.                                    */
int fqlpop_debug_timeline(void* handle, long long* host_out, int32_t max_records, int32_t* n_out);
/* Measurement: `iters` back-to-back launches on `stream` of one layer-chain launch of the step over every active
 * member (which = 0: Euler flow chain, 1: one-step policy forward pass, 2: critic forward pass with LayerNorm and
 * saved tensors, 3: critic input-gradient chain, 4: target-critic forward pass). Call after fqlpop_step.        */
int fqlpop_debug_chain_launch(void* handle, int32_t which, int32_t iters, void* stream);
/* Number of kernels one replay of the captured step launches (bench.py's gpu_launches).                   */
int fqlpop_launches_per_step(void* handle);

/* ---- learned world model + fused simulated evaluation ----------------------------------------------------
 * Replaces load_model(...)->apply_fn for the state predictor and the termination predictor
 * (utils/envmodel.py:9-56; modules envmodel/baseline.py:25-37, envmodel/termination_predictor.py:14-21).
 * Kernels are flax layout [in][out] fp32 on the HOST, one pointer per Dense layer (hidden..., output);
 * ln_scale / ln_bias: the state predictor's input LayerNorm over concat(obs, action), [ob_dim+action_dim].  */
typedef struct {
  int32_t ob_dim, action_dim;
  int32_t n_state_hidden, state_hidden[6];
  int32_t n_term_hidden, term_hidden[6];
  float ln_eps;                               /* flax LayerNorm default 1e-6                               */
} fqlpop_envmodel_config_t;
int fqlpop_envmodel_create(const fqlpop_envmodel_config_t* cfg, const float* const* state_kernels,
                           const float* const* state_biases, const float* ln_scale, const float* ln_bias,
                           const float* const* term_kernels, const float* const* term_biases, void** out_handle);
int fqlpop_envmodel_destroy(void* envmodel);

/* One simulated evaluation of every ACTIVE member: evaluate_actor_fn (evaluator/evaluation.py:67-114) driving
 * OfflineTaskWithSimulatedEvaluations.step (task/offline_task_simulated.py:85-107) with the member's one-step
 * policy (FQLAgent.sample_actions). All episodes of all members advance in ONE kernel launch.
 * Arrays marked [P] / [P][n] are indexed by member id (population order), inactive members are left untouched. */
typedef struct {
  int32_t n_episodes;           /* n: parallel episodes per member (num_evaluation_envs)                    */
  int32_t max_episode_steps;    /* truncation (offline_task_simulated.py:23,96-97)                          */
  int32_t horizon;              /* > 0: stop after this many lockstep steps even if episodes are unfinished */
  const float* init_obs;        /* (dev) [n][ob_dim] (+ member * init_obs_member_stride elements)           */
  int64_t init_obs_member_stride;
  const float* noise;           /* (dev) [T][n][action_dim] policy noise (+ member * stride) or NULL=Philox */
  int64_t noise_member_stride;
  uint64_t eval_seed;           /* Philox stream of this evaluation when noise == NULL: the noise of member m,
                                   lockstep step t, episode e, action slot j is the normal draw (fqlpop_philox_fill
                                   kind 3) of key = member seed, counter = (2^63 | (eval_seed & 0x7FFFFFFF) << 32) + t,
                                   row e, slot j                                                            */
  uint8_t* out_success;         /* (dev, optional) [P][n] terminated at the first done step                 */
  int32_t* out_length;          /* (dev, optional) [P][n] step index of the first done step, 0 = none       */
  float* out_success_rate;      /* (dev, optional) [P]  == stats["success"] of evaluate_actor_fn            */
  float* out_return_mean;       /* (dev, optional) [P]  mean of -(length) + success                         */
  int32_t* out_steps;           /* (dev, optional) [P]  lockstep steps executed (max over the member's CTAs)*/
  float* traj_obs;              /* (dev, optional) [P][T][n][ob_dim] observation fed to the policy          */
  float* traj_act;              /* (dev, optional) [P][T][n][action_dim] clipped action                     */
  float* traj_logit;            /* (dev, optional) [P][T][n] termination logit                              */
} fqlpop_rollout_t;             /* T = horizon if 0 < horizon < max_episode_steps else max_episode_steps     */
int fqlpop_rollout(void* handle, void* envmodel, const fqlpop_rollout_t* args, void* stream);

/* FQLAgent.sample_actions(observations, seed) of one member (evaluator/evaluation.py:58-63,94):
 * out[n][action_dim] = clip(pi(obs, z), -1, 1); z = noise[n][action_dim] (dev) or Philox(seed) when NULL.    */
int fqlpop_sample_actions(void* handle, int32_t member, const float* obs, int32_t n, const float* noise,
                          uint64_t seed, float* out, void* stream);
/* Task.step of the simulated task (task/offline_task_simulated.py:85-91): next_obs[n][ob_dim] and the
 * termination logits[n] (terminated = logit > 0) for obs[n][ob_dim], actions[n][action_dim] (all dev).       */
int fqlpop_env_step(void* envmodel, const float* obs, const float* actions, int32_t n, float* next_obs,
                    float* logits, void* stream);
/* Kernels launched so far by fqlpop_rollout / fqlpop_sample_actions on this handle (bench.py's gpu_launches). */
int64_t fqlpop_rollout_launches(void* handle);

/* ---- world-model training ----------------------------------------------------------------
 * Replaces StatePredictorTrainer.train_step / eval_step (envmodel/state_predictor_trainer.py:70-125) for the
 * BaselineStatePredictor (envmodel/baseline.py:25-37) and TerminationPredictorTrainer.train_step / eval_step
 * (envmodel/termination_predictor_trainer.py:58-126) for the TerminationPredictor, with the losses of
 * envmodel/loss.py:7-111 and optax.adam over optax.cosine_decay_schedule(init_learning_rate, steps).
 * Parameters travel as ONE flat fp32 vector: [LayerNorm scale, LayerNorm bias (state predictor only, ob_dim+action_dim
 * each)], then per Dense layer (hidden..., output) kernel [in][out] (flax layout) followed by bias [out].             */
typedef struct {
  int32_t kind;                       /* 0 state predictor, 1 termination predictor, 2 multistep state predictor  *
                                       * (envmodel/multistep.py:34-54: the kind-0 cell scanned over a window; batch   *
                                       * arrays are then obs[batch][T][ob_dim] (only t = 0 is read), act[batch][T]    *
                                       * [action_dim], next_obs[batch][T][ob_dim], rewards[batch][T]; same flat       *
                                       * parameter vector as kind 0; fqlpop_wm_train draws an episode and a window start   *
                                       * per batch row from a resident dataset of whole 1000-step episodes)            */
  int32_t ob_dim, action_dim, batch;
  int32_t n_hidden, hidden[8];
  int32_t term_n_hidden, term_hidden[8]; /* kind 0, termination_weight > 0: the frozen termination predictor      */
  int32_t decay_steps;                /* TrainerConfig.steps (envmodel/config.py:8)                               */
  float lr, ln_eps;                   /* init_learning_rate (config.py:15); flax LayerNorm eps 1e-6               */
  float true_termination_weight, termination_weight, reconstruction_weight;   /* config.py:11-13                 */
  float focal_alpha, focal_gamma;     /* kind 1: focal_loss defaults 0.25 / 2.0 (loss.py:34-35)                  */
  uint64_t seed;                      /* Philox stream of fqlpop_wm_train's sampler                               */
  int32_t seq_len;                    /* kind 2: window length T (TrainerConfig.sequence_length, config.py:17)    */
  int32_t reserved;
} fqlpop_wm_config_t;
/* Number of floats of the flat parameter vector of a network with this config (same for m, v, gradients).  */
int64_t fqlpop_wm_num_params(const fqlpop_wm_config_t* cfg);
int fqlpop_wm_create(const fqlpop_wm_config_t* cfg, const float* params_host, const float* term_params_host,
                     void** out_handle);
int fqlpop_wm_destroy(void* wm);
/* One optimiser step (train != 0) or one forward-only evaluation on an explicit batch (all dev, fp32):
 * obs[batch][ob_dim], act[batch][action_dim] (kind 0 only), next_obs[batch][ob_dim], rewards[batch]
 * (terminated = rewards == 0), dropout_mask[batch][ob_dim] (kind 1 training, optional): the multiplier
 * keep / (1 - 0.1) of nn.Dropout(0.1) -- the reference passes the SAME rng every step, i.e. one fixed mask.       */
int fqlpop_wm_step(void* wm, int32_t train, const float* obs, const float* act, const float* next_obs,
                   const float* rewards, const float* dropout_mask, void* stream);
/* StepLoader.sample on the device (utils/data_loader.py:42-47): borrow a resident dataset of n_rows transitions,
 * then run n_steps optimiser steps, each on `batch` rows drawn uniformly with replacement (Philox(seed, step)).  */
int fqlpop_wm_set_dataset(void* wm, const float* obs, const float* act, const float* next_obs, const float* rewards,
                          int64_t n_rows);
int fqlpop_wm_train(void* wm, int32_t n_steps, const float* dropout_mask, void* stream);
/* out_host[16]: 0 loss, 1 next_observation_loss, 2 termination_loss, 3 true_(termination_)loss, 4 false_(termination_)loss,
 * 5 reconstruction_loss, 6 accuracy, 7 precision, 8 recall, 9 f1, 10 learning rate of the last step, 11 steps taken. */
int fqlpop_wm_logs(void* wm, float* out_host, void* stream);
/* which: 0 parameters, 1 Adam m, 2 Adam v, 3 gradients of the last training step; n = fqlpop_wm_num_params.     */
int fqlpop_wm_export(void* wm, int32_t which, float* out_host, int64_t n, void* stream);
int64_t fqlpop_wm_launches(void* wm);

/* ---- initial-observation generator -------------------------------------------------------
 * The decoder half of InitialObservationEnvModel (envmodel/initial_observation.py:21-30): z -> [Dense, relu]* ->
 * Dense(ob_dim), sampled with z ~ N(0, I) as at distribution_visualizer.py:96-99; replaces real-environment resets
 * (task/offline_task_simulated.py:60-83) as the source of the rollout's initial observations.
 * params_host: per Dense layer (hidden..., output) kernel [in][out] then bias [out], fp32.                          */
int fqlpop_decoder_create(int32_t latent_dim, int32_t n_hidden, const int32_t* hidden, int32_t out_dim,
                          const float* params_host, void** out_handle);
int fqlpop_decoder_destroy(void* decoder);
/* out[n][out_dim] (dev) = decoder(z); z = noise[n][latent_dim] (dev) or Philox normal(seed, 0, row, col) when NULL. */
int fqlpop_decoder_sample(void* decoder, int32_t n, const float* noise, uint64_t seed, float* out, void* stream);

/* Training of that VAE: InitialObservationEnvModel.__call__ + vae_loss (envmodel/initial_observation.py:32-78) --
 *   mu, logvar = encoder(obs); z = mu + eps * exp(logvar / 2); recon = decoder(z);
 *   loss = (reconstruction_weight * mean((obs - recon)^2) + kl) / (reconstruction_weight + 1),
 *   kl = -0.5 * mean(1 + logvar - mu^2 - exp(logvar))
 * -- with backward pass and optimiser on the device. The reference ships the model and the loss but no trainer for it
 * (train_env_model.py does not know the model): the optimiser is the one of its other world-model trainers,
 * optax.adam over optax.cosine_decay_schedule(lr, decay_steps) (envmodel/state_predictor_trainer.py:57-61).
 * Flat parameter vector (params_host, fqlpop_vae_export), fp32, kernel [in][out] then bias per Dense layer: encoder
 * hidden layers, ONE head layer of 2 * latent_dim columns (columns [0, L) = the reference's mean head Dense_h, columns
 * [L, 2L) = its logvar head Dense_{h+1}), then the decoder (hidden dims reversed, output ob_dim), which is exactly the
 * vector fqlpop_decoder_create takes.                                                                              */
typedef struct {
  int32_t ob_dim, latent_dim, batch;
  int32_t n_hidden, hidden[8];          /* encoder order, as initial_observation_config.yaml stores hidden_dims */
  int32_t decay_steps;
  float lr, reconstruction_weight;
  uint64_t seed;                        /* Philox stream of the resident-dataset sampler and of the latent noise */
} fqlpop_vae_config_t;
int64_t fqlpop_vae_num_params(const fqlpop_vae_config_t* cfg);
int fqlpop_vae_create(const fqlpop_vae_config_t* cfg, const float* params_host, void** out_handle);
int fqlpop_vae_destroy(void* vae);
/* One optimiser step (train != 0) or a forward-only evaluation on obs[batch][ob_dim] (dev). eps[batch][latent_dim]
 * (dev) is the reparameterisation noise; NULL draws it from Philox normal(seed, step, row, 1 + column).          */
int fqlpop_vae_step(void* vae, int32_t train, const float* obs, const float* eps, void* stream);
/* Resident observations (borrowed device pointer, n_rows rows) + n_steps optimiser steps, each on `batch` rows drawn
 * uniformly with replacement (Philox(seed, step, row, 0)).                                                       */
int fqlpop_vae_set_dataset(void* vae, const float* obs, int64_t n_rows);
int fqlpop_vae_train(void* vae, int32_t n_steps, void* stream);
/* out_host[16]: 0 loss, 5 reconstruction_loss, 12 kl, 10 learning rate of the last step, 11 steps taken.         */
int fqlpop_vae_logs(void* vae, float* out_host, void* stream);
/* which: 0 parameters, 1 Adam m, 2 Adam v, 3 gradients of the last training step; n = fqlpop_vae_num_params.     */
int fqlpop_vae_export(void* vae, int32_t which, float* out_host, int64_t n, void* stream);
int64_t fqlpop_vae_launches(void* vae);

#ifdef __cplusplus
}
#endif
#endif /* FQLPOP_H_ */
