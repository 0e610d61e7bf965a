// World-model training on the device: the state predictor and the termination predictor that the
// fused rollout (rollout.cu) evaluates are trained here, so that the simulated task has weights without JAX.
//
// Reference path this replaces: StatePredictorTrainer.train_step / eval_step
// (/root/reference/envmodel/state_predictor_trainer.py:70-125), TerminationPredictorTrainer.train_step / eval_step
// (/root/reference/envmodel/termination_predictor_trainer.py:58-126), the losses of /root/reference/envmodel/loss.py:7-111,
// optax.adam over optax.cosine_decay_schedule (state_predictor_trainer.py:57-61), and StepLoader.sample
// (/root/reference/utils/data_loader.py:42-47) when the dataset is resident.
//
// The networks are tiny (50 -> 128 -> 256 -> 128 -> 42 at batch 256: 0.12 GFLOP per step), so a step is bound by
// launch latency and not by any pipe: the kernels are plain fp32 (bit-for-bit the reference's arithmetic type, no
// bf16 rounding), one CUDA graph per step, parameters / moments / gradients in three flat arenas so that the optimiser
// is one elementwise launch.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

namespace fqlpop {

constexpr int kWmMaxLayers = 8;  // Dense layers of one network, output layer included
constexpr int kWmLogs = 16;

enum WmKind : int { WM_STATE = 0, WM_TERMINATION = 1, WM_MULTISTEP = 2 };

// logs[] slots (float). State predictor: the keys of state_prediction_loss's logs (envmodel/loss.py:98-109).
// Termination predictor: focal_loss's logs + the eval_step metrics (termination_predictor_trainer.py:88-126).
enum WmLog : int {
  WL_LOSS = 0,
  WL_NEXT_OBS_LOSS = 1,     // state
  WL_TERM_LOSS = 2,         // state
  WL_TRUE_LOSS = 3,         // state: true_termination_loss; termination: true_loss
  WL_FALSE_LOSS = 4,
  WL_RECON_LOSS = 5,        // state (0: the baseline model returns its input as the reconstruction)
  WL_ACCURACY = 6,          // termination, eval
  WL_PRECISION = 7,
  WL_RECALL = 8,
  WL_F1 = 9,
  WL_LR = 10,               // learning rate the last train step used
  WL_STEP = 11,             // optimiser steps taken (as float)
  WL_KL = 12,               // VAE: KL term (WL_RECON_LOSS holds its reconstruction loss)
};

struct WmNet {
  int n_layers;                  // Dense layers incl. the output layer
  int dims[kWmMaxLayers + 1];    // dims[0] = input width ... dims[n_layers] = output width
  long long off_ln_scale, off_ln_bias;  // -1: no input LayerNorm
  long long offW[kWmMaxLayers], offB[kWmMaxLayers];  // kernel [in][out] (flax layout), bias [out]
  long long n_params;
};

struct WmConfig {
  int kind, O, A, batch;
  int seq_len;  // WM_MULTISTEP: window length T (batch arrays are [batch][T][...]); 1 otherwise
  int n_hidden, hidden[kWmMaxLayers];
  int term_n_hidden, term_hidden[kWmMaxLayers];  // frozen termination predictor inside the state loss (0 = none)
  int decay_steps;
  float lr, ln_eps;
  float true_termination_weight, termination_weight, reconstruction_weight;
  float focal_alpha, focal_gamma;
  unsigned long long seed;  // Philox stream of the resident-dataset sampler
};

struct WmTrainer;

// Layout of the flat parameter vector of a network (what params_host / fqlpop_wm_export use).
WmNet wm_layout(int in_dim, int n_hidden, const int* hidden, int out_dim, bool input_ln);

int wm_create_impl(const WmConfig& cfg, const float* params_host, const float* term_params_host, WmTrainer** out,
                   std::string* err);
void wm_destroy_impl(WmTrainer* t);
// One optimiser step (train != 0) or a forward-only evaluation on an explicit device batch. dropout_mask: (dev, optional,
// termination training) [batch][O] multiplier already holding keep / (1 - rate).
int wm_step_impl(WmTrainer* t, int train, const float* obs, const float* act, const float* next_obs,
                 const float* rewards, const float* dropout_mask, cudaStream_t stream, std::string* err);
// Resident dataset (borrowed device pointers, N rows) + n_steps optimiser steps with on-device uniform sampling.
int wm_set_dataset_impl(WmTrainer* t, const float* obs, const float* act, const float* next_obs, const float* rewards,
                        long long n_rows, std::string* err);
int wm_train_impl(WmTrainer* t, int n_steps, const float* dropout_mask, cudaStream_t stream, std::string* err);
int wm_logs_impl(WmTrainer* t, float* out_host, cudaStream_t stream, std::string* err);
// which: 0 params, 1 Adam m, 2 Adam v, 3 gradients of the last train step
int wm_export_impl(WmTrainer* t, int which, float* out_host, long long n, cudaStream_t stream, std::string* err);
long long wm_num_params(const WmTrainer* t);

// Initial-observation generator: the VAE decoder z -> [Dense, relu]* -> Dense(O)
// (/root/reference/envmodel/initial_observation.py:21-30, sampled as at /root/reference/distribution_visualizer.py:96-99).
struct WmDecoder;
int wm_decoder_create_impl(int latent_dim, int n_hidden, const int* hidden, int out_dim, const float* params_host,
                           WmDecoder** out, std::string* err);
void wm_decoder_destroy_impl(WmDecoder* d);
// out[n][out_dim] = decoder(z); z = noise[n][latent_dim] (dev) or N(0, I) from Philox(seed) when noise == nullptr
int wm_decoder_sample_impl(WmDecoder* d, int n, const float* noise, unsigned long long seed, float* out,
                           cudaStream_t stream, std::string* err);
long long wm_launches(const WmTrainer* t);

// Training of the initial-observation VAE (/root/reference/envmodel/initial_observation.py:9-78; wm_vae.cuh).
// Flat parameter vector: encoder hidden layers, ONE head layer of 2 * L columns (mean | logvar), then the decoder
// (hidden dims reversed).
struct VaeConfig {
  int O, L, batch;
  int n_hidden, hidden[kWmMaxLayers];   // encoder order, as the reference's config stores them
  int decay_steps;
  float lr, reconstruction_weight;
  unsigned long long seed;              // Philox stream of the resident-dataset sampler and of the latent noise
};
struct WmVae;
long long vae_num_params(const VaeConfig& c);
int vae_create_impl(const VaeConfig& c, const float* params_host, WmVae** out, std::string* err);
void vae_destroy_impl(WmVae* t);
// explicit device batch obs[batch][O]; eps[batch][L] (dev) or nullptr -> Philox normal
int vae_step_impl(WmVae* t, int train, const float* obs, const float* eps, cudaStream_t stream, std::string* err);
int vae_set_dataset_impl(WmVae* t, const float* obs, long long n_rows, std::string* err);
int vae_train_impl(WmVae* t, int n_steps, cudaStream_t stream, std::string* err);
int vae_logs_impl(WmVae* t, float* out_host, cudaStream_t stream, std::string* err);
int vae_export_impl(WmVae* t, int which, float* out_host, long long n, cudaStream_t stream, std::string* err);
long long vae_launches(const WmVae* t);

}  // namespace fqlpop
