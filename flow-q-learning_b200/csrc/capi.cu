// C-ABI of libfqlpop.so (declared in include/fqlpop.h).
#include "../../include/fqlpop.h"
#include "gemm.h"

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {
thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return 1;
}
}  // namespace

namespace fqlpop {
void set_last_error(const std::string& m) { g_err = m; }
}

extern "C" const char* fqlpop_last_error(void) { return g_err.c_str(); }
extern "C" int fqlpop_version(void) { return 1; }

extern "C" int fqlpop_gemm_test(const fqlpop_gemm_test_t* d, void* stream) {
  using namespace fqlpop;
  if (!d) return fail("null desc");
  if (d->N % 64 || d->K % 64) return fail("N and K must be multiples of 64");
  if (d->n_sub != 1 && d->n_sub != 4) return fail("n_sub must be 1 or 4");
  if (cudaError_t e = gemm_init(); e != cudaSuccess) return fail(std::string("gemm_init: ") + cudaGetErrorString(e));
  static GemmLaunch L;  // ~8 KB: keep off the stack
  memset(&L, 0, sizeof(L));
  L.num_groups = 1;
  L.P = d->P;
  L.n_active = d->n_active;
  L.member_list = d->member_list;
  GemmGroup& g = L.g[0];
  g.M = d->M;
  g.N = d->N;
  g.k_blocks = d->K / 64;
  g.a_mn = d->a_mn;
  g.b_mn = d->b_mn;
  int rc;
  if (!d->a_mn) rc = encode_tmap_bf16(&g.mapA, d->A, d->K, d->M, d->P, d->a_ld, d->a_ms, 128);
  else rc = encode_tmap_bf16(&g.mapA, d->A, d->M, d->K, d->P, d->a_ld, d->a_ms, 64);
  if (rc) return fail("encode mapA failed rc=" + std::to_string(rc));
  if (!d->b_mn) rc = encode_tmap_bf16(&g.mapB, d->B, d->K, d->N, d->P, d->b_ld, d->b_ms, 128);
  else rc = encode_tmap_bf16(&g.mapB, d->B, d->N, d->K, d->P, d->b_ld, d->b_ms, 64);
  if (rc) return fail("encode mapB failed rc=" + std::to_string(rc));
  EpiParams& E = g.epi;
  E.kind = d->epi_kind;
  E.act = d->act;
  E.ln_eps = d->ln_eps;
  const long long MN = (long long)d->M * d->N;
  const int mt = (d->M + 127) / 128;
  E.bias = {d->bias, d->N, d->N, 0};
  E.gamma = {d->gamma, d->N, d->N, 0};
  E.beta = {d->beta, d->N, d->N, 0};
  E.out_h = {(__nv_bfloat16*)d->out_h, MN, d->N, 0};
  E.out_gp = {(__nv_bfloat16*)d->out_gp, MN, d->N, 0};
  E.out_xhat = {(__nv_bfloat16*)d->out_xhat, MN, d->N, 0};
  E.out_rstd = {d->out_rstd, d->M, 1, 0};
  E.gp = {(const __nv_bfloat16*)d->gp, MN, d->N, 0};
  E.xhat = {(const __nv_bfloat16*)d->xhat, MN, d->N, 0};
  E.rstd = {d->rstd, d->M, 1, 0};
  E.out_f32 = {d->out_f32, MN, d->N, 0};
  E.out_dbias_part = {d->dbias_part, (long long)mt * d->N, d->N, 0};
  E.out_dgamma_part = {d->dgamma_part, (long long)mt * d->N, d->N, 0};
  E.out_dbeta_part = {d->dbeta_part, (long long)mt * d->N, d->N, 0};
  const int grid = finalize_launch(&L, d->n_sub);
  cudaError_t e = launch_gemm(L, d->n_sub, grid, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(std::string("launch: ") + cudaGetErrorString(e));
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// population
// ---------------------------------------------------------------------------------------------------------
#include "pop.h"

#define POP(h)                                   \
  fqlpop::Population* P = (fqlpop::Population*)(h); \
  if (!P) return fail("null handle")

extern "C" int fqlpop_create(const fqlpop_config_t* c, const float* alphas, const uint64_t* seeds, void** out) {
  if (!c || !alphas || !seeds || !out) return fail("null argument");
  fqlpop::PopParams p{};
  p.O = c->ob_dim; p.A = c->action_dim; p.H_actor = c->hidden_actor; p.H_value = c->hidden_value;
  p.n_hidden_actor = c->n_hidden_actor; p.n_hidden_value = c->n_hidden_value; p.E = c->num_ensembles;
  p.F = c->flow_steps; p.B = c->batch_size; p.P = c->population;
  p.lr = c->lr; p.discount = c->discount; p.tau = c->tau; p.adam_b1 = c->adam_b1; p.adam_b2 = c->adam_b2;
  p.adam_eps = c->adam_eps; p.ln_eps = c->ln_eps;
  p.q_agg_min = c->q_agg_min; p.normalize_q_loss = c->normalize_q_loss; p.layer_norm = c->layer_norm;
  p.actor_layer_norm = c->actor_layer_norm;
  fqlpop::Population* P = nullptr;
  std::string err;
  if (fqlpop::pop_create_impl(p, alphas, (const unsigned long long*)seeds, &P, &err)) return fail(err);
  *out = P;
  return 0;
}
extern "C" int fqlpop_destroy(void* h) {
  fqlpop::pop_destroy_impl((fqlpop::Population*)h);
  return 0;
}
extern "C" int fqlpop_set_dataset(void* h, const float* obs, const float* act, const float* next_obs,
                                  const float* rew, const float* mask, int64_t n) {
  POP(h);
  if (!obs || !act || !next_obs || !rew || !mask || n <= 0 || n >= (1ll << 32)) return fail("bad dataset");
  fqlpop::pop_invalidate_plan(P);
  P->ds = fqlpop::Dataset{obs, act, next_obs, rew, mask, (long long)n};
  return 0;
}
extern "C" int fqlpop_set_noise(void* h, const int64_t* indices, const float* z_critic, const float* x0,
                                const float* t, const float* z_actor, const float* z_mse) {
  POP(h);
  fqlpop::pop_invalidate_plan(P);
  P->noise.indices = (const long long*)indices;
  P->noise.z_critic = z_critic; P->noise.x0 = x0; P->noise.t = t; P->noise.z_actor = z_actor; P->noise.z_mse = z_mse;
  return 0;
}
extern "C" int fqlpop_set_active(void* h, const int32_t* members, int32_t n) {
  POP(h);
  std::string err;
  if (fqlpop::pop_set_active_impl(P, members, n, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_step(void* h, int32_t n_steps, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_step_impl(P, n_steps, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_total_loss(void* h, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_total_loss_impl(P, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_info(void* h, float* out_host, int32_t with_grad_stats, void* stream) {
  POP(h);
  if (!out_host) return fail("null output");
  cudaStream_t s = (cudaStream_t)stream;
  if (with_grad_stats) fqlpop::launch_grad_stats(P, s);
  cudaError_t e = cudaMemcpyAsync(out_host, P->info, sizeof(float) * P->c.P * fqlpop::kInfoSlots,
                                  cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) return fail(std::string("fqlpop_info: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" int fqlpop_info_async(void* h, float* out_pinned, void* stream) {
  POP(h);
  if (!out_pinned) return fail("null output");
  cudaError_t e = cudaMemcpyAsync(out_pinned, P->info, sizeof(float) * P->c.P * fqlpop::kInfoSlots,
                                  cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e != cudaSuccess) return fail(std::string("fqlpop_info_async: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" int fqlpop_leaf(void* h, int32_t net, int32_t layer, int32_t kind, int32_t which, int32_t member,
                           float* host, int32_t to_host) {
  POP(h);
  if (!host) return fail("null host buffer");
  cudaDeviceSynchronize();
  std::string err;
  if (fqlpop::pop_leaf_io(P, net, layer, kind, which, member, host, to_host, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_refresh_shadow(void* h, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_refresh_shadow_impl(P, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int64_t fqlpop_member_state_words(void* h) {
  if (!h) return -1;
  return fqlpop::pop_member_state_words(static_cast<fqlpop::Population*>(h));
}
extern "C" int fqlpop_member_export(void* h, int32_t member, float* blob_dev, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_member_io(P, member, blob_dev, 0, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_member_import(void* h, int32_t member, const float* blob_dev, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_member_io(P, member, const_cast<float*>(blob_dev), 1, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_counters(void* h, int32_t* adam_steps, uint64_t* step_counter, int32_t set) {
  POP(h);
  cudaDeviceSynchronize();
  cudaError_t e = cudaSuccess;
  const cudaMemcpyKind k = set ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  if (adam_steps) e = set ? cudaMemcpy(P->adam_step, adam_steps, sizeof(int) * P->c.P, k)
                          : cudaMemcpy(adam_steps, P->adam_step, sizeof(int) * P->c.P, k);
  // the Philox stream position is kept per member (member groups step independently; a resumed member continues its
  // own stream): [population] in both directions
  if (e == cudaSuccess && step_counter)
    e = set ? cudaMemcpy(P->step_counter, step_counter, sizeof(uint64_t) * P->c.P, k)
            : cudaMemcpy(step_counter, P->step_counter, sizeof(uint64_t) * P->c.P, k);
  if (e != cudaSuccess) return fail(std::string("fqlpop_counters: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" int fqlpop_last_batch(void* h, float* rewards_host, float* masks_host) {
  POP(h);
  cudaDeviceSynchronize();
  const size_t bytes = sizeof(float) * (size_t)P->c.P * P->c.B;
  cudaError_t e = cudaSuccess;
  if (rewards_host) e = cudaMemcpy(rewards_host, P->rew, bytes, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && masks_host) e = cudaMemcpy(masks_host, P->mask, bytes, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return fail(std::string("fqlpop_last_batch: ") + cudaGetErrorString(e));
  return 0;
}
extern "C" int fqlpop_philox_fill(uint64_t seed, uint64_t step, int32_t rows, int32_t slots, int32_t kind, uint32_t n,
                                  void* out_host) {
  if (!out_host || rows < 1 || slots < 1 || kind < 0 || kind > 3) return fail("philox_fill: bad arguments");
  if (fqlpop::philox_fill(seed, step, rows, slots, kind, n, out_host)) return fail("philox_fill: CUDA error");
  return 0;
}
extern "C" int fqlpop_philox_fill_noalloc(uint64_t seed, uint64_t step, int32_t rows, int32_t slots, int32_t kind, uint32_t n,
                                          void* out_host) {
  if (!out_host || rows < 1 || slots < 1 || kind < 0 || kind > 3) return fail("philox_fill_noalloc: bad arguments");
  std::string err;
  if (fqlpop::philox_fill_noalloc(seed, step, rows, slots, kind, n, out_host, &err)) return fail(err);
  return 0;
}
extern "C" int32_t fqlpop_philox_fill_noalloc_max_words(void) { return fqlpop::kPhiloxNoAllocMaxWords; }
extern "C" int fqlpop_debug_chain_trace(void* h, long long* host_out, int32_t n) {
  POP(h);
  if (!P->chain_trace) return fail("chain trace disabled (set FQLPOP_CHAIN_TRACE=1 before fqlpop_create)");
  cudaDeviceSynchronize();
  if (cudaMemcpy(host_out, P->chain_trace, sizeof(long long) * (n < 8192 ? n : 8192), cudaMemcpyDeviceToHost) != cudaSuccess)
    return fail("chain trace copy failed");
  return 0;
}
extern "C" int fqlpop_debug_timeline(void* h, long long* host_out, int32_t max_records, int32_t* n_out) {
  POP(h);
  if (!P->timeline) return fail("timeline disabled (set FQLPOP_TIMELINE=1 before fqlpop_create)");
  cudaDeviceSynchronize();
  long long n = 0;
  if (cudaMemcpy(&n, P->timeline, sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) return fail("timeline copy failed");
  if (n > 65536) n = 65536;
  if (n > max_records) n = max_records;
  if (n > 0 && cudaMemcpy(host_out, P->timeline + 8, sizeof(long long) * 4 * n, cudaMemcpyDeviceToHost) != cudaSuccess)
    return fail("timeline copy failed");
  cudaMemset(P->timeline, 0, sizeof(long long));
  if (n_out) *n_out = (int32_t)n;
  return 0;
}
extern "C" int fqlpop_debug_chain_launch(void* h, int32_t which, int32_t iters, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::pop_debug_chain_launch(P, which, iters, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_launches_per_step(void* h) {
  fqlpop::Population* P = (fqlpop::Population*)h;
  return P ? fqlpop::pop_launches_per_step(P) : 0;
}

// ---------------------------------------------------------------------------------------------------------
// world model + fused rollout
// ---------------------------------------------------------------------------------------------------------
#include "rollout.h"

extern "C" int fqlpop_envmodel_create(const fqlpop_envmodel_config_t* c, const float* const* state_kernels,
                                      const float* const* state_biases, const float* ln_scale, const float* ln_bias,
                                      const float* const* term_kernels, const float* const* term_biases, void** out) {
  if (!c || !state_kernels || !state_biases || !ln_scale || !ln_bias || !term_kernels || !term_biases || !out)
    return fail("null argument");
  if (c->n_state_hidden < 1 || c->n_state_hidden > 6 || c->n_term_hidden < 1 || c->n_term_hidden > 6)
    return fail("world model: between 1 and 6 hidden layers per network");
  fqlpop::EnvModel* m = nullptr;
  std::string err;
  if (fqlpop::envmodel_create_impl(c->ob_dim, c->action_dim, c->n_state_hidden, c->state_hidden, c->n_term_hidden,
                                   c->term_hidden, c->ln_eps, state_kernels, state_biases, ln_scale, ln_bias,
                                   term_kernels, term_biases, &m, &err))
    return fail(err);
  *out = m;
  return 0;
}
extern "C" int fqlpop_envmodel_destroy(void* m) {
  fqlpop::envmodel_destroy_impl((fqlpop::EnvModel*)m);
  return 0;
}
extern "C" int fqlpop_rollout(void* h, void* envmodel, const fqlpop_rollout_t* a, void* stream) {
  POP(h);
  if (!envmodel || !a) return fail("null argument");
  fqlpop::RolloutArgs r;
  r.n_episodes = a->n_episodes; r.max_episode_steps = a->max_episode_steps; r.horizon = a->horizon;
  r.init_obs = a->init_obs; r.init_obs_ms = a->init_obs_member_stride;
  r.noise = a->noise; r.noise_ms = a->noise_member_stride;
  r.eval_seed = a->eval_seed;
  r.success = a->out_success; r.length = a->out_length;
  r.success_rate = a->out_success_rate; r.return_mean = a->out_return_mean; r.steps_run = a->out_steps;
  r.traj_obs = a->traj_obs; r.traj_act = a->traj_act; r.traj_logit = a->traj_logit;
  std::string err;
  if (fqlpop::rollout_impl(P, (fqlpop::EnvModel*)envmodel, r, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_sample_actions(void* h, int32_t member, const float* obs, int32_t n, const float* noise,
                                     uint64_t seed, float* out, void* stream) {
  POP(h);
  std::string err;
  if (fqlpop::sample_actions_impl(P, member, obs, n, noise, seed, out, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_env_step(void* envmodel, const float* obs, const float* actions, int32_t n, float* next_obs,
                               float* logits, void* stream) {
  std::string err;
  if (fqlpop::env_step_impl((fqlpop::EnvModel*)envmodel, obs, actions, n, next_obs, logits, (cudaStream_t)stream, &err))
    return fail(err);
  return 0;
}
extern "C" int64_t fqlpop_rollout_launches(void* h) {
  fqlpop::Population* P = (fqlpop::Population*)h;
  return P ? P->ro.launches : 0;
}


// ---------------------------------------------------------------------------------------------------------
// world-model training
// ---------------------------------------------------------------------------------------------------------
#include "wm.h"

namespace {
bool wm_cfg(const fqlpop_wm_config_t* c, fqlpop::WmConfig* o) {
  if (!c || c->n_hidden < 1 || c->n_hidden >= fqlpop::kWmMaxLayers || c->term_n_hidden < 0 ||
      c->term_n_hidden >= fqlpop::kWmMaxLayers)
    return false;
  memset(o, 0, sizeof(*o));
  o->kind = c->kind; o->O = c->ob_dim; o->A = c->action_dim; o->batch = c->batch;
  o->n_hidden = c->n_hidden; o->term_n_hidden = c->term_n_hidden;
  for (int i = 0; i < c->n_hidden; ++i) o->hidden[i] = c->hidden[i];
  for (int i = 0; i < c->term_n_hidden; ++i) o->term_hidden[i] = c->term_hidden[i];
  o->decay_steps = c->decay_steps; o->lr = c->lr; o->ln_eps = c->ln_eps;
  o->true_termination_weight = c->true_termination_weight; o->termination_weight = c->termination_weight;
  o->reconstruction_weight = c->reconstruction_weight;
  o->focal_alpha = c->focal_alpha; o->focal_gamma = c->focal_gamma; o->seed = c->seed;
  o->seq_len = c->seq_len;
  return true;
}
}  // namespace
#define WM(h)                                      \
  fqlpop::WmTrainer* T = (fqlpop::WmTrainer*)(h);  \
  if (!T) return fail("null world-model handle");  \
  std::string err

extern "C" int64_t fqlpop_wm_num_params(const fqlpop_wm_config_t* c) {
  fqlpop::WmConfig w;
  if (!wm_cfg(c, &w)) return -1;
  return (w.kind != fqlpop::WM_TERMINATION ? fqlpop::wm_layout(w.O + w.A, w.n_hidden, w.hidden, w.O, true)
                                     : fqlpop::wm_layout(w.O, w.n_hidden, w.hidden, 1, false)).n_params;
}
extern "C" int fqlpop_wm_create(const fqlpop_wm_config_t* c, const float* params_host, const float* term_params_host,
                                void** out) {
  fqlpop::WmConfig w;
  if (!out || !wm_cfg(c, &w)) return fail("world model: null argument or layer count outside [1, 7]");
  fqlpop::WmTrainer* t = nullptr;
  std::string err;
  if (fqlpop::wm_create_impl(w, params_host, term_params_host, &t, &err)) return fail(err);
  *out = t;
  return 0;
}
extern "C" int fqlpop_wm_destroy(void* h) {
  fqlpop::wm_destroy_impl((fqlpop::WmTrainer*)h);
  return 0;
}
extern "C" int fqlpop_wm_step(void* h, int32_t train, const float* obs, const float* act, const float* next_obs,
                              const float* rewards, const float* dropout_mask, void* stream) {
  WM(h);
  if (fqlpop::wm_step_impl(T, train, obs, act, next_obs, rewards, dropout_mask, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_wm_set_dataset(void* h, const float* obs, const float* act, const float* next_obs,
                                     const float* rewards, int64_t n_rows) {
  WM(h);
  if (fqlpop::wm_set_dataset_impl(T, obs, act, next_obs, rewards, n_rows, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_wm_train(void* h, int32_t n_steps, const float* dropout_mask, void* stream) {
  WM(h);
  if (fqlpop::wm_train_impl(T, n_steps, dropout_mask, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_wm_logs(void* h, float* out_host, void* stream) {
  WM(h);
  if (!out_host) return fail("null argument");
  if (fqlpop::wm_logs_impl(T, out_host, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_wm_export(void* h, int32_t which, float* out_host, int64_t n, void* stream) {
  WM(h);
  if (!out_host) return fail("null argument");
  if (fqlpop::wm_export_impl(T, which, out_host, n, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int64_t fqlpop_wm_launches(void* h) { return h ? fqlpop::wm_launches((fqlpop::WmTrainer*)h) : 0; }

// initial-observation generator (VAE decoder)
extern "C" int fqlpop_decoder_create(int32_t latent_dim, int32_t n_hidden, const int32_t* hidden, int32_t out_dim,
                                     const float* params_host, void** out) {
  if (!out || !hidden) return fail("null argument");
  int h[fqlpop::kWmMaxLayers] = {};
  if (n_hidden < 1 || n_hidden >= fqlpop::kWmMaxLayers) return fail("decoder: between 1 and 7 hidden layers");
  for (int i = 0; i < n_hidden; ++i) h[i] = hidden[i];
  fqlpop::WmDecoder* d = nullptr;
  std::string err;
  if (fqlpop::wm_decoder_create_impl(latent_dim, n_hidden, h, out_dim, params_host, &d, &err)) return fail(err);
  *out = d;
  return 0;
}
extern "C" int fqlpop_decoder_destroy(void* h) {
  fqlpop::wm_decoder_destroy_impl((fqlpop::WmDecoder*)h);
  return 0;
}
extern "C" int fqlpop_decoder_sample(void* h, int32_t n, const float* noise, uint64_t seed, float* out, void* stream) {
  if (!h) return fail("null decoder handle");
  std::string err;
  if (fqlpop::wm_decoder_sample_impl((fqlpop::WmDecoder*)h, n, noise, seed, out, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}

// initial-observation VAE training (wm_vae.cuh)
static fqlpop::VaeConfig vae_cfg(const fqlpop_vae_config_t* c) {
  fqlpop::VaeConfig v{};
  v.O = c->ob_dim; v.L = c->latent_dim; v.batch = c->batch;
  v.n_hidden = c->n_hidden;   // (range-checked by the callers below)
  for (int i = 0; i < v.n_hidden && i < fqlpop::kWmMaxLayers; ++i) v.hidden[i] = c->hidden[i];
  v.decay_steps = c->decay_steps; v.lr = c->lr; v.reconstruction_weight = c->reconstruction_weight; v.seed = c->seed;
  return v;
}
extern "C" int64_t fqlpop_vae_num_params(const fqlpop_vae_config_t* c) {
  if (!c || c->n_hidden < 1 || c->n_hidden >= fqlpop::kWmMaxLayers || c->ob_dim < 1 || c->latent_dim < 1) return -1;
  return fqlpop::vae_num_params(vae_cfg(c));
}
extern "C" int fqlpop_vae_create(const fqlpop_vae_config_t* c, const float* params_host, void** out) {
  if (!c || !out) return fail("null argument");
  if (c->n_hidden < 1 || c->n_hidden >= fqlpop::kWmMaxLayers) return fail("VAE: between 1 and 7 hidden layers");
  fqlpop::WmVae* t = nullptr;
  std::string err;
  if (fqlpop::vae_create_impl(vae_cfg(c), params_host, &t, &err)) return fail(err);
  *out = t;
  return 0;
}
extern "C" int fqlpop_vae_destroy(void* h) {
  fqlpop::vae_destroy_impl((fqlpop::WmVae*)h);
  return 0;
}
#define VAE(h)                                \
  fqlpop::WmVae* T = (fqlpop::WmVae*)(h);     \
  if (!T) return fail("null VAE handle");     \
  std::string err
extern "C" int fqlpop_vae_step(void* h, int32_t train, const float* obs, const float* eps, void* stream) {
  VAE(h);
  if (fqlpop::vae_step_impl(T, train, obs, eps, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_vae_set_dataset(void* h, const float* obs, int64_t n_rows) {
  VAE(h);
  if (fqlpop::vae_set_dataset_impl(T, obs, n_rows, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_vae_train(void* h, int32_t n_steps, void* stream) {
  VAE(h);
  if (fqlpop::vae_train_impl(T, n_steps, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_vae_logs(void* h, float* out_host, void* stream) {
  VAE(h);
  if (!out_host) return fail("null output");
  if (fqlpop::vae_logs_impl(T, out_host, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int fqlpop_vae_export(void* h, int32_t which, float* out_host, int64_t n, void* stream) {
  VAE(h);
  if (!out_host) return fail("null output");
  if (fqlpop::vae_export_impl(T, which, out_host, n, (cudaStream_t)stream, &err)) return fail(err);
  return 0;
}
extern "C" int64_t fqlpop_vae_launches(void* h) { return h ? fqlpop::vae_launches((fqlpop::WmVae*)h) : 0; }
