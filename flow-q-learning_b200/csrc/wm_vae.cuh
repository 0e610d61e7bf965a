// Training of the initial-observation VAE (included by wm.cu, inside namespace fqlpop).
//
// Reference: envmodel/initial_observation.py:9-52 (Encoder: [Dense, relu]* then two heads mean / logvar; reparameterize
// z = mu + eps * exp(logvar / 2); Decoder: [Dense, relu]* over hidden_dims[::-1], Dense(O)) and vae_loss :55-78:
//   reconstruction = mean((obs - recon)^2), kl = -0.5 * mean(1 + logvar - mu^2 - exp(logvar)),
//   loss = (reconstruction_weight * reconstruction + kl) / (reconstruction_weight + 1).
// The reference ships the model and its loss but no trainer for it (train_env_model.py does not know the model): the
// optimiser here is the one its other world-model trainers use, optax.adam over a cosine schedule
// (state_predictor_trainer.py:57-61), i.e. wm_adam_kernel.
//
// Layout: ONE flat parameter vector = encoder (hidden layers, then a single head layer of 2 L columns whose first L are
// the mean head and whose last L the logvar head -- one product instead of two) followed by the decoder, so that Adam is
// one launch; the host splits the head into the two flax leaves. The step is the same fp32 SIMT kernels as the other
// world-model trainers (Launcher::gemm / colsum), captured in a CUDA graph; the only VAE-specific kernels are the three
// elementwise ones below.
#pragma once

struct WmVae {
  int O = 0, L = 0, B = 0, device = 0, decay_steps = 1;
  float lr = 1e-3f, rw = 1.f;
  unsigned long long seed = 0;
  WmNet enc, dec;
  long long n_params = 0;
  float *params = nullptr, *m = nullptr, *v = nullptr, *grads = nullptr;
  float* xe[kWmMaxLayers + 1] = {};   // encoder activations: xe[0] aliases the batch, xe[n] = (mu | logvar)
  float* xd[kWmMaxLayers + 1] = {};   // decoder activations: xd[0] = z, xd[n] = reconstruction
  float* ge[kWmMaxLayers + 1] = {};   // gradients wrt xe[l] (ge[n] = wrt the head outputs)
  float* gd[kWmMaxLayers + 1] = {};   // gradients wrt xd[l] (gd[n] = wrt the reconstruction, gd[0] = wrt z)
  float *eps = nullptr, *b_obs = nullptr, *logs = nullptr;
  int* step = nullptr;
  const float* ds_obs = nullptr;
  long long ds_rows = 0;
  struct Entry { int mode; const void *obs, *eps; cudaGraphExec_t exec; int launches; };
  std::vector<Entry> graphs;
  long long launches = 0;
  std::vector<void*> owned;
};

namespace {

// resident-dataset sampler: uniform rows with replacement, Philox(seed, step, row, 0) -- the stream of wm_gather_kernel
__global__ void vae_gather_kernel(const float* __restrict__ obs, long long n_rows, int O, int batch, unsigned long long seed,
                                  const int* __restrict__ step, float* __restrict__ b_obs) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= batch) return;
  const long long idx = philox_index(seed, (unsigned long long)*step, (uint32_t)row, 0u, (uint32_t)n_rows);
  for (int j = lane; j < O; j += 32) b_obs[(long long)row * O + j] = obs[idx * O + j];
}

// reparameterize (initial_observation.py:41-44). eps_in == nullptr: eps = Philox normal (seed, step, row, 1 + j).
__global__ void vae_latent_kernel(const float* __restrict__ head, const float* __restrict__ eps_in, int B, int L,
                                  unsigned long long seed, const int* __restrict__ step, float* __restrict__ eps_out,
                                  float* __restrict__ z) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * L) return;
  const int b = i / L, j = i - b * L;
  const float mu = head[(long long)b * 2 * L + j], lv = head[(long long)b * 2 * L + L + j];
  const float e = eps_in ? eps_in[i] : philox_normal(seed, (unsigned long long)*step, (uint32_t)b, (uint32_t)(1 + j));
  eps_out[i] = e;
  z[i] = mu + e * expf(0.5f * lv);
}

// vae_loss (initial_observation.py:55-78) + d loss / d reconstruction. One CTA, fixed summation order.
__global__ void __launch_bounds__(256) vae_loss_kernel(const float* __restrict__ obs, const float* __restrict__ recon,
                                                       const float* __restrict__ head, int B, int O, int L, float rw,
                                                       float* __restrict__ grecon, float* __restrict__ logs,
                                                       int* __restrict__ step) {
  __shared__ float sm[8][2];
  float v[2] = {0.f, 0.f};
  const float gs = rw / (rw + 1.f) * 2.f / ((float)B * O);
  for (int i = threadIdx.x; i < B * O; i += 256) {
    const float d = recon[i] - obs[i];
    v[0] += d * d;
    if (grecon) grecon[i] = gs * d;
  }
  for (int i = threadIdx.x; i < B * L; i += 256) {
    const int b = i / L, j = i - b * L;
    const float mu = head[(long long)b * 2 * L + j], lv = head[(long long)b * 2 * L + L + j];
    v[1] += 1.f + lv - mu * mu - expf(lv);
  }
  wm_block_reduce<2>(v, sm);
  if (threadIdx.x == 0) {
    const float rl = v[0] / ((float)B * O), kl = -0.5f * v[1] / ((float)B * L);
    logs[WL_RECON_LOSS] = rl;
    logs[WL_KL] = kl;
    logs[WL_LOSS] = (rw * rl + kl) / (rw + 1.f);
    wm_tick(step, logs);
  }
}

// d loss / d (mu | logvar): through z = mu + eps * exp(logvar / 2) and the KL term
__global__ void vae_latent_bwd_kernel(const float* __restrict__ dz, const float* __restrict__ head,
                                      const float* __restrict__ eps, int B, int L, float rw, float* __restrict__ ghead) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * L) return;
  const int b = i / L, j = i - b * L;
  const float mu = head[(long long)b * 2 * L + j], lv = head[(long long)b * 2 * L + L + j];
  const float c = 1.f / ((rw + 1.f) * (float)B * L);
  ghead[(long long)b * 2 * L + j] = dz[i] + c * mu;
  ghead[(long long)b * 2 * L + L + j] = dz[i] * eps[i] * 0.5f * expf(0.5f * lv) + c * 0.5f * (expf(lv) - 1.f);
}

// backward pass of a Dense / relu stack: weight gradients and bias sums branch off to the side streams, the input-gradient
// chain stays on the main one (as wm_issue does). g = gradient wrt x[n_layers]; gx[l] receives the gradient wrt x[l].
const float* vae_backward(Launcher& L, const WmNet& net, const float* params, float* grads, float* const* x,
                          const float* g, float* const* gx, int B, bool input_grad) {
  for (int l = net.n_layers - 1; l >= 0; --l) {
    const int din = net.dims[l], dout = net.dims[l + 1];
    L.fork();
    L.cur = L.w;
    L.gemm(true, false, x[l], din, g, dout, grads + net.offW[l], dout, din, dout, B, WE_NONE, nullptr, nullptr, 0, 0);
    L.cur = L.c;
    L.colsum(g, nullptr, dout, B, dout, grads + net.offB[l], nullptr);
    L.cur = L.s;
    if (l > 0 || input_grad) {
      L.gemm(false, true, g, dout, params + net.offW[l], dout, gx[l], din, B, din, dout, l > 0 ? WE_MASK : WE_NONE, nullptr,
             l > 0 ? x[l] : nullptr, din, 0);
      g = gx[l];
    }
  }
  return g;
}

// mode: bit 0 train, bit 1 gather the batch from the resident dataset
void vae_issue(WmVae* t, Launcher& L, int mode, const float* obs, const float* eps_in) {
  const int B = t->B, O = t->O, Ld = t->L, ne = t->enc.n_layers, nd = t->dec.n_layers;
  const bool train = mode & 1;
  cudaStream_t s = L.s;
  if (mode & 2) {
    vae_gather_kernel<<<(B * 32 + 255) / 256, 256, 0, s>>>(t->ds_obs, t->ds_rows, O, B, t->seed, t->step, t->b_obs);
    ++L.n;
    obs = t->b_obs;
  }
  float* xe[kWmMaxLayers + 1];
  for (int l = 0; l <= ne; ++l) xe[l] = t->xe[l];
  xe[0] = const_cast<float*>(obs);   // read only
  const float* pe = t->params;
  const float* pd = t->params + t->enc.n_params;
  wm_forward(L, t->enc, pe, xe, B, nullptr, 0);
  vae_latent_kernel<<<(B * Ld + 255) / 256, 256, 0, s>>>(xe[ne], eps_in, B, Ld, t->seed, t->step, t->eps, t->xd[0]);
  ++L.n;
  wm_forward(L, t->dec, pd, t->xd, B, nullptr, 0);
  vae_loss_kernel<<<1, 256, 0, s>>>(obs, t->xd[nd], xe[ne], B, O, Ld, t->rw, train ? t->gd[nd] : nullptr, t->logs,
                                    train ? t->step : nullptr);
  ++L.n;
  if (!train) return;
  const float* dz = vae_backward(L, t->dec, pd, t->grads + t->enc.n_params, t->xd, t->gd[nd], t->gd, B, true);
  vae_latent_bwd_kernel<<<(B * Ld + 255) / 256, 256, 0, s>>>(dz, xe[ne], t->eps, B, Ld, t->rw, t->ge[ne]);
  ++L.n;
  vae_backward(L, t->enc, pe, t->grads, xe, t->ge[ne], t->ge, B, false);
  L.join();
  wm_adam_kernel<<<(int)((t->n_params + 255) / 256), 256, 0, s>>>(t->params, t->m, t->v, t->grads, t->n_params, t->step,
                                                                 t->lr, t->decay_steps, t->logs);
  ++L.n;
}

int vae_alloc(WmVae* t, float** p, long long n, std::string* err) {
  WM_CUDA(cudaMalloc((void**)p, sizeof(float) * (size_t)(n > 0 ? n : 1)));
  WM_CUDA(cudaMemset(*p, 0, sizeof(float) * (size_t)(n > 0 ? n : 1)));
  t->owned.push_back(*p);
  return 0;
}

int vae_run(WmVae* t, int mode, int repeats, const float* obs, const float* eps, cudaStream_t stream, std::string* err) {
  WM_CUDA(cudaSetDevice(t->device));
  WmVae::Entry* e = nullptr;
  for (auto& g : t->graphs)
    if (g.mode == mode && g.obs == obs && g.eps == eps) e = &g;
  if (!e) {
    if (t->graphs.size() >= 8) {
      cudaGraphExecDestroy(t->graphs.front().exec);
      t->graphs.erase(t->graphs.begin());
    }
    cudaStream_t cs, sw, sc;
    WM_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    WM_CUDA(cudaStreamCreateWithFlags(&sw, cudaStreamNonBlocking));
    WM_CUDA(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking));
    WM_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    Launcher L{cs, sw, sc, cs};
    vae_issue(t, L, mode, obs, eps);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(cs, &graph), le = cudaGetLastError();
    for (cudaEvent_t ev : L.evs) cudaEventDestroy(ev);
    cudaStreamDestroy(sw);
    cudaStreamDestroy(sc);
    cudaStreamDestroy(cs);
    if (ce != cudaSuccess || le != cudaSuccess) {
      if (err) *err = std::string("VAE step capture: ") + cudaGetErrorString(ce != cudaSuccess ? ce : le);
      return 1;
    }
    cudaGraphExec_t exec;
    WM_CUDA(cudaGraphInstantiate(&exec, graph, 0));
    cudaGraphDestroy(graph);
    t->graphs.push_back({mode, obs, eps, exec, L.n});
    e = &t->graphs.back();
  }
  for (int i = 0; i < repeats; ++i) WM_CUDA(cudaGraphLaunch(e->exec, stream));
  t->launches += (long long)repeats * e->launches;
  return 0;
}

}  // namespace

void vae_layouts(const VaeConfig& c, WmNet* enc, WmNet* dec) {
  int rev[kWmMaxLayers];
  for (int i = 0; i < c.n_hidden; ++i) rev[i] = c.hidden[c.n_hidden - 1 - i];
  *enc = wm_layout(c.O, c.n_hidden, c.hidden, 2 * c.L, false);
  *dec = wm_layout(c.L, c.n_hidden, rev, c.O, false);
}

long long vae_num_params(const VaeConfig& c) {
  WmNet e, d;
  vae_layouts(c, &e, &d);
  return e.n_params + d.n_params;
}

void vae_destroy_impl(WmVae* t) {
  if (!t) return;
  cudaSetDevice(t->device);
  cudaDeviceSynchronize();
  for (auto& g : t->graphs) cudaGraphExecDestroy(g.exec);
  for (void* p : t->owned) cudaFree(p);
  delete t;
}

int vae_create_impl(const VaeConfig& c, const float* params_host, WmVae** out, std::string* err) {
  auto bad = [&](const char* m) { if (err) *err = m; return 1; };
  if (c.O < 1 || c.L < 1 || c.batch < 1) return bad("VAE: observation / latent dimension and batch must be positive");
  if (c.n_hidden < 1 || c.n_hidden >= kWmMaxLayers) return bad("VAE: between 1 and 7 hidden layers");
  if (c.decay_steps < 1) return bad("VAE: decay_steps must be positive");
  if (!(c.reconstruction_weight >= 0.f)) return bad("VAE: reconstruction_weight must be >= 0");
  if (!params_host) return bad("VAE: initial parameters are required");
  WmVae* t = new WmVae();
  t->O = c.O; t->L = c.L; t->B = c.batch; t->decay_steps = c.decay_steps; t->lr = c.lr; t->rw = c.reconstruction_weight;
  t->seed = c.seed;
  cudaGetDevice(&t->device);
  vae_layouts(c, &t->enc, &t->dec);
  t->n_params = t->enc.n_params + t->dec.n_params;
  const long long n = t->n_params, B = c.batch;
  int rc = vae_alloc(t, &t->params, n, err) || vae_alloc(t, &t->m, n, err) || vae_alloc(t, &t->v, n, err) ||
           vae_alloc(t, &t->grads, n, err);
  for (int l = 1; l <= t->enc.n_layers && !rc; ++l)
    rc = vae_alloc(t, &t->xe[l], B * t->enc.dims[l], err) || vae_alloc(t, &t->ge[l], B * t->enc.dims[l], err);
  for (int l = 0; l <= t->dec.n_layers && !rc; ++l)
    rc = vae_alloc(t, &t->xd[l], B * t->dec.dims[l], err) || vae_alloc(t, &t->gd[l], B * t->dec.dims[l], err);
  rc = rc || vae_alloc(t, &t->eps, B * c.L, err) || vae_alloc(t, &t->b_obs, B * c.O, err) ||
       vae_alloc(t, &t->logs, kWmLogs, err) || vae_alloc(t, (float**)&t->step, 2, err);
  if (!rc && cudaMemcpy(t->params, params_host, sizeof(float) * n, cudaMemcpyHostToDevice) != cudaSuccess)
    rc = bad("VAE: copying the initial parameters failed");
  if (rc) {
    vae_destroy_impl(t);
    return 1;
  }
  *out = t;
  return 0;
}

int vae_step_impl(WmVae* t, int train, const float* obs, const float* eps, cudaStream_t stream, std::string* err) {
  if (!obs) {
    if (err) *err = "VAE step: null batch pointer";
    return 1;
  }
  return vae_run(t, train ? 1 : 0, 1, obs, eps, stream, err);
}

int vae_set_dataset_impl(WmVae* t, const float* obs, long long n_rows, std::string* err) {
  if (!obs || n_rows < 1 || n_rows > 0xffffffffLL) {
    if (err) *err = "VAE dataset: null pointer or row count outside [1, 2^32)";
    return 1;
  }
  t->ds_obs = obs;
  t->ds_rows = n_rows;
  for (auto& g : t->graphs) cudaGraphExecDestroy(g.exec);   // captured gathers hold the old pointer
  t->graphs.clear();
  return 0;
}

int vae_train_impl(WmVae* t, int n_steps, cudaStream_t stream, std::string* err) {
  if (!t->ds_rows) {
    if (err) *err = "VAE train: no resident dataset (fqlpop_vae_set_dataset)";
    return 1;
  }
  return vae_run(t, 3, n_steps < 0 ? 0 : n_steps, t->ds_obs, nullptr, stream, err);
}

int vae_logs_impl(WmVae* t, float* out_host, cudaStream_t stream, std::string* err) {
  WM_CUDA(cudaSetDevice(t->device));
  WM_CUDA(cudaMemcpyAsync(out_host, t->logs, sizeof(float) * kWmLogs, cudaMemcpyDeviceToHost, stream));
  WM_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

int vae_export_impl(WmVae* t, int which, float* out_host, long long n, cudaStream_t stream, std::string* err) {
  if (which < 0 || which > 3 || n != t->n_params) {
    if (err) *err = "VAE export: which must be 0..3 (params, Adam mu, Adam nu, grads) and n the parameter count";
    return 1;
  }
  const float* src = which == 0 ? t->params : which == 1 ? t->m : which == 2 ? t->v : t->grads;
  WM_CUDA(cudaSetDevice(t->device));
  WM_CUDA(cudaMemcpyAsync(out_host, src, sizeof(float) * n, cudaMemcpyDeviceToHost, stream));
  WM_CUDA(cudaStreamSynchronize(stream));
  return 0;
}

long long vae_launches(const WmVae* t) { return t->launches; }
