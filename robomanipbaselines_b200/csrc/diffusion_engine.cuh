// B200-native ConditionalUnet1D denoising engine + DP3 point-cloud encoder behind the C ABI of include/rmb_b200.h.
// Included at the end of act_engine.cu (one translation unit: shares fail(), RMB_CUDA, grid_for, the GEMM launchers).
//
// Replaces, for B trajectories at once, what the reference calls at
//   policy/diffusion_policy/RolloutDiffusionPolicy.py:74       policy.predict_action -> conditional_sample (DDPM 100)
//   policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:90  DP3.predict_action   -> conditional_sample (DDIM 10)
// The module arithmetic is restated in oracle/diffusion_oracle.py (sources not in the reference mount).
//
// Layout: activations [B, T, C] 16-bit channel-minor, so Conv1d(k) is an implicit GEMM with k taps shifted along T by
// the TMA box (out-of-range rows zero-filled = conv padding) and rows = (sample, t). Everything that depends only on
// the diffusion step (the timestep MLP and its share of every FiLM projection) is tabulated at load; the
// global_cond share of all 12 FiLM projections is ONE GEMM per inference.
#pragma once
#include "diffusion_kernels.cuh"
#include "gemm_gn.cuh"

namespace {

// gemm_tc_kernel with the fused GroupNorm + Mish + FiLM + residual epilogue (gemm_gn.cuh)
template <int BN, int ST, int EPI, bool PAIR>
cudaError_t launch_gemm_gn_instance(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  using S = GemmSmem<BN, 64, ST, PAIR>;
  static_assert(S::kTotal <= 232448, "shared memory");
  auto kern = PAIR ? gemm_tc_pair_kernel<BN, 64, ST, false, 0, EPI> : gemm_tc_kernel<BN, 64, ST, false, 0, EPI>;
  static bool attr_set = false;
  static int max_pairs = 0;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
    if (e != cudaSuccess) return e;
    max_pairs = PAIR ? max_active_pairs(kern, S::kTotal, num_sms) : 0;
    attr_set = true;
  }
  return launch_pdl(kern, dim3(gemm_grid(g, num_sms, max_pairs)), dim3(kGemmThreads), S::kTotal, stream, g.ta, g.tb, g.ta, g.p);
}

template <int EPI>
cudaError_t launch_gemm_gn_epi(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  if (g.pair) {
    if (g.block_n == 64) return launch_gemm_gn_instance<64, 8, EPI, true>(g, num_sms, stream);
    if (g.block_n == 128) return launch_gemm_gn_instance<128, 8, EPI, true>(g, num_sms, stream);
    if (g.block_n == 256) return launch_gemm_gn_instance<256, 6, EPI, true>(g, num_sms, stream);
    return cudaErrorInvalidValue;
  }
  if (g.block_n == 64) return launch_gemm_gn_instance<64, 6, EPI, false>(g, num_sms, stream);
  if (g.block_n == 128) return launch_gemm_gn_instance<128, 6, EPI, false>(g, num_sms, stream);
  if (g.block_n == 256) return launch_gemm_gn_instance<256, 4, EPI, false>(g, num_sms, stream);
  return cudaErrorInvalidValue;
}

cudaError_t launch_gemm_gn(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  return g.p.gn_film_gc != nullptr ? launch_gemm_gn_epi<1>(g, num_sms, stream) : launch_gemm_gn_epi<2>(g, num_sms, stream);
}

// can this conv plan normalise its own output? (whole samples per row tile, whole groups per column tile)
bool gn_fusable(const GemmPlan& g, int T, int C) {
  const int cg = C / 8;
  return g.p.tiles_w == 1 && g.p.tiles_h == 1 && g.p.box_w == T && g.p.box_h == 1 && T >= 4 && 128 % T == 0 && cg % 32 == 0 &&
         g.block_n % cg == 0 && g.p.bias != nullptr && g.block_k == 64;
}

// Conv1d-style implicit GEMM over a [B, W_view, c_extent] view (row stride / batch stride in elements).
bool plan_conv1d(GemmPlan* g, const void* in, int c_extent, long long row_stride, int W_view, long long batch_stride,
                 int B, int Cin, int ntaps, const int* tap_c0, const int* tap_d1, const void* Wt, int Cout,
                 const float* bias, void* out, int ld_out, bool out_f32, std::string* err) {
  *g = GemmPlan();
  g->block_k = 64;
  g->block_n = pick_block_n(Cout);
  if (g->block_n == 0 || Cin % 64 != 0) { *err = "plan_conv1d: Cout and Cin must be multiples of 64"; return false; }
  if (ntaps > kMaxTaps) { *err = "plan_conv1d: too many taps"; return false; }
  GemmParams& p = g->p;
  pick_box(W_view, 1, B, &p.box_w, &p.box_h, &p.box_b);
  p.W = W_view; p.H = 1; p.B = B;
  p.tiles_w = (W_view + p.box_w - 1) / p.box_w; p.tiles_h = 1; p.tiles_b = (B + p.box_b - 1) / p.box_b;
  p.n_col_tiles = Cout / g->block_n;
  p.num_taps = ntaps; p.chunks_per_tap = Cin / 64;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  for (int t = 0; t < ntaps; ++t) { p.tap_c0[t] = tap_c0[t]; p.tap_d1[t] = tap_d1[t]; }
  fill_common(p, bias, nullptr, 0, false, out, ld_out, out_f32, false);
  const uint64_t e = 2;
  const uint64_t dims[5] = {(uint64_t)c_extent, (uint64_t)W_view, 1, 1, (uint64_t)B};
  const uint64_t strides[4] = {(uint64_t)row_stride * e, (uint64_t)batch_stride * e, (uint64_t)batch_stride * e, (uint64_t)batch_stride * e};
  const uint32_t box[5] = {64, (uint32_t)p.box_w, 1, 1, (uint32_t)p.box_b};
  if (!make_tmap_act(&g->ta, in, 5, dims, strides, box, 128, err)) return false;
  const uint64_t K = (uint64_t)ntaps * Cin;
  const uint64_t wdims[2] = {K, (uint64_t)Cout};
  const uint64_t wstr[1] = {K * e};
  g->pair = gemm_use_pair(p);
  const uint32_t wbox[2] = {64, (uint32_t)(g->pair ? g->block_n / 2 : g->block_n)};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  g->flops = 2ll * B * W_view * Cout * (long long)K;
  return true;
}

float mish_host(float x) {
  if (x > 20.0f) return x;
  return x * tanhf(log1pf(expf(x)));
}

}  // namespace

// =====================================================================================================================
struct rmb_unet_engine {
  rmb_unet_config_t cfg;
  int num_sms = 148;
  bool finalized = false;
  std::unordered_map<std::string, std::vector<float>> host_w;
  std::vector<void*> allocs;

  int L = 0, d[4] = {}, Tl[4] = {};
  bool fuse_gn = true;                      // RMB_UNET_NO_FUSE=1 keeps GroupNorm in its own kernel (A/B measurements)
  struct Crb {
    std::string name;
    int cin = 0, cout = 0, T = 0, film_off = 0;
    bool has_res = false, first = false;
    act_t *w0 = nullptr, *w1 = nullptr, *wres = nullptr;
    float *b0 = nullptr, *b1 = nullptr, *bres = nullptr, *g0 = nullptr, *be0 = nullptr, *g1 = nullptr, *be1 = nullptr;
  };
  std::vector<Crb> crbs;                    // down (2 per level), mid (2), up (2 per level)
  struct Samp { act_t* w[2] = {nullptr, nullptr}; float* b = nullptr; };
  Samp down[3], up[3];
  act_t* w_fin = nullptr;
  float *b_fin = nullptr, *g_fin = nullptr, *be_fin = nullptr, *w_head = nullptr, *b_head = nullptr;
  int film_total = 0, Gp = 0;
  float* film_t = nullptr;                  // [num_train_timesteps, film_total]
  act_t* w_film = nullptr;                  // [film_total, Gp]
  float* b_film = nullptr;

  // workspace
  int xcol_w = 64;          // im2col width of the network input: kernel_size * input_dim rounded up to 64 or 128
  act_t *xcol = nullptr, *gcm = nullptr, *h1 = nullptr, *resb = nullptr, *xa = nullptr, *xb = nullptr;
  act_t* cat[4] = {};
  float *film_gc = nullptr, *conv_out = nullptr, *x_state = nullptr;

  struct Op {
    int kind = 0;                           // 0 GEMM, 1 GroupNorm(+Mish+FiLM+residual), 2 GEMM with that fused in
    GemmPlan g;
    GnParams gn;
    std::string name;
    double bytes = 0;
  };
  struct Plans {
    GemmPlan film;
    std::vector<Op> ops;
    FinalParams fin;
    double fin_bytes = 0;
  };
  std::map<int, std::unique_ptr<Plans>> plans;

  template <typename T>
  T* dalloc(size_t n) {
    void* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
    allocs.push_back(p);
    return static_cast<T*>(p);
  }
  const std::vector<float>* find(const std::string& name, size_t numel, std::string* err) {
    auto it = host_w.find(name);
    if (it == host_w.end()) { *err = "missing checkpoint tensor: " + name; return nullptr; }
    if (it->second.size() != numel) {
      *err = "checkpoint tensor " + name + " has " + std::to_string(it->second.size()) + " elements, expected " + std::to_string(numel);
      return nullptr;
    }
    return &it->second;
  }
  bool up_f32(const std::vector<float>& h, float** dptr, std::string* err) {
    *dptr = dalloc<float>(h.size());
    if (!*dptr || cudaMemcpy(*dptr, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
  bool up_act(const std::vector<float>& h, act_t** dptr, std::string* err) {
    std::vector<act_t> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = to_act(h[i]);
    *dptr = dalloc<act_t>(h.size());
    if (!*dptr || cudaMemcpy(*dptr, t.data(), t.size() * sizeof(act_t), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
  // Conv1d weight [Cout, Cin, k] -> [Cout, k*Cin] tap-major
  bool up_conv1d(const std::string& name, int cout, int cin, int k, act_t** w, float** b, std::string* err) {
    const auto* W = find(name + ".weight", (size_t)cout * cin * k, err); if (!W) return false;
    const auto* B = find(name + ".bias", cout, err); if (!B) return false;
    std::vector<float> r((size_t)cout * k * cin);
    for (int co = 0; co < cout; ++co)
      for (int ci = 0; ci < cin; ++ci)
        for (int kk = 0; kk < k; ++kk) r[((size_t)co * k + kk) * cin + ci] = (*W)[((size_t)co * cin + ci) * k + kk];
    return up_act(r, w, err) && up_f32(*B, b, err);
  }
};

extern "C" {

int rmb_unet_engine_create(const rmb_unet_config_t* cfg, rmb_unet_engine_t** out) {
  if (!cfg || !out) return fail("null argument");
  if (cfg->n_levels < 2 || cfg->n_levels > 4) return fail("rmb_unet: n_levels must be 2..4");
  if (cfg->n_groups != 8) return fail("rmb_unet: n_groups must be 8");
  if (cfg->kernel_size < 1 || cfg->kernel_size % 2 == 0 || cfg->kernel_size > kMaxTaps) return fail("rmb_unet: kernel_size must be odd and <= 9");
  if (cfg->input_dim < 1 || cfg->kernel_size * cfg->input_dim > kXcolMaxWidth) return fail("rmb_unet: kernel_size * input_dim must be <= 128 (action dim <= 25 at kernel size 5)");
  if (cfg->horizon < 1 || (cfg->horizon & (cfg->horizon - 1)) != 0 || (cfg->horizon >> (cfg->n_levels - 1)) < 1) return fail("rmb_unet: horizon must be a power of two >= 2^(n_levels-1)");
  if (cfg->max_batch < 1 || cfg->global_cond_dim < 1 || cfg->dsed < 2 || cfg->dsed % 2 != 0 || cfg->num_train_timesteps < 1) return fail("rmb_unet: bad config");
  for (int i = 0; i < cfg->n_levels; ++i) {
    if (cfg->down_dims[i] < 64 || cfg->down_dims[i] % 64 != 0) return fail("rmb_unet: down_dims must be multiples of 64");
    if ((long long)(cfg->horizon >> i) * cfg->down_dims[i] > 16384) return fail("rmb_unet: horizon_level * dim must be <= 16384 (GroupNorm register tile)");
  }
  RMB_CUDA(cudaSetDevice(cfg->device));
  if (int r = rmb_check_device(cfg->device)) return r;
  auto* e = new rmb_unet_engine();
  e->cfg = *cfg;
  e->xcol_w = (cfg->kernel_size * cfg->input_dim <= 64) ? 64 : 128;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  e->num_sms = prop.multiProcessorCount;
  e->L = cfg->n_levels;
  { const char* nf = getenv("RMB_UNET_NO_FUSE"); e->fuse_gn = !(nf && nf[0] == '1'); }
  for (int i = 0; i < e->L; ++i) { e->d[i] = cfg->down_dims[i]; e->Tl[i] = cfg->horizon >> i; }
  *out = e;
  return 0;
}

void rmb_unet_engine_destroy(rmb_unet_engine_t* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : e->allocs) cudaFree(p);
  delete e;
}

int rmb_unet_load_tensor(rmb_unet_engine_t* e, const char* name, const float* host_data, int64_t numel) {
  if (!e || !name || (!host_data && numel > 0) || numel < 0) return fail("rmb_unet_load_tensor: bad argument");
  if (e->finalized) return fail("rmb_unet_load_tensor: engine already finalized");
  e->host_w[name] = std::vector<float>(host_data, host_data + numel);
  return 0;
}

int rmb_unet_finalize(rmb_unet_engine_t* e) {
  if (!e) return fail("null engine");
  if (e->finalized) return fail("already finalized");
  RMB_CUDA(cudaSetDevice(e->cfg.device));
  std::string err;
  const rmb_unet_config_t& c = e->cfg;
  const int L = e->L, K = c.kernel_size, Da = c.input_dim, dsed = c.dsed, G = c.global_cond_dim, cond_dim = dsed + G;
  const int pad = K / 2;

  // ---- block list in execution order
  auto add_crb = [&](const std::string& name, int cin, int cout, int T) {
    rmb_unet_engine::Crb b;
    b.name = name; b.cin = cin; b.cout = cout; b.T = T; b.has_res = cin != cout;
    b.film_off = e->film_total;
    e->film_total += 2 * cout;
    e->crbs.push_back(b);
  };
  for (int i = 0; i < L; ++i) {
    add_crb("down_modules." + std::to_string(i) + ".0", i == 0 ? Da : e->d[i - 1], e->d[i], e->Tl[i]);
    add_crb("down_modules." + std::to_string(i) + ".1", e->d[i], e->d[i], e->Tl[i]);
  }
  e->crbs[0].first = true;
  add_crb("mid_modules.0", e->d[L - 1], e->d[L - 1], e->Tl[L - 1]);
  add_crb("mid_modules.1", e->d[L - 1], e->d[L - 1], e->Tl[L - 1]);
  for (int j = 0; j + 1 < L; ++j) {
    const int l = L - 1 - j;
    add_crb("up_modules." + std::to_string(j) + ".0", 2 * e->d[l], e->d[l - 1], e->Tl[l]);
    add_crb("up_modules." + std::to_string(j) + ".1", e->d[l - 1], e->d[l - 1], e->Tl[l]);
  }

  // ---- timestep MLP table: m[t] = Mish(diffusion_step_encoder(t)), the first dsed inputs of every cond_encoder
  std::vector<float> mt((size_t)c.num_train_timesteps * dsed);
  {
    const auto* w1 = e->find("diffusion_step_encoder.1.weight", (size_t)4 * dsed * dsed, &err); if (!w1) return fail(err);
    const auto* b1 = e->find("diffusion_step_encoder.1.bias", 4 * dsed, &err); if (!b1) return fail(err);
    const auto* w2 = e->find("diffusion_step_encoder.3.weight", (size_t)4 * dsed * dsed, &err); if (!w2) return fail(err);
    const auto* b2 = e->find("diffusion_step_encoder.3.bias", dsed, &err); if (!b2) return fail(err);
    const int half = dsed / 2;
    const float kf = logf(10000.0f) / (float)(half - 1);
    std::vector<float> emb(dsed), hid(4 * dsed);
    for (int t = 0; t < c.num_train_timesteps; ++t) {
      for (int i = 0; i < half; ++i) {
        const float a = (float)t * expf((float)i * -kf);
        emb[i] = sinf(a); emb[half + i] = cosf(a);
      }
      for (int o = 0; o < 4 * dsed; ++o) {
        double acc = (*b1)[o];
        for (int k = 0; k < dsed; ++k) acc += (double)(*w1)[(size_t)o * dsed + k] * emb[k];
        hid[o] = mish_host((float)acc);
      }
      for (int o = 0; o < dsed; ++o) {
        double acc = (*b2)[o];
        for (int k = 0; k < 4 * dsed; ++k) acc += (double)(*w2)[(size_t)o * 4 * dsed + k] * hid[k];
        mt[(size_t)t * dsed + o] = mish_host((float)acc);
      }
    }
  }

  // ---- per-block weights + FiLM tables
  e->Gp = (G + 63) / 64 * 64;
  std::vector<float> film_t((size_t)c.num_train_timesteps * e->film_total);
  std::vector<float> w_film((size_t)e->film_total * e->Gp, 0.f), b_film(e->film_total);
  for (auto& b : e->crbs) {
    const std::string p = b.name + ".";
    if (b.first) {
      // Conv1d(Da -> cout, K) and the 1x1 residual conv both read the im2col row [tap*Da + a] (64 columns)
      const auto* W = e->find(p + "blocks.0.block.0.weight", (size_t)b.cout * Da * K, &err); if (!W) return fail(err);
      const auto* B0 = e->find(p + "blocks.0.block.0.bias", b.cout, &err); if (!B0) return fail(err);
      std::vector<float> r((size_t)b.cout * e->xcol_w, 0.f);
      for (int co = 0; co < b.cout; ++co)
        for (int a = 0; a < Da; ++a)
          for (int kk = 0; kk < K; ++kk) r[(size_t)co * e->xcol_w + kk * Da + a] = (*W)[((size_t)co * Da + a) * K + kk];
      if (!e->up_act(r, &b.w0, &err) || !e->up_f32(*B0, &b.b0, &err)) return fail(err);
      const auto* WR = e->find(p + "residual_conv.weight", (size_t)b.cout * Da, &err); if (!WR) return fail(err);
      const auto* BR = e->find(p + "residual_conv.bias", b.cout, &err); if (!BR) return fail(err);
      std::vector<float> rr((size_t)b.cout * e->xcol_w, 0.f);
      for (int co = 0; co < b.cout; ++co)
        for (int a = 0; a < Da; ++a) rr[(size_t)co * e->xcol_w + pad * Da + a] = (*WR)[(size_t)co * Da + a];
      if (!e->up_act(rr, &b.wres, &err) || !e->up_f32(*BR, &b.bres, &err)) return fail(err);
    } else {
      if (!e->up_conv1d(p + "blocks.0.block.0", b.cout, b.cin, K, &b.w0, &b.b0, &err)) return fail(err);
      if (b.has_res && !e->up_conv1d(p + "residual_conv", b.cout, b.cin, 1, &b.wres, &b.bres, &err)) return fail(err);
    }
    if (!e->up_conv1d(p + "blocks.1.block.0", b.cout, b.cout, K, &b.w1, &b.b1, &err)) return fail(err);
    const auto* g0 = e->find(p + "blocks.0.block.1.weight", b.cout, &err); if (!g0) return fail(err);
    const auto* be0 = e->find(p + "blocks.0.block.1.bias", b.cout, &err); if (!be0) return fail(err);
    const auto* g1 = e->find(p + "blocks.1.block.1.weight", b.cout, &err); if (!g1) return fail(err);
    const auto* be1 = e->find(p + "blocks.1.block.1.bias", b.cout, &err); if (!be1) return fail(err);
    if (!e->up_f32(*g0, &b.g0, &err) || !e->up_f32(*be0, &b.be0, &err) || !e->up_f32(*g1, &b.g1, &err) || !e->up_f32(*be1, &b.be1, &err)) return fail(err);
    const auto* wc = e->find(p + "cond_encoder.1.weight", (size_t)2 * b.cout * cond_dim, &err); if (!wc) return fail(err);
    const auto* bc = e->find(p + "cond_encoder.1.bias", 2 * b.cout, &err); if (!bc) return fail(err);
    for (int o = 0; o < 2 * b.cout; ++o) {
      const float* wr = wc->data() + (size_t)o * cond_dim;
      for (int k = 0; k < G; ++k) w_film[(size_t)(b.film_off + o) * e->Gp + k] = wr[dsed + k];
      b_film[b.film_off + o] = (*bc)[o];
      for (int t = 0; t < c.num_train_timesteps; ++t) {
        const float* m = mt.data() + (size_t)t * dsed;
        double acc = 0;
        for (int k = 0; k < dsed; ++k) acc += (double)wr[k] * m[k];
        film_t[(size_t)t * e->film_total + b.film_off + o] = (float)acc;
      }
    }
  }
  if (!e->up_f32(film_t, &e->film_t, &err) || !e->up_act(w_film, &e->w_film, &err) || !e->up_f32(b_film, &e->b_film, &err)) return fail(err);

  // ---- Downsample1d Conv1d(C, C, 3, 2, 1) / Upsample1d ConvTranspose1d(C, C, 4, 2, 1)
  for (int i = 0; i + 1 < L; ++i)
    if (!e->up_conv1d("down_modules." + std::to_string(i) + ".2.conv", e->d[i], e->d[i], 3, &e->down[i].w[0], &e->down[i].b, &err)) return fail(err);
  for (int j = 0; j + 1 < L; ++j) {
    const int C = e->d[L - 2 - j];
    const std::string n = "up_modules." + std::to_string(j) + ".2.conv";
    const auto* W = e->find(n + ".weight", (size_t)C * C * 4, &err); if (!W) return fail(err);
    const auto* B = e->find(n + ".bias", C, &err); if (!B) return fail(err);
    // out[2m] = x[m-1] w3 + x[m] w1 ; out[2m+1] = x[m] w2 + x[m+1] w0   (weight layout [Cin, Cout, 4])
    const int kk_of[2][2] = {{3, 1}, {2, 0}};
    for (int ph = 0; ph < 2; ++ph) {
      std::vector<float> r((size_t)C * 2 * C);
      for (int co = 0; co < C; ++co)
        for (int tp = 0; tp < 2; ++tp)
          for (int ci = 0; ci < C; ++ci) r[((size_t)co * 2 + tp) * C + ci] = (*W)[((size_t)ci * C + co) * 4 + kk_of[ph][tp]];
      if (!e->up_act(r, &e->up[j].w[ph], &err)) return fail(err);
    }
    if (!e->up_f32(*B, &e->up[j].b, &err)) return fail(err);
  }
  // ---- final_conv
  {
    const int C = e->d[0];
    if (!e->up_conv1d("final_conv.0.block.0", C, C, K, &e->w_fin, &e->b_fin, &err)) return fail(err);
    const auto* g = e->find("final_conv.0.block.1.weight", C, &err); if (!g) return fail(err);
    const auto* be = e->find("final_conv.0.block.1.bias", C, &err); if (!be) return fail(err);
    const auto* wh = e->find("final_conv.1.weight", (size_t)Da * C, &err); if (!wh) return fail(err);
    const auto* bh = e->find("final_conv.1.bias", Da, &err); if (!bh) return fail(err);
    if (!e->up_f32(*g, &e->g_fin, &err) || !e->up_f32(*be, &e->be_fin, &err) || !e->up_f32(*wh, &e->w_head, &err) || !e->up_f32(*bh, &e->b_head, &err)) return fail(err);
  }

  // ---- workspace
  {
    size_t maxTC = 0;
    for (int l = 0; l < L; ++l) {
      maxTC = std::max(maxTC, (size_t)e->Tl[l] * e->d[l]);
      if (l > 0) maxTC = std::max(maxTC, (size_t)e->Tl[l] * e->d[l - 1]);
    }
    const size_t B = c.max_batch;
    bool ok = true;
    ok &= (e->xcol = e->dalloc<act_t>(B * c.horizon * e->xcol_w)) != nullptr;
    ok &= (e->gcm = e->dalloc<act_t>(B * e->Gp)) != nullptr;
    ok &= (e->film_gc = e->dalloc<float>(B * e->film_total)) != nullptr;
    ok &= (e->conv_out = e->dalloc<float>(B * maxTC)) != nullptr;
    ok &= (e->h1 = e->dalloc<act_t>(B * maxTC)) != nullptr;
    ok &= (e->resb = e->dalloc<act_t>(B * maxTC)) != nullptr;
    ok &= (e->xa = e->dalloc<act_t>(B * maxTC)) != nullptr;
    ok &= (e->xb = e->dalloc<act_t>(B * maxTC)) != nullptr;
    ok &= (e->x_state = e->dalloc<float>(B * c.horizon * Da)) != nullptr;
    for (int l = 1; l < L; ++l) ok &= (e->cat[l] = e->dalloc<act_t>(B * e->Tl[l] * 2 * e->d[l])) != nullptr;
    if (!ok) return fail("cudaMalloc failed (unet workspace)");
  }
  e->host_w.clear();
  e->finalized = true;
  RMB_CUDA(cudaDeviceSynchronize());
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// Launch plan of one UNet evaluation for `nb` trajectories.
static int unet_build_plans(rmb_unet_engine* e, int nb, rmb_unet_engine::Plans** out) {
  auto it = e->plans.find(nb);
  if (it != e->plans.end()) { *out = it->second.get(); return 0; }
  std::unique_ptr<rmb_unet_engine::Plans> pl(new rmb_unet_engine::Plans());
  std::string err;
  const rmb_unet_config_t& c = e->cfg;
  const int L = e->L, K = c.kernel_size, pad = K / 2, T0 = c.horizon;
  using Op = rmb_unet_engine::Op;

  int taps_c0[kMaxTaps] = {}, taps_d1[kMaxTaps] = {};
  for (int t = 0; t < K; ++t) taps_d1[t] = t - pad;
  const int zero1[1] = {0};

  // stride-1 Conv1d over [nb, T, Cin] at `in` (row stride ld_in) -> out (fp32 conv_out or act)
  auto conv_s1 = [&](const std::string& name, const act_t* in, int ld_in, int Cin, int T, int ntaps, const act_t* W,
                     int Cout, const float* bias, void* outp, int ld_out, bool out_f32) -> bool {
    Op op; op.kind = 0; op.name = name;
    if (!plan_conv1d(&op.g, in, Cin, ld_in, T, (long long)T * ld_in, nb, Cin, ntaps, ntaps == 1 ? zero1 : taps_c0,
                     ntaps == 1 ? zero1 : taps_d1, W, Cout, bias, outp, ld_out, out_f32, &err)) return false;
    op.bytes = (double)nb * T * Cin * 2 + (double)Cout * ntaps * Cin * 2 + (double)nb * T * Cout * (out_f32 ? 4 : 2);
    pl->ops.push_back(op);
    return true;
  };
  auto gn = [&](const std::string& name, int C, int T, const float* gamma, const float* beta, int film_off,
                const act_t* res, int ld_res, act_t* outp, int ld_out) {
    Op op; op.kind = 1; op.name = name;
    GnParams& g = op.gn;
    g.in = e->conv_out; g.in_f32 = 1; g.ld_in = C; g.gamma = gamma; g.beta = beta; g.C = C; g.T = T;
    g.film_gc = film_off >= 0 ? e->film_gc : nullptr; g.film_t = e->film_t; g.film_ld = e->film_total;
    g.film_off = film_off >= 0 ? film_off : 0;
    g.res = res; g.ld_res = ld_res; g.out = outp; g.ld_out = ld_out;
    op.bytes = (double)nb * T * C * (4 + 2 + (res ? 2 : 0)) + (film_off >= 0 ? (double)nb * 2 * C * 4 : 0);
    pl->ops.push_back(op);
  };
  // conv + GroupNorm + Mish (+ FiLM) (+ residual): one launch when the tile holds whole (sample, group)s, else two
  auto conv_gn = [&](const std::string& name, const act_t* in, int ld_in, int Cin, int T, int ntaps, const act_t* W,
                     int Cout, const float* bias, const float* gamma, const float* beta, int film_off, const act_t* res,
                     int ld_res, act_t* outp, int ld_out) -> bool {
    if (!conv_s1(name, in, ld_in, Cin, T, ntaps, W, Cout, bias, e->conv_out, Cout, true)) return false;
    Op& op = pl->ops.back();
    if (e->fuse_gn && gn_fusable(op.g, T, Cout)) {
      GemmParams& p = op.g.p;
      op.kind = 2;
      op.name += film_off >= 0 ? "+gn+mish+film" : (res ? "+gn+mish+res" : "+gn+mish");
      p.out = outp; p.ld_out = ld_out; p.out_f32 = 0;
      p.gn_gamma = gamma; p.gn_beta = beta; p.gn_cg = Cout / 8;
      p.gn_film_gc = film_off >= 0 ? e->film_gc : nullptr; p.gn_film_t = e->film_t; p.gn_film_ld = e->film_total;
      p.gn_film_off = film_off >= 0 ? film_off : 0;
      p.gn_res = res; p.gn_ld_res = ld_res;
      op.bytes = (double)nb * T * Cin * 2 + (double)Cout * ntaps * Cin * 2 + (double)nb * T * Cout * (2 + (res ? 2 : 0)) +
                 (film_off >= 0 ? (double)nb * 2 * Cout * 4 : 0);
      return true;
    }
    gn(name + " gn+mish", Cout, T, gamma, beta, film_off, res, ld_res, outp, ld_out);
    return true;
  };
  // ConditionalResidualBlock1D: x_in [nb, T, cin] (ld_x) -> x_out (ld_o)
  auto crb = [&](const rmb_unet_engine::Crb& b, const act_t* x_in, int ld_x, act_t* x_out, int ld_o) -> bool {
    const act_t* res = x_in; int ld_res = ld_x;
    const act_t* in0 = b.first ? e->xcol : x_in;
    const int ld0 = b.first ? e->xcol_w : ld_x, cin0 = b.first ? e->xcol_w : b.cin, taps0 = b.first ? 1 : K;
    if (!conv_gn(b.name + " conv0", in0, ld0, cin0, b.T, taps0, b.w0, b.cout, b.b0, b.g0, b.be0, b.film_off, nullptr, 0, e->h1, b.cout)) return false;
    if (b.has_res) {
      if (!conv_s1(b.name + " res1x1", in0, ld0, cin0, b.T, 1, b.wres, b.cout, b.bres, e->resb, b.cout, false)) return false;
      res = e->resb; ld_res = b.cout;
    }
    return conv_gn(b.name + " conv1", e->h1, b.cout, b.cout, b.T, K, b.w1, b.cout, b.b1, b.g1, b.be1, -1, res, ld_res, x_out, ld_o);
  };

  // once per inference: FiLM(global_cond) for all blocks
  if (!plan_linear(&pl->film, e->gcm, nb, e->Gp, e->Gp, e->w_film, e->film_total, e->b_film, nullptr, 0, false,
                   e->film_gc, e->film_total, true, false, &err)) return fail(err);

  int bi = 0;
  // ---- down path
  const act_t* cur = nullptr; int ld_cur = 0;
  for (int i = 0; i < L; ++i) {
    const int C = e->d[i], T = e->Tl[i];
    if (!crb(e->crbs[bi++], cur, ld_cur, e->xb, C)) return fail(err);
    act_t* skip = (i == 0) ? e->xa : e->cat[i] + C;
    const int ld_skip = (i == 0) ? C : 2 * C;
    if (!crb(e->crbs[bi++], e->xb, C, skip, ld_skip)) return fail(err);
    cur = skip; ld_cur = ld_skip;
    if (i + 1 < L) {
      // Downsample1d: Conv1d(C, C, 3, stride 2, pad 1) on the phase view [nb, T/2, 2*ld]
      Op op; op.kind = 0; op.name = "down_modules." + std::to_string(i) + ".2 conv3/2";
      const int tc0[3] = {ld_skip, 0, ld_skip}, td1[3] = {-1, 0, 0};
      if (!plan_conv1d(&op.g, skip, ld_skip + C, 2ll * ld_skip, T / 2, (long long)T * ld_skip, nb, C, 3, tc0, td1,
                       e->down[i].w[0], C, e->down[i].b, e->xb, C, false, &err)) return fail(err);
      op.bytes = (double)nb * T * C * 2 + 3.0 * C * C * 2 + (double)nb * (T / 2) * C * 2;
      pl->ops.push_back(op);
      cur = e->xb; ld_cur = C;
      // the next level's first block reads xb and writes xb: safe, a block writes x_out only in its last kernel, and
      // by then x_in is either dead (residual conv already in resb) or read element-aligned by the writing thread
    }
  }
  // ---- mid
  {
    const int C = e->d[L - 1];
    if (!crb(e->crbs[bi++], cur, ld_cur, e->xa == cur ? e->xb : e->xa, C)) return fail(err);
    const act_t* m0 = (e->xa == cur) ? e->xb : e->xa;
    if (!crb(e->crbs[bi++], m0, C, e->cat[L - 1], 2 * C)) return fail(err);
  }
  // ---- up path
  for (int j = 0; j + 1 < L; ++j) {
    const int l = L - 1 - j, Cin2 = 2 * e->d[l], Co = e->d[l - 1], T = e->Tl[l];
    if (!crb(e->crbs[bi++], e->cat[l], Cin2, e->xa, Co)) return fail(err);
    if (!crb(e->crbs[bi++], e->xa, Co, e->xb, Co)) return fail(err);
    // Upsample1d: ConvTranspose1d(Co, Co, 4, 2, 1) = two 2-tap GEMMs, one per output phase
    act_t* dst = (l - 1 >= 1) ? e->cat[l - 1] : e->xa;
    const int ld_dst = (l - 1 >= 1) ? 2 * Co : Co;
    for (int ph = 0; ph < 2; ++ph) {
      Op op; op.kind = 0; op.name = "up_modules." + std::to_string(j) + ".2 convT phase" + std::to_string(ph);
      const int tc0[2] = {0, 0};
      const int td1[2] = {ph == 0 ? -1 : 0, ph == 0 ? 0 : 1};
      if (!plan_conv1d(&op.g, e->xb, Co, Co, T, (long long)T * Co, nb, Co, 2, tc0, td1, e->up[j].w[ph], Co, e->up[j].b,
                       dst + ph * ld_dst, 2 * ld_dst, false, &err)) return fail(err);
      op.bytes = (double)nb * T * Co * 2 + 2.0 * Co * Co * 2 + (double)nb * T * Co * 2;
      pl->ops.push_back(op);
    }
  }
  // ---- final_conv.0 conv (its GroupNorm + Mish + the 1x1 head + scheduler step are unet_final_step_kernel)
  {
    const int C = e->d[0];
    if (!conv_s1("final_conv.0 conv", e->xa, C, C, T0, K, e->w_fin, C, e->b_fin, e->conv_out, C, true)) return fail(err);
    FinalParams& f = pl->fin;
    memset(&f, 0, sizeof(f));
    f.gn.in = e->conv_out; f.gn.in_f32 = 1; f.gn.ld_in = C; f.gn.gamma = e->g_fin; f.gn.beta = e->be_fin; f.gn.C = C; f.gn.T = T0;
    f.w_head = e->w_head; f.b_head = e->b_head; f.Da = c.input_dim; f.ksize = K; f.xcol = e->xcol; f.xcol_w = e->xcol_w;
    pl->fin_bytes = (double)nb * T0 * C * 4 + (double)nb * T0 * (c.input_dim * 12 + e->xcol_w * 2);
  }
  *out = pl.get();
  e->plans[nb] = std::move(pl);
  return 0;
}

static int unet_run_eval(rmb_unet_engine* e, rmb_unet_engine::Plans* pl, int nb, int t, const FinalParams& fin,
                         cudaStream_t s, LaunchProf* prof) {
  auto begin = [&](const std::string& name, const char* kernel, double flops, double bytes) {
    if (!prof) return;
    rmb_launch_prof_t r;
    memset(&r, 0, sizeof(r));
    snprintf(r.name, sizeof(r.name), "%s", name.c_str());
    snprintf(r.kernel, sizeof(r.kernel), "%s", kernel);
    r.flops = flops; r.bytes = bytes;
    prof->rec.push_back(r);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    prof->ev.push_back(a); prof->ev.push_back(b);
    cudaEventRecord(a, s);
  };
  auto end = [&]() { if (prof) cudaEventRecord(prof->ev.back(), s); };
  const float* film_t_row = e->film_t + (size_t)t * e->film_total;
  for (auto& op : pl->ops) {
    if (op.kind == 0) {
      begin(op.name, op.g.pair ? "gemm_tc_pair_kernel" : "gemm_tc_kernel", (double)op.g.flops, op.bytes);
      RMB_CUDA(launch_gemm(op.g, e->num_sms, s));
      end();
    } else if (op.kind == 2) {
      GemmPlan g = op.g;
      g.p.gn_film_t = film_t_row;
      begin(op.name, g.pair ? "gemm_tc_pair_kernel<gn>" : "gemm_tc_kernel<gn>", (double)op.g.flops, op.bytes);
      RMB_CUDA(launch_gemm_gn(g, e->num_sms, s));
      end();
    } else {
      GnParams g = op.gn;
      g.film_t = film_t_row;
      begin(op.name, "unet_gn_mish_kernel", 0, op.bytes);
      unet_gn_mish_kernel<<<nb, 256, 0, s>>>(g);
      end();
    }
  }
  const int smem = (fin.gn.T * fin.gn.C + 3 * fin.gn.T * fin.Da) * (int)sizeof(float);
  static bool attr_set = false;
  if (!attr_set) {
    RMB_CUDA(cudaFuncSetAttribute(unet_final_step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    attr_set = true;
  }
  begin("final gn+mish+head+sched step", "unet_final_step_kernel", 2.0 * nb * fin.gn.T * fin.gn.C * fin.Da, pl->fin_bytes);
  unet_final_step_kernel<<<nb, 256, smem, s>>>(fin);
  end();
  RMB_CUDA(cudaGetLastError());
  return 0;
}

// global_cond -> Mish -> 16-bit -> FiLM GEMM (all blocks at once)
static int unet_prepare_cond(rmb_unet_engine* e, rmb_unet_engine::Plans* pl, const float* d_gc, int nb, cudaStream_t s) {
  mish_cast_kernel<<<grid_for((long long)nb * e->Gp, 256, e->num_sms), 256, 0, s>>>(d_gc, e->gcm, nb, e->cfg.global_cond_dim, e->Gp);
  RMB_CUDA(launch_gemm(pl->film, e->num_sms, s));
  return 0;
}

static int unet_check(rmb_unet_engine* e, int batch, const char* who) {
  if (!e) return fail(std::string(who) + ": null engine");
  if (!e->finalized) return fail(std::string(who) + ": weights not finalized");
  if (batch < 1 || batch > e->cfg.max_batch) return fail(std::string(who) + ": batch out of range");
  return 0;
}

extern "C" {

int rmb_unet_launches_per_step(const rmb_unet_engine_t* e) {
  if (!e) return -1;
  if (!e->plans.empty()) return (int)e->plans.begin()->second->ops.size() + 1;   // as planned (+ the final kernel)
  int n = 1 + 1;                                  // final conv + final kernel
  for (auto& b : e->crbs) n += (e->fuse_gn ? 2 : 4) + (b.has_res ? 1 : 0);
  n += (e->L - 1) * 3;                            // down conv + 2 convT phases per level
  return n;
}

double rmb_unet_step_flops(const rmb_unet_engine_t* e, int batch) {
  if (!e) return -1.0;
  const int K = e->cfg.kernel_size, Da = e->cfg.input_dim;
  double f = 0;
  for (auto& b : e->crbs) {
    f += 2.0 * b.T * b.cout * K * b.cin + 2.0 * b.T * b.cout * K * b.cout;
    if (b.has_res) f += 2.0 * b.T * b.cout * b.cin;
    f += 2.0 * 2 * b.cout * (e->cfg.dsed + e->cfg.global_cond_dim);      // FiLM projection (reference: every step)
  }
  for (int i = 0; i + 1 < e->L; ++i) {
    f += 2.0 * e->Tl[i + 1] * e->d[i] * 3 * e->d[i];                       // stride-2 conv
    f += 2.0 * e->Tl[i] * e->d[i] * 2 * e->d[i];                           // ConvTranspose1d: 4 taps / 2 per output
  }
  f += 2.0 * e->Tl[0] * e->d[0] * K * e->d[0] + 2.0 * e->Tl[0] * Da * e->d[0];
  return f * batch;
}

int rmb_scheduler_step(const float* d_x, const float* d_model_out, const float* d_noise, const rmb_sched_step_t* step,
                       float* d_out, int64_t n, void* stream_) {
  if (!d_x || !d_model_out || !step || !d_out || n < 0) return fail("rmb_scheduler_step: null argument");
  if (step->kind != 0 && step->kind != 1) return fail("rmb_scheduler_step: scheduler kind must be 0 (DDPM) or 1 (DDIM)");
  if (step->kind == 0 && step->sigma > 0.f && !d_noise) return fail("rmb_scheduler_step: sigma > 0 needs d_noise");
  if (n == 0) return 0;
  SchedStep st;
  st.kind = step->kind; st.pred_sample = step->pred_sample; st.clip = step->clip; st.derive_eps = step->derive_eps;
  st.s_beta = step->s_beta; st.s_alpha = step->s_alpha; st.c_x0 = step->c_x0; st.c_xt = step->c_xt; st.sigma = step->sigma;
  sched_step_kernel<<<grid_for((n + 3) / 4, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(d_x, d_model_out, d_noise, d_out, n, st);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_unet1d_denoise(rmb_unet_engine_t* e, const float* d_global_cond, const float* d_x_init, const float* d_noise,
                       const rmb_sched_step_t* steps, int n_steps, float* d_x_out, int batch, void* stream_) {
  if (int r = unet_check(e, batch, "rmb_unet1d_denoise")) return r;
  if (!d_global_cond || !d_x_init || !steps || !d_x_out || n_steps < 1) return fail("rmb_unet1d_denoise: null argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const rmb_unet_config_t& c = e->cfg;
  for (int i = 0; i < n_steps; ++i) {
    if (steps[i].t < 0 || steps[i].t >= c.num_train_timesteps) return fail("rmb_unet1d_denoise: timestep out of range");
    if (steps[i].kind != 0 && steps[i].kind != 1) return fail("rmb_unet1d_denoise: scheduler kind must be 0 (DDPM) or 1 (DDIM)");
    if (steps[i].kind == 0 && steps[i].sigma > 0.f && !d_noise) return fail("rmb_unet1d_denoise: DDPM steps with sigma > 0 need d_noise");
  }
  rmb_unet_engine::Plans* pl = nullptr;
  if (int r = unet_build_plans(e, batch, &pl)) return r;
  if (int r = unet_prepare_cond(e, pl, d_global_cond, batch, s)) return r;
  unet_init_kernel<<<batch, 256, 0, s>>>(d_x_init, d_x_out, e->xcol, c.horizon, c.input_dim, c.kernel_size, e->xcol_w);
  const size_t per_step = (size_t)batch * c.horizon * c.input_dim;
  for (int i = 0; i < n_steps; ++i) {
    FinalParams fin = pl->fin;
    fin.x = d_x_out;
    fin.noise = d_noise ? d_noise + (size_t)i * per_step : nullptr;
    fin.st.kind = steps[i].kind; fin.st.pred_sample = steps[i].pred_sample; fin.st.clip = steps[i].clip;
    fin.st.derive_eps = steps[i].derive_eps; fin.st.s_beta = steps[i].s_beta; fin.st.s_alpha = steps[i].s_alpha;
    fin.st.c_x0 = steps[i].c_x0; fin.st.c_xt = steps[i].c_xt; fin.st.sigma = steps[i].sigma;
    if (int r = unet_run_eval(e, pl, batch, steps[i].t, fin, s, nullptr)) return r;
  }
  return 0;
}

static int unet_forward_impl(rmb_unet_engine_t* e, const float* d_sample, int timestep, const float* d_global_cond,
                             float* d_out, int batch, cudaStream_t s, LaunchProf* prof) {
  if (int r = unet_check(e, batch, "rmb_unet1d_forward")) return r;
  if (!d_sample || !d_global_cond || !d_out) return fail("rmb_unet1d_forward: null argument");
  if (timestep < 0 || timestep >= e->cfg.num_train_timesteps) return fail("rmb_unet1d_forward: timestep out of range");
  const rmb_unet_config_t& c = e->cfg;
  rmb_unet_engine::Plans* pl = nullptr;
  if (int r = unet_build_plans(e, batch, &pl)) return r;
  if (int r = unet_prepare_cond(e, pl, d_global_cond, batch, s)) return r;
  unet_init_kernel<<<batch, 256, 0, s>>>(d_sample, e->x_state, e->xcol, c.horizon, c.input_dim, c.kernel_size, e->xcol_w);
  FinalParams fin = pl->fin;
  fin.x = e->x_state; fin.noise = nullptr; fin.model_out = d_out;
  fin.st.kind = 2;
  return unet_run_eval(e, pl, batch, timestep, fin, s, prof);
}

int rmb_unet1d_forward(rmb_unet_engine_t* e, const float* d_sample, int timestep, const float* d_global_cond,
                       float* d_out, int batch, void* stream_) {
  return unet_forward_impl(e, d_sample, timestep, d_global_cond, d_out, batch, static_cast<cudaStream_t>(stream_), nullptr);
}

int rmb_unet1d_forward_profile(rmb_unet_engine_t* e, const float* d_sample, int timestep, const float* d_global_cond,
                               float* d_out, int batch, void* stream_, rmb_launch_prof_t* out, int cap) {
  if (!out || cap < 1) return fail("rmb_unet1d_forward_profile: null output");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  LaunchProf prof;
  const int r = unet_forward_impl(e, d_sample, timestep, d_global_cond, d_out, batch, s, &prof);
  cudaError_t ce = cudaStreamSynchronize(s);
  int n = 0;
  if (r == 0 && ce == cudaSuccess) {
    for (size_t i = 0; i < prof.rec.size() && n < cap; ++i, ++n) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, prof.ev[2 * i], prof.ev[2 * i + 1]);
      prof.rec[i].ms = ms;
      out[n] = prof.rec[i];
    }
  }
  for (cudaEvent_t ev : prof.ev) cudaEventDestroy(ev);
  if (r != 0) return r;
  if (ce != cudaSuccess) return fail(std::string("rmb_unet1d_forward_profile: ") + cudaGetErrorString(ce), -2);
  return n;
}

}  // extern "C"

// =====================================================================================================================
// DP3Encoder: PointNetEncoderXYZ (Linear 3-64-128-256 with LayerNorm + ReLU, max over points, Linear 256-Co, LayerNorm)
// + state MLP. Layer 1 (K = 3) is an elementwise kernel; layers 2 and 3 are tcgen05 GEMMs over all points.
struct rmb_pointnet_engine {
  rmb_pointnet_config_t cfg;
  int num_sms = 148;
  bool finalized = false;
  std::unordered_map<std::string, std::vector<float>> host_w;
  std::vector<void*> allocs;
  float *w1 = nullptr, *b1 = nullptr, *g1 = nullptr, *be1 = nullptr;
  act_t *w2 = nullptr, *w3 = nullptr;
  float *b2 = nullptr, *g2 = nullptr, *be2 = nullptr, *b3 = nullptr, *g3 = nullptr, *be3 = nullptr;
  float *wp = nullptr, *bp = nullptr, *gp = nullptr, *bep = nullptr;
  float *ws1 = nullptr, *bs1 = nullptr, *ws2 = nullptr, *bs2 = nullptr;
  act_t *h1 = nullptr, *h2 = nullptr;
  float *y = nullptr, *feat = nullptr;
  struct Plans { GemmPlan l2, l3; };
  std::map<int, std::unique_ptr<Plans>> plans;
  template <typename T>
  T* dalloc(size_t n) {
    void* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
    allocs.push_back(p);
    return static_cast<T*>(p);
  }
  bool up(const std::string& name, size_t numel, float** d, std::string* err) {
    auto it = host_w.find(name);
    if (it == host_w.end()) { *err = "missing checkpoint tensor: " + name; return false; }
    if (it->second.size() != numel) { *err = "checkpoint tensor " + name + " has the wrong size"; return false; }
    *d = dalloc<float>(numel);
    if (!*d || cudaMemcpy(*d, it->second.data(), numel * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
  bool up_act(const std::string& name, size_t numel, act_t** d, std::string* err) {
    auto it = host_w.find(name);
    if (it == host_w.end()) { *err = "missing checkpoint tensor: " + name; return false; }
    if (it->second.size() != numel) { *err = "checkpoint tensor " + name + " has the wrong size"; return false; }
    std::vector<act_t> t(numel);
    for (size_t i = 0; i < numel; ++i) t[i] = to_act(it->second[i]);
    *d = dalloc<act_t>(numel);
    if (!*d || cudaMemcpy(*d, t.data(), numel * sizeof(act_t), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
};

extern "C" {

int rmb_pointnet_engine_create(const rmb_pointnet_config_t* cfg, rmb_pointnet_engine_t** out) {
  if (!cfg || !out) return fail("null argument");
  if (cfg->max_clouds < 1 || cfg->num_points < 1) return fail("rmb_pointnet: bad sizes");
  if (cfg->in_channels < 1 || cfg->in_channels > 6) return fail("rmb_pointnet: in_channels must be 1..6");
  if (cfg->out_channels < 1 || cfg->out_channels > 64) return fail("rmb_pointnet: out_channels must be <= 64");
  if (cfg->state_dim < 0) return fail("rmb_pointnet: bad state_dim");
  RMB_CUDA(cudaSetDevice(cfg->device));
  if (int r = rmb_check_device(cfg->device)) return r;
  auto* e = new rmb_pointnet_engine();
  e->cfg = *cfg;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  e->num_sms = prop.multiProcessorCount;
  *out = e;
  return 0;
}

void rmb_pointnet_engine_destroy(rmb_pointnet_engine_t* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : e->allocs) cudaFree(p);
  delete e;
}

int rmb_pointnet_load_tensor(rmb_pointnet_engine_t* e, const char* name, const float* host_data, int64_t numel) {
  if (!e || !name || (!host_data && numel > 0) || numel < 0) return fail("rmb_pointnet_load_tensor: bad argument");
  if (e->finalized) return fail("rmb_pointnet_load_tensor: engine already finalized");
  e->host_w[name] = std::vector<float>(host_data, host_data + numel);
  return 0;
}

int rmb_pointnet_finalize(rmb_pointnet_engine_t* e) {
  if (!e) return fail("null engine");
  if (e->finalized) return fail("already finalized");
  RMB_CUDA(cudaSetDevice(e->cfg.device));
  std::string err;
  const rmb_pointnet_config_t& c = e->cfg;
  const size_t ci = c.in_channels, co = c.out_channels;
  bool ok = e->up("extractor.mlp.0.weight", 64 * ci, &e->w1, &err) && e->up("extractor.mlp.0.bias", 64, &e->b1, &err) &&
            e->up("extractor.mlp.1.weight", 64, &e->g1, &err) && e->up("extractor.mlp.1.bias", 64, &e->be1, &err) &&
            e->up_act("extractor.mlp.3.weight", 128 * 64, &e->w2, &err) && e->up("extractor.mlp.3.bias", 128, &e->b2, &err) &&
            e->up("extractor.mlp.4.weight", 128, &e->g2, &err) && e->up("extractor.mlp.4.bias", 128, &e->be2, &err) &&
            e->up_act("extractor.mlp.6.weight", 256 * 128, &e->w3, &err) && e->up("extractor.mlp.6.bias", 256, &e->b3, &err) &&
            e->up("extractor.mlp.7.weight", 256, &e->g3, &err) && e->up("extractor.mlp.7.bias", 256, &e->be3, &err) &&
            e->up("extractor.final_projection.0.weight", co * 256, &e->wp, &err) && e->up("extractor.final_projection.0.bias", co, &e->bp, &err) &&
            e->up("extractor.final_projection.1.weight", co, &e->gp, &err) && e->up("extractor.final_projection.1.bias", co, &e->bep, &err);
  if (ok && c.state_dim > 0)
    ok = e->up("state_mlp.0.weight", 64 * (size_t)c.state_dim, &e->ws1, &err) && e->up("state_mlp.0.bias", 64, &e->bs1, &err) &&
         e->up("state_mlp.2.weight", 64 * 64, &e->ws2, &err) && e->up("state_mlp.2.bias", 64, &e->bs2, &err);
  if (!ok) return fail(err);
  const size_t rows = (size_t)c.max_clouds * c.num_points;
  ok &= (e->h1 = e->dalloc<act_t>(rows * 64)) != nullptr;
  ok &= (e->h2 = e->dalloc<act_t>(rows * 128)) != nullptr;
  ok &= (e->y = e->dalloc<float>(rows * 256)) != nullptr;
  ok &= (e->feat = e->dalloc<float>((size_t)c.max_clouds * 256)) != nullptr;
  if (!ok) return fail("cudaMalloc failed (pointnet workspace)");
  e->host_w.clear();
  e->finalized = true;
  RMB_CUDA(cudaDeviceSynchronize());
  return 0;
}

int rmb_pointnet_encode(rmb_pointnet_engine_t* e, const float* d_points, int point_stride, const float* d_state,
                        float* d_out, int n_clouds, void* stream_) {
  if (!e || !d_points || !d_out) return fail("rmb_pointnet_encode: null argument");
  if (!e->finalized) return fail("rmb_pointnet_encode: weights not finalized");
  const rmb_pointnet_config_t& c = e->cfg;
  if (n_clouds < 1 || n_clouds > c.max_clouds) return fail("rmb_pointnet_encode: n_clouds out of range");
  if (point_stride < c.in_channels) return fail("rmb_pointnet_encode: point_stride < in_channels");
  if (c.state_dim > 0 && !d_state) return fail("rmb_pointnet_encode: state required");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const long long rows = (long long)n_clouds * c.num_points;
  auto it = e->plans.find(n_clouds);
  if (it == e->plans.end()) {
    std::unique_ptr<rmb_pointnet_engine::Plans> pl(new rmb_pointnet_engine::Plans());
    std::string err;
    if (!plan_linear(&pl->l2, e->h1, rows, 64, 64, e->w2, 128, e->b2, nullptr, 0, false, e->y, 128, true, false, &err)) return fail(err);
    if (!plan_linear(&pl->l3, e->h2, rows, 128, 128, e->w3, 256, e->b3, nullptr, 0, false, e->y, 256, true, false, &err)) return fail(err);
    it = e->plans.emplace(n_clouds, std::move(pl)).first;
  }
  const int rows_per_cta = 64;
  const int chunks = (c.num_points + rows_per_cta - 1) / rows_per_cta;
  pointnet_layer1_kernel<<<grid_for(rows * 8, 256, e->num_sms), 256, 0, s>>>(d_points, point_stride, c.in_channels, e->w1, e->b1, e->g1, e->be1, e->h1, rows);
  RMB_CUDA(launch_gemm(it->second->l2, e->num_sms, s));
  pointnet_ln_relu_kernel<128><<<n_clouds * chunks, 256, 0, s>>>(e->y, e->g2, e->be2, e->h2, nullptr, c.num_points, rows_per_cta);
  RMB_CUDA(launch_gemm(it->second->l3, e->num_sms, s));
  RMB_CUDA(cudaMemsetAsync(e->feat, 0, (size_t)n_clouds * 256 * sizeof(float), s));
  pointnet_ln_relu_kernel<256><<<n_clouds * chunks, 256, 0, s>>>(e->y, e->g3, e->be3, nullptr, e->feat, c.num_points, rows_per_cta);
  const int ld_out = c.out_channels + (c.state_dim > 0 ? 64 : 0);
  pointnet_final_kernel<<<n_clouds, 128, 0, s>>>(e->feat, e->wp, e->bp, e->gp, e->bep, c.out_channels,
                                                 c.state_dim > 0 ? d_state : nullptr, c.state_dim, e->ws1, e->bs1, e->ws2, e->bs2, d_out, ld_out);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
