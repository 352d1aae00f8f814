// Diffusion Policy image observation encoder (one camera key) behind the C ABI of include/rmb_b200.h.
// Included at the end of act_engine.cu (same translation unit as the GEMM / halo-conv launchers).
//
// robomimic 0.2.0 VisualCore as diffusion_policy's DiffusionUnetHybridImagePolicy builds it (obs_encoder_group_norm,
// eval_fixed_crop; arguments at policy/diffusion_policy/TrainDiffusionPolicy.py:114-126):
//   centre crop -> ResNet18Conv (torchvision resnet18 children[:-2], every BatchNorm2d replaced by
//   GroupNorm(C/16, C)) -> SpatialSoftmax(num_kp 32) -> Flatten -> Linear(64, 64).
// Source not in the reference mount; restated in oracle/diffusion_oracle.py (DpImageEncoder), parity unpinned.
//
// Maps are NHWC act_t. The cropped image is zero-extended to multiples of 32 (216 x 288 -> 224 x 288) so that every
// stride-2 stage halves exactly; the extension behaves as the convs' own zero padding because gn2d_apply_kernel keeps
// everything outside the valid extent at zero, and (post-ReLU, values >= 0) as the max-pool's -inf padding.
#pragma once
#include "diffusion_kernels.cuh"

struct rmb_imgenc_engine {
  rmb_imgenc_config_t cfg;
  int num_sms = 148;
  bool finalized = false;
  std::unordered_map<std::string, std::vector<float>> host_w;
  std::vector<void*> allocs;
  int Hp = 0, Wp = 0;              // padded crop
  int Hv[6] = {}, Wv[6] = {};      // valid extents: [0] crop, [1] stem out, [2..5] layer1..4
  int Ha[6] = {}, Wa[6] = {};      // allocated extents

  struct ConvG { act_t* w = nullptr; float *g = nullptr, *b = nullptr; };   // conv weight + the GroupNorm after it
  ConvG stem, blk[4][2][2], down[4];
  act_t* w_kp = nullptr;           // [64, 512] (rows >= num_kp zero)
  float *b_kp = nullptr, *w_lin = nullptr, *b_lin = nullptr;

  act_t *x0 = nullptr, *c1 = nullptr, *c1n = nullptr;
  act_t* st[4][4] = {};            // per level: raw, t, y, z
  float2* partial = nullptr;
  float* logits = nullptr;

  struct Plans {
    HaloPlan stem, l1[2][2];
    GemmPlan conv[4][2][2], down[4], kp;
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
    if (it->second.size() != numel) { *err = "checkpoint tensor " + name + " has the wrong size"; return nullptr; }
    return &it->second;
  }
  bool up_f32(const std::vector<float>& h, float** d, std::string* err) {
    *d = dalloc<float>(h.size());
    if (!*d || cudaMemcpy(*d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
  bool up_act(const std::vector<float>& h, act_t** d, std::string* err) {
    std::vector<act_t> t(h.size());
    for (size_t i = 0; i < h.size(); ++i) t[i] = to_act(h[i]);
    *d = dalloc<act_t>(h.size());
    if (!*d || cudaMemcpy(*d, t.data(), t.size() * sizeof(act_t), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "weight upload failed"; return false; }
    return true;
  }
  // Conv2d weight [Cout, Cin, k, k] (no bias) -> [Cout, k*k*Cin] tap-major + GroupNorm gamma / beta
  bool up_conv(const std::string& conv, const std::string& gn, int cin, int cout, int k, ConvG* out, std::string* err) {
    const auto* w = find(conv + ".weight", (size_t)cout * cin * k * k, err); if (!w) return false;
    const auto* g = find(gn + ".weight", cout, err); if (!g) return false;
    const auto* b = find(gn + ".bias", cout, err); if (!b) return false;
    std::vector<float> r((size_t)cout * k * k * cin);
    for (int co = 0; co < cout; ++co)
      for (int ci = 0; ci < cin; ++ci)
        for (int kh = 0; kh < k; ++kh)
          for (int kw = 0; kw < k; ++kw)
            r[(((size_t)co * k + kh) * k + kw) * cin + ci] = (*w)[(((size_t)co * cin + ci) * k + kh) * k + kw];
    return up_act(r, &out->w, err) && up_f32(*g, &out->g, err) && up_f32(*b, &out->b, err);
  }
};

static const int kImgencChans[4] = {64, 128, 256, 512};
static const int kImgencMaxChunks = 64;

extern "C" {

int rmb_imgenc_engine_create(const rmb_imgenc_config_t* cfg, rmb_imgenc_engine_t** out) {
  if (!cfg || !out) return fail("null argument");
  if (cfg->max_images < 1) return fail("rmb_imgenc: bad max_images");
  if (cfg->crop_h < 32 || cfg->crop_w < 32 || cfg->crop_h > cfg->in_h || cfg->crop_w > cfg->in_w) return fail("rmb_imgenc: crop must be >= 32 and fit the image");
  if (cfg->num_kp < 1 || cfg->num_kp > 32 || cfg->feat_dim < 1 || cfg->feat_dim > 64) return fail("rmb_imgenc: num_kp <= 32 and feat_dim <= 64");
  RMB_CUDA(cudaSetDevice(cfg->device));
  if (int r = rmb_check_device(cfg->device)) return r;
  auto* e = new rmb_imgenc_engine();
  e->cfg = *cfg;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  e->num_sms = prop.multiProcessorCount;
  e->Hp = (cfg->crop_h + 31) / 32 * 32; e->Wp = (cfg->crop_w + 31) / 32 * 32;
  e->Hv[0] = cfg->crop_h; e->Wv[0] = cfg->crop_w; e->Ha[0] = e->Hp; e->Wa[0] = e->Wp;
  e->Hv[1] = (e->Hv[0] + 1) / 2; e->Wv[1] = (e->Wv[0] + 1) / 2; e->Ha[1] = e->Hp / 2; e->Wa[1] = e->Wp / 2;
  e->Hv[2] = (e->Hv[1] + 1) / 2; e->Wv[2] = (e->Wv[1] + 1) / 2; e->Ha[2] = e->Hp / 4; e->Wa[2] = e->Wp / 4;   // max-pool
  for (int l = 3; l < 6; ++l) {
    e->Hv[l] = (e->Hv[l - 1] + 1) / 2; e->Wv[l] = (e->Wv[l - 1] + 1) / 2;
    e->Ha[l] = e->Ha[l - 1] / 2; e->Wa[l] = e->Wa[l - 1] / 2;
  }
  *out = e;
  return 0;
}

void rmb_imgenc_engine_destroy(rmb_imgenc_engine_t* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : e->allocs) cudaFree(p);
  delete e;
}

int rmb_imgenc_load_tensor(rmb_imgenc_engine_t* e, const char* name, const float* host_data, int64_t numel) {
  if (!e || !name || (!host_data && numel > 0) || numel < 0) return fail("rmb_imgenc_load_tensor: bad argument");
  if (e->finalized) return fail("rmb_imgenc_load_tensor: engine already finalized");
  e->host_w[name] = std::vector<float>(host_data, host_data + numel);
  return 0;
}

int rmb_imgenc_finalize(rmb_imgenc_engine_t* e) {
  if (!e) return fail("null engine");
  if (e->finalized) return fail("already finalized");
  RMB_CUDA(cudaSetDevice(e->cfg.device));
  std::string err;
  const rmb_imgenc_config_t& c = e->cfg;
  const std::string bb = "backbone.nets.";
  {
    // stem [64,3,7,7] -> [64][kh][slot 0..7][4] (slot j >= 1 <-> kw = j-1; slot 0 and channel 3 zero), see plan_conv1_halo
    const auto* w = e->find(bb + "0.weight", 64 * 3 * 49, &err); if (!w) return fail(err);
    const auto* g = e->find(bb + "1.weight", 64, &err); if (!g) return fail(err);
    const auto* b = e->find(bb + "1.bias", 64, &err); if (!b) return fail(err);
    std::vector<float> r(64 * 7 * 32, 0.f);
    for (int co = 0; co < 64; ++co)
      for (int ci = 0; ci < 3; ++ci)
        for (int kh = 0; kh < 7; ++kh)
          for (int kw = 0; kw < 7; ++kw) r[((co * 7 + kh) * 8 + (kw + 1)) * 4 + ci] = (*w)[((co * 3 + ci) * 7 + kh) * 7 + kw];
    if (!e->up_act(r, &e->stem.w, &err) || !e->up_f32(*g, &e->stem.g, &err) || !e->up_f32(*b, &e->stem.b, &err)) return fail(err);
  }
  for (int li = 0; li < 4; ++li)
    for (int bi = 0; bi < 2; ++bi) {
      const std::string p = bb + std::to_string(4 + li) + "." + std::to_string(bi) + ".";
      const int cin = (bi == 0 && li > 0) ? kImgencChans[li - 1] : kImgencChans[li], co = kImgencChans[li];
      if (!e->up_conv(p + "conv1", p + "bn1", cin, co, 3, &e->blk[li][bi][0], &err)) return fail(err);
      if (!e->up_conv(p + "conv2", p + "bn2", co, co, 3, &e->blk[li][bi][1], &err)) return fail(err);
      if (bi == 0 && li > 0 && !e->up_conv(p + "downsample.0", p + "downsample.1", cin, co, 1, &e->down[li], &err)) return fail(err);
    }
  {
    const int K = c.num_kp, F = c.feat_dim;
    const auto* w = e->find("pool.nets.weight", (size_t)K * 512, &err); if (!w) return fail(err);
    const auto* b = e->find("pool.nets.bias", K, &err); if (!b) return fail(err);
    const auto* wl = e->find("nets.3.weight", (size_t)F * 2 * K, &err); if (!wl) return fail(err);
    const auto* bl = e->find("nets.3.bias", F, &err); if (!bl) return fail(err);
    std::vector<float> wp((size_t)64 * 512, 0.f), bp(64, 0.f);
    std::copy(w->begin(), w->end(), wp.begin());
    std::copy(b->begin(), b->end(), bp.begin());
    if (!e->up_act(wp, &e->w_kp, &err) || !e->up_f32(bp, &e->b_kp, &err) || !e->up_f32(*wl, &e->w_lin, &err) || !e->up_f32(*bl, &e->b_lin, &err)) return fail(err);
  }
  {
    const size_t NI = c.max_images;
    bool ok = true;
    ok &= (e->x0 = e->dalloc<act_t>(NI * e->Hp * (e->Wp + 8) * 4)) != nullptr;
    ok &= (e->c1 = e->dalloc<act_t>(NI * e->Ha[1] * e->Wa[1] * 64)) != nullptr;
    ok &= (e->c1n = e->dalloc<act_t>(NI * e->Ha[1] * e->Wa[1] * 64)) != nullptr;
    for (int li = 0; li < 4; ++li)
      for (int k = 0; k < 4; ++k)
        ok &= (e->st[li][k] = e->dalloc<act_t>(NI * e->Ha[2 + li] * e->Wa[2 + li] * kImgencChans[li])) != nullptr;
    ok &= (e->partial = e->dalloc<float2>(NI * kImgencMaxChunks * 32)) != nullptr;
    ok &= (e->logits = e->dalloc<float>(NI * e->Ha[5] * e->Wa[5] * 64)) != nullptr;
    if (!ok) return fail("cudaMalloc failed (imgenc workspace)");
  }
  e->host_w.clear();
  e->finalized = true;
  RMB_CUDA(cudaDeviceSynchronize());
  return 0;
}

double rmb_imgenc_flops(const rmb_imgenc_engine_t* e, int n_images) {
  if (!e) return -1.0;
  double f = 2.0 * e->Hv[1] * e->Wv[1] * 64 * 147;
  for (int li = 0; li < 4; ++li) {
    const double px = (double)e->Hv[2 + li] * e->Wv[2 + li], co = kImgencChans[li], cin = li == 0 ? 64 : kImgencChans[li - 1];
    f += 2.0 * px * co * 9 * cin + 3 * 2.0 * px * co * 9 * co;
    if (li > 0) f += 2.0 * px * co * cin;
  }
  f += 2.0 * e->Hv[5] * e->Wv[5] * e->cfg.num_kp * 512 + 2.0 * e->cfg.feat_dim * 2 * e->cfg.num_kp;
  return f * n_images;
}

}  // extern "C"

static int imgenc_build_plans(rmb_imgenc_engine* e, int NI, rmb_imgenc_engine::Plans** out) {
  auto it = e->plans.find(NI);
  if (it != e->plans.end()) { *out = it->second.get(); return 0; }
  std::unique_ptr<rmb_imgenc_engine::Plans> pl(new rmb_imgenc_engine::Plans());
  std::string err;
  if (!plan_conv1_halo(&pl->stem, e->x0, NI, e->Hp, e->Wp, e->stem.w, nullptr, e->c1, &err)) return fail(err);
  pl->stem.hp.g.relu = 0;                                       // GroupNorm comes before the ReLU
  for (int li = 0; li < 4; ++li) {
    const int h = e->Ha[2 + li], w = e->Wa[2 + li], co = kImgencChans[li];
    act_t *raw = e->st[li][0], *t = e->st[li][1], *y = e->st[li][2], *z = e->st[li][3];
    const act_t* x_in = (li == 0) ? e->st[0][3] : e->st[li - 1][3];
    if (li == 0) {
      if (!plan_conv3x3_halo(&pl->l1[0][0], x_in, NI, h, w, e->blk[0][0][0].w, nullptr, nullptr, raw, false, &err)) return fail(err);
      if (!plan_conv3x3_halo(&pl->l1[0][1], t, NI, h, w, e->blk[0][0][1].w, nullptr, nullptr, raw, false, &err)) return fail(err);
      if (!plan_conv3x3_halo(&pl->l1[1][0], y, NI, h, w, e->blk[0][1][0].w, nullptr, nullptr, raw, false, &err)) return fail(err);
      if (!plan_conv3x3_halo(&pl->l1[1][1], t, NI, h, w, e->blk[0][1][1].w, nullptr, nullptr, raw, false, &err)) return fail(err);
      continue;
    }
    const int cin = kImgencChans[li - 1];
    if (!plan_conv(&pl->down[li], x_in, NI, 2 * h, 2 * w, cin, 1, 2, e->down[li].w, co, nullptr, nullptr, false, raw, co, false, false, &err)) return fail(err);
    if (!plan_conv(&pl->conv[li][0][0], x_in, NI, 2 * h, 2 * w, cin, 3, 2, e->blk[li][0][0].w, co, nullptr, nullptr, false, raw, co, false, false, &err)) return fail(err);
    if (!plan_conv(&pl->conv[li][0][1], t, NI, h, w, co, 3, 1, e->blk[li][0][1].w, co, nullptr, nullptr, false, raw, co, false, false, &err)) return fail(err);
    if (!plan_conv(&pl->conv[li][1][0], y, NI, h, w, co, 3, 1, e->blk[li][1][0].w, co, nullptr, nullptr, false, raw, co, false, false, &err)) return fail(err);
    if (!plan_conv(&pl->conv[li][1][1], t, NI, h, w, co, 3, 1, e->blk[li][1][1].w, co, nullptr, nullptr, false, raw, co, false, false, &err)) return fail(err);
  }
  // SpatialSoftmax's 1x1 conv 512 -> num_kp (padded to 64 columns), fp32 logits
  if (!plan_conv(&pl->kp, e->st[3][3], NI, e->Ha[5], e->Wa[5], 512, 1, 1, e->w_kp, 64, e->b_kp, nullptr, false, e->logits, 64, true, false, &err)) return fail(err);
  *out = pl.get();
  e->plans[NI] = std::move(pl);
  return 0;
}

extern "C" {

int rmb_imgenc_encode(rmb_imgenc_engine_t* e, const float* d_images, float* d_out, int64_t ld_out, int n_images,
                      void* stream_) {
  if (!e || !d_images || !d_out) return fail("rmb_imgenc_encode: null argument");
  if (!e->finalized) return fail("rmb_imgenc_encode: weights not finalized");
  const rmb_imgenc_config_t& c = e->cfg;
  if (n_images < 1 || n_images > c.max_images) return fail("rmb_imgenc_encode: n_images out of range");
  if (ld_out < c.feat_dim) return fail("rmb_imgenc_encode: ld_out < feat_dim");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  rmb_imgenc_engine::Plans* pl = nullptr;
  const int NI = n_images, sms = e->num_sms;
  if (int r = imgenc_build_plans(e, NI, &pl)) return r;

  // GroupNorm (+ residual) (+ ReLU) of a raw conv output at pyramid level `lv` (index into Hv / Ha)
  auto group_norm = [&](const act_t* raw, int lv, int C, const rmb_imgenc_engine::ConvG& w, const act_t* res, act_t* out, bool relu) {
    const int H = e->Ha[lv], W = e->Wa[lv], Hv = e->Hv[lv], Wv = e->Wv[lv];
    const int nchunk = std::max(1, std::min(kImgencMaxChunks, (Hv * Wv * C / 16 + 8191) / 8192));
    gn2d_stats_kernel<<<dim3(nchunk, NI), 256, 0, s>>>(raw, e->partial, H, W, Hv, Wv, C, nchunk);
    const int napply = std::max(1, std::min(256, (H * W * (C / 8) + 8191) / 8192));   // 32 x 16 B per thread: the per-CTA statistics prologue is amortised
    gn2d_apply_kernel<<<dim3(napply, NI), 256, 0, s>>>(raw, e->partial, nchunk, w.g, w.b, res, out, H, W, Hv, Wv, C, relu ? 1 : 0, napply);
  };

  {
    const long long items = (long long)NI * e->Hp * (e->Wp / 4 + 2);
    preprocess_crop_f32_kernel<<<grid_for(items, 256, sms), 256, 0, s>>>(d_images, e->x0, NI, c.in_h, c.in_w, (c.in_h - c.crop_h) / 2,
                                                                        (c.in_w - c.crop_w) / 2, c.crop_h, c.crop_w, e->Hp, e->Wp);
  }
  RMB_CUDA(launch_halo(pl->stem, sms, s));
  group_norm(e->c1, 1, 64, e->stem, nullptr, e->c1n, true);
  {
    const long long items = (long long)NI * e->Ha[2] * e->Wa[2] * 8;
    maxpool3x3s2_kernel<<<grid_for(items, 256, sms), 256, 0, s>>>(e->c1n, e->st[0][3], NI, e->Ha[1], e->Wa[1], 64);
  }
  for (int li = 0; li < 4; ++li) {
    const int lv = 2 + li, co = kImgencChans[li];
    act_t *raw = e->st[li][0], *t = e->st[li][1], *y = e->st[li][2], *z = e->st[li][3];
    const act_t* identity = (li == 0) ? e->st[0][3] : z;
    if (li == 0) {
      RMB_CUDA(launch_halo(pl->l1[0][0], sms, s));
      group_norm(raw, lv, co, e->blk[0][0][0], nullptr, t, true);
      RMB_CUDA(launch_halo(pl->l1[0][1], sms, s));
      group_norm(raw, lv, co, e->blk[0][0][1], identity, y, true);
      RMB_CUDA(launch_halo(pl->l1[1][0], sms, s));
      group_norm(raw, lv, co, e->blk[0][1][0], nullptr, t, true);
      RMB_CUDA(launch_halo(pl->l1[1][1], sms, s));
      group_norm(raw, lv, co, e->blk[0][1][1], y, z, true);
      continue;
    }
    // conv1 first (reads the previous level), then the downsample branch into z (its raw output reuses `raw`)
    RMB_CUDA(launch_gemm(pl->conv[li][0][0], sms, s));
    group_norm(raw, lv, co, e->blk[li][0][0], nullptr, t, true);
    RMB_CUDA(launch_gemm(pl->down[li], sms, s));
    group_norm(raw, lv, co, e->down[li], nullptr, z, false);
    RMB_CUDA(launch_gemm(pl->conv[li][0][1], sms, s));
    group_norm(raw, lv, co, e->blk[li][0][1], z, y, true);
    RMB_CUDA(launch_gemm(pl->conv[li][1][0], sms, s));
    group_norm(raw, lv, co, e->blk[li][1][0], nullptr, t, true);
    RMB_CUDA(launch_gemm(pl->conv[li][1][1], sms, s));
    group_norm(raw, lv, co, e->blk[li][1][1], y, z, true);
  }
  RMB_CUDA(launch_gemm(pl->kp, sms, s));
  spatial_softmax_linear_kernel<<<NI, 64, 0, s>>>(e->logits, 64, e->Hv[5], e->Wv[5], c.num_kp, e->w_lin, e->b_lin, c.feat_dim, d_out, ld_out);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

__global__ void imgenc_act_to_f32_kernel(const act_t* __restrict__ in, float* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = static_cast<float>(in[i]);
}

int64_t rmb_imgenc_debug_read(rmb_imgenc_engine_t* e, float* d_out, int n_images, void* stream_) {
  if (!e || !d_out) return fail("rmb_imgenc_debug_read: null argument");
  const long long n = (long long)n_images * e->Ha[5] * e->Wa[5] * 512;
  imgenc_act_to_f32_kernel<<<grid_for(n, 256, e->num_sms), 256, 0, static_cast<cudaStream_t>(stream_)>>>(e->st[3][3], d_out, n);
  RMB_CUDA(cudaGetLastError());
  return n;
}

}  // extern "C"
