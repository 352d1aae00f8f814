// B200-native ACT inference engine + the C ABI declared in include/rmb_b200.h.
// Replaces, for B episodes at once, the reference call chain
//   RolloutAct.infer_policy (policy/act/RolloutAct.py:68-101) -> ACTPolicy.__call__ (third_party/act/policy.py:18-38)
//   -> DETRVAE.forward inference branch (third_party/act/detr/models/detr_vae.py:89-150)
//   -> Transformer.forward (third_party/act/detr/models/transformer.py:49-77).
// Only the live graph is executed: decoder layers 1..6 and the CVAE encoder never influence a_hat (detr_vae.py:142
// takes hs[0]); decoder layer 0's self-attention sees tgt = 0, so its output is the constant W_o b_v + b_o.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rmb_b200.h"
#include "attention_tc.cuh"
#include "attention_v2.cuh"
#include "attention_v3.cuh"
#include "attention_v4.cuh"
#include "elementwise.cuh"
#include "gemm_plan.h"

using namespace rmb;

static thread_local std::string g_last_error;
static int fail(const std::string& msg, int code = -1) { g_last_error = msg; return code; }
#define RMB_CUDA(expr)                                                                             \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess) return fail(std::string(#expr) + ": " + cudaGetErrorString(_e), -2);    \
  } while (0)

namespace {

struct AttnPlan {
  CUtensorMap tq, tk, tv;
  CUtensorMap tk64, tv64;   // 64-key boxes of K / V (attention_v4_kernel)
  AttnParams p;
  int grid = 0;
};

bool plan_attention(AttnPlan* a, const void* Q, int ldq, int q_col0, bool q_shared, const void* K, int ldk, int k_col0,
                    const void* V, int ldv, int v_col0, void* out, int ld_out, int B, int n_heads, int Lq, int Lk,
                    std::string* err) {
  const uint32_t box[3] = {64, 128, 1};
  {
    const uint64_t dims[3] = {(uint64_t)ldq, (uint64_t)Lq, (uint64_t)(q_shared ? 1 : B)};
    const uint64_t str[2] = {(uint64_t)ldq * 2, (uint64_t)ldq * 2 * Lq};
    if (!make_tmap_act(&a->tq, Q, 3, dims, str, box, 128, err)) return false;
  }
  {
    const uint64_t dims[3] = {(uint64_t)ldk, (uint64_t)Lk, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldk * 2, (uint64_t)ldk * 2 * Lk};
    if (!make_tmap_act(&a->tk, K, 3, dims, str, box, 128, err)) return false;
  }
  {
    const uint64_t dims[3] = {(uint64_t)ldv, (uint64_t)Lk, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldv * 2, (uint64_t)ldv * 2 * Lk};
    if (!make_tmap_act(&a->tv, V, 3, dims, str, box, 128, err)) return false;
    const uint32_t box64[3] = {64, 64, 1};
    if (!make_tmap_act(&a->tv64, V, 3, dims, str, box64, 128, err)) return false;
  }
  {
    const uint64_t dims[3] = {(uint64_t)ldk, (uint64_t)Lk, (uint64_t)B};
    const uint64_t str[2] = {(uint64_t)ldk * 2, (uint64_t)ldk * 2 * Lk};
    const uint32_t box64[3] = {64, 64, 1};
    if (!make_tmap_act(&a->tk64, K, 3, dims, str, box64, 128, err)) return false;
  }
  AttnParams& p = a->p;
  p.Lq = Lq; p.Lk = Lk; p.n_q_tiles = (Lq + 127) / 128; p.n_heads = n_heads;
  p.q_col0 = q_col0; p.k_col0 = k_col0; p.v_col0 = v_col0; p.q_shared = q_shared ? 1 : 0;
  p.scale_log2e = 0.125f * 1.4426950408889634f;
  p.out = static_cast<act_t*>(out); p.ld_out = ld_out;
  a->grid = B * n_heads * p.n_q_tiles;
  return true;
}

// Attention kernel variants (same results within the 16-bit P rounding; tests run all three):
//   1 attention_tc_kernel  two sweeps, one tile per CTA, two CTAs per SM (round 1)
//   2 attention_v2_kernel  persistent, two streams per CTA, one sweep, P through shared memory (round 2, not faster)
//   3 attention_v3_kernel  v2's structure with P in TMEM as the A operand of P V, S released before the arithmetic,
//                          packed fp32x2 softmax arithmetic
//   4 attention_v4_kernel  v3 with three streams per CTA and 64-key blocks (three softmax warps per scheduler); default
// RMB_ATTN=1|2|3|4 (or RMB_ATTN_V2=1) picks one at start-up; rmb_set_attention_variant() switches at run time.
#ifndef RMB_ATTN_DEFAULT
#define RMB_ATTN_DEFAULT 4
#endif
cudaError_t launch_attention_v1(const AttnPlan& a, cudaStream_t stream);
static int g_attn_variant = [] {
  const char* e = getenv("RMB_ATTN");
  if (e != nullptr && e[0] >= '1' && e[0] <= '4') return e[0] - '0';
  const char* e2 = getenv("RMB_ATTN_V2");
  if (e2 != nullptr && e2[0] == '1') return 2;
  return RMB_ATTN_DEFAULT;
}();
int attention_variant() { return g_attn_variant; }
const char* attention_kernel_name() {
  return g_attn_variant == 4 ? "attention_v4_kernel" : g_attn_variant == 3 ? "attention_v3_kernel" : g_attn_variant == 2 ? "attention_v2_kernel" : "attention_tc_kernel";
}
cudaError_t launch_attention(const AttnPlan& a, cudaStream_t stream) {
  if (g_attn_variant == 1) return launch_attention_v1(a, stream);
  static bool attr_set = false;
  static int sms = 148;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttn2SmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(attention_v3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttn3SmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(attention_v4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttn4SmemBytes);
    if (e != cudaSuccess) return e;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    attr_set = true;
  }
  const int n_items = a.grid;                       // (episode, head, query tile) work items
  if (g_attn_variant == 4)
    return launch_pdl(attention_v4_kernel, dim3(std::min(sms, (n_items + kAttn4Streams - 1) / kAttn4Streams)), dim3(kAttn4Threads),
                      kAttn4SmemBytes, stream, a.tq, a.tk64, a.tv64, a.p, n_items);
  const int grid = std::min(sms, (n_items + 1) / 2);
  if (g_attn_variant == 2)
    return launch_pdl(attention_v2_kernel, dim3(grid), dim3(kAttn2Threads), kAttn2SmemBytes, stream, a.tq, a.tk, a.tv, a.p, n_items);
  return launch_pdl(attention_v3_kernel, dim3(grid), dim3(kAttn3Threads), kAttn3SmemBytes, stream, a.tq, a.tk, a.tv, a.p, n_items);
}

cudaError_t launch_attention_v1(const AttnPlan& a, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmemBytes);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  return launch_pdl(attention_tc_kernel, dim3(a.grid), dim3(kAttnThreads), kAttnSmemBytes, stream, a.tq, a.tk, a.tv, a.p);
}

int grid_for(long long work_items, int threads, int num_sms) {
  long long blocks = (work_items + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms) * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace

// =====================================================================================================================
struct rmb_act_engine {
  rmb_act_config_t cfg;
  int num_sms = 148;
  bool finalized = false;
  bool attn_debug = false;          // opt-in: head-averaged encoder attention maps of one episode (RolloutAct.draw_plot)
  int attn_debug_episode = 0;
  float* attn_maps = nullptr;       // [enc_layers, L, L]
  bool check_saturation = false;   // diagnostic mode: scan every 16-bit tensor a forward stores (rmb_act_set_check_saturation)
  std::unordered_map<std::string, std::vector<float>> host_w;
  std::vector<void*> allocs;

  // derived
  int Hf = 0, Wf = 0, L = 0, n_img_max = 0;
  int n_prefix = 2;        // tokens in front of the image tokens: latent, proprio (+ task token if cfg.task_token)

  // ---- MT-ACT (cfg.multi_task): text-conditioned FiLM of the ResNet trunk; ARCHITECTURE HYPOTHESIS, see DESIGN.md
  static constexpr int kFilmLd = 2 * 2 * (128 + 256 + 512);   // per episode: layers 2..4 x [gamma|beta] x 2 blocks x planes
  act_t *w_text = nullptr, *w_film = nullptr;                 // proj_text_emb [512, E]; film generators [kFilmLd, 512]
  float *b_text = nullptr, *b_film = nullptr;
  act_t *task_emb_act = nullptr, *task_tok_act = nullptr;     // [max_batch, E], [max_batch, 512]
  float *task_tok = nullptr, *film = nullptr;                 // [max_batch, 512], [max_batch, kFilmLd]
  int task_batch = 0;                                         // episodes covered by the last rmb_act_set_task_embedding

  // ---- device weights
  struct ConvW { act_t* w = nullptr; float* b = nullptr; int cin = 0, cout = 0, k = 0, stride = 1; };
  ConvW stem;
  ConvW blk[4][2][2];      // [layer][block][conv]
  ConvW down[4];           // layer 1..3 (index = layer), layer 0 unused
  ConvW c1down[4];         // conv1 of block 0 with the downsample conv as side columns (make_conv1_with_downsample)
  // RMB_DOWN_FUSE=1: the downsample convs as side column tiles of conv1's launch. Correct and bit-identical, but measured
  // SLOWER on B200 (layer2.0: 0.95 ms against 0.25 + 0.45 ms; step 19.8 against 19.1 ms): a side tile is one k-block of
  // MMA work and a full epilogue, and the plain epilogue (~1 us per 128 x 128 tile) then runs twice per nine k-blocks,
  // so the launch becomes epilogue-bound; in their own launches the two epilogues overlap different MMAs. Opt-in.
  bool fuse_down = false;
  ConvW input_proj;
  struct EncW {
    act_t *w_qk, *w_v, *w_o, *w_1, *w_2;
    float *b_qk, *b_v, *b_o, *b_1, *b_2, *g1, *be1, *g2, *be2;
  };
  std::vector<EncW> enc;
  struct DecW {
    act_t *q_const;                 // [C, 512] projected queries (input independent)
    act_t *w_k, *w_v, *w_o, *w_1, *w_2;
    float *b_k, *b_v, *b_o_fold, *b_1, *b_2, *g2, *be2, *g3, *be3, *gn, *ben;
  } dec{};
  float *w_head = nullptr, *b_head = nullptr, *w_state = nullptr, *b_state = nullptr, *latent_bias = nullptr;
  float* pos = nullptr;    // [L, 512]

  // ---- workspace (sized for max_batch)
  act_t *x0 = nullptr, *c1 = nullptr;
  act_t* stage[4][3] = {};
  float *tok_f32 = nullptr, *y_f32 = nullptr, *x1_f32 = nullptr;
  act_t *tok_b = nullptr, *tok_pb = nullptr, *qk = nullptr, *v = nullptr, *ao = nullptr, *x1_b = nullptr, *hh = nullptr;
  act_t *kd = nullptr, *vd = nullptr, *aod = nullptr, *t_b = nullptr, *hd = nullptr;
  float *yd = nullptr, *t_f32 = nullptr, *yd2 = nullptr;

  // RMB_LN_FUSE: never (default) = residual GEMM + layernorm512_kernel; always = the encoder's residual GEMMs carry
  // their LayerNorm (gemm_ln.cuh); auto = fused when there is a 256-row pair tile for every SM pair. The fused launch
  // is correct (tests) but measured SLOWER on B200 (out-proj 0.40 vs 0.20 + 0.13 ms, FFN-down 0.64 vs 0.39 + 0.16 ms at
  // 256 episodes, same box): one TMEM accumulator serialises the MMAs with the first epilogue phase and the epilogue is
  // memory-latency bound (ncu: profiles/r02_gemm_ln_out_proj_ncu_full_summary.txt), so it is opt-in.
  int fuse_ln = 0;              // 0 never, 1 auto, 2 always
  // Residual operand of the encoder's two residual GEMMs: the 16-bit copy of the LayerNorm output (the very tensor the
  // next GEMMs read as their A operand) instead of a separate fp32 copy. The LayerNorm kernels then write no fp32 output
  // (6 -> 4 / 5 -> 3 KB per token) and the out-projection reads 1 KB less per token; the pre-norm sums stay fp32. Default
  // in the fp16 build (one more 2^-11 rounding per sublayer); the bf16 build keeps the fp32 copy. RMB_RES_F32=1 / 0 overrides.
#ifdef RMB_ACT_BF16
  bool res16 = false;
#else
  bool res16 = true;
#endif
  bool fuse_stem_pool = true;   // RMB_NO_STEM_POOL=1: stem and max-pool as two kernels through a [img, H/2, W/2, 64] buffer
  struct BatchPlans {
    HaloPlan stem;
    StemPoolPlan stem_pool;
    GemmPlan conv[4][2][2];
    HaloPlan conv_l1[2][2];   // layer1 (Cin = Cout = 64): halo-reuse kernel
    GemmPlan down[4];
    GemmPlan input_proj;
    struct Enc { GemmPlan qk, v, o, ff1, ff2, o_ln, ff2_ln; AttnPlan attn; };
    bool ln_fused = false;    // enough rows for a 256 x 512 pair tile per SM pair: the two residual GEMMs carry their LayerNorm
    std::vector<Enc> enc;
    GemmPlan dk, dv, dout, dff1, dff2;
    AttnPlan dattn;
  };
  std::map<int, std::unique_ptr<BatchPlans>> plans;

  template <typename T>
  T* dalloc(size_t n) {
    void* p = nullptr;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) return nullptr;
    allocs.push_back(p);
    return static_cast<T*>(p);
  }
};

namespace {

const std::vector<float>* find_w(rmb_act_engine* e, const std::string& name, size_t numel, std::string* err) {
  auto it = e->host_w.find(name);
  if (it == e->host_w.end()) { *err = "missing checkpoint tensor: " + name; return nullptr; }
  if (it->second.size() != numel) {
    *err = "checkpoint tensor " + name + " has " + std::to_string(it->second.size()) + " elements, expected " + std::to_string(numel);
    return nullptr;
  }
  return &it->second;
}

bool upload_f32(rmb_act_engine* e, const std::vector<float>& h, float** d, std::string* err) {
  *d = e->dalloc<float>(h.size());
  if (!*d) { *err = "cudaMalloc failed (weights)"; return false; }
  if (cudaMemcpy(*d, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "cudaMemcpy H2D failed"; return false; }
  return true;
}
bool upload_act(rmb_act_engine* e, const std::vector<float>& h, act_t** d, std::string* err) {
  std::vector<act_t> t(h.size());
  for (size_t i = 0; i < h.size(); ++i) t[i] = to_act(h[i]);
  *d = e->dalloc<act_t>(h.size());
  if (!*d) { *err = "cudaMalloc failed (weights)"; return false; }
  if (cudaMemcpy(*d, t.data(), t.size() * sizeof(act_t), cudaMemcpyHostToDevice) != cudaSuccess) { *err = "cudaMemcpy H2D failed"; return false; }
  return true;
}

// FrozenBatchNorm2d (third_party/act/detr/models/backbone.py:21-57): y = x*scale + bias,
// scale = w * rsqrt(rv + 1e-5), bias = b - rm*scale  -> folded into the conv that precedes it.
bool bn_fold(rmb_act_engine* e, const std::string& bn, int c, std::vector<float>* scale, std::vector<float>* bias, std::string* err) {
  const auto* w = find_w(e, bn + ".weight", c, err); if (!w) return false;
  const auto* b = find_w(e, bn + ".bias", c, err); if (!b) return false;
  const auto* rm = find_w(e, bn + ".running_mean", c, err); if (!rm) return false;
  const auto* rv = find_w(e, bn + ".running_var", c, err); if (!rv) return false;
  scale->resize(c); bias->resize(c);
  for (int i = 0; i < c; ++i) {
    const float s = (*w)[i] * (1.0f / sqrtf((*rv)[i] + 1e-5f));
    (*scale)[i] = s;
    (*bias)[i] = (*b)[i] - (*rm)[i] * s;
  }
  return true;
}

// BN-folded conv weights on the host: r [cout][k][k][cin] (tap-major, channel-minor), bias [cout]
bool fold_conv_host(rmb_act_engine* e, const std::string& conv, const std::string& bn, int cin, int cout, int k,
                    std::vector<float>* r, std::vector<float>* bias, std::string* err) {
  const auto* w = find_w(e, conv + ".weight", (size_t)cout * cin * k * k, err); if (!w) return false;
  std::vector<float> scale;
  if (!bn_fold(e, bn, cout, &scale, bias, err)) return false;
  r->assign((size_t)cout * k * k * cin, 0.f);
  for (int co = 0; co < cout; ++co)
    for (int ci = 0; ci < cin; ++ci)
      for (int kh = 0; kh < k; ++kh)
        for (int kw = 0; kw < k; ++kw)
          (*r)[(((size_t)co * k + kh) * k + kw) * cin + ci] = (*w)[(((size_t)co * cin + ci) * k + kh) * k + kw] * scale[co];
  return true;
}

bool make_conv(rmb_act_engine* e, const std::string& conv, const std::string& bn, int cin, int cout, int k, int stride,
               rmb_act_engine::ConvW* out, std::string* err) {
  std::vector<float> r, bias;
  if (!fold_conv_host(e, conv, bn, cin, cout, k, &r, &bias, err)) return false;
  out->cin = cin; out->cout = cout; out->k = k; out->stride = stride;
  return upload_act(e, r, &out->w, err) && upload_f32(e, bias, &out->b, err);
}

// conv1 (3x3 stride 2) of a down-sampling BasicBlock with the block's 1x1 stride-2 downsample conv as extra output
// rows: [2 * cout, 9 * cin], the downsample rows non-zero in the centre tap only (backbone.py:86-96 -> torchvision
// BasicBlock: identity = downsample(x) reads x[2h, 2w], the pixel conv1's centre tap reads); bias [2 * cout].
bool make_conv1_with_downsample(rmb_act_engine* e, const std::string& blk, int cin, int cout, rmb_act_engine::ConvW* out,
                                std::string* err) {
  std::vector<float> r1, b1, rd, bd;
  if (!fold_conv_host(e, blk + "conv1", blk + "bn1", cin, cout, 3, &r1, &b1, err)) return false;
  if (!fold_conv_host(e, blk + "downsample.0", blk + "downsample.1", cin, cout, 1, &rd, &bd, err)) return false;
  std::vector<float> rc((size_t)2 * cout * 9 * cin, 0.f), bc(2 * (size_t)cout);
  std::copy(r1.begin(), r1.end(), rc.begin());
  for (int co = 0; co < cout; ++co) {
    for (int ci = 0; ci < cin; ++ci) rc[(((size_t)(cout + co)) * 9 + 4) * cin + ci] = rd[(size_t)co * cin + ci];
    bc[co] = b1[co]; bc[cout + co] = bd[co];
  }
  out->cin = cin; out->cout = 2 * cout; out->k = 3; out->stride = 2;
  return upload_act(e, rc, &out->w, err) && upload_f32(e, bc, &out->b, err);
}

void layernorm_host(const std::vector<double>& x, const std::vector<float>& g, const std::vector<float>& b, std::vector<double>* y) {
  const size_t n = x.size();
  double mu = 0; for (double v : x) mu += v; mu /= n;
  double var = 0; for (double v : x) var += (v - mu) * (v - mu); var /= n;
  const double rstd = 1.0 / sqrt(var + 1e-5);
  y->resize(n);
  for (size_t i = 0; i < n; ++i) (*y)[i] = (x[i] - mu) * rstd * g[i] + b[i];
}

}  // namespace

// =====================================================================================================================
extern "C" {

const char* rmb_last_error(void) { return g_last_error.c_str(); }
const char* rmb_version(void) { return "rmb_b200 0.1 (sm_100a, " RMB_ACT_DTYPE_NAME " operands, fp32 accumulate)"; }
const char* rmb_act_dtype(void) { return RMB_ACT_DTYPE_NAME; }

int rmb_device_fault(int clear) {
  int v = 0;
  RMB_CUDA(cudaDeviceSynchronize());
  RMB_CUDA(cudaMemcpyFromSymbol(&v, rmb::g_rmb_fault, sizeof(int)));
  if (v != 0 && clear) {
    const int zero = 0;
    RMB_CUDA(cudaMemcpyToSymbol(rmb::g_rmb_fault, &zero, sizeof(int)));
  }
  if (v != 0) g_last_error = "device fault flag set: an mbarrier wait timed out inside a kernel (pipeline wedged)";
  return v;
}

int rmb_check_device(int dev) {
  cudaDeviceProp prop;
  RMB_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) return fail(std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) + ", this library is sm_100a only", -3);
  return 0;
}

int rmb_act_engine_create(const rmb_act_config_t* cfg, rmb_act_engine_t** out) {
  if (!cfg || !out) return fail("null argument");
  if (cfg->hidden_dim != 512) return fail("hidden_dim must be 512");
  if (cfg->nheads != 8) return fail("nheads must be 8 (head dim 64)");
  if (cfg->dim_feedforward % 64 != 0) return fail("dim_feedforward must be a multiple of 64");
  if (cfg->img_h % 32 != 0 || cfg->img_w % 32 != 0) return fail("image size must be a multiple of 32");
  if (cfg->max_batch < 1 || cfg->num_cams < 1 || cfg->chunk_size < 1 || cfg->enc_layers < 1 || cfg->dec_layers < 1) return fail("bad config");
  if (cfg->state_dim < 1 || cfg->action_dim < 1) return fail("bad state/action dim");
  if (cfg->multi_task && (cfg->task_emb_dim < 64 || cfg->task_emb_dim % 64 != 0)) return fail("task_emb_dim must be a multiple of 64 (384 for all-MiniLM-L6-v2)");
  if (!cfg->multi_task && cfg->task_token) return fail("task_token needs multi_task");
  RMB_CUDA(cudaSetDevice(cfg->device));
  if (int r = rmb_check_device(cfg->device)) return r;
  auto* e = new rmb_act_engine();
  e->cfg = *cfg;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  e->num_sms = prop.multiProcessorCount;
  { const char* ev = getenv("RMB_NO_STEM_POOL"); e->fuse_stem_pool = !(ev != nullptr && ev[0] == '1'); }
  { const char* ev = getenv("RMB_LN_FUSE"); e->fuse_ln = (ev == nullptr) ? 0 : (ev[0] == 'n' ? 0 : (ev[0] == 'a' && ev[1] == 'l') ? 2 : 1); }
  { const char* ev = getenv("RMB_DOWN_FUSE"); e->fuse_down = (ev != nullptr && ev[0] == '1'); }
  { const char* ev = getenv("RMB_RES_F32"); if (ev != nullptr && (ev[0] == '0' || ev[0] == '1')) e->res16 = (ev[0] == '0'); }
  e->Hf = cfg->img_h / 32; e->Wf = cfg->img_w / 32;
  e->n_prefix = 2 + (cfg->task_token ? 1 : 0);
  e->L = e->n_prefix + e->Hf * e->Wf * cfg->num_cams;
  e->n_img_max = cfg->max_batch * cfg->num_cams;
  *out = e;
  return 0;
}

void rmb_act_engine_destroy(rmb_act_engine_t* e) {
  if (!e) return;
  cudaSetDevice(e->cfg.device);
  cudaDeviceSynchronize();
  for (void* p : e->allocs) cudaFree(p);
  delete e;
}

int rmb_act_load_tensor(rmb_act_engine_t* e, const char* name, const float* host_data, int64_t numel) {
  if (!e || !name || (!host_data && numel > 0) || numel < 0) return fail("rmb_act_load_tensor: bad argument");
  if (e->finalized) return fail("rmb_act_load_tensor: engine already finalized");
  e->host_w[name] = std::vector<float>(host_data, host_data + numel);
  return 0;
}

int rmb_act_finalize(rmb_act_engine_t* e) {
  if (!e) return fail("null engine");
  if (e->finalized) return fail("already finalized");
  RMB_CUDA(cudaSetDevice(e->cfg.device));
  std::string err;
  const rmb_act_config_t& c = e->cfg;
  const int D = 512, FF = c.dim_feedforward, C = c.chunk_size;
  const std::string bb = "model.backbones.0.0.body.";

  // ---- ResNet-18 trunk
  {
    // stem: [64,3,7,7] -> [64][kh][slot 0..7][4], slot j>=1 <-> kw = j-1, channel 3 and slot 0 are zero
    const auto* w = find_w(e, bb + "conv1.weight", 64 * 3 * 49, &err); if (!w) return fail(err);
    std::vector<float> scale, bias;
    if (!bn_fold(e, bb + "bn1", 64, &scale, &bias, &err)) return fail(err);
    std::vector<float> r(64 * 7 * 32, 0.f);
    for (int co = 0; co < 64; ++co)
      for (int ci = 0; ci < 3; ++ci)
        for (int kh = 0; kh < 7; ++kh)
          for (int kw = 0; kw < 7; ++kw)
            r[((co * 7 + kh) * 8 + (kw + 1)) * 4 + ci] = (*w)[((co * 3 + ci) * 7 + kh) * 7 + kw] * scale[co];
    e->stem.cin = 3; e->stem.cout = 64; e->stem.k = 7; e->stem.stride = 2;
    if (!upload_act(e, r, &e->stem.w, &err) || !upload_f32(e, bias, &e->stem.b, &err)) return fail(err);
  }
  const int chans[4] = {64, 128, 256, 512};
  for (int li = 0; li < 4; ++li)
    for (int bi = 0; bi < 2; ++bi) {
      const std::string p = bb + "layer" + std::to_string(li + 1) + "." + std::to_string(bi) + ".";
      const int cin = (bi == 0 && li > 0) ? chans[li - 1] : chans[li];
      const int stride = (bi == 0 && li > 0) ? 2 : 1;
      if (!make_conv(e, p + "conv1", p + "bn1", cin, chans[li], 3, stride, &e->blk[li][bi][0], &err)) return fail(err);
      if (!make_conv(e, p + "conv2", p + "bn2", chans[li], chans[li], 3, 1, &e->blk[li][bi][1], &err)) return fail(err);
      if (bi == 0 && li > 0) {
        if (!make_conv(e, p + "downsample.0", p + "downsample.1", cin, chans[li], 1, 2, &e->down[li], &err)) return fail(err);
        if (e->fuse_down && !make_conv1_with_downsample(e, p, cin, chans[li], &e->c1down[li], &err)) return fail(err);
      }
    }
  {
    // input_proj: Conv2d 1x1 512 -> hidden with bias, no BN (detr_vae.py:67,135)
    const auto* w = find_w(e, "model.input_proj.weight", (size_t)D * 512, &err); if (!w) return fail(err);
    const auto* b = find_w(e, "model.input_proj.bias", D, &err); if (!b) return fail(err);
    e->input_proj.cin = 512; e->input_proj.cout = D; e->input_proj.k = 1; e->input_proj.stride = 1;
    if (!upload_act(e, *w, &e->input_proj.w, &err) || !upload_f32(e, *b, &e->input_proj.b, &err)) return fail(err);
  }

  // ---- encoder layers (transformer.py:143-197)
  e->enc.resize(c.enc_layers);
  for (int l = 0; l < c.enc_layers; ++l) {
    const std::string p = "model.transformer.encoder.layers." + std::to_string(l) + ".";
    const auto* wi = find_w(e, p + "self_attn.in_proj_weight", (size_t)3 * D * D, &err); if (!wi) return fail(err);
    const auto* bi = find_w(e, p + "self_attn.in_proj_bias", 3 * D, &err); if (!bi) return fail(err);
    const auto* wo = find_w(e, p + "self_attn.out_proj.weight", (size_t)D * D, &err); if (!wo) return fail(err);
    const auto* bo = find_w(e, p + "self_attn.out_proj.bias", D, &err); if (!bo) return fail(err);
    const auto* w1 = find_w(e, p + "linear1.weight", (size_t)FF * D, &err); if (!w1) return fail(err);
    const auto* b1 = find_w(e, p + "linear1.bias", FF, &err); if (!b1) return fail(err);
    const auto* w2 = find_w(e, p + "linear2.weight", (size_t)D * FF, &err); if (!w2) return fail(err);
    const auto* b2 = find_w(e, p + "linear2.bias", D, &err); if (!b2) return fail(err);
    const auto* g1 = find_w(e, p + "norm1.weight", D, &err); if (!g1) return fail(err);
    const auto* be1 = find_w(e, p + "norm1.bias", D, &err); if (!be1) return fail(err);
    const auto* g2 = find_w(e, p + "norm2.weight", D, &err); if (!g2) return fail(err);
    const auto* be2 = find_w(e, p + "norm2.bias", D, &err); if (!be2) return fail(err);
    rmb_act_engine::EncW& w = e->enc[l];
    std::vector<float> wqk(wi->begin(), wi->begin() + (size_t)2 * D * D), wv(wi->begin() + (size_t)2 * D * D, wi->end());
    std::vector<float> bqk(bi->begin(), bi->begin() + 2 * D), bv(bi->begin() + 2 * D, bi->end());
    if (!upload_act(e, wqk, &w.w_qk, &err) || !upload_act(e, wv, &w.w_v, &err) || !upload_act(e, *wo, &w.w_o, &err) ||
        !upload_act(e, *w1, &w.w_1, &err) || !upload_act(e, *w2, &w.w_2, &err) || !upload_f32(e, bqk, &w.b_qk, &err) ||
        !upload_f32(e, bv, &w.b_v, &err) || !upload_f32(e, *bo, &w.b_o, &err) || !upload_f32(e, *b1, &w.b_1, &err) ||
        !upload_f32(e, *b2, &w.b_2, &err) || !upload_f32(e, *g1, &w.g1, &err) || !upload_f32(e, *be1, &w.be1, &err) ||
        !upload_f32(e, *g2, &w.g2, &err) || !upload_f32(e, *be2, &w.be2, &err))
      return fail(err);
  }

  // ---- decoder layer 0 (transformer.py:229-250) with its input-independent parts folded
  {
    const std::string p = "model.transformer.decoder.layers.0.";
    const auto* sa_b = find_w(e, p + "self_attn.in_proj_bias", 3 * D, &err); if (!sa_b) return fail(err);
    const auto* sa_wo = find_w(e, p + "self_attn.out_proj.weight", (size_t)D * D, &err); if (!sa_wo) return fail(err);
    const auto* sa_bo = find_w(e, p + "self_attn.out_proj.bias", D, &err); if (!sa_bo) return fail(err);
    const auto* g1 = find_w(e, p + "norm1.weight", D, &err); if (!g1) return fail(err);
    const auto* be1 = find_w(e, p + "norm1.bias", D, &err); if (!be1) return fail(err);
    const auto* wi = find_w(e, p + "multihead_attn.in_proj_weight", (size_t)3 * D * D, &err); if (!wi) return fail(err);
    const auto* bi = find_w(e, p + "multihead_attn.in_proj_bias", 3 * D, &err); if (!bi) return fail(err);
    const auto* wo = find_w(e, p + "multihead_attn.out_proj.weight", (size_t)D * D, &err); if (!wo) return fail(err);
    const auto* bo = find_w(e, p + "multihead_attn.out_proj.bias", D, &err); if (!bo) return fail(err);
    const auto* w1 = find_w(e, p + "linear1.weight", (size_t)FF * D, &err); if (!w1) return fail(err);
    const auto* b1 = find_w(e, p + "linear1.bias", FF, &err); if (!b1) return fail(err);
    const auto* w2 = find_w(e, p + "linear2.weight", (size_t)D * FF, &err); if (!w2) return fail(err);
    const auto* b2 = find_w(e, p + "linear2.bias", D, &err); if (!b2) return fail(err);
    const auto* g2 = find_w(e, p + "norm2.weight", D, &err); if (!g2) return fail(err);
    const auto* be2 = find_w(e, p + "norm2.bias", D, &err); if (!be2) return fail(err);
    const auto* g3 = find_w(e, p + "norm3.weight", D, &err); if (!g3) return fail(err);
    const auto* be3 = find_w(e, p + "norm3.bias", D, &err); if (!be3) return fail(err);
    const auto* gn = find_w(e, "model.transformer.decoder.norm.weight", D, &err); if (!gn) return fail(err);
    const auto* ben = find_w(e, "model.transformer.decoder.norm.bias", D, &err); if (!ben) return fail(err);
    const auto* qe = find_w(e, "model.query_embed.weight", (size_t)C * D, &err); if (!qe) return fail(err);
    // self-attention with value = tgt = 0: every value row is b_v, softmax rows sum to 1 -> sa = W_o b_v + b_o
    std::vector<double> sa(D);
    for (int i = 0; i < D; ++i) {
      double acc = (*sa_bo)[i];
      for (int k = 0; k < D; ++k) acc += (double)(*sa_wo)[(size_t)i * D + k] * (double)(*sa_b)[2 * D + k];
      sa[i] = acc;
    }
    std::vector<double> tgt1;
    layernorm_host(sa, *g1, *be1, &tgt1);                       // tgt after norm1: one vector for all queries/episodes
    std::vector<float> qc((size_t)C * D);
    for (int qi = 0; qi < C; ++qi)
      for (int i = 0; i < D; ++i) {
        double acc = (*bi)[i];
        for (int k = 0; k < D; ++k) acc += (double)(*wi)[(size_t)i * D + k] * (tgt1[k] + (double)(*qe)[(size_t)qi * D + k]);
        qc[(size_t)qi * D + i] = (float)acc;
      }
    std::vector<float> wk(wi->begin() + (size_t)D * D, wi->begin() + (size_t)2 * D * D), wv(wi->begin() + (size_t)2 * D * D, wi->end());
    std::vector<float> bk(bi->begin() + D, bi->begin() + 2 * D), bv(bi->begin() + 2 * D, bi->end());
    std::vector<float> bo_fold(D);
    for (int i = 0; i < D; ++i) bo_fold[i] = (float)((double)(*bo)[i] + tgt1[i]);   // residual tgt1 folded into the bias
    auto& d = e->dec;
    if (!upload_act(e, qc, &d.q_const, &err) || !upload_act(e, wk, &d.w_k, &err) || !upload_act(e, wv, &d.w_v, &err) ||
        !upload_act(e, *wo, &d.w_o, &err) || !upload_act(e, *w1, &d.w_1, &err) || !upload_act(e, *w2, &d.w_2, &err) ||
        !upload_f32(e, bk, &d.b_k, &err) || !upload_f32(e, bv, &d.b_v, &err) || !upload_f32(e, bo_fold, &d.b_o_fold, &err) ||
        !upload_f32(e, *b1, &d.b_1, &err) || !upload_f32(e, *b2, &d.b_2, &err) || !upload_f32(e, *g2, &d.g2, &err) ||
        !upload_f32(e, *be2, &d.be2, &err) || !upload_f32(e, *g3, &d.g3, &err) || !upload_f32(e, *be3, &d.be3, &err) ||
        !upload_f32(e, *gn, &d.gn, &err) || !upload_f32(e, *ben, &d.ben, &err))
      return fail(err);
  }
  {
    const auto* wh = find_w(e, "model.action_head.weight", (size_t)c.action_dim * D, &err); if (!wh) return fail(err);
    const auto* bh = find_w(e, "model.action_head.bias", c.action_dim, &err); if (!bh) return fail(err);
    const auto* ws = find_w(e, "model.input_proj_robot_state.weight", (size_t)D * c.state_dim, &err); if (!ws) return fail(err);
    const auto* bs = find_w(e, "model.input_proj_robot_state.bias", D, &err); if (!bs) return fail(err);
    const auto* lb = find_w(e, "model.latent_out_proj.bias", D, &err); if (!lb) return fail(err);
    const auto* ap = find_w(e, "model.additional_pos_embed.weight", (size_t)e->n_prefix * D, &err); if (!ap) return fail(err);
    if (!upload_f32(e, *wh, &e->w_head, &err) || !upload_f32(e, *bh, &e->b_head, &err) || !upload_f32(e, *ws, &e->w_state, &err) ||
        !upload_f32(e, *bs, &e->b_state, &err) || !upload_f32(e, *lb, &e->latent_bias, &err))
      return fail(err);
    // position table [L, 512]: rows 0,1 = additional_pos_embed; then PositionEmbeddingSine(256, normalize=True)
    // (position_encoding.py:30-52) of a (Hf, Wf) map, identical for every camera, cameras folded along width
    // (detr_vae.py:140-141): token 2 + h*(Wf*Ncam) + cam*Wf + w.
    std::vector<float> pos((size_t)e->L * D);
    for (int i = 0; i < e->n_prefix * D; ++i) pos[i] = (*ap)[i];
    const float two_pi = 2.0f * 3.14159265358979323846f;
    for (int h = 0; h < e->Hf; ++h)
      for (int w = 0; w < e->Wf; ++w) {
        const float y_embed = (float)(h + 1) / ((float)e->Hf + 1e-6f) * two_pi;
        const float x_embed = (float)(w + 1) / ((float)e->Wf + 1e-6f) * two_pi;
        float row[512];
        for (int i = 0; i < 256; ++i) {
          const float dim_t = powf(10000.0f, 2.0f * (float)(i / 2) / 256.0f);
          const float ay = y_embed / dim_t, ax = x_embed / dim_t;
          row[i] = (i % 2 == 0) ? sinf(ay) : cosf(ay);
          row[256 + i] = (i % 2 == 0) ? sinf(ax) : cosf(ax);
        }
        for (int cam = 0; cam < c.num_cams; ++cam) {
          const size_t t = e->n_prefix + (size_t)h * (e->Wf * c.num_cams) + (size_t)cam * e->Wf + w;
          memcpy(&pos[t * D], row, sizeof(row));
        }
      }
    if (!upload_f32(e, pos, &e->pos, &err)) return fail(err);
  }

  // ---- MT-ACT: text projection and the FiLM generators of layers 2..4 (hypothesised keys, see policy/mt_act/policy.py)
  if (c.multi_task) {
    const int E = c.task_emb_dim;
    const auto* wt = find_w(e, "model.proj_text_emb.weight", (size_t)D * E, &err); if (!wt) return fail(err);
    const auto* bt = find_w(e, "model.proj_text_emb.bias", D, &err); if (!bt) return fail(err);
    std::vector<float> wf, bf;
    const int planes[3] = {128, 256, 512};
    for (int i = 0; i < 3; ++i) {
      const std::string p = "model.backbones.0.0.film_models." + std::to_string(i + 1) + ".";
      const auto* w = find_w(e, p + "weight", (size_t)4 * planes[i] * D, &err); if (!w) return fail(err);
      const auto* b = find_w(e, p + "bias", (size_t)4 * planes[i], &err); if (!b) return fail(err);
      wf.insert(wf.end(), w->begin(), w->end());
      bf.insert(bf.end(), b->begin(), b->end());
    }
    if (!upload_act(e, *wt, &e->w_text, &err) || !upload_f32(e, *bt, &e->b_text, &err) ||
        !upload_act(e, wf, &e->w_film, &err) || !upload_f32(e, bf, &e->b_film, &err))
      return fail(err);
    bool ok = true;
    ok &= (e->task_emb_act = e->dalloc<act_t>((size_t)c.max_batch * E)) != nullptr;
    ok &= (e->task_tok_act = e->dalloc<act_t>((size_t)c.max_batch * D)) != nullptr;
    ok &= (e->task_tok = e->dalloc<float>((size_t)c.max_batch * D)) != nullptr;
    ok &= (e->film = e->dalloc<float>((size_t)c.max_batch * rmb_act_engine::kFilmLd)) != nullptr;
    if (!ok) return fail("cudaMalloc failed (task embedding buffers)");
  }

  // ---- workspace
  {
    const size_t NI = e->n_img_max, H = c.img_h, W = c.img_w;
    const size_t BL = (size_t)c.max_batch * e->L, BC = (size_t)c.max_batch * C;
    bool ok = true;
    ok &= (e->x0 = e->dalloc<act_t>(NI * H * (W + 8) * 4)) != nullptr;
    if (!e->fuse_stem_pool) ok &= (e->c1 = e->dalloc<act_t>(NI * (H / 2) * (W / 2) * 64)) != nullptr;
    for (int li = 0; li < 4; ++li)
      for (int k = 0; k < 3; ++k)
        ok &= (e->stage[li][k] = e->dalloc<act_t>(NI * (H >> (2 + li)) * (W >> (2 + li)) * chans[li])) != nullptr;
    ok &= (e->tok_f32 = e->dalloc<float>(BL * D)) != nullptr;
    ok &= (e->y_f32 = e->dalloc<float>(BL * D)) != nullptr;
    ok &= (e->x1_f32 = e->dalloc<float>(BL * D)) != nullptr;
    ok &= (e->tok_b = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->tok_pb = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->qk = e->dalloc<act_t>(BL * 2 * D)) != nullptr;
    ok &= (e->v = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->ao = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->x1_b = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->hh = e->dalloc<act_t>(BL * FF)) != nullptr;
    ok &= (e->kd = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->vd = e->dalloc<act_t>(BL * D)) != nullptr;
    ok &= (e->aod = e->dalloc<act_t>(BC * D)) != nullptr;
    ok &= (e->t_b = e->dalloc<act_t>(BC * D)) != nullptr;
    ok &= (e->hd = e->dalloc<act_t>(BC * FF)) != nullptr;
    ok &= (e->yd = e->dalloc<float>(BC * D)) != nullptr;
    ok &= (e->t_f32 = e->dalloc<float>(BC * D)) != nullptr;
    ok &= (e->yd2 = e->dalloc<float>(BC * D)) != nullptr;
    if (!ok) return fail("cudaMalloc failed (workspace)");
  }
  e->host_w.clear();
  e->finalized = true;
  RMB_CUDA(cudaDeviceSynchronize());
  return 0;
}

static int build_plans(rmb_act_engine* e, int nb, rmb_act_engine::BatchPlans** out) {
  auto it = e->plans.find(nb);
  if (it != e->plans.end()) { *out = it->second.get(); return 0; }
  std::unique_ptr<rmb_act_engine::BatchPlans> bp(new rmb_act_engine::BatchPlans());
  std::string err;
  const rmb_act_config_t& c = e->cfg;
  const int NI = nb * c.num_cams, H = c.img_h, W = c.img_w, D = 512, FF = c.dim_feedforward, C = c.chunk_size, L = e->L;
  const int chans[4] = {64, 128, 256, 512};
  if (e->fuse_stem_pool) {
    if (!plan_stem_pool(&bp->stem_pool, e->x0, NI, H, W, e->stem.w, e->stem.b, e->stage[0][2], e->num_sms, &err)) return fail(err);
  } else {
    if (!plan_conv1_halo(&bp->stem, e->x0, NI, H, W, e->stem.w, e->stem.b, e->c1, &err)) return fail(err);
  }
  // stage buffers: [0] = t (mid), [1] = y (block 0 out), [2] = d (downsample) then z (block 1 out)
  for (int li = 0; li < 4; ++li) {
    const int h = H >> (2 + li), w = W >> (2 + li);
    const act_t* x_in = (li == 0) ? e->stage[0][2] /* maxpool output lives in stage[0][2] */ : e->stage[li - 1][2];
    act_t *t = e->stage[li][0], *y = e->stage[li][1], *z = e->stage[li][2];
    const int hin = (li == 0) ? h : h * 2, win = (li == 0) ? w : w * 2;
    const int cin = (li == 0) ? 64 : chans[li - 1];
    const void* identity = x_in;
    if (li > 0) {
      if (!e->fuse_down && !plan_conv(&bp->down[li], x_in, NI, hin, win, cin, 1, 2, e->down[li].w, chans[li], e->down[li].b, nullptr, false, z, chans[li], false, false, &err)) return fail(err);
      identity = z;
    }
    if (li == 0) {
      if (!plan_conv3x3_halo(&bp->conv_l1[0][0], x_in, NI, h, w, e->blk[0][0][0].w, e->blk[0][0][0].b, nullptr, t, true, &err)) return fail(err);
      if (!plan_conv3x3_halo(&bp->conv_l1[0][1], t, NI, h, w, e->blk[0][0][1].w, e->blk[0][0][1].b, identity, y, true, &err)) return fail(err);
      if (!plan_conv3x3_halo(&bp->conv_l1[1][0], y, NI, h, w, e->blk[0][1][0].w, e->blk[0][1][0].b, nullptr, t, true, &err)) return fail(err);
      if (!plan_conv3x3_halo(&bp->conv_l1[1][1], t, NI, h, w, e->blk[0][1][1].w, e->blk[0][1][1].b, y, z, true, &err)) return fail(err);
      continue;
    }
    if (li > 0 && e->fuse_down) {   // conv1 + the downsample conv (-> z, no ReLU) in one launch
      if (!plan_conv(&bp->conv[li][0][0], x_in, NI, hin, win, cin, 3, 2, e->c1down[li].w, chans[li], e->c1down[li].b, nullptr, false, t, chans[li], false, true, &err,
                     chans[li], z, chans[li], false)) return fail(err);
    } else
    if (!plan_conv(&bp->conv[li][0][0], x_in, NI, hin, win, cin, 3, (li == 0) ? 1 : 2, e->blk[li][0][0].w, chans[li], e->blk[li][0][0].b, nullptr, false, t, chans[li], false, true, &err)) return fail(err);
    if (!plan_conv(&bp->conv[li][0][1], t, NI, h, w, chans[li], 3, 1, e->blk[li][0][1].w, chans[li], e->blk[li][0][1].b, identity, false, y, chans[li], false, true, &err)) return fail(err);
    if (!plan_conv(&bp->conv[li][1][0], y, NI, h, w, chans[li], 3, 1, e->blk[li][1][0].w, chans[li], e->blk[li][1][0].b, nullptr, false, t, chans[li], false, true, &err)) return fail(err);
    if (!plan_conv(&bp->conv[li][1][1], t, NI, h, w, chans[li], 3, 1, e->blk[li][1][1].w, chans[li], e->blk[li][1][1].b, y, false, z, chans[li], false, true, &err)) return fail(err);
    if (c.multi_task) {
      // per-episode layout of e->film: layer base, then [gamma | beta] x [block 0 | block 1] x planes
      const int base = (li == 1) ? 0 : (li == 2) ? 4 * 128 : 4 * 128 + 4 * 256;
      for (int bi = 0; bi < 2; ++bi) {
        GemmParams& fp = bp->conv[li][bi][1].p;
        fp.film = e->film; fp.film_ld = rmb_act_engine::kFilmLd; fp.film_div = c.num_cams;
        fp.film_g_off = base + (0 * 2 + bi) * chans[li];
        fp.film_b_off = base + (1 * 2 + bi) * chans[li];
      }
    }
  }
  // input_proj: 1x1 conv whose output rows are scattered straight into token order, fp32 residual stream
  if (!plan_conv(&bp->input_proj, e->stage[3][2], NI, e->Hf, e->Wf, 512, 1, 1, e->input_proj.w, D, e->input_proj.b, nullptr, false, e->tok_f32, D, true, false, &err)) return fail(err);
  bp->input_proj.p.remap_ncam = c.num_cams; bp->input_proj.p.remap_L = L; bp->input_proj.p.remap_off = e->n_prefix;

  const long long BL = (long long)nb * L, BC = (long long)nb * C;
  bp->enc.resize(c.enc_layers);
  bp->ln_fused = e->fuse_ln == 2 || (e->fuse_ln == 1 && (BL + 255) / 256 >= e->num_sms / 2);
  for (int l = 0; l < c.enc_layers; ++l) {
    auto& w = e->enc[l];
    auto& pl = bp->enc[l];
    if (bp->ln_fused) {
      if (!plan_linear_ln(&pl.o_ln, e->ao, BL, D, D, w.w_o, w.b_o, e->tok_f32, e->x1_f32, w.g1, w.be1, nullptr, L, e->x1_b, nullptr, 1e-5f, &err)) return fail(err);
      if (!plan_linear_ln(&pl.ff2_ln, e->hh, BL, FF, FF, w.w_2, w.b_2, e->x1_f32, e->tok_f32, w.g2, w.be2, e->pos, L, e->tok_b, e->tok_pb, 1e-5f, &err)) return fail(err);
    }
    if (!plan_linear(&pl.qk, e->tok_pb, BL, D, D, w.w_qk, 2 * D, w.b_qk, nullptr, 0, false, e->qk, 2 * D, false, false, &err)) return fail(err);
    if (!plan_linear(&pl.v, e->tok_b, BL, D, D, w.w_v, D, w.b_v, nullptr, 0, false, e->v, D, false, false, &err)) return fail(err);
    if (!plan_attention(&pl.attn, e->qk, 2 * D, 0, false, e->qk, 2 * D, D, e->v, D, 0, e->ao, D, nb, c.nheads, L, L, &err)) return fail(err);
    const bool r16 = e->res16 && !bp->ln_fused;
    // r16: residual = the 16-bit LayerNorm output, pre-norm sum y written in 16 bits into the same (fp32-sized) buffer
    if (!plan_linear(&pl.o, e->ao, BL, D, D, w.w_o, D, w.b_o, r16 ? (const void*)e->tok_b : (const void*)e->tok_f32, D, !r16, e->y_f32, D, !r16, false, &err)) return fail(err);
    if (!plan_linear(&pl.ff1, e->x1_b, BL, D, D, w.w_1, FF, w.b_1, nullptr, 0, false, e->hh, FF, false, true, &err)) return fail(err);
    if (!plan_linear(&pl.ff2, e->hh, BL, FF, FF, w.w_2, D, w.b_2, r16 ? (const void*)e->x1_b : (const void*)e->x1_f32, D, !r16, e->y_f32, D, !r16, false, &err)) return fail(err);
  }
  auto& d = e->dec;
  if (!plan_linear(&bp->dk, e->tok_pb, BL, D, D, d.w_k, D, d.b_k, nullptr, 0, false, e->kd, D, false, false, &err)) return fail(err);
  if (!plan_linear(&bp->dv, e->tok_b, BL, D, D, d.w_v, D, d.b_v, nullptr, 0, false, e->vd, D, false, false, &err)) return fail(err);
  if (!plan_attention(&bp->dattn, d.q_const, D, 0, true, e->kd, D, 0, e->vd, D, 0, e->aod, D, nb, c.nheads, C, L, &err)) return fail(err);
  if (!plan_linear(&bp->dout, e->aod, BC, D, D, d.w_o, D, d.b_o_fold, nullptr, 0, false, e->yd, D, true, false, &err)) return fail(err);
  if (!plan_linear(&bp->dff1, e->t_b, BC, D, D, d.w_1, FF, d.b_1, nullptr, 0, false, e->hd, FF, false, true, &err)) return fail(err);
  if (!plan_linear(&bp->dff2, e->hd, BC, FF, FF, d.w_2, D, d.b_2, e->t_f32, D, true, e->yd2, D, true, false, &err)) return fail(err);
  *out = bp.get();
  e->plans[nb] = std::move(bp);
  return 0;
}

int rmb_act_forward_num_launches(const rmb_act_engine_t* e) {
  if (!e) return -1;
  // preprocess, stem, maxpool, 16 convs + 3 downsamples, input_proj, extra tokens, cast(+pos), 8 per encoder layer,
  // decoder: K, V, attention, out, LN, FF1, FF2, head
  return (e->fuse_stem_pool ? 2 : 3) + (e->fuse_down ? 16 : 19) + 3 + 8 * e->cfg.enc_layers + 8;
}

double rmb_act_forward_flops(const rmb_act_engine_t* e, int batch) {
  if (!e) return -1.0;
  const rmb_act_config_t& c = e->cfg;
  const double NI = (double)batch * c.num_cams, H = c.img_h, W = c.img_w, D = 512, FF = c.dim_feedforward, L = e->L, C = c.chunk_size;
  double f = 0;
  f += 2.0 * NI * (H / 2) * (W / 2) * 64 * 147;                                       // stem
  const double ch[4] = {64, 128, 256, 512};
  for (int li = 0; li < 4; ++li) {
    const double px = (H / (4 << li)) * (W / (4 << li));
    const double cin = li == 0 ? 64 : ch[li - 1];
    f += 2.0 * NI * px * ch[li] * 9 * cin + 3 * 2.0 * NI * px * ch[li] * 9 * ch[li];
    if (li > 0) f += 2.0 * NI * px * ch[li] * cin;
  }
  f += 2.0 * NI * e->Hf * e->Wf * D * 512;                                            // input_proj
  const double BL = (double)batch * L;
  f += c.enc_layers * (2.0 * BL * D * (4 * D + 2 * FF) + 4.0 * batch * L * L * D);    // QKV+O, FFN, QK^T + PV
  f += 2.0 * BL * D * 2 * D + 4.0 * batch * C * L * D + 2.0 * batch * C * D * (D + 2 * FF) + 2.0 * batch * C * D * c.action_dim;
  return f;
}

// ---------------------------------------------------------------------------------------------------------------------
// Debug output for RolloutAct.draw_plot (robo_manip_baselines/policy/act/RolloutAct.py:36-43,115-125): the reference
// hooks every encoder layer's nn.MultiheadAttention and keeps `attn_output_weights[0]`, the head-AVERAGED softmax map
// (L, L) of episode 0 (need_weights=True, average_attn_weights=True). The engine's attention kernel never materialises
// the map; with the debug switch on, this kernel recomputes it for ONE episode from the layer's Q / K projections
// (one CTA per query row, warp h = head h). Off in the timed path.
__global__ void attn_map_kernel(const act_t* __restrict__ qk, int ld, int L, float scale, float* __restrict__ out) {
  extern __shared__ float am_smem[];
  float* q = am_smem;                 // [512]
  float* sc = am_smem + 512;          // [8][L]
  const int i = blockIdx.x;
  const act_t* qrow = qk + static_cast<long long>(i) * ld;
  for (int c = threadIdx.x; c < 512; c += blockDim.x) q[c] = static_cast<float>(qrow[c]);
  __syncthreads();
  for (int idx = threadIdx.x; idx < 8 * L; idx += blockDim.x) {
    const int h = idx / L, j = idx - h * L;
    const act_t* krow = qk + static_cast<long long>(j) * ld + 512 + h * 64;
    float acc = 0.f;
#pragma unroll 8
    for (int d = 0; d < 64; ++d) acc += q[h * 64 + d] * static_cast<float>(krow[d]);
    sc[idx] = acc * scale;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 8) {
    float* s = sc + warp * L;
    float mx = -INFINITY;
    for (int j = lane; j < L; j += 32) mx = fmaxf(mx, s[j]);
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int j = lane; j < L; j += 32) { const float ev = expf(s[j] - mx); s[j] = ev; sum += ev; }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.0f / sum;
    for (int j = lane; j < L; j += 32) s[j] *= inv;
  }
  __syncthreads();
  for (int j = threadIdx.x; j < L; j += blockDim.x) {
    float acc = 0.f;
#pragma unroll
    for (int h = 0; h < 8; ++h) acc += sc[h * L + j];
    out[static_cast<long long>(i) * L + j] = acc * 0.125f;
  }
}

int rmb_act_set_attention_debug(rmb_act_engine_t* e, int on, int episode) {
  if (!e) return fail("rmb_act_set_attention_debug: null engine");
  if (!e->finalized) return fail("rmb_act_set_attention_debug: weights not finalized");
  if (episode < 0 || episode >= e->cfg.max_batch) return fail("rmb_act_set_attention_debug: episode out of range");
  if (on && e->attn_maps == nullptr) {
    RMB_CUDA(cudaSetDevice(e->cfg.device));
    e->attn_maps = e->dalloc<float>((size_t)e->cfg.enc_layers * e->L * e->L);
    if (!e->attn_maps) return fail("rmb_act_set_attention_debug: cudaMalloc failed");
    RMB_CUDA(cudaMemset(e->attn_maps, 0, (size_t)e->cfg.enc_layers * e->L * e->L * sizeof(float)));
  }
  e->attn_debug = on != 0;
  e->attn_debug_episode = episode;
  return 0;
}

int64_t rmb_act_read_attention_maps(rmb_act_engine_t* e, float* d_out, void* stream_) {
  if (!e || !d_out) return fail("rmb_act_read_attention_maps: null argument");
  if (e->attn_maps == nullptr) return fail("rmb_act_read_attention_maps: attention debug output was never switched on");
  const int64_t n = (int64_t)e->cfg.enc_layers * e->L * e->L;
  RMB_CUDA(cudaMemcpyAsync(d_out, e->attn_maps, n * sizeof(float), cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream_)));
  return n;
}

// ---------------------------------------------------------------------------------------------------------------------
// Saturation diagnostics. fp16 conversions on this path clamp to +-65504 (cvt.satfinite) instead of producing inf, so
// a checkpoint whose activations leave the fp16 range would be silently wrong. In check mode every 16-bit tensor that
// a forward stores is scanned right after the launch that wrote it, and elements at the clamp value (fp16: |v| ==
// 65504; bf16: inf / nan, it has fp32's range) are counted into a device counter read by rmb_saturation_count().
// Off by default: the timed path launches nothing extra.
__device__ unsigned long long g_rmb_sat_count = 0ull;
__global__ void sat_scan_kernel(const uint16_t* __restrict__ p, long long n) {
  unsigned int local = 0;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint16_t v = p[i];
#ifdef RMB_ACT_BF16
    local += ((v & 0x7f80u) == 0x7f80u) ? 1u : 0u;
#else
    local += ((v & 0x7fffu) >= 0x7bffu) ? 1u : 0u;
#endif
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0 && local != 0) atomicAdd(&g_rmb_sat_count, static_cast<unsigned long long>(local));
}

int rmb_act_set_check_saturation(rmb_act_engine_t* e, int on) {
  if (!e) return fail("rmb_act_set_check_saturation: null engine");
  e->check_saturation = on != 0;
  return 0;
}

int64_t rmb_saturation_count(int clear) {
  unsigned long long v = 0;
  RMB_CUDA(cudaDeviceSynchronize());
  RMB_CUDA(cudaMemcpyFromSymbol(&v, g_rmb_sat_count, sizeof(v)));
  if (clear) {
    const unsigned long long zero = 0;
    RMB_CUDA(cudaMemcpyToSymbol(g_rmb_sat_count, &zero, sizeof(zero)));
  }
  return static_cast<int64_t>(v);
}

// One record per kernel launch of a forward (filled by rmb_act_forward_profile).
struct LaunchProf {
  std::vector<cudaEvent_t> ev;      // 2 per launch
  std::vector<rmb_launch_prof_t> rec;
};

static int forward_impl(rmb_act_engine_t* e, const void* d_images, int image_format, const float* d_qpos, float* d_a_hat,
                        int batch, cudaStream_t s, LaunchProf* prof) {
  if (!e || !d_images || !d_qpos || !d_a_hat) return fail("rmb_act_forward: null argument");
  if (!e->finalized) return fail("rmb_act_forward: weights not finalized");
  if (batch < 1 || batch > e->cfg.max_batch) return fail("rmb_act_forward: batch out of range");
  if (image_format != 0 && image_format != 1) return fail("rmb_act_forward: image_format must be 0 (u8 HWC) or 1 (f32 CHW)");
  if (e->cfg.multi_task && e->task_batch < batch)
    return fail("rmb_act_forward: multi-task engine without a task embedding for every episode (call rmb_act_set_task_embedding)");
  PdlScope pdl;   // every launch below is a pdl_wait()-aware kernel (launch.h)
  rmb_act_engine::BatchPlans* bp = nullptr;
  if (int r = build_plans(e, batch, &bp)) return r;
  const rmb_act_config_t& c = e->cfg;
  const int NI = batch * c.num_cams, H = c.img_h, W = c.img_w, L = e->L, C = c.chunk_size, FF = c.dim_feedforward;
  const int sms = e->num_sms;
  const float3 mean = make_float3(0.485f, 0.456f, 0.406f);
  const float3 istd = make_float3(1.0f / 0.229f, 1.0f / 0.224f, 1.0f / 0.225f);
  const long long BL = (long long)batch * L, BC = (long long)batch * C;

  // begin/end bracket every launch with CUDA events when profiling (no-ops otherwise)
  auto begin = [&](const char* name, const char* kernel, double flops, double bytes) {
    if (!prof) return;
    rmb_launch_prof_t r;
    memset(&r, 0, sizeof(r));
    snprintf(r.name, sizeof(r.name), "%s", name);
    snprintf(r.kernel, sizeof(r.kernel), "%s", kernel);
    r.flops = flops; r.bytes = bytes;
    prof->rec.push_back(r);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    prof->ev.push_back(a); prof->ev.push_back(b);
    cudaEventRecord(a, s);
  };
  auto end = [&]() { if (prof) cudaEventRecord(prof->ev.back(), s); };
  // algorithmic HBM bytes of an implicit GEMM: A read once, weights once, output written once (+ residual read)
  auto gemm_bytes = [&](const GemmPlan& g, double a_bytes) {
    const double rows = (double)g.p.W * g.p.H * g.p.B, N = (double)g.p.n_col_tiles * g.block_n;
    const double K = (double)g.p.num_taps * g.p.chunks_per_tap * g.block_k;
    const double ln_extra = (g.p.ln_y_act ? rows * N * 2 : 0) + (g.p.ln_ypos_act ? rows * N * 2 : 0);
    return a_bytes + N * K * 2 + rows * N * (g.p.out_f32 ? 4 : 2) + (g.p.residual ? rows * N * (g.p.res_f32 ? 4 : 2) : 0) + ln_extra;
  };
  // check mode: count clamped values in the 16-bit tensor the launch just stored
  auto scan = [&](const void* ptr, long long n_elems) {
    if (!e->check_saturation || n_elems <= 0) return;
    sat_scan_kernel<<<grid_for(n_elems, 256, sms), 256, 0, s>>>(static_cast<const uint16_t*>(ptr), n_elems);
  };
  auto run_halo = [&](const char* name, const HaloPlan& g, double a_bytes) -> int {
    const double rows = (double)g.hp.g.W * g.hp.g.H * g.hp.g.B;
    const double K = (double)g.hp.n_taps * g.block_k;
    begin(name, "conv_halo_kernel", (double)g.flops, a_bytes + 64 * K * 2 + rows * 64 * 2 + (g.hp.g.residual ? rows * 64 * 2 : 0));
    RMB_CUDA(launch_halo(g, sms, s));
    end();
    scan(g.hp.g.out, (long long)rows * 64);
    return 0;
  };
  auto run_gemm = [&](const char* name, const GemmPlan& g, double a_bytes) -> int {
    begin(name, g.ln ? "gemm_tc_pair_kernel<LN>" : g.pair ? "gemm_tc_pair_kernel" : "gemm_tc_kernel", (double)g.flops, gemm_bytes(g, a_bytes));
    RMB_CUDA(launch_gemm(g, sms, s));
    end();
    if (!g.p.out_f32 && g.p.remap_ncam == 0) scan(g.p.out, (long long)g.p.W * g.p.H * g.p.B * g.p.ld_out);
    if (g.p.n_main_ct > 0) scan(g.p.out2, (long long)g.p.W * g.p.H * g.p.B * g.p.ld_out2);
    return 0;
  };
#define RMB_RUN(x) do { if (int _r = (x)) return _r; } while (0)

  {
    const long long items = (long long)NI * H * (W / 4 + 2);
    const double out_b = (double)NI * H * (W + 8) * 8;
    if (image_format == 0) {
      begin("preprocess_u8", "preprocess_u8_kernel", 0, (double)NI * H * W * 3 + out_b);
      RMB_CUDA(launch_pdl(preprocess_u8_kernel, dim3(grid_for(items, 256, sms)), dim3(256), 0, s, static_cast<const uint8_t*>(d_images), e->x0, NI, H, W, mean, istd));
    } else {
      begin("preprocess_f32", "preprocess_f32_kernel", 0, (double)NI * H * W * 12 + out_b);
      RMB_CUDA(launch_pdl(preprocess_f32_kernel, dim3(grid_for(items, 256, sms)), dim3(256), 0, s, static_cast<const float*>(d_images), e->x0, NI, H, W, mean, istd));
    }
    end();
    scan(e->x0, (long long)NI * H * (W + 8) * 4);
  }
  if (e->fuse_stem_pool) {
    // algorithmic bytes: the padded 16-bit image in, the pooled map out (the conv output never reaches HBM)
    begin("stem conv7x7/2 + maxpool3x3/2", "stem_pool_kernel", (double)bp->stem_pool.flops,
          (double)NI * H * (W + 8) * 8 + 64 * 224 * 2 + (double)NI * (H / 4) * (W / 4) * 128);
    RMB_CUDA(launch_stem_pool(bp->stem_pool, s));
    end();
    scan(e->stage[0][2], (long long)NI * (H / 4) * (W / 4) * 64);
  } else {
    RMB_RUN(run_halo("stem conv7x7/2", bp->stem, (double)NI * H * (W + 8) * 8));
    const long long items = (long long)NI * (H / 4) * (W / 4) * 8;
    begin("maxpool3x3/2", "maxpool3x3s2_kernel", 0, (double)NI * (H / 2) * (W / 2) * 128 + (double)NI * (H / 4) * (W / 4) * 128);
    RMB_CUDA(launch_pdl(maxpool3x3s2_kernel, dim3(grid_for(items, 256, sms)), dim3(256), 0, s, (const act_t*)e->c1, e->stage[0][2], NI, H / 2, W / 2, 64));
    end();
  }
  const int chans[4] = {64, 128, 256, 512};
  for (int li = 0; li < 4; ++li) {
    char nm[48];
    const double px = (double)NI * (H >> (2 + li)) * (W >> (2 + li));
    const double cin = li == 0 ? 64 : chans[li - 1];
    const double in_b = (li == 0 ? px : px * 4) * cin * 2, mid_b = px * chans[li] * 2;
    if (li > 0 && !e->fuse_down) { snprintf(nm, sizeof(nm), "layer%d.0.downsample 1x1/2", li + 1); RMB_RUN(run_gemm(nm, bp->down[li], in_b)); }
    if (li == 0) {
      RMB_RUN(run_halo("layer1.0.conv1 3x3", bp->conv_l1[0][0], in_b));
      RMB_RUN(run_halo("layer1.0.conv2 3x3", bp->conv_l1[0][1], mid_b));
      RMB_RUN(run_halo("layer1.1.conv1 3x3", bp->conv_l1[1][0], mid_b));
      RMB_RUN(run_halo("layer1.1.conv2 3x3", bp->conv_l1[1][1], mid_b));
      continue;
    }
    snprintf(nm, sizeof(nm), (li > 0 && e->fuse_down) ? "layer%d.0.conv1 3x3 + downsample" : "layer%d.0.conv1 3x3", li + 1);
    RMB_RUN(run_gemm(nm, bp->conv[li][0][0], in_b));
    snprintf(nm, sizeof(nm), "layer%d.0.conv2 3x3", li + 1); RMB_RUN(run_gemm(nm, bp->conv[li][0][1], mid_b));
    snprintf(nm, sizeof(nm), "layer%d.1.conv1 3x3", li + 1); RMB_RUN(run_gemm(nm, bp->conv[li][1][0], mid_b));
    snprintf(nm, sizeof(nm), "layer%d.1.conv2 3x3", li + 1); RMB_RUN(run_gemm(nm, bp->conv[li][1][1], mid_b));
  }
  RMB_RUN(run_gemm("input_proj 1x1 -> tokens", bp->input_proj, (double)NI * e->Hf * e->Wf * 512 * 2));
  begin("latent+proprio tokens", "act_extra_tokens_kernel", 0, (double)batch * e->n_prefix * 512 * 4);
  RMB_CUDA(launch_pdl(act_extra_tokens_kernel, dim3(batch), dim3(256), 0, s, d_qpos, c.state_dim, (const float*)e->latent_bias, (const float*)e->w_state, (const float*)e->b_state, (const float*)(c.task_token ? e->task_tok : nullptr), e->tok_f32, L));
  end();
  const int ln_grid = grid_for(BL * 32, 256, sms);
  begin("tokens cast(+pos)", "layernorm512_kernel", 0, (double)BL * 512 * (4 + 2 + 2));
  RMB_CUDA(launch_pdl(layernorm512_kernel, dim3(ln_grid), dim3(256), 0, s, e->tok_f32, nullptr, nullptr, e->pos, L, nullptr, e->tok_b, e->tok_pb, BL, 0, 1e-5f));
  end();
  const double tokb = (double)BL * 512 * 2;
  const double attn_flops = 4.0 * batch * (double)L * L * 512, attn_bytes = (double)BL * 512 * 2 * 4;
  for (int l = 0; l < c.enc_layers; ++l) {
    auto& w = e->enc[l];
    auto& pl = bp->enc[l];
    char nm[48];
    snprintf(nm, sizeof(nm), "enc%d QK proj", l); RMB_RUN(run_gemm(nm, pl.qk, tokb));
    if (e->attn_debug && e->attn_debug_episode < batch) {
      const size_t smem = (512 + 8 * (size_t)L) * sizeof(float);
      static bool am_attr = false;
      if (!am_attr) { RMB_CUDA(cudaFuncSetAttribute(attn_map_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); am_attr = true; }
      if (smem > 200 * 1024) return fail("attention debug output: sequence too long for the debug kernel");
      attn_map_kernel<<<L, 256, smem, s>>>(e->qk + (size_t)e->attn_debug_episode * L * 1024, 1024, L, 0.125f,
                                           e->attn_maps + (size_t)l * L * L);
    }
    snprintf(nm, sizeof(nm), "enc%d V proj", l); RMB_RUN(run_gemm(nm, pl.v, tokb));
    snprintf(nm, sizeof(nm), "enc%d self-attention", l);
    begin(nm, attention_kernel_name(), attn_flops, attn_bytes);
    RMB_CUDA(launch_attention(pl.attn, s));
    end();
    if (bp->ln_fused) {
      snprintf(nm, sizeof(nm), "enc%d out proj +res +norm1", l); RMB_RUN(run_gemm(nm, pl.o_ln, tokb));
      snprintf(nm, sizeof(nm), "enc%d linear1+relu", l); RMB_RUN(run_gemm(nm, pl.ff1, tokb));
      snprintf(nm, sizeof(nm), "enc%d linear2 +res +norm2", l); RMB_RUN(run_gemm(nm, pl.ff2_ln, (double)BL * FF * 2));
      continue;
    }
    snprintf(nm, sizeof(nm), "enc%d out proj +res", l); RMB_RUN(run_gemm(nm, pl.o, tokb));   // y = ao Wo^T + bo + x
    snprintf(nm, sizeof(nm), "enc%d norm1", l);
    begin(nm, "layernorm512_kernel", 0, (double)BL * 512 * (e->res16 ? 2 + 2 : 4 + 4 + 2));
    if (e->res16) RMB_CUDA(launch_pdl(layernorm512_kernel_t<act_t>, dim3(ln_grid), dim3(256), 0, s, (const act_t*)e->y_f32, w.g1, w.be1, nullptr, L, (float*)nullptr, e->x1_b, nullptr, BL, 1, 1e-5f));
    else RMB_CUDA(launch_pdl(layernorm512_kernel, dim3(ln_grid), dim3(256), 0, s, e->y_f32, w.g1, w.be1, nullptr, L, e->x1_f32, e->x1_b, nullptr, BL, 1, 1e-5f));
    end();
    snprintf(nm, sizeof(nm), "enc%d linear1+relu", l); RMB_RUN(run_gemm(nm, pl.ff1, tokb));
    snprintf(nm, sizeof(nm), "enc%d linear2 +res", l); RMB_RUN(run_gemm(nm, pl.ff2, (double)BL * FF * 2));   // y = relu(.) W2^T + b2 + x1
    snprintf(nm, sizeof(nm), "enc%d norm2", l);
    // the last layer's fp32 output is the encoder memory read by rmb_act_debug_read
    const bool f32_out = !e->res16 || l == c.enc_layers - 1;
    begin(nm, "layernorm512_kernel", 0, (double)BL * 512 * ((e->res16 ? 2 : 4) + (f32_out ? 4 : 0) + 2 + 2));
    if (e->res16) RMB_CUDA(launch_pdl(layernorm512_kernel_t<act_t>, dim3(ln_grid), dim3(256), 0, s, (const act_t*)e->y_f32, w.g2, w.be2, e->pos, L, f32_out ? e->tok_f32 : (float*)nullptr, e->tok_b, e->tok_pb, BL, 1, 1e-5f));
    else RMB_CUDA(launch_pdl(layernorm512_kernel, dim3(ln_grid), dim3(256), 0, s, e->y_f32, w.g2, w.be2, e->pos, L, e->tok_f32, e->tok_b, e->tok_pb, BL, 1, 1e-5f));
    end();
  }
  RMB_RUN(run_gemm("dec0 K proj", bp->dk, tokb));
  RMB_RUN(run_gemm("dec0 V proj", bp->dv, tokb));
  begin("dec0 cross-attention", attention_kernel_name(), 4.0 * batch * (double)C * L * 512, (double)BL * 512 * 2 * 2 + (double)BC * 512 * 2);
  RMB_CUDA(launch_attention(bp->dattn, s));
  end();
  RMB_RUN(run_gemm("dec0 out proj +tgt", bp->dout, (double)BC * 512 * 2));
  const int lnd_grid = grid_for(BC * 32, 256, sms);
  begin("dec0 norm2", "layernorm512_kernel", 0, (double)BC * 512 * (4 + 4 + 2));
  RMB_CUDA(launch_pdl(layernorm512_kernel, dim3(lnd_grid), dim3(256), 0, s, e->yd, e->dec.g2, e->dec.be2, nullptr, C, e->t_f32, e->t_b, nullptr, BC, 1, 1e-5f));
  end();
  RMB_RUN(run_gemm("dec0 linear1+relu", bp->dff1, (double)BC * 512 * 2));
  RMB_RUN(run_gemm("dec0 linear2 +res", bp->dff2, (double)BC * FF * 2));
  begin("norm3+decoder.norm+action_head", "act_head_kernel", 2.0 * BC * 512 * c.action_dim, (double)BC * 512 * 4 + (double)BC * c.action_dim * 4);
  RMB_CUDA(launch_pdl(act_head_kernel, dim3(lnd_grid), dim3(256), 0, s, e->yd2, e->dec.g3, e->dec.be3, e->dec.gn, e->dec.ben, e->w_head, e->b_head, d_a_hat, BC, c.action_dim, 1e-5f));
  end();
#undef RMB_RUN
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_act_set_task_embedding(rmb_act_engine_t* e, const float* d_task_emb, int batch, void* stream_) {
  if (!e || !d_task_emb) return fail("rmb_act_set_task_embedding: null argument");
  if (!e->finalized) return fail("rmb_act_set_task_embedding: weights not finalized");
  if (!e->cfg.multi_task) return fail("rmb_act_set_task_embedding: engine was not created with multi_task");
  if (batch < 1 || batch > e->cfg.max_batch) return fail("rmb_act_set_task_embedding: batch out of range");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const int E = e->cfg.task_emb_dim, D = 512, F = rmb_act_engine::kFilmLd;
  std::string err;
  GemmPlan proj, film;
  if (!plan_linear(&proj, e->task_emb_act, batch, E, E, e->w_text, D, e->b_text, nullptr, 0, false, e->task_tok, D, true, false, &err)) return fail(err);
  if (!plan_linear(&film, e->task_tok_act, batch, D, D, e->w_film, F, e->b_film, nullptr, 0, false, e->film, F, true, false, &err)) return fail(err);
  cast_f32_to_act_kernel<<<grid_for((long long)batch * E, 256, e->num_sms), 256, 0, s>>>(d_task_emb, e->task_emb_act, (long long)batch * E);
  RMB_CUDA(launch_gemm(proj, e->num_sms, s));
  cast_f32_to_act_kernel<<<grid_for((long long)batch * D, 256, e->num_sms), 256, 0, s>>>(e->task_tok, e->task_tok_act, (long long)batch * D);
  RMB_CUDA(launch_gemm(film, e->num_sms, s));
  RMB_CUDA(cudaGetLastError());
  e->task_batch = batch;
  return 0;
}

int rmb_act_forward(rmb_act_engine_t* e, const void* d_images, int image_format, const float* d_qpos, float* d_a_hat,
                    int batch, void* stream_) {
  return forward_impl(e, d_images, image_format, d_qpos, d_a_hat, batch, static_cast<cudaStream_t>(stream_), nullptr);
}

int rmb_act_forward_profile(rmb_act_engine_t* e, const void* d_images, int image_format, const float* d_qpos,
                            float* d_a_hat, int batch, void* stream_, rmb_launch_prof_t* out, int cap) {
  if (!out || cap < 1) return fail("rmb_act_forward_profile: null output");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  LaunchProf prof;
  const int r = forward_impl(e, d_images, image_format, d_qpos, d_a_hat, batch, s, &prof);
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
  if (ce != cudaSuccess) return fail(std::string("rmb_act_forward_profile: ") + cudaGetErrorString(ce), -2);
  return n;
}

int64_t rmb_act_debug_read(rmb_act_engine_t* e, int which, float* d_out, int batch, void* stream_) {
  if (!e || !d_out) return fail("rmb_act_debug_read: null argument");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const int64_t n = (int64_t)batch * e->L * 512;
  if (which == 1) {
    RMB_CUDA(cudaMemcpyAsync(d_out, e->tok_f32, n * sizeof(float), cudaMemcpyDeviceToDevice, s));
    return n;
  }
  return fail("rmb_act_debug_read: unknown tap");
}

// ---------------------------------------------------------------------------------------------------------------------
int rmb_ensemble_step(float* d_ring, const float* d_chunks, int32_t* d_step, const uint8_t* d_reset,
                      const double* d_wtab, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                      double* d_action, float* d_action_f32, double* d_action_norm, int batch, int C, int Da,
                      void* stream_) {
  if (!d_ring || !d_chunks || !d_step || !d_wtab || !d_p0 || !d_p1) return fail("rmb_ensemble_step: null argument");
  if (batch < 1 || C < 1 || Da < 1 || Da > 128) return fail("rmb_ensemble_step: bad sizes");
  if (norm_type != 0 && norm_type != 1) return fail("rmb_ensemble_step: norm_type must be 0 (gaussian) or 1 (limits)");
  DenormParams dn{d_p0, d_p1, out_min, norm_type};
  PdlScope pdl;
  const size_t diag_bytes = static_cast<size_t>(C) * Da * sizeof(float);
  const int use_smem = diag_bytes <= 48 * 1024 ? 1 : 0;
  RMB_CUDA(launch_pdl(ensemble_step_kernel, dim3(batch), dim3(128), use_smem ? diag_bytes : 0, static_cast<cudaStream_t>(stream_), d_ring, d_chunks, d_step, d_reset, d_wtab, dn, d_action, d_action_f32, d_action_norm, C, Da, use_smem));
  return 0;
}

int rmb_normalize(const double* d_x, int norm_type, const double* d_p0, const double* d_p1, double out_min, float* d_y,
                  int64_t rows, int D, void* stream_) {
  if (!d_x || !d_p0 || !d_p1 || !d_y) return fail("rmb_normalize: null argument");
  if (norm_type != 0 && norm_type != 1) return fail("rmb_normalize: norm_type must be 0 or 1");
  if (rows * D == 0) return 0;
  PdlScope pdl;
  RMB_CUDA(launch_pdl(normalize_state_kernel, dim3(grid_for(rows * D, 256, 148)), dim3(256), 0, static_cast<cudaStream_t>(stream_), d_x, d_p0, d_p1, out_min, norm_type, d_y, rows * D, D));
  return 0;
}

int rmb_denormalize(const double* d_x, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                    double* d_y, int64_t rows, int D, void* stream_) {
  if (!d_x || !d_p0 || !d_p1 || !d_y) return fail("rmb_denormalize: null argument");
  if (norm_type != 0 && norm_type != 1) return fail("rmb_denormalize: norm_type must be 0 or 1");
  if (rows * D == 0) return 0;
  DenormParams dn{d_p0, d_p1, out_min, norm_type};
  denormalize_kernel<<<grid_for(rows * D, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(d_x, dn, d_y, rows * D, D);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_preprocess_images_u8(const uint8_t* d_images, void* d_out, int n_img, int H, int W, void* stream_) {
  if (!d_images || !d_out) return fail("rmb_preprocess_images_u8: null argument");
  if (W % 4 != 0) return fail("rmb_preprocess_images_u8: W must be a multiple of 4");
  const long long items = (long long)n_img * H * (W / 4 + 2);
  preprocess_u8_kernel<<<grid_for(items, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_images, static_cast<act_t*>(d_out), n_img, H, W, make_float3(0.485f, 0.456f, 0.406f),
      make_float3(1.0f / 0.229f, 1.0f / 0.224f, 1.0f / 0.225f));
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_preprocess_images_f32(const float* d_images, void* d_out, int n_img, int H, int W, void* stream_) {
  if (!d_images || !d_out) return fail("rmb_preprocess_images_f32: null argument");
  if (W % 4 != 0) return fail("rmb_preprocess_images_f32: W must be a multiple of 4");
  const long long items = (long long)n_img * H * (W / 4 + 2);
  preprocess_f32_kernel<<<grid_for(items, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_images, static_cast<act_t*>(d_out), n_img, H, W, make_float3(0.485f, 0.456f, 0.406f),
      make_float3(1.0f / 0.229f, 1.0f / 0.224f, 1.0f / 0.225f));
  RMB_CUDA(cudaGetLastError());
  return 0;
}

static int device_sms() {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

int rmb_gemm(const void* d_A, const void* d_Wt, const float* d_bias, const void* d_residual, void* d_out,
                  int64_t M, int N, int K, int flags, void* stream_) {
  if (!d_A || !d_Wt || !d_out) return fail("rmb_gemm: null argument");
  GemmPlan g;
  std::string err;
  if (!plan_linear(&g, d_A, M, K, K, d_Wt, N, d_bias, d_residual, N, (flags & 4) != 0, d_out, N, (flags & 2) != 0, (flags & 1) != 0, &err)) return fail(err);
  RMB_CUDA(launch_gemm(g, device_sms(), static_cast<cudaStream_t>(stream_)));
  return 0;
}

int rmb_gemm_residual_layernorm(const void* d_A, const void* d_Wt, const float* d_bias, const float* d_residual,
                                const float* d_gamma, const float* d_beta, const float* d_pos, int L, float* d_out,
                                void* d_y_act, void* d_ypos_act, int64_t M, int K, void* stream_) {
  if (!d_A || !d_Wt) return fail("rmb_gemm_residual_layernorm: null argument");
  GemmPlan g;
  std::string err;
  if (!plan_linear_ln(&g, d_A, M, K, K, d_Wt, d_bias, d_residual, d_out, d_gamma, d_beta, d_pos, L, d_y_act, d_ypos_act, 1e-5f, &err)) return fail(err);
  RMB_CUDA(launch_gemm(g, device_sms(), static_cast<cudaStream_t>(stream_)));
  return 0;
}

int rmb_conv2d_nhwc(const void* d_in, const void* d_Wt, const float* d_bias, const void* d_residual, void* d_out,
                         int B, int H, int W, int Cin, int Cout, int ksize, int stride, int relu, void* stream_) {
  if (!d_in || !d_Wt || !d_out) return fail("rmb_conv2d_nhwc: null argument");
  std::string err;
  if (conv3x3_halo_eligible(W, Cin, Cout, ksize, stride)) {
    HaloPlan hg;
    if (!plan_conv3x3_halo(&hg, d_in, B, H, W, d_Wt, d_bias, d_residual, d_out, relu != 0, &err)) return fail(err);
    RMB_CUDA(launch_halo(hg, device_sms(), static_cast<cudaStream_t>(stream_)));
    return 0;
  }
  GemmPlan g;
  if (!plan_conv(&g, d_in, B, H, W, Cin, ksize, stride, d_Wt, Cout, d_bias, d_residual, false, d_out, Cout, false, relu != 0, &err)) return fail(err);
  RMB_CUDA(launch_gemm(g, device_sms(), static_cast<cudaStream_t>(stream_)));
  return 0;
}

int rmb_conv1_stem(const void* d_in_padded, const void* d_Wt, const float* d_bias, void* d_out, int B, int H, int W,
                        void* stream_) {
  if (!d_in_padded || !d_Wt || !d_out) return fail("rmb_conv1_stem: null argument");
  HaloPlan g;
  std::string err;
  if (!plan_conv1_halo(&g, d_in_padded, B, H, W, d_Wt, d_bias, d_out, &err)) return fail(err);
  RMB_CUDA(launch_halo(g, device_sms(), static_cast<cudaStream_t>(stream_)));
  return 0;
}

int rmb_stem_pool(const void* d_in_padded, const void* d_Wt, const float* d_bias, void* d_out, int B, int H, int W,
                  void* stream_) {
  if (!d_in_padded || !d_Wt || !d_out) return fail("rmb_stem_pool: null argument");
  StemPoolPlan g;
  std::string err;
  if (!plan_stem_pool(&g, d_in_padded, B, H, W, d_Wt, d_bias, d_out, device_sms(), &err)) return fail(err);
  RMB_CUDA(launch_stem_pool(g, static_cast<cudaStream_t>(stream_)));
  return 0;
}

int rmb_maxpool3x3s2_nhwc(const void* d_in, void* d_out, int B, int H, int W, int C, void* stream_) {
  if (!d_in || !d_out) return fail("rmb_maxpool3x3s2_nhwc: null argument");
  if (C % 8 != 0) return fail("rmb_maxpool3x3s2_nhwc: C must be a multiple of 8");
  const long long items = (long long)B * ((H + 1) / 2) * ((W + 1) / 2) * (C / 8);
  maxpool3x3s2_kernel<<<grid_for(items, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const act_t*>(d_in), static_cast<act_t*>(d_out), B, H, W, C);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_attention(const void* d_Q, int ldq, int q_col0, int q_shared, const void* d_K, int ldk, int k_col0,
                       const void* d_V, int ldv, int v_col0, void* d_out, int ld_out, int B, int n_heads, int Lq, int Lk,
                       void* stream_) {
  if (!d_Q || !d_K || !d_V || !d_out) return fail("rmb_attention: null argument");
  AttnPlan a;
  std::string err;
  if (!plan_attention(&a, d_Q, ldq, q_col0, q_shared != 0, d_K, ldk, k_col0, d_V, ldv, v_col0, d_out, ld_out, B, n_heads, Lq, Lk, &err)) return fail(err);
  RMB_CUDA(launch_attention(a, static_cast<cudaStream_t>(stream_)));
  return 0;
}

#ifdef RMB_ATTN_TRACE
// debug builds only (scripts/attn_trace.py): copies the (event, clock) pairs logged by attention_v3_kernel and resets the log
extern "C" int rmb_attention_trace_dump(long long* h_words, int max_words) {
  RMB_CUDA(cudaDeviceSynchronize());
  const int n = max_words < 2 * 16384 ? max_words : 2 * 16384;
  RMB_CUDA(cudaMemcpyFromSymbol(h_words, rmb::g_attn_trace, sizeof(long long) * n));
  static long long zeros[2 * 16384] = {};
  RMB_CUDA(cudaMemcpyToSymbol(rmb::g_attn_trace, zeros, sizeof(zeros)));
  return n;
}
#endif

int rmb_set_attention_variant(int variant) {
  if (variant == 0) return g_attn_variant;
  if (variant < 1 || variant > 4) return fail("rmb_set_attention_variant: variant must be 1, 2, 3 or 4 (0 = query)");
  g_attn_variant = variant;
  return variant;
}

int rmb_layernorm512(const float* d_x, const float* d_gamma, const float* d_beta, const float* d_pos, int L,
                     float* d_y_f32, void* d_y_act, void* d_ypos_act, int64_t rows, int do_norm, void* stream_) {
  if (!d_x) return fail("rmb_layernorm512: null argument");
  if (do_norm && (!d_gamma || !d_beta)) return fail("rmb_layernorm512: gamma/beta required");
  if (d_ypos_act && (!d_pos || L < 1)) return fail("rmb_layernorm512: pos table required");
  layernorm512_kernel<<<grid_for(rows * 32, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_x, d_gamma, d_beta, d_pos, L < 1 ? 1 : L, d_y_f32, static_cast<act_t*>(d_y_act),
      static_cast<act_t*>(d_ypos_act), rows, do_norm, 1e-5f);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_layernorm512_act(const void* d_x_act, const float* d_gamma, const float* d_beta, const float* d_pos, int L,
                         float* d_y_f32, void* d_y_act, void* d_ypos_act, int64_t rows, int do_norm, void* stream_) {
  if (!d_x_act) return fail("rmb_layernorm512_act: null argument");
  if (do_norm && (!d_gamma || !d_beta)) return fail("rmb_layernorm512_act: gamma/beta required");
  if (d_ypos_act && (!d_pos || L < 1)) return fail("rmb_layernorm512_act: pos table required");
  layernorm512_kernel_t<act_t><<<grid_for(rows * 32, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      static_cast<const act_t*>(d_x_act), d_gamma, d_beta, d_pos, L < 1 ? 1 : L, d_y_f32, static_cast<act_t*>(d_y_act),
      static_cast<act_t*>(d_ypos_act), rows, do_norm, 1e-5f);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"

#include "diffusion_engine.cuh"
#include "imgenc_engine.cuh"
#include "pointcloud.cuh"
#include "resize.cuh"
#include "rollout_glue.cuh"
#include "ik.cuh"
