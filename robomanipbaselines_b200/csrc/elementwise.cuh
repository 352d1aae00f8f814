// Memory-bound kernels of the ACT rollout path (coalesced, vectorised, warp-shuffle; no tensor cores).
//   preprocess_u8 / preprocess_f32 : RolloutBase.get_images (common/base/RolloutBase.py:479-490) + ACTPolicy's ImageNet
//                                    normalise (third_party/act/policy.py:20-22), written as padded NHWC4 act_t for conv1
//   maxpool3x3s2                   : torchvision resnet18 maxpool (via backbone.py:86-96)
//   layernorm_rows                 : nn.LayerNorm after each transformer sub-block (transformer.py:166-179,229-250)
//   act_head                       : norm3 -> decoder.norm -> action_head (transformer.py:129-139, detr_vae.py:148)
//   ensemble_step                  : temporal ensembling + denormalise (policy/act/RolloutAct.py:68-101, DataUtils.py:26-40)
//   normalize_state                : normalize_data (common/utils/DataUtils.py:9-23) + f64->f32 cast (RolloutBase.py:474-475)
#pragma once
#include <type_traits>
#include <cuda_runtime.h>
#include <stdint.h>

#include "ptx.cuh"

namespace rmb {

constexpr int kConv1PadPixels = 4;  // zero pixels on each side of a row in the padded NHWC4 conv1 input

// ---------------------------------------------------------------------------------------------------------------------
// uint8 HWC image -> normalised 16-bit (act_t), 4 channels (4th = 0), row pitch (W + 8) pixels with 4 zero pixels either side.
// One thread = 4 pixels (12 input bytes as 3 x u32, 32 output bytes as 2 x uint4).
__global__ void preprocess_u8_kernel(const uint8_t* __restrict__ in, act_t* __restrict__ out, int n_img, int H,
                                     int W, float3 mean, float3 inv_std) {
  pdl_trigger();
  pdl_wait();
  const int quads_per_row = W / 4 + 2;  // incl. one zero quad on each side
  const long long total = static_cast<long long>(n_img) * H * quads_per_row;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(idx % quads_per_row);
    const long long row = idx / quads_per_row;  // img*H + y
    uint4 o0 = make_uint4(0, 0, 0, 0), o1 = make_uint4(0, 0, 0, 0);
    if (q >= 1 && q <= W / 4) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(in + (row * W + (q - 1) * 4) * 3);
      const uint32_t a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
      const uint8_t px[12] = {static_cast<uint8_t>(a),       static_cast<uint8_t>(a >> 8),  static_cast<uint8_t>(a >> 16),
                              static_cast<uint8_t>(a >> 24), static_cast<uint8_t>(b),       static_cast<uint8_t>(b >> 8),
                              static_cast<uint8_t>(b >> 16), static_cast<uint8_t>(b >> 24), static_cast<uint8_t>(c),
                              static_cast<uint8_t>(c >> 8),  static_cast<uint8_t>(c >> 16), static_cast<uint8_t>(c >> 24)};
      uint32_t w[8];
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        // same arithmetic as the reference: x * (1/255) in fp32 (v2.ToDtype scale=True), then (x - mean)/std
        const float k255 = 1.0f / 255.0f;
        const float r = (static_cast<float>(px[3 * p + 0]) * k255 - mean.x) * inv_std.x;
        const float g = (static_cast<float>(px[3 * p + 1]) * k255 - mean.y) * inv_std.y;
        const float bl = (static_cast<float>(px[3 * p + 2]) * k255 - mean.z) * inv_std.z;
        w[2 * p] = pack_act2(r, g);
        w[2 * p + 1] = pack_act2(bl, 0.0f);
      }
      o0 = make_uint4(w[0], w[1], w[2], w[3]);
      o1 = make_uint4(w[4], w[5], w[6], w[7]);
    }
    uint4* dst = reinterpret_cast<uint4*>(out + (row * (W + 2 * kConv1PadPixels) + q * 4) * 4);
    dst[0] = o0;
    dst[1] = o1;
  }
}

// float32 CHW image in [0,1] (the reference nn.Module's input) -> same padded NHWC4 act_t layout.
__global__ void preprocess_f32_kernel(const float* __restrict__ in, act_t* __restrict__ out, int n_img, int H,
                                      int W, float3 mean, float3 inv_std) {
  pdl_trigger();
  pdl_wait();
  const int quads_per_row = W / 4 + 2;
  const long long total = static_cast<long long>(n_img) * H * quads_per_row;
  const long long plane = static_cast<long long>(H) * W;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(idx % quads_per_row);
    const long long row = idx / quads_per_row;
    const long long img = row / H;
    const int y = static_cast<int>(row % H);
    uint4 o0 = make_uint4(0, 0, 0, 0), o1 = make_uint4(0, 0, 0, 0);
    if (q >= 1 && q <= W / 4) {
      const float* base = in + img * 3 * plane + static_cast<long long>(y) * W + (q - 1) * 4;
      const float4 r4 = __ldg(reinterpret_cast<const float4*>(base));
      const float4 g4 = __ldg(reinterpret_cast<const float4*>(base + plane));
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(base + 2 * plane));
      const float rr[4] = {r4.x, r4.y, r4.z, r4.w}, gg[4] = {g4.x, g4.y, g4.z, g4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
      uint32_t w[8];
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        w[2 * p] = pack_act2((rr[p] - mean.x) * inv_std.x, (gg[p] - mean.y) * inv_std.y);
        w[2 * p + 1] = pack_act2((bb[p] - mean.z) * inv_std.z, 0.0f);
      }
      o0 = make_uint4(w[0], w[1], w[2], w[3]);
      o1 = make_uint4(w[4], w[5], w[6], w[7]);
    }
    uint4* dst = reinterpret_cast<uint4*>(out + (row * (W + 2 * kConv1PadPixels) + q * 4) * 4);
    dst[0] = o0;
    dst[1] = o1;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// 3x3 stride-2 pad-1 max-pool, NHWC act_t. One thread = 8 channels (16 B) of one output pixel.
__global__ void maxpool3x3s2_kernel(const act_t* __restrict__ in, act_t* __restrict__ out, int n_img,
                                    int H, int W, int C) {
  pdl_trigger();
  pdl_wait();
  const int OH = (H + 1) / 2, OW = (W + 1) / 2, CV = C / 8;
  const long long total = static_cast<long long>(n_img) * OH * OW * CV;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(idx % CV);
    long long t = idx / CV;
    const int ox = static_cast<int>(t % OW); t /= OW;
    const int oy = static_cast<int>(t % OH);
    const long long img = t / OH;
    uint4 m = make_uint4(kActNegInf2, kActNegInf2, kActNegInf2, kActNegInf2);  // -inf pairs
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
      const int y = 2 * oy + dy;
      if (y < 0 || y >= H) continue;
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        const int x = 2 * ox + dx;
        if (x < 0 || x >= W) continue;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + ((img * H + y) * W + x) * C) + cv);
        m.x = act2_max(m.x, v.x); m.y = act2_max(m.y, v.y); m.z = act2_max(m.z, v.z); m.w = act2_max(m.w, v.w);
      }
    }
    reinterpret_cast<uint4*>(out + ((img * OH + oy) * OW + ox) * C)[cv] = m;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Row-wise LayerNorm over D = 512 (one warp per row; fp32 statistics, eps inside the sqrt like torch).
//   x_in  [rows, 512] fp32
//   outputs (each optional): y_f32 (residual stream), y_act (GEMM A operand), ypos_act = act_t(y + pos[row % L])
//   do_norm = 0 turns it into the plain cast(+pos) pass used once on the assembled input tokens.
constexpr int kD = 512;
template <typename TIn>
__global__ void layernorm512_kernel_t(const TIn* __restrict__ x_in, const float* __restrict__ gamma,
                                    const float* __restrict__ beta, const float* __restrict__ pos, int L,
                                    float* __restrict__ y_f32, act_t* __restrict__ y_act,
                                    act_t* __restrict__ ypos_act, long long rows, int do_norm, float eps) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long warp_global = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  // the next row's loads are issued before this row is reduced and written (two rows in flight per warp: a 16-bit
  // row is only 1 KB, and one row per warp left the kernel at 4.6-5.0 TB/s)
  using raw_t = typename std::conditional<sizeof(TIn) == 4, float4, uint2>::type;
  raw_t nxt[4];
  if (warp_global < rows) {
    const raw_t* xp = reinterpret_cast<const raw_t*>(x_in + warp_global * kD);
#pragma unroll
    for (int i = 0; i < 4; ++i) nxt[i] = __ldg(xp + i * 32 + lane);
  }
  for (long long row = warp_global; row < rows; row += warps_total) {
    raw_t cur[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) cur[i] = nxt[i];
    if (row + warps_total < rows) {
      const raw_t* xp = reinterpret_cast<const raw_t*>(x_in + (row + warps_total) * kD);
#pragma unroll
      for (int i = 0; i < 4; ++i) nxt[i] = __ldg(xp + i * 32 + lane);
    }
    float v[16];
    if constexpr (sizeof(TIn) == 4) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 t = *reinterpret_cast<const float4*>(&cur[i]);
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
      }
    } else {                                          // 16-bit pre-norm sums (same element -> lane mapping)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint2 t = *reinterpret_cast<const uint2*>(&cur[i]);
        const float2 a = unpack_act2(t.x), b = unpack_act2(t.y);
        v[4 * i] = a.x; v[4 * i + 1] = a.y; v[4 * i + 2] = b.x; v[4 * i + 3] = b.y;
      }
    }
    if (do_norm) {
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) s += v[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mu = s * (1.0f / kD);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) { const float d = v[i] - mu; q += d * d; }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float rstd = rsqrtf(q * (1.0f / kD) + eps);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i * 32 + lane);
        const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + i * 32 + lane);
        v[4 * i + 0] = (v[4 * i + 0] - mu) * rstd * g.x + b.x;
        v[4 * i + 1] = (v[4 * i + 1] - mu) * rstd * g.y + b.y;
        v[4 * i + 2] = (v[4 * i + 2] - mu) * rstd * g.z + b.z;
        v[4 * i + 3] = (v[4 * i + 3] - mu) * rstd * g.w + b.w;
      }
    }
    if (y_f32 != nullptr) {
      float4* yp = reinterpret_cast<float4*>(y_f32 + row * kD);
#pragma unroll
      for (int i = 0; i < 4; ++i) yp[i * 32 + lane] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
    }
    if (y_act != nullptr) {
      uint2* yp = reinterpret_cast<uint2*>(y_act + row * kD);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        yp[i * 32 + lane] = make_uint2(pack_act2(v[4 * i], v[4 * i + 1]), pack_act2(v[4 * i + 2], v[4 * i + 3]));
      }
    }
    if (ypos_act != nullptr) {
      const float4* pp = reinterpret_cast<const float4*>(pos + (row % L) * kD);
      uint2* yp = reinterpret_cast<uint2*>(ypos_act + row * kD);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 pv = __ldg(pp + i * 32 + lane);
        yp[i * 32 + lane] = make_uint2(pack_act2(v[4 * i] + pv.x, v[4 * i + 1] + pv.y),
                                       pack_act2(v[4 * i + 2] + pv.z, v[4 * i + 3] + pv.w));
      }
    }
  }
}
// the two instances in use: fp32 pre-norm rows (C ABI, decoder, bf16 build) and 16-bit pre-norm rows (encoder, fp16 build)
#define layernorm512_kernel layernorm512_kernel_t<float>

// ---------------------------------------------------------------------------------------------------------------------
// Tokens 0 (latent) and 1 (proprio) of every episode (detr_vae.py:124-125,138; transformer.py:62-63), fp32 [B, L, 512].
//   latent token = latent_out_proj.bias (latent sample is zeros); proprio = input_proj_robot_state(qpos).
//   task_tok (MT-ACT with the task token switched on, else nullptr): third prefix token = proj_text_emb(task_emb).
__global__ void act_extra_tokens_kernel(const float* __restrict__ qpos, int Ds, const float* __restrict__ latent_bias,
                                        const float* __restrict__ w_state,  // [512, Ds]
                                        const float* __restrict__ b_state, const float* __restrict__ task_tok,
                                        float* __restrict__ x_f32, int L) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < kD; c += blockDim.x) {
    float acc = b_state[c];
    for (int k = 0; k < Ds; ++k) acc += w_state[c * Ds + k] * qpos[b * Ds + k];
    x_f32[(static_cast<long long>(b) * L + 0) * kD + c] = latent_bias[c];
    x_f32[(static_cast<long long>(b) * L + 1) * kD + c] = acc;
    if (task_tok != nullptr) x_f32[(static_cast<long long>(b) * L + 2) * kD + c] = task_tok[static_cast<long long>(b) * kD + c];
  }
}

// fp32 -> 16-bit operand cast (task embeddings of MT-ACT in front of their projection GEMMs)
__global__ void cast_f32_to_act_kernel(const float* __restrict__ x, act_t* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    y[i] = to_act(x[i]);
}

// ---------------------------------------------------------------------------------------------------------------------
// Decoder tail: y = decoder.norm(norm3(x)); a_hat = action_head(y).   One warp per (episode, query) row.
__global__ void act_head_kernel(const float* __restrict__ x_in, const float* __restrict__ g3, const float* __restrict__ b3,
                                const float* __restrict__ gd, const float* __restrict__ bd,
                                const float* __restrict__ w_head,  // [Da, 512]
                                const float* __restrict__ b_head, float* __restrict__ a_hat, long long rows, int Da,
                                float eps) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const long long warp_global = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  for (long long row = warp_global; row < rows; row += warps_total) {
    float v[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(x_in + row * kD) + i * 32 + lane);
      v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
#pragma unroll
    for (int pass = 0; pass < 2; ++pass) {
      const float* g = pass == 0 ? g3 : gd;
      const float* bta = pass == 0 ? b3 : bd;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) s += v[i];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mu = s * (1.0f / kD);
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) { const float d = v[i] - mu; q += d * d; }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float rstd = rsqrtf(q * (1.0f / kD) + eps);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 gg = __ldg(reinterpret_cast<const float4*>(g) + i * 32 + lane);
        const float4 bb = __ldg(reinterpret_cast<const float4*>(bta) + i * 32 + lane);
        v[4 * i + 0] = (v[4 * i + 0] - mu) * rstd * gg.x + bb.x;
        v[4 * i + 1] = (v[4 * i + 1] - mu) * rstd * gg.y + bb.y;
        v[4 * i + 2] = (v[4 * i + 2] - mu) * rstd * gg.z + bb.z;
        v[4 * i + 3] = (v[4 * i + 3] - mu) * rstd * gg.w + bb.w;
      }
    }
    for (int a = 0; a < Da; ++a) {
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 wv = __ldg(reinterpret_cast<const float4*>(w_head + a * kD) + i * 32 + lane);
        acc += v[4 * i] * wv.x + v[4 * i + 1] * wv.y + v[4 * i + 2] * wv.z + v[4 * i + 3] * wv.w;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) a_hat[row * Da + a] = acc + b_head[a];
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Temporal ensembling ring buffer (policy/act/RolloutAct.py:68-101).
//   ring   [B, C, C, Da] fp32 : slot (n mod C) holds the chunk predicted at policy call n of that episode
//   chunks [B, C, Da]   fp32 : chunk predicted at this call
//   step   [B] int32          : n = 0-based count of policy calls in the episode (advanced by this kernel)
//   reset  [B] uint8 or null  : 1 -> episode restarts: n := 0 before this call (RolloutAct.reset_variables :62-66)
//   wtab   [C, C] fp64        : wtab[(m-1)*C + r] = exp(-k r) / sum_{s<m} exp(-k s), made on the host with numpy so
//                               that it is bit-identical to the reference's np.exp / np.sum
//   action_n = sum_{i=0}^{m-1} wtab[m-1][m-1-i] * A[n-i][i],  m = min(n+1, C), accumulated in fp64 in the order
//   i = 0,1,... with separately rounded multiply and add (numpy semantics; no FMA contraction), then denormalised:
//   gaussian: std*a + mean ; limits: scale*(a - out_min) + min       (common/utils/DataUtils.py:26-40)
// One thread per (episode, action dim); the chunk store is cooperative over the block.
struct DenormParams {
  const double* a;   // gaussian: std   | limits: scale = range / (out_max - out_min)
  const double* b;   // gaussian: mean  | limits: min
  double out_min;    // limits only
  int type;          // 0 gaussian, 1 limits
};
__global__ void ensemble_step_kernel(float* __restrict__ ring, const float* __restrict__ chunks,
                                     int32_t* __restrict__ step, const uint8_t* __restrict__ reset,
                                     const double* __restrict__ wtab, DenormParams dn, double* __restrict__ action_out,
                                     float* __restrict__ action_out_f32, double* __restrict__ action_norm_out, int C,
                                     int Da, int use_smem) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  int n = step[b];
  if (reset != nullptr && reset[b]) n = 0;
  const int slot = n % C;
  float* ring_b = ring + static_cast<long long>(b) * C * C * Da;
  const float* chunk_b = chunks + static_cast<long long>(b) * C * Da;
  for (int i = threadIdx.x; i < C * Da; i += blockDim.x) ring_b[static_cast<long long>(slot) * C * Da + i] = chunk_b[i];
  __syncthreads();
  const int m = n + 1 < C ? n + 1 : C;
  // stage the diagonal A[n-i][i] (i < m) in shared memory with all threads (m independent 28-byte gathers in flight at
  // once instead of m dependent ones); the fp64 sum below keeps the reference's order
  extern __shared__ float ens_diag[];
  if (use_smem) {
    for (int idx = threadIdx.x; idx < m * Da; idx += blockDim.x) {
      const int i = idx / Da, d = idx - i * Da;
      int s = (n - i) % C;
      if (s < 0) s += C;
      ens_diag[idx] = ring_b[(static_cast<long long>(s) * C + i) * Da + d];
    }
    __syncthreads();
  }
  if (threadIdx.x < Da) {
    const int d = threadIdx.x;
    const double* w = wtab + static_cast<long long>(m - 1) * C;
    double acc = 0.0;
    for (int i = 0; i < m; ++i) {
      float av;
      if (use_smem) {
        av = ens_diag[i * Da + d];
      } else {
        int s = (n - i) % C;
        if (s < 0) s += C;
        av = ring_b[(static_cast<long long>(s) * C + i) * Da + d];
      }
      acc = __dadd_rn(acc, __dmul_rn(w[m - 1 - i], static_cast<double>(av)));
    }
    double o;
    if (dn.type == 0) o = __dadd_rn(__dmul_rn(dn.a[d], acc), dn.b[d]);
    else o = __dadd_rn(__dmul_rn(dn.a[d], __dsub_rn(acc, dn.out_min)), dn.b[d]);
    if (action_norm_out != nullptr) action_norm_out[b * Da + d] = acc;
    if (action_out != nullptr) action_out[b * Da + d] = o;
    if (action_out_f32 != nullptr) action_out_f32[b * Da + d] = static_cast<float>(o);
  }
  __syncthreads();
  if (threadIdx.x == 0) step[b] = n + 1;
}

// ---------------------------------------------------------------------------------------------------------------------
// normalize_data on the device: fp64 in (as numpy), fp32 out (torch.tensor(..., dtype=float32)).
//   gaussian: (x - mean)/std ; limits: scale*(x - min) + out_min with scale = (out_max - out_min)/range
__global__ void normalize_state_kernel(const double* __restrict__ x, const double* __restrict__ p0,
                                       const double* __restrict__ p1, double out_min, int type,
                                       float* __restrict__ y, long long total, int D) {
  pdl_trigger();
  pdl_wait();
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int d = static_cast<int>(idx % D);
    double r;
    if (type == 0) r = __ddiv_rn(__dsub_rn(x[idx], p0[d]), p1[d]);                       // p0 = mean, p1 = std
    else r = __dadd_rn(__dmul_rn(p1[d], __dsub_rn(x[idx], p0[d])), out_min);             // p0 = min, p1 = scale
    y[idx] = static_cast<float>(r);
  }
}

// denormalize_data on the device for plain (non-ensembled) actions: fp64 in/out.
__global__ void denormalize_kernel(const double* __restrict__ x, DenormParams dn, double* __restrict__ y,
                                   long long total, int D) {
  pdl_trigger();
  pdl_wait();
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int d = static_cast<int>(idx % D);
    if (dn.type == 0) y[idx] = __dadd_rn(__dmul_rn(dn.a[d], x[idx]), dn.b[d]);
    else y[idx] = __dadd_rn(__dmul_rn(dn.a[d], __dsub_rn(x[idx], dn.out_min)), dn.b[d]);
  }
}

}  // namespace rmb
