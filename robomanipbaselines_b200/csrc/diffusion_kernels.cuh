// Memory-bound kernels of the Diffusion Policy / DP3 denoising path (coalesced, vectorised, warp-shuffle).
// The dense contractions (Conv1d / ConvTranspose1d / FiLM Linear / PointNet Linear) run on gemm_tc_kernel.
//   unet_gn_mish      : Conv1dBlock's GroupNorm(8) + Mish, then the FiLM affine and the block residual of
//                       ConditionalResidualBlock1D (upstream diffusion_policy conditional_unet1d.py; restated in
//                       oracle/diffusion_oracle.py:57-94)
//   unet_final_step   : final_conv's GroupNorm + Mish + Conv1d(C -> Da, 1), fused with DDPMScheduler.step /
//                       DDIMScheduler.step (diffusers 0.11.1; oracle/diffusion_oracle.py:179-264) and with the
//                       im2col of x_{t-1} that feeds the next UNet call
//   unet_init / mish_cast : x_T -> state + im2col;  Mish(global_cond) -> 16-bit GEMM operand
//   pointnet_*        : DP3 PointNetEncoderXYZ + state MLP (oracle/diffusion_oracle.py:280-316)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ptx.cuh"

namespace rmb {

constexpr int kGnMaxVec = 16;      // float4 vectors a lane may hold: T * (C / 8) / 128 <= 16
constexpr int kXcolMaxWidth = 128;  // im2col row of the UNet input: kernel_size * input_dim padded to 64 or 128 columns (rest zero)

__device__ __forceinline__ float mish_f(float x) {
  // x * tanh(softplus(x)) with tanh(log(1 + e)) = n / (n + 2), n = e * (e + 2), e = exp(x). Branch-free: the exponent is
  // clamped at 20 (torch's softplus threshold), where n / (n + 2) already rounds to 1 in fp32, so x > 20 returns x.
  float e, rc;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(x, 20.0f) * 1.4426950408889634f));
  const float n = e * (e + 2.0f);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(n + 2.0f));
  return x * (n * rc);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct GnParams {
  const void* in;        // conv output [B*T, ld_in]: fp32 (in_f32) or act_t
  int in_f32, ld_in;
  const float* gamma;    // [C]
  const float* beta;     // [C]
  int C, T;
  const float* film_gc;  // [B, film_ld] step-independent part of the FiLM embedding (+ cond_encoder bias), or nullptr
  const float* film_t;   // [>= film_off + 2C] part that depends on the diffusion step only (row of the table)
  int film_ld, film_off; // scale at film_off + c, bias at film_off + C + c
  const act_t* res;      // residual [B*T, ld_res] or nullptr
  int ld_res;
  act_t* out;            // [B*T, ld_out]
  int ld_out;
};

// One CTA per sample, one warp per group (n_groups = 8): the group's T x C/8 values live in registers.
__device__ __forceinline__ void gn_load_group(const GnParams& p, int b, int g, int lane, float4 (&v)[kGnMaxVec],
                                              float* mean_out, float* rstd_out) {
  const int Cg = p.C >> 3, vpr = Cg >> 2, nvec = p.T * vpr;
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kGnMaxVec; ++i) {
    const int idx = lane + 32 * i;
    v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (idx < nvec) {
      const int row = idx / vpr, col = g * Cg + (idx - row * vpr) * 4;
      const long long off = (static_cast<long long>(b) * p.T + row) * p.ld_in + col;
      if (p.in_f32) {
        v[i] = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(p.in) + off));
      } else {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(static_cast<const act_t*>(p.in) + off));
        const float2 lo = unpack_act2(u.x), hi = unpack_act2(u.y);
        v[i] = make_float4(lo.x, lo.y, hi.x, hi.y);
      }
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    }
  }
  const float n = static_cast<float>(p.T * Cg);
  const float mean = warp_sum(s) / n;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kGnMaxVec; ++i) {
    if (lane + 32 * i < nvec) {
      const float a = v[i].x - mean, b2 = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b2 * b2) + (c * c + d * d);
    }
  }
  *mean_out = mean;
  *rstd_out = rsqrtf(warp_sum(q) / n + 1e-5f);
}

__global__ void __launch_bounds__(256) unet_gn_mish_kernel(const GnParams p) {
  const int b = blockIdx.x, g = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Cg = p.C >> 3, vpr = Cg >> 2, nvec = p.T * vpr;
  float4 v[kGnMaxVec];
  float mean, rstd;
  gn_load_group(p, b, g, lane, v, &mean, &rstd);
  const float* fg = p.film_gc ? p.film_gc + static_cast<long long>(b) * p.film_ld + p.film_off : nullptr;
  const float* ft = p.film_gc ? p.film_t + p.film_off : nullptr;
#pragma unroll
  for (int i = 0; i < kGnMaxVec; ++i) {
    const int idx = lane + 32 * i;
    if (idx < nvec) {
      const int row = idx / vpr, col = g * Cg + (idx - row * vpr) * 4;
      const float4 ga = __ldg(reinterpret_cast<const float4*>(p.gamma + col));
      const float4 be = __ldg(reinterpret_cast<const float4*>(p.beta + col));
      float4 y;
      y.x = mish_f((v[i].x - mean) * rstd * ga.x + be.x);
      y.y = mish_f((v[i].y - mean) * rstd * ga.y + be.y);
      y.z = mish_f((v[i].z - mean) * rstd * ga.z + be.z);
      y.w = mish_f((v[i].w - mean) * rstd * ga.w + be.w);
      if (fg != nullptr) {
        const float4 s1 = __ldg(reinterpret_cast<const float4*>(fg + col)), s2 = __ldg(reinterpret_cast<const float4*>(ft + col));
        const float4 b1 = __ldg(reinterpret_cast<const float4*>(fg + p.C + col)), b2 = __ldg(reinterpret_cast<const float4*>(ft + p.C + col));
        y.x = (s1.x + s2.x) * y.x + (b1.x + b2.x);
        y.y = (s1.y + s2.y) * y.y + (b1.y + b2.y);
        y.z = (s1.z + s2.z) * y.z + (b1.z + b2.z);
        y.w = (s1.w + s2.w) * y.w + (b1.w + b2.w);
      }
      const long long r = static_cast<long long>(b) * p.T + row;
      if (p.res != nullptr) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(p.res + r * p.ld_res + col));
        const float2 lo = unpack_act2(u.x), hi = unpack_act2(u.y);
        y.x += lo.x; y.y += lo.y; y.z += hi.x; y.w += hi.y;
      }
      *reinterpret_cast<uint2*>(p.out + r * p.ld_out + col) = make_uint2(pack_act2(y.x, y.y), pack_act2(y.z, y.w));
    }
  }
}

// One denoising step's scalars (host-made with the scheduler's own fp32 arithmetic; see include/rmb_b200.h).
struct SchedStep {
  int kind;            // 0 DDPM, 1 DDIM, 2 none (write the model output only)
  int pred_sample;     // prediction_type: 0 epsilon, 1 sample
  int clip;            // clip_sample
  int derive_eps;      // DDIM: direction from the derived epsilon (later diffusers) instead of the raw model output
  float s_beta, s_alpha, c_x0, c_xt, sigma;
};

struct FinalParams {
  GnParams gn;              // in / gamma / beta / C / T of final_conv.0 (no FiLM, no residual, out unused)
  const float* w_head;      // [Da, C] final_conv.1.weight
  const float* b_head;      // [Da]
  int Da, ksize;
  float* x;                 // [B, T, Da] fp32 trajectory, updated in place
  const float* noise;       // [B, T, Da] or nullptr
  float* model_out;         // kind 2: [B, T, Da]
  act_t* xcol;              // [B*T, xcol_w] im2col of the new x for the first conv of the next step
  int xcol_w;               // 64 or 128 (kernel_size * Da rounded up)
  SchedStep st;
};

// torch evaluates the scheduler formulas op by op in fp32: no fused multiply-add here, same op order.
__device__ __forceinline__ float sched_step(const SchedStep& s, float x, float o, float nz) {
  float x0 = s.pred_sample ? o : __fdiv_rn(__fsub_rn(x, __fmul_rn(s.s_beta, o)), s.s_alpha);
  if (s.clip) x0 = fminf(fmaxf(x0, -1.0f), 1.0f);
  if (s.kind == 0) {
    float prev = __fadd_rn(__fmul_rn(s.c_x0, x0), __fmul_rn(s.c_xt, x));
    if (s.sigma > 0.0f) prev = __fadd_rn(prev, __fmul_rn(s.sigma, nz));
    return prev;
  }
  float dir = o;
  if (s.derive_eps && s.pred_sample) dir = __fdiv_rn(__fsub_rn(x, __fmul_rn(s.s_alpha, x0)), s.s_beta);
  return __fadd_rn(__fmul_rn(s.c_x0, x0), __fmul_rn(s.c_xt, dir));
}

// Stand-alone scheduler step over n elements (the same sched_step() the fused kernel applies): used by the bit-exact
// parity test against the oracle and by the HBM micro-benchmark (3 reads + 1 write of 4 B per element).
__global__ void __launch_bounds__(256) sched_step_kernel(const float* __restrict__ x, const float* __restrict__ model_out,
                                                         const float* __restrict__ noise, float* __restrict__ out,
                                                         long long n, const SchedStep st) {
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(x) + i), o = __ldg(reinterpret_cast<const float4*>(model_out) + i);
    const float4 z = noise ? __ldg(reinterpret_cast<const float4*>(noise) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(out)[i] = make_float4(sched_step(st, a.x, o.x, z.x), sched_step(st, a.y, o.y, z.y),
                                                    sched_step(st, a.z, o.z, z.z), sched_step(st, a.w, o.w, z.w));
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    out[i] = sched_step(st, x[i], model_out[i], noise ? noise[i] : 0.f);
  }
}

// dynamic smem: T*C floats (normalised activations) + 3 * T*Da floats (new x, current x, noise)
__global__ void __launch_bounds__(256) unet_final_step_kernel(const FinalParams p) {
  extern __shared__ float fs_smem[];
  float* ys = fs_smem;
  float* xs = fs_smem + p.gn.T * p.gn.C;
  const int b = blockIdx.x, g = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = p.gn.C, T = p.gn.T, Da = p.Da;
  const int Cg = C >> 3, vpr = Cg >> 2, nvec = T * vpr;
  float* x_in = xs + T * Da;
  float* nz_in = x_in + T * Da;
  // the sample's current x and this step's noise are fetched now, under the GroupNorm pass: the head loop below would
  // otherwise pay one HBM round trip per output element in its single-lane tail
  for (int i = threadIdx.x; i < T * Da; i += 256) {
    const long long gi = static_cast<long long>(b) * T * Da + i;
    x_in[i] = p.x[gi];
    nz_in[i] = (p.noise != nullptr && p.st.kind != 2) ? __ldg(p.noise + gi) : 0.0f;
  }
  {
    float4 v[kGnMaxVec];
    float mean, rstd;
    gn_load_group(p.gn, b, g, lane, v, &mean, &rstd);
#pragma unroll
    for (int i = 0; i < kGnMaxVec; ++i) {
      const int idx = lane + 32 * i;
      if (idx < nvec) {
        const int row = idx / vpr, col = g * Cg + (idx - row * vpr) * 4;
        const float4 ga = __ldg(reinterpret_cast<const float4*>(p.gn.gamma + col));
        const float4 be = __ldg(reinterpret_cast<const float4*>(p.gn.beta + col));
        float4 y;
        y.x = mish_f((v[i].x - mean) * rstd * ga.x + be.x);
        y.y = mish_f((v[i].y - mean) * rstd * ga.y + be.y);
        y.z = mish_f((v[i].z - mean) * rstd * ga.z + be.z);
        y.w = mish_f((v[i].w - mean) * rstd * ga.w + be.w);
        *reinterpret_cast<float4*>(ys + row * C + col) = y;
      }
    }
  }
  __syncthreads();
  for (int o = g; o < T * Da; o += 8) {
    const int t = o / Da, a = o - t * Da;
    const float* yr = ys + t * C;
    const float* wr = p.w_head + a * C;
    float acc = 0.f;
    for (int c = lane; c < C; c += 32) acc = fmaf(yr[c], __ldg(wr + c), acc);
    acc = warp_sum(acc);
    if (lane == 0) {
      const float out = acc + __ldg(p.b_head + a);
      const long long gi = (static_cast<long long>(b) * T + t) * Da + a;
      if (p.st.kind == 2) {
        p.model_out[gi] = out;
        xs[o] = x_in[o];
      } else {
        const float xn = sched_step(p.st, x_in[o], out, nz_in[o]);
        p.x[gi] = xn;
        xs[o] = xn;
      }
    }
  }
  __syncthreads();
  const int pad = p.ksize >> 1, kd = p.ksize * Da;
  const int xw = p.xcol_w;
  for (int i = threadIdx.x; i < T * xw; i += 256) {
    const int t = i / xw, j = i - t * xw;
    float val = 0.f;
    if (j < kd) {
      const int tap = j / Da, a = j - tap * Da, tt = t + tap - pad;
      if (tt >= 0 && tt < T) val = xs[tt * Da + a];
    }
    p.xcol[(static_cast<long long>(b) * T + t) * xw + j] = to_act(val);
  }
}

// x_T (initial noise trajectory) -> fp32 state + im2col. One CTA per sample.
__global__ void unet_init_kernel(const float* __restrict__ x_init, float* __restrict__ x, act_t* __restrict__ xcol,
                                 int T, int Da, int ksize, int xw) {
  const int b = blockIdx.x;
  const float* src = x_init + static_cast<long long>(b) * T * Da;
  if (x != x_init)
    for (int i = threadIdx.x; i < T * Da; i += blockDim.x) x[static_cast<long long>(b) * T * Da + i] = src[i];
  const int pad = ksize >> 1, kd = ksize * Da;
  for (int i = threadIdx.x; i < T * xw; i += blockDim.x) {
    const int t = i / xw, j = i - t * xw;
    float val = 0.f;
    if (j < kd) {
      const int tap = j / Da, a = j - tap * Da, tt = t + tap - pad;
      if (tt >= 0 && tt < T) val = src[tt * Da + a];
    }
    xcol[(static_cast<long long>(b) * T + t) * xw + j] = to_act(val);
  }
}

// Mish(global_cond) [B, G] fp32 -> act_t [B, Gp] (zero padded): the A operand of the once-per-inference FiLM GEMM
__global__ void mish_cast_kernel(const float* __restrict__ gc, act_t* __restrict__ out, long long B, int G, int Gp) {
  const long long total = B * Gp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / Gp;
    const int c = static_cast<int>(i - r * Gp);
    out[i] = to_act(c < G ? mish_f(__ldg(gc + r * G + c)) : 0.0f);
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// DP3 PointNetEncoderXYZ layer 1: Linear(in_ch -> 64) + LayerNorm(64) + ReLU per point, 8 lanes per point (8 outputs
// each), so a warp writes 4 points x 128 B contiguous.
__global__ void __launch_bounds__(256) pointnet_layer1_kernel(const float* __restrict__ pts, int pt_stride, int in_ch,
                                                              const float* __restrict__ W, const float* __restrict__ bias,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              act_t* __restrict__ out, long long rows) {
  const int sub = threadIdx.x & 7;
  float w[8][6], bb[8], ga[8], be[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = sub * 8 + j;
#pragma unroll
    for (int k = 0; k < 6; ++k) w[j][k] = (k < in_ch) ? __ldg(W + c * in_ch + k) : 0.f;
    bb[j] = __ldg(bias + c); ga[j] = __ldg(gamma + c); be[j] = __ldg(beta + c);
  }
  const long long stride = static_cast<long long>(gridDim.x) * (blockDim.x >> 3);
  for (long long r = blockIdx.x * static_cast<long long>(blockDim.x >> 3) + (threadIdx.x >> 3);
       r < ((rows + 3) & ~3ll); r += stride) {
    const bool ok = r < rows;
    float x[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) x[k] = (ok && k < in_ch) ? __ldg(pts + r * pt_stride + k) : 0.f;
    float y[8], s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float a = bb[j];
#pragma unroll
      for (int k = 0; k < 6; ++k) a = fmaf(x[k], w[j][k], a);
      y[j] = a; s += a;
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    const float mean = s * (1.0f / 64.0f);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float d = y[j] - mean; q += d * d; }
    q += __shfl_xor_sync(0xffffffffu, q, 1); q += __shfl_xor_sync(0xffffffffu, q, 2); q += __shfl_xor_sync(0xffffffffu, q, 4);
    const float rstd = rsqrtf(q * (1.0f / 64.0f) + 1e-5f);
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float a = fmaxf((y[2 * j] - mean) * rstd * ga[2 * j] + be[2 * j], 0.f);
      const float b2 = fmaxf((y[2 * j + 1] - mean) * rstd * ga[2 * j + 1] + be[2 * j + 1], 0.f);
      o[j] = pack_act2(a, b2);
    }
    if (ok) *reinterpret_cast<uint4*>(out + r * 64 + sub * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// LayerNorm(N) + ReLU over rows of a GEMM output (fp32 [rows, N]), one warp per row. Writes act_t rows (layer 2) and/or
// folds the rows of one cloud into a running max (layer 3: torch.max(x, 1)[0]); ReLU output >= 0, so max on the int
// bit pattern against a zero-initialised buffer is exact.
template <int N>
__global__ void __launch_bounds__(256) pointnet_ln_relu_kernel(const float* __restrict__ in, const float* __restrict__ gamma,
                                                               const float* __restrict__ beta, act_t* __restrict__ out,
                                                               float* __restrict__ maxout, int Np, int rows_per_cta) {
  constexpr int V = N / 128;
  __shared__ float smax[8][N];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chunks = (Np + rows_per_cta - 1) / rows_per_cta;
  const int cloud = blockIdx.x / chunks, r0 = (blockIdx.x % chunks) * rows_per_cta;
  const int r1 = min(r0 + rows_per_cta, Np);
  float4 ga[V], be[V], mx[V];
#pragma unroll
  for (int k = 0; k < V; ++k) {
    ga[k] = __ldg(reinterpret_cast<const float4*>(gamma + k * 128 + lane * 4));
    be[k] = __ldg(reinterpret_cast<const float4*>(beta + k * 128 + lane * 4));
    mx[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int r = r0 + warp; r < r1; r += 8) {
    const long long row = static_cast<long long>(cloud) * Np + r;
    float4 v[V];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      v[k] = __ldg(reinterpret_cast<const float4*>(in + row * N + k * 128 + lane * 4));
      s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
    }
    const float mean = warp_sum(s) * (1.0f / N);
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float a = v[k].x - mean, b2 = v[k].y - mean, c = v[k].z - mean, d = v[k].w - mean;
      q += (a * a + b2 * b2) + (c * c + d * d);
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / N) + 1e-5f);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      float4 y;
      y.x = fmaxf((v[k].x - mean) * rstd * ga[k].x + be[k].x, 0.f);
      y.y = fmaxf((v[k].y - mean) * rstd * ga[k].y + be[k].y, 0.f);
      y.z = fmaxf((v[k].z - mean) * rstd * ga[k].z + be[k].z, 0.f);
      y.w = fmaxf((v[k].w - mean) * rstd * ga[k].w + be[k].w, 0.f);
      if (out != nullptr)
        *reinterpret_cast<uint2*>(out + row * N + k * 128 + lane * 4) = make_uint2(pack_act2(y.x, y.y), pack_act2(y.z, y.w));
      mx[k].x = fmaxf(mx[k].x, y.x); mx[k].y = fmaxf(mx[k].y, y.y); mx[k].z = fmaxf(mx[k].z, y.z); mx[k].w = fmaxf(mx[k].w, y.w);
    }
  }
  if (maxout == nullptr) return;
#pragma unroll
  for (int k = 0; k < V; ++k) *reinterpret_cast<float4*>(&smax[warp][k * 128 + lane * 4]) = mx[k];
  __syncthreads();
  for (int c = threadIdx.x; c < N; c += 256) {
    float m = smax[0][c];
#pragma unroll
    for (int w2 = 1; w2 < 8; ++w2) m = fmaxf(m, smax[w2][c]);
    atomicMax(reinterpret_cast<int*>(maxout + static_cast<long long>(cloud) * N + c), __float_as_int(m));
  }
}

// Per cloud: final_projection Linear(256 -> Co) + LayerNorm(Co) -> out[cloud, 0:Co]; DP3Encoder.state_mlp
// Linear(Ds, 64) + ReLU + Linear(64, 64) -> out[cloud, Co:Co+64] (skipped when state == nullptr). blockDim = 128.
__global__ void __launch_bounds__(128) pointnet_final_kernel(const float* __restrict__ feat, const float* __restrict__ Wp,
                                                             const float* __restrict__ bp, const float* __restrict__ gp,
                                                             const float* __restrict__ bep, int Co,
                                                             const float* __restrict__ state, int Ds,
                                                             const float* __restrict__ Ws1, const float* __restrict__ bs1,
                                                             const float* __restrict__ Ws2, const float* __restrict__ bs2,
                                                             float* __restrict__ out, int ld_out) {
  __shared__ float f[256], y[128], h[64], red[2];
  const int cloud = blockIdx.x, t = threadIdx.x;
  f[t] = feat[static_cast<long long>(cloud) * 256 + t];
  f[t + 128] = feat[static_cast<long long>(cloud) * 256 + 128 + t];
  __syncthreads();
  if (t < Co) {
    float a = __ldg(bp + t);
    for (int k = 0; k < 256; ++k) a = fmaf(f[k], __ldg(Wp + t * 256 + k), a);
    y[t] = a;
  }
  if (state != nullptr && t >= 64) {
    const int j = t - 64;
    float a = __ldg(bs1 + j);
    for (int k = 0; k < Ds; ++k) a = fmaf(__ldg(state + static_cast<long long>(cloud) * Ds + k), __ldg(Ws1 + j * Ds + k), a);
    h[j] = fmaxf(a, 0.f);
  }
  __syncthreads();
  if (t == 0) {
    float s = 0.f;
    for (int k = 0; k < Co; ++k) s += y[k];
    const float mean = s / Co;
    float q = 0.f;
    for (int k = 0; k < Co; ++k) q += (y[k] - mean) * (y[k] - mean);
    red[0] = mean; red[1] = rsqrtf(q / Co + 1e-5f);
  }
  __syncthreads();
  if (t < Co) out[static_cast<long long>(cloud) * ld_out + t] = (y[t] - red[0]) * red[1] * __ldg(gp + t) + __ldg(bep + t);
  if (state != nullptr && t >= 64) {
    const int j = t - 64;
    float a = __ldg(bs2 + j);
    for (int k = 0; k < 64; ++k) a = fmaf(h[k], __ldg(Ws2 + j * 64 + k), a);
    out[static_cast<long long>(cloud) * ld_out + Co + j] = a;
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// Diffusion Policy image encoder (robomimic ResNet18Conv with BatchNorm -> GroupNorm(C/16, C), SpatialSoftmax(32)):
// GroupNorm over (16 channels x H x W) per image on NHWC act_t maps. A map may be allocated one row / column larger
// than its valid extent (Hv, Wv) so that the next stride-2 conv sees even sizes; the extra row / column is kept at
// zero (= the conv's own zero padding) by gn2d_apply and excluded from the statistics.
//   gn2d_stats : partial[img][chunk][group] = (sum, sum of squares) over the chunk's valid pixels (deterministic)
//   gn2d_apply : y = relu?((x - mean) * rstd * gamma + beta (+ residual)), zero outside the valid extent
__global__ void __launch_bounds__(256) gn2d_stats_kernel(const act_t* __restrict__ in, float2* __restrict__ partial,
                                                         int H, int W, int Hv, int Wv, int C, int nchunk) {
  __shared__ float2 red[256];
  const int img = blockIdx.y, chunk = blockIdx.x, tid = threadIdx.x;
  const int octs = C >> 3, ppp = 256 / octs, oct = tid % octs, pslot = tid / octs;
  const int npix = Hv * Wv, per = (npix + nchunk - 1) / nchunk;
  const int p0 = chunk * per, p1 = min(p0 + per, npix);
  float s = 0.f, q = 0.f;
  if (pslot < ppp) {
    for (int pi = p0 + pslot; pi < p1; pi += ppp) {
      const int y = pi / Wv, x = pi - y * Wv;
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + ((static_cast<long long>(img) * H + y) * W + x) * C) + oct);
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = unpack_act2(u[k]);
        s += f.x + f.y;
        q += f.x * f.x + f.y * f.y;
      }
    }
  }
  red[tid] = make_float2(s, q);
  __syncthreads();
  const int G = C >> 4;
  if (tid < G) {
    float ss = 0.f, qq = 0.f;
    for (int ps = 0; ps < ppp; ++ps) {
      const float2 a = red[ps * octs + 2 * tid], b = red[ps * octs + 2 * tid + 1];
      ss += a.x + b.x; qq += a.y + b.y;
    }
    partial[(static_cast<long long>(img) * nchunk + chunk) * G + tid] = make_float2(ss, qq);
  }
}

__global__ void __launch_bounds__(256) gn2d_apply_kernel(const act_t* __restrict__ in, const float2* __restrict__ partial,
                                                         int nchunk, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, const act_t* __restrict__ res,
                                                         act_t* __restrict__ out, int H, int W, int Hv, int Wv, int C,
                                                         int relu, int napply) {
  __shared__ float2 st[32];                       // (mean, rstd) per group, G = C/16 <= 32
  const int img = blockIdx.y, tid = threadIdx.x, G = C >> 4;
  if (tid < G) {
    double s = 0, q = 0;
    for (int c = 0; c < nchunk; ++c) {
      const float2 a = partial[(static_cast<long long>(img) * nchunk + c) * G + tid];
      s += a.x; q += a.y;
    }
    const double n = static_cast<double>(Hv) * Wv * 16.0, mean = s / n;
    const double var = fmax(q / n - mean * mean, 0.0);
    st[tid] = make_float2(static_cast<float>(mean), static_cast<float>(1.0 / sqrt(var + 1e-5)));
  }
  __syncthreads();
  const int octs = C >> 3, total = H * W * octs, per = (total + napply - 1) / napply;
  const int i1 = min((blockIdx.x + 1) * per, total);
  const float floor_v = relu ? 0.0f : -INFINITY;
  for (int i = blockIdx.x * per + tid; i < i1; i += 256) {
    const int oct = i % octs, pix = i / octs, y = pix / W, x = pix - y * W;
    const long long off = ((static_cast<long long>(img) * H + y) * W + x) * C + oct * 8;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);
    if (y < Hv && x < Wv) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(in + off));
      const float2 ms = st[oct >> 1];
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + oct * 8)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + oct * 8 + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + oct * 8)), b1 = __ldg(reinterpret_cast<const float4*>(beta + oct * 8 + 4));
      const float ga[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w}, be[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
      uint32_t r[4] = {0u, 0u, 0u, 0u};
      if (res != nullptr) {
        const uint4 rv = __ldg(reinterpret_cast<const uint4*>(res + off));
        r[0] = rv.x; r[1] = rv.y; r[2] = rv.z; r[3] = rv.w;
      }
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = unpack_act2(u[k]), rr = unpack_act2(r[k]);
        const float a = fmaxf((f.x - ms.x) * ms.y * ga[2 * k] + be[2 * k] + rr.x, floor_v);
        const float b = fmaxf((f.y - ms.x) * ms.y * ga[2 * k + 1] + be[2 * k + 1] + rr.y, floor_v);
        w[k] = pack_act2(a, b);
      }
      o = make_uint4(w[0], w[1], w[2], w[3]);
    }
    *reinterpret_cast<uint4*>(out + off) = o;
  }
}

// float32 CHW image [n_img, 3, in_H, in_W] -> centre crop (off_y, off_x, Hv x Wv) -> padded NHWC4 act_t [H, W + 8] for the
// stem, zero outside the crop (H, W = crop rounded up to multiples of 32)
// (values copied as they are: the reference already scaled them to [-1, 1], RolloutDiffusionPolicy.py:121-123).
__global__ void preprocess_crop_f32_kernel(const float* __restrict__ in, act_t* __restrict__ out, int n_img, int in_H,
                                           int in_W, int off_y, int off_x, int Hv, int Wv, int H, int W) {
  const int quads_per_row = W / 4 + 2;
  const long long total = static_cast<long long>(n_img) * H * quads_per_row;
  const long long plane = static_cast<long long>(in_H) * in_W;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int q = static_cast<int>(idx % quads_per_row);
    const long long row = idx / quads_per_row;
    const long long img = row / H;
    const int y = static_cast<int>(row % H);
    uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    if (q >= 1 && q <= W / 4 && y < Hv) {
      const float* base = in + img * 3 * plane + static_cast<long long>(y + off_y) * in_W + off_x + (q - 1) * 4;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        if ((q - 1) * 4 + p < Wv) {
          w[2 * p] = pack_act2(__ldg(base + p), __ldg(base + plane + p));
          w[2 * p + 1] = pack_act2(__ldg(base + 2 * plane + p), 0.0f);
        }
      }
    }
    uint4* dst = reinterpret_cast<uint4*>(out + (row * (W + 2 * 4) + q * 4) * 4);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
  }
}

// robomimic SpatialSoftmax (num_kp K <= 32, temperature 1) on the 1x1-conv logits [img*H*W, ld] (fp32, first K columns),
// then Flatten + Linear(2K -> F): one CTA (64 threads) per image. Keypoint k = (E[x], E[y]) over softmax(H*W) with
// x, y in linspace(-1, 1). out[img * ld_out + j], j < F <= 64.
__global__ void __launch_bounds__(64) spatial_softmax_linear_kernel(const float* __restrict__ logits, int ld, int H, int W,
                                                                    int K, const float* __restrict__ Wl,
                                                                    const float* __restrict__ bl, int F,
                                                                    float* __restrict__ out, long long ld_out) {
  __shared__ float kp[64];
  const int img = blockIdx.x, t = threadIdx.x, HW = H * W;
  if (t < K) {
    const float* lg = logits + static_cast<long long>(img) * HW * ld + t;
    float m = -INFINITY;
    for (int i = 0; i < HW; ++i) m = fmaxf(m, lg[static_cast<long long>(i) * ld]);
    float s = 0.f, ex = 0.f, ey = 0.f;
    for (int i = 0; i < HW; ++i) {
      const float e = __expf(lg[static_cast<long long>(i) * ld] - m);
      const int y = i / W, x = i - y * W;
      const float px = (W > 1) ? -1.0f + 2.0f * x / (W - 1) : 0.f, py = (H > 1) ? -1.0f + 2.0f * y / (H - 1) : 0.f;
      s += e; ex += e * px; ey += e * py;
    }
    kp[2 * t] = ex / s;
    kp[2 * t + 1] = ey / s;
  }
  __syncthreads();
  if (t < F) {
    float a = __ldg(bl + t);
    for (int i = 0; i < 2 * K; ++i) a = fmaf(kp[i], __ldg(Wl + t * 2 * K + i), a);
    out[static_cast<long long>(img) * ld_out + t] = a;
  }
}

}  // namespace rmb
