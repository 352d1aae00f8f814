// Host-side planning of gemm_tc_kernel launches: tensor maps, tap tables, row tiling.
#pragma once
#include <algorithm>
#include <string>

#include "gemm_tc.cuh"
#include "tma_host.h"

namespace rmb {

struct GemmPlan {
  CUtensorMap ta, tb;
  GemmParams p;
  int block_n = 0, block_k = 0;
  long long flops = 0;   // 2*rows*N*K of useful work (for reporting)
};

inline int pick_block_n(int N) {
  if (N % 256 == 0) return 256;
  if (N % 128 == 0) return 128;
  if (N % 64 == 0) return 64;
  return 0;
}

// Choose the pixel box (bw x bh x bb <= 128 rows) that wastes the fewest MMA rows for an OW x OH x B output.
inline void pick_box(int OW, int OH, int B, int* bw, int* bh, int* bb) {
  double best = -1.0;
  int best_w = 1, best_h = 1, best_b = 1;
  for (int w = std::min(OW, 128); w >= 1; --w) {
    if (w < 8 && OW >= 8) break;
    for (int h = std::min(OH, 128 / w); h >= 1; --h) {
      const int bmax = std::min(B, 128 / (w * h));
      for (int b = bmax; b >= 1; --b) {
        const long long tw = (OW + w - 1) / w, th = (OH + h - 1) / h, tb = (B + b - 1) / b;
        const double eff = static_cast<double>(OW) * OH * B / (static_cast<double>(tw * th * tb) * 128.0);
        // prefer higher efficiency; ties -> wider boxes (longer contiguous runs per TMA row group)
        if (eff > best + 1e-9) { best = eff; best_w = w; best_h = h; best_b = b; }
      }
    }
  }
  *bw = best_w; *bh = best_h; *bb = best_b;
}

inline void fill_common(GemmParams& p, const float* bias, const void* residual, int ld_res, bool res_f32, void* out,
                        int ld_out, bool out_f32, bool relu) {
  p.bias = bias; p.residual = residual; p.ld_res = ld_res; p.res_f32 = res_f32 ? 1 : 0;
  p.out = out; p.ld_out = ld_out; p.out_f32 = out_f32 ? 1 : 0; p.relu = relu ? 1 : 0;
  p.remap_ncam = 0; p.remap_L = 0; p.remap_off = 0;
}

// out[M, N] = A[M, K] (row-major, leading dim lda) * Wt[N, K]^T
inline bool plan_linear(GemmPlan* g, const void* A, long long M, int K, int lda, const void* Wt, int N,
                        const float* bias, const void* residual, int ld_res, bool res_f32, void* out, int ld_out,
                        bool out_f32, bool relu, std::string* err) {
  *g = GemmPlan();
  g->block_k = 64;
  g->block_n = pick_block_n(N);
  if (g->block_n == 0 || K % 64 != 0) { if (err) *err = "plan_linear: N must be a multiple of 64 and K of 64"; return false; }
  GemmParams& p = g->p;
  p.box_w = 128; p.box_h = 1; p.box_b = 1;
  p.W = static_cast<int>(M); p.H = 1; p.B = 1;
  p.tiles_w = static_cast<int>((M + 127) / 128); p.tiles_h = 1; p.tiles_b = 1;
  p.n_col_tiles = N / g->block_n;
  p.num_taps = 1; p.chunks_per_tap = K / 64;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  fill_common(p, bias, residual, ld_res, res_f32, out, ld_out, out_f32, relu);
  const uint64_t dims[5] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M), 1, 1, 1};
  const uint64_t row = static_cast<uint64_t>(lda) * 2;
  const uint64_t strides[4] = {row, row * M, row * M, row * M};
  const uint32_t box[5] = {64, 128, 1, 1, 1};
  if (!make_tmap_act(&g->ta, A, 5, dims, strides, box, 128, err)) return false;
  const uint64_t wdims[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(N)};
  const uint64_t wstr[1] = {static_cast<uint64_t>(K) * 2};
  const uint32_t wbox[2] = {64, static_cast<uint32_t>(g->block_n)};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  g->flops = 2ll * M * N * K;
  return true;
}

// NHWC 16-bit conv, k x k (k = 1 or 3), stride 1 or 2, pad = k/2, as an implicit GEMM.
//   in [B, H, W, Cin], Wt [Cout, k*k*Cin] (tap-major, channel-minor), out rows = B*OH*OW, ld = Cout (or ld_out)
inline bool plan_conv(GemmPlan* g, const void* in, int B, int H, int W, int Cin, int ksize, int stride, const void* Wt,
                      int Cout, const float* bias, const void* residual, bool res_f32, void* out, int ld_out,
                      bool out_f32, bool relu, std::string* err) {
  *g = GemmPlan();
  g->block_k = 64;
  g->block_n = pick_block_n(Cout);
  if (g->block_n == 0 || Cin % 64 != 0) { if (err) *err = "plan_conv: Cout and Cin must be multiples of 64"; return false; }
  if ((ksize != 1 && ksize != 3) || (stride != 1 && stride != 2)) { if (err) *err = "plan_conv: unsupported k/stride"; return false; }
  if (stride == 2 && (H % 2 != 0 || W % 2 != 0)) { if (err) *err = "plan_conv: stride-2 conv needs even H and W"; return false; }
  const int pad = ksize / 2;
  const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
  GemmParams& p = g->p;
  pick_box(OW, OH, B, &p.box_w, &p.box_h, &p.box_b);
  p.W = OW; p.H = OH; p.B = B;
  p.tiles_w = (OW + p.box_w - 1) / p.box_w; p.tiles_h = (OH + p.box_h - 1) / p.box_h; p.tiles_b = (B + p.box_b - 1) / p.box_b;
  p.n_col_tiles = Cout / g->block_n;
  p.num_taps = ksize * ksize; p.chunks_per_tap = Cin / 64;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  for (int kh = 0; kh < ksize; ++kh)
    for (int kw = 0; kw < ksize; ++kw) {
      const int t = kh * ksize + kw;
      const int oh = kh - pad, ow = kw - pad;
      if (stride == 1) {
        p.tap_c0[t] = 0; p.tap_d1[t] = ow; p.tap_d2[t] = 0; p.tap_d3[t] = oh;
      } else {
        const int w2 = (ow < 0) ? -1 : 0, pw = ow - 2 * w2;   // w_in = 2*(ow_out + w2) + pw
        const int h2 = (oh < 0) ? -1 : 0, ph = oh - 2 * h2;
        p.tap_c0[t] = pw * Cin; p.tap_d1[t] = w2; p.tap_d2[t] = ph; p.tap_d3[t] = h2;
      }
    }
  fill_common(p, bias, residual, Cout, res_f32, out, ld_out, out_f32, relu);
  const uint64_t e = 2;
  if (stride == 1) {
    const uint64_t dims[5] = {(uint64_t)Cin, (uint64_t)W, 1, (uint64_t)H, (uint64_t)B};
    const uint64_t strides[4] = {Cin * e, (uint64_t)W * Cin * e, (uint64_t)W * Cin * e, (uint64_t)H * W * Cin * e};
    const uint32_t box[5] = {64, (uint32_t)p.box_w, 1, (uint32_t)p.box_h, (uint32_t)p.box_b};
    if (!make_tmap_act(&g->ta, in, 5, dims, strides, box, 128, err)) return false;
  } else {
    // phase view: [B, H/2, 2, W/2, 2*Cin]
    const uint64_t dims[5] = {(uint64_t)2 * Cin, (uint64_t)W / 2, 2, (uint64_t)H / 2, (uint64_t)B};
    const uint64_t strides[4] = {2 * Cin * e, (uint64_t)W * Cin * e, 2 * (uint64_t)W * Cin * e, (uint64_t)H * W * Cin * e};
    const uint32_t box[5] = {64, (uint32_t)p.box_w, 1, (uint32_t)p.box_h, (uint32_t)p.box_b};
    if (!make_tmap_act(&g->ta, in, 5, dims, strides, box, 128, err)) return false;
  }
  const uint64_t K = (uint64_t)ksize * ksize * Cin;
  const uint64_t wdims[2] = {K, (uint64_t)Cout};
  const uint64_t wstr[1] = {K * e};
  const uint32_t wbox[2] = {64, (uint32_t)g->block_n};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  g->flops = 2ll * B * OH * OW * Cout * (long long)K;
  return true;
}

// ResNet stem: 7x7 stride-2 pad-3 conv over the padded NHWC4 16-bit image written by preprocess_*_kernel
// ([B, H, W + 8, 4], 4 zero pixels either side). For output pixel (oh, ow) and filter row kh the 8 input pixels
// starting at padded column 2*ow are ONE contiguous 64-byte run = one 32-element K chunk (slot 0 carries a zero
// weight, slots 1..7 are kw = 0..6), so the A tile is a plain TMA box over an overlapping-row view of the image:
//   dims (32 elems, OW [stride 16 B], 2 [row parity], H/2, B).  Wt is [64, 7*32] with the zero slots baked in.
inline bool plan_conv1(GemmPlan* g, const void* in_padded, int B, int H, int W, const void* Wt, const float* bias,
                       void* out, std::string* err) {
  *g = GemmPlan();
  g->block_k = 32;
  g->block_n = 64;
  if (H % 2 != 0 || W % 4 != 0) { if (err) *err = "plan_conv1: H must be even and W a multiple of 4"; return false; }
  const int OH = H / 2, OW = W / 2, Wp = W + 8;
  GemmParams& p = g->p;
  pick_box(OW, OH, B, &p.box_w, &p.box_h, &p.box_b);
  p.W = OW; p.H = OH; p.B = B;
  p.tiles_w = (OW + p.box_w - 1) / p.box_w; p.tiles_h = (OH + p.box_h - 1) / p.box_h; p.tiles_b = (B + p.box_b - 1) / p.box_b;
  p.n_col_tiles = 1;
  p.num_taps = 7; p.chunks_per_tap = 1;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  for (int kh = 0; kh < 7; ++kh) {
    const int o = kh - 3;                                   // h_in = 2*oh + o = 2*(oh + h2) + ph
    const int h2 = (o >= 0) ? o / 2 : -((-o + 1) / 2);
    const int ph = o - 2 * h2;
    p.tap_d2[kh] = ph; p.tap_d3[kh] = h2;
  }
  fill_common(p, bias, nullptr, 0, false, out, 64, false, true);
  const uint64_t rowb = (uint64_t)Wp * 8;                    // bytes per padded image row
  const uint64_t dims[5] = {32, (uint64_t)OW, 2, (uint64_t)H / 2, (uint64_t)B};
  const uint64_t strides[4] = {16, rowb, 2 * rowb, (uint64_t)H * rowb};
  const uint32_t box[5] = {32, (uint32_t)p.box_w, 1, (uint32_t)p.box_h, (uint32_t)p.box_b};
  if (!make_tmap_act(&g->ta, in_padded, 5, dims, strides, box, 64, err)) return false;
  const uint64_t wdims[2] = {224, 64};
  const uint64_t wstr[1] = {224 * 2};
  const uint32_t wbox[2] = {32, 64};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 64, err)) return false;
  g->flops = 2ll * B * OH * OW * 64 * 147;
  return true;
}

template <int BN, int BK, int ST>
inline cudaError_t launch_gemm_variant(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  using S = GemmSmem<BN, BK, ST>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN, BK, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const long long tiles = 1ll * g.p.tiles_w * g.p.tiles_h * g.p.tiles_b * g.p.n_col_tiles;
  const int grid = static_cast<int>(std::min<long long>(tiles, num_sms));
  gemm_tc_kernel<BN, BK, ST><<<grid, kGemmThreads, S::kTotal, stream>>>(g.ta, g.tb, g.p);
  return cudaGetLastError();
}

inline cudaError_t launch_gemm(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  if (g.block_k == 32 && g.block_n == 64) return launch_gemm_variant<64, 32, 8>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 64) return launch_gemm_variant<64, 64, 6>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 128) return launch_gemm_variant<128, 64, 6>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 256) return launch_gemm_variant<256, 64, 4>(g, num_sms, stream);
  return cudaErrorInvalidValue;
}

}  // namespace rmb
