// Host-side planning of gemm_tc_kernel launches: tensor maps, tap tables, row tiling.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>

#include "conv_halo.cuh"
#include "stem_pool.cuh"
#include "gemm_tc.cuh"
#include "gemm_ln.cuh"
#include "launch.h"
#include "tma_host.h"

namespace rmb {

struct GemmPlan {
  CUtensorMap ta, tb, tout;   // tout: [rows, N] 16-bit output in 64-column x 128-row boxes (tma_out launches only)
  GemmParams p;
  int block_n = 0, block_k = 0;
  bool tma_out = false;  // epilogue stores through tout (plain matrices, 16-bit output, no residual, no row remap)
  bool pair = false;     // launch gemm_tc_pair_kernel (tb's box then holds block_n / 2 weight rows)
  bool ln = false;       // 512-wide tiles with the fused residual + LayerNorm epilogue (gemm_ln.cuh); always a pair launch
  long long flops = 0;   // 2*rows*N*K of useful work (for reporting)
};

// CTA pairs pay when the launch is throughput-bound (a tile for every SM): a pair loads 2/3 of the operand bytes per
// output tile, and the L2 -> SM path is what bounds the single-CTA 128 x 256 tile (48 KB per k-block against 524 MMA
// clocks; ncu: 61 B/clk/SM ingest with the tensor pipe 47 % busy). RMB_GEMM_NO_PAIR=1 keeps single-CTA tiles (A/B runs).
inline bool gemm_use_pair(const GemmParams& p) {
  static const bool off = [] { const char* e = getenv("RMB_GEMM_NO_PAIR"); return e != nullptr && e[0] == '1'; }();
  const long long row_tiles = 1ll * p.tiles_w * p.tiles_h * p.tiles_b;
  return !off && (row_tiles % 2 == 0 || row_tiles >= 32) && row_tiles * p.n_col_tiles >= 148;
}

// Column-tile width. Wide tiles re-read the A operand least, so they win whenever there is a tile for every SM; a
// launch with fewer tiles than SMs (batch-1 latency: 5 row tiles) is bound by the serial k-loop of its few CTAs, and
// narrower tiles spread the same work over up to 4x as many SMs (K = 3200 FFN-down at batch 1: 31 us with 10 tiles).
inline int pick_block_n(int N, long long row_tiles = 1 << 20) {
  const int cand[3] = {256, 128, 64};
  int smallest = 0;
  for (int bn : cand) {
    if (N % bn != 0) continue;
    if (row_tiles * (N / bn) >= 148) return bn;
    smallest = bn;
  }
  return smallest;
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
  p.n_valid = 1 << 30;
  p.film = nullptr; p.film_ld = 0; p.film_g_off = 0; p.film_b_off = 0; p.film_div = 1;
  p.ln_gamma = nullptr; p.ln_beta = nullptr; p.ln_pos = nullptr; p.ln_L = 1; p.ln_y_act = nullptr; p.ln_ypos_act = nullptr;
  p.ln_eps = 1e-5f;
}

// out[M, N] = A[M, K] (row-major, leading dim lda) * Wt[N, K]^T
inline bool plan_linear(GemmPlan* g, const void* A, long long M, int K, int lda, const void* Wt, int N,
                        const float* bias, const void* residual, int ld_res, bool res_f32, void* out, int ld_out,
                        bool out_f32, bool relu, std::string* err) {
  *g = GemmPlan();
  g->block_k = 64;
  g->block_n = pick_block_n(N, (M + 127) / 128);
  // wide N that 256 does not divide (dim_feedforward 3200): 256-wide tiles with a ragged last one beat 128-wide tiles,
  // whose operand bytes per flop are 4/3 higher; the weight rows past N are TMA zero fill, their columns are not stored
  const bool ragged = N >= 1024 && N % 256 != 0 && N % 64 == 0 && ((M + 127) / 128) * ((N + 255) / 256) >= 148;
  if (ragged) g->block_n = 256;
  if (g->block_n == 0 || K % 64 != 0) { if (err) *err = "plan_linear: N must be a multiple of 64 and K of 64"; return false; }
  GemmParams& p = g->p;
  p.box_w = 128; p.box_h = 1; p.box_b = 1;
  p.W = static_cast<int>(M); p.H = 1; p.B = 1;
  p.tiles_w = static_cast<int>((M + 127) / 128); p.tiles_h = 1; p.tiles_b = 1;
  p.n_col_tiles = (N + g->block_n - 1) / g->block_n;
  p.num_taps = 1; p.chunks_per_tap = K / 64;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  fill_common(p, bias, residual, ld_res, res_f32, out, ld_out, out_f32, relu);
  if (ragged) p.n_valid = N;
  const uint64_t dims[5] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M), 1, 1, 1};
  const uint64_t row = static_cast<uint64_t>(lda) * 2;
  const uint64_t strides[4] = {row, row * M, row * M, row * M};
  const uint32_t box[5] = {64, 128, 1, 1, 1};
  if (!make_tmap_act(&g->ta, A, 5, dims, strides, box, 128, err)) return false;
  const uint64_t wdims[2] = {static_cast<uint64_t>(K), static_cast<uint64_t>(N)};
  const uint64_t wstr[1] = {static_cast<uint64_t>(K) * 2};
  g->pair = gemm_use_pair(p);
  const uint32_t wbox[2] = {64, static_cast<uint32_t>(g->pair ? g->block_n / 2 : g->block_n)};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  static const bool no_tma_out = [] { const char* e = getenv("RMB_GEMM_NO_TMA_OUT"); return e != nullptr && e[0] == '1'; }();
  if (!out_f32 && residual == nullptr && !no_tma_out && M >= 128) {
    const uint64_t odims[2] = {static_cast<uint64_t>(N), static_cast<uint64_t>(M)};
    const uint64_t ostr[1] = {static_cast<uint64_t>(ld_out) * 2};
    const uint32_t obox[2] = {64, 128};
    if (!make_tmap_act(&g->tout, out, 2, odims, ostr, obox, 128, err)) return false;
    g->tma_out = true;
  }
  g->flops = 2ll * M * N * K;
  return true;
}

// y = LayerNorm(A[M, K] * Wt[512, K]^T + bias + residual) with the outputs of layernorm512_kernel, in one launch
// (gemm_ln.cuh). residual / out are fp32 [M, 512]; y_act / ypos_act (optional) 16-bit [M, 512].
inline bool plan_linear_ln(GemmPlan* g, const void* A, long long M, int K, int lda, const void* Wt, const float* bias,
                           const float* residual, float* out, const float* gamma, const float* beta, const float* pos,
                           int L, void* y_act, void* ypos_act, float eps, std::string* err) {
  *g = GemmPlan();
  g->block_k = 64; g->block_n = 512; g->pair = true; g->ln = true;
  if (K % 64 != 0 || M < 1) { if (err) *err = "plan_linear_ln: K must be a multiple of 64"; return false; }
  if (!residual || !out || !gamma || !beta || (ypos_act && (!pos || L < 1))) { if (err) *err = "plan_linear_ln: null argument"; return false; }
  GemmParams& p = g->p;
  p.box_w = 128; p.box_h = 1; p.box_b = 1;
  p.W = static_cast<int>(M); p.H = 1; p.B = 1;
  p.tiles_w = static_cast<int>((M + 127) / 128); p.tiles_h = 1; p.tiles_b = 1;
  p.n_col_tiles = 1;
  p.num_taps = 1; p.chunks_per_tap = K / 64;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
  fill_common(p, bias, residual, 512, true, out, 512, true, false);
  p.ln_gamma = gamma; p.ln_beta = beta; p.ln_pos = pos; p.ln_L = L < 1 ? 1 : L;
  p.ln_y_act = y_act; p.ln_ypos_act = ypos_act; p.ln_eps = eps;
  const uint64_t dims[5] = {static_cast<uint64_t>(K), static_cast<uint64_t>(M), 1, 1, 1};
  const uint64_t row = static_cast<uint64_t>(lda) * 2;
  const uint64_t strides[4] = {row, row * M, row * M, row * M};
  const uint32_t box[5] = {64, 128, 1, 1, 1};
  if (!make_tmap_act(&g->ta, A, 5, dims, strides, box, 128, err)) return false;
  const uint64_t wdims[2] = {static_cast<uint64_t>(K), 512};
  const uint64_t wstr[1] = {static_cast<uint64_t>(K) * 2};
  const uint32_t wbox[2] = {64, 128};       // a CTA holds two 128-row boxes per k-block (one per N = 256 half)
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  g->flops = 2ll * M * 512 * K;
  return true;
}

// NHWC 16-bit conv, k x k (k = 1 or 3), stride 1 or 2, pad = k/2, as an implicit GEMM.
//   in [B, H, W, Cin], Wt [Cout, k*k*Cin] (tap-major, channel-minor), out rows = B*OH*OW, ld = Cout (or ld_out)
// side_cout > 0 (3x3 stride-2 convs only): Wt is [Cout + side_cout, 9*Cin] whose rows >= Cout are a 1x1 stride-2 conv on
// the same input (non-zero in the centre tap's columns only), bias is [Cout + side_cout]; those columns go to out2
// (GemmParams::n_main_ct).
inline bool plan_conv(GemmPlan* g, const void* in, int B, int H, int W, int Cin, int ksize, int stride, const void* Wt,
                      int Cout, const float* bias, const void* residual, bool res_f32, void* out, int ld_out,
                      bool out_f32, bool relu, std::string* err, int side_cout = 0, void* out2 = nullptr, int ld_out2 = 0,
                      bool relu2 = false) {
  *g = GemmPlan();
  g->block_k = 64;
  if (Cout % 64 != 0 || Cin % 64 != 0) { if (err) *err = "plan_conv: Cout and Cin must be multiples of 64"; return false; }
  if ((ksize != 1 && ksize != 3) || (stride != 1 && stride != 2)) { if (err) *err = "plan_conv: unsupported k/stride"; return false; }
  if (stride == 2 && (H % 2 != 0 || W % 2 != 0)) { if (err) *err = "plan_conv: stride-2 conv needs even H and W"; return false; }
  const int pad = ksize / 2;
  const int OH = (H + 2 * pad - ksize) / stride + 1, OW = (W + 2 * pad - ksize) / stride + 1;
  GemmParams& p = g->p;
  pick_box(OW, OH, B, &p.box_w, &p.box_h, &p.box_b);
  p.W = OW; p.H = OH; p.B = B;
  p.tiles_w = (OW + p.box_w - 1) / p.box_w; p.tiles_h = (OH + p.box_h - 1) / p.box_h; p.tiles_b = (B + p.box_b - 1) / p.box_b;
  g->block_n = pick_block_n(Cout, 1ll * p.tiles_w * p.tiles_h * p.tiles_b);
  p.n_col_tiles = Cout / g->block_n;
  if (side_cout > 0) {
    if (ksize != 3 || stride != 2 || residual != nullptr || out_f32 || out2 == nullptr || side_cout % g->block_n != 0) {
      if (err) *err = "plan_conv: side columns need a 3x3 stride-2 conv without residual, a 16-bit output and side_cout % block_n == 0";
      return false;
    }
    p.n_main_ct = p.n_col_tiles;
    p.n_col_tiles += side_cout / g->block_n;
    p.out2 = out2; p.ld_out2 = ld_out2; p.relu2 = relu2 ? 1 : 0;
  }
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
  const uint64_t wdims[2] = {K, (uint64_t)(Cout + side_cout)};
  const uint64_t wstr[1] = {K * e};
  g->pair = gemm_use_pair(p);
  // halo3 (gemm_tc.cuh, GemmParams::halo3): 3x3 stride-1 convs with 128-wide column tiles launched as CTA pairs (the
  // ResNet layer2 convs). Needs a w x h = 128-pixel box with w % 8 == 0 inside one image that tiles the map at least
  // as well as the generic box; a stage is the (h + 2)-row input box + three half-tiles of weights, and the launch
  // uses as many stages as fit in the variant's shared memory (4 of 44 KB for the 16 x 8 box). RMB_GEMM_NO_HALO3=1: off.
  static const bool halo_off = [] { const char* ev = getenv("RMB_GEMM_NO_HALO3"); return ev != nullptr && ev[0] == '1'; }();
  if (!halo_off && ksize == 3 && stride == 1 && g->block_n == 128 && g->pair) {
    const double eff0 = static_cast<double>(OW) * OH * B / (128.0 * p.tiles_w * p.tiles_h * p.tiles_b);
    int bw = 0, bh = 0;
    double best = 0.0;
    for (int w : {8, 16, 32, 64}) {                        // taller boxes first: (h + 2) / h rows are fetched per tile
      const int h = 128 / w;
      if (w > OW + 7 || h > OH + 7) continue;
      const long long tw = (OW + w - 1) / w, th = (OH + h - 1) / h;
      const double eff = static_cast<double>(OW) * OH / (128.0 * tw * th);
      if (eff > best + 1e-9) { best = eff; bw = w; bh = h; }
    }
    const int a_bytes = (bh + 2) * bw * 128, stage_bytes = a_bytes + 3 * (g->block_n / 2) * 128;
    const int stages = std::min(8, (8 * (128 * 128 + (g->block_n / 2) * 128)) / std::max(stage_bytes, 1));   // GemmSmem<128, 64, 8, true>
    if (bw > 0 && best >= eff0 - 0.07 && stages >= 3) {
      p.box_w = bw; p.box_h = bh; p.box_b = 1;
      p.tiles_w = (OW + bw - 1) / bw; p.tiles_h = (OH + bh - 1) / bh; p.tiles_b = B;
      p.halo3 = 1; p.halo_stages = stages; p.halo_a_bytes = a_bytes;
      const uint64_t dims[5] = {(uint64_t)Cin, (uint64_t)W, 1, (uint64_t)H, (uint64_t)B};
      const uint64_t strides[4] = {Cin * e, (uint64_t)W * Cin * e, (uint64_t)W * Cin * e, (uint64_t)H * W * Cin * e};
      const uint32_t box[5] = {64, (uint32_t)bw, 1, (uint32_t)(bh + 2), 1};
      if (!make_tmap_act(&g->ta, in, 5, dims, strides, box, 128, err)) return false;
      g->pair = gemm_use_pair(p);
      if (!g->pair) { if (err) *err = "plan_conv: halo3 box lost the pair launch"; return false; }
    }
  }
  const uint32_t wbox[2] = {64, (uint32_t)(g->pair ? g->block_n / 2 : g->block_n)};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  // 1x1 convs with a 16-bit NHWC output and no residual (the stride-2 downsample convs: one or a few k-blocks per tile,
  // bound by the epilogue): the TMA-store epilogue (EPI 3) over a rank-5 output map. Measured at 256 episodes: 0.252 /
  // 0.134 / 0.080 -> 0.213 / 0.106 / 0.066 ms; on the 3x3 stride-2 conv1 launches the same epilogue measured SLOWER
  // (layer2.0.conv1 0.46 -> 0.53 ms: boxes of two images, 2 KB runs), so they keep the staged-transpose epilogue.
  // RMB_GEMM_NO_TMA_OUT=1 keeps the plain epilogue everywhere.
  static const bool no_tma_out = [] { const char* ev = getenv("RMB_GEMM_NO_TMA_OUT"); return ev != nullptr && ev[0] == '1'; }();
  if (!no_tma_out && ksize == 1 && !out_f32 && residual == nullptr && side_cout == 0 && ld_out == Cout &&
      p.box_w * p.box_h * p.box_b <= 128 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
    const uint64_t odims[5] = {(uint64_t)Cout, (uint64_t)OW, 1, (uint64_t)OH, (uint64_t)B};
    const uint64_t ostr[4] = {Cout * e, (uint64_t)OW * Cout * e, (uint64_t)OW * Cout * e, (uint64_t)OH * OW * Cout * e};
    const uint32_t obox[5] = {64, (uint32_t)p.box_w, 1, (uint32_t)p.box_h, (uint32_t)p.box_b};
    if (!make_tmap_act(&g->tout, out, 5, odims, ostr, obox, 128, err)) return false;
    g->tma_out = true;
    p.out_5d = 1;
  }
  g->flops = 2ll * B * OH * OW * Cout * (long long)K + 2ll * B * OH * OW * side_cout * (long long)Cin;
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

// grid of a persistent GEMM launch: one CTA per SM, or one cluster of two per co-schedulable SM pair
inline int gemm_grid(const GemmPlan& g, int num_sms, int max_pairs) {
  const long long row_tiles = 1ll * g.p.tiles_w * g.p.tiles_h * g.p.tiles_b;
  if (!g.pair) return static_cast<int>(std::min<long long>(row_tiles * g.p.n_col_tiles, num_sms));
  return 2 * static_cast<int>(std::min<long long>((row_tiles + 1) / 2 * g.p.n_col_tiles, max_pairs));
}

// How many 2-CTA clusters of `kern` the device can hold at once (a GPC with an odd number of usable SMs leaves one out):
// the persistent pair kernels must not launch more, or the surplus clusters would run as a second wave.
template <typename K>
inline int max_active_pairs(K kern, int smem_bytes, int num_sms) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(num_sms & ~1), 1, 1);
  cfg.blockDim = dim3(kGemmThreads, 1, 1);
  cfg.dynamicSmemBytes = static_cast<size_t>(smem_bytes);
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr; cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) { (void)cudaGetLastError(); n = num_sms / 2; }
  if (getenv("RMB_VERBOSE")) fprintf(stderr, "[rmb] co-resident 2-CTA clusters: %d (%d SMs, %d B smem)\n", n, num_sms, smem_bytes);
  return std::min(n, num_sms / 2);
}

template <int BN, int BK, int ST, bool OUT_F32, int RES, bool PAIR, int EPI = 0>
inline cudaError_t launch_gemm_instance(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  using S = GemmSmem<BN, BK, ST, PAIR, EPI>;
  static_assert(S::kTotal <= 232448, "shared memory");
  auto kern = PAIR ? gemm_tc_pair_kernel<BN, BK, ST, OUT_F32, RES, EPI> : gemm_tc_kernel<BN, BK, ST, OUT_F32, RES, EPI>;
  static bool attr_set = false;
  static int max_pairs = 0;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
    if (e != cudaSuccess) return e;
    max_pairs = PAIR ? max_active_pairs(kern, S::kTotal, num_sms) : 0;
    attr_set = true;
  }
  return launch_pdl(kern, dim3(gemm_grid(g, num_sms, max_pairs)), dim3(kGemmThreads), S::kTotal, stream, g.ta, g.tb,
                    g.tma_out ? g.tout : g.ta, g.p);
}

// epilogue modes in use: act out (+ act residual) for the convs and QKV / FFN-up projections; fp32 out (+ fp32
// residual) for the transformer's residual stream
template <int BN, int BK, int ST, bool PAIR>
inline cudaError_t launch_gemm_variant(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  const bool has_res = g.p.residual != nullptr;
  if (!g.p.out_f32) {
    if constexpr (BK == 64) {
      if (!has_res && g.tma_out) return launch_gemm_instance<BN, BK, ST, false, 0, PAIR, 3>(g, num_sms, stream);
    }
    if (!has_res) return launch_gemm_instance<BN, BK, ST, false, 0, PAIR>(g, num_sms, stream);
    if (!g.p.res_f32) {
      if constexpr (BK == 64) {
        if (g.p.film != nullptr) return launch_gemm_instance<BN, BK, ST, false, 1, PAIR, 4>(g, num_sms, stream);
      }
      return launch_gemm_instance<BN, BK, ST, false, 1, PAIR>(g, num_sms, stream);
    }
  } else {
    if (!has_res) return launch_gemm_instance<BN, BK, ST, true, 0, PAIR>(g, num_sms, stream);
    if (g.p.res_f32) return launch_gemm_instance<BN, BK, ST, true, 2, PAIR>(g, num_sms, stream);
    return launch_gemm_instance<BN, BK, ST, true, 1, PAIR>(g, num_sms, stream);
  }
  return cudaErrorInvalidValue;   // act out + fp32 residual: not needed on this path
}

// the LayerNorm-fused mode exists as a pair kernel only (launch_gemm_instance would also name the single-CTA kernel)
inline cudaError_t launch_gemm_ln(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  using S = GemmSmem<512, 64, 4, true, 5>;
  static_assert(S::kTotal <= 232448, "shared memory");
  auto kern = gemm_tc_pair_kernel<512, 64, 4, true, 2, 5>;
  static bool attr_set = false;
  static int max_pairs = 0;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kTotal);
    if (e != cudaSuccess) return e;
    max_pairs = max_active_pairs(kern, S::kTotal, num_sms);
    attr_set = true;
  }
  return launch_pdl(kern, dim3(gemm_grid(g, num_sms, max_pairs)), dim3(kGemmThreads), S::kTotal, stream, g.ta, g.tb, g.ta, g.p);
}

inline cudaError_t launch_gemm(const GemmPlan& g, int num_sms, cudaStream_t stream) {
  if (g.ln) return launch_gemm_ln(g, num_sms, stream);
  if (g.pair) {
    if (g.block_k == 64 && g.block_n == 64) return launch_gemm_variant<64, 64, 8, true>(g, num_sms, stream);
    if (g.block_k == 64 && g.block_n == 128) return launch_gemm_variant<128, 64, 8, true>(g, num_sms, stream);
    if (g.block_k == 64 && g.block_n == 256) return launch_gemm_variant<256, 64, 6, true>(g, num_sms, stream);
    return cudaErrorInvalidValue;
  }
  if (g.block_k == 32 && g.block_n == 64) return launch_gemm_variant<64, 32, 8, false>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 64) return launch_gemm_variant<64, 64, 6, false>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 128) return launch_gemm_variant<128, 64, 6, false>(g, num_sms, stream);
  if (g.block_k == 64 && g.block_n == 256) return launch_gemm_variant<256, 64, 4, false>(g, num_sms, stream);
  return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------------------------------------------------------------
// Halo-reuse convs (conv_halo.cuh)
struct HaloPlan {
  CUtensorMap ta, tb, tout, tres;   // input boxes, weights, output tile store, residual tile load
  HaloParams hp;
  int block_k = 0;
  int smem = 0;
  long long flops = 0;
};

constexpr int kMaxDynSmem = 232448;   // 227 KB per CTA on sm_100

// staging tiles per epilogue group of the halo convs: which = 0 stem, 1 layer1. RMB_HALO_OUT_BUFS="<stem><layer1>" (digits 1/2).
inline int halo_out_bufs(int which) {
  const char* e = getenv("RMB_HALO_OUT_BUFS");
  if (e != nullptr && strlen(e) >= 2 && (e[which] == '1' || e[which] == '2')) return e[which] - '0';
  (void)which;
  return 2;   // measured: stem 2.15 -> 1.56-1.74 ms, layer1 neutral
}

// pixel box for a halo conv: box_w a multiple of 8 (tap shifts must be whole 8-row swizzle atoms), 128 rows per tile
inline void pick_halo_box(int OW, int OH, int halo_rows, int* bw, int* bh) {
  double best = -1.0; long long best_bytes = 0;
  for (int w = 8; w <= 128; w *= 2) {
    const int h = 128 / w;
    const long long tw = (OW + w - 1) / w, th = (OH + h - 1) / h;
    const double eff = static_cast<double>(OW) * OH / (static_cast<double>(tw * th) * 128.0);
    const long long bytes = static_cast<long long>(h + halo_rows) * w;
    if (eff > best + 1e-9 || (eff > best - 1e-9 && bytes < best_bytes)) { best = eff; best_bytes = bytes; *bw = w; *bh = h; }
  }
}

inline void halo_fill_tiles(GemmParams& p, int OW, int OH, int B, int bw, int bh) {
  p.box_w = bw; p.box_h = bh; p.box_b = 1;
  p.W = OW; p.H = OH; p.B = B;
  p.tiles_w = (OW + bw - 1) / bw; p.tiles_h = (OH + bh - 1) / bh; p.tiles_b = B;
  p.n_col_tiles = 1; p.num_taps = 0; p.chunks_per_tap = 0;
  for (int t = 0; t < kMaxTaps; ++t) { p.tap_c0[t] = p.tap_d1[t] = p.tap_d2[t] = p.tap_d3[t] = 0; }
}

inline bool conv3x3_halo_eligible(int W, int Cin, int Cout, int ksize, int stride) {
  return Cin == 64 && Cout == 64 && ksize == 3 && stride == 1 && W >= 8;
}

// 3x3 stride-1 pad-1 conv, Cin = Cout = 64, NHWC act_t; Wt [64, 9*64] tap-major; residual (optional) act_t [rows, 64]
inline bool plan_conv3x3_halo(HaloPlan* g, const void* in, int B, int H, int W, const void* Wt, const float* bias,
                              const void* residual, void* out, bool relu, std::string* err) {
  *g = HaloPlan();
  g->block_k = 64;
  HaloParams& hp = g->hp;
  memset(&hp, 0, sizeof(hp));
  int bw = 8, bh = 16;
  pick_halo_box(W, H, 2, &bw, &bh);
  halo_fill_tiles(hp.g, W, H, B, bw, bh);
  fill_common(hp.g, bias, residual, 64, false, out, 64, false, relu);
  hp.n_loads = 3; hp.n_taps = 9;
  hp.box_bytes = static_cast<uint32_t>((bh + 2) * bw * 128);
  hp.stage_bytes = hp.box_bytes;                      // bw % 8 == 0 -> already a multiple of 1024
  for (int kw = 0; kw < 3; ++kw) {
    hp.ld_c0[kw] = 0; hp.ld_d1[kw] = kw - 1; hp.ld_d2[kw] = 0; hp.ld_d3[kw] = -1;
    hp.ld_tap0[kw] = 3 * kw; hp.ld_ntaps[kw] = 3;
    for (int kh = 0; kh < 3; ++kh) {
      hp.tap_aoff[3 * kw + kh] = static_cast<uint32_t>(kh * bw * 128);
      hp.tap_w[3 * kw + kh] = kh * 3 + kw;
    }
  }
  hp.out_bufs = halo_out_bufs(1);
  const int fixed = 9 * HaloSmem<64>::kWTapBytes + 2 * hp.out_bufs * HaloSmem<64>::kStagingTileBytes + HaloSmem<64>::kBarrierBytes + 1024;
  hp.n_stages = std::min<int>(kHaloMaxStages, (kMaxDynSmem - fixed) / static_cast<int>(hp.stage_bytes));
  if (hp.n_stages < 2) { if (err) *err = "plan_conv3x3_halo: tile does not fit shared memory"; return false; }
  g->smem = HaloSmem<64>::total(9, hp.n_stages, hp.stage_bytes, hp.out_bufs);
  const uint64_t e = 2, Cin = 64;
  const uint64_t dims[5] = {Cin, (uint64_t)W, 1, (uint64_t)H, (uint64_t)B};
  const uint64_t strides[4] = {Cin * e, (uint64_t)W * Cin * e, (uint64_t)W * Cin * e, (uint64_t)H * W * Cin * e};
  const uint32_t box[5] = {64, (uint32_t)bw, 1, (uint32_t)(bh + 2), 1};
  if (!make_tmap_act(&g->ta, in, 5, dims, strides, box, 128, err)) return false;
  const uint64_t wdims[2] = {576, 64};
  const uint64_t wstr[1] = {576 * e};
  const uint32_t wbox[2] = {64, 64};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 128, err)) return false;
  const uint32_t obox[5] = {64, (uint32_t)bw, 1, (uint32_t)bh, 1};
  if (!make_tmap_act(&g->tout, out, 5, dims, strides, obox, 128, err)) return false;
  if (!make_tmap_act(&g->tres, residual ? residual : out, 5, dims, strides, obox, 128, err)) return false;
  g->flops = 2ll * B * H * W * 64 * 576;
  return true;
}

// ResNet stem 7x7/2 pad 3 on the padded NHWC4 image (see plan_conv1 for the overlapping-window view):
// box 0 = odd input rows (filter rows kh = 0,2,4,6), box 1 = even input rows (kh = 1,3,5).
inline bool plan_conv1_halo(HaloPlan* g, const void* in_padded, int B, int H, int W, const void* Wt, const float* bias,
                            void* out, std::string* err) {
  *g = HaloPlan();
  g->block_k = 32;
  if (H % 2 != 0 || W % 4 != 0) { if (err) *err = "plan_conv1_halo: H must be even and W a multiple of 4"; return false; }
  const int OH = H / 2, OW = W / 2, Wp = W + 8;
  HaloParams& hp = g->hp;
  memset(&hp, 0, sizeof(hp));
  int bw = 8, bh = 16;
  pick_halo_box(OW, OH, 3, &bw, &bh);
  halo_fill_tiles(hp.g, OW, OH, B, bw, bh);
  fill_common(hp.g, bias, nullptr, 0, false, out, 64, false, true);
  hp.n_loads = 2; hp.n_taps = 7;
  hp.box_bytes = static_cast<uint32_t>((bh + 3) * bw * 64);
  hp.stage_bytes = (hp.box_bytes + 1023u) & ~1023u;
  // input row of filter row kh for output row oh: 2*oh + kh - 3 = 2*(oh + h2) + ph
  int t = 0;
  for (int ph = 1; ph >= 0; --ph) {
    const int l = 1 - ph;
    const int d3 = (ph == 1) ? -2 : -1;               // first H/2-row of the box relative to the tile's first output row
    hp.ld_c0[l] = 0; hp.ld_d1[l] = 0; hp.ld_d2[l] = ph; hp.ld_d3[l] = d3;
    hp.ld_tap0[l] = t; hp.ld_ntaps[l] = 0;
    for (int kh = 0; kh < 7; ++kh) {
      const int o = kh - 3;
      const int h2 = (o >= 0) ? o / 2 : -((-o + 1) / 2);
      if (o - 2 * h2 != ph) continue;
      hp.tap_aoff[t] = static_cast<uint32_t>((h2 - d3) * bw * 64);
      hp.tap_w[t] = kh;
      ++t; ++hp.ld_ntaps[l];
    }
  }
  hp.out_bufs = halo_out_bufs(0);
  const int fixed = 7 * HaloSmem<32>::kWTapBytes + 2 * hp.out_bufs * HaloSmem<32>::kStagingTileBytes + HaloSmem<32>::kBarrierBytes + 1024;
  hp.n_stages = std::min<int>(kHaloMaxStages, (kMaxDynSmem - fixed) / static_cast<int>(hp.stage_bytes));
  if (hp.n_stages < 2) { if (err) *err = "plan_conv1_halo: tile does not fit shared memory"; return false; }
  g->smem = HaloSmem<32>::total(7, hp.n_stages, hp.stage_bytes, hp.out_bufs);
  const uint64_t rowb = (uint64_t)Wp * 8;                    // bytes per padded image row
  const uint64_t dims[5] = {32, (uint64_t)OW, 2, (uint64_t)H / 2, (uint64_t)B};
  const uint64_t strides[4] = {16, rowb, 2 * rowb, (uint64_t)H * rowb};
  const uint32_t box[5] = {32, (uint32_t)bw, 1, (uint32_t)(bh + 3), 1};
  if (!make_tmap_act(&g->ta, in_padded, 5, dims, strides, box, 64, err)) return false;
  const uint64_t wdims[2] = {224, 64};
  const uint64_t wstr[1] = {224 * 2};
  const uint32_t wbox[2] = {32, 64};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 64, err)) return false;
  const uint64_t odims[5] = {64, (uint64_t)OW, 1, (uint64_t)OH, (uint64_t)B};
  const uint64_t ostr[4] = {128, (uint64_t)OW * 128, (uint64_t)OW * 128, (uint64_t)OH * OW * 128};
  const uint32_t obox[5] = {64, (uint32_t)bw, 1, (uint32_t)bh, 1};
  if (!make_tmap_act(&g->tout, out, 5, odims, ostr, obox, 128, err)) return false;
  g->tres = g->tout;
  g->flops = 2ll * B * OH * OW * 64 * 147;
  return true;
}

// ---------------------------------------------------------------------------------------------------------------------
// Fused stem + max-pool (stem_pool.cuh)
struct StemPoolPlan {
  CUtensorMap ta, tb;
  StemPoolParams p;
  int smem = 0, grid = 0;
  long long flops = 0;
};

// in_padded: the NHWC4 image of preprocess_*_kernel [B, H, W + 8, 4]; out: pooled [B, H/4, W/4, 64]
inline bool plan_stem_pool(StemPoolPlan* g, const void* in_padded, int B, int H, int W, const void* Wt, const float* bias,
                           void* out, int num_sms, std::string* err) {
  *g = StemPoolPlan();
  if (H % 8 != 0 || W % 4 != 0 || B < 1) { if (err) *err = "plan_stem_pool: H must be a multiple of 8 and W of 4"; return false; }
  const int OH = H / 2, OW = W / 2, Wp = W + 8;
  StemPoolParams& p = g->p;
  memset(&p, 0, sizeof(p));
  p.n_img = B; p.OH = OH; p.OW = OW;
  p.strips = (OW + kSpBoxW - 1) / kSpBoxW;
  const int TR = OH / kSpBoxH;
  // bands per image: a divisor of the tile rows; more bands = better balance over the 2 x SMs epilogue groups, fewer =
  // less warm-up recomputation (one tile row per band that does not start at the image top)
  double best = 1e30; int best_nbi = 0;
  for (int nbi = 1; nbi <= TR; ++nbi) {
    if (TR % nbi != 0 || TR / nbi > kSpMaxBandRows) continue;
    const long long bands = 1ll * B * nbi;
    const long long groups = 2 * std::min<long long>(num_sms, (bands + 1) / 2);
    const long long rounds = (bands + groups - 1) / groups;
    const double rows = TR / nbi + (nbi > 1 ? static_cast<double>(nbi - 1) / nbi : 0.0);
    const double cost = static_cast<double>(rounds) * rows;
    if (cost < best - 1e-9) { best = cost; best_nbi = nbi; }
  }
  if (best_nbi == 0) { if (err) *err = "plan_stem_pool: no band split fits"; return false; }
  p.bands_per_img = best_nbi; p.rows_band = TR / best_nbi; p.n_bands = B * best_nbi;
  p.bias = bias; p.out = static_cast<act_t*>(out);
  const int fixed = kSpWBytes + 2 * kSpGroupBytes + 256 + 1024;
  p.n_stages = std::min<int>(kSpMaxStages, (kMaxDynSmem - fixed) / kSpBoxBytes);
  if (p.n_stages < 2) { if (err) *err = "plan_stem_pool: shared memory"; return false; }
  g->smem = fixed + p.n_stages * kSpBoxBytes;
  g->grid = static_cast<int>(std::min<long long>(num_sms, (1ll * p.n_bands + 1) / 2));
  const uint64_t rowb = (uint64_t)Wp * 8;                    // bytes per padded image row
  const uint64_t dims[5] = {32, (uint64_t)OW, 2, (uint64_t)H / 2, (uint64_t)B};
  const uint64_t strides[4] = {16, rowb, 2 * rowb, (uint64_t)H * rowb};
  const uint32_t box[5] = {32, (uint32_t)kSpBoxW, 1, (uint32_t)(kSpBoxH + 3), 1};
  if (!make_tmap_act(&g->ta, in_padded, 5, dims, strides, box, 64, err)) return false;
  const uint64_t wdims[2] = {224, 64};
  const uint64_t wstr[1] = {224 * 2};
  const uint32_t wbox[2] = {32, 64};
  if (!make_tmap_act(&g->tb, Wt, 2, wdims, wstr, wbox, 64, err)) return false;
  g->flops = 2ll * B * OH * OW * 64 * 147;
  return true;
}

inline cudaError_t launch_stem_pool(const StemPoolPlan& g, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(stem_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  return launch_pdl(stem_pool_kernel, dim3(g.grid), dim3(kGemmThreads), g.smem, stream, g.ta, g.tb, g.p);
}

template <int BK, int RES>
inline cudaError_t launch_halo_instance(const HaloPlan& g, int num_sms, cudaStream_t stream) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_halo_kernel<BK, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  const long long tiles = 1ll * g.hp.g.tiles_w * g.hp.g.tiles_h * g.hp.g.tiles_b;
  const int grid = static_cast<int>(std::min<long long>(tiles, num_sms));
  return launch_pdl(conv_halo_kernel<BK, RES>, dim3(grid), dim3(kGemmThreads), g.smem, stream, g.ta, g.tb, g.tout, g.tres, g.hp);
}

inline cudaError_t launch_halo(const HaloPlan& g, int num_sms, cudaStream_t stream) {
  if (g.block_k == 32) return launch_halo_instance<32, 0>(g, num_sms, stream);
  if (g.hp.g.residual != nullptr) return launch_halo_instance<64, 1>(g, num_sms, stream);
  return launch_halo_instance<64, 0>(g, num_sms, stream);
}

}  // namespace rmb
