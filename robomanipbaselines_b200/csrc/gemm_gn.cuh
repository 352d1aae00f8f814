// Fused epilogue of the ConditionalUnet1D convs: conv bias -> GroupNorm(8) -> Mish -> FiLM (scale * y + bias) ->
// + residual -> 16-bit store, computed on the TMEM accumulator of gemm_tc_kernel<..., EPI = 1>, so the Conv1dBlock
// output never makes a round trip through HBM before it is normalised
// (upstream diffusion_policy Conv1dBlock / ConditionalResidualBlock1D; oracle/diffusion_oracle.py:57-94).
//
// Why it fits one tile: rows of a tile are box_b whole samples x T time steps (128 % T == 0) and a group is
// C/8 channels wide, a multiple of 32 that divides BLOCK_N, so every (sample, group) the tile touches is complete
// inside it. Two sweeps over the accumulator (tcgen05.ld is cheap): sweep 1 leaves per-(row, 32-column chunk) sums
// in shared memory, a 256-thread named barrier, each thread reduces the T rows x Cg/32 chunks of its own sample's
// groups, sweep 2 normalises and stores. The MMA warp is already working on the next tile (double-buffered TMEM).
#pragma once
#include "diffusion_kernels.cuh"
#include "gemm_tc.cuh"

namespace rmb {

// FILM = true: conv0 of a block (FiLM after the Mish, no residual); FILM = false: conv1 (residual or nothing).
template <int BLOCK_N, bool FILM, bool PAIR>
__device__ __forceinline__ void gemm_gn_epilogue(const GemmParams& p, uint8_t* smem_stage, uint32_t tmem_base,
                                                 uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                                 const TileSched<PAIR>& ts, int warp, int lane) {
  constexpr int kChunks = BLOCK_N / 32;   // 32-column chunks of a tile
  constexpr int kMine = kChunks / 2;      // chunks handled by this warp: columns half * BLOCK_N/2 .. +BLOCK_N/2
  float2* stat = reinterpret_cast<float2*>(smem_stage);          // [128 rows][kChunks] (sum, sum of squares): <= 8 KB
  // per-warp 2 KB FiLM staging: [samples of this warp][16 float4] = 8 x scale, 8 x bias of the current 32-column chunk
  float4* film_s = reinterpret_cast<float4*>(smem_stage + 8192 + warp * 2048);
  const int q = warp & 3, half = warp >> 2;
  const int r = q * 32 + lane;
  const int T = p.box_w, rows_per_tile = p.box_w * p.box_b;
  const int wl = r % T, bl = r / T;
  const int S = (T < 32) ? 32 / T : 1;                           // samples covered by one warp
  const int bl0 = (q * 32) / T, ls = bl - bl0;                   // first sample of the warp, this lane's sample slot
  const int cpg = p.gn_cg >> 5;                                  // chunks per group
  const float inv_n = 1.0f / static_cast<float>(T * p.gn_cg);
  const int N = p.n_col_tiles * BLOCK_N;
  int acc = 0; uint32_t acc_phase = 0;
  for (int ti = ts.first; ti < ts.count; ti += ts.step) {
    const int tile = ts.tile(ti);
    const int ct = tile % p.n_col_tiles, bt = tile / p.n_col_tiles;
    const int b = bt * p.box_b + bl;
    const bool valid = (r < rows_per_tile) && (b < p.B);
    const long long grow = static_cast<long long>(b) * T + wl;
    const int n0 = ct * BLOCK_N;
    // FiLM rows of this warp's samples for one chunk, (gc + t) summed: lane k-th load = float4 #(lane + 32k) of S*16
    float4 F[4];
    auto load_film = [&](int c) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int idx = lane + 32 * k;
        F[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (idx < S * 16) {
          const int ss = idx >> 4, part = idx & 15;
          const int sb = bt * p.box_b + bl0 + ss;
          const int off = p.gn_film_off + n0 + c * 32 + (part & 7) * 4 + ((part & 8) ? N : 0);
          const float4 t4 = __ldg(reinterpret_cast<const float4*>(p.gn_film_t + off));
          if (sb < p.B && bl0 + ss < p.box_b) {
            const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.gn_film_gc + static_cast<long long>(sb) * p.gn_film_ld + off));
            F[k] = make_float4(g4.x + t4.x, g4.y + t4.y, g4.z + t4.z, g4.w + t4.w);
          }
        }
      }
    };
    const act_t* rp = (!FILM && p.gn_res != nullptr && valid) ? static_cast<const act_t*>(p.gn_res) + grow * p.gn_ld_res + n0 : nullptr;
    uint4 R[4];
    auto load_res = [&](int c) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        R[k] = (rp != nullptr) ? __ldg(reinterpret_cast<const uint4*>(rp + c * 32) + k) : make_uint4(0u, 0u, 0u, 0u);
    };
    if (FILM) load_film(half * kMine); else load_res(half * kMine);   // in flight while the accumulator is still being built

    mbar_wait(&tmem_full_bar[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_row = tmem_base + acc * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t v[32];
    // ---- sweep 1: per-row partial sums of (acc + bias) per 32-column chunk
#pragma unroll
    for (int j = 0; j < kMine; ++j) {
      const int c = half * kMine + j;
      tmem_ld_32x32b_x32(t_row + c * 32, v);
      tmem_ld_wait();
      const float4* bp = reinterpret_cast<const float4*>(p.bias + n0 + c * 32);
      float s = 0.f, qq = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 b4 = __ldg(bp + i);
        const float x0 = __uint_as_float(v[4 * i]) + b4.x, x1 = __uint_as_float(v[4 * i + 1]) + b4.y;
        const float x2 = __uint_as_float(v[4 * i + 2]) + b4.z, x3 = __uint_as_float(v[4 * i + 3]) + b4.w;
        s += (x0 + x1) + (x2 + x3);
        qq += (x0 * x0 + x1 * x1) + (x2 * x2 + x3 * x3);
      }
      stat[r * kChunks + c] = make_float2(s, qq);
    }
    named_bar_sync(1, 256);
    // ---- statistics of the groups this warp's columns belong to (<= kMine distinct groups)
    float g_mean[kMine], g_rstd[kMine];
#pragma unroll
    for (int j = 0; j < kMine; ++j) {
      const int c = half * kMine + j, gi = c / cpg;
      if (j > 0 && gi == (c - 1) / cpg) { g_mean[j] = g_mean[j - 1]; g_rstd[j] = g_rstd[j - 1]; continue; }
      float s = 0.f, qq = 0.f;
      for (int rr = bl * T; rr < bl * T + T; ++rr)
        for (int cc = gi * cpg; cc < gi * cpg + cpg; ++cc) {
          const float2 a = stat[rr * kChunks + cc];
          s += a.x; qq += a.y;
        }
      const float mean = s * inv_n;
      g_mean[j] = mean;
      g_rstd[j] = rsqrtf(fmaxf(qq * inv_n - mean * mean, 0.0f) + 1e-5f);
    }
    named_bar_sync(1, 256);                                       // stat[] may be overwritten by the next tile
    // ---- sweep 2: normalise, Mish, FiLM | residual, store
    act_t* op = static_cast<act_t*>(p.out) + grow * p.ld_out + n0;
#pragma unroll
    for (int j = 0; j < kMine; ++j) {
      const int c = half * kMine + j;
      tmem_ld_32x32b_x32(t_row + c * 32, v);
      uint4 Rc[4];
      if (FILM) {
        __syncwarp();                                             // the previous chunk's FiLM reads are done
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (lane + 32 * k < S * 16) film_s[lane + 32 * k] = F[k];
        __syncwarp();
        if (j + 1 < kMine) load_film(c + 1);                      // next chunk's FiLM rows fly during this chunk's math
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) Rc[k] = R[k];
        if (j + 1 < kMine) load_res(c + 1);
      }
      tmem_ld_wait();
      if (j + 1 == kMine) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) ts.release_acc(tmem_empty_bar, acc);       // 8 arrivals (per CTA) free the accumulator
      }
      const float mean = g_mean[j], rstd = g_rstd[j];
#pragma unroll
      for (int i = 0; i < 4; ++i) {                               // 8 columns per iteration
        const int col = c * 32 + i * 8;
        float x[8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + col) + h);
          const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.gn_gamma + n0 + col) + h);
          const float4 e4 = __ldg(reinterpret_cast<const float4*>(p.gn_beta + n0 + col) + h);
          // ((v + b) - mean) * rstd * g + e  =  v * G + ((b - mean) * G + e),  G = rstd * g
          const float G0 = rstd * g4.x, G1 = rstd * g4.y, G2 = rstd * g4.z, G3 = rstd * g4.w;
          x[4 * h + 0] = mish_f(fmaf(__uint_as_float(v[8 * i + 4 * h + 0]), G0, fmaf(b4.x - mean, G0, e4.x)));
          x[4 * h + 1] = mish_f(fmaf(__uint_as_float(v[8 * i + 4 * h + 1]), G1, fmaf(b4.y - mean, G1, e4.y)));
          x[4 * h + 2] = mish_f(fmaf(__uint_as_float(v[8 * i + 4 * h + 2]), G2, fmaf(b4.z - mean, G2, e4.z)));
          x[4 * h + 3] = mish_f(fmaf(__uint_as_float(v[8 * i + 4 * h + 3]), G3, fmaf(b4.w - mean, G3, e4.w)));
        }
        if (FILM) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const float4 sc = film_s[ls * 16 + i * 2 + h], bi = film_s[ls * 16 + 8 + i * 2 + h];
            x[4 * h + 0] = sc.x * x[4 * h + 0] + bi.x;
            x[4 * h + 1] = sc.y * x[4 * h + 1] + bi.y;
            x[4 * h + 2] = sc.z * x[4 * h + 2] + bi.z;
            x[4 * h + 3] = sc.w * x[4 * h + 3] + bi.w;
          }
        } else {
          const uint4 rv = Rc[i];
          const float2 r0 = unpack_act2(rv.x), r1 = unpack_act2(rv.y), r2 = unpack_act2(rv.z), r3 = unpack_act2(rv.w);
          x[0] += r0.x; x[1] += r0.y; x[2] += r1.x; x[3] += r1.y; x[4] += r2.x; x[5] += r2.y; x[6] += r3.x; x[7] += r3.y;
        }
        if (valid)
          *reinterpret_cast<uint4*>(op + col) = make_uint4(pack_act2(x[0], x[1]), pack_act2(x[2], x[3]), pack_act2(x[4], x[5]), pack_act2(x[6], x[7]));
      }
    }
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
  }
}

}  // namespace rmb
