// Fused epilogue of the transformer's two residual GEMMs (attention out-projection, FFN down-projection):
//   v = A Wt^T + bias + residual      (fp32, residual stream)          transformer.py:172-173,176-177 (src + src2)
//   y = LayerNorm(v) * gamma + beta   (eps inside the sqrt, as torch)  transformer.py:173,178 (norm1 / norm2)
// written as fp32 (next residual), as the 16-bit A operand of the next GEMM, and as 16-bit (y + pos) for the next
// layer's Q/K projection -- everything layernorm512_kernel produced, without v ever being a separate HBM round trip
// (unfused: out-proj + residual writes v, LayerNorm reads it back: 632 MB per launch at 256 episodes, HBM-bound).
//
// The tile is 128 rows x 512 columns = all of TMEM (gemm_tc_body's EPI == 5 mode), so rows are complete in the tile.
// Warp e of the eight epilogue warps owns TMEM lanes 32 (e % 4).. and the 32-column chunks with (chunk % 2) == e / 4,
// exactly as gemm_epilogue: pass 1 thread = row (tcgen05.ld -> swizzled staging), pass 2 eight lanes = one row, so
// every residual load and output store is a full 128-byte row segment.
//   phase A  per chunk: v = acc + bias + residual; row sums of v and v^2; v is PARKED in the fp32 output (it is
//            rewritten below by the same lane, so it only ever lives in L2). The accumulator is released at the end of
//            phase A: the next tile's MMAs run under phase B.
//   exchange the two warps that share a row quarter swap their half-row sums through shared memory (64-thread barrier)
//   phase B  per chunk: re-read v (L2 hit, same lane that wrote it), normalise, store fp32 + 16-bit (+ pos) outputs.
#pragma once
#include "gemm_tc.cuh"

namespace rmb {

template <bool PAIR>
__device__ __forceinline__ void gemm_ln_epilogue(const GemmParams& p, uint8_t* smem_stage, uint32_t tmem_base,
                                                 uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                                 const TileSched<PAIR>& ts, int warp, int lane) {
  constexpr int BLOCK_N = 512;
  constexpr int kChunks = BLOCK_N / 64;            // 32-column chunks per warp
  const int q = warp & 3, half = warp >> 2;
  const int sub = lane & 7, rsel = lane >> 3;
  const uint32_t stg = smem_u32(smem_stage + warp * 4096);
  uint32_t st_addr[8];
#pragma unroll
  for (int g = 0; g < 8; ++g) st_addr[g] = stg + lane * 128 + ((g ^ (lane & 7)) << 4);
  const uint32_t ld_even = stg + rsel * 128 + ((sub ^ rsel) << 4);
  const uint32_t ld_odd = stg + (rsel + 4) * 128 + ((sub ^ (rsel + 4)) << 4);
  // half-row sums of this warp's 32 rows live at the start of its own staging tile between phase A and phase B
  float2* my_stats = reinterpret_cast<float2*>(smem_stage + warp * 4096);
  const float2* peer_stats = reinterpret_cast<const float2*>(smem_stage + (warp ^ 4) * 4096);
  const int col0 = half * 32 + sub * 4;            // first of this lane's 4 columns inside chunk 0 of this warp
  const long long M = p.W;
  uint32_t acc_phase = 0;
  // The residual rows of a tile (256 KB per CTA, straight from HBM: the previous LayerNorm wrote them a whole launch ago)
  // are requested into L2 long before phase A reads them: with eight float4 loads per lane in flight a warp covers only
  // 4 KB per HBM round trip, which alone would bound the epilogue at ~20 GB/s per SM.
  auto prefetch_residual = [&](int t) {
    const long long grow = static_cast<long long>(t) * 128 + q * 32 + lane;      // one 2 KB row per thread of warps 0..3
    if (half == 0 && grow < M) prefetch_l2_bulk(static_cast<const float*>(p.residual) + grow * BLOCK_N, BLOCK_N * 4);
  };
  for (int ti = ts.first; ti < ts.count; ti += ts.step) {
    const int tile = ts.tile(ti);                  // n_col_tiles == 1: tile == row tile
    const long long row_q = static_cast<long long>(tile) * 128 + q * 32;
    // rows of this lane in pass 2: row_q + rsel + 4 it; residual / out are dense [M, 512] (plan_linear_ln), so row `it`
    // is a compile-time offset from one base pointer; rows at or past M are never dereferenced
    const int rows_left = static_cast<int>(min(M - row_q - rsel, static_cast<long long>(128)));   // row `it` is valid iff 4 it < rows_left
    const float* rbase = static_cast<const float*>(p.residual) + (row_q + rsel) * BLOCK_N + col0;
    float* obase = static_cast<float*>(p.out) + (row_q + rsel) * BLOCK_N + col0;
#define RMB_LN_OK(it) (4 * (it) < rows_left)
#define RMB_LN_R(it) (rbase + (it) * 4 * BLOCK_N)
#define RMB_LN_O(it) (obase + (it) * 4 * BLOCK_N)
    if (ti == ts.first) prefetch_residual(tile);   // later tiles are prefetched one tile ahead, under phase B (below)
    mbar_wait(&tmem_full_bar[0], acc_phase);
    tc_fence_after();
    const uint32_t t_row = tmem_base + half * 32 + (static_cast<uint32_t>(q * 32) << 16);
    float sum[8], sq[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) { sum[it] = 0.f; sq[it] = 0.f; }
    uint32_t v[32];
    tmem_ld_32x32b_x32(t_row, v);
    // ---------------------------------------------------------------- phase A
#pragma unroll
    for (int j = 0; j < kChunks; ++j) {
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 8; ++g) st_shared_v4(st_addr[g], v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
      if (j + 1 < kChunks) tmem_ld_32x32b_x32(t_row + (j + 1) * 64, v);
      else tc_fence_before();
      __syncwarp();
      if (j + 1 == kChunks && lane == 0) ts.release_acc(tmem_empty_bar, 0);
      const float4 bv = (p.bias != nullptr) ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j * 64)) : make_float4(0.f, 0.f, 0.f, 0.f);
      float4 f[8], rv[8];
#pragma unroll
      for (int it = 0; it < 8; ++it) f[it] = ld_shared_v4f(((it & 1) ? ld_odd : ld_even) + (it >> 1) * 1024);
#pragma unroll
      for (int it = 0; it < 8; ++it)
        rv[it] = RMB_LN_OK(it) ? __ldg(reinterpret_cast<const float4*>(RMB_LN_R(it) + j * 64)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const float x = f[it].x + bv.x + rv[it].x, y = f[it].y + bv.y + rv[it].y;
        const float z = f[it].z + bv.z + rv[it].z, w = f[it].w + bv.w + rv[it].w;
        sum[it] += (x + y) + (z + w);
        sq[it] += (x * x + y * y) + (z * z + w * w);
        if (RMB_LN_OK(it)) *reinterpret_cast<float4*>(RMB_LN_O(it) + j * 64) = make_float4(x, y, z, w);
      }
      __syncwarp();                                // staging tile is rewritten by the next chunk's pass 1
    }
    // ---------------------------------------------------------------- half-row sums -> whole-row statistics
#pragma unroll
    for (int it = 0; it < 8; ++it) {
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        sum[it] += __shfl_xor_sync(0xffffffffu, sum[it], o);
        sq[it] += __shfl_xor_sync(0xffffffffu, sq[it], o);
      }
      if (sub == 0) my_stats[it * 4 + rsel] = make_float2(sum[it], sq[it]);
    }
    named_bar_sync(1 + q, 64);                     // the two warps of this row quarter
    float mean[8], rstd[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const float2 o = peer_stats[it * 4 + rsel];
      const float s = sum[it] + o.x, ss = sq[it] + o.y;
      mean[it] = s * (1.0f / BLOCK_N);
      const float var = fmaxf(ss * (1.0f / BLOCK_N) - mean[it] * mean[it], 0.f);
      rstd[it] = rsqrtf(var + p.ln_eps);
    }
    // ---------------------------------------------------------------- phase B
    if (ti + ts.step < ts.count) prefetch_residual(ts.tile(ti + ts.step));
    act_t* yact = static_cast<act_t*>(p.ln_y_act);
    act_t* ypos = static_cast<act_t*>(p.ln_ypos_act);
    // The parked rows come back from L2: all eight loads of a chunk are issued before anything is stored (a store in
    // between would order the loads behind it), and the next chunk's loads are in flight under this chunk's arithmetic.
    float4 x[8], xn[8];
#pragma unroll
    for (int it = 0; it < 8; ++it)
      x[it] = RMB_LN_OK(it) ? *reinterpret_cast<const float4*>(RMB_LN_O(it)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < kChunks; ++j) {
      if (j + 1 < kChunks) {
#pragma unroll
        for (int it = 0; it < 8; ++it)
          xn[it] = RMB_LN_OK(it) ? *reinterpret_cast<const float4*>(RMB_LN_O(it) + (j + 1) * 64) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      const float4 g4 = __ldg(reinterpret_cast<const float4*>(p.ln_gamma + col0 + j * 64));
      const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.ln_beta + col0 + j * 64));
      float4 pv[8];
      if (ypos != nullptr) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const long long grow = row_q + it * 4 + rsel;
          pv[it] = RMB_LN_OK(it) ? __ldg(reinterpret_cast<const float4*>(p.ln_pos + (grow % p.ln_L) * BLOCK_N + col0 + j * 64))
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        if (!RMB_LN_OK(it)) continue;
        float4 y;
        y.x = (x[it].x - mean[it]) * rstd[it] * g4.x + b4.x; y.y = (x[it].y - mean[it]) * rstd[it] * g4.y + b4.y;
        y.z = (x[it].z - mean[it]) * rstd[it] * g4.z + b4.z; y.w = (x[it].w - mean[it]) * rstd[it] * g4.w + b4.w;
        *reinterpret_cast<float4*>(RMB_LN_O(it) + j * 64) = y;
        const long long eoff = (row_q + it * 4 + rsel) * BLOCK_N + col0 + j * 64;
        if (yact != nullptr)
          *reinterpret_cast<uint2*>(yact + eoff) = make_uint2(pack_act2(y.x, y.y), pack_act2(y.z, y.w));
        if (ypos != nullptr)
          *reinterpret_cast<uint2*>(ypos + eoff) = make_uint2(pack_act2(y.x + pv[it].x, y.y + pv[it].y), pack_act2(y.z + pv[it].z, y.w + pv[it].w));
      }
#pragma unroll
      for (int it = 0; it < 8; ++it) x[it] = xn[it];
    }
    named_bar_sync(1 + q, 64);                     // peer has read my_stats before the next tile's pass 1 overwrites it
    acc_phase ^= 1;
  }
#undef RMB_LN_OK
#undef RMB_LN_R
#undef RMB_LN_O
}

}  // namespace rmb
