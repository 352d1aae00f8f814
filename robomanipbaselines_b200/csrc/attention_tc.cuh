// Multi-head attention (head dim 64) on tcgen05/TMEM for the DETR encoder self-attention and decoder cross-attention
// (nn.MultiheadAttention inside third_party/act/detr/models/transformer.py:166-179,229-250; 8 heads x 64, no masks,
// dropout off at eval). One CTA = one (episode, head, 128-query tile), 8 softmax warps + 1 issuer warp, two CTAs per SM.
//
// Sequence lengths here are short (S = 2 + 300*Ncam = 302..902 keys), so the softmax is done in two sweeps over the
// keys instead of an online-rescaled single sweep:
//   sweep 0: S = Q K^T per 128-key block (tcgen05.mma, fp32 in TMEM) -> exact row maximum
//   sweep 1: S again, P = exp2((S - max) * scale*log2e) -> 16-bit P tile in swizzled smem -> O += P V (tcgen05.mma)
// No accumulator rescaling is ever needed. The kernel is bound by the softmax instruction stream, so:
//   * warp w owns TMEM lanes 32*(w%4).. (one query row per thread) and key columns 64*(w/4)..+63 of each S tile: two
//     threads per row, row max / row sum combined through shared memory once per sweep;
//   * the per-element work is FFMA + EX2 + FADD + half a pack; key masking only runs on the last (ragged) key block;
//   * a ninth warp issues every TMA load and MMA, driven by two mbarriers the softmax warps arrive on ("S is in
//     registers", "P is in shared memory"): S(g+1) goes to the tensor pipe the moment S(g) has left TMEM, under the
//     softmax arithmetic of block g, and no softmax warp ever waits for another one's issue work (with thread 0 as the
//     issuer, 31 % of the stall samples sat on the two __syncthreads per block: ncu, profiles/r01n_*);
//   * K / V tiles arrive by TMA two blocks ahead.
//   Q: K-major A operand [128 q x 64 d];  K: K-major B operand [128 kv x 64 d];
//   P: K-major A operand [128 q x 128 kv] (two 64-wide swizzle atoms);  V: MN-major B operand [64 d x 128 kv].
#pragma once
#include <cuda.h>
#include "ptx.cuh"

namespace rmb {

struct AttnParams {
  int Lq, Lk;          // valid queries / keys per episode
  int n_q_tiles;       // ceil(Lq / 128)
  int n_heads;
  int q_col0, k_col0, v_col0;   // first column of head 0 inside the Q / K / V source matrices
  int q_shared;        // 1: Q is the same [Lq, *] matrix for every episode (decoder: input-independent queries)
  float scale_log2e;   // (1/sqrt(64)) * log2(e)
  act_t* out;  // [B*Lq, ld_out], head h at columns h*64
  int ld_out;
};

constexpr int kAttnThreads = 288;   // warps 0-7: softmax (one query row x 64 keys per thread), warp 8: TMA + MMA issuer
// 112 KB of tiles + barriers: two CTAs must fit one SM's 228 KB (each also pays 1 KB of system-reserved shared memory),
// so there is no alignment slack (the dynamic window is declared 1024-byte aligned and checked) and the row statistics
// exchanged between the two column halves alias tile buffers that are idle at that moment.
constexpr int kAttnSmemBytes = 16384 /*Q*/ + 2 * 16384 /*K*/ + 2 * 16384 /*V*/ + 32768 /*P*/ + 256 /*barriers*/;

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}

__global__ void __launch_bounds__(kAttnThreads, 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnParams p) {
  pdl_trigger();
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) {   // swizzled tiles need 1024-byte alignment; never observed, but do not guess
    if (threadIdx.x == 0 && atomicCAS(&g_rmb_fault, 0, 2) == 0) printf("[rmb] attention: shared memory window misaligned\n");
    return;
  }
  uint8_t* sQ = smem;
  uint8_t* sK = smem + 16384;           // 2 stages
  uint8_t* sV = smem + 16384 * 3;       // 2 stages
  uint8_t* sP = smem + 16384 * 5;       // 32 KB
  float* red_max = reinterpret_cast<float*>(sV + 16384);   // [2][128] on V stage 1: idle until block 1 of sweep 1 loads
  float* red_sum = reinterpret_cast<float*>(sK);           // [2][128] on K stage 0: idle once the last S is complete
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 16384 * 7);
  uint64_t* q_bar = bars;               // Q landed
  uint64_t* k_bar = bars + 1;           // [2] K block landed (stage = g & 1)
  uint64_t* v_bar = bars + 3;           // [2] V block landed (stage = kt & 1)
  uint64_t* s_bar = bars + 5;           // S = QK^T complete
  uint64_t* o_bar = bars + 6;           // O += PV complete
  uint64_t* sc_bar = bars + 7;          // S read out of TMEM by the eight softmax warps
  uint64_t* pr_bar = bars + 8;          // P tile written by the eight softmax warps
  uint64_t* kz_bar = bars + 9;          // [6] K blocks of the max sweep (fast path: every block has its own buffer)
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars + 15);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int quarter = warp & 3, half = warp >> 2;
  const int row = quarter * 32 + lane;                 // query row of the tile owned by this thread
  const int qt = blockIdx.x % p.n_q_tiles;
  const int head = (blockIdx.x / p.n_q_tiles) % p.n_heads;
  const int b = blockIdx.x / (p.n_q_tiles * p.n_heads);
  const int nkt = (p.Lk + 127) / 128;
  const int total = 2 * nkt;
  // The max sweep does almost no arithmetic per block, so with two K stages it runs at the latency of a TMA load per
  // block. V and P buffers are idle until the second sweep: with <= 6 key blocks every block of the max sweep gets its
  // own 16 KB buffer (K0 K1 V0 V1 P0 P1 are contiguous) and all of them are requested up front; blocks 0 and 1 are
  // still in K0 / K1 when the second sweep starts.
  const bool fast = nkt <= 6;

  if (tid == 256) {   // issuer thread: barriers, then Q and the first two K blocks go out under the TMEM allocation + CTA sync
    mbar_init(q_bar, 1);
    mbar_init(&k_bar[0], 1); mbar_init(&k_bar[1], 1);
    mbar_init(&v_bar[0], 1); mbar_init(&v_bar[1], 1);
    mbar_init(s_bar, 1); mbar_init(o_bar, 1);
    mbar_init(sc_bar, 8); mbar_init(pr_bar, 8);
    for (int j = 0; j < 6; ++j) mbar_init(&kz_bar[j], 1);
    fence_mbar_init();
    pdl_wait();                                                      // Q / K / V are the previous kernels' outputs
    mbar_arrive_expect_tx(q_bar, 16384u);
    tma_load_3d(sQ, &tmap_q, q_bar, p.q_col0 + head * 64, qt * 128, p.q_shared ? 0 : b);
    if (fast) {
      for (int j = 0; j < nkt; ++j) {
        mbar_arrive_expect_tx(&kz_bar[j], 16384u);
        tma_load_3d(sK + j * 16384, &tmap_k, &kz_bar[j], p.k_col0 + head * 64, j * 128, b);
      }
    } else {
      mbar_arrive_expect_tx(&k_bar[0], 16384u);
      tma_load_3d(sK, &tmap_k, &k_bar[0], p.k_col0 + head * 64, 0, b);
      mbar_arrive_expect_tx(&k_bar[1], 16384u);                      // total >= 2 always
      tma_load_3d(sK + 16384, &tmap_k, &k_bar[1], p.k_col0 + head * 64, (1 % nkt) * 128, b);
    }
  }
  if (warp == 0) { __syncwarp(); tmem_alloc(tmem_ptr_smem, 256); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                                                        // (the issuer thread has already waited)
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_S = tmem_base;          // 128 fp32 columns
  const uint32_t tmem_O = tmem_base + 128;    // 64 fp32 columns
  const uint32_t lane_off = static_cast<uint32_t>(quarter * 32) << 16;

  constexpr uint32_t kIdescS = make_idesc_f16(128, 128, false, false);   // S[128 q,128 kv] = Q * K^T
  constexpr uint32_t kIdescO = make_idesc_f16(128, 64, false, true);     // O[128 q, 64 d] = P * V (V is MN-major)

  // ---- single-thread helpers (warp 8, lane 0): TMA loads and MMA issue
  auto load_k = [&](int g) {                  // K block of step g (both sweeps walk the same blocks) -> stage g & 1
    const int st = g & 1;
    mbar_arrive_expect_tx(&k_bar[st], 16384u);
    tma_load_3d(sK + st * 16384, &tmap_k, &k_bar[st], p.k_col0 + head * 64, (g % nkt) * 128, b);
  };
  auto load_v = [&](int kt) {                 // V block kt -> stage kt & 1
    const int st = kt & 1;
    mbar_arrive_expect_tx(&v_bar[st], 16384u);
    tma_load_3d(sV + st * 16384, &tmap_v, &v_bar[st], p.v_col0 + head * 64, kt * 128, b);
  };
  auto issue_s = [&](int g) {
    int st = g & 1;
    if (!fast) {
      mbar_wait(&k_bar[st], (g >> 1) & 1);
    } else if (g < nkt) {
      st = g;                                  // max sweep: block g lives in buffer g
      mbar_wait(&kz_bar[g], 0);
    } else {
      const int kt = g - nkt;                  // second sweep: blocks 0, 1 are resident, later ones come through the ring
      st = kt & 1;
      if (kt >= 2) mbar_wait(&k_bar[st], ((kt >> 1) - 1) & 1);
    }
    tc_fence_after();
    const uint64_t dq = make_smem_desc(smem_u32(sQ), 0, 1024, 2);
    const uint64_t dk = make_smem_desc(smem_u32(sK + st * 16384), 0, 1024, 2);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_S, dq + 2 * k, dk + 2 * k, kIdescS, k != 0);
    umma_commit(s_bar);
  };
  auto issue_pv = [&](int kt) {
    const int st = kt & 1;
    mbar_wait(&v_bar[st], (kt >> 1) & 1);
    tc_fence_after();
    const uint64_t dp0 = make_smem_desc(smem_u32(sP), 0, 1024, 2);
    const uint64_t dv0 = make_smem_desc(smem_u32(sV + st * 16384), 0, 1024, 2);
#pragma unroll
    for (int k = 0; k < 8; ++k)   // P: atom k>>2, 32 B per 16 keys inside it; V (MN-major): 16 keys = 2 groups of 8 x 128 B rows
      umma_f16_ss(tmem_O, dp0 + ((k >> 2) * 1024 + (k & 3) * 2), dv0 + k * 128, kIdescO, (kt | k) != 0);
    umma_commit(o_bar);
  };

  if (warp == 8) {
    // ------------------------------------------------------------ issuer: TMA loads and MMAs, off the softmax warps' path
    if (lane == 0) {
      mbar_wait(q_bar, 0);
      issue_s(0);
      for (int g = 0; g < total; ++g) {
        mbar_wait(sc_bar, g & 1);              // S(g) is in registers in all eight warps
        tc_fence_after();
        if (g + 1 < total) issue_s(g + 1);     // runs under this block's softmax arithmetic
        if (!fast) {
          if (g + 2 < total) load_k(g + 2);    // stage g & 1: S(g) is complete
        } else if (g >= nkt && g + 2 < total) {
          const int kt2 = g + 2 - nkt, st = kt2 & 1;   // second sweep, block kt2 -> the stage S(g) has just finished with
          mbar_arrive_expect_tx(&k_bar[st], 16384u);
          tma_load_3d(sK + st * 16384, &tmap_k, &k_bar[st], p.k_col0 + head * 64, kt2 * 128, b);
        }
        if (g == nkt - 1) load_v(0);
        if (g >= nkt) {
          const int kt = g - nkt;
          mbar_wait(pr_bar, kt & 1);           // P(kt) is in shared memory; its writers had waited for PV(kt - 1)
          if (kt + 1 < nkt) load_v(kt + 1);    // stage (kt + 1) & 1, last read by PV(kt - 1)
          issue_pv(kt);
        }
      }
    }
  } else {
    // ------------------------------------------------------------ softmax warps 0..7
    float row_max = -INFINITY, row_sum = 0.f, mscaled = 0.f;
    const uint32_t p_row = smem_u32(sP) + half * 16384 + row * 128;   // this thread's 64 keys = swizzle atom `half`
    for (int g = 0; g < total; ++g) {
      const int sweep = (g >= nkt) ? 1 : 0;
      const int kt = sweep ? g - nkt : g;
      const bool ragged = (kt == nkt - 1) && ((p.Lk & 127) != 0);
      const int kcol0 = kt * 128 + half * 64;    // first key of this thread's 64 columns
      __syncwarp();
      mbar_wait(s_bar, g & 1);
      tc_fence_after();
      uint32_t v[64];
      tmem_ld_32x32b_x32(tmem_S + lane_off + half * 64, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      tmem_ld_32x32b_x32(tmem_S + lane_off + half * 64 + 32, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(sc_bar);        // 8 arrivals: the S buffer may be overwritten
      if (ragged) {                              // keys past Lk (TMA zero fill) drop out of both the max and the sum
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (kcol0 + i >= p.Lk) v[i] = 0xff800000u;   // -inf: exp2 gives exactly 0
      }
      if (sweep == 0) {
#pragma unroll
        for (int i = 0; i < 64; i += 2) row_max = max3(row_max, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
        if (kt == nkt - 1) {                     // exact row maximum over both column halves
          red_max[half * 128 + row] = row_max;
          named_bar_sync(1, 256);
          row_max = fmaxf(row_max, red_max[(half ^ 1) * 128 + row]);
          mscaled = row_max * p.scale_log2e;
        }
      } else {
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 64; i += 2) {
          const float p0 = fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -mscaled));
          const float p1 = fast_exp2(fmaf(__uint_as_float(v[i + 1]), p.scale_log2e, -mscaled));
          row_sum += p0 + p1;
          pk[i >> 1] = pack_act2(p0, p1);
        }
        if (kt > 0) {                            // PV of the previous block has finished reading the P tile
          mbar_wait(o_bar, (kt - 1) & 1);
          tc_fence_after();
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          st_shared_v4(p_row + ((j ^ (row & 7)) << 4), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        fence_proxy_async_smem();                // generic-proxy smem writes -> visible to the tensor core (async proxy)
        __syncwarp();
        if (lane == 0) mbar_arrive(pr_bar);      // 8 arrivals: PV(kt) may be issued
      }
    }
    __syncwarp();
    red_sum[half * 128 + row] = row_sum;
    mbar_wait(o_bar, (nkt - 1) & 1);
    tc_fence_after();
    named_bar_sync(1, 256);
    const float inv = 1.0f / (row_sum + red_sum[(half ^ 1) * 128 + row]);

  // O tile: this thread writes 32 of the head's 64 output columns of its query row
  const int q = qt * 128 + row;
  {
    uint32_t o[32];
    tmem_ld_32x32b_x32(tmem_O + lane_off + half * 32, o);
    tmem_ld_wait();
    if (q < p.Lq) {
      act_t* orow = p.out + (static_cast<long long>(b) * p.Lq + q) * p.ld_out + head * 64 + half * 32;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        reinterpret_cast<uint4*>(orow)[i] =
            make_uint4(pack_act2(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv),
                       pack_act2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv),
                       pack_act2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv),
                       pack_act2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv));
    }
  }
  }   // softmax warps
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 256); }
}

}  // namespace rmb
