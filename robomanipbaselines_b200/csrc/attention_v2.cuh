// Multi-head attention (head dim 64) on tcgen05/TMEM, second version: persistent CTAs, two independent query-tile
// STREAMS per CTA, one-sweep online softmax with one thread per query row.
// (nn.MultiheadAttention inside third_party/act/detr/models/transformer.py:166-179,229-250; 8 heads x 64, no masks.)
//
// What bounds this op on B200 is the exponential unit: B*H*L^2 = 742 M exp2 per encoder launch at 256 episodes against
// 16 MUFU results per clock per SM = 0.19 ms, more than the two MMAs need (0.15 ms at the sustained tensor rate). The
// first version (attention_tc.cuh: one 128-query tile per CTA, two CTAs per SM, two sweeps over the keys) kept the MUFU
// 39 % busy (ncu): every tile paid the TMEM allocation / barrier set-up / first-load latency, the max sweep was a chain
// of MMA latencies with idle softmax warps, and row statistics crossed warps through shared memory.
// Here:
//   * one CTA per SM walks work items (episode, head, 128-query tile) for the whole launch; the next item's Q and first
//     K blocks are requested as soon as the last S of the current item has left TMEM;
//   * the CTA runs TWO items at a time ("streams"): warps 0-3 are the softmax warps of stream 0, warps 4-7 of stream 1,
//     warps 8 / 9 their issuers (TMA loads + MMAs, one elected thread each). Each stream has its own Q, K/V ring,
//     P tile, S and O accumulators (TMEM columns 192 s .. 192 s + 191), so while one stream's warps are between blocks
//     (waiting for S, reading TMEM, writing P) the other stream's warps keep the exponential unit busy;
//   * thread = query row (TMEM lane), all 128 keys of a block: row maximum and row sum never leave the thread;
//   * ONE sweep: S is computed once per key block. The running maximum m only has to be *some* reference point of the
//     row, so it is moved (and O, row_sum rescaled by exp2(m_old - m_new)) only when a block's maximum exceeds it by more
//     than 2^8: P then stays <= 256, far inside the 16-bit range, and the final O / row_sum is the exact softmax. With
//     LayerNorm-bounded logits that happens in the first block only, where nothing has been accumulated yet;
//   * 32-key chunks past Lk are neither read nor exponentiated, and a warp whose 32 query rows are all past Lq only
//     takes part in the barriers: the 602 -> 640 padding costs MMA time but no MUFU time.
//   Q: K-major A operand [128 q x 64 d];  K: K-major B operand [128 kv x 64 d];
//   P: K-major A operand [128 q x 128 kv] (two 64-wide swizzle atoms);  V: MN-major B operand [64 d x 128 kv].
#pragma once
#include <cuda.h>
#include "attention_tc.cuh"

namespace rmb {

constexpr int kAttn2Threads = 320;
constexpr int kAttn2StreamBytes = 16384 /*Q*/ + 2 * 16384 /*K*/ + 2 * 16384 /*V*/ + 32768 /*P*/;   // 112 KB
constexpr int kAttn2SmemBytes = 2 * kAttn2StreamBytes + 512 /*barriers*/ + 1024 /*alignment slack*/;
constexpr float kAttn2RescaleLog2 = 8.0f;

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,"
      "%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__global__ void __launch_bounds__(kAttn2Threads, 1)
attention_v2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnParams p, int n_items) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars_all = reinterpret_cast<uint64_t*>(smem + 2 * kAttn2StreamBytes);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars_all + 24);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stream = (warp < 8) ? (warp >> 2) : (warp - 8);
  uint8_t* sbase = smem + stream * kAttn2StreamBytes;
  uint8_t* sQ = sbase;
  uint8_t* sK = sbase + 16384;            // 2 stages
  uint8_t* sV = sbase + 16384 * 3;        // 2 stages
  uint8_t* sP = sbase + 16384 * 5;        // 32 KB
  uint64_t* bars = bars_all + stream * 12;
  uint64_t* q_bar = bars;                 // Q landed
  uint64_t* k_bar = bars + 1;             // [2] K block landed
  uint64_t* v_bar = bars + 3;             // [2] V block landed
  uint64_t* s_bar = bars + 5;             // S = Q K^T complete
  uint64_t* o_bar = bars + 6;             // O += P V complete
  uint64_t* sc_bar = bars + 7;            // S read out of TMEM by the stream's four softmax warps
  uint64_t* pr_bar = bars + 8;            // P tile written by the stream's four softmax warps

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      uint64_t* b = bars_all + s * 12;
      mbar_init(b, 1);
      mbar_init(b + 1, 1); mbar_init(b + 2, 1);
      mbar_init(b + 3, 1); mbar_init(b + 4, 1);
      mbar_init(b + 5, 1); mbar_init(b + 6, 1);
      mbar_init(b + 7, 4); mbar_init(b + 8, 4);
    }
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) { prefetch_tmap(&tmap_q); prefetch_tmap(&tmap_k); prefetch_tmap(&tmap_v); }
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                                     // Q / K / V are the previous kernels' outputs
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_S = tmem_base + stream * 192;          // 128 fp32 columns
  const uint32_t tmem_O = tmem_base + stream * 192 + 128;    // 64 fp32 columns

  const int nkt = (p.Lk + 127) / 128;
  const int n_streams = 2 * gridDim.x;
  const int first = 2 * blockIdx.x + stream;

  constexpr uint32_t kIdescS = make_idesc_f16(128, 128, false, false);   // S[128 q,128 kv] = Q * K^T
  constexpr uint32_t kIdescO = make_idesc_f16(128, 64, false, true);     // O[128 q, 64 d] = P * V (V is MN-major)

  if (warp >= 8) {
    // ------------------------------------------------------------ issuer of this stream: TMA loads and MMAs
    if (lane == 0 && first < n_items) {
      // K and V blocks are numbered through the whole item sequence (kseq / vseq): stage = seq & 1, parity = (seq >> 1) & 1
      uint32_t kload = 0, kuse = 0, vload = 0, vuse = 0, q_cnt = 0, sc_cnt = 0, pr_cnt = 0;
      auto decode = [&](int item, int* b, int* head, int* qt) {
        *qt = item % p.n_q_tiles;
        *head = (item / p.n_q_tiles) % p.n_heads;
        *b = item / (p.n_q_tiles * p.n_heads);
      };
      auto load_q = [&](int item) {
        int b, head, qt; decode(item, &b, &head, &qt);
        mbar_arrive_expect_tx(q_bar, 16384u);
        tma_load_3d(sQ, &tmap_q, q_bar, p.q_col0 + head * 64, qt * 128, p.q_shared ? 0 : b);
      };
      // K blocks form one sequence through all items of the stream: block n belongs to item first + (n / nkt) * n_streams
      auto load_next_k = [&]() {
        const int item = first + static_cast<int>(kload / nkt) * n_streams;
        if (item >= n_items) return;
        int b, head, qt; decode(item, &b, &head, &qt);
        const int st = kload & 1;
        mbar_arrive_expect_tx(&k_bar[st], 16384u);
        tma_load_3d(sK + st * 16384, &tmap_k, &k_bar[st], p.k_col0 + head * 64, static_cast<int>(kload % nkt) * 128, b);
        ++kload;
      };
      auto load_v = [&](int item, int kt) {
        int b, head, qt; decode(item, &b, &head, &qt);
        const int st = vload & 1;
        mbar_arrive_expect_tx(&v_bar[st], 16384u);
        tma_load_3d(sV + st * 16384, &tmap_v, &v_bar[st], p.v_col0 + head * 64, kt * 128, b);
        ++vload;
      };
      auto issue_s = [&]() {
        const int st = kuse & 1;
        mbar_wait(&k_bar[st], (kuse >> 1) & 1);
        ++kuse;
        tc_fence_after();
        const uint64_t dq = make_smem_desc(smem_u32(sQ), 0, 1024, 2);
        const uint64_t dk = make_smem_desc(smem_u32(sK + st * 16384), 0, 1024, 2);
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_S, dq + 2 * k, dk + 2 * k, kIdescS, k != 0);
        umma_commit(s_bar);
      };
      auto issue_pv = [&](int kt) {
        const int st = vuse & 1;
        mbar_wait(&v_bar[st], (vuse >> 1) & 1);
        ++vuse;
        tc_fence_after();
        const uint64_t dp0 = make_smem_desc(smem_u32(sP), 0, 1024, 2);
        const uint64_t dv0 = make_smem_desc(smem_u32(sV + st * 16384), 0, 1024, 2);
#pragma unroll
        for (int k = 0; k < 8; ++k)   // P: atom k>>2, 32 B per 16 keys inside it; V (MN-major): 16 keys = 2 groups of 8 x 128 B rows
          umma_f16_ss(tmem_O, dp0 + ((k >> 2) * 1024 + (k & 3) * 2), dv0 + k * 128, kIdescO, (kt | k) != 0);
        umma_commit(o_bar);
      };
      load_q(first);
      load_next_k();
      load_next_k();
      for (int item = first; item < n_items; item += n_streams) {
        mbar_wait(q_bar, q_cnt & 1); ++q_cnt;
        issue_s();                                   // S(0); an S(nkt-1) of the previous item has been consumed
        for (int g = 0; g < nkt; ++g) {
          mbar_wait(sc_bar, sc_cnt & 1); ++sc_cnt;   // S(g) is in registers: the S buffer and K stage of block g are free
          tc_fence_after();
          if (g + 1 < nkt) issue_s();                // runs under this block's softmax arithmetic
          else if (item + n_streams < n_items) load_q(item + n_streams);   // last S of the item: Q buffer is free
          load_next_k();                             // two blocks ahead, into the stage S(g) has finished with
          if (g == 0) {                              // the stream's warps are past the previous item's epilogue: every
            load_v(item, 0);                         // earlier P V has completed, both V stages are free
            if (nkt > 1) load_v(item, 1);
          }
          mbar_wait(pr_bar, pr_cnt & 1); ++pr_cnt;   // P(g) is in shared memory; its writers had waited for P V (g - 1)
          if (g >= 1 && g + 1 < nkt) load_v(item, g + 1);   // stage last read by P V (g - 1)
          issue_pv(g);
        }
      }
    }
  } else {
    // ------------------------------------------------------------ softmax warps of this stream: thread = query row
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(wq * 32) << 16;
    const uint32_t p_row = smem_u32(sP) + row * 128;
    const uint32_t swz = static_cast<uint32_t>(row & 7);
    uint32_t s_cnt = 0, o_cnt = 0;                  // completed S / P V phases this thread has accounted for
    for (int item = first; item < n_items; item += n_streams) {
      const int qt = item % p.n_q_tiles;
      const int head = (item / p.n_q_tiles) % p.n_heads;
      const int b = item / (p.n_q_tiles * p.n_heads);
      const bool warp_live = qt * 128 + wq * 32 < p.Lq;    // warp-uniform
      float m = -INFINITY, msc = 0.f, row_sum = 0.f;       // reference maximum (raw logits), m * scale_log2e, sum of P
      for (int g = 0; g < nkt; ++g) {
        const int nvalid = min(128, p.Lk - g * 128);
        const int nchunks = (nvalid + 31) >> 5;
        mbar_wait(s_bar, s_cnt & 1); ++s_cnt;
        tc_fence_after();
        if (warp_live) {
          // ONE pass over S(g), 32 keys at a time, the next chunk's tcgen05.ld in flight under this chunk's arithmetic.
          // Each chunk is max-scanned first; P is formed against the current reference m. When the last chunk has been
          // seen and the block maximum turns out to exceed m by more than 2^8 (rare: see the header), the block is
          // redone against a new reference in the slow path below, before S(g) is released.
          const float row_sum_prev = row_sum;
          float bm = -INFINITY;
          bool redo = false;
          uint32_t va[32], vb[32];
          tmem_ld_32x32b_x32(tmem_S + lane_off, va);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            uint32_t (&cur)[32] = (c & 1) ? vb : va;
            uint32_t (&nxt)[32] = (c & 1) ? va : vb;
            uint32_t pk[16];
            if (c < nchunks) {
              tmem_ld_wait();
              if (c + 1 < nchunks) tmem_ld_32x32b_x32(tmem_S + lane_off + (c + 1) * 32, nxt);
              if ((c + 1) * 32 > nvalid) {            // keys past Lk: -inf drops out of the maximum and gives P = 0
#pragma unroll
                for (int i = 0; i < 32; ++i)
                  if (c * 32 + i >= nvalid) cur[i] = 0xff800000u;
              }
#pragma unroll
              for (int i = 0; i < 32; i += 2) bm = max3(bm, __uint_as_float(cur[i]), __uint_as_float(cur[i + 1]));
              if (g == 0 && c == 0) { m = bm; msc = m * p.scale_log2e; }   // first reference point: the first 32 keys
              if (c + 1 == nchunks) {
                redo = __any_sync(0xffffffffu, (bm - m) * p.scale_log2e > kAttn2RescaleLog2);
                if (!redo) {                          // last read of S(g): the issuer may overwrite it
                  tc_fence_before();
                  __syncwarp();
                  if (lane == 0) mbar_arrive(sc_bar);
                }
              }
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                const float p0 = fast_exp2(fmaf(__uint_as_float(cur[i]), p.scale_log2e, -msc));
                const float p1 = fast_exp2(fmaf(__uint_as_float(cur[i + 1]), p.scale_log2e, -msc));
                row_sum += p0 + p1;
                pk[i >> 1] = pack_act2(p0, p1);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 16; ++i) pk[i] = 0u;
            }
            if (c == 0 && g > 0) {                    // P V (g - 1) has finished reading the P tile (and writing O)
              mbar_wait(o_bar, o_cnt & 1); ++o_cnt;
              tc_fence_after();
            }
            const uint32_t pa = p_row + (c >> 1) * 16384;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_v4(pa + (((((c & 1) << 2) + j) ^ swz) << 4), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
          }
          if (redo) {
            // ---- slow path (warp-uniform): move the reference point of the rows that need it, rescale what they have
            // accumulated, and form this block's P again
            const bool need = (bm - m) * p.scale_log2e > kAttn2RescaleLog2;
            const float alpha = (need && g > 0) ? fast_exp2((m - bm) * p.scale_log2e) : 1.0f;
            if (need) { m = bm; msc = m * p.scale_log2e; }
            row_sum = row_sum_prev * alpha;
            if (g > 0) {
#pragma unroll
              for (int hcol = 0; hcol < 2; ++hcol) {
                uint32_t o[32];
                tmem_ld_32x32b_x32(tmem_O + lane_off + hcol * 32, o);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                tmem_st_32x32b_x32(tmem_O + lane_off + hcol * 32, o);
              }
              tmem_st_wait();
            }
            for (int c = 0; c < nchunks; ++c) {
              uint32_t v[32], pk[16];
              tmem_ld_32x32b_x32(tmem_S + lane_off + c * 32, v);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; i += 2) {
                float p0 = fast_exp2(fmaf(__uint_as_float(v[i]), p.scale_log2e, -msc));
                float p1 = fast_exp2(fmaf(__uint_as_float(v[i + 1]), p.scale_log2e, -msc));
                if (c * 32 + i >= nvalid) p0 = 0.f;
                if (c * 32 + i + 1 >= nvalid) p1 = 0.f;
                row_sum += p0 + p1;
                pk[i >> 1] = pack_act2(p0, p1);
              }
              const uint32_t pa = p_row + (c >> 1) * 16384;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                st_shared_v4(pa + (((((c & 1) << 2) + j) ^ swz) << 4), pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sc_bar);
          }
        } else {
          // all 32 query rows of this warp are past Lq: no arithmetic, only the protocol
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(sc_bar);
          if (g > 0) ++o_cnt;                         // (never waited on by this warp; the count only tracks the phase)
        }
        fence_proxy_async_smem();                     // generic-proxy smem writes -> visible to the tensor core
        __syncwarp();
        if (lane == 0) mbar_arrive(pr_bar);           // 4 arrivals: P V (g) may be issued
      }
      // ---- epilogue: O / row_sum -> 64 columns of this head
      if (warp_live) {
        mbar_wait(o_bar, o_cnt & 1); ++o_cnt;         // P V (nkt - 1) complete
        tc_fence_after();
        const int q = qt * 128 + row;
        const float inv = 1.0f / row_sum;
        act_t* orow = p.out + (static_cast<long long>(b) * p.Lq + q) * p.ld_out + head * 64;
#pragma unroll
        for (int hcol = 0; hcol < 2; ++hcol) {
          uint32_t o[32];
          tmem_ld_32x32b_x32(tmem_O + lane_off + hcol * 32, o);
          tmem_ld_wait();
          if (q < p.Lq) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<uint4*>(orow + hcol * 32)[i] =
                  make_uint4(pack_act2(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv),
                             pack_act2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv),
                             pack_act2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv),
                             pack_act2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv));
          }
        }
        tc_fence_before();                            // O is read out before this warp's next pr_bar arrival lets P V (0) run
      } else {
        ++o_cnt;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace rmb
