// Multi-head attention (head dim 64) on tcgen05/TMEM, third version: the structure of attention_v2.cuh (persistent CTAs,
// two independent query-tile STREAMS per CTA, one sweep, thread = query row) with the hand-offs and the instruction
// stream that ncu showed bounding v2 (profiles/r02_attention_v2_ncu_full_summary.txt: issue slots 32 % busy, 6.9
// instructions per score, softmax warps waiting on S / P hand-offs) rebuilt:
// (nn.MultiheadAttention inside third_party/act/detr/models/transformer.py:166-179,229-250; 8 heads x 64, no masks.)
//   * P never touches shared memory: the softmax warps write it with tcgen05.st into 64 TMEM columns of their stream
//     (row = lane, two 16-bit probabilities per 32-bit column) and O += P V runs as a tcgen05.mma with the A operand IN
//     TMEM. That removes 32 KB of st.shared + a proxy fence per block and the shared-memory read of A that kept the
//     N = 64 P V MMAs at 2/3 of the tensor rate (192 B/clk wanted, 128 available);
//   * the whole 128-key S block of a row is pulled into registers up front and the S buffer is released before any
//     arithmetic, so S(g+1) is on the tensor pipe during ALL of block g's softmax (v2 released it after the last
//     32-key chunk had been scanned); K blocks come three ahead, Q of the stream's next work item is double-buffered;
//   * the reference maximum is settled before the exponentials (one max pass over the registers, 4 independent
//     3-input-max chains), so there is no redo path; it only moves (O and the row sum rescaled in TMEM) when a block
//     exceeds it by more than 2^8, which with LayerNorm-bounded logits is never after block 0;
//   * per score: half an FFMA2 (packed fp32x2 scale-and-shift), one EX2, half an FADD2, half a pack, half a 3-input
//     max = 3 issue slots (v2: 6.9); masking only touches the ragged last chunk;
//   * the ragged last key block runs S with N = ceil16(valid keys) and P V over that many keys only.
//   TMEM per stream (256 columns): S fp32 [128 x 128] | P 16-bit [128 x 128] in 64 columns | O fp32 [128 x 64].
//   Shared memory per stream (112 KB): Q x 2, K x 3, V x 2 tiles of 16 KB (128 rows x 64 d, 128B swizzle).
#pragma once
#include <cuda.h>
#include <type_traits>
#include "attention_v2.cuh"

namespace rmb {

constexpr int kAttn3Threads = 384;   // warps 0-3 / 4-7: softmax of stream 0 / 1; warps 8, 9: their issuers; 10, 11 idle
constexpr int kAttn3StreamBytes = 7 * 16384;                                        // Q0 Q1 K0 K1 K2 V0 V1
constexpr int kAttn3SmemBytes = 2 * kAttn3StreamBytes + 512 /*barriers*/ + 1024 /*alignment slack*/;
constexpr int kAttn3BarsPerStream = 16;

// D[tmem] (+)= A[tmem] * B[smem]: A is [128 x 16] 16-bit, row = lane, 8 columns of two packed values each
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
// packed fp32x2 arithmetic (sm_100): one issue slot for two lanes
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)), "l"(reinterpret_cast<uint64_t&>(c)));
  return d;
}
// Register reallocation between warpgroups: the kernel is launched at 168 registers per thread (65536 / 384); the
// issuer warpgroup gives back what it does not need and the two softmax warpgroups take it (a row's 128 scores stay in
// registers). All four warps of a warpgroup execute the instruction.
__device__ __forceinline__ void setmaxnreg_dec_56() { asm volatile("setmaxnreg.dec.sync.aligned.u32 56;" ::: "memory"); }
__device__ __forceinline__ void setmaxnreg_inc_224() { asm volatile("setmaxnreg.inc.sync.aligned.u32 224;" ::: "memory"); }
// mbar_wait without the printf of ptx.cuh's (its argument set-up costs registers at every inlined site, and the issuer
// thread runs on a small budget): same time-out, same fault flag (code 3), same "carry on" behaviour.
// Barriers are addressed by their 32-bit shared-window address (one register per stream + immediate offsets): a
// generic pointer costs a cvta sequence of ~10 uniform-datapath instructions at every use.
__device__ __forceinline__ bool mbar_try_wait_a(uint32_t bar_addr, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}\n"
      : "=r"(ok)
      : "r"(bar_addr), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait_lite(uint32_t bar_addr, uint32_t parity) {
  if (mbar_try_wait_a(bar_addr, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait_a(bar_addr, parity)) {
    if (clock64() - t0 > RMB_MBAR_TIMEOUT_CYCLES) {
      atomicCAS(&g_rmb_fault, 0, 3);
      return;
    }
  }
}
// keeps ptxas from re-deriving a value from threadIdx / the shared window base at every use (it rematerialises such
// "cheap" chains of ~12 instructions rather than hold one register)
__device__ __forceinline__ uint32_t opaque(uint32_t x) {
  uint32_t y;
  asm volatile("mov.u32 %0, %1;" : "=r"(y) : "r"(x));
  return y;
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_a(uint32_t bar_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_3d_a(uint32_t smem_dst, const void* tmap, uint32_t bar_addr, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_addr), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// Part of the exponentials runs on the FMA pipe instead of the MUFU unit (16 ex2 per clock per SM is what bounds the
// kernel: one exponential per score against 256 tensor flops at head dim 64). 2^t for t <= ~8: n = round(t) by the
// magic-number add, f = t - n in [-0.5, 0.5], degree-3 minimax polynomial (max relative error 7.5e-5, below the 16-bit
// rounding of P), n added into the exponent field. Of every 16 score pairs of a chunk, RMB_ATTN_POLY_NUM * 4 of the
// odd ones take this path. Measured on B200 at 256 episodes x 602 keys: 0.3645 ms without, 0.3674 / 0.3752 / 0.3802 ms
// with 25 / 37.5 / 50 % of the scores on the polynomial (an FFMA2 occupies the FMA pipe for as long as two EX2 occupy
// the MUFU unit, and the kernel was latency- not MUFU-bound), so the default is 0 (off).
#ifndef RMB_ATTN_PINGPONG
#define RMB_ATTN_PINGPONG 1
#endif
#ifndef RMB_ATTN_STAGGER
#define RMB_ATTN_STAGGER 0
#endif
#ifndef RMB_ATTN_MAX_CHAINS
#define RMB_ATTN_MAX_CHAINS 4
#endif
#ifndef RMB_ATTN_POLY_NUM
#define RMB_ATTN_POLY_NUM 0
#endif
__device__ __forceinline__ float2 poly_exp2_x2(float2 t) {
  t.x = fmaxf(t.x, -125.0f); t.y = fmaxf(t.y, -125.0f);
  const float2 magic = make_float2(12582912.0f, 12582912.0f), nmagic = make_float2(-12582912.0f, -12582912.0f);
  const float2 none = make_float2(-1.0f, -1.0f);
  float2 j, r, f, q;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(j)) : "l"(reinterpret_cast<uint64_t&>(t)), "l"(reinterpret_cast<const uint64_t&>(magic)));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(reinterpret_cast<uint64_t&>(r)) : "l"(reinterpret_cast<uint64_t&>(j)), "l"(reinterpret_cast<const uint64_t&>(nmagic)));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(f)) : "l"(reinterpret_cast<uint64_t&>(r)), "l"(reinterpret_cast<const uint64_t&>(none)), "l"(reinterpret_cast<uint64_t&>(t)));
  const float2 c3 = make_float2(0.0551716685295105f, 0.0551716685295105f), c2 = make_float2(0.2426111251115799f, 0.2426111251115799f);
  const float2 c1 = make_float2(0.6932609677314758f, 0.6932609677314758f), c0 = make_float2(0.9999280571937561f, 0.9999280571937561f);
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(q)) : "l"(reinterpret_cast<uint64_t&>(f)), "l"(reinterpret_cast<const uint64_t&>(c3)), "l"(reinterpret_cast<const uint64_t&>(c2)));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(q)) : "l"(reinterpret_cast<uint64_t&>(q)), "l"(reinterpret_cast<uint64_t&>(f)), "l"(reinterpret_cast<const uint64_t&>(c1)));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(reinterpret_cast<uint64_t&>(q)) : "l"(reinterpret_cast<uint64_t&>(q)), "l"(reinterpret_cast<uint64_t&>(f)), "l"(reinterpret_cast<const uint64_t&>(c0)));
  return make_float2(__uint_as_float(__float_as_uint(q.x) + (__float_as_uint(j.x) << 23)),
                     __uint_as_float(__float_as_uint(q.y) + (__float_as_uint(j.y) << 23)));
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<uint64_t&>(d))
      : "l"(reinterpret_cast<uint64_t&>(a)), "l"(reinterpret_cast<uint64_t&>(b)));
  return d;
}

// RMB_ATTN_TRACE (scripts/attn_trace.py builds a separate library with it): CTA 0 / stream 0 logs (event, clock64) pairs
// of its first softmax warp and of its issuer thread; never compiled into the product libraries.
#ifdef RMB_ATTN_TRACE
__device__ long long g_attn_trace[2 * 16384];     // [0, 16384): softmax warp, [16384, 32768): issuer; (id << 48) | clock
__device__ int g_attn_trace_n;
#define ATRACE(cond, id)                                                                                   \
  do {                                                                                                     \
    if (blockIdx.x == 0 && stream == 0 && lane == 0 && (cond) && trace_k < 16384) {                        \
      g_attn_trace[((id) >= 100 ? 16384 : 0) + trace_k] = (static_cast<long long>(id) << 48) | (clock64() & 0xffffffffffffll); \
      ++trace_k;                                                                                           \
    }                                                                                                      \
  } while (0)
#else
#define ATRACE(cond, id)
#endif

__global__ void __launch_bounds__(kAttn3Threads, 1)
attention_v3_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnParams p, int n_items) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars_all = reinterpret_cast<uint64_t*>(smem + 2 * kAttn3StreamBytes);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars_all + 2 * kAttn3BarsPerStream);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stream = (warp < 8) ? (warp >> 2) : (warp & 1);
#ifdef RMB_ATTN_TRACE
  int trace_k = 0;
#endif
  const uint32_t sbase = opaque(smem_u32(smem) + stream * kAttn3StreamBytes);   // 32-bit shared-window addresses from here on
  const uint32_t sQ = sbase;              // 2 buffers
  const uint32_t sK = sbase + 2 * 16384;  // 3 stages
  const uint32_t sV = sbase + 5 * 16384;  // 2 stages
  const uint32_t bars = opaque(smem_u32(bars_all) + stream * kAttn3BarsPerStream * 8);
  const uint32_t q_bar = bars;            // [2] Q landed
  const uint32_t k_bar = bars + 2 * 8;    // [3] K block landed
  const uint32_t v_bar = bars + 5 * 8;    // [2] V block landed
  const uint32_t s_bar = bars + 7 * 8;    // S = Q K^T complete
  const uint32_t o_bar = bars + 8 * 8;    // O += P V complete
  const uint32_t sc_bar = bars + 9 * 8;   // S pulled into registers by the stream's four softmax warps
  const uint32_t pr_bar = bars + 10 * 8;  // P written to TMEM by the stream's four softmax warps

  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      uint64_t* b = bars_all + s * kAttn3BarsPerStream;
      for (int j = 0; j < 9; ++j) mbar_init(b + j, 1);
      mbar_init(b + 9, 4);
      mbar_init(b + 10, 4);
      for (int j = 11; j < 15; ++j) mbar_init(b + j, 1);   // exponential-section tokens, one per softmax warp
    }
    fence_mbar_init();
#if RMB_ATTN_PINGPONG
    for (int j = 11; j < 15; ++j) mbar_arrive(bars_all + j);   // stream 0 goes first: its round-0 grants
#endif
  }
  if (warp == 8 && lane == 0) { prefetch_tmap(&tmap_q); prefetch_tmap(&tmap_k); prefetch_tmap(&tmap_v); }
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                                     // Q / K / V are the previous kernels' outputs
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_S = opaque(tmem_base + stream * 256);  // 128 fp32 columns
  const uint32_t tmem_P = tmem_S + 128;                      // 64 columns of packed 16-bit pairs
  const uint32_t tmem_O = tmem_S + 192;                      // 64 fp32 columns

  const int nkt = (p.Lk + 127) / 128;
  const int n_streams = 2 * gridDim.x;
  const int first = 2 * blockIdx.x + stream;
  const int my_items = (first < n_items) ? (n_items - first + n_streams - 1) / n_streams : 0;
  const int last_valid = p.Lk - (nkt - 1) * 128;             // valid keys of the last block, 1..128
  const int last_n16 = (last_valid + 15) & ~15;              // MMA width / P V depth of the last block

  if (warp >= 8) {
    setmaxnreg_dec_56();
    // ------------------------------------------------------------ issuer of this stream: TMA loads and MMAs
    // One thread; the two MMA issues of a block sit directly behind the barrier they wait for (everything they need
    // is precomputed), loads and their index arithmetic come after them, off the softmax warps' critical path.
    if (warp < 10 && lane == 0 && my_items > 0) {
      const int total = my_items * nkt;              // key blocks of this stream, numbered through all its items
      struct Cursor { int it, g, b, head, qt; };     // a position in the stream's block sequence
      auto set_item = [&](Cursor& c, int it) {
        const int item = first + it * n_streams;
        c.it = it; c.g = 0;
        c.qt = item % p.n_q_tiles;
        c.head = (item / p.n_q_tiles) % p.n_heads;
        c.b = item / (p.n_q_tiles * p.n_heads);
      };
      auto advance = [&](Cursor& c) { if (++c.g == nkt) set_item(c, c.it + 1); };
      Cursor kc, vc, qc;
      set_item(kc, 0); set_item(vc, 0); set_item(qc, 0);
      int kl = 0, vl = 0, kl_st = 0;                 // K / V blocks requested so far, K stage of the next request
      auto load_q = [&]() {                          // Q of item qc.it -> buffer it & 1
        const int buf = qc.it & 1;
        mbar_expect_tx_a(q_bar + buf * 8, 16384u);
        tma_load_3d_a(sQ + buf * 16384, &tmap_q, q_bar + buf * 8, p.q_col0 + qc.head * 64, qc.qt * 128, p.q_shared ? 0 : qc.b);
        set_item(qc, qc.it + 1);
      };
      auto load_k = [&]() {
        mbar_expect_tx_a(k_bar + kl_st * 8, 16384u);
        tma_load_3d_a(sK + kl_st * 16384, &tmap_k, k_bar + kl_st * 8, p.k_col0 + kc.head * 64, kc.g * 128, kc.b);
        advance(kc); ++kl;
        if (++kl_st == 3) kl_st = 0;
      };
      auto load_v = [&]() {
        const int st = vl & 1;
        mbar_expect_tx_a(v_bar + st * 8, 16384u);
        tma_load_3d_a(sV + st * 16384, &tmap_v, v_bar + st * 8, p.v_col0 + vc.head * 64, vc.g * 128, vc.b);
        advance(vc); ++vl;
      };
      const uint64_t dq0 = make_smem_desc(sQ, 0, 1024, 2);
      const uint64_t dk0 = make_smem_desc(sK, 0, 1024, 2);
      const uint64_t dv0 = make_smem_desc(sV, 0, 1024, 2);
      constexpr uint32_t kIdescSFull = make_idesc_f16(128, 128, false, false);
      const uint32_t idesc_s_last = make_idesc_f16(128, last_n16, false, false);
      constexpr uint32_t kIdescO = make_idesc_f16(128, 64, false, true);       // V is MN-major
      const int last_ksteps = last_n16 >> 4;
      auto issue_s = [&](int qbuf, int kst, bool last) {                        // descriptors advance 1024 (x16 B) per tile
        const uint64_t dq = dq0 + static_cast<uint64_t>(qbuf * 1024);
        const uint64_t dk = dk0 + static_cast<uint64_t>(kst * 1024);
        const uint32_t idesc = last ? idesc_s_last : kIdescSFull;
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_f16_ss(tmem_S, dq + 2 * k, dk + 2 * k, idesc, k != 0);
        umma_commit_u32(s_bar);
      };
      load_q();
      if (my_items > 1) load_q();
      for (int j = 0; j < 3 && j < total; ++j) load_k();
      load_v();
      if (total > 1) load_v();
      mbar_wait_lite(q_bar, 0);
      mbar_wait_lite(k_bar, 0);
      tc_fence_after();
      issue_s(0, 0, nkt == 1);
      int g = 0, it = 0, ks = 0, kpar = 0;           // block n: index in its item, item ordinal, K stage, K parity
      for (int n = 0; n < total; ++n) {
        int g1 = g + 1, it1 = it, ks1 = ks + 1, kpar1 = kpar;
        if (g1 == nkt) { g1 = 0; it1 = it + 1; }
        if (ks1 == 3) { ks1 = 0; kpar1 ^= 1; }
        mbar_wait_lite(sc_bar, n & 1);                    // S(n) is in registers: the S buffer and K stage ks are free
        ATRACE(warp == 8, 101);
        if (n + 1 < total) {                         // S(n+1): on the tensor pipe during the whole of block n's softmax
          if (g1 == 0) mbar_wait_lite(q_bar + (it1 & 1) * 8, (it1 >> 1) & 1);
          mbar_wait_lite(k_bar + ks1 * 8, kpar1);
          tc_fence_after();
          issue_s(it1 & 1, ks1, g1 == nkt - 1);
        }
        ATRACE(warp == 8, 102);
        if (kl < total) load_k();                    // K(n+3) -> stage ks
        if (g == nkt - 1 && it + 2 < my_items) load_q();   // last S of the item is complete: its Q buffer is free
        mbar_wait_lite(pr_bar, n & 1);                    // P(n) is in TMEM; its writers had waited for P V (n - 1)
        ATRACE(warp == 8, 103);
        mbar_wait_lite(v_bar + (n & 1) * 8, (n >> 1) & 1);
        ATRACE(warp == 8, 104);
        tc_fence_after();
        {
          const uint64_t dv = dv0 + static_cast<uint64_t>((n & 1) * 1024);
          if (g == nkt - 1) {
            for (int k = 0; k < last_ksteps; ++k) umma_f16_ts(tmem_O, tmem_P + 8 * k, dv + k * 128, kIdescO, (g | k) != 0);
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k)   // P: 8 TMEM columns per 16 keys; V: 16 keys = 2 groups of 8 x 128 B rows
              umma_f16_ts(tmem_O, tmem_P + 8 * k, dv + k * 128, kIdescO, (g | k) != 0);
          }
          umma_commit_u32(o_bar);
        }
        ATRACE(warp == 8, 105);
        if (n >= 1 && vl < total) load_v();          // V(n+1) -> the stage P V (n - 1) has finished with
        g = g1; it = it1; ks = ks1; kpar = kpar1;
      }
    }
  } else {
    setmaxnreg_inc_224();
    // ------------------------------------------------------------ softmax warps of this stream: thread = query row
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t tmem_row = opaque(tmem_S + (static_cast<uint32_t>(wq * 32) << 16));   // this warp's 32 lanes of S
    const float scale = p.scale_log2e;
    const float2 scale2 = make_float2(scale, scale);
#if RMB_ATTN_PINGPONG
    // The two softmax warps of a scheduler (same wq, one per stream) take turns at the exponential section of a block.
    // Free-running they fall into lock-step (scripts/attn_trace.py): both in the section at once, each getting half
    // of the MUFU unit, then both outside it with the unit idle. With a token that alternates between them one warp
    // has the unit to itself while the other loads S, takes its maximum and writes O. Round r of a warp = its r-th key
    // block; a stream with fewer blocks (or none) keeps passing the token until the longer one is done.
    const uint32_t tok_mine = bars + (11 + wq) * 8;
    const uint32_t tok_other = opaque(smem_u32(bars_all) + (1 - stream) * kAttn3BarsPerStream * 8) + (11 + wq) * 8;
    const int first_o = 2 * blockIdx.x + (1 - stream);
    const int items_o = (first_o < n_items) ? (n_items - first_o + n_streams - 1) / n_streams : 0;
    const int rounds_total = (my_items > items_o ? my_items : items_o) * nkt;
#endif
    uint32_t n = 0;                                   // key blocks of this stream seen so far (= S / P V phase counter)
    // The output of an item is written one block late: its last P V is still on the tensor pipe when the item's
    // softmax ends, so the warp goes on to block 0 of its next item and reads O just before it lets that block's P V
    // (which overwrites O) be issued.
    bool pend = false;
    float pend_inv = 0.f;
    act_t* pend_out = nullptr;
    auto write_out = [&]() {                          // O of the pending item -> 64 columns of its head
#pragma unroll
      for (int hcol = 0; hcol < 2; ++hcol) {
        uint32_t o[32];
        tmem_ld_32x32b_x32(tmem_row + 192 + hcol * 32, o);
        tmem_ld_wait();
        if (pend_out != nullptr) {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            reinterpret_cast<uint4*>(pend_out + hcol * 32)[i] =
                make_uint4(pack_act2(__uint_as_float(o[8 * i]) * pend_inv, __uint_as_float(o[8 * i + 1]) * pend_inv),
                           pack_act2(__uint_as_float(o[8 * i + 2]) * pend_inv, __uint_as_float(o[8 * i + 3]) * pend_inv),
                           pack_act2(__uint_as_float(o[8 * i + 4]) * pend_inv, __uint_as_float(o[8 * i + 5]) * pend_inv),
                           pack_act2(__uint_as_float(o[8 * i + 6]) * pend_inv, __uint_as_float(o[8 * i + 7]) * pend_inv));
        }
      }
      pend = false;
    };
    float m = 0.f, row_sum = 0.f;                          // reference maximum (raw logits), sum of P of the current item
#if RMB_ATTN_STAGGER > 0
    if (stream == 1) {                                // start the second stream half a block late (see RMB_ATTN_STAGGER)
      const long long t_start = clock64();
      while (clock64() - t_start < RMB_ATTN_STAGGER) {}
    }
#endif
    // One key block of a live warp. The common case (all 128 keys valid) runs straight through: the ragged last block
    // of an item takes ONE side branch that masks its dead keys and skips the 32-key chunks without valid keys.
    auto do_block = [&](const int g, const int nvalid) {
      constexpr bool kFull = false;
      const int nchunks = (nvalid + 31) >> 5;
      uint32_t v[128];
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (kFull || c < nchunks) tmem_ld_32x32b_x32(tmem_row + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&v[c * 32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(sc_bar);             // the issuer may overwrite S: S(g+1) runs under this block's softmax
      ATRACE(wq == 0, 2);
      // ---- block maximum (keys past Lk: -inf drops out of the maximum and gives P = 0)
      float bm0 = -INFINITY, bm1 = -INFINITY, bm2 = -INFINITY, bm3 = -INFINITY;
      if (nvalid < 128) {
#pragma unroll
        for (int i = 0; i < 128; ++i)
          if (i >= nvalid) v[i] = 0xff800000u;        // (also the chunks that were not loaded)
      }
#if RMB_ATTN_MAX_CHAINS == 8
      float bm4 = -INFINITY, bm5 = -INFINITY, bm6 = -INFINITY, bm7 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 128; i += 16) {
        bm0 = max3(bm0, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
        bm1 = max3(bm1, __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        bm2 = max3(bm2, __uint_as_float(v[i + 4]), __uint_as_float(v[i + 5]));
        bm3 = max3(bm3, __uint_as_float(v[i + 6]), __uint_as_float(v[i + 7]));
        bm4 = max3(bm4, __uint_as_float(v[i + 8]), __uint_as_float(v[i + 9]));
        bm5 = max3(bm5, __uint_as_float(v[i + 10]), __uint_as_float(v[i + 11]));
        bm6 = max3(bm6, __uint_as_float(v[i + 12]), __uint_as_float(v[i + 13]));
        bm7 = max3(bm7, __uint_as_float(v[i + 14]), __uint_as_float(v[i + 15]));
      }
      bm0 = max3(bm0, bm4, bm5); bm1 = max3(bm1, bm6, bm7);
#else
#pragma unroll
      for (int i = 0; i < 128; i += 8) {
        bm0 = max3(bm0, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
        bm1 = max3(bm1, __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        bm2 = max3(bm2, __uint_as_float(v[i + 4]), __uint_as_float(v[i + 5]));
        bm3 = max3(bm3, __uint_as_float(v[i + 6]), __uint_as_float(v[i + 7]));
      }
#endif
      const float bm = fmaxf(fmaxf(bm0, bm1), fmaxf(bm2, bm3));
      bool o_waited = false;
      if (g == 0) {
        m = bm;                                       // nothing accumulated yet
      } else {
        const bool need = (bm - m) * scale > kAttn2RescaleLog2;
        if (__any_sync(0xffffffffu, need)) {
          // ---- rare (warp-uniform): move the reference point of the rows that need it and rescale what they have
          // accumulated. P V (n - 1) must have completed; P V (n) is not issued before this warp's pr_bar arrival.
          mbar_wait_lite(o_bar, (n - 1) & 1);
          tc_fence_after();
          o_waited = true;
          const float alpha = need ? fast_exp2((m - bm) * scale) : 1.0f;
          if (need) m = bm;
          row_sum *= alpha;
#pragma unroll 1
          for (int hcol = 0; hcol < 2; ++hcol) {
            uint32_t o[32];
            tmem_ld_32x32b_x32(tmem_row + 192 + hcol * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x32(tmem_row + 192 + hcol * 32, o);
          }
          tmem_st_wait();
        }
      }
      ATRACE(wq == 0, 3);
#if RMB_ATTN_PINGPONG
      mbar_wait_lite(tok_mine, n & 1);
      ATRACE(wq == 0, 9);
#endif
      const float nmsc = -m * scale;
      const float2 nmsc2 = make_float2(nmsc, nmsc);
      float2 acc0 = make_float2(0.f, 0.f), acc1 = make_float2(0.f, 0.f);
      // One 32-key chunk in three passes over its registers -- scale-and-shift all, exponentiate all, then sum and pack
      // all -- so that the 32 EX2 go to the MUFU unit back to back. Interleaved per pair (FFMA2, 2 x EX2, FADD2, pack)
      // a warp issues in order and the FADD2 behind a pair waits out the MUFU latency: the block timeline
      // (scripts/attn_trace.py) showed ~16 clocks per exponential per warp, twice the unit's 8, with two warps a scheduler.
      auto exp_chunk = [&](auto c_tag, uint32_t (&pk)[16]) {
        constexpr int c = decltype(c_tag)::value;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          const float2 t = ffma2(make_float2(__uint_as_float(v[c * 32 + i]), __uint_as_float(v[c * 32 + i + 1])), scale2, nmsc2);
          v[c * 32 + i] = __float_as_uint(t.x);
          v[c * 32 + i + 1] = __float_as_uint(t.y);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
#if RMB_ATTN_POLY_NUM > 0
          if ((i & 3) >= 2 && ((i >> 2) & 3) < RMB_ATTN_POLY_NUM) {
            if (i & 1) continue;
            const float2 e = poly_exp2_x2(make_float2(__uint_as_float(v[c * 32 + i]), __uint_as_float(v[c * 32 + i + 1])));
            v[c * 32 + i] = __float_as_uint(e.x);
            v[c * 32 + i + 1] = __float_as_uint(e.y);
            continue;
          }
#endif
          v[c * 32 + i] = __float_as_uint(fast_exp2(__uint_as_float(v[c * 32 + i])));
        }
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const float2 e0 = make_float2(__uint_as_float(v[c * 32 + 2 * i]), __uint_as_float(v[c * 32 + 2 * i + 1]));
          const float2 e1 = make_float2(__uint_as_float(v[c * 32 + 2 * i + 2]), __uint_as_float(v[c * 32 + 2 * i + 3]));
          acc0 = fadd2(acc0, e0);
          acc1 = fadd2(acc1, e1);
          pk[i] = pack_act2(e0.x, e0.y);
          pk[i + 1] = pack_act2(e1.x, e1.y);
        }
      };
      {
        // chunks 0 and 1 are formed before the wait for P V (n - 1) (which still reads the P columns): its latency
        // hides under 64 exponentials per row
        uint32_t pk0[16], pk1[16];
        exp_chunk(std::integral_constant<int, 0>{}, pk0);
        if (kFull || nchunks > 1) exp_chunk(std::integral_constant<int, 1>{}, pk1);
        ATRACE(wq == 0, 4);
        if (n > 0 && !o_waited) {
          mbar_wait_lite(o_bar, (n - 1) & 1);
          tc_fence_after();
        }
        ATRACE(wq == 0, 5);
        tmem_st_32x32b_x16(tmem_row + 128, pk0);
        if (kFull || nchunks > 1) tmem_st_32x32b_x16(tmem_row + 128 + 16, pk1);
      }
      if (kFull || 2 < nchunks) {
        uint32_t pk[16];
        exp_chunk(std::integral_constant<int, 2>{}, pk);
        tmem_st_32x32b_x16(tmem_row + 128 + 32, pk);
      }
      if (kFull || 3 < nchunks) {
        uint32_t pk[16];
        exp_chunk(std::integral_constant<int, 3>{}, pk);
        tmem_st_32x32b_x16(tmem_row + 128 + 48, pk);
      }
#if RMB_ATTN_PINGPONG
      __syncwarp();
      if (lane == 0) mbar_arrive_a(tok_other);
#endif
      row_sum += (acc0.x + acc0.y) + (acc1.x + acc1.y);
      ATRACE(wq == 0, 6);
      if (pend) write_out();                          // g == 0: O of the previous item, complete since the o_bar wait above
      ATRACE(wq == 0, 8);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(pr_bar);             // 4 arrivals: P V (n) may be issued
      ATRACE(wq == 0, 7);
    };
    for (int it = 0; it < my_items; ++it) {
      const int item = first + it * n_streams;
      const int qt = item % p.n_q_tiles;
      const int head = (item / p.n_q_tiles) % p.n_heads;
      const int b = item / (p.n_q_tiles * p.n_heads);
      if (qt * 128 + wq * 32 >= p.Lq) {               // warp-uniform
        // all 32 query rows of this warp are past Lq: no arithmetic, only the protocol. The arrival on pr_bar for
        // block n must not land in phase n - 1, which the live warps may still be working towards.
        for (int g = 0; g < nkt; ++g, ++n) {
          mbar_wait_lite(s_bar, n & 1);
          tc_fence_after();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_a(sc_bar);
          if (n > 0) {
            if (pend) {                               // (this warp was live in the previous item)
              mbar_wait_lite(o_bar, (n - 1) & 1);
              tc_fence_after();
              write_out();
              tc_fence_before();
            } else {
              mbar_wait_lite(pr_bar, (n - 1) & 1);
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive_a(pr_bar);
          __syncwarp();
#if RMB_ATTN_PINGPONG
          mbar_wait_lite(tok_mine, n & 1);
          __syncwarp();
          if (lane == 0) mbar_arrive_a(tok_other);
#endif
        }
        continue;
      }
      row_sum = 0.f;
      for (int g = 0; g < nkt; ++g, ++n) {
        mbar_wait_lite(s_bar, n & 1);
        tc_fence_after();
        ATRACE(wq == 0, 1);
        do_block(g, (g == nkt - 1) ? last_valid : 128);
      }
      const int q = qt * 128 + row;
      pend = true;
      pend_inv = 1.0f / row_sum;
      pend_out = (q < p.Lq) ? p.out + (static_cast<long long>(b) * p.Lq + q) * p.ld_out + head * 64 : nullptr;
    }
    if (pend) {                                       // the stream's last item
      mbar_wait_lite(o_bar, (n - 1) & 1);
      tc_fence_after();
      write_out();
    }
#if RMB_ATTN_PINGPONG
    for (uint32_t r = n; r < static_cast<uint32_t>(rounds_total); ++r) {   // keep the partner's token moving
      mbar_wait_lite(tok_mine, r & 1);
      __syncwarp();
      if (lane == 0) mbar_arrive_a(tok_other);
    }
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace rmb
