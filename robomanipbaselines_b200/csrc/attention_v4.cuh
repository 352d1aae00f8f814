// Multi-head attention (head dim 64) on tcgen05/TMEM, fourth version: attention_v3.cuh with THREE query-tile streams per
// CTA and 64-key blocks. What bounds v3 is the MUFU unit inside the exponential section of a block and the latency of
// one warp outside it (scripts/attn_trace.py): with two streams a scheduler has two softmax warps, and whenever one of
// them loads S, takes its maximum, waits for P V or writes O the unit runs at the single-warp rate (12.5 clocks per ex2)
// or idles. Three streams put three softmax warps on every scheduler; they only fit with 64-key blocks:
//   TMEM per stream (160 columns, 480 of 512): S fp32 [128 x 64] | P 16-bit [128 x 64] in 32 columns | O fp32 [128 x 64]
//   shared memory per stream (72 KB): Q x 2 (16 KB), K x 3, V x 2 (8 KB: 64 keys x 64 d, 128B swizzle)
//   512 threads: warps 0-3 / 4-7 / 8-11 softmax of stream 0 / 1 / 2 (64 scores of a row in registers), warps 12-14 their
//   issuers, warp 15 idle; 128 registers per thread, no reallocation.
// Everything else (one sweep, lazily moved reference maximum, P as the TMEM A operand of P V, packed fp32x2 arithmetic,
// deferred output write, ragged last block) is v3's. (nn.MultiheadAttention inside
// third_party/act/detr/models/transformer.py:166-179,229-250; 8 heads x 64, no masks.)
#pragma once
#include "attention_v3.cuh"

namespace rmb {

constexpr int kAttn4Streams = 3;
constexpr int kAttn4Threads = 512;
constexpr uint32_t kAttn4KvBytes = 8192;                                            // 64 keys x 64 d x 2 B
constexpr int kAttn4StreamBytes = 2 * 16384 + 5 * 8192;                             // Q0 Q1 | K0 K1 K2 | V0 V1
constexpr int kAttn4SmemBytes = kAttn4Streams * kAttn4StreamBytes + 512 /*barriers*/ + 1024 /*alignment slack*/;

__global__ void __launch_bounds__(kAttn4Threads, 1)
attention_v4_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ AttnParams p, int n_items) {
  pdl_trigger();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars_all = reinterpret_cast<uint64_t*>(smem + kAttn4Streams * kAttn4StreamBytes);
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bars_all + kAttn4Streams * kAttn3BarsPerStream);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stream = (warp < 12) ? (warp >> 2) : (warp - 12);   // warp 15: no stream (idle)
#ifdef RMB_ATTN_TRACE
  int trace_k = 0;
#endif
  const uint32_t sbase = opaque(smem_u32(smem) + (stream % kAttn4Streams) * kAttn4StreamBytes);   // 32-bit shared-window addresses from here on
  const uint32_t sQ = sbase;              // 2 buffers
  const uint32_t sK = sbase + 2 * 16384;  // 3 stages
  const uint32_t sV = sbase + 2 * 16384 + 3 * kAttn4KvBytes;  // 2 stages
  const uint32_t bars = opaque(smem_u32(bars_all) + (stream % kAttn4Streams) * kAttn3BarsPerStream * 8);
  const uint32_t q_bar = bars;            // [2] Q landed
  const uint32_t k_bar = bars + 2 * 8;    // [3] K block landed
  const uint32_t v_bar = bars + 5 * 8;    // [2] V block landed
  const uint32_t s_bar = bars + 7 * 8;    // S = Q K^T complete
  const uint32_t o_bar = bars + 8 * 8;    // O += P V complete
  const uint32_t sc_bar = bars + 9 * 8;   // S pulled into registers by the stream's four softmax warps
  const uint32_t pr_bar = bars + 10 * 8;  // P written to TMEM by the stream's four softmax warps

  if (tid == 0) {
    for (int s = 0; s < kAttn4Streams; ++s) {
      uint64_t* b = bars_all + s * kAttn3BarsPerStream;
      for (int j = 0; j < 9; ++j) mbar_init(b + j, 1);
      mbar_init(b + 9, 4);
      mbar_init(b + 10, 4);
      for (int j = 11; j < 15; ++j) mbar_init(b + j, 1);   // exponential-section tokens, one per softmax warp
    }
    fence_mbar_init();
  }
  if (warp == 12 && lane == 0) { prefetch_tmap(&tmap_q); prefetch_tmap(&tmap_k); prefetch_tmap(&tmap_v); }
  if (warp == 0) { tmem_alloc(tmem_ptr_smem, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_wait();                                     // Q / K / V are the previous kernels' outputs
  const uint32_t tmem_base = *tmem_ptr_smem;
  const uint32_t tmem_S = opaque(tmem_base + (stream % kAttn4Streams) * 160);  // 64 fp32 columns
  const uint32_t tmem_P = tmem_S + 64;                       // 32 columns of packed 16-bit pairs
  const uint32_t tmem_O = tmem_S + 96;                       // 64 fp32 columns

  const int nkt = (p.Lk + 63) / 64;
  const int n_streams = kAttn4Streams * gridDim.x;
  const int first = kAttn4Streams * blockIdx.x + stream;
  const int my_items = (first < n_items) ? (n_items - first + n_streams - 1) / n_streams : 0;
  const int last_valid = p.Lk - (nkt - 1) * 64;              // valid keys of the last block, 1..64
  const int last_n16 = (last_valid + 15) & ~15;              // MMA width / P V depth of the last block

  if (warp >= 12) {
    // ------------------------------------------------------------ issuer of this stream: TMA loads and MMAs
    // One thread; the two MMA issues of a block sit directly behind the barrier they wait for (everything they need
    // is precomputed), loads and their index arithmetic come after them, off the softmax warps' critical path.
    if (warp < 15 && lane == 0 && my_items > 0) {
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
        mbar_expect_tx_a(k_bar + kl_st * 8, kAttn4KvBytes);
        tma_load_3d_a(sK + kl_st * kAttn4KvBytes, &tmap_k, k_bar + kl_st * 8, p.k_col0 + kc.head * 64, kc.g * 64, kc.b);
        advance(kc); ++kl;
        if (++kl_st == 3) kl_st = 0;
      };
      auto load_v = [&]() {
        const int st = vl & 1;
        mbar_expect_tx_a(v_bar + st * 8, kAttn4KvBytes);
        tma_load_3d_a(sV + st * kAttn4KvBytes, &tmap_v, v_bar + st * 8, p.v_col0 + vc.head * 64, vc.g * 64, vc.b);
        advance(vc); ++vl;
      };
      const uint64_t dq0 = make_smem_desc(sQ, 0, 1024, 2);
      const uint64_t dk0 = make_smem_desc(sK, 0, 1024, 2);
      const uint64_t dv0 = make_smem_desc(sV, 0, 1024, 2);
      constexpr uint32_t kIdescSFull = make_idesc_f16(128, 64, false, false);
      const uint32_t idesc_s_last = make_idesc_f16(128, last_n16, false, false);
      constexpr uint32_t kIdescO = make_idesc_f16(128, 64, false, true);       // V is MN-major
      const int last_ksteps = last_n16 >> 4;
      auto issue_s = [&](int qbuf, int kst, bool last) {                        // descriptors advance 1024 (x16 B) per tile
        const uint64_t dq = dq0 + static_cast<uint64_t>(qbuf * 1024);
        const uint64_t dk = dk0 + static_cast<uint64_t>(kst * (kAttn4KvBytes >> 4));
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
        ATRACE(warp == 12, 101);
        if (n + 1 < total) {                         // S(n+1): on the tensor pipe during the whole of block n's softmax
          if (g1 == 0) mbar_wait_lite(q_bar + (it1 & 1) * 8, (it1 >> 1) & 1);
          mbar_wait_lite(k_bar + ks1 * 8, kpar1);
          tc_fence_after();
          issue_s(it1 & 1, ks1, g1 == nkt - 1);
        }
        ATRACE(warp == 12, 102);
        if (kl < total) load_k();                    // K(n+3) -> stage ks
        if (g == nkt - 1 && it + 2 < my_items) load_q();   // last S of the item is complete: its Q buffer is free
        mbar_wait_lite(pr_bar, n & 1);                    // P(n) is in TMEM; its writers had waited for P V (n - 1)
        ATRACE(warp == 12, 103);
        mbar_wait_lite(v_bar + (n & 1) * 8, (n >> 1) & 1);
        ATRACE(warp == 12, 104);
        tc_fence_after();
        {
          const uint64_t dv = dv0 + static_cast<uint64_t>((n & 1) * (kAttn4KvBytes >> 4));
          if (g == nkt - 1) {
            for (int k = 0; k < last_ksteps; ++k) umma_f16_ts(tmem_O, tmem_P + 8 * k, dv + k * 128, kIdescO, (g | k) != 0);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)   // P: 8 TMEM columns per 16 keys; V: 16 keys = 2 groups of 8 x 128 B rows
              umma_f16_ts(tmem_O, tmem_P + 8 * k, dv + k * 128, kIdescO, (g | k) != 0);
          }
          umma_commit_u32(o_bar);
        }
        ATRACE(warp == 12, 105);
        if (n >= 1 && vl < total) load_v();          // V(n+1) -> the stage P V (n - 1) has finished with
        g = g1; it = it1; ks = ks1; kpar = kpar1;
      }
    }
  } else {
    // ------------------------------------------------------------ softmax warps of this stream: thread = query row
    const int wq = warp & 3;
    const int row = wq * 32 + lane;
    const uint32_t tmem_row = opaque(tmem_S + (static_cast<uint32_t>(wq * 32) << 16));   // this warp's 32 lanes of S
    const float scale = p.scale_log2e;
    const float2 scale2 = make_float2(scale, scale);
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
         tmem_ld_32x32b_x32(tmem_row + 96 + hcol * 32, o);
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
    // One key block of a live warp. The common case (all 128 keys valid) runs straight through: the ragged last block
    // of an item takes ONE side branch that masks its dead keys and skips the 32-key chunks without valid keys.
    auto do_block = [&](const int g, const int nvalid) {
      constexpr bool kFull = false;
      const int nchunks = (nvalid + 31) >> 5;
      uint32_t v[64];
#pragma unroll
      for (int c = 0; c < 2; ++c)
        if (kFull || c < nchunks) tmem_ld_32x32b_x32(tmem_row + c * 32, *reinterpret_cast<uint32_t(*)[32]>(&v[c * 32]));
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_a(sc_bar);             // the issuer may overwrite S: S(g+1) runs under this block's softmax
      ATRACE(wq == 0, 2);
      // ---- block maximum (keys past Lk: -inf drops out of the maximum and gives P = 0)
      float bm0 = -INFINITY, bm1 = -INFINITY, bm2 = -INFINITY, bm3 = -INFINITY;
      if (nvalid < 64) {
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (i >= nvalid) v[i] = 0xff800000u;        // (also the chunks that were not loaded)
      }
#pragma unroll
      for (int i = 0; i < 64; i += 8) {
        bm0 = max3(bm0, __uint_as_float(v[i]), __uint_as_float(v[i + 1]));
        bm1 = max3(bm1, __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        bm2 = max3(bm2, __uint_as_float(v[i + 4]), __uint_as_float(v[i + 5]));
        bm3 = max3(bm3, __uint_as_float(v[i + 6]), __uint_as_float(v[i + 7]));
      }
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
             tmem_ld_32x32b_x32(tmem_row + 96 + hcol * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32b_x32(tmem_row + 96 + hcol * 32, o);
          }
          tmem_st_wait();
        }
      }
      ATRACE(wq == 0, 3);
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
        // chunk 0 is formed before the wait for P V (n - 1) (which still reads the P columns)
        uint32_t pk0[16];
        exp_chunk(std::integral_constant<int, 0>{}, pk0);
        ATRACE(wq == 0, 4);
        if (n > 0 && !o_waited) {
          mbar_wait_lite(o_bar, (n - 1) & 1);
          tc_fence_after();
        }
        ATRACE(wq == 0, 5);
        tmem_st_32x32b_x16(tmem_row + 64, pk0);
      }
      if (kFull || 1 < nchunks) {
        uint32_t pk[16];
        exp_chunk(std::integral_constant<int, 1>{}, pk);
        tmem_st_32x32b_x16(tmem_row + 64 + 16, pk);
      }
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
        }
        continue;
      }
      row_sum = 0.f;
      for (int g = 0; g < nkt; ++g, ++n) {
        mbar_wait_lite(s_bar, n & 1);
        tc_fence_after();
        ATRACE(wq == 0, 1);
        do_block(g, (g == nkt - 1) ? last_valid : 64);
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
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace rmb
