// tcgen05/TMEM implicit-GEMM kernel shared by every dense contraction on the ACT path:
//   * ResNet-18 convs (NHWC 16-bit, FrozenBN folded into weights/bias)  -- reference: third_party/act/detr/models/backbone.py:21-57,86-96
//   * 1x1 input_proj + token scatter                                   -- third_party/act/detr/models/detr_vae.py:131-141
//   * transformer linear layers                                        -- third_party/act/detr/models/transformer.py:166-179,229-250
//
// D[rows, N] = A[rows, K] * Wt[N, K]^T  (+bias, +residual, ReLU), fp32 accumulation in TMEM.
// A rows are output pixels. The A tile of one k-block is ONE TMA box of a rank-5 tensor map: a rectangle of
// (box_w x box_h x box_b) pixels x BLOCK_K channels, shifted by the filter tap (out-of-bounds -> zero = conv padding).
// A plain matrix is the degenerate case (1 tap, box_w = 128 rows).
#pragma once
#include <cuda.h>
#include "ptx.cuh"

namespace rmb {

constexpr int kMaxTaps = 9;
constexpr int kGemmThreads = 320;  // warps 0-7: epilogue, warp 8: TMA producer, warp 9: MMA issuer + TMEM owner

struct GemmParams {
  // row (pixel) tiling
  int tiles_w, tiles_h, tiles_b;
  int box_w, box_h, box_b;   // rows per tile = box_w*box_h*box_b <= 128
  int W, H, B;               // output extents (rows = B*H*W)
  int n_col_tiles;           // ceil(N / BLOCK_N)
  int n_valid;               // N: columns at or past it (ragged last column tile, plain epilogue only) are not stored
  // K loop: k-block kb -> (tap = kb / chunks_per_tap, chunk = kb % chunks_per_tap)
  int num_taps, chunks_per_tap;
  int tap_c0[kMaxTaps];      // added to coordinate 0 (channel / phase-channel offset)
  int tap_d1[kMaxTaps];      // added to coordinate 1 (w)
  int tap_d2[kMaxTaps];      // coordinate 2 (h phase for stride-2 views, else 0)
  int tap_d3[kMaxTaps];      // added to coordinate 3 (h)
  // halo3 (3x3 stride-1 convs, box_w * box_h == 128, box_b == 1, box_w % 8 == 0): a pipeline stage holds, for one filter
  // COLUMN kw and one 64-channel chunk, the (box_h + 2)-row input box ONCE plus the three weight tiles of kh = 0..2; the
  // three row taps are MMAs over views of that box shifted by kh * box_w rows (a multiple of the 1024-byte swizzle
  // period). A is then fetched 3 * (box_h + 2) / box_h times instead of 9 -- what bounded the N = 128 convs was the
  // L2 -> SM operand stream (96 B/clk per SM wanted at 128-wide tiles, ~61 available).
  int halo3, halo_stages, halo_a_bytes;
  // side column tiles (n_main_ct > 0): column tiles ct >= n_main_ct belong to a SECOND, 1x1 conv on the same input pixels
  // -- the stride-2 downsample of a ResNet block riding in the launch of the block's 3x3 stride-2 conv1, whose centre
  // tap reads exactly the pixels the 1x1 conv reads. Their weight rows (>= n_main_ct * BLOCK_N of the same matrix) are
  // zero outside the centre tap, so the k-loop of a side tile is the centre tap's k-blocks only; they are stored to
  // out2 (columns counted from 0) with their own ReLU switch; bias rows continue the main conv's.
  int n_main_ct, ld_out2, relu2;
  void* out2;
  // EPI 3 on a conv: tmap_out is the rank-5 NHWC output map (channels, W, 1, H, B) and a slab is stored as one
  // (64, box_w, 1, box_h, box_b) box at the tile's pixel origin instead of a 128-row box of a plain matrix
  int out_5d;
  // epilogue
  const float* bias;         // [N] or nullptr
  const void* residual;      // [rows, ld_res] or nullptr
  void* out;                 // [rows, ld_out]
  int ld_out, ld_res;
  int out_f32, res_f32, relu;
  // optional output-row remap: rows (img, h, w) -> token rows of DETRVAE's flattened multi-camera feature map
  //   out_row = (img / remap_ncam) * remap_L + remap_off + h * (W * remap_ncam) + (img % remap_ncam) * W + w
  int remap_ncam, remap_L, remap_off;
  // fused GroupNorm(8) + Mish (+ FiLM) (+ residual) epilogue of the ConditionalUnet1D convs (EPI == 1, gemm_gn.cuh):
  // rows of a tile are whole samples (box_w = T rows each), `bias` is the conv bias, `out` is act_t
  const float* gn_gamma;     // [N]
  const float* gn_beta;      // [N]
  const float* gn_film_gc;   // [B, gn_film_ld] or nullptr: scale at gn_film_off + c, bias at gn_film_off + N + c
  const float* gn_film_t;    // the diffusion step's row of the FiLM table (same indexing); set per launch
  int gn_film_ld, gn_film_off;
  const void* gn_res;        // act_t [rows, gn_ld_res] or nullptr
  int gn_ld_res;
  int gn_cg;                 // channels per group = N / 8
  // per-episode FiLM of the MT-ACT backbone (EPI == 4): out = relu((1 + gamma) * (acc + bias) + beta + residual), with
  // gamma = film[ep * film_ld + film_g_off + c], beta = film[ep * film_ld + film_b_off + c], ep = image / film_div
  const float* film;
  int film_ld, film_g_off, film_b_off, film_div;
  // fused residual + LayerNorm(512) epilogue (EPI == 5, gemm_ln.cuh): v = acc + bias + residual (fp32); y = LN(v);
  // `out` gets y in fp32 (the residual stream), ln_y_act y in 16 bits, ln_ypos_act (y + ln_pos[row % ln_L]) in 16 bits
  const float* ln_gamma;
  const float* ln_beta;
  const float* ln_pos;       // [ln_L, 512] or nullptr
  int ln_L;
  void* ln_y_act;            // act_t [rows, 512] or nullptr
  void* ln_ypos_act;         // act_t [rows, 512] or nullptr (needs ln_pos)
  float ln_eps;
};

// PAIR: two CTAs of a cluster work on one 256-row tile (tcgen05 cta_group::2); each holds its 128 rows of A and half
// of the BLOCK_N rows of B, so a k-block costs a CTA 2/3 of the L2 -> SM bytes of the single-CTA tile.
template <int BLOCK_N, int BLOCK_K, int STAGES, bool PAIR = false, int EPI = 0>
struct GemmSmem {
  static constexpr int kABytes = 128 * BLOCK_K * 2;
  static constexpr int kBBytes = (PAIR ? BLOCK_N / 2 : BLOCK_N) * BLOCK_K * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarrierBytes = 256;
  // epilogue staging: one 32 x 128 B tile per warp, or (EPI 3) one 128 x 128 B slab per 4-warp group
  static constexpr int kStagingBytes = 8 * 4096;
  static constexpr int kTotal = STAGES * kStageBytes + kStagingBytes + kBarrierBytes + 1024;  // +1024: alignment slack
};

// Persistent tile loop of one CTA. Single CTA: tiles blockIdx.x, +gridDim.x, ... PAIR: cluster c walks pair tiles
// c, +clusters, ...; pair tile (prt, ct) is row tiles 2*prt (leader) and 2*prt + 1 (peer) of column tile ct. tile(i)
// is the (row tile * n_col_tiles + column tile) id the roles decode as before; a row tile past the end is all padding.
template <bool PAIR>
struct TileSched {
  int first, step, count, rank, n_col_tiles;
  uint32_t acc_empty_addr;   // PAIR: shared::cluster address of the leader's tmem_empty_bar[0]
  __device__ __forceinline__ int tile(int i) const {
    if (!PAIR) return i;
    return (2 * (i / n_col_tiles) + rank) * n_col_tiles + (i % n_col_tiles);
  }
  // accumulator `acc` has been read out by this warp
  __device__ __forceinline__ void release_acc(uint64_t* tmem_empty_bar, int acc) const {
    if (PAIR) mbar_arrive_cluster(acc_empty_addr + acc * 8);
    else mbar_arrive(&tmem_empty_bar[acc]);
  }
};

// ---------------------------------------------------------------------------------------------------------------------
// Epilogue shared by the GEMM / conv kernels (warps 0..7 of the CTA).
template <int BLOCK_N, bool OUT_F32, int RES, bool PAIR, bool FILM = false>
__device__ __forceinline__ void gemm_epilogue(const GemmParams& p, uint8_t* smem_stage, uint32_t tmem_base,
                                              uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                              const TileSched<PAIR>& ts, int warp, int lane) {
  const int rows_per_tile = p.box_w * p.box_h * p.box_b;
  // ------------------------------------------------------------ epilogue warps 0..7
  // Warp e reads TMEM lanes 32*(e%4)..+31 (hardware rule: lane quarter = warp id % 4) and the 32-column chunks
  // ci with ci % 2 == e/4. Two passes per chunk so that global traffic is coalesced:
  //   pass 1: thread = accumulator row; tcgen05.ld 32 fp32 columns -> per-warp smem staging tile (32 rows x 128 B,
  //           16-byte groups XOR-swizzled by row & 7: conflict-free for both passes)
  //   pass 2: 8 lanes = one row (8 x 16 B = the row's 128 B), 4 rows per instruction; bias / residual / ReLU here,
  //           so residual loads and output stores are full 64-128 B row segments.
  // The next chunk's tcgen05.ld is in flight while pass 2 runs. Output / residual modes are template parameters and
  // row pointers are built once per tile: a single warp per scheduler is instruction-latency bound, so the
  // per-chunk instruction count is what sets the epilogue rate.
  constexpr int kChunksPerWarp = BLOCK_N / 64;
  constexpr int kOutEsz = OUT_F32 ? 4 : 2;
  constexpr int kResEsz = (RES == 2) ? 4 : 2;
  int acc = 0; uint32_t acc_phase = 0;
  const int q = warp & 3, half = warp >> 2;
  const int r = q * 32 + lane;                    // row of the tile owned by this thread in pass 1
  const int wl = r % p.box_w;
  const int hl = (r / p.box_w) % p.box_h;
  const int bl = r / (p.box_w * p.box_h);
  const int sub = lane & 7, rsel = lane >> 3;
  const uint32_t stg = smem_u32(smem_stage + warp * 4096);
  uint32_t st_addr[8];
  #pragma unroll
  for (int g = 0; g < 8; ++g) st_addr[g] = stg + lane * 128 + ((g ^ (lane & 7)) << 4);
  const uint32_t ld_even = stg + rsel * 128 + ((sub ^ rsel) << 4);
  const uint32_t ld_odd = stg + (rsel + 4) * 128 + ((sub ^ (rsel + 4)) << 4);
  const float relu_floor = p.relu ? 0.0f : -INFINITY;
  const int col_in_tile = half * 32 + sub * 4;    // first of this lane's 4 columns inside chunk 0 of this warp
  for (int ti = ts.first; ti < ts.count; ti += ts.step) {
    const int tile = ts.tile(ti);
    const int ct = tile % p.n_col_tiles;
    const int rt = tile / p.n_col_tiles;
    const int wt = rt % p.tiles_w;
    const int ht = (rt / p.tiles_w) % p.tiles_h;
    const int bt = rt / (p.tiles_w * p.tiles_h);
    const int w = wt * p.box_w + wl, h = ht * p.box_h + hl, b = bt * p.box_b + bl;
    const bool valid = (r < rows_per_tile) && (w < p.W) && (h < p.H) && (b < p.B);
    int my_grow = -1, my_orow = -1;               // source row (residual, natural pixel order) / destination row
    int my_film = 0;                              // FiLM row offset of this row's episode
    if (valid) {
      if (FILM) my_film = (b / p.film_div) * p.film_ld;
      my_grow = (b * p.H + h) * p.W + w;
      my_orow = my_grow;
      if (p.remap_ncam > 0)
        my_orow = (b / p.remap_ncam) * p.remap_L + p.remap_off + h * (p.W * p.remap_ncam) + (b % p.remap_ncam) * p.W + w;
    }
    const int col0 = ct * BLOCK_N + col_in_tile;
    const bool side = p.n_main_ct > 0 && ct >= p.n_main_ct;          // (RES == 0 launches only: plan_conv)
    char* const out_base = static_cast<char*>(side ? p.out2 : p.out);
    const int out_ld = side ? p.ld_out2 : p.ld_out;
    const int out_col0 = side ? col0 - p.n_main_ct * BLOCK_N : col0;
    const float relu_floor_t = side ? (p.relu2 ? 0.0f : -INFINITY) : relu_floor;
    if (RES != 0 && my_grow >= 0) {
      // this row's residual segments of the tile go to L2 now, while the accumulator is still being built: the loads
      // in pass 2 then see L2 latency instead of HBM latency (no registers to hold them across the chunk loop)
      const char* rrow = static_cast<const char*>(p.residual) +
                         (static_cast<long long>(my_grow) * p.ld_res + ct * BLOCK_N + half * 32) * kResEsz;
#pragma unroll
      for (int j = 0; j < kChunksPerWarp; ++j)
        if (ct * BLOCK_N + half * 32 + j * 64 < p.n_valid) prefetch_l2(rrow + j * 64 * kResEsz);
    }
    char* optr[8];
    const char* rptr[8];
    const float* fptr[8];
  #pragma unroll
    for (int it = 0; it < 8; ++it) {
      fptr[it] = FILM ? p.film + __shfl_sync(0xffffffffu, my_film, it * 4 + rsel) + col0 : nullptr;
      const int orow = __shfl_sync(0xffffffffu, my_orow, it * 4 + rsel);
      optr[it] = (orow >= 0) ? out_base + (static_cast<long long>(orow) * out_ld + out_col0) * kOutEsz : nullptr;
      if (RES != 0) {
        const int grow = __shfl_sync(0xffffffffu, my_grow, it * 4 + rsel);
        rptr[it] = static_cast<const char*>(p.residual) + (static_cast<long long>(grow < 0 ? 0 : grow) * p.ld_res + col0) * kResEsz;
      } else {
        rptr[it] = nullptr;
      }
    }
    const float* bias_ptr = (p.bias != nullptr) ? p.bias + col0 : nullptr;
    // 16-bit residual rows are loaded into registers one chunk ahead: chunk 0's before the wait for the accumulator
    // (the tile is still on the tensor pipe), chunk j + 1's at the top of chunk j -- behind the accumulator they cost a
    // full L2 round trip per chunk, which set the pace of the N = 128 convs (two chunks per warp and tile)
    uint2 rvn[8];
    if (RES == 1) {
  #pragma unroll
      for (int it = 0; it < 8; ++it)
        rvn[it] = (optr[it] != nullptr && col0 < p.n_valid) ? __ldg(reinterpret_cast<const uint2*>(rptr[it])) : make_uint2(0u, 0u);
    }
    mbar_wait(&tmem_full_bar[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_row = tmem_base + acc * BLOCK_N + half * 32 + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t v[32];
    tmem_ld_32x32b_x32(t_row, v);
  #pragma unroll
    for (int j = 0; j < kChunksPerWarp; ++j) {      // this warp's chunk j = columns half*32 + j*64 .. +31 of the tile
      uint2 rv1[8];
      if (RES == 1) {
  #pragma unroll
        for (int it = 0; it < 8; ++it) rv1[it] = rvn[it];
        if (j + 1 < kChunksPerWarp) {
          const bool nok = col0 + (j + 1) * 64 < p.n_valid;
  #pragma unroll
          for (int it = 0; it < 8; ++it)
            rvn[it] = (optr[it] != nullptr && nok) ? __ldg(reinterpret_cast<const uint2*>(rptr[it] + (j + 1) * 64 * kResEsz))
                                                    : make_uint2(0u, 0u);
        }
      }
      tmem_ld_wait();
  #pragma unroll
      for (int g = 0; g < 8; ++g) st_shared_v4(st_addr[g], v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
      if (j + 1 < kChunksPerWarp) tmem_ld_32x32b_x32(t_row + (j + 1) * 64, v);
      else tc_fence_before();
      __syncwarp();
      if (j + 1 == kChunksPerWarp && lane == 0) ts.release_acc(tmem_empty_bar, acc);   // 8 arrivals (per CTA) free the accumulator
      const bool col_ok = col0 + j * 64 < p.n_valid;   // warp-uniform: N is a multiple of 64, a chunk is 32 columns
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (bias_ptr != nullptr && col_ok) bv = __ldg(reinterpret_cast<const float4*>(bias_ptr + j * 64));
      float4 f[8];
  #pragma unroll
      for (int it = 0; it < 8; ++it) f[it] = ld_shared_v4f(((it & 1) ? ld_odd : ld_even) + (it >> 1) * 1024);
      if (FILM) {
        // bn2 output (acc + folded bias) modulated per episode BEFORE the identity branch is added; the bias is
        // consumed here, so the common tail below adds zero
  #pragma unroll
        for (int it = 0; it < 8; ++it) {
          if (optr[it] == nullptr) continue;
          const float4 g4 = __ldg(reinterpret_cast<const float4*>(fptr[it] + p.film_g_off + j * 64));
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(fptr[it] + p.film_b_off + j * 64));
          f[it].x = (1.0f + g4.x) * (f[it].x + bv.x) + b4.x; f[it].y = (1.0f + g4.y) * (f[it].y + bv.y) + b4.y;
          f[it].z = (1.0f + g4.z) * (f[it].z + bv.z) + b4.z; f[it].w = (1.0f + g4.w) * (f[it].w + bv.w) + b4.w;
        }
        bv = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      if (RES == 2) {
        float4 rv[8];
  #pragma unroll
        for (int it = 0; it < 8; ++it)
          rv[it] = (optr[it] != nullptr && col_ok) ? __ldg(reinterpret_cast<const float4*>(rptr[it] + j * 64 * kResEsz))
                                         : make_float4(0.f, 0.f, 0.f, 0.f);
  #pragma unroll
        for (int it = 0; it < 8; ++it) { f[it].x += rv[it].x; f[it].y += rv[it].y; f[it].z += rv[it].z; f[it].w += rv[it].w; }
      } else if (RES == 1) {
  #pragma unroll
        for (int it = 0; it < 8; ++it) {
          const float2 lo = unpack_act2(rv1[it].x), hi = unpack_act2(rv1[it].y);
          f[it].x += lo.x; f[it].y += lo.y; f[it].z += hi.x; f[it].w += hi.y;
        }
      }
  #pragma unroll
      for (int it = 0; it < 8; ++it) {
        const float ox = fmaxf(f[it].x + bv.x, relu_floor_t), oy = fmaxf(f[it].y + bv.y, relu_floor_t);
        const float oz = fmaxf(f[it].z + bv.z, relu_floor_t), ow = fmaxf(f[it].w + bv.w, relu_floor_t);
        if (optr[it] != nullptr && col_ok) {
          if (OUT_F32) *reinterpret_cast<float4*>(optr[it] + j * 64 * kOutEsz) = make_float4(ox, oy, oz, ow);
          else *reinterpret_cast<uint2*>(optr[it] + j * 64 * kOutEsz) = make_uint2(pack_act2(ox, oy), pack_act2(oz, ow));
        }
      }
      __syncwarp();                                // staging tile is rewritten by the next chunk's pass 1
    }
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
  }
}

// Epilogue for plain matrices with a 16-bit output and no residual (EPI == 3: QKV / FFN-up projections, the short-K
// GEMMs whose rate the epilogue sets: with the staged-transpose epilogue above they run at 1.0 PFLOP/s, with its global
// stores removed at 1.35). The eight warps work as two groups of four; a group owns 64-column slabs of the tile:
// thread = one accumulator row, 2 x tcgen05.ld of 32 columns, bias / ReLU / pack, eight swizzled 16-byte st.shared into
// the group's [128 rows x 128 B] staging tile, one TMA store per slab (rows past M and columns past N are clipped by
// the tensor map). No ld.shared, no per-lane global stores, two named barriers per slab.
template <int BLOCK_N, bool PAIR>
__device__ __forceinline__ void gemm_epilogue_tma(const CUtensorMap* tmap_out, const GemmParams& p, uint8_t* smem_stage,
                                                  uint32_t tmem_base, uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                                  const TileSched<PAIR>& ts, int warp, int lane) {
  constexpr int kSlabs = BLOCK_N / 64;                 // 64-column slabs per tile, dealt alternately to the two groups
  const int q = warp & 3, grp = warp >> 2;
  const int r = q * 32 + lane;
  const bool issuer = (q == 0 && lane == 0);
  const uint32_t stage = smem_u32(smem_stage + grp * 16384);    // (a second slab per group measured no faster)
  uint32_t off[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) off[c] = r * 128 + ((c ^ (r & 7)) << 4);
  const float relu_floor = p.relu ? 0.0f : -INFINITY;
  int acc = 0; uint32_t acc_phase = 0;
  for (int ti = ts.first; ti < ts.count; ti += ts.step) {
    const int tile = ts.tile(ti);
    const int ct = tile % p.n_col_tiles, rt = tile / p.n_col_tiles;
    const int row0 = rt * 128;
    mbar_wait(&tmem_full_bar[acc], acc_phase);
    tc_fence_after();
    const uint32_t t_row = tmem_base + acc * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll
    for (int sl = grp; sl < kSlabs; sl += 2) {
      const int col0 = ct * BLOCK_N + sl * 64;
      uint32_t v0[32], v1[32];
      tmem_ld_32x32b_x32(t_row + sl * 64, v0);
      tmem_ld_32x32b_x32(t_row + sl * 64 + 32, v1);
      tmem_ld_wait();
      if (sl + 2 >= kSlabs) {                          // this group's last slab of the tile is in registers
        tc_fence_before();
        __syncwarp();
        if (lane == 0) ts.release_acc(tmem_empty_bar, acc);
      }
      if (issuer) bulk_wait_group_read0();             // the previous store has read the staging slab
      named_bar_sync(1 + grp, 128);
      const bool col_ok = col0 < p.n_valid;            // group-uniform (ragged last column tile)
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint32_t* v = (c < 4) ? &v0[8 * c] : &v1[8 * (c - 4)];
        float f[8];
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          const float4 b4 = (p.bias != nullptr && col_ok) ? __ldg(reinterpret_cast<const float4*>(p.bias + col0 + 8 * c) + h2)
                                                          : make_float4(0.f, 0.f, 0.f, 0.f);
          f[4 * h2 + 0] = fmaxf(__uint_as_float(v[4 * h2 + 0]) + b4.x, relu_floor);
          f[4 * h2 + 1] = fmaxf(__uint_as_float(v[4 * h2 + 1]) + b4.y, relu_floor);
          f[4 * h2 + 2] = fmaxf(__uint_as_float(v[4 * h2 + 2]) + b4.z, relu_floor);
          f[4 * h2 + 3] = fmaxf(__uint_as_float(v[4 * h2 + 3]) + b4.w, relu_floor);
        }
        st_shared_v4(stage + off[c], pack_act2(f[0], f[1]), pack_act2(f[2], f[3]), pack_act2(f[4], f[5]), pack_act2(f[6], f[7]));
      }
      fence_proxy_async_smem();                        // generic-proxy writes -> visible to the TMA store
      named_bar_sync(1 + grp, 128);
      if (p.out_5d) {
        const int wt = rt % p.tiles_w, ht = (rt / p.tiles_w) % p.tiles_h, bt = rt / (p.tiles_w * p.tiles_h);
        if (issuer && col_ok && bt < p.tiles_b) {      // (a pair's odd last row tile is all padding)
          tma_store_5d(tmap_out, stage, col0, wt * p.box_w, 0, ht * p.box_h, bt * p.box_b);
          bulk_commit_group();
        }
      } else if (issuer && col_ok && row0 < p.W) {
        tma_store_2d(tmap_out, stage, col0, row0);
        bulk_commit_group();
      }
    }
    if (grp >= kSlabs) {                               // BLOCK_N = 64: the second group has no slab, only its arrivals
      tc_fence_before();
      __syncwarp();
      if (lane == 0) ts.release_acc(tmem_empty_bar, acc);
    }
    if (++acc == 2) { acc = 0; acc_phase ^= 1; }
  }
  if (issuer) bulk_wait_group0();
}

template <int BLOCK_N, bool FILM, bool PAIR>
__device__ __forceinline__ void gemm_gn_epilogue(const GemmParams& p, uint8_t* smem_stage, uint32_t tmem_base,
                                                 uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                                 const TileSched<PAIR>& ts, int warp, int lane);
template <bool PAIR>
__device__ __forceinline__ void gemm_ln_epilogue(const GemmParams& p, uint8_t* smem_stage, uint32_t tmem_base,
                                                 uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                                 const TileSched<PAIR>& ts, int warp, int lane);

// OUT_F32: fp32 output (else act_t). RES: 0 no residual, 1 act_t residual, 2 fp32 residual.
// EPI: 0 plain epilogue (bias / residual / ReLU); fused GroupNorm + Mish (gemm_gn.cuh) followed by 1: FiLM, 2: residual;
// 3: plain 16-bit output through a TMA store (tmap_out; the other modes ignore it); 4: plain + per-episode FiLM (MT-ACT);
// 5: BLOCK_N = 512 = the whole row, fp32 residual + LayerNorm fused (gemm_ln.cuh). A 128 x 512 fp32 tile is all of TMEM,
//    so this mode has ONE accumulator and issues every k-step as two N = 256 MMAs; PAIR only (each CTA then holds two
//    128-row boxes of the weight tile, rows 128 r.. and 256 + 128 r..).
// PAIR: launched as clusters of two CTAs (gemm_tc_pair_kernel); see GemmSmem / TileSched.
template <int BLOCK_N, int BLOCK_K, int STAGES, bool OUT_F32, int RES, int EPI, bool PAIR>
__device__ __forceinline__ void gemm_tc_body(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const CUtensorMap& tmap_out,
                                             const GemmParams& p) {
  static_assert(BLOCK_K == 64 || BLOCK_K == 32, "BLOCK_K selects the 128B / 64B swizzle");
  static_assert(BLOCK_N == 64 || BLOCK_N == 128 || BLOCK_N == 256 || BLOCK_N == 512, "BLOCK_N");
  constexpr bool kWide = (BLOCK_N == 512);
  static_assert(!kWide || (PAIR && EPI == 5), "512-wide tiles: CTA pairs with the LayerNorm epilogue only");
  static_assert(kWide || EPI != 5, "the LayerNorm epilogue needs the whole 512-wide row in the tile");
  constexpr int kMmaN = kWide ? 256 : BLOCK_N;                      // N of one tcgen05.mma
  constexpr int kAccs = kWide ? 1 : 2;                              // accumulators in TMEM
  pdl_trigger();
  using S = GemmSmem<BLOCK_N, BLOCK_K, STAGES, PAIR, EPI>;
  constexpr uint32_t kSwizzleLayout = (BLOCK_K == 64) ? 2u : 4u;   // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint32_t kSBO = 8u * BLOCK_K * 2u;                      // 8-row core-matrix group stride
  constexpr uint32_t kTmemCols = kAccs * BLOCK_N;                   // double-buffered accumulator (one when 512 wide)
  constexpr uint32_t kIdesc = make_idesc_f16(PAIR ? 256 : 128, kMmaN, false, false);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * S::kABytes;
  uint8_t* smem_stage = smem + STAGES * S::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + S::kStagingBytes);
  uint64_t* full_bar = bars;                    // [STAGES]   PAIR: the leader's are used (both CTAs' TMA bytes land on it)
  uint64_t* empty_bar = bars + STAGES;          // [STAGES]   PAIR: the leader's commit arrives on both CTAs' copies
  uint64_t* tmem_full_bar = bars + 2 * STAGES;  // [2]        PAIR: likewise
  uint64_t* tmem_empty_bar = tmem_full_bar + 2; // [2]        PAIR: the leader's collects both CTAs' epilogue warps
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = PAIR ? static_cast<int>(cluster_ctarank()) : 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], PAIR ? 16 : 8); }
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) { prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_b); }
  if (warp == 9) {
    if (PAIR) { tmem_alloc_pair(tmem_ptr_smem, kTmemCols); tmem_relinquish_pair(); }
    else { tmem_alloc(tmem_ptr_smem, kTmemCols); tmem_relinquish(); }
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();   // PAIR: the peer's barriers are live before anything lands on them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();                                           // everything above ran under the previous kernel's tail

  const int rows_per_tile = p.box_w * p.box_h * p.box_b;
  const int num_kb = p.num_taps * p.chunks_per_tap;
  const int row_tiles = p.tiles_w * p.tiles_h * p.tiles_b;
  TileSched<PAIR> ts;
  ts.rank = rank; ts.n_col_tiles = p.n_col_tiles;
  if (PAIR) {
    ts.first = blockIdx.x >> 1; ts.step = gridDim.x >> 1; ts.count = ((row_tiles + 1) >> 1) * p.n_col_tiles;
    ts.acc_empty_addr = mapa_u32(smem_u32(tmem_empty_bar), 0);
  } else {
    ts.first = blockIdx.x; ts.step = gridDim.x; ts.count = row_tiles * p.n_col_tiles;
    ts.acc_empty_addr = 0;
  }
  const uint32_t cta_tx_bytes = static_cast<uint32_t>(rows_per_tile * BLOCK_K * 2 + S::kBBytes);

  // The producer and MMA warps run their loops warp-uniformly (all 32 lanes, identical control flow) and only the
  // issuing instruction is predicated on one elected lane: loop state stays in uniform registers and the per-k-block
  // instruction count -- the throughput limit of a single issuing thread -- is a few dozen instead of ~150.
  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer
    const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
    const uint32_t full_base = PAIR ? mapa_u32(smem_u32(full_bar), 0) : smem_u32(full_bar);
    const int n_half = PAIR ? rank * (BLOCK_N / 2) : 0;
    int stage = 0; uint32_t phase = 0;
    for (int ti = ts.first; ti < ts.count; ti += ts.step) {
      const int tile = ts.tile(ti);
      const int ct = tile % p.n_col_tiles;
      const int rt = tile / p.n_col_tiles;
      const int wt = rt % p.tiles_w;
      const int ht = (rt / p.tiles_w) % p.tiles_h;
      const int bt = rt / (p.tiles_w * p.tiles_h);
      const int w0 = wt * p.box_w, h0 = ht * p.box_h, b0 = bt * p.box_b;
      const int n0 = ct * BLOCK_N + n_half;
      if (p.halo3) {
        const uint32_t a_stride = static_cast<uint32_t>(p.halo_a_bytes), b_stride = 3u * S::kBBytes;
        const uint32_t bh_base = a_base + p.halo_stages * a_stride;
        const uint32_t tx = a_stride + b_stride;
        for (int kw = 0; kw < 3; ++kw) {
          for (int ch = 0; ch < p.chunks_per_tap; ++ch) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            if (elect_one()) {
              const uint32_t fb = full_base + stage * 8;
              if (PAIR) {
                if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * tx);
                tma_load_5d_pair(a_base + stage * a_stride, &tmap_a, fb, ch * BLOCK_K, w0 + kw - 1, 0, h0 - 1, b0);
#pragma unroll
                for (int kh = 0; kh < 3; ++kh)
                  tma_load_2d_pair(bh_base + stage * b_stride + kh * S::kBBytes, &tmap_b, fb,
                                   ((kh * 3 + kw) * p.chunks_per_tap + ch) * BLOCK_K, n0);
              } else {
                mbar_arrive_expect_tx_u32(fb, tx);
                tma_load_5d_u32(a_base + stage * a_stride, &tmap_a, fb, ch * BLOCK_K, w0 + kw - 1, 0, h0 - 1, b0);
#pragma unroll
                for (int kh = 0; kh < 3; ++kh)
                  tma_load_2d_u32(bh_base + stage * b_stride + kh * S::kBBytes, &tmap_b, fb,
                                  ((kh * 3 + kw) * p.chunks_per_tap + ch) * BLOCK_K, n0);
              }
            }
            __syncwarp();
            if (++stage == p.halo_stages) { stage = 0; phase ^= 1; }
          }
        }
        continue;
      }
      // k-block order. 9-tap (3x3) convs walk (kw, channel chunk, kh) -- the order of the halo3 stages above -- so that
      // an output element accumulates its products in ONE order whichever path a launch takes (results stay
      // bit-identical across batch sizes); everything else walks (tap, chunk) = ascending weight column.
      const bool kw_major = (p.num_taps == 9);
      const bool side_tile = p.n_main_ct > 0 && ct >= p.n_main_ct;
      const int outer_n = kw_major ? 3 : p.num_taps, inner_n = kw_major ? 3 : 1;
      for (int outer = 0; outer < outer_n; ++outer) {
        for (int ch = 0; ch < p.chunks_per_tap; ++ch) {
         for (int inner = 0; inner < inner_n; ++inner) {
          const int tap = kw_major ? inner * 3 + outer : outer;
          if (side_tile && tap != 4) continue;
          const int c0 = p.tap_c0[tap], d1 = w0 + p.tap_d1[tap], d2 = p.tap_d2[tap], d3 = h0 + p.tap_d3[tap];
          const int kcol = (tap * p.chunks_per_tap + ch) * BLOCK_K;   // column of the weight matrix
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            const uint32_t fb = full_base + stage * 8;
            if (PAIR) {
              if (rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * cta_tx_bytes);   // the peer's boxes land here too
              tma_load_5d_pair(a_base + stage * S::kABytes, &tmap_a, fb, c0 + ch * BLOCK_K, d1, d2, d3, b0);
              if (kWide) {   // this CTA's share of both N = 256 halves: weight rows 128 r.. and 256 + 128 r..
                tma_load_2d_pair(b_base + stage * S::kBBytes, &tmap_b, fb, kcol, ct * BLOCK_N + rank * 128);
                tma_load_2d_pair(b_base + stage * S::kBBytes + S::kBBytes / 2, &tmap_b, fb, kcol, ct * BLOCK_N + 256 + rank * 128);
              } else {
                tma_load_2d_pair(b_base + stage * S::kBBytes, &tmap_b, fb, kcol, n0);
              }
            } else {
              mbar_arrive_expect_tx_u32(fb, cta_tx_bytes);
              tma_load_5d_u32(a_base + stage * S::kABytes, &tmap_a, fb, c0 + ch * BLOCK_K, d1, d2, d3, b0);
              tma_load_2d_u32(b_base + stage * S::kBBytes, &tmap_b, fb, kcol, n0);
            }
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
         }
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer (one elected lane issues; PAIR: leader only)
    if (rank == 0) {
      const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), 0, kSBO, kSwizzleLayout);
      const uint64_t b_desc0 = make_smem_desc(smem_u32(smem_b), 0, kSBO, kSwizzleLayout);
      const uint32_t empty_base = smem_u32(empty_bar);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int ti = ts.first; ti < ts.count; ti += ts.step) {
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        const int tile_kb = (p.n_main_ct > 0 && ts.tile(ti) % p.n_col_tiles >= p.n_main_ct) ? p.chunks_per_tap : num_kb;
        if (p.halo3) {
          const uint32_t a_stride16 = static_cast<uint32_t>(p.halo_a_bytes) >> 4, b_stride16 = (3u * S::kBBytes) >> 4;
          const uint64_t bh_desc0 = a_desc0 + static_cast<uint64_t>(p.halo_stages * a_stride16);
          const uint32_t row_shift16 = static_cast<uint32_t>(p.box_w * BLOCK_K * 2) >> 4;   // one input row of the box
          const int n_it = 3 * p.chunks_per_tap;
          for (int it = 0; it < n_it; ++it) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            if (elect_one()) {
              const uint64_t da = a_desc0 + static_cast<uint64_t>(stage * a_stride16);
              const uint64_t db = bh_desc0 + static_cast<uint64_t>(stage * b_stride16);
#pragma unroll
              for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / 16; ++k) {
                  const uint64_t dak = da + static_cast<uint64_t>(kh * row_shift16 + 2 * k);
                  const uint64_t dbk = db + static_cast<uint64_t>(kh * (S::kBBytes >> 4) + 2 * k);
                  if (PAIR) umma_f16_ss_pair(d_tmem, dak, dbk, kIdesc, (it | kh | k) != 0 ? 1u : 0u);
                  else umma_f16_ss(d_tmem, dak, dbk, kIdesc, (it | kh | k) != 0 ? 1u : 0u);
                }
              }
              if (PAIR) umma_commit_pair_u32(empty_base + stage * 8, 3u);
              else umma_commit_u32(empty_base + stage * 8);
            }
            __syncwarp();
            if (++stage == p.halo_stages) { stage = 0; phase ^= 1; }
          }
        } else
        for (int kb = 0; kb < tile_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            // the 14-bit start-address field (bytes >> 4) never carries: all of shared memory is < 256 KB
            const uint64_t da = a_desc0 + static_cast<uint64_t>(stage * (S::kABytes >> 4));
            const uint64_t db = b_desc0 + static_cast<uint64_t>(stage * (S::kBBytes >> 4));
#pragma unroll
            for (int k = 0; k < BLOCK_K / 16; ++k) {
              if (kWide) {
                umma_f16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, kIdesc, (kb | k) != 0 ? 1u : 0u);
                umma_f16_ss_pair(d_tmem + 256, da + 2 * k, db + 2 * k + (S::kBBytes >> 5), kIdesc, (kb | k) != 0 ? 1u : 0u);
              } else if (PAIR) {
                umma_f16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, kIdesc, (kb | k) != 0 ? 1u : 0u);
              } else {
                umma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, kIdesc, (kb | k) != 0 ? 1u : 0u);
              }
            }
            // frees this smem stage (in both CTAs) when the MMAs above retire
            if (PAIR) umma_commit_pair_u32(empty_base + stage * 8, 3u);
            else umma_commit_u32(empty_base + stage * 8);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) {                                   // accumulator complete -> epilogue (of both CTAs)
          if (PAIR) umma_commit_pair_u32(smem_u32(&tmem_full_bar[acc]), 3u);
          else umma_commit(&tmem_full_bar[acc]);
        }
        __syncwarp();
        if (++acc == kAccs) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    if constexpr (EPI == 5)
      gemm_ln_epilogue<PAIR>(p, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, ts, warp, lane);
    else if constexpr (EPI == 3)
      gemm_epilogue_tma<BLOCK_N, PAIR>(&tmap_out, p, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, ts, warp, lane);
    else if constexpr (EPI == 4)
      gemm_epilogue<BLOCK_N, OUT_F32, RES, PAIR, true>(p, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, ts, warp, lane);
    else if constexpr (EPI != 0)
      gemm_gn_epilogue<BLOCK_N, EPI == 1, PAIR>(p, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, ts, warp, lane);
    else
      gemm_epilogue<BLOCK_N, OUT_F32, RES, PAIR>(p, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, ts, warp, lane);
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all(); else __syncthreads();   // PAIR: nothing of the peer is touched after this point
  if (warp == 9) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
  }
}

template <int BLOCK_N, int BLOCK_K, int STAGES, bool OUT_F32, int RES, int EPI = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ GemmParams p) {
  gemm_tc_body<BLOCK_N, BLOCK_K, STAGES, OUT_F32, RES, EPI, false>(tmap_a, tmap_b, tmap_out, p);
}

// CTA-pair version: grid = 2 x clusters; tmap_b's box holds BLOCK_N / 2 weight rows.
template <int BLOCK_N, int BLOCK_K, int STAGES, bool OUT_F32, int RES, int EPI = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm_tc_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ GemmParams p) {
  gemm_tc_body<BLOCK_N, BLOCK_K, STAGES, OUT_F32, RES, EPI, true>(tmap_a, tmap_b, tmap_out, p);
}

}  // namespace rmb
