// Halo-reuse implicit-GEMM convolution for the two bandwidth-critical stages of the ResNet-18 trunk
// (reference: torchvision resnet18 conv1 / layer1 via third_party/act/detr/models/backbone.py:86-96), Cout = 64.
//
// gemm_tc_kernel fetches one TMA box per filter tap, i.e. it pulls every input pixel kh*kw times from L2 and pays one
// producer/MMA barrier round trip per tap; with only N = 64 output channels per tile that is >190 B of L2 traffic per
// tensor-core cycle. Here a tile's input is fetched ONCE per filter COLUMN (3x3 conv) or per input-row PARITY (7x7/2
// stem) as a box that carries the vertical halo, laid out pixel-major [h][w][K] with w fastest. A filter tap that moves
// by one input row is then just the same shared-memory buffer advanced by box_w rows of K elements = a whole number of
// 8-row swizzle atoms, so all taps of a buffer are tcgen05.mma instructions on shifted descriptors -- no extra loads.
// The 64 x K weight slices of all taps stay resident in shared memory for the lifetime of the (persistent) CTA.
//   3x3/1 (Cin 64): 3 boxes of (box_h + 2) x box_w pixels per tile instead of 9 of box_h x box_w;  9 x 4 MMAs
//   7x7/2 stem     : 2 boxes (even / odd input rows) of <= 19 x box_w windows instead of 7;          7 x 2 MMAs
#pragma once
#include "gemm_tc.cuh"

namespace rmb {

constexpr int kHaloMaxLoads = 3;
constexpr int kHaloMaxStages = 8;

struct HaloParams {
  GemmParams g;                       // tiling (box_w x box_h x 1 pixels, n_col_tiles = 1) and epilogue; g.tap_* unused
  int n_loads;                        // TMA boxes (= pipeline stages consumed) per tile
  int ld_c0[kHaloMaxLoads];           // coordinate offsets of each box, as in GemmParams' tap tables
  int ld_d1[kHaloMaxLoads];
  int ld_d2[kHaloMaxLoads];
  int ld_d3[kHaloMaxLoads];
  int ld_tap0[kHaloMaxLoads];         // taps [ld_tap0, ld_tap0 + ld_ntaps) read this box
  int ld_ntaps[kHaloMaxLoads];
  uint32_t box_bytes;                 // bytes one box deposits (expect_tx)
  uint32_t stage_bytes;               // box_bytes rounded up to 1024
  int n_stages;
  int out_bufs;                       // staging tiles per epilogue group (2: the TMA store of tile i overlaps the packing of tile i+1)
  int n_taps;
  uint32_t tap_aoff[kMaxTaps];        // byte offset of the tap's first A row inside its box (multiple of 8 rows)
  int tap_w[kMaxTaps];                // index of the tap's 64 x BLOCK_K weight slice in the resident weights
};

template <int BLOCK_K>
struct HaloSmem {
  static constexpr int kWTapBytes = 64 * BLOCK_K * 2;
  static constexpr int kStagingTileBytes = 16384;   // epilogue: one output tile (128 px x 64 ch act_t); 2 groups x out_bufs of them
  static constexpr int kBarrierBytes = 256;
  static int total(int n_taps, int n_stages, uint32_t stage_bytes, int out_bufs) {
    return n_taps * kWTapBytes + n_stages * static_cast<int>(stage_bytes) + 2 * out_bufs * kStagingTileBytes + kBarrierBytes + 1024;
  }
};

// Epilogue of the halo conv (warps 0..7): TMEM -> registers (+bias, +residual, ReLU) -> 16-bit tile in shared memory in
// the TMA layout -> ONE cp.async.bulk.tensor store per tile (coalesced by hardware, clipped at the image border).
// These convs have short K (7 or 27 MMAs per tile), so a tile's epilogue costs about as much as its MMAs: the eight
// epilogue warps therefore work as TWO independent groups of four (group g owns accumulator g, output staging tile g and
// named barrier 1 + g) and drain two tiles concurrently. Thread = accumulator row (TMEM lane quarter = warp % 4), all 64
// output channels. The residual (identity branch of the BasicBlock) is the thread's own 128-byte row, prefetched into
// registers before the accumulator is waited for.
template <int RES>
__device__ __forceinline__ void halo_epilogue(const CUtensorMap* tmap_out, const GemmParams& p, int out_bufs, uint8_t* out_stage_base,
                                              uint32_t tmem_base, uint64_t* tmem_full_bar, uint64_t* tmem_empty_bar,
                                              int total_tiles, int warp, int lane) {
  const int q = warp & 3, grp = warp >> 2;
  const int r = q * 32 + lane;
  const bool issuer = (q == 0 && lane == 0);
  const int wl = r % p.box_w, hl = r / p.box_w;
  const float relu_floor = p.relu ? 0.0f : -INFINITY;
  const uint32_t out_base0 = smem_u32(out_stage_base + grp * out_bufs * 16384);
  uint32_t buf = 0;
  uint32_t off[8];                                   // this thread's eight 16-byte chunks inside its 128-byte row
#pragma unroll
  for (int c = 0; c < 8; ++c) off[c] = r * 128 + ((c ^ (r & 7)) << 4);
  uint32_t acc_phase = 0;
  // residual row of tile `t` for this thread (nullptr outside the image or past the last tile)
  auto res_row = [&](int t) -> const uint4* {
    if (RES == 0 || t >= total_tiles) return nullptr;
    const int w = (t % p.tiles_w) * p.box_w + wl, h = ((t / p.tiles_w) % p.tiles_h) * p.box_h + hl;
    const int bt2 = t / (p.tiles_w * p.tiles_h);
    if (w >= p.W || h >= p.H) return nullptr;
    return reinterpret_cast<const uint4*>(static_cast<const act_t*>(p.residual) + ((static_cast<long long>(bt2) * p.H + h) * p.W + w) * 64);
  };
  uint4 rv[8];
  if (RES != 0) {
    const uint4* rp = res_row(blockIdx.x + grp * gridDim.x);
#pragma unroll
    for (int c = 0; c < 8; ++c) rv[c] = rp ? __ldg(rp + c) : make_uint4(0u, 0u, 0u, 0u);
  }
  for (int tile = blockIdx.x + grp * gridDim.x; tile < total_tiles; tile += 2 * gridDim.x) {
    const int w0 = (tile % p.tiles_w) * p.box_w;
    const int h0 = ((tile / p.tiles_w) % p.tiles_h) * p.box_h;
    const int bt = tile / (p.tiles_w * p.tiles_h);
    const uint4* rp_next = res_row(tile + 2 * gridDim.x);       // reloaded chunk by chunk as soon as a chunk is consumed
    mbar_wait(&tmem_full_bar[grp], acc_phase);
    tc_fence_after();
    uint32_t v0[32], v1[32];
    const uint32_t t_row = tmem_base + grp * 64 + (static_cast<uint32_t>(q * 32) << 16);
    tmem_ld_32x32b_x32(t_row, v0);
    tmem_ld_32x32b_x32(t_row + 32, v1);
    tmem_ld_wait();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(&tmem_empty_bar[grp]);           // 4 arrivals free this group's accumulator
    const uint32_t out_base = out_base0 + buf * 16384;
    if (issuer) { if (out_bufs == 2) bulk_wait_group_read1(); else bulk_wait_group_read0(); }   // the store that last used this staging tile has read it
    named_bar_sync(1 + grp, 128);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      const uint32_t* v = (c < 4) ? &v0[8 * c] : &v1[8 * (c - 4)];
      float f[8];
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const float4 b4 = (p.bias != nullptr) ? __ldg(reinterpret_cast<const float4*>(p.bias + 8 * c) + h2) : make_float4(0.f, 0.f, 0.f, 0.f);
        f[4 * h2 + 0] = __uint_as_float(v[4 * h2 + 0]) + b4.x;
        f[4 * h2 + 1] = __uint_as_float(v[4 * h2 + 1]) + b4.y;
        f[4 * h2 + 2] = __uint_as_float(v[4 * h2 + 2]) + b4.z;
        f[4 * h2 + 3] = __uint_as_float(v[4 * h2 + 3]) + b4.w;
      }
      if (RES != 0) {
        const uint32_t u[4] = {rv[c].x, rv[c].y, rv[c].z, rv[c].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 t = unpack_act2(u[j]);
          f[2 * j] += t.x; f[2 * j + 1] += t.y;
        }
        rv[c] = rp_next ? __ldg(rp_next + c) : make_uint4(0u, 0u, 0u, 0u);   // next tile's residual, in flight from here
      }
      st_shared_v4(out_base + off[c], pack_act2(fmaxf(f[0], relu_floor), fmaxf(f[1], relu_floor)),
                   pack_act2(fmaxf(f[2], relu_floor), fmaxf(f[3], relu_floor)),
                   pack_act2(fmaxf(f[4], relu_floor), fmaxf(f[5], relu_floor)),
                   pack_act2(fmaxf(f[6], relu_floor), fmaxf(f[7], relu_floor)));
    }
    fence_proxy_async_smem();                                   // generic-proxy writes -> visible to the TMA store
    named_bar_sync(1 + grp, 128);
    if (issuer) {
      tma_store_5d(tmap_out, out_base, 0, w0, 0, h0, bt);
      bulk_commit_group();
    }
    acc_phase ^= 1;
    if (out_bufs == 2) buf ^= 1;
  }
  if (issuer) bulk_wait_group0();
}

template <int BLOCK_K, int RES>
__global__ void __launch_bounds__(kGemmThreads, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                 const __grid_constant__ HaloParams hp) {
  pdl_trigger();
  using S = HaloSmem<BLOCK_K>;
  constexpr int BLOCK_N = 64;
  constexpr uint32_t kSwizzleLayout = (BLOCK_K == 64) ? 2u : 4u;   // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint32_t kSBO = 8u * BLOCK_K * 2u;
  constexpr uint32_t kTmemCols = 2 * BLOCK_N;
  constexpr uint32_t kIdesc = make_idesc_f16(128, BLOCK_N, false, false);
  const GemmParams& p = hp.g;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_w = smem;
  uint8_t* smem_a = smem_w + hp.n_taps * S::kWTapBytes;
  uint8_t* smem_stage = smem_a + hp.n_stages * hp.stage_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_stage + 2 * hp.out_bufs * S::kStagingTileBytes);
  uint64_t* full_bar = bars;                              // [kHaloMaxStages]
  uint64_t* empty_bar = bars + kHaloMaxStages;            // [kHaloMaxStages]
  uint64_t* tmem_full_bar = bars + 2 * kHaloMaxStages;    // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;           // [2]
  uint64_t* w_bar = tmem_empty_bar + 2;                   // weights resident
  uint64_t* res_full = w_bar + 1;                         // residual tile landed
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(res_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kHaloMaxStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 4); }
    mbar_init(w_bar, 1);
    mbar_init(res_full, 1);
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) { prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_b); prefetch_tmap(&tmap_out); prefetch_tmap(&tmap_res); }
  if (warp == 9) { tmem_alloc(tmem_ptr_smem, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();                                           // everything above ran under the previous kernel's tail

  const int total_tiles = p.tiles_w * p.tiles_h * p.tiles_b;

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer
    const uint32_t a_base = smem_u32(smem_a), full_base = smem_u32(full_bar);
    if (elect_one()) {
      const uint32_t wb = smem_u32(w_bar);
      mbar_arrive_expect_tx_u32(wb, static_cast<uint32_t>(hp.n_taps * S::kWTapBytes));
      for (int t = 0; t < hp.n_taps; ++t)
        tma_load_2d_u32(smem_u32(smem_w) + t * S::kWTapBytes, &tmap_b, wb, t * BLOCK_K, 0);
    }
    __syncwarp();
    int stage = 0; uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int wt = tile % p.tiles_w;
      const int ht = (tile / p.tiles_w) % p.tiles_h;
      const int bt = tile / (p.tiles_w * p.tiles_h);
      const int w0 = wt * p.box_w, h0 = ht * p.box_h;
      for (int l = 0; l < hp.n_loads; ++l) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        if (elect_one()) {
          const uint32_t fb = full_base + stage * 8;
          mbar_arrive_expect_tx_u32(fb, hp.box_bytes);
          tma_load_5d_u32(a_base + stage * hp.stage_bytes, &tmap_a, fb, hp.ld_c0[l], w0 + hp.ld_d1[l], hp.ld_d2[l],
                          h0 + hp.ld_d3[l], bt);
        }
        __syncwarp();
        if (++stage == hp.n_stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer (one elected lane issues)
    const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), 0, kSBO, kSwizzleLayout);
    const uint64_t w_desc0 = make_smem_desc(smem_u32(smem_w), 0, kSBO, kSwizzleLayout);
    const uint32_t empty_base = smem_u32(empty_bar);
    mbar_wait(w_bar, 0);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
      // The tap structure is compile-time (BLOCK_K 64: 3x3 conv, box l = filter column, taps = filter rows;
      // BLOCK_K 32: stem, box 0 = odd input rows with kh = 0,2,4,6, box 1 = even rows with kh = 1,3,5), so each
      // tcgen05.mma costs the single issuing thread only its descriptor adds. A tap's A operand starts
      // kh * (box_w rows) into the box; its weights are slice `w` of the resident array.
      constexpr int kLoads = (BLOCK_K == 64) ? 3 : 2;
      const uint32_t rb16 = static_cast<uint32_t>(p.box_w * BLOCK_K * 2) >> 4;   // one box row of pixels, in 16-byte units
#pragma unroll
      for (int l = 0; l < kLoads; ++l) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da0 = a_desc0 + static_cast<uint64_t>((stage * hp.stage_bytes) >> 4);
          const int kTaps = (BLOCK_K == 64) ? 3 : (4 - l);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (i < kTaps) {
              const int wslice = (BLOCK_K == 64) ? (i * 3 + l) : (2 * i + l);
              const uint64_t da = da0 + static_cast<uint64_t>(i * rb16);
              const uint64_t db = w_desc0 + static_cast<uint64_t>(wslice * (S::kWTapBytes >> 4));
#pragma unroll
              for (int k = 0; k < BLOCK_K / 16; ++k)
                umma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, kIdesc, (l | i | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit_u32(empty_base + stage * 8);
        }
        __syncwarp();
        if (++stage == hp.n_stages) { stage = 0; phase ^= 1; }
      }
      if (elect_one()) umma_commit(&tmem_full_bar[acc]);
      __syncwarp();
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else {
    halo_epilogue<RES>(&tmap_out, p, hp.out_bufs, smem_stage, tmem_base, tmem_full_bar, tmem_empty_bar, total_tiles, warp, lane);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

}  // namespace rmb
