// ResNet stem fused with its max-pool: conv1 7x7/2 (+ folded FrozenBN + ReLU) -> maxpool 3x3/2 pad 1 in ONE kernel
// (reference: torchvision resnet18 conv1/bn1/relu/maxpool via third_party/act/detr/models/backbone.py:86-96).
//
// Unfused, the stem writes its [img, H/2, W/2, 64] output to HBM (5 GB at 512 images of 480x640) and the pool reads it
// back: 12.6 GB moved for 2.5 GB of algorithmic bytes. Here the conv output never leaves the SM.
//
// MMA side: as conv_halo_kernel<32> (one TMA box per input-row parity carrying the vertical halo, 7 taps = 14
// tcgen05.mma on shifted descriptors, weights resident), with 32 x 4 conv-pixel tiles.
// Pooling side: a pooling window straddles tile borders, so tiles are walked in an order that makes the halo a CARRY
// instead of a recompute: each of the two 4-warp epilogue groups owns a BAND (an image's rows [4 t0, 4 t1)) and walks it
// strip by strip (32 conv columns), top to bottom inside a strip.
//   * vertical 3-max first: V_i[col] for the tile's two pooled rows needs conv rows 4t-1 .. 4t+3; row 4t-1 is the last
//     row of the tile above = `prow` (4 KB of shared memory per group);
//   * horizontal 3-max second: pooled column j needs V[2j-1 .. 2j+1]; the column left of the tile is the last V column
//     of the previous strip at the same rows = `vcarry` (128 B per pooled row of the band).
// Zero stands in for the pool's -inf padding (all values are post-ReLU, every window holds a real element). A band that
// does not start at the image top first runs the tile row above it with stores off (1 extra tile per strip), which is
// the only recomputation: 1/16 at 4 bands per 480x640 image.
// max commutes with the monotone bias + ReLU + fp16/bf16 rounding, so results equal the unfused kernels bit for bit.
#pragma once
#include "conv_halo.cuh"

namespace rmb {

constexpr int kSpBoxW = 32, kSpBoxH = 4;                 // conv pixels per tile: 32 x 4 = 128 MMA rows
constexpr int kSpBoxBytes = (kSpBoxH + 3) * kSpBoxW * 64;  // one parity box: 7 rows x 32 windows x 64 B = 14336
constexpr int kSpWBytes = 7 * 64 * 32 * 2;               // resident weights, 7 taps
constexpr int kSpStagingBytes = 16384;                   // 128 px x 64 ch act_t, per group
constexpr int kSpProwBytes = kSpBoxW * 128;              // last conv row of the tile above, per group
constexpr int kSpVbufBytes = 2 * (kSpBoxW + 1) * 128;    // V for 2 pooled rows x (carry col + 32 cols), per group
constexpr int kSpMaxBandRows = 32;                       // tile rows per band (vcarry: 2 pooled rows each)
constexpr int kSpVcarryBytes = 2 * kSpMaxBandRows * 128;
constexpr int kSpGroupBytes = kSpStagingBytes + kSpProwBytes + ((kSpVbufBytes + 1023) & ~1023) + kSpVcarryBytes;
constexpr int kSpMaxStages = 8;

struct StemPoolParams {
  int n_img, OH, OW;            // conv output extents (H/2, W/2); pooled output is OH/2 x OW/2
  int strips;                   // ceil(OW / 32)
  int bands_per_img, rows_band; // band = rows_band tile rows (4 conv rows each); bands_per_img * rows_band = OH / 4
  int n_bands;                  // n_img * bands_per_img
  int n_stages;
  const float* bias;            // [64]
  act_t* out;                   // [n_img, OH/2, OW/2, 64]
};

// Tile sequence of one epilogue group: bands gid, gid + G, ...; inside a band strip-major, top to bottom; a band with
// t0 > 0 starts every strip one tile row early (stores off) to rebuild the `prow` carry. Warp-uniform.
struct BandWalk {
  int band, band_step, n_bands, nbi, rows_band, strips;
  int img, t0, t_end, s, t;
  bool valid;
  __device__ __forceinline__ void start_band() {
    img = band / nbi;
    t0 = (band - img * nbi) * rows_band;
    t_end = t0 + rows_band;
    s = 0;
    t = t0 > 0 ? t0 - 1 : 0;
  }
  __device__ __forceinline__ void init(const StemPoolParams& p, int gid, int n_groups) {
    band = gid; band_step = n_groups; n_bands = p.n_bands; nbi = p.bands_per_img; rows_band = p.rows_band; strips = p.strips;
    valid = band < n_bands;
    if (valid) start_band();
  }
  __device__ __forceinline__ void next() {
    if (++t < t_end) return;
    if (++s < strips) { t = t0 > 0 ? t0 - 1 : 0; return; }
    band += band_step;
    valid = band < n_bands;
    if (valid) start_band();
  }
};

__device__ __forceinline__ uint4 act8_max3(uint4 a, uint4 b, uint4 c) {
  uint4 r;
  r.x = act2_max(act2_max(a.x, b.x), c.x); r.y = act2_max(act2_max(a.y, b.y), c.y);
  r.z = act2_max(act2_max(a.z, b.z), c.z); r.w = act2_max(act2_max(a.w, b.w), c.w);
  return r;
}
__device__ __forceinline__ void st_shared_u4(uint32_t addr, uint4 v) { st_shared_v4(addr, v.x, v.y, v.z, v.w); }

__global__ void __launch_bounds__(kGemmThreads, 1)
stem_pool_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ StemPoolParams p) {
  pdl_trigger();
  constexpr int BLOCK_K = 32, BLOCK_N = 64;
  constexpr uint32_t kSBO = 8u * BLOCK_K * 2u;
  constexpr uint32_t kTmemCols = 2 * BLOCK_N;
  constexpr uint32_t kIdesc = make_idesc_f16(128, BLOCK_N, false, false);
  constexpr int kWTap = 64 * BLOCK_K * 2;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_w = smem;
  uint8_t* smem_a = smem_w + kSpWBytes;
  uint8_t* smem_grp = smem_a + p.n_stages * kSpBoxBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_grp + 2 * kSpGroupBytes);
  uint64_t* full_bar = bars;                          // [kSpMaxStages]
  uint64_t* empty_bar = bars + kSpMaxStages;          // [kSpMaxStages]
  uint64_t* tmem_full_bar = bars + 2 * kSpMaxStages;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;       // [2]
  uint64_t* w_bar = tmem_empty_bar + 2;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(w_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kSpMaxStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 4); }
    mbar_init(w_bar, 1);
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) { prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_b); }
  if (warp == 9) { tmem_alloc(tmem_ptr_smem, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  pdl_wait();

  const int n_groups = 2 * gridDim.x;

  if (warp == 8) {
    // ------------------------------------------------------------ TMA producer: 2 parity boxes per tile
    const uint32_t a_base = smem_u32(smem_a), full_base = smem_u32(full_bar);
    if (elect_one()) {
      const uint32_t wb = smem_u32(w_bar);
      mbar_arrive_expect_tx_u32(wb, static_cast<uint32_t>(kSpWBytes));
      for (int t = 0; t < 7; ++t) tma_load_2d_u32(smem_u32(smem_w) + t * kWTap, &tmap_b, wb, t * BLOCK_K, 0);
    }
    __syncwarp();
    BandWalk w[2];
    w[0].init(p, 2 * blockIdx.x, n_groups);
    w[1].init(p, 2 * blockIdx.x + 1, n_groups);
    int stage = 0; uint32_t phase = 0;
    while (w[0].valid || w[1].valid) {
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        if (!w[g].valid) continue;
        const int w0 = w[g].s * kSpBoxW, h0 = w[g].t * kSpBoxH, img = w[g].img;
#pragma unroll
        for (int l = 0; l < 2; ++l) {                 // l = 0: odd input rows (kh 0,2,4,6), l = 1: even rows (kh 1,3,5)
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (elect_one()) {
            const uint32_t fb = full_base + stage * 8;
            mbar_arrive_expect_tx_u32(fb, kSpBoxBytes);
            tma_load_5d_u32(a_base + stage * kSpBoxBytes, &tmap_a, fb, 0, w0, 1 - l, h0 + (l == 0 ? -2 : -1), img);
          }
          __syncwarp();
          if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
        }
        w[g].next();
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------ MMA issuer: accumulator g belongs to epilogue group g
    const uint64_t a_desc0 = make_smem_desc(smem_u32(smem_a), 0, kSBO, 4u);   // SWIZZLE_64B
    const uint64_t w_desc0 = make_smem_desc(smem_u32(smem_w), 0, kSBO, 4u);
    const uint32_t empty_base = smem_u32(empty_bar);
    constexpr uint32_t rb16 = static_cast<uint32_t>(kSpBoxW * BLOCK_K * 2) >> 4;   // one box row of windows, 16-byte units
    mbar_wait(w_bar, 0);
    BandWalk w[2];
    w[0].init(p, 2 * blockIdx.x, n_groups);
    w[1].init(p, 2 * blockIdx.x + 1, n_groups);
    int stage = 0; uint32_t phase = 0;
    uint32_t acc_phase[2] = {0, 0};
    while (w[0].valid || w[1].valid) {
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        if (!w[g].valid) continue;
        mbar_wait(&tmem_empty_bar[g], acc_phase[g] ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + g * BLOCK_N;
#pragma unroll
        for (int l = 0; l < 2; ++l) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t da0 = a_desc0 + static_cast<uint64_t>((stage * kSpBoxBytes) >> 4);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              if (i < 4 - l) {                         // tap i of this box is filter row kh = 2 i + l, i box rows down
                const uint64_t da = da0 + static_cast<uint64_t>(i * rb16);
                const uint64_t db = w_desc0 + static_cast<uint64_t>((2 * i + l) * (kWTap >> 4));
#pragma unroll
                for (int k = 0; k < BLOCK_K / 16; ++k)
                  umma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, kIdesc, (l | i | k) != 0 ? 1u : 0u);
              }
            }
            umma_commit_u32(empty_base + stage * 8);
          }
          __syncwarp();
          if (++stage == p.n_stages) { stage = 0; phase ^= 1; }
        }
        if (elect_one()) umma_commit(&tmem_full_bar[g]);
        __syncwarp();
        acc_phase[g] ^= 1;
        w[g].next();
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue + pooling: group grp = warps 4 grp .. 4 grp + 3
    const int q = warp & 3, grp = warp >> 2;
    const int tid = q * 32 + lane;                      // 0..127 inside the group; also the accumulator row (conv pixel)
    const uint32_t g_base = smem_u32(smem_grp + grp * kSpGroupBytes);
    const uint32_t S = g_base;                                            // [128 px][128 B], 16-byte chunks XOR (px & 7)
    const uint32_t PROW = g_base + kSpStagingBytes;                       // [32 cols][128 B], chunks XOR (col & 7)
    const uint32_t VBUF = PROW + kSpProwBytes;                            // [2][33 cols][128 B], chunks XOR (k & 7)
    const uint32_t VCARRY = VBUF + ((kSpVbufBytes + 1023) & ~1023);       // [2 * rows_band][128 B]
    uint32_t s_off[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) s_off[c] = S + tid * 128 + ((c ^ (tid & 7)) << 4);
    const int POH = p.OH >> 1, POW = p.OW >> 1;
    BandWalk w;
    w.init(p, 2 * blockIdx.x + grp, n_groups);
    uint32_t acc_phase = 0;
    while (w.valid) {
      const int t = w.t, s = w.s;
      const bool store = t >= w.t0;                     // false: warm-up tile row above the band (only prow is kept)
      // ---- accumulator -> bias + ReLU -> 16-bit -> staging tile
      mbar_wait(&tmem_full_bar[grp], acc_phase);
      tc_fence_after();
      uint32_t v0[32], v1[32];
      const uint32_t t_row = tmem_base + grp * BLOCK_N + (static_cast<uint32_t>(q * 32) << 16);
      tmem_ld_32x32b_x32(t_row, v0);
      tmem_ld_32x32b_x32(t_row + 32, v1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[grp]);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint32_t* v = (c < 4) ? &v0[8 * c] : &v1[8 * (c - 4)];
        const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + 8 * c));
        const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + 8 * c) + 1);
        st_shared_v4(s_off[c],
                     pack_act2(fmaxf(__uint_as_float(v[0]) + b0.x, 0.f), fmaxf(__uint_as_float(v[1]) + b0.y, 0.f)),
                     pack_act2(fmaxf(__uint_as_float(v[2]) + b0.z, 0.f), fmaxf(__uint_as_float(v[3]) + b0.w, 0.f)),
                     pack_act2(fmaxf(__uint_as_float(v[4]) + b1.x, 0.f), fmaxf(__uint_as_float(v[5]) + b1.y, 0.f)),
                     pack_act2(fmaxf(__uint_as_float(v[6]) + b1.z, 0.f), fmaxf(__uint_as_float(v[7]) + b1.w, 0.f)));
      }
      named_bar_sync(1 + grp, 128);
      // ---- pass A, vertical: thread = conv column `lane`, chunks 2q and 2q + 1 (16 channels)
      {
        const bool top = (t == 0);                      // image top: the row above is the pool's padding
#pragma unroll
        for (int cc = 0; cc < 2; ++cc) {
          const int c = 2 * q + cc;
          const uint32_t sw = static_cast<uint32_t>((c ^ (lane & 7)) << 4);
          const uint32_t pr_addr = PROW + lane * 128 + sw;
          uint4 r[4];
#pragma unroll
          for (int h = 0; h < 4; ++h) r[h] = ld_shared_v4u(S + (h * 32 + lane) * 128 + sw);   // (h*32 + lane) & 7 == lane & 7
          const uint4 pr = top ? make_uint4(0u, 0u, 0u, 0u) : ld_shared_v4u(pr_addr);
          const uint32_t swv = static_cast<uint32_t>((c ^ ((lane + 1) & 7)) << 4);
          st_shared_u4(VBUF + (lane + 1) * 128 + swv, act8_max3(pr, r[0], r[1]));
          st_shared_u4(VBUF + (kSpBoxW + 1 + lane + 1) * 128 + swv, act8_max3(r[1], r[2], r[3]));
          st_shared_u4(pr_addr, r[3]);
        }
        if (tid < 16 && store) {                        // column left of the tile: last V column of the previous strip
          const int i = tid >> 3, c = tid & 7;
          const uint4 cv = (s == 0) ? make_uint4(0u, 0u, 0u, 0u)
                                    : ld_shared_v4u(VCARRY + (2 * (t - w.t0) + i) * 128 + c * 16);
          st_shared_u4(VBUF + (i * (kSpBoxW + 1)) * 128 + (c << 4), cv);   // k = 0: swizzle term is c ^ 0
        }
      }
      named_bar_sync(1 + grp, 128);
      // ---- pass B, horizontal + store: item = (pooled row i, pooled col j, chunk c), two items per thread
      if (store) {
        const int c = tid & 7, j = (tid >> 3) & 15;
        const int pcol = (s * kSpBoxW >> 1) + j;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const uint32_t vb = VBUF + (i * (kSpBoxW + 1)) * 128;
          const int k0 = 2 * j;
          const uint4 a = ld_shared_v4u(vb + k0 * 128 + ((c ^ (k0 & 7)) << 4));
          const uint4 b = ld_shared_v4u(vb + (k0 + 1) * 128 + ((c ^ ((k0 + 1) & 7)) << 4));
          const uint4 d = ld_shared_v4u(vb + (k0 + 2) * 128 + ((c ^ ((k0 + 2) & 7)) << 4));
          const uint4 m = act8_max3(a, b, d);
          const int prow_i = 2 * t + i;
          if (pcol < POW && prow_i < POH)
            reinterpret_cast<uint4*>(p.out + ((static_cast<long long>(w.img) * POH + prow_i) * POW + pcol) * 64)[c] = m;
        }
        if (tid < 16) {                                 // this strip's last V column -> carry for the next strip
          const int i = tid >> 3, cc = tid & 7;
          const int k = kSpBoxW;                        // conv column 31 of the tile
          const uint4 cv = ld_shared_v4u(VBUF + (i * (kSpBoxW + 1) + k) * 128 + ((cc ^ (k & 7)) << 4));
          st_shared_u4(VCARRY + (2 * (t - w.t0) + i) * 128 + cc * 16, cv);
        }
      }
      acc_phase ^= 1;
      w.next();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

}  // namespace rmb
