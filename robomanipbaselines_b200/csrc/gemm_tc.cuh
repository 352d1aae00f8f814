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
constexpr int kGemmThreads = 192;  // warps 0-3: epilogue, warp 4: TMA producer, warp 5: MMA issuer + TMEM owner

struct GemmParams {
  // row (pixel) tiling
  int tiles_w, tiles_h, tiles_b;
  int box_w, box_h, box_b;   // rows per tile = box_w*box_h*box_b <= 128
  int W, H, B;               // output extents (rows = B*H*W)
  int n_col_tiles;           // N / BLOCK_N
  // K loop: k-block kb -> (tap = kb / chunks_per_tap, chunk = kb % chunks_per_tap)
  int num_taps, chunks_per_tap;
  int tap_c0[kMaxTaps];      // added to coordinate 0 (channel / phase-channel offset)
  int tap_d1[kMaxTaps];      // added to coordinate 1 (w)
  int tap_d2[kMaxTaps];      // coordinate 2 (h phase for stride-2 views, else 0)
  int tap_d3[kMaxTaps];      // added to coordinate 3 (h)
  // epilogue
  const float* bias;         // [N] or nullptr
  const void* residual;      // [rows, ld_res] or nullptr
  void* out;                 // [rows, ld_out]
  int ld_out, ld_res;
  int out_f32, res_f32, relu;
  // optional output-row remap: rows (img, h, w) -> token rows of DETRVAE's flattened multi-camera feature map
  //   out_row = (img / remap_ncam) * remap_L + remap_off + h * (W * remap_ncam) + (img % remap_ncam) * W + w
  int remap_ncam, remap_L, remap_off;
};

template <int BLOCK_N, int BLOCK_K, int STAGES>
struct GemmSmem {
  static constexpr int kABytes = 128 * BLOCK_K * 2;
  static constexpr int kBBytes = BLOCK_N * BLOCK_K * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarrierBytes = 256;
  static constexpr int kTotal = STAGES * kStageBytes + kBarrierBytes + 1024;  // +1024: manual alignment slack
};

template <int BLOCK_N, int BLOCK_K, int STAGES>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ GemmParams p) {
  static_assert(BLOCK_K == 64 || BLOCK_K == 32, "BLOCK_K selects the 128B / 64B swizzle");
  static_assert(BLOCK_N == 64 || BLOCK_N == 128 || BLOCK_N == 256, "BLOCK_N");
  using S = GemmSmem<BLOCK_N, BLOCK_K, STAGES>;
  constexpr uint32_t kSwizzleLayout = (BLOCK_K == 64) ? 2u : 4u;   // SWIZZLE_128B : SWIZZLE_64B
  constexpr uint32_t kSBO = 8u * BLOCK_K * 2u;                      // 8-row core-matrix group stride
  constexpr uint32_t kTmemCols = 2 * BLOCK_N;                       // double-buffered accumulator
  constexpr uint32_t kIdesc = make_idesc_f16(128, BLOCK_N, false, false);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * S::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * S::kStageBytes);
  uint64_t* full_bar = bars;                    // [STAGES]
  uint64_t* empty_bar = bars + STAGES;          // [STAGES]
  uint64_t* tmem_full_bar = bars + 2 * STAGES;  // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2; // [2]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full_bar[a], 1); mbar_init(&tmem_empty_bar[a], 4); }
    fence_mbar_init();
  }
  if (warp == 4 && lane == 0) { prefetch_tmap(&tmap_a); prefetch_tmap(&tmap_b); }
  if (warp == 5) { tmem_alloc(tmem_ptr_smem, kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  const int rows_per_tile = p.box_w * p.box_h * p.box_b;
  const int num_kb = p.num_taps * p.chunks_per_tap;
  const int row_tiles = p.tiles_w * p.tiles_h * p.tiles_b;
  const int total_tiles = row_tiles * p.n_col_tiles;
  const uint32_t tx_bytes = static_cast<uint32_t>(rows_per_tile * BLOCK_K * 2 + S::kBBytes);

  if (warp == 4) {
    // ------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int ct = tile % p.n_col_tiles;
        const int rt = tile / p.n_col_tiles;
        const int wt = rt % p.tiles_w;
        const int ht = (rt / p.tiles_w) % p.tiles_h;
        const int bt = rt / (p.tiles_w * p.tiles_h);
        const int w0 = wt * p.box_w, h0 = ht * p.box_h, b0 = bt * p.box_b;
        for (int kb = 0; kb < num_kb; ++kb) {
          const int tap = kb / p.chunks_per_tap;
          const int ch = kb - tap * p.chunks_per_tap;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
          tma_load_5d(smem_a + stage * S::kABytes, &tmap_a, &full_bar[stage], p.tap_c0[tap] + ch * BLOCK_K,
                      w0 + p.tap_d1[tap], p.tap_d2[tap], h0 + p.tap_d3[tap], b0);
          tma_load_2d(smem_b + stage * S::kBBytes, &tmap_b, &full_bar[stage], kb * BLOCK_K, ct * BLOCK_N);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 5) {
    // ------------------------------------------------------------ MMA issuer (single thread)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * S::kABytes);
          const uint32_t b_addr = smem_u32(smem_b + stage * S::kBBytes);
#pragma unroll
          for (int k = 0; k < BLOCK_K / 16; ++k) {
            const uint64_t da = make_smem_desc(a_addr + k * 32, 0, kSBO, kSwizzleLayout);
            const uint64_t db = make_smem_desc(b_addr + k * 32, 0, kSBO, kSwizzleLayout);
            umma_f16_ss(d_tmem, da, db, kIdesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);   // frees this smem stage when the MMAs above retire
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full_bar[acc]);   // accumulator complete -> epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue warps 0..3 (TMEM lanes 32*warp..+31)
    int acc = 0; uint32_t acc_phase = 0;
    const int r = warp * 32 + lane;                 // row of the tile owned by this thread
    const int wl = r % p.box_w;
    const int hl = (r / p.box_w) % p.box_h;
    const int bl = r / (p.box_w * p.box_h);
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int ct = tile % p.n_col_tiles;
      const int rt = tile / p.n_col_tiles;
      const int wt = rt % p.tiles_w;
      const int ht = (rt / p.tiles_w) % p.tiles_h;
      const int bt = rt / (p.tiles_w * p.tiles_h);
      const int w = wt * p.box_w + wl, h = ht * p.box_h + hl, b = bt * p.box_b + bl;
      const bool valid = (r < rows_per_tile) && (w < p.W) && (h < p.H) && (b < p.B);
      long long grow;                               // source row (residual) in natural pixel order
      long long orow;                               // destination row
      grow = (static_cast<long long>(b) * p.H + h) * p.W + w;
      if (p.remap_ncam > 0) {
        orow = static_cast<long long>(b / p.remap_ncam) * p.remap_L + p.remap_off +
               static_cast<long long>(h) * (p.W * p.remap_ncam) + (b % p.remap_ncam) * p.W + w;
      } else {
        orow = grow;
      }
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + acc * BLOCK_N + (static_cast<uint32_t>(warp * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BLOCK_N; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(t_row + c, v);
        tmem_ld_wait();
        if (valid) {
          const int col = ct * BLOCK_N + c;
          float f[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) f[i] = __uint_as_float(v[i]);
          if (p.bias != nullptr) {
            const float4* bp = reinterpret_cast<const float4*>(p.bias + col);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 bv = __ldg(bp + i);
              f[4 * i + 0] += bv.x; f[4 * i + 1] += bv.y; f[4 * i + 2] += bv.z; f[4 * i + 3] += bv.w;
            }
          }
          if (p.residual != nullptr) {
            if (p.res_f32) {
              const float4* rp = reinterpret_cast<const float4*>(static_cast<const float*>(p.residual) +
                                                                  grow * p.ld_res + col);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 rv = __ldg(rp + i);
                f[4 * i + 0] += rv.x; f[4 * i + 1] += rv.y; f[4 * i + 2] += rv.z; f[4 * i + 3] += rv.w;
              }
            } else {
              const uint4* rp = reinterpret_cast<const uint4*>(static_cast<const act_t*>(p.residual) +
                                                                grow * p.ld_res + col);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint4 rv = __ldg(rp + i);
                const uint32_t u[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 t = unpack_act2(u[j]);
                  f[8 * i + 2 * j + 0] += t.x;
                  f[8 * i + 2 * j + 1] += t.y;
                }
              }
            }
          }
          if (p.relu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = fmaxf(f[i], 0.0f);
          }
          if (p.out_f32) {
            float4* op = reinterpret_cast<float4*>(static_cast<float*>(p.out) + orow * p.ld_out + col);
#pragma unroll
            for (int i = 0; i < 8; ++i) op[i] = make_float4(f[4 * i], f[4 * i + 1], f[4 * i + 2], f[4 * i + 3]);
          } else {
            uint4* op = reinterpret_cast<uint4*>(static_cast<act_t*>(p.out) + orow * p.ld_out + col);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              op[i] = make_uint4(pack_act2(f[8 * i], f[8 * i + 1]), pack_act2(f[8 * i + 2], f[8 * i + 3]),
                                 pack_act2(f[8 * i + 4], f[8 * i + 5]), pack_act2(f[8 * i + 6], f[8 * i + 7]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);   // 4 arrivals (one per epilogue warp) free the accumulator
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem_base, kTmemCols); }
}

}  // namespace rmb
