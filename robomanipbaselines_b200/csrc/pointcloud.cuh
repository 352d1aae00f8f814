// Device-side DP3 observation pipeline: the step immediately before DP3.predict_action.
//   convert_depth_image_to_pointcloud  robo_manip_baselines/common/utils/VisionUtils.py:55-87
//   crop_pointcloud_bb                 robo_manip_baselines/common/utils/Vision3dUtils.py:6-14
//   downsample_pointcloud_fps          robo_manip_baselines/common/utils/Vision3dUtils.py:17-25 (pytorch3d FPS, start index 0)
//   normalize_data + float32 cast      policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:152-153
// One CTA per cloud. Arithmetic is fp64 with the reference's dtype flow and op order (see oracle/pointcloud_oracle.py):
// separately rounded multiplies / adds, so back-projection, crop, FPS indices and the normalised float32 output are
// bit-exact with the numpy path. Included at the end of act_engine.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <mutex>
#include <string>

namespace rmb {

constexpr int kPcThreads = 512;
constexpr int kPcPpt = 14;                            // points per thread whose FPS state lives in registers
constexpr int kPcGroup = 2;                           // points of a thread swept together (instruction-level parallelism)
constexpr int kPcCap = kPcThreads * kPcPpt;           // clouds of up to 7168 kept points run the resident FPS loop
// dynamic shared memory of the resident path: (x, y) double2 [cap] | z float [cap] | rgb packed u32 [cap] | u8/255 table
constexpr int kPcSmemBytes = kPcCap * 24 + 256 * 4;

// float -> double inside the sweep: volatile so that ptxas does not hoist the conversions of the register-resident
// colours out of the loop (it then keeps them as doubles and spills)
__device__ __forceinline__ double pc_widen(float f) {
  double d;
  asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d) : "f"(f));
  return d;
}

struct PcParams {
  int H, W, K;
  double focal, near_clip, far_clip;
  int has_min, has_max;
  double mn[3], mx[3];
  int norm_type;            // -1 none, 0 gaussian ((x - p0) / p1), 1 limits (p1 * (x - p0) + out_min)
  double p0[6], p1[6], out_min;
};

// scratch per cloud: pts [HW][6] double, mind [HW] double
// Farthest point sampling is 1023 dependent sweeps over the kept points. Round 2: a cloud of up to kPcCap points (the
// DP3 input is 84 x 84 = 7056 pixels) is swept from ON-CHIP state -- coordinates in shared memory (x, y as doubles; z and
// the colours are float32-valued, so they are kept as floats and widened, which is exact), the running minimum distance
// and the point's colour in registers (thread t owns points t, t + 512, ...: the same assignment and so the same
// first-maximum tie-break as the sweep over global memory), one __syncthreads per sweep (double-buffered partial
// maxima, every warp finishes the reduction itself). Same fp64 operations in the same order => same indices bit for
// bit; larger clouds take the sweep over the scratch buffer.
__global__ void __launch_bounds__(kPcThreads, 1) pointcloud_from_depth_kernel(const float* __restrict__ depth,
                                                                              const uint8_t* __restrict__ rgb,
                                                                              const PcParams p, double* __restrict__ scratch,
                                                                              float* __restrict__ out, int* __restrict__ count,
                                                                              int* __restrict__ indices) {
  extern __shared__ __align__(16) uint8_t pc_smem[];
  double2* s_xy = reinterpret_cast<double2*>(pc_smem);
  float* s_z = reinterpret_cast<float*>(pc_smem + kPcCap * 16);
  uint32_t* s_rgb = reinterpret_cast<uint32_t*>(pc_smem + kPcCap * 20);
  float* s_lut = reinterpret_cast<float*>(pc_smem + kPcCap * 24);
  __shared__ int warp_cnt[2][kPcThreads / 32];
  __shared__ double red_v[2][kPcThreads / 32];
  __shared__ int red_i[2][kPcThreads / 32];
  __shared__ int s_sel;
  __shared__ double sel_pt[6];
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int HW = p.H * p.W;
  double* pts = scratch + static_cast<long long>(b) * HW * 7;
  double* mind_g = pts + static_cast<long long>(HW) * 6;
  const float* dep = depth + static_cast<long long>(b) * HW;
  const uint8_t* col = rgb + static_cast<long long>(b) * HW * 3;
  if (tid < 256) s_lut[tid] = __fdiv_rn(static_cast<float>(tid), 255.0f);
  // ---- back-projection + clip + bounding box, stable compaction in pixel order (one barrier per 512 pixels)
  const float half_h = 0.5f * static_cast<float>(p.H), half_w = 0.5f * static_cast<float>(p.W);
  int base = 0;                                     // kept points so far (the same value in every thread)
  for (int start = 0, r = 0; start < HW; start += kPcThreads, r ^= 1) {
    const int pix = start + tid;
    bool keep = false;
    double x = 0, y = 0, z = 0;
    float dz = 0.f;
    if (pix < HW) {
      const int i = pix / p.W, j = pix - i * p.W;
      dz = dep[pix];
      z = static_cast<double>(dz);
      // ((float(j) - 0.5*W) / focal) * depth : float32 grid, float64 from the division on (VisionUtils.py:60-71)
      x = __dmul_rn(__ddiv_rn(static_cast<double>(static_cast<float>(j) - half_w), p.focal), z);
      y = __dmul_rn(__ddiv_rn(static_cast<double>(static_cast<float>(i) - half_h), p.focal), z);
      keep = (p.near_clip < z) && (z < p.far_clip);
      if (p.has_min) keep = keep && x > p.mn[0] && y > p.mn[1] && z > p.mn[2];
      if (p.has_max) keep = keep && x < p.mx[0] && y < p.mx[1] && z < p.mx[2];
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_cnt[r][warp] = __popc(bal);
    __syncthreads();
    int off = base, tot = 0;
#pragma unroll
    for (int w2 = 0; w2 < kPcThreads / 32; ++w2) {
      const int c = warp_cnt[r][w2];
      if (w2 < warp) off += c;
      tot += c;
    }
    base += tot;
    if (keep) {
      const int at = off + __popc(bal & ((1u << lane) - 1u));
      const uint32_t c0 = col[pix * 3 + 0], c1 = col[pix * 3 + 1], c2 = col[pix * 3 + 2];
      double* q = pts + static_cast<long long>(at) * 6;
      q[0] = x; q[1] = y; q[2] = z;
      q[3] = static_cast<double>(__fdiv_rn(static_cast<float>(c0), 255.0f));
      q[4] = static_cast<double>(__fdiv_rn(static_cast<float>(c1), 255.0f));
      q[5] = static_cast<double>(__fdiv_rn(static_cast<float>(c2), 255.0f));
      if (at < kPcCap) {
        s_xy[at] = make_double2(x, y);
        s_z[at] = dz;
        s_rgb[at] = c0 | (c1 << 8) | (c2 << 16);
      }
    }
  }
  const int N = base;
  if (tid == 0 && count != nullptr) count[b] = N;
  int* idx_out = indices + static_cast<long long>(b) * p.K;
  for (int k = tid; k < p.K; k += kPcThreads) idx_out[k] = (k == 0 && N > 0) ? 0 : -1;
  const int kmax = p.K < N ? p.K : N;
  __syncthreads();
  // ---- farthest point sampling on the 6-D points (the reference passes xyz and rgb columns)
  if (N <= kPcCap) {
    // Points past N never win: their running minimum starts at -inf and fmin keeps it there whatever the (stale)
    // shared-memory contents are, so the sweep needs no per-point branch and the chains of different points interleave.
    double mind[kPcPpt];
    float cr[kPcPpt], cg[kPcPpt], cb[kPcPpt];
    const int jmax = (N + kPcThreads - 1) / kPcThreads;
#pragma unroll
    for (int j = 0; j < kPcPpt; ++j) {
      const int i = j * kPcThreads + tid;
      mind[j] = (i < N) ? INFINITY : -INFINITY;
      cr[j] = cg[j] = cb[j] = 0.f;
      if (i < N) {
        const uint32_t u = s_rgb[i];
        cr[j] = s_lut[u & 255u]; cg[j] = s_lut[(u >> 8) & 255u]; cb[j] = s_lut[(u >> 16) & 255u];
      } else {
        s_xy[i] = make_double2(0.0, 0.0);
        s_z[i] = 0.f;
      }
    }
    __syncthreads();
    int sel = 0, par = 0;
    for (int k = 1; k < kmax; ++k, par ^= 1) {
      const double2 sxy = s_xy[sel];
      const double sz = static_cast<double>(s_z[sel]);
      const uint32_t su = s_rgb[sel];
      const double sr = static_cast<double>(s_lut[su & 255u]), sg = static_cast<double>(s_lut[(su >> 8) & 255u]),
                   sb = static_cast<double>(s_lut[(su >> 16) & 255u]);
      // two (value, index) accumulators (even / odd j), each over ascending indices with a strict >, merged with the
      // full rule (larger value, then smaller index) = the first maximum of the thread's points
      double best = -1.0, best1 = -1.0; int best_i = 0x7fffffff, best1_i = 0x7fffffff;
#pragma unroll
      for (int j = 0; j < kPcPpt; j += kPcGroup) {
        if (j < jmax) {                                       // block-uniform; a group of points per branch
          constexpr int kW = kPcGroup;
          double m2[kW];
#pragma unroll
          for (int u = 0; u < kW; ++u) {
            if (j + u < kPcPpt) {
              const int i = (j + u) * kPcThreads + tid;
              const double2 q = s_xy[i];
              double df = __dsub_rn(q.x, sxy.x);
              double d2 = __dmul_rn(df, df);                      // 0.0 + v == v for v >= +0
              df = __dsub_rn(q.y, sxy.y);                          d2 = __dadd_rn(d2, __dmul_rn(df, df));
              df = __dsub_rn(static_cast<double>(s_z[i]), sz);    d2 = __dadd_rn(d2, __dmul_rn(df, df));
              df = __dsub_rn(pc_widen(cr[j + u]), sr);             d2 = __dadd_rn(d2, __dmul_rn(df, df));
              df = __dsub_rn(pc_widen(cg[j + u]), sg);             d2 = __dadd_rn(d2, __dmul_rn(df, df));
              df = __dsub_rn(pc_widen(cb[j + u]), sb);             d2 = __dadd_rn(d2, __dmul_rn(df, df));
              m2[u] = fmin(mind[j + u], d2);
              mind[j + u] = m2[u];
            }
          }
#pragma unroll
          for (int u = 0; u < kW; u += 2) {
            if (j + u < kPcPpt) {
              if (m2[u] > best) { best = m2[u]; best_i = (j + u) * kPcThreads + tid; }
              if (m2[u + 1] > best1) { best1 = m2[u + 1]; best1_i = (j + u + 1) * kPcThreads + tid; }
            }
          }
        }
      }
      if (best1 > best || (best1 == best && best1_i < best_i)) { best = best1; best_i = best1_i; }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
      }
      if (lane == 0) { red_v[par][warp] = best; red_i[par][warp] = best_i; }
      __syncthreads();
      best = (lane < kPcThreads / 32) ? red_v[par][lane] : -2.0;
      best_i = (lane < kPcThreads / 32) ? red_i[par][lane] : 0x7fffffff;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
      }
      sel = __shfl_sync(0xffffffffu, best_i, 0);
      if (tid == 0) idx_out[k] = sel;
    }
  } else {
    for (int i = tid; i < N; i += kPcThreads) mind_g[i] = INFINITY;
    if (tid == 0) s_sel = 0;
    __syncthreads();
    for (int k = 1; k < kmax; ++k) {
      if (tid < 6) sel_pt[tid] = pts[static_cast<long long>(s_sel) * 6 + tid];
      __syncthreads();
      double best = -1.0; int best_i = 0x7fffffff;
      for (int i = tid; i < N; i += kPcThreads) {
        const double* q = pts + static_cast<long long>(i) * 6;
        double d2 = 0.0;
#pragma unroll
        for (int d = 0; d < 6; ++d) {
          const double df = __dsub_rn(q[d], sel_pt[d]);
          d2 = __dadd_rn(d2, __dmul_rn(df, df));
        }
        const double m = fmin(mind_g[i], d2);
        mind_g[i] = m;
        if (m > best) { best = m; best_i = i; }              // ascending i per thread: strict > keeps the first
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
        if (ov > best || (ov == best && oi < best_i)) { best = ov; best_i = oi; }
      }
      if (lane == 0) { red_v[0][warp] = best; red_i[0][warp] = best_i; }
      __syncthreads();
      if (tid == 0) {
        double bv = red_v[0][0]; int bi = red_i[0][0];
        for (int w2 = 1; w2 < kPcThreads / 32; ++w2)
          if (red_v[0][w2] > bv || (red_v[0][w2] == bv && red_i[0][w2] < bi)) { bv = red_v[0][w2]; bi = red_i[0][w2]; }
        s_sel = bi;
        idx_out[k] = bi;
      }
      __syncthreads();
    }
  }
  __syncthreads();
  // ---- gather (-1 picks the last point, as numpy indexing does), normalise in fp64, cast to float32
  for (int e = tid; e < p.K * 6; e += kPcThreads) {
    const int k = e / 6, d = e - k * 6;
    float v = 0.f;
    if (N > 0) {
      int src = idx_out[k];
      if (src < 0) src = N - 1;
      double a = pts[static_cast<long long>(src) * 6 + d];
      if (p.norm_type == 0) a = __ddiv_rn(__dsub_rn(a, p.p0[d]), p.p1[d]);
      else if (p.norm_type == 1) a = __dadd_rn(__dmul_rn(p.p1[d], __dsub_rn(a, p.p0[d])), p.out_min);
      v = static_cast<float>(a);
    }
    out[(static_cast<long long>(b) * p.K + k) * 6 + d] = v;
  }
}

}  // namespace rmb

// one launch path for both entry points; the dynamic shared-memory opt-in is set once per device (it allocates nothing)
static int pc_launch(const float* d_depth, const uint8_t* d_rgb, const rmb::PcParams& p, double* scratch, float* d_out,
                     int32_t* d_count, int* idx, int batch, void* stream_) {
  constexpr int kMaxDevices = 64;
  static bool attr_set[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  RMB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail("rmb_pointcloud_from_depth: device ordinal out of range");
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!attr_set[dev]) {
      RMB_CUDA(cudaFuncSetAttribute(rmb::pointcloud_from_depth_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rmb::kPcSmemBytes));
      attr_set[dev] = true;
    }
  }
  rmb::pointcloud_from_depth_kernel<<<batch, rmb::kPcThreads, rmb::kPcSmemBytes, static_cast<cudaStream_t>(stream_)>>>(
      d_depth, d_rgb, p, scratch, d_out, d_count, idx);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

extern "C" {

int64_t rmb_pointcloud_scratch_bytes(int batch, int H, int W, int num_points) {
  if (batch < 1 || H < 1 || W < 1 || num_points < 1) return fail("rmb_pointcloud_scratch_bytes: bad sizes");
  return static_cast<int64_t>(batch) * H * W * 7 * sizeof(double) + static_cast<int64_t>(batch) * num_points * sizeof(int);
}

int rmb_pointcloud_from_depth(const float* d_depth, const uint8_t* d_rgb, const rmb_pointcloud_params_t* prm, void* d_scratch,
                              float* d_out, int32_t* d_count, int32_t* d_indices, int batch, void* stream_) {
  if (!d_depth || !d_rgb || !prm || !d_scratch || !d_out) return fail("rmb_pointcloud_from_depth: null argument");
  if (batch < 1 || prm->H < 1 || prm->W < 1 || prm->num_points < 1) return fail("rmb_pointcloud_from_depth: bad sizes");
  if (prm->norm_type < -1 || prm->norm_type > 1) return fail("rmb_pointcloud_from_depth: norm_type must be -1, 0 or 1");
  if (!(prm->focal_scaling > 0.0)) return fail("rmb_pointcloud_from_depth: focal_scaling must be positive");
  rmb::PcParams p;
  p.H = prm->H; p.W = prm->W; p.K = prm->num_points;
  p.focal = prm->focal_scaling; p.near_clip = prm->near_clip; p.far_clip = prm->far_clip;
  p.has_min = prm->has_min_bound; p.has_max = prm->has_max_bound;
  for (int i = 0; i < 3; ++i) { p.mn[i] = prm->min_bound[i]; p.mx[i] = prm->max_bound[i]; }
  p.norm_type = prm->norm_type; p.out_min = prm->norm_out_min;
  for (int i = 0; i < 6; ++i) { p.p0[i] = prm->norm_p0[i]; p.p1[i] = prm->norm_p1[i]; }
  double* scratch = static_cast<double*>(d_scratch);
  int* idx = d_indices ? d_indices : reinterpret_cast<int*>(scratch + static_cast<long long>(batch) * p.H * p.W * 7);
  return pc_launch(d_depth, d_rgb, p, scratch, d_out, d_count, idx, batch, stream_);
}


// ---- rmb_pointcloud_from_depth_static: the same kernel with its scratch in a file-scope device array, for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the
// reference. The entry point allocates nothing (no cudaMalloc / cudaFree /
// allocating library routine). Every GPU that loads the library holds the array (32 MiB) until exit.
// Every call shares the array and it keeps the last call's contents: the kernel itself (re)initialises whatever it reads
// (compacted points are written before they are read, the FPS distances are set to +inf, the index list is rewritten), so
// nothing has to start at zero -- but two calls that overlap on the GPU (different streams, or graph branches) RACE.
#define RMB_PC_STATIC_SCRATCH_BYTES (32ll << 20)
static __device__ __align__(256) unsigned char g_pc_static_scratch[RMB_PC_STATIC_SCRATCH_BYTES];

// In host code `g_pc_static_scratch` is a host-side stand-in, not a device address: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int pc_static_scratch_ptr(void** out) {
  constexpr int kMaxDevices = 64;
  static void* ptrs[kMaxDevices] = {};
  static std::mutex mu;
  int dev = 0;
  RMB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail("rmb_pointcloud_from_depth_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* ptr = nullptr;
    RMB_CUDA(cudaGetSymbolAddress(&ptr, g_pc_static_scratch));
    if (ptr == nullptr) return fail("rmb_pointcloud_from_depth_static: cudaGetSymbolAddress returned a null address");
    ptrs[dev] = ptr;
  }
  *out = ptrs[dev];
  return 0;
}

int rmb_pointcloud_from_depth_static(const float* d_depth, const uint8_t* d_rgb, const rmb_pointcloud_params_t* prm,
                                     float* d_out, int32_t* d_count, int32_t* d_indices, int batch, void* stream_) {
  if (!d_depth || !d_rgb || !prm || !d_out) return fail("rmb_pointcloud_from_depth_static: null argument");
  if (batch < 1 || prm->H < 1 || prm->W < 1 || prm->num_points < 1) return fail("rmb_pointcloud_from_depth_static: bad sizes");
  if (prm->norm_type < -1 || prm->norm_type > 1) return fail("rmb_pointcloud_from_depth_static: norm_type must be -1, 0 or 1");
  if (!(prm->focal_scaling > 0.0)) return fail("rmb_pointcloud_from_depth_static: focal_scaling must be positive");
  const long long need = static_cast<long long>(batch) * prm->H * prm->W * 7 * static_cast<long long>(sizeof(double)) +
                         static_cast<long long>(batch) * prm->num_points * static_cast<long long>(sizeof(int));
  if (need > RMB_PC_STATIC_SCRATCH_BYTES)
    return fail("rmb_pointcloud_from_depth_static: batch * H * W * 56 + batch * num_points * 4 = " + std::to_string(need) +
                " bytes of scratch exceed the static maximum of " + std::to_string(RMB_PC_STATIC_SCRATCH_BYTES) +
                " bytes (use rmb_pointcloud_from_depth with caller-provided scratch)");
  void* scratch_v = nullptr;
  if (int r = pc_static_scratch_ptr(&scratch_v)) return r;
  rmb::PcParams p;
  p.H = prm->H; p.W = prm->W; p.K = prm->num_points;
  p.focal = prm->focal_scaling; p.near_clip = prm->near_clip; p.far_clip = prm->far_clip;
  p.has_min = prm->has_min_bound; p.has_max = prm->has_max_bound;
  for (int i = 0; i < 3; ++i) { p.mn[i] = prm->min_bound[i]; p.mx[i] = prm->max_bound[i]; }
  p.norm_type = prm->norm_type; p.out_min = prm->norm_out_min;
  for (int i = 0; i < 6; ++i) { p.p0[i] = prm->norm_p0[i]; p.p1[i] = prm->norm_p1[i]; }
  double* scratch = static_cast<double*>(scratch_v);
  int* idx = d_indices ? d_indices : reinterpret_cast<int*>(scratch + static_cast<long long>(batch) * p.H * p.W * 7);
  return pc_launch(d_depth, d_rgb, p, scratch, d_out, d_count, idx, batch, stream_);
}

}  // extern "C"
