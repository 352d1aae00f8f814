// Per-episode rollout state of the diffusion policies, kept on the device for B episodes at once. Replaces the Python
// lists of the reference's rollout classes:
//   * observation ring buffers   -- RolloutDiffusionPolicy.get_state / get_images
//                                   (robo_manip_baselines/policy/diffusion_policy/RolloutDiffusionPolicy.py:87-143),
//                                   RolloutDiffusionPolicy3d.get_state / get_pointcloud
//                                   (robo_manip_baselines/policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:103-171):
//                                   first call fills all n_obs_steps slots, later calls pop(0) + append
//   * the float64 action FIFO    -- infer_policy (RolloutDiffusionPolicy.py:66-85, RolloutDiffusionPolicy3d.py:83-101):
//                                   list(action.astype(float64)), one pop(0) + denormalize_data per policy step
// plus the row gather that compacts the episodes whose FIFO is empty into a dense sub-batch for predict_action.
// All three are pure data movement (HBM-bound); the only arithmetic is denormalize_data in fp64 (numpy op order, no FMA).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "elementwise.cuh"

namespace rmb {

// mode[b]: 0 = episode b does not call the policy at this step (buffer untouched), 1 = first call since reset (every
// slot <- new), 2 = later call (slot k <- slot k+1, last slot <- new). One thread owns one element position of one
// episode and walks the To slots in order, so the in-place shift needs no scratch.
template <typename V>
__global__ void obs_ring_push_kernel(V* __restrict__ buf, const V* __restrict__ new_obs, const uint8_t* __restrict__ mode,
                                     int To, long long E) {
  const int b = blockIdx.y;
  const uint8_t m = mode[b];
  if (m == 0) return;
  V* bb = buf + static_cast<long long>(b) * To * E;
  const V* nb = new_obs + static_cast<long long>(b) * E;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < E;
       e += static_cast<long long>(gridDim.x) * blockDim.x) {
    const V v = nb[e];
    if (m == 1) {
      for (int k = 0; k < To; ++k) bb[k * E + e] = v;
    } else {
      for (int k = 0; k + 1 < To; ++k) bb[k * E + e] = bb[(k + 1) * E + e];
      bb[(To - 1) * E + e] = v;
    }
  }
}

// dst row i <- src row idx[i]; rows are `row_vec` 16-byte vectors (or 4-byte words when the row is not 16-byte sized)
template <typename V>
__global__ void gather_rows_kernel(const V* __restrict__ src, const int32_t* __restrict__ idx, V* __restrict__ dst,
                                   long long row_vec) {
  const int i = blockIdx.y;
  const V* s = src + static_cast<long long>(idx[i]) * row_vec;
  V* d = dst + static_cast<long long>(i) * row_vec;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < row_vec;
       e += static_cast<long long>(gridDim.x) * blockDim.x)
    d[e] = s[e];
}

// One thread per episode. sub_row[b] >= 0: this episode called the policy; its FIFO is refilled from row sub_row[b] of
// the predicted actions (float32 -> float64, RolloutDiffusionPolicy.py:75-77) and its read position reset. Then
// pop(0) + denormalize_data (:80-82). An episode whose FIFO is empty and was not refilled keeps its last action.
__global__ void action_fifo_step_kernel(double* __restrict__ fifo, int32_t* __restrict__ pos,
                                        const float* __restrict__ act_new, long long act_row_stride,
                                        const int32_t* __restrict__ sub_row, DenormParams dn,
                                        double* __restrict__ action_out, double* __restrict__ action_norm_out, int B,
                                        int Na, int Da) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double* f = fifo + static_cast<long long>(b) * Na * Da;
  int p = pos[b];
  const int r = (sub_row != nullptr) ? sub_row[b] : -1;
  if (r >= 0) {
    const float* a = act_new + static_cast<long long>(r) * act_row_stride;
    for (int i = 0; i < Na * Da; ++i) f[i] = static_cast<double>(a[i]);
    p = 0;
  }
  if (p >= Na) return;
  for (int d = 0; d < Da; ++d) {
    const double x = f[p * Da + d];
    double o;
    if (dn.type == 0) o = __dadd_rn(__dmul_rn(dn.a[d], x), dn.b[d]);
    else o = __dadd_rn(__dmul_rn(dn.a[d], __dsub_rn(x, dn.out_min)), dn.b[d]);
    if (action_norm_out != nullptr) action_norm_out[b * Da + d] = x;
    action_out[b * Da + d] = o;
  }
  pos[b] = p + 1;
}

}  // namespace rmb

extern "C" {

int rmb_obs_ring_push(void* d_buf, const void* d_new, const uint8_t* d_mode, int batch, int n_obs_steps,
                      int64_t row_bytes, void* stream_) {
  if (!d_buf || !d_new || !d_mode) return fail("rmb_obs_ring_push: null argument");
  if (batch < 1 || n_obs_steps < 1 || row_bytes < 1 || row_bytes % 4 != 0) return fail("rmb_obs_ring_push: bad sizes (row_bytes must be a multiple of 4)");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const bool v16 = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(d_buf) % 16 == 0) && (reinterpret_cast<uintptr_t>(d_new) % 16 == 0);
  const long long E = v16 ? row_bytes / 16 : row_bytes / 4;
  const int bx = static_cast<int>(std::min<long long>((E + 255) / 256, 64));
  if (v16) rmb::obs_ring_push_kernel<uint4><<<dim3(bx, batch), 256, 0, s>>>(static_cast<uint4*>(d_buf), static_cast<const uint4*>(d_new), d_mode, n_obs_steps, E);
  else rmb::obs_ring_push_kernel<uint32_t><<<dim3(bx, batch), 256, 0, s>>>(static_cast<uint32_t*>(d_buf), static_cast<const uint32_t*>(d_new), d_mode, n_obs_steps, E);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_gather_rows(const void* d_src, const int32_t* d_idx, void* d_dst, int n_rows, int64_t row_bytes, void* stream_) {
  if (!d_src || !d_idx || !d_dst) return fail("rmb_gather_rows: null argument");
  if (n_rows < 1 || row_bytes < 1 || row_bytes % 4 != 0) return fail("rmb_gather_rows: bad sizes (row_bytes must be a multiple of 4)");
  cudaStream_t s = static_cast<cudaStream_t>(stream_);
  const bool v16 = row_bytes % 16 == 0 && (reinterpret_cast<uintptr_t>(d_src) % 16 == 0) && (reinterpret_cast<uintptr_t>(d_dst) % 16 == 0);
  const long long E = v16 ? row_bytes / 16 : row_bytes / 4;
  const int bx = static_cast<int>(std::min<long long>((E + 255) / 256, 64));
  if (v16) rmb::gather_rows_kernel<uint4><<<dim3(bx, n_rows), 256, 0, s>>>(static_cast<const uint4*>(d_src), d_idx, static_cast<uint4*>(d_dst), E);
  else rmb::gather_rows_kernel<uint32_t><<<dim3(bx, n_rows), 256, 0, s>>>(static_cast<const uint32_t*>(d_src), d_idx, static_cast<uint32_t*>(d_dst), E);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_action_fifo_step(double* d_fifo, int32_t* d_pos, const float* d_act_new, int64_t act_row_stride,
                         const int32_t* d_sub_row, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                         double* d_action, double* d_action_norm, int batch, int n_action_steps, int Da, void* stream_) {
  if (!d_fifo || !d_pos || !d_p0 || !d_p1 || !d_action) return fail("rmb_action_fifo_step: null argument");
  if (d_sub_row && !d_act_new) return fail("rmb_action_fifo_step: sub_row given without predicted actions");
  if (norm_type != 0 && norm_type != 1) return fail("rmb_action_fifo_step: norm_type must be 0 or 1");
  if (batch < 1 || n_action_steps < 1 || Da < 1) return fail("rmb_action_fifo_step: bad sizes");
  rmb::DenormParams dn{d_p0, d_p1, out_min, norm_type};
  rmb::action_fifo_step_kernel<<<(batch + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_fifo, d_pos, d_act_new, act_row_stride, d_sub_row, dn, d_action, d_action_norm, batch, n_action_steps, Da);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
