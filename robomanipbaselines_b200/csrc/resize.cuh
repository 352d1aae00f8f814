// cv2.resize(..., interpolation=INTER_LINEAR) on the device: the reference shrinks every camera frame
// on the host before the policy sees it
//   RolloutDiffusionPolicy.get_images      policy/diffusion_policy/RolloutDiffusionPolicy.py:117  (uint8 rgb)
//   RolloutDiffusionPolicy3d.get_pointcloud policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:134-137 (uint8 rgb, float32 depth)
// uint8: OpenCV's fixed-point bilinear path restated exactly -- source coordinate (float)((d + 0.5) * scale - 0.5),
// coefficients cvRound(w * 2048) as int16, horizontal pass in int32, vertical pass
// (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2 -- bit-exact with cv2 4.13 for the down-scaling the
// reference does (tests compare against cv2 itself). float32 (depth): plain bilinear with float coefficients
// (cv2's float path differs in the last bits of its coefficients: tolerance 1e-4, written in the test).
// The uint8 kernel can also emit the policy's float32 CHW input directly: (v * (1/255)) * scale + offset
// (RolloutBase.image_transforms, common/base/RolloutBase.py:353, then `image * 2.0 - 1.0`, RolloutDiffusionPolicy.py:123).
// Included at the end of act_engine.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rmb {

// source index and fractional weight of destination index d (OpenCV resize.cpp, INTER_LINEAR, no FMA contraction)
__device__ __forceinline__ void resize_coord(int d, double scale, int src_size, int* s0, float* f) {
  float fx = static_cast<float>(__dadd_rn(__dmul_rn(static_cast<double>(d) + 0.5, scale), -0.5));
  int s = static_cast<int>(floorf(fx));
  fx = __fsub_rn(fx, static_cast<float>(s));
  if (s < 0) { fx = 0.f; s = 0; }
  if (s >= src_size - 1) { fx = 0.f; s = src_size - 1; }
  *s0 = s; *f = fx;
}

// out_mode 0: uint8 HWC [n, H2, W2, 3]; 1: float32 CHW [n, 3, H2, W2] = (v * (1/255)) * scale + offset
__global__ void __launch_bounds__(256) resize_u8_bilinear_kernel(const uint8_t* __restrict__ src, int n, int H, int W, int H2,
                                                                 int W2, double scale_y, double scale_x, int out_mode,
                                                                 float oscale, float ooffset, void* __restrict__ dst) {
  const long long total = static_cast<long long>(n) * H2 * W2;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int dx = static_cast<int>(idx % W2);
    const long long t = idx / W2;
    const int dy = static_cast<int>(t % H2);
    const long long img = t / H2;
    int sx, sy; float fx, fy;
    resize_coord(dx, scale_x, W, &sx, &fx);
    resize_coord(dy, scale_y, H, &sy, &fy);
    const int a0 = __float2int_rn(__fmul_rn(__fsub_rn(1.0f, fx), 2048.0f)), a1 = __float2int_rn(__fmul_rn(fx, 2048.0f));
    const int b0 = __float2int_rn(__fmul_rn(__fsub_rn(1.0f, fy), 2048.0f)), b1 = __float2int_rn(__fmul_rn(fy, 2048.0f));
    const int sx1 = min(sx + 1, W - 1), sy1 = min(sy + 1, H - 1);
    const uint8_t* r0 = src + (img * H + sy) * static_cast<long long>(W) * 3;
    const uint8_t* r1 = src + (img * H + sy1) * static_cast<long long>(W) * 3;
    uint8_t o[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const int h0 = r0[sx * 3 + c] * a0 + r0[sx1 * 3 + c] * a1;
      const int h1 = r1[sx * 3 + c] * a0 + r1[sx1 * 3 + c] * a1;
      o[c] = static_cast<uint8_t>((((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2);
    }
    if (out_mode == 0) {
      uint8_t* q = static_cast<uint8_t*>(dst) + idx * 3;
      q[0] = o[0]; q[1] = o[1]; q[2] = o[2];
    } else {
      float* q = static_cast<float*>(dst) + (img * 3 * H2 + dy) * static_cast<long long>(W2) + dx;
      const long long plane = static_cast<long long>(H2) * W2;
#pragma unroll
      for (int c = 0; c < 3; ++c)
        q[c * plane] = __fadd_rn(__fmul_rn(__fmul_rn(static_cast<float>(o[c]), 1.0f / 255.0f), oscale), ooffset);
    }
  }
}

__global__ void __launch_bounds__(256) resize_f32_bilinear_kernel(const float* __restrict__ src, int n, int H, int W, int H2,
                                                                  int W2, double scale_y, double scale_x,
                                                                  float* __restrict__ dst) {
  const long long total = static_cast<long long>(n) * H2 * W2;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int dx = static_cast<int>(idx % W2);
    const long long t = idx / W2;
    const int dy = static_cast<int>(t % H2);
    const long long img = t / H2;
    int sx, sy; float fx, fy;
    resize_coord(dx, scale_x, W, &sx, &fx);
    resize_coord(dy, scale_y, H, &sy, &fy);
    const int sx1 = min(sx + 1, W - 1), sy1 = min(sy + 1, H - 1);
    const float* r0 = src + (img * H + sy) * static_cast<long long>(W);
    const float* r1 = src + (img * H + sy1) * static_cast<long long>(W);
    const float a0 = 1.0f - fx, b0 = 1.0f - fy;
    const float h0 = __fadd_rn(__fmul_rn(r0[sx], a0), __fmul_rn(r0[sx1], fx));
    const float h1 = __fadd_rn(__fmul_rn(r1[sx], a0), __fmul_rn(r1[sx1], fx));
    dst[idx] = __fadd_rn(__fmul_rn(h0, b0), __fmul_rn(h1, fy));
  }
}

}  // namespace rmb

extern "C" {

int rmb_resize_u8_bilinear(const uint8_t* d_src, int n, int H, int W, int H2, int W2, int out_mode, float scale,
                           float offset, void* d_dst, void* stream_) {
  if (!d_src || !d_dst) return fail("rmb_resize_u8_bilinear: null argument");
  if (n < 1 || H < 1 || W < 1 || H2 < 1 || W2 < 1) return fail("rmb_resize_u8_bilinear: bad sizes");
  if (out_mode != 0 && out_mode != 1) return fail("rmb_resize_u8_bilinear: out_mode must be 0 (uint8 HWC) or 1 (float32 CHW)");
  const long long total = static_cast<long long>(n) * H2 * W2;
  rmb::resize_u8_bilinear_kernel<<<grid_for(total, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_src, n, H, W, H2, W2, static_cast<double>(H) / H2, static_cast<double>(W) / W2, out_mode, scale, offset, d_dst);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

int rmb_resize_f32_bilinear(const float* d_src, int n, int H, int W, int H2, int W2, float* d_dst, void* stream_) {
  if (!d_src || !d_dst) return fail("rmb_resize_f32_bilinear: null argument");
  if (n < 1 || H < 1 || W < 1 || H2 < 1 || W2 < 1) return fail("rmb_resize_f32_bilinear: bad sizes");
  const long long total = static_cast<long long>(n) * H2 * W2;
  rmb::resize_f32_bilinear_kernel<<<grid_for(total, 256, 148), 256, 0, static_cast<cudaStream_t>(stream_)>>>(
      d_src, n, H, W, H2, W2, static_cast<double>(H) / H2, static_cast<double>(W) / W2, d_dst);
  RMB_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
