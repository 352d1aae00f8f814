// Kernel launch helper: every kernel on the policy-step path is launched with programmatic stream serialization
// (PDL), so its prologue (barrier init, TMEM allocation, tensor-map prefetch, CTA scheduling) overlaps the tail of its
// predecessor; the kernels call pdl_wait() before they touch global memory (ptx.cuh). RMB_NO_PDL=1 launches normally.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>

#include <utility>

namespace rmb {

inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("RMB_NO_PDL"); return !(e != nullptr && e[0] == '1'); }();
  return on;
}

// PDL is applied inside a PdlScope only: the ACT policy step, where every launch in the stream is one of the kernels
// that call pdl_wait(). The diffusion engines share the GEMM launchers but interleave kernels that do not, and their
// CUDA-graph replay measured 2.7 % slower with programmatic edges, so they launch normally.
inline bool& pdl_scope_flag() {
  static thread_local bool on = false;
  return on;
}
struct PdlScope {
  bool prev;
  PdlScope() : prev(pdl_scope_flag()) { pdl_scope_flag() = true; }
  ~PdlScope() { pdl_scope_flag() = prev; }
};

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = (pdl_enabled() && pdl_scope_flag()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace rmb
