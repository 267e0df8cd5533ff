// wavefront.cuh — placeholder until the wavefront pipeline lands (next commit).
#pragma once
#include <cuda_runtime.h>
#include "../../include/vpt_cuda.h"
#include "shaders.cuh"
#include "upload.cuh"
namespace vpt {
struct wavefront_buffers {
  bool live = false;
  void release() {}
};
inline cudaError_t run_wavefront(const flat_scene&, wavefront_buffers&, const DParams&, f4*,
    rng_t*, int, int*, vpt_trace_stats*) {
  return cudaErrorNotSupported;
}
}  // namespace vpt
