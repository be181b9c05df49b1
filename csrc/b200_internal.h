// Internal declarations shared between the .cu translation units of libb200bench.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/b200bench.h"

#include <utility>

namespace b200 {
// <<<>>> with the programmatic-stream-serialization attribute: the kernel may be scheduled while its
// predecessor in the stream is still draining.  Every kernel launched this way calls pdl_wait() first.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols);
int device_sm_count();
// Device word mixed into every dropout seed (advanced once per micro-step, graph-replay safe).
const uint32_t* step_seed_ptr();
}  // namespace b200
