// Internal declarations shared between the .cu translation units of libb200bench.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/b200bench.h"

namespace b200 {
int make_tmap_2d_bf16(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_cols);
int device_sm_count();
// Device word mixed into every dropout seed (advanced once per micro-step, graph-replay safe).
const uint32_t* step_seed_ptr();
}  // namespace b200
