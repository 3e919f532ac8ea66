// Host-side interface of the tcgen05 grouped GEMM (umma_gemm.cu).
#pragma once
#include <stdint.h>

namespace glm4moe {

struct UmmaLaunch {
  const void* tm_w;        // CUtensorMap (128 bytes) over the stacked weights
  const void* tm_x;        // CUtensorMap over the activation rows, box rows == tile_rows
  const int32_t* tiles;    // int4 {expert, row_start, nrows, 0}
  const int32_t* num_tiles;
  int dense_rows;          // tiles == nullptr: rows [0, dense_rows) in tile_rows chunks, expert 0
  void* out;
  int ld_out;
  const float* row_weight;
  int n_routed_rows;
  int n_feat_tiles;        // ceil(output features / 128)
  int n_feat_valid;        // plain mode: features >= this are not stored
  int n_kblocks;           // K / 64
  int w_rows_per_expert;
  int up_row_offset;
  int tile_rows;           // BN: 16 / 32 / 64 / 128 / 256
  int mode;                // 0 down (* routing weight), 1 gate_up (SiLU*mul), 2 plain (router logits)
  int dtype;
  int grid;
  int32_t* sched;          // 2 zero-initialised ints owned by this launch site (self re-arming)
};

// 2-D row-major [rows, cols] two-byte tensor, box {64 cols, box_rows}, 128B swizzle.
int umma_make_tensor_map(void* map_out, const void* base, int dtype, uint64_t rows, uint64_t cols,
                         uint32_t box_rows);
int umma_launch(const UmmaLaunch& args, void* stream);
bool umma_shape_ok(int H, int I, int dtype);

}  // namespace glm4moe
