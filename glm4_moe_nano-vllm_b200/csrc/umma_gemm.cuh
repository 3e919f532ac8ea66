// Host-side interface of the tcgen05 grouped GEMM (umma_gemm.cu).
#pragma once
#include <stdint.h>

#include "ep_combine.cuh"

namespace glm4moe {

// Decode "direct" mode of the gate_up GEMM (T <= tile_rows <= 32): the MMA token tile is the layer
// input x itself (tm_x maps x, every token is a column of every expert's MMA), so no gathered copy of
// the token rows exists; the epilogue writes column t of expert e to the expert-major row that
// row_token assigns it.  The shared expert's items do not depend on the routing and run before the
// programmatic dependency wait.  tokens == 0: off.
struct DirectArgs {
  int tokens;                // T
  int shared_expert;         // stack index of the shared expert's weights
  int shared_row0;           // T*k: first shared-expert row of the row space
  int sh_lo, sh_hi;          // tokens the shared expert is applied to
  const int32_t* row_token;  // [T*k] token of every routed row
};

struct UmmaLaunch {
  const void* tm_w;        // CUtensorMap (128 bytes) over the stacked weights
  const void* tm_x;        // CUtensorMap over the activation rows, box rows == tile_rows
  const void* tm_x_tail[3];   // optional: the same rows with smaller boxes (ascending tail_rows, multiples of 16):
  int tail_rows[3];           // a tile whose rows fit a smaller box is fetched through it (ragged expert tails)
  const int32_t* tiles;    // int4 {expert, row_start, nrows, index in expert | tiles of expert << 16}
  const int32_t* num_tiles;
  int dense_rows;          // tiles == nullptr: rows [0, dense_rows) in tile_rows chunks, expert 0
  void* out;
  int ld_out;
  const float* row_weight;
  int n_routed_rows;
  int n_feat_tiles;        // ceil(output features / 128)
  int n_feat_valid;        // plain mode: features >= this are not stored
  int n_kblocks;           // K / bk
  int bk;                  // 0 / 64: 128-byte swizzle atoms; 32: 64-byte atoms, half-size stages (tile_rows 256 only)
  int w_rows_per_expert;
  int up_row_offset;
  int tile_rows;           // BN: 16 / 32 / 64 / 128 / 256
  int mode;                // 0 down (* routing weight), 1 gate_up (SiLU*mul), 2 plain (router logits)
  int dtype;
  int grid;
  int32_t* sched;          // 2 zero-initialised ints owned by this launch site (self re-arming)
  DirectArgs direct;       // gate_up only
};

// 2-D row-major [rows, cols] two-byte tensor, box {64 cols, box_rows}, 128B swizzle.
int umma_make_tensor_map(void* map_out, const void* base, int dtype, uint64_t rows, uint64_t cols,
                         uint32_t box_rows);
// same with box {bk cols, box_rows}: bk = 64 (128B swizzle) or 32 (64B swizzle)
int umma_make_tensor_map_bk(void* map_out, const void* base, int dtype, uint64_t rows, uint64_t cols,
                            uint32_t box_rows, int bk);
int umma_launch(const UmmaLaunch& args, void* stream);
bool umma_shape_ok(int H, int I, int dtype);

// ---- decode regime: gate_up + down in one persistent kernel (decode_fused.cu) ----
// e4m3 tier of the fused kernel: 128x128 fp32 weight block scales, per-token 1x128 activation scales
// (k-block major, [K/128, scale_ld]); the gate_up epilogue quantises its activation tile itself.
struct Fp8Args {
  const float* s_gate_up;    // [n_stack, 2I/128, H/128]
  const float* s_down;       // [n_stack, H/128, I/128]
  const float* sx_t;         // [H/128, scale_ld]: scales of the quantised input tokens (row = token)
  float* sact_t;             // [I/128, scale_ld]: scales of the quantised activation rows
  unsigned char* actq;       // [R, I] e4m3 activation rows
  int scale_ld;
  int pow2;                  // activation scales rounded up to a power of two (GLM4MOE_W_FP8_BLOCK128_POW2)
};

// T <= 2 (bf16 / fp16, single rank): the router front-end runs INSIDE the persistent kernel instead of as its
// own launch.  The epilogue warps of CTA c compute the logits of expert c (the summation order of
// route_small_kernel, bit for bit), the last CTA to finish selects the experts of the <= 2 tokens (the same
// `select_token`) and builds the permutation / work list with one warp, then releases a flag the producers
// acquire before they fetch routed work.  Saves the grid dependency between two launches (~5 us of a 70 us
// layer at T = 1).
struct RouteMerged {
  int enabled;
  const void* x;              // [T, H] layer input
  const void* gate_w;         // [E, H]
  const float* bias;          // [E] fp32
  void* logits;               // [T, E] model dtype (kept: tests and callers read it)
  int32_t* idx;               // [T, k]
  float* w;                   // [T, k]
  int32_t* expert_offsets;    // [E + 1]
  int32_t* row_token;         // the same buffers the kernel reads through its const parameters
  float* row_weight;
  int32_t* slot_row;
  int32_t* tiles;
  int tile_capacity;
  int32_t* num_tiles;
  int E, k, e_lo, e_hi;
  float scale;
  int32_t* sync;              // [0] ticket, [1] flag: zero-initialised, re-armed by the kernel
};

struct DecodeFusedLaunch {
  const void* tm_wgu;        // gate_up weight stack, box {64, 128}
  const void* tm_wd;         // down weight stack, box {64, 128}
  const void* tm_x;          // the layer input [T, H], box {64, tile_rows}
  const void* tm_act;        // activation rows [R, I], box {64, tile_rows}
  const int32_t* tiles;      // one tile per hit expert (rows <= tile_rows), the shared tile last
  const int32_t* num_tiles;
  const int32_t* row_token;
  const float* row_weight;
  void* act;                 // [R, I]
  void* y;                   // [R, H]
  int H, I;
  int tokens;                // T <= tile_rows
  int n_routed_rows;         // T * k
  int shared_expert;         // stack index of the shared expert
  int sh_lo, sh_hi;
  int tile_rows;             // 16 / 32
  int dtype;
  int grid;                  // <= SM count (co-residency is required by the in-kernel dependency)
  int32_t* sched;            // 2 zero-initialised ints (self re-arming)
  int32_t* ready;            // 1 + tile capacity zero-initialised ints (self re-arming)
  const int32_t* slot_row;   // [T, k] row of every (token, slot), -1 = expert not local
  void* out;                 // [T, H] layer output (combine fused into the last down item per feature tile)
  int k;
  int32_t* fdone;            // H/128 zero-initialised ints (self re-arming)
  const void* pf0;           // shared expert's gate_up weights (L2 prefetch while the router runs)
  size_t pf0_bytes;
  const void* pf1;           // shared expert's down weights
  size_t pf1_bytes;
  const void* pf2;           // speculative: gate_up weights of the first local experts (see api.cu, spec_prefetch_experts)
  size_t pf2_bytes;
  const void* pf3;           // speculative: their down weights
  size_t pf3_bytes;
  int fp8;                   // 1: e4m3 weights / activations (tm_* are byte maps), q is set
  Fp8Args q;
  // expert-parallel decode fast path (out == nullptr): rows go to the epoch-parity buffer and the kernel
  // itself publishes "rows of this call complete" to every rank when its last down item has landed
  EpFastSignal ep;
  RouteMerged rt;            // T <= 2: router front-end inside this kernel (no route_small launch)
};
int decode_fused_launch(const DecodeFusedLaunch& args, void* stream);

// ---- FP8 block-scaled tier (umma_fp8.cu) ----
struct UmmaFp8Launch {
  const void* tm_w;          // CUtensorMap over the e4m3 weight stack (uint8), box {128 bytes, 128 rows}
  const void* tm_x;          // CUtensorMap over the e4m3 activation rows, box {128 bytes, tile_rows}
  const float* w_scale;      // [n_stack, w_rows_per_expert/128, K/128]
  const float* x_scale_t;    // [K/128, x_scale_ld]
  int x_scale_ld;
  const int32_t* tiles;
  const int32_t* num_tiles;
  void* out;
  int ld_out;
  const float* row_weight;
  int n_routed_rows;
  int n_feat_tiles;
  int n_kblocks;             // K / 128
  int w_rows_per_expert;
  int up_row_offset;
  int tile_rows;             // 16 / 32 / 64
  int gated;
  int dtype;                 // output / activation model dtype
  int grid;
  int32_t* sched;
};
int umma_make_tensor_map_u8(void* map_out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows);
int umma_fp8_launch(const UmmaFp8Launch& args, void* stream);
// rows -> e4m3 + per-128 fp32 scales (scale_t is k-block major); row_token != nullptr gathers tokens
int quant_rows_fp8(const void* src, int src_ld, const int32_t* row_token, const int32_t* expert_offsets, int E,
                   int T, int k, int K, int sh_lo, int sh_hi, void* dst, float* scale_t, int scale_ld, int dtype,
                   int pow2, void* sf, void* stream);
// layer input -> e4m3 rows, each token quantised once and written to its k expert rows + shared row
int quant_scatter_fp8(const void* x, int K, const int32_t* slot_row, int T, int k, int sh_lo, int sh_hi, void* dst,
                      float* scale_t, int scale_ld, int dtype, int pow2, void* sf, void* stream);
// pow2 = 1: scales rounded up to a power of two; sf (nullable): packed UE8M0 copy [ceil(K/512), scale_ld] x 4 bytes

// ---- FP8 tier, prefill geometry (umma_fp8_prefill.cu): token rows = MMA M (128), weight rows = N (256) ----
struct UmmaFp8PrefillLaunch {
  const void* tm_w;          // e4m3 weight stack (uint8), box {128 bytes, 128 rows}
  const void* tm_x;          // e4m3 activation rows, box {128 bytes, 128 rows}
  const float* w_scale;      // [n_stack, w_rows_per_expert/128, K/128]
  const float* x_scale_t;    // [K/128, scale_ld]
  int scale_ld;
  const int32_t* tiles;      // 128-row tiles (permute with tile_rows = 128)
  const int32_t* num_tiles;
  void* out;                 // down: y rows [R, H] model dtype
  int ld_out;                // down: H; gate_up: I (row pitch of out_q in bytes)
  void* out_q;               // gate_up: e4m3 activation rows [R, I]
  float* out_scale_t;        // gate_up: [I/128, scale_ld]
  const float* row_weight;
  int n_routed_rows;
  int n_feat_tiles;          // gate_up: I/128; down: H/256
  int n_kblocks;             // K / 128
  int w_rows_per_expert;     // gate_up: 2I; down: H
  int up_row_offset;         // gate_up: I
  int gated;
  int dtype;
  int grid;
  int32_t* sched;
  int pow2;                  // gate_up: power-of-two activation scales
  void* out_sf;              // gate_up, nullable: packed UE8M0 copy of out_scale_t
};
bool umma_fp8_prefill_shape_ok(int H, int I);
int umma_fp8_prefill_launch(const UmmaFp8PrefillLaunch& args, void* stream);

// ---- FP8 tier, prefill geometry with hardware block scaling (umma_fp8_mx.cu) ----
struct UmmaFp8MxLaunch {
  const void* tm_w;          // e4m3 weight stack (uint8), box {128 bytes, 128 rows}
  const void* tm_x;          // e4m3 activation rows, box {128 bytes, 128 rows}
  const uint32_t* w_sf;      // [n_stack, w_rows_per_expert/128, ceil(K/512)] packed UE8M0 block scales
  const uint32_t* x_sf;      // [ceil(K/512), scale_ld] packed UE8M0 activation scales
  int scale_ld;
  const int32_t* tiles;      // 128-row tiles
  const int32_t* num_tiles;
  void* out;                 // down: y rows [R, H] model dtype
  int ld_out;
  void* out_q;               // gate_up: e4m3 activation rows [R, I]
  float* out_scale_t;        // gate_up: [I/128, scale_ld] fp32 copy of the scales (decode-size kernels read these)
  void* out_sf;              // gate_up: [ceil(I/512), scale_ld] packed UE8M0
  const float* row_weight;
  int n_routed_rows;
  int n_feat_tiles;          // gate_up: I/128; down: H/256
  int n_kblocks;             // K / 128
  int w_rows_per_expert;
  int up_row_offset;
  int gated;
  int dtype;
  int grid;
  int32_t* sched;
};
int umma_fp8_mx_launch(const UmmaFp8MxLaunch& args, void* stream);
// token-major form (umma_fp8_mx_tm.cu): tokens = MMA M, one N = 256 instruction per K slice; used for gate_up
int umma_fp8_mx_tm_launch(const UmmaFp8MxLaunch& args, void* stream);
// CTA-pair form (umma_fp8_mx2.cu, tcgen05 cta_group::2): 256-token x 256-weight-row tiles, `tiles` = 256-row tiles
int umma_fp8_mx2_launch(const UmmaFp8MxLaunch& args, void* stream);
// fp32 128x128 block scales [n_blocks, nkb] -> packed UE8M0 words [n_blocks, ceil(nkb/4)]; *bad (device int)
// is set when a scale is not a power of two in UE8M0 range
int pack_weight_sf(const float* scales, uint32_t* out, size_t n_blocks, int nkb, int* bad, void* stream);

}  // namespace glm4moe
