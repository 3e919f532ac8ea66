// Host-side interface of the peer-memory expert-parallel combine (ep_combine.cu).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace glm4moe {

constexpr int kEpMaxRanks = 8;     // one NVSwitch box
constexpr int kEpMaxTopK = 8;
constexpr int kEpThreads = 128;

struct EpCombineArgs {
  unsigned char* ws[kEpMaxRanks];   // workspace base of every rank as mapped in THIS process (own = local)
  size_t off_y, off_slot_row, off_idx, off_ctl;   // identical workspace layout on every rank
  size_t off_stage;                 // [T_max, H] staging rows peers store their token slices into (sliced mode)
  int sliced;                       // 1: reduce own token slice, store it to every peer (ep_combine.cu)
  int rank, world;
  int tokens, k, H;
  int experts_per_rank;             // E / world (contiguous expert blocks)
  int with_shared;                  // 1: shared-expert rows T*k + t exist on the rank owning token t's slice
  int shared_per_rank;              // ceil(T / world): token slice size of the shared expert
};

// control block inside the workspace (int32): [0] epoch, [1] blocks finished, [16 + r] ready flag
// written by rank r, [32 + r] done flag written by rank r
constexpr size_t kEpCtlBytes = 256;

int ep_combine_p2p_launch(const EpCombineArgs& args, void* out, int dtype, void* stream);

}  // namespace glm4moe
