// Host-side interface of the peer-memory expert-parallel combine (ep_combine.cu).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace glm4moe {

constexpr int kEpMaxRanks = 8;     // one NVSwitch box
constexpr int kEpMaxTopK = 8;
constexpr int kEpThreads = 128;

// What a cross-GPU wait that gives up leaves behind.  The record sits in pinned, device-mapped HOST memory (allocated
// once per process by glm4moe_p2p_alloc): the wait still ends in a trap -- the other blocks of the grid cannot be
// unwound -- and a trap poisons the CUDA context, but host memory stays readable, so glm4moe_ep_last_error can tell
// the operator which rank waited for what.  `spin_limit`: SM clocks a wait may spin (glm4moe_ep_set_timeout_ms).
struct EpErrorRecord {
  int32_t code;        // 0 none, kEpErrTimeout, kEpErrModeMismatch (first failure wins)
  int32_t rank;        // the rank that gave up
  int32_t expected;    // call number it was waiting for
  int32_t observed;    // flag word it last read (call * 4 + mode)
  int32_t site;        // kEpSite*
  int32_t peer;        // index of the flag within its group (the rank being waited for), -1 unknown
  int32_t pad[10];
};
struct EpWatch {
  EpErrorRecord* rec;  // device-visible address of the host record, nullptr = none
  long long spin_limit;
  int rank;
  int site;
};
constexpr int kEpSiteCombine = 1, kEpSiteAllReduce = 2;
EpWatch ep_watch(int rank, int site);     // api.cu: current record and limit

struct EpCombineArgs {
  EpWatch watch;
  unsigned char* ws[kEpMaxRanks];   // workspace base of every rank as mapped in THIS process (own = local)
  size_t off_y, off_slot_row, off_idx, off_ctl;   // identical workspace layout on every rank
  size_t off_stage;                 // [T_max, H] staging rows peers store their token slices into (sliced mode)
  int sliced;                       // 1: reduce own token slice, store it to every peer (ep_combine.cu)
  int rank, world;
  int tokens, k, H;
  int experts_per_rank;             // E / world (contiguous expert blocks)
  int with_shared;                  // 1: shared-expert rows T*k + t exist on the rank owning token t's slice
  int shared_per_rank;              // ceil(T / world): token slice size of the shared expert
  // decode fast path (T <= 32, fused expert kernel): the rows and slot_row of call e live in the
  // epoch-parity buffers below, the producer kernel itself has sent the ready flags, and there is no
  // done round (see ep_combine.cu)
  int fast;
  size_t off_y_alt[2], off_slot_row_alt[2];
};

// control block inside the workspace (int32): [0] epoch (calls completed), [1] blocks finished,
// [2] error code of a timed-out / mismatched wait, [16 + r] ready flag written by rank r,
// [32 + r] done flag written by rank r.  Flag value = epoch * 4 + mode (kEpMode*).
constexpr size_t kEpCtlBytes = 256;
constexpr int kEpModePull = 0, kEpModeSliced = 1, kEpModeFast = 2;
constexpr int kEpErrTimeout = 1, kEpErrModeMismatch = 2;

// What the fused decode kernel needs to publish "my rows of call e are complete" to every rank itself.
struct EpFastSignal {
  int enabled;
  int world, rank;
  int32_t* ctl;                      // own control block (epoch)
  int32_t* peer_ready[kEpMaxRanks];  // rank r's ctl + 16 + my rank, as mapped in this process
  void* y_alt[2];                    // epoch-parity row buffers, [32 * (k + 1), H] each
};

// route_small.cu, expert-parallel decode fast path: slot_row goes to the buffer of the call's parity
struct RouteSmallEp {
  const int32_t* epoch;             // own ctl[0]; nullptr = off
  int32_t* slot_row_alt[2];
  int act_pow2;                     // e4m3 tier: token scales rounded up to a power of two (not an EP matter; rides here)
};
int route_small_ep(const void* h, const void* gate_w, const float* bias, void* logits, int T, int H, int E, int k,
                   float scale, int dtype, int e_lo, int e_hi, int tile_rows, int shared_lo, int shared_hi,
                   int32_t* idx, float* w, int32_t* expert_offsets, int32_t* row_token, float* row_weight,
                   int32_t* slot_row, int32_t* tiles, int32_t tile_capacity, int32_t* num_tiles, int32_t* ticket,
                   void* xq, float* x_scale_t, int x_scale_ld, const RouteSmallEp& ep, void* stream);

int ep_combine_p2p_launch(const EpCombineArgs& args, void* out, int dtype, void* stream);

}  // namespace glm4moe
