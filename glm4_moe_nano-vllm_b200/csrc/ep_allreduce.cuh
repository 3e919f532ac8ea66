// Host-side interface of the fused all-reduce + residual add + RMSNorm over peer memory (ep_allreduce.cu).
#pragma once
#include <stddef.h>
#include <stdint.h>

#include "ep_combine.cuh"

namespace glm4moe {

// Communication buffer of one rank (one per process, shared by every layer; identical layout on every rank):
//   ctl      1 KB of int32: [0] epoch (calls completed), [1] [3] [4] block tickets, [2] error code,
//            [16 + r] two-shot: "my partial rows are in your inbox" flag written by rank r,
//            [32 + r] two-shot: "the sums of my token slice are in your inbox" flag written by rank r
//   os[p][src]   one-shot inbox of call parity p: rank src's whole partial [T, H]   (T*H*es <= os_slot)
//   rs[p][src]   two-shot reduce-scatter inbox: rank src's partial rows of MY token slice [per_cap, H]
//   ag[p]        two-shot all-gather inbox: y = the all-reduced rows of the other ranks' slices, [max_tokens, H]
//   rowflag[src][row]   one-shot: epoch*4 + mode of the last call in which rank src pushed `row`
struct EpCommLayout {
  size_t ctl, os[2], os_slot, rs[2], rs_slot, ag[2], rowflag, total;
};
constexpr size_t kEpArOneShotSlot = 1u << 20;     // one-shot up to 1 MiB of partial per rank (T = 128 at H = 4096, 2 B)
constexpr int kEpArOneShotRows = 1024;            // ... and up to this many rows (one flag word per row and source)
constexpr int kEpArModeOne = 0, kEpArModeTwo = 1;
constexpr int kEpArThreads = 256;

EpCommLayout ep_comm_layout(int max_tokens, int H, int dtype, int world);

struct EpArArgs {
  EpWatch watch;
  unsigned char* comm[kEpMaxRanks];   // comm buffer of every rank as mapped in THIS process (own = local)
  EpCommLayout L;
  int rank, world, tokens, H;
  int per;                            // ceil(tokens / world): token slice of a rank (two-shot)
  float eps;
  const void* partial;                // [T, H] this rank's partial sums (e.g. o_proj of its head shard)
  const void* residual;               // [T, H] or nullptr
  const void* weight;                 // [H] RMSNorm weight, or nullptr: plain all-reduce
  void* out;                          // [T, H]
  void* residual_out;                 // [T, H] (residual != nullptr)
};

int ep_allreduce_launch(const EpArArgs& a, int dtype, int two_shot, void* stream);

}  // namespace glm4moe
