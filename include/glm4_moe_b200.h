/*
 * glm4_moe_b200.h — C ABI of the B200-native GLM4-MoE sparse-expert block.
 *
 * Drop-in boundary for the reference's MoE layer (a Python nn.Module; the reference has no FFI
 * of its own).  Each entry point names the reference code it replaces.
 * All pointers are raw device pointers unless the name says `_host`; sizes are element counts;
 * `stream` is a `cudaStream_t` passed as `void*`.  Device memory: the forward / kernel entry points allocate
 * nothing (the caller passes workspace, staging and scratch).  Two set-up entries do allocate:
 * `glm4moe_p2p_alloc` (that is its job; `glm4moe_p2p_free` releases it; the first call of a process also pins a
 * 64-byte host record for glm4moe_ep_last_error) and `glm4moe_layer_create` with
 * weight_format = GLM4MOE_W_FP8_BLOCK128_POW2 (two packed UE8M0 copies of the block scales, 1 byte per
 * 128 x 128 weight block, owned by the layer and released by `glm4moe_layer_destroy`, plus a 4-byte flag freed
 * before it returns).
 * Synchronisation: `glm4moe_layer_forward_host` waits for its stream; `glm4moe_layer_create` synchronises the
 * device twice (around zeroing the workspace counters and the power-of-two scale check); nothing else does.
 * Return value: 0 = ok, <0 = argument / support error (GLM4MOE_ERR_*), >0 = `cudaError_t`.
 * ABI history: 1 = round 1; 2 = weight-format enum + FP8 tiers, debug knobs, p2p / prenorm / all-reduce entries,
 * `glm4moe_permute_static`, H <= GLM4MOE_EP_COMM_MAX_HIDDEN enforced by `glm4moe_ep_comm_bytes`,
 * `glm4moe_ep_last_error` / `glm4moe_ep_set_timeout_ms`.
 *
 * Row space used by permute / expert GEMMs / combine:
 *   rows [0, T*k)        routed rows, expert-major, token-ascending inside an expert
 *                        (the order the reference loop visits them, moe.py:135-143)
 *   rows [T*k, T*k + T)  the always-on shared expert, token order (moe.py:165)
 */
#ifndef GLM4_MOE_B200_H
#define GLM4_MOE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GLM4MOE_ABI_VERSION 2

enum { GLM4MOE_F32 = 0, GLM4MOE_F16 = 1, GLM4MOE_BF16 = 2 };          /* model dtype */
/* expert weight format.  FP8_BLOCK128: e4m3 bytes + one fp32 scale per 128 x 128 block, activations e4m3 with
 * scale = amax / 448 per token per 128 columns.  FP8_BLOCK128_POW2: the same layout, but every weight block scale
 * must be a power of two (checked at layer_create) and activation scales are rounded UP to a power of two (UE8M0):
 * the scale format the tcgen05 block-scaled MMA applies in hardware, used by the prefill kernel (umma_fp8_mx.cu). */
enum { GLM4MOE_W_NATIVE = 0, GLM4MOE_W_FP8_BLOCK128 = 1, GLM4MOE_W_FP8_BLOCK128_POW2 = 2 };
enum { GLM4MOE_IMPL_AUTO = 0, GLM4MOE_IMPL_SIMT = 1, GLM4MOE_IMPL_UMMA = 2 };

#define GLM4MOE_OK 0
#define GLM4MOE_ERR_ARG (-1)          /* bad pointer / size / alignment */
#define GLM4MOE_ERR_UNSUPPORTED (-2)  /* shape or dtype outside what the kernels cover */
#define GLM4MOE_ERR_WORKSPACE (-3)    /* workspace too small */
#define GLM4MOE_ERR_DEVICE (-4)       /* not an sm_100 device / driver entry point missing */
#define GLM4MOE_ERR_EP_TIMEOUT (-5)   /* glm4moe_ep_last_error: a cross-GPU wait gave up, a peer never signalled */
#define GLM4MOE_ERR_EP_MODE (-6)      /* glm4moe_ep_last_error: the ranks disagree on the form of the call */

typedef struct glm4moe_config {
  int32_t hidden_size;            /* H   (Glm4MoeConfig.hidden_size) */
  int32_t n_routed_experts;       /* E   */
  int32_t num_experts_per_tok;    /* k   */
  int32_t moe_intermediate_size;  /* I   (also the shared expert's width, moe.py:65-69) */
  float routed_scaling_factor;    /* moe.py:24,46 */
  int32_t dtype;                  /* GLM4MOE_F32 / F16 / BF16: activations, router, outputs */
  int32_t weight_format;          /* GLM4MOE_W_* for the expert stacks */
  int32_t max_tokens;             /* largest T any forward will see (workspace sizing) */
  int32_t expert_begin;           /* expert-parallel: this rank owns experts [begin, end) */
  int32_t expert_end;             /*   (0, E) on a single GPU                              */
  int32_t with_shared;            /* 1: shared expert on all tokens (moe.py:165,171); 0: routed partial only;
                                     2: shared expert on this rank's 1/R slice of the tokens (expert-parallel) */
  int32_t impl;                   /* GLM4MOE_IMPL_* (AUTO picks tcgen05 when the shape allows) */
} glm4moe_config_t;

typedef struct glm4moe_layer glm4moe_layer_t;   /* opaque */

int glm4moe_abi_version(void);
/* Process-wide tuning / test knobs (parity tests compare the paths against each other).
 * GLM4MOE_KNOB_DIRECT: 1 (default) = decode batches T <= 32 use the fused front-end and read the
 * input tokens in place; 0 = always the general router / permute / gather sequence.
 * GLM4MOE_KNOB_FUSED: 1 (default) = inside direct mode, gate_up and down run as one persistent kernel
 * (decode_fused.cu); 0 = two grouped-GEMM launches.
 * GLM4MOE_KNOB_FP8_PREFILL: 1 (default) = the e4m3 tier uses token-major 128 x 256 tiles once an expert
 * averages >= 48 rows (umma_fp8_prefill.cu); 0 = always the 16..64-token swap-AB tiles (umma_fp8.cu).
 * GLM4MOE_KNOB_BK32: 1 = K blocks of 32 elements (64-byte swizzle, 7-deep ring) for the 256-token bf16 prefill tiles
 * (default 0: measured slower).  GLM4MOE_KNOB_DOWN2: 1 (default) = the prefill down projection runs 256 features x
 * 256 tokens per CTA tile.  GLM4MOE_KNOB_FORCE_TILE_ROWS: 0 (default) = token rows per tile follow the mean rows per
 * expert; 16/32/64/128/256 force one tile size (parity tests walk every tile geometry with it). */
enum { GLM4MOE_KNOB_DIRECT = 0, GLM4MOE_KNOB_FUSED = 1, GLM4MOE_KNOB_FP8_PREFILL = 2,
       GLM4MOE_KNOB_ROUTER_ROWS = 3 /* token rows per tensor-core router tile: 0 = by T, 16/32/64/128 */,
       GLM4MOE_KNOB_BK32 = 4, GLM4MOE_KNOB_DOWN2 = 5, GLM4MOE_KNOB_FORCE_TILE_ROWS = 6,
       GLM4MOE_KNOB_EP_FAST = 7 /* 1 (default): decode-size expert-parallel calls use ONE cross-rank rendezvous (rows in
                                   epoch-parity buffers, ready flags from the expert kernel, no done round); must be
                                   the same on every rank (a mismatch traps with an error code instead of hanging) */,
       GLM4MOE_KNOB_MERGE_FRONT = 9 /* 1 (default): at T <= 2 (bf16 / fp16, single rank) the router front-end runs
                                       inside the persistent decode kernel: ONE launch per layer; 0 = route_small +
                                       decode_fused as for larger decode batches */,
       GLM4MOE_KNOB_TAIL_BOX = 10 /* 1 (default): a ragged token tile of the tensor-core grouped GEMM is fetched
                                     through a TMA box just large enough for its rows; 0 = always a whole tile box */,
       GLM4MOE_KNOB_SPEC_PREFETCH = 12 /* decode (T <= 32): number of local experts whose weights the persistent kernel
                                          pulls towards L2 while the router still runs, i.e. before the routing is
                                          known; 0 (default) = off, -1 = as many as fit ~70 MB once an expert is hit
                                          with probability >= 0.8 (T >= ~24 at k/E = 1/16).  Results unchanged; measured
                                          neutral to 2 % slower (the decode stream leaves no idle HBM time to fill) */,
       GLM4MOE_KNOB_EP_AR_MODE = 11 /* fused all-reduce + residual add + RMSNorm over peer memory
                                       (glm4moe_ep_allreduce_add_rmsnorm): -1 (default) one-shot or two-shot by size,
                                       0 one-shot whenever the batch fits its inbox, 1 always two-shot (reduce-scatter ->
                                       norm on the token shard -> all-gather); must be the same on every rank */,
       GLM4MOE_KNOB_FP8_MX = 8 /* FP8_BLOCK128_POW2 layers at prefill: 1 (default) = hardware block scaling, one CTA
                                  per tile (gate_up token-major umma_fp8_mx_tm.cu, down feature-major
                                  umma_fp8_mx.cu); 2 = hardware block scaling on CTA pairs, 256-token x 256-row tiles
                                  (umma_fp8_mx2.cu, tcgen05 cta_group::2); 3 = feature-major for both GEMMs;
                                  0 = the software-scaled kernel on the same operands */ };
int glm4moe_debug_set(int knob, int value);
const char* glm4moe_error_string(int code);

/* ---- router: replaces GlmMoeSelectTopk.forward (moe.py:28-50) ------------------------------ */
/* logits[T,E] = h[T,H] @ gate_w[E,H]^T, fp32 accumulate, rounded to the model dtype (moe.py:32) */
int glm4moe_router_logits(const void* h, const void* gate_w, void* logits, int T, int H, int E,
                          int dtype, void* stream);
/* sigmoid -> +bias(fp32) -> top-k (higher score, then lower expert id) -> gather un-biased
 * scores -> renormalise (+1e-20) -> * scale, with the model-dtype roundings of moe.py:37-46.
 * idx[T,k] int32, slots sorted by ascending expert id; w[T,k] fp32 holding model-dtype values. */
int glm4moe_router_select(const void* logits, const float* bias, int T, int E, int k, float scale,
                          int dtype, int32_t* idx, float* w, void* stream);

/* Decode front-end, T <= GLM4MOE_ROUTE_SMALL_MAX_TOKENS: glm4moe_router_logits + glm4moe_router_select +
 * glm4moe_permute in ONE launch with bit-identical results (the last block to finish its logits runs
 * the selection and the permutation).  `ticket`: 40 zero-initialised int32 owned by the call site
 * (left at zero).  Replaces moe.py:28-50 and the bookkeeping of moe.py:135-144 at decode sizes.
 * FP8 tier: xq != NULL additionally writes the e4m3 copy of the tokens, xq[T,H], with per-token 1x128
 * scales x_scale_t[H/128, x_scale_ld] (scale = amax/448), the token tile of the fused expert kernel.  */
#define GLM4MOE_ROUTE_SMALL_MAX_TOKENS 64
int glm4moe_route_small(const void* h, const void* gate_w, const float* bias, void* logits, int T, int H,
                        int E, int k, float scale, int dtype, int e_lo, int e_hi, int tile_rows,
                        int shared_lo, int shared_hi, int32_t* idx, float* w, int32_t* expert_offsets,
                        int32_t* row_token, float* row_weight, int32_t* slot_row, int32_t* tiles,
                        int32_t tile_capacity, int32_t* num_tiles, int32_t* ticket, void* xq,
                        float* x_scale_t, int x_scale_ld, void* stream);

/* ---- permutation: replaces the mask / where / gather loop (moe.py:135-147) ------------------ */
/* Stable expert-major order of the (token, slot) pairs whose expert is in [e_lo, e_hi).
 *   expert_offsets[E+1]  rows of expert e are [offsets[e], offsets[e+1])
 *   row_token[T*k]       token feeding row r          row_weight[T*k]  its routing weight
 *   slot_row[T*k]        row of (t, slot), -1 if that expert is not local
 *   tiles[cap] int4      {expert - e_lo, row_start, nrows, 0} work list in `tile_rows` chunks
 *                        (index into the local weight stack), the shared expert's tiles last
 *                        (stack index e_hi - e_lo, tokens [shared_lo, shared_hi)); *num_tiles = its length
 * scratch: int32[(ceil(T/256)+1) * E + 8].                                                     */
int glm4moe_permute(const int32_t* idx, const float* w, int T, int k, int E, int e_lo, int e_hi,
                    int tile_rows, int shared_lo, int shared_hi, int32_t* expert_offsets, int32_t* row_token,
                    float* row_weight, int32_t* slot_row, int32_t* tiles, int32_t tile_capacity,
                    int32_t* num_tiles, int32_t* scratch, void* stream);
size_t glm4moe_permute_scratch_elems(int T, int E);
int32_t glm4moe_tile_capacity(int T, int k, int E, int tile_rows);

/* glm4moe_permute for callers that cannot supply scratch memory.  (No counterpart in the reference: the reference
 * is eager PyTorch and never has to think about where temporaries come from.)
 * Who needs it: code that is handed only the input and output buffers of its op and has no allocator to turn to --
 * a plugin inside a runtime that owns all device memory and forbids cudaMalloc from plugins, a process whose GPU is
 * already fully pledged (KV cache, other tenants) when this library is loaded, a call site where an allocation
 * would stall or deadlock the device.  Same arguments and same results as glm4moe_permute, minus `scratch`: the
 * per-block count table lives in a fixed array inside the library's device image instead (about 65 KB, present on
 * every GPU that loads the library until the process exits).
 *   - Limit: T <= GLM4MOE_PERMUTE_STATIC_MAX_TOKENS (and E <= 256 as everywhere).  A larger T is refused with
 *     GLM4MOE_ERR_WORKSPACE before anything is launched; use glm4moe_permute with your own scratch for those.
 *   - Allocates nothing: no cudaMalloc / cudaFree, no allocating library call.
 *   - The first call on each GPU looks up that GPU's address of the array (with the calling thread's current device,
 *     which must be the device `stream` and all buffers belong to).  That look-up can make the CUDA runtime load
 *     the library's module onto the GPU, so make the first call per GPU outside any stream capture.
 *   - All calls on a GPU share the one array, from every stream and every thread.  Calls that run one after the
 *     other (same stream, or ordered by events) are safe -- nothing in the array has to be cleared between them,
 *     each call overwrites what it later reads.  Two calls with T > 256 that may overlap on the GPU race on it and
 *     give wrong permutations: order them, or give the second one its own scratch through glm4moe_permute.
 *     (T <= 256 runs in a single block and never touches the array.)                                          */
#define GLM4MOE_PERMUTE_STATIC_MAX_TOKENS 16384
int glm4moe_permute_static(const int32_t* idx, const float* w, int T, int k, int E, int e_lo, int e_hi,
                           int tile_rows, int shared_lo, int shared_hi, int32_t* expert_offsets, int32_t* row_token,
                           float* row_weight, int32_t* slot_row, int32_t* tiles, int32_t tile_capacity,
                           int32_t* num_tiles, void* stream);

/* x_rows[r,:] = x[row_token[r],:] for routed rows, x_rows[T*k + t,:] = x[t,:] for t in
 * [shared_lo, shared_hi)  (moe.py:147) */
int glm4moe_gather_rows(const void* x, const int32_t* row_token, const int32_t* expert_offsets,
                        void* x_rows, int T, int k, int E, int H, int shared_lo, int shared_hi, int dtype,
                        void* stream);

/* ---- expert MLP: replaces Glm4MoeMLP.forward over every hit expert + the shared expert ------ */
/* act[r, :I] = silu(x_rows[r] @ Wg^T) * (x_rows[r] @ Wu^T)   (mlp.py:50-51, activation.py:13-14)
 * w_gate_up: stacked [E+1, 2I, H], rows [0,I) gate, [I,2I) up (mlp.py:118); index E = shared.  */
int glm4moe_expert_gate_up_simt(const void* x_rows, const void* w_gate_up, const int32_t* tiles,
                                const int32_t* num_tiles, void* act, int H, int I, int tile_rows,
                                int dtype, void* stream);
/* y[r, :H] = (act[r] @ Wd^T) * row_weight[r]   (mlp.py:52, moe.py:149); shared rows weight 1 */
int glm4moe_expert_down_simt(const void* act, const void* w_down, const int32_t* tiles,
                             const int32_t* num_tiles, const float* row_weight, void* y, int H,
                             int I, int T, int k, int tile_rows, int dtype, void* stream);

/* Dense SiLU-gated MLP on the tensor cores: replaces Glm4MoeMLP.forward (mlp.py:49-54) for ONE weight pair, e.g.
 * the dense layer-0 MLP of the decoder stack (decode_layer.py:31-36; Air: I = 10944) or a single expert.
 * x [T, H], w_gate_up [2I, H] (rows [0,I) gate, [I,2I) up), w_down [H, I], out [T, H], all in `dtype` (bf16 / fp16);
 * act_ws: [T, I] scratch; sched: 4 int32 zero-initialised once by the caller (the kernels re-arm them).
 * H % 128 == 0, I % 8 == 0 (I need not be a multiple of 128 or 64).  GLM4MOE_ERR_UNSUPPORTED otherwise: the
 * caller keeps the exact CUDA-core entries above for fp32 / odd shapes. */
int glm4moe_mlp_forward(const void* x, const void* w_gate_up, const void* w_down, void* act_ws, void* out, int T,
                        int H, int I, int dtype, int32_t* sched, void* stream);

/* ---- combine: replaces index_add_ + shared add (moe.py:151,171) ----------------------------- */
/* out[t] = sum_{slot, ascending expert} y[slot_row[t,slot]] (+ y[T*k+t] for t in
 * [shared_lo, shared_hi)); fp32 accumulate.                                                     */
int glm4moe_combine(const void* y, const int32_t* slot_row, void* out, int T, int k, int H,
                    int shared_lo, int shared_hi, int dtype, void* stream);

/* ---- whole layer: replaces Glm4MoeMoE.forward (moe.py:128-176) ------------------------------ */
size_t glm4moe_workspace_bytes(const glm4moe_config_t* cfg);
/* w_gate_up [E_local+1, 2I, H], w_down [E_local+1, H, I] hold experts [begin,end) then the
 * shared expert.  s_* are the fp32 128x128 block scales of the FP8 format (NULL otherwise).     */
int glm4moe_layer_create(const glm4moe_config_t* cfg, const void* gate_w, const float* bias,
                         const void* w_gate_up, const void* w_down, const float* s_gate_up,
                         const float* s_down, void* workspace, size_t workspace_bytes,
                         glm4moe_layer_t** out);
int glm4moe_layer_destroy(glm4moe_layer_t* layer);
int glm4moe_layer_forward(glm4moe_layer_t* layer, const void* x, void* out, int T, void* stream);
/* As glm4moe_layer_forward, with the shared expert applied to tokens [shared_lo, shared_hi) only
 * (expert-parallel dispatch: the owner's own tokens come first in the batch).                  */
int glm4moe_layer_forward_ex(glm4moe_layer_t* layer, const void* x, void* out, int T, int shared_lo,
                             int shared_hi, void* stream);
/* Same launches, with a CUDA event recorded on `stream` before every stage; after the last event
 * completes, stage_ms[0..6] = {router_logits, router_select, permute, gather, gate_up, down,
 * combine} durations and stage_ms[7] = the whole sequence (bench.py's per-kernel roofline).     */
int glm4moe_layer_forward_timed(glm4moe_layer_t* layer, const void* x, void* out, int T, void* stream,
                                float* stage_ms);
/* Host-buffer entry: H2D of x_host (pinned or pageable), forward, D2H into out_host, stream sync. */
int glm4moe_layer_forward_host(glm4moe_layer_t* layer, const void* x_host, void* out_host, int T,
                               void* x_dev_staging, void* out_dev_staging, void* stream);
/* Debug views into the workspace after a forward (parity triage; mirrors the reference's
 * experts_output / shared_output dump idea, moe.py:160-172).
 * which: 0 idx[T,k] i32, 1 w[T,k] f32, 2 expert_offsets[E+1] i32, 3 row_token[T*k] i32,
 *        4 slot_row[T*k] i32, 5 logits[T,E] model dtype, 6 act rows, 7 y rows, 8 num_tiles i32,
 *        9 tiles int4, 10 x_rows, 11 row_weight f32.                                           */
int glm4moe_layer_view(glm4moe_layer_t* layer, int which, void** ptr, size_t* bytes);
/* number of kernels the last forward launched, and which expert-GEMM implementation it used */
int glm4moe_layer_last_launches(const glm4moe_layer_t* layer);
int glm4moe_layer_last_impl(const glm4moe_layer_t* layer);
/* 1 when the last forward ran gate_up + down + combine as the single persistent decode kernel */
int glm4moe_layer_last_fused(const glm4moe_layer_t* layer);
int glm4moe_layer_tile_rows(const glm4moe_layer_t* layer, int T);

/* ---- the step in front of the block: residual add + RMSNorm (layers/layernorm.py:29-40, called as
 * post_attention_layernorm at models/glm4_moe/decode_layer.py:98) --------------------------------
 * x32 = x + residual (fp32); residual_out = round(x32); out = round(round(x32 * rsqrt(mean(x32^2) + eps))
 * * weight).  residual == NULL: plain RMSNorm of x (layernorm.py:17-26), residual_out unused.
 * out may alias x; residual_out may alias residual.                                               */
int glm4moe_add_rmsnorm(const void* x, const void* residual, const void* weight, float eps, int T, int H,
                        int dtype, void* out, void* residual_out, void* stream);

/* ---- expert parallelism over NVLink peer memory (replaces the dead all_reduce of linear.py:153-154
 * for the sparse block; the reference itself has no expert parallelism) -----------------
 * Every rank allocates its layer workspace with glm4moe_p2p_alloc (cudaMalloc + cudaIpc handle, zeroed),
 * ships the 64-byte handle to its peers over any host channel, and maps theirs with glm4moe_p2p_open.
 * glm4moe_layer_forward_ep_p2p then runs the local experts of a replicated token batch and ONE kernel
 * that, per token, reads the k weighted expert rows (+ shared-expert row) from the ranks that own them
 * and sums them in ascending expert order in fp32: bit-identical to the single-GPU forward, result
 * replicated on every rank, no NCCL call.  peer_workspaces[r] = rank r's workspace as mapped in this
 * process (peer_workspaces[rank] = the layer's own workspace).  The layer must have been created with
 * with_shared = 2 (world > 1) and contiguous expert blocks; all ranks call it the same number of times.
 * Large batches (T*(k+1)*H*elt >= 16 MiB) use the sliced form of the same kernel: every rank reduces
 * only its ceil(T/world) token slice and stores the finished rows into `out` and into a staging region
 * of every peer's workspace (reduce-scatter + all-gather in one launch; same sums, same bits).
 * glm4moe_set_ep_sliced: -1 choose by size (default), 0 never, 1 always; process-wide, must be set
 * identically on every rank. */
int glm4moe_p2p_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int glm4moe_p2p_open(const unsigned char* handle64, void** peer_ptr);
int glm4moe_p2p_close(void* peer_ptr);
int glm4moe_p2p_free(void* ptr);
int glm4moe_layer_forward_ep_p2p(glm4moe_layer_t* layer, const void* x, void* out, int T,
                                 void* const* peer_workspaces, int rank, int world, void* stream);
int glm4moe_set_ep_sliced(int mode);

/* When a cross-GPU wait gives up.  Every wait inside the peer-memory kernels (glm4moe_layer_forward_ep_p2p*,
 * glm4moe_ep_allreduce_add_rmsnorm) is bounded: a rank that never arrives, or arrives in another form of the call,
 * must not hang the GPU.  The kernel cannot unwind the rest of its grid, so it still ends in a trap, and a trap
 * makes every later CUDA call of the process fail (cudaErrorLaunchFailed).  Before trapping it writes what it was
 * waiting for into a record in pinned HOST memory (allocated by the first glm4moe_p2p_alloc of the process), and
 * these two entries make no CUDA call, so they still work afterwards:
 * glm4moe_ep_last_error: 0 = no failure recorded, else GLM4MOE_ERR_EP_TIMEOUT / GLM4MOE_ERR_EP_MODE; info6 (may be
 *   NULL) = {raw code, rank that gave up, call number it waited for, flag word it last saw (call * 4 + mode),
 *   site (1 = expert combine, 2 = all-reduce + norm), peer rank it was waiting for}; clear != 0 resets the record.
 * glm4moe_ep_set_timeout_ms: how long a wait may spin; ms <= 0 restores the default (2^35 SM clocks, about 18 s).
 *   Process-wide, read at every launch; env GLM4MOE_EP_TIMEOUT_MS sets the initial value.  Raise it when a peer may
 *   legitimately be late (first-call module load, a debugger); lower it in tests. */
int glm4moe_ep_last_error(int32_t* info6, int clear);
int glm4moe_ep_set_timeout_ms(int ms);


/* ---- the norm in front of the block fused to the block (decode_layer.py:98-99 as one call) --------------------
 * out = Glm4MoeMoE.forward(post_attention_layernorm(x, residual)[0]), residual_out = round(x + residual): the
 * arithmetic of glm4moe_add_rmsnorm followed by glm4moe_layer_forward, one launch chain, the normalised tokens
 * never leave the layer workspace.  residual == NULL: plain RMSNorm (layernorm.py:17-26). */
int glm4moe_layer_forward_prenorm(glm4moe_layer_t* layer, const void* x, const void* residual,
                                  const void* norm_weight, float eps, void* residual_out, void* out, int T,
                                  void* stream);

/* ---- the tensor-parallel boundary in front of the block, fused over NVLink peer memory -----------
 * Replaces `dist.all_reduce(y)` of RowParallelLinear.forward (linear.py:153-154; the attention o_proj) TOGETHER
 * with the residual add + RMSNorm that follows it (layernorm.py:29-40, decode_layer.py:98): every rank passes its
 * PARTIAL sums [T, H]; ONE kernel per rank produces, identically on every rank,
 *   y = round(sum_r partial_r)  (fp32, ascending rank)      residual_out = round(y + residual)
 *   out = round(round((y + residual) * rsqrt(mean((y + residual)^2) + eps)) * weight)
 * weight == NULL (then residual must be NULL too): out = y, a plain all-reduce.  No NCCL call, graph capturable.
 * Small batches are one-shot (partials pushed to every peer, one flag round); large ones run reduce-scatter ->
 * norm on the rank's token shard -> all-gather inside the same launch (csrc/ep_allreduce.cu).
 * peer_comm[r]: rank r's communication buffer as mapped in this process (glm4moe_p2p_alloc / _open, zero-filled,
 * glm4moe_ep_comm_bytes(max_tokens, H, dtype, world) bytes, one per process, shared by all layers); every rank
 * passes the same max_tokens and calls the entry the same number of times with the same T. */
#define GLM4MOE_EP_COMM_MAX_HIDDEN 8192   /* one row = at most 4 16-byte vectors per thread of a 256-thread block */
size_t glm4moe_ep_comm_bytes(int max_tokens, int H, int dtype, int world);
int glm4moe_ep_allreduce_add_rmsnorm(void* const* peer_comm, int rank, int world, int max_tokens,
                                     const void* partial, const void* residual, const void* weight, float eps,
                                     int T, int H, int dtype, void* out, void* residual_out, void* stream);
/* o_proj partial sums -> fused all-reduce + add + norm -> expert-parallel block (glm4moe_layer_forward_ep_p2p):
 * the reference's decode_layer.py:96-99 from the output of the o_proj GEMM to the block's output. */
int glm4moe_layer_forward_ep_p2p_prenorm(glm4moe_layer_t* layer, const void* partial, const void* residual,
                                         const void* norm_weight, float eps, void* residual_out, void* out, int T,
                                         void* const* peer_workspaces, void* const* peer_comm,
                                         int comm_max_tokens, int rank, int world, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GLM4_MOE_B200_H */
