// C-ABI layer object: replaces Glm4MoeMoE.forward (reference moe.py:128-176) with the kernel
// sequence  router logits -> select -> permute (rank, scan, scatter) -> gather -> gate_up ->
// down -> combine, all on the caller's stream, no host sync, no allocation (CUDA-graph safe;
// the reference's decode path replays graphs, model_runner.py:197-213,239-247).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "common.cuh"
#include "umma_gemm.cuh"
#include "ep_combine.cuh"
#include "ep_allreduce.cuh"

using namespace glm4moe;

namespace {

constexpr int kSimtRows = 8;
constexpr int kNumBn = 5;
constexpr int kBnChoices[kNumBn] = {16, 32, 64, 128, 256};
constexpr int kRouterUmmaMinTokens = 64;   // below this the K-split CUDA-core router wins (latency)
constexpr int kDirectMaxTokens = 32;       // decode "direct" mode (fused front-end, no gather): T <= this

int g_fused = -1;    // decode_fused kernel inside direct mode; GLM4MOE_NO_FUSED=1 / glm4moe_debug_set(1, v)
bool fused_enabled() {
  if (g_fused < 0) {
    const char* e = getenv("GLM4MOE_NO_FUSED");
    g_fused = (e && e[0] == '1') ? 0 : 1;
  }
  return g_fused == 1;
}
int g_fp8_prefill = -1;   // token-major FP8 kernel at prefill; GLM4MOE_NO_FP8_PREFILL=1 / glm4moe_debug_set(2, v)
bool fp8_prefill_enabled() {
  if (g_fp8_prefill < 0) {
    const char* e = getenv("GLM4MOE_NO_FP8_PREFILL");
    g_fp8_prefill = (e && e[0] == '1') ? 0 : 1;
  }
  return g_fp8_prefill == 1;
}
int g_direct = -1;   // -1: read GLM4MOE_NO_DIRECT on first use; glm4moe_debug_set(0, v) overrides
bool direct_enabled() {
  if (g_direct < 0) {
    const char* e = getenv("GLM4MOE_NO_DIRECT");   // tuning aid: 1 = always use the general path
    g_direct = (e && e[0] == '1') ? 0 : 1;
  }
  return g_direct == 1;
}

int g_router_rows = 0;   // 0: pick by T; 16/32/64/128: token rows per router tile (glm4moe_debug_set(3, v))
int g_bk32 = -1;         // see bk32_enabled()
int g_down2 = -1;        // see down2_enabled()
int g_tail_box = 1;      // ragged token tiles fetched through smaller TMA boxes (umma_gemm.cu TailMaps)
int g_force_bn = -2;     // see forced_bn_index()
int g_fp8_mx = -1;       // hardware block scaling for FP8_BLOCK128_POW2 prefill: 0 off, 1 one CTA per tile
                         // (gate_up umma_fp8_mx_tm.cu, down umma_fp8_mx.cu), 2 CTA pairs (umma_fp8_mx2.cu),
                         // 3 feature-major for both GEMMs; KNOB_FP8_MX / GLM4MOE_FP8_MX
int fp8_mx_mode() {
  if (g_fp8_mx < 0) {
    const char* e = getenv("GLM4MOE_FP8_MX");
    g_fp8_mx = (e && e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : 1;
  }
  return g_fp8_mx;
}
int g_merge_front = 1;   // T <= 2: router front-end inside the persistent decode kernel (one launch per layer), KNOB_MERGE_FRONT
int g_spec_prefetch = 0;  // decode: local experts whose weights are L2-prefetched before the routing is known (KNOB_SPEC_PREFETCH;
                          // -1 = as many as fit 70 MB once P(hit) >= 0.8).  Measured neutral to -2 % (profiles/r2.md section 11): off
int g_ep_ar_mode = -1;    // fused all-reduce + norm (ep_allreduce.cu): -1 by size, 0 one-shot, 1 two-shot, KNOB_EP_AR_MODE
int g_ep_fast = 1;       // decode fast path of the peer-memory combine (one rendezvous per step), KNOB_EP_FAST

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WsLayout {
  size_t logits, idx, w, offsets, row_token, row_weight, slot_row, num_tiles, sched, ready, fdone, tiles, scratch, x_rows,
      act, y, xq, sx, actq, sact, sx_sf, sact_sf, ctl, stage, y_alt[2], slot_row_alt[2], xn, total;
  int tile_capacity;
};

bool is_fp8(int weight_format) {
  return weight_format == GLM4MOE_W_FP8_BLOCK128 || weight_format == GLM4MOE_W_FP8_BLOCK128_POW2;
}

WsLayout make_layout(const glm4moe_config_t& c) {
  WsLayout L{};
  const size_t T = c.max_tokens, k = c.num_experts_per_tok, E = c.n_routed_experts;
  const size_t H = c.hidden_size, I = c.moe_intermediate_size, es = dtype_size(c.dtype);
  const size_t R = T * (k + 1);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = align_up(off + bytes, 1024);
    return o;
  };
  L.logits = take(T * E * es);
  L.idx = take(T * k * 4);
  L.w = take(T * k * 4);
  L.offsets = take((E + 1) * 4);
  L.row_token = take(T * k * 4);
  L.row_weight = take(T * k * 4);
  L.slot_row = take(T * k * 4);
  L.num_tiles = take(16);
  L.sched = take(256);
  L.fdone = take((H / 128 + 8) * 4);
  L.ready = take((E + 8) * 4);   // decode_fused: gate_up completion counter per tile (<= E + 1 tiles)
  L.tile_capacity = glm4moe_tile_capacity(static_cast<int>(T), static_cast<int>(k), static_cast<int>(E),
                                          kSimtRows);
  L.tiles = take(static_cast<size_t>(L.tile_capacity) * 16);
  L.scratch = take(glm4moe_permute_scratch_elems(static_cast<int>(T), static_cast<int>(E)) * 4);
  L.x_rows = take(R * H * es);
  L.act = take(R * I * es);
  L.y = take(R * H * es);
  L.xq = L.sx = L.actq = L.sact = L.sx_sf = L.sact_sf = 0;
  if (is_fp8(c.weight_format)) {
    L.xq = take(R * H);                 // e4m3 activation rows
    L.sx = take((H / 128) * R * 4);     // their per-128 scales, k-block major [H/128, R]
    L.actq = take(R * I);
    L.sact = take(((I + 127) / 128) * R * 4);
  }
  if (c.weight_format == GLM4MOE_W_FP8_BLOCK128_POW2) {
    // the same scales as packed UE8M0 exponent bytes, 4 k-blocks per word: [ceil(K/512), R] (umma_fp8_mx.cu)
    L.sx_sf = take(((H / 128 + 3) / 4) * R * 4);
    L.sact_sf = take(((I / 128 + 3) / 4) * R * 4);
  }
  L.ctl = take(kEpCtlBytes);   // peer-memory combine: epoch, ticket, ready / done flags (ep_combine.cu)
  // sliced peer-memory combine: the rows peers store their reduced token slices into.  Only layers
  // that can run expert-parallel (with_shared == 2 or a strict subset of the experts) pay for it.
  const bool ep_capable = c.with_shared == 2 || c.expert_end - c.expert_begin < c.n_routed_experts;
  L.stage = take(ep_capable && c.dtype != GLM4MOE_F32 ? T * H * es : 0);
  // decode fast path of the peer-memory combine: weighted rows and slot_row of a call live in the buffer
  // of the call's parity (<= 32 tokens; ep_combine.cu).  Small (2 x 32 * (k + 1) rows), so every layer the
  // peer-memory entry accepts has them (a world of one rank included).
  const bool ep_alt = c.dtype != GLM4MOE_F32;
  for (int i = 0; i < 2; ++i) {
    L.y_alt[i] = take(ep_alt ? static_cast<size_t>(kDirectMaxTokens) * (k + 1) * H * es : 0);
    L.slot_row_alt[i] = take(ep_alt ? static_cast<size_t>(kDirectMaxTokens) * k * 4 : 0);
  }
  // glm4moe_layer_forward_prenorm: the normalised tokens (the block's actual input), [T, H]
  L.xn = take(T * H * es);
  L.total = off;
  return L;
}

}  // namespace

struct glm4moe_layer {
  glm4moe_config_t cfg;
  WsLayout L;
  unsigned char* ws;
  const void* gate_w;
  const float* bias;
  const void* w_gate_up;
  const void* w_down;
  const float* s_gate_up;
  const float* s_down;
  uint32_t* sf_gate_up;   // FP8_BLOCK128_POW2: packed UE8M0 copies of the block scales (owned, cudaMalloc)
  uint32_t* sf_down;
  int sm_count;
  bool umma_ok;
  int last_launches;
  int last_impl;
  int last_fused;
  bool events_ready;
  cudaEvent_t ev[8];
  alignas(64) CUtensorMap tm_w_gu;
  alignas(64) CUtensorMap tm_w_d;
  alignas(64) CUtensorMap tm_gate;
  alignas(64) CUtensorMap tm_x[kNumBn];
  alignas(64) CUtensorMap tm_act[kNumBn];
  // prefill (256-token tiles), half-size pipeline stages: box {32 cols, rows}, 64-byte swizzle
  bool bk32_ok;
  alignas(64) CUtensorMap tm_w_gu32;
  alignas(64) CUtensorMap tm_w_d32;
  alignas(64) CUtensorMap tm_x32;
  alignas(64) CUtensorMap tm_act32;
};

namespace glm4moe {
static thread_local int g_launches = 0;
void note_launch() { ++g_launches; }
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("GLM4MOE_NO_PDL");
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}
}  // namespace glm4moe

extern "C" int glm4moe_debug_set(int knob, int value) {
  switch (knob) {
    case GLM4MOE_KNOB_DIRECT: g_direct = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_FUSED: g_fused = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_FP8_PREFILL: g_fp8_prefill = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_ROUTER_ROWS:
      if (value != 0 && value != 16 && value != 32 && value != 64 && value != 128) return GLM4MOE_ERR_ARG;
      g_router_rows = value;
      return GLM4MOE_OK;
    case GLM4MOE_KNOB_EP_FAST: g_ep_fast = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_SPEC_PREFETCH: g_spec_prefetch = value < 0 ? -1 : value; return GLM4MOE_OK;
    case GLM4MOE_KNOB_EP_AR_MODE: g_ep_ar_mode = value < 0 ? -1 : value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_MERGE_FRONT: g_merge_front = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_FP8_MX: g_fp8_mx = value < 0 || value > 3 ? 1 : value; return GLM4MOE_OK;
    case GLM4MOE_KNOB_BK32: g_bk32 = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_DOWN2: g_down2 = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_TAIL_BOX: g_tail_box = value ? 1 : 0; return GLM4MOE_OK;
    case GLM4MOE_KNOB_FORCE_TILE_ROWS:
      g_force_bn = -1;
      if (value == 0) return GLM4MOE_OK;
      for (int i = 0; i < kNumBn; ++i) if (value == kBnChoices[i]) g_force_bn = i;
      return g_force_bn >= 0 ? GLM4MOE_OK : GLM4MOE_ERR_ARG;
    default: return GLM4MOE_ERR_ARG;
  }
}

extern "C" int glm4moe_abi_version(void) { return GLM4MOE_ABI_VERSION; }

extern "C" const char* glm4moe_error_string(int code) {
  switch (code) {
    case GLM4MOE_OK: return "ok";
    case GLM4MOE_ERR_ARG: return "invalid argument";
    case GLM4MOE_ERR_UNSUPPORTED: return "unsupported shape or dtype";
    case GLM4MOE_ERR_WORKSPACE: return "workspace too small";
    case GLM4MOE_ERR_DEVICE: return "device/driver does not provide the required sm_100 features";
    case GLM4MOE_ERR_EP_TIMEOUT: return "expert-parallel wait timed out: a peer rank never signalled (see glm4moe_ep_last_error)";
    case GLM4MOE_ERR_EP_MODE: return "expert-parallel ranks disagree on the call form (sliced / fast / all-reduce mode set differently)";
    default: return code > 0 ? cudaGetErrorString(static_cast<cudaError_t>(code)) : "unknown error";
  }
}

static int check_cfg(const glm4moe_config_t* c) {
  if (!c) return GLM4MOE_ERR_ARG;
  if (c->hidden_size <= 0 || c->n_routed_experts <= 0 || c->num_experts_per_tok <= 0 ||
      c->moe_intermediate_size <= 0 || c->max_tokens <= 0)
    return GLM4MOE_ERR_ARG;
  if (c->num_experts_per_tok > c->n_routed_experts || c->n_routed_experts > 256) return GLM4MOE_ERR_UNSUPPORTED;
  if (c->dtype != GLM4MOE_F32 && c->dtype != GLM4MOE_F16 && c->dtype != GLM4MOE_BF16) return GLM4MOE_ERR_ARG;
  if (c->expert_begin < 0 || c->expert_end > c->n_routed_experts || c->expert_begin > c->expert_end)
    return GLM4MOE_ERR_ARG;
  if (is_fp8(c->weight_format)) {
    // e4m3 weights with 128x128 fp32 block scales run only on the tensor-core path
    if (c->dtype == GLM4MOE_F32 || c->hidden_size % 128 != 0 || c->moe_intermediate_size % 128 != 0)
      return GLM4MOE_ERR_UNSUPPORTED;
  } else if (c->weight_format != GLM4MOE_W_NATIVE) {
    return GLM4MOE_ERR_UNSUPPORTED;
  }
  return GLM4MOE_OK;
}

extern "C" size_t glm4moe_workspace_bytes(const glm4moe_config_t* cfg) {
  if (check_cfg(cfg) != GLM4MOE_OK) return 0;
  return make_layout(*cfg).total;
}

extern "C" int glm4moe_layer_create(const glm4moe_config_t* cfg, const void* gate_w, const float* bias,
                                    const void* w_gate_up, const void* w_down, const float* s_gate_up,
                                    const float* s_down, void* workspace, size_t workspace_bytes,
                                    glm4moe_layer_t** out) {
  int rc = check_cfg(cfg);
  if (rc != GLM4MOE_OK) return rc;
  if (!gate_w || !bias || !w_gate_up || !w_down || !workspace || !out) return GLM4MOE_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(workspace) % 1024 != 0) return GLM4MOE_ERR_ARG;
  WsLayout L = make_layout(*cfg);
  if (workspace_bytes < L.total) return GLM4MOE_ERR_WORKSPACE;
  glm4moe_layer* l = new (std::nothrow) glm4moe_layer();
  if (!l) return GLM4MOE_ERR_ARG;
  l->cfg = *cfg;
  l->L = L;
  l->ws = static_cast<unsigned char*>(workspace);
  l->gate_w = gate_w;
  l->bias = bias;
  l->w_gate_up = w_gate_up;
  l->w_down = w_down;
  l->s_gate_up = s_gate_up;
  l->s_down = s_down;
  l->sf_gate_up = l->sf_down = nullptr;
  l->last_launches = 0;
  l->last_impl = GLM4MOE_IMPL_AUTO;
  l->last_fused = 0;
  l->events_ready = false;
  // The self-re-arming counters start at zero.  The workspace usually comes from a stream-ordered
  // caching allocator and forwards run on the caller's (non-blocking) stream, which is not ordered
  // against the legacy stream these memsets run on: fence the device on both sides.  Creation is
  // rare (once per layer / capacity change) and never happens inside a graph capture.
  if (cudaDeviceSynchronize() != cudaSuccess ||
      cudaMemset(l->ws + L.sched, 0, 256) != cudaSuccess ||
      cudaMemset(l->ws + L.ready, 0, (cfg->n_routed_experts + 8) * 4) != cudaSuccess ||
      cudaMemset(l->ws + L.fdone, 0, (cfg->hidden_size / 128 + 8) * 4) != cudaSuccess ||
      cudaMemset(l->ws + L.ctl, 0, kEpCtlBytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) {
    delete l;
    return GLM4MOE_ERR_DEVICE;
  }
  int dev = 0;
  l->sm_count = 148;
  if (cudaGetDevice(&dev) == cudaSuccess)
    cudaDeviceGetAttribute(&l->sm_count, cudaDevAttrMultiProcessorCount, dev);
  const int H = cfg->hidden_size, I = cfg->moe_intermediate_size;
  l->umma_ok = false;
  const bool fp8 = is_fp8(cfg->weight_format);
  if (fp8) {
    if (!s_gate_up || !s_down || cfg->impl == GLM4MOE_IMPL_SIMT) {
      delete l;
      return GLM4MOE_ERR_ARG;
    }
    const uint64_t n_stack = static_cast<uint64_t>(cfg->expert_end - cfg->expert_begin) + 1;
    if (cfg->weight_format == GLM4MOE_W_FP8_BLOCK128_POW2) {
      // packed exponent bytes of the block scales for the block-scaled MMA; refuses scales that are
      // not powers of two (this format is a promise about the checkpoint, not a conversion)
      const size_t nb_gu = n_stack * (2 * I / 128), nb_d = n_stack * (H / 128);
      const int kb_gu = H / 128, kb_d = I / 128;
      int* bad = nullptr;
      int bad_h = 0;
      bool ok = cudaMalloc(&l->sf_gate_up, nb_gu * ((kb_gu + 3) / 4) * 4) == cudaSuccess &&
                cudaMalloc(&l->sf_down, nb_d * ((kb_d + 3) / 4) * 4) == cudaSuccess &&
                cudaMalloc(&bad, 4) == cudaSuccess && cudaMemset(bad, 0, 4) == cudaSuccess &&
                pack_weight_sf(s_gate_up, l->sf_gate_up, nb_gu, kb_gu, bad, nullptr) == GLM4MOE_OK &&
                pack_weight_sf(s_down, l->sf_down, nb_d, kb_d, bad, nullptr) == GLM4MOE_OK &&
                cudaMemcpy(&bad_h, bad, 4, cudaMemcpyDeviceToHost) == cudaSuccess;
      cudaFree(bad);
      if (!ok || bad_h) {
        cudaFree(l->sf_gate_up);
        cudaFree(l->sf_down);
        delete l;
        return ok ? GLM4MOE_ERR_ARG : GLM4MOE_ERR_DEVICE;
      }
    }
    const uint64_t R = static_cast<uint64_t>(cfg->max_tokens) * (cfg->num_experts_per_tok + 1);
    rc = umma_make_tensor_map_u8(&l->tm_w_gu, w_gate_up, n_stack * 2 * I, H, 128);
    if (rc == GLM4MOE_OK) rc = umma_make_tensor_map_u8(&l->tm_w_d, w_down, n_stack * H, I, 128);
    for (int i = 0; i < 4 && rc == GLM4MOE_OK; ++i) {
      rc = umma_make_tensor_map_u8(&l->tm_x[i], l->ws + L.xq, R, H, kBnChoices[i]);
      if (rc == GLM4MOE_OK) rc = umma_make_tensor_map_u8(&l->tm_act[i], l->ws + L.actq, R, I, kBnChoices[i]);
    }
    // the router stays in the model dtype
    if (rc == GLM4MOE_OK && H % 64 == 0)
      rc = umma_make_tensor_map(&l->tm_gate, gate_w, cfg->dtype, cfg->n_routed_experts, H, 128);
    if (rc != GLM4MOE_OK) {
      delete l;
      return rc;
    }
    l->umma_ok = true;
  } else if (cfg->impl != GLM4MOE_IMPL_SIMT && umma_shape_ok(H, I, cfg->dtype)) {
    const uint64_t n_stack = static_cast<uint64_t>(cfg->expert_end - cfg->expert_begin) + 1;
    const uint64_t R = static_cast<uint64_t>(cfg->max_tokens) * (cfg->num_experts_per_tok + 1);
    rc = umma_make_tensor_map(&l->tm_w_gu, w_gate_up, cfg->dtype, n_stack * 2 * I, H, 128);
    if (rc == GLM4MOE_OK) rc = umma_make_tensor_map(&l->tm_w_d, w_down, cfg->dtype, n_stack * H, I, 128);
    if (rc == GLM4MOE_OK)
      rc = umma_make_tensor_map(&l->tm_gate, gate_w, cfg->dtype, cfg->n_routed_experts, H, 128);
    for (int i = 0; i < kNumBn && rc == GLM4MOE_OK; ++i) {
      rc = umma_make_tensor_map(&l->tm_x[i], l->ws + L.x_rows, cfg->dtype, R, H, kBnChoices[i]);
      if (rc == GLM4MOE_OK)
        rc = umma_make_tensor_map(&l->tm_act[i], l->ws + L.act, cfg->dtype, R, I, kBnChoices[i]);
    }
    if (rc == GLM4MOE_OK) l->umma_ok = true;
    if (rc == GLM4MOE_OK && H % 32 == 0 && I % 32 == 0) {
      int r2 = umma_make_tensor_map_bk(&l->tm_w_gu32, w_gate_up, cfg->dtype, n_stack * 2 * I, H, 128, 32);
      if (!r2) r2 = umma_make_tensor_map_bk(&l->tm_w_d32, w_down, cfg->dtype, n_stack * H, I, 128, 32);
      if (!r2) r2 = umma_make_tensor_map_bk(&l->tm_x32, l->ws + L.x_rows, cfg->dtype, R, H, 256, 32);
      if (!r2) r2 = umma_make_tensor_map_bk(&l->tm_act32, l->ws + L.act, cfg->dtype, R, I, 256, 32);
      l->bk32_ok = r2 == GLM4MOE_OK;
    }
  }
  if (cfg->impl == GLM4MOE_IMPL_UMMA && !l->umma_ok) {
    delete l;
    return rc != GLM4MOE_OK ? rc : GLM4MOE_ERR_UNSUPPORTED;
  }
  *out = l;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_layer_destroy(glm4moe_layer_t* layer) {
  if (layer && layer->events_ready)
    for (int i = 0; i < 8; ++i) cudaEventDestroy(layer->ev[i]);
  if (layer) {
    cudaFree(layer->sf_gate_up);
    cudaFree(layer->sf_down);
  }
  delete layer;
  return GLM4MOE_OK;
}

// ---- tuning / test knobs of the tensor-core path (glm4moe_debug_set; the environment variable only
// gives the initial value) ----
// GLM4MOE_BK32: K blocks of 32 with a 7-deep ring for the 256-token prefill tiles.  Off by default:
// measured SLOWER than the 3 x 64 KB ring (Air T=4096 gate_up 852 -> 881 us, T=8192 1490 -> 1625 us;
// profiles/r1d.md) although it keeps 1.5x the bytes in flight.
static bool bk32_enabled() {
  if (g_bk32 < 0) {
    const char* e = getenv("GLM4MOE_BK32");
    g_bk32 = (e && atoi(e) != 0) ? 1 : 0;
  }
  return g_bk32 == 1;
}
// GLM4MOE_NO_DOWN2=1: plain 128-feature down tiles at prefill instead of 256 features x 256 tokens
static bool tail_box_enabled() { return g_tail_box != 0; }

static bool down2_enabled() {
  if (g_down2 < 0) {
    const char* e = getenv("GLM4MOE_NO_DOWN2");
    g_down2 = (e && atoi(e) != 0) ? 0 : 1;
  }
  return g_down2 == 1;
}
// GLM4MOE_FORCE_BN=16/32/64/128/256: token rows per tile regardless of the batch (-1 = by mean rows per expert)
static int forced_bn_index() {
  if (g_force_bn == -2) {
    const char* e = getenv("GLM4MOE_FORCE_BN");
    g_force_bn = -1;
    if (e) for (int i = 0; i < kNumBn; ++i) if (atoi(e) == kBnChoices[i]) g_force_bn = i;
  }
  return g_force_bn;
}

// rows of one token tile on the tensor-core path: sized to the mean rows per local expert so a
// typical expert is one tile (weights cross HBM once; a second tile re-reads them from L2).
static int pick_bn_index(const glm4moe_config_t& c, int T) {
  const int forced = forced_bn_index();
  if (forced >= 0) return forced;
  const int e_local = c.expert_end - c.expert_begin;
  const double mean = e_local > 0 ? static_cast<double>(T) * c.num_experts_per_tok / c.n_routed_experts : T;
  if (mean <= 6.0) return 0;
  if (mean <= 20.0) return 1;
  if (mean <= 48.0) return 2;
  if (mean <= 96.0) return 3;
  return 4;
}

extern "C" int glm4moe_layer_tile_rows(const glm4moe_layer_t* l, int T) {
  if (!l) return GLM4MOE_ERR_ARG;
  return l->umma_ok ? kBnChoices[pick_bn_index(l->cfg, T)] : kSimtRows;
}

static int forward_impl(glm4moe_layer_t* l, const void* x, void* out, int T, void* stream, cudaEvent_t* ev,
                        int sh_lo_arg = -1, int sh_hi_arg = -1, bool rows_only = false,
                        const EpFastSignal* epf = nullptr, bool* ep_fast_used = nullptr) {
  // epf (rows_only): when this call runs the fused decode kernel, use the expert-parallel fast path (rows and
  // slot_row in the parity buffers, ready flags published by the expert kernel) and report it in *ep_fast_used
  // rows_only: stop after the weighted expert rows y (no combine, `out` unused): the peer-memory
  // expert-parallel combine (ep_combine.cu) reads the rows of every rank
  if (!l || !x || (!out && !rows_only)) return GLM4MOE_ERR_ARG;
  cudaStream_t cst = static_cast<cudaStream_t>(stream);
#define STAGE_MARK(i)                                        \
  do {                                                       \
    if (ev) GLM4MOE_CUDA_CHECK(cudaEventRecord(ev[i], cst)); \
  } while (0)
  const glm4moe_config_t& c = l->cfg;
  if (T < 0 || T > c.max_tokens) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  const int H = c.hidden_size, E = c.n_routed_experts, k = c.num_experts_per_tok,
            I = c.moe_intermediate_size;
  unsigned char* ws = l->ws;
  const WsLayout& L = l->L;
  void* logits = ws + L.logits;
  int32_t* idx = reinterpret_cast<int32_t*>(ws + L.idx);
  float* w = reinterpret_cast<float*>(ws + L.w);
  int32_t* offsets = reinterpret_cast<int32_t*>(ws + L.offsets);
  int32_t* row_token = reinterpret_cast<int32_t*>(ws + L.row_token);
  float* row_weight = reinterpret_cast<float*>(ws + L.row_weight);
  int32_t* slot_row = reinterpret_cast<int32_t*>(ws + L.slot_row);
  int32_t* num_tiles = reinterpret_cast<int32_t*>(ws + L.num_tiles);
  int32_t* tiles = reinterpret_cast<int32_t*>(ws + L.tiles);
  int32_t* scratch = reinterpret_cast<int32_t*>(ws + L.scratch);
  void* x_rows = ws + L.x_rows;
  void* act = ws + L.act;
  void* y = ws + L.y;
  // shared-expert rows this rank computes: all tokens, none, or (with_shared == 2) its 1/R slice
  // of the tokens so expert-parallel ranks stay balanced (the partial outputs are summed later)
  int sh_lo = 0, sh_hi = 0;
  if (c.with_shared == 1) {
    sh_hi = T;
  } else if (c.with_shared == 2) {
    const int n_local = c.expert_end - c.expert_begin;
    const int R = n_local > 0 ? E / n_local : 1, r = n_local > 0 ? c.expert_begin / n_local : 0;
    const int per = (T + R - 1) / R;
    sh_lo = r * per < T ? r * per : T;
    sh_hi = sh_lo + per < T ? sh_lo + per : T;
  }
  if (sh_lo_arg >= 0) {   // explicit range (glm4moe_layer_forward_ex)
    if (sh_hi_arg > T || sh_lo_arg > sh_hi_arg) return GLM4MOE_ERR_ARG;
    sh_lo = sh_lo_arg;
    sh_hi = sh_hi_arg;
  }
  const bool use_umma = l->umma_ok;
  const bool fp8 = is_fp8(c.weight_format);
  const int pow2 = c.weight_format == GLM4MOE_W_FP8_BLOCK128_POW2 ? 1 : 0;
  // decode "direct" mode: fused front-end + gate_up reading the input tokens in place
  // (the e4m3 tier has no un-fused direct variant)
  const bool direct = use_umma && T <= kDirectMaxTokens && H % 64 == 0 && direct_enabled() &&
                      (!fp8 || fused_enabled());
  // e4m3 tier at prefill (>= ~48 rows per expert): token-major 128 x 256 tiles (umma_fp8_prefill.cu)
  const bool fp8_prefill = fp8 && !direct && pick_bn_index(c, T) >= 3 && fp8_prefill_enabled() &&
                           umma_fp8_prefill_shape_ok(H, I);
  // block-scaled pair kernel (cta_group::2): 256-row tiles, each CTA of a pair loads 128-row boxes
  const bool fp8_mx2 = fp8_prefill && pow2 && fp8_mx_mode() == 2;
  const int bi = direct ? (T <= 16 ? 0 : 1)
                 : fp8_mx2 ? 4
                 : fp8_prefill ? 3
                 : fp8 ? (pick_bn_index(c, T) < 2 ? pick_bn_index(c, T) : 2) : pick_bn_index(c, T);
  const int tile_rows = use_umma ? kBnChoices[bi] : kSimtRows;
  const int launches0 = g_launches;
  int rc;
  alignas(64) CUtensorMap tm_xdirect;
  bool combine_done = false;
  const bool ep_fast = epf != nullptr && rows_only && direct && fused_enabled();
  // T <= 2: the front-end runs inside the persistent kernel (decode_fused.cu, RouteMerged): one launch per layer
  const bool merged = direct && fused_enabled() && !fp8 && !rows_only && T <= 2 && T * k <= 32 && E <= 256 &&
                      g_merge_front != 0 && ev == nullptr;
  if (ep_fast_used) *ep_fast_used = ep_fast;
  STAGE_MARK(0);
  if (direct) {
    if (!fp8 && (rc = umma_make_tensor_map(&tm_xdirect, x, c.dtype, static_cast<uint64_t>(T), H, tile_rows)))
      return rc;
    RouteSmallEp rep{};
    rep.act_pow2 = pow2;
    if (!merged) {
    if (ep_fast) {
      rep.epoch = reinterpret_cast<const int32_t*>(ws + L.ctl);
      for (int i = 0; i < 2; ++i) rep.slot_row_alt[i] = reinterpret_cast<int32_t*>(ws + L.slot_row_alt[i]);
    }
    if ((rc = route_small_ep(x, l->gate_w, l->bias, logits, T, H, E, k, c.routed_scaling_factor, c.dtype,
                             c.expert_begin, c.expert_end, tile_rows, sh_lo, sh_hi, idx, w, offsets, row_token,
                             row_weight, slot_row, tiles, L.tile_capacity, num_tiles,
                             reinterpret_cast<int32_t*>(ws + L.sched) + 24, fp8 ? ws + L.xq : nullptr,
                             fp8 ? reinterpret_cast<float*>(ws + L.sx) : nullptr,
                             c.max_tokens * (k + 1), rep, stream)))
      return rc;
    }
    STAGE_MARK(1);
    STAGE_MARK(2);
    STAGE_MARK(3);
  } else {
  if (use_umma && T >= kRouterUmmaMinTokens && H % 64 == 0) {
    // router logits on the tensor cores: D[expert, token] = Wg . x^T, dense token tiles of x itself
    alignas(64) CUtensorMap tm_xin;
    // token-tile rows: the gate has a single 128-expert feature tile, so the tile count is the only
    // source of parallelism; small tiles until the tiles alone cover the SMs
    // The gate's weight tile (E x H, ~1 MB) is re-read from L2 by every token tile, so at prefill the
    // kernel is bound by L2 bandwidth in proportion to the tile count: 128-token tiles from T = 4096
    // (32 tiles: each SM then streams its own 2 MB instead of 148 SMs fighting for 160 MB).
    int rbn = T >= 4096 ? 128 : T >= 3200 ? 32 : 16;
    if (g_router_rows) rbn = g_router_rows;
    if ((rc = umma_make_tensor_map(&tm_xin, x, c.dtype, static_cast<uint64_t>(T), H, rbn))) return rc;
    UmmaLaunch r{};
    r.tm_w = &l->tm_gate;
    r.tm_x = &tm_xin;
    r.tiles = nullptr;
    r.num_tiles = nullptr;
    r.dense_rows = T;
    r.out = logits;
    r.ld_out = E;
    r.n_feat_tiles = (E + 127) / 128;
    r.n_feat_valid = E;
    r.n_kblocks = H / 64;
    r.w_rows_per_expert = 0;
    r.tile_rows = rbn;
    r.mode = 2;
    r.dtype = c.dtype;
    r.grid = l->sm_count;
    r.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 4;
    if ((rc = umma_launch(r, stream))) return rc;
  } else if ((rc = glm4moe_router_logits(x, l->gate_w, logits, T, H, E, c.dtype, stream))) {
    return rc;
  }
  STAGE_MARK(1);
  if ((rc = glm4moe_router_select(logits, l->bias, T, E, k, c.routed_scaling_factor, c.dtype, idx, w, stream)))
    return rc;
  STAGE_MARK(2);
  if ((rc = glm4moe_permute(idx, w, T, k, E, c.expert_begin, c.expert_end, tile_rows, sh_lo, sh_hi, offsets,
                            row_token, row_weight, slot_row, tiles, L.tile_capacity, num_tiles, scratch, stream)))
    return rc;
  STAGE_MARK(3);
  }
  const int R_cap = c.max_tokens * (k + 1);
  if (direct) {
    // no gathered copy of the token rows
  } else if (fp8) {
    // gather fused with the per-token 1x128 activation quantiser
    if ((rc = quant_scatter_fp8(x, H, slot_row, T, k, sh_lo, sh_hi, ws + L.xq,
                                reinterpret_cast<float*>(ws + L.sx), R_cap, c.dtype, pow2,
                                pow2 ? ws + L.sx_sf : nullptr, stream)))
      return rc;
  } else if ((rc = glm4moe_gather_rows(x, row_token, offsets, x_rows, T, k, E, H, sh_lo, sh_hi, c.dtype, stream))) {
    return rc;
  }
  STAGE_MARK(4);
  if (direct && fused_enabled()) {
    DecodeFusedLaunch f{};
    f.tm_wgu = &l->tm_w_gu;
    f.tm_wd = &l->tm_w_d;
    f.tm_x = fp8 ? &l->tm_x[bi] : &tm_xdirect;
    f.tm_act = &l->tm_act[bi];
    f.tiles = tiles;
    f.num_tiles = num_tiles;
    f.row_token = row_token;
    f.row_weight = row_weight;
    f.act = act;
    f.y = y;
    f.H = H;
    f.I = I;
    f.tokens = T;
    f.n_routed_rows = T * k;
    f.shared_expert = c.expert_end - c.expert_begin;
    f.sh_lo = sh_lo;
    f.sh_hi = sh_hi;
    f.tile_rows = tile_rows;
    f.dtype = c.dtype;
    f.grid = l->sm_count;
    f.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 16;
    f.ready = reinterpret_cast<int32_t*>(ws + L.ready);
    f.slot_row = slot_row;
    f.out = rows_only ? nullptr : out;
    f.k = k;
    f.fdone = reinterpret_cast<int32_t*>(ws + L.fdone);
    {
      const size_t es = fp8 ? 1 : dtype_size(c.dtype), n_local = static_cast<size_t>(c.expert_end - c.expert_begin);
      f.pf0 = static_cast<const unsigned char*>(l->w_gate_up) + n_local * 2 * I * H * es;
      f.pf0_bytes = static_cast<size_t>(2) * I * H * es;
      f.pf1 = static_cast<const unsigned char*>(l->w_down) + n_local * H * I * es;
      f.pf1_bytes = static_cast<size_t>(H) * I * es;
      // speculative prefetch of the first local experts (decode_fused.cu): on when an expert is hit with
      // probability >= 0.8 (1 - (1 - k/E)^T), as many experts as fit ~70 MB of L2 next to the shared expert
      int n_spec = g_spec_prefetch;
      if (n_spec < 0) {
        float miss = 1.f;
        for (int t = 0; t < T; ++t) miss *= 1.f - static_cast<float>(k) / static_cast<float>(E);
        const size_t per_expert = static_cast<size_t>(3) * I * H * es;
        n_spec = miss <= 0.2f ? static_cast<int>((70u << 20) / per_expert) : 0;
      }
      if (n_spec > static_cast<int>(n_local)) n_spec = static_cast<int>(n_local);
      if (merged) n_spec = 0;
      f.pf2 = l->w_gate_up;
      f.pf2_bytes = static_cast<size_t>(n_spec) * 2 * I * H * es;
      f.pf3 = l->w_down;
      f.pf3_bytes = static_cast<size_t>(n_spec) * H * I * es;
    }
    if (fp8) {
      f.fp8 = 1;
      f.q.s_gate_up = l->s_gate_up;
      f.q.s_down = l->s_down;
      f.q.sx_t = reinterpret_cast<const float*>(ws + L.sx);
      f.q.sact_t = reinterpret_cast<float*>(ws + L.sact);
      f.q.actq = ws + L.actq;
      f.q.scale_ld = R_cap;
      f.q.pow2 = pow2;
    }
    if (ep_fast) f.ep = *epf;
    if (merged) {
      f.rt.enabled = 1;
      f.rt.x = x;
      f.rt.gate_w = l->gate_w;
      f.rt.bias = l->bias;
      f.rt.logits = logits;
      f.rt.idx = idx;
      f.rt.w = w;
      f.rt.expert_offsets = offsets;
      f.rt.row_token = row_token;
      f.rt.row_weight = row_weight;
      f.rt.slot_row = slot_row;
      f.rt.tiles = tiles;
      f.rt.tile_capacity = L.tile_capacity;
      f.rt.num_tiles = num_tiles;
      f.rt.E = E;
      f.rt.k = k;
      f.rt.e_lo = c.expert_begin;
      f.rt.e_hi = c.expert_end;
      f.rt.scale = c.routed_scaling_factor;
      f.rt.sync = reinterpret_cast<int32_t*>(ws + L.sched) + 40;
    }
    if ((rc = decode_fused_launch(f, stream))) return rc;
    STAGE_MARK(5);
    combine_done = true;
  } else if (fp8_prefill && pow2 && fp8_mx_mode() != 0) {
    // scales applied by the tensor pipe (tcgen05 block-scaled MMA), one accumulator read per tile
    UmmaFp8MxLaunch a{};
    a.tiles = tiles;
    a.num_tiles = num_tiles;
    a.dtype = c.dtype;
    a.grid = l->sm_count;
    a.row_weight = row_weight;
    a.n_routed_rows = T * k;
    a.scale_ld = R_cap;
    a.tm_w = &l->tm_w_gu;
    a.tm_x = &l->tm_x[3];
    a.w_sf = l->sf_gate_up;
    a.x_sf = reinterpret_cast<const uint32_t*>(ws + L.sx_sf);
    a.out = nullptr;
    a.ld_out = I;
    a.out_q = ws + L.actq;
    a.out_scale_t = reinterpret_cast<float*>(ws + L.sact);
    a.out_sf = ws + L.sact_sf;
    a.n_feat_tiles = I / 128;
    a.n_kblocks = H / 128;
    a.w_rows_per_expert = 2 * I;
    a.up_row_offset = I;
    a.gated = 1;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched);
    // gate_up: token-major (thread-local re-quantiser); down: feature-major (ragged tails cost their rows)
    if ((rc = fp8_mx2 ? umma_fp8_mx2_launch(a, stream)
              : fp8_mx_mode() == 3 ? umma_fp8_mx_launch(a, stream) : umma_fp8_mx_tm_launch(a, stream)))
      return rc;
    STAGE_MARK(5);
    a.tm_w = &l->tm_w_d;
    a.tm_x = &l->tm_act[3];
    a.w_sf = l->sf_down;
    a.x_sf = reinterpret_cast<const uint32_t*>(ws + L.sact_sf);
    a.out = y;
    a.ld_out = H;
    a.out_q = nullptr;
    a.out_scale_t = nullptr;
    a.out_sf = nullptr;
    a.n_feat_tiles = H / 256;
    a.n_kblocks = I / 128;
    a.w_rows_per_expert = H;
    a.up_row_offset = 0;
    a.gated = 0;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 2;
    if ((rc = fp8_mx2 ? umma_fp8_mx2_launch(a, stream) : umma_fp8_mx_launch(a, stream))) return rc;
  } else if (fp8_prefill) {
    UmmaFp8PrefillLaunch a{};
    a.pow2 = pow2;
    a.tiles = tiles;
    a.num_tiles = num_tiles;
    a.dtype = c.dtype;
    a.grid = l->sm_count;
    a.row_weight = row_weight;
    a.n_routed_rows = T * k;
    a.scale_ld = R_cap;
    // gate_up + SiLU*mul + e4m3 re-quantisation (the activation rows exist only in e4m3)
    a.tm_w = &l->tm_w_gu;
    a.tm_x = &l->tm_x[bi];
    a.w_scale = l->s_gate_up;
    a.x_scale_t = reinterpret_cast<const float*>(ws + L.sx);
    a.out = nullptr;
    a.ld_out = I;
    a.out_q = ws + L.actq;
    a.out_scale_t = reinterpret_cast<float*>(ws + L.sact);
    a.out_sf = nullptr;
    a.n_feat_tiles = I / 128;
    a.n_kblocks = H / 128;
    a.w_rows_per_expert = 2 * I;
    a.up_row_offset = I;
    a.gated = 1;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched);
    if ((rc = umma_fp8_prefill_launch(a, stream))) return rc;
    STAGE_MARK(5);
    // down * routing weight
    a.tm_w = &l->tm_w_d;
    a.tm_x = &l->tm_act[bi];
    a.w_scale = l->s_down;
    a.x_scale_t = reinterpret_cast<const float*>(ws + L.sact);
    a.out = y;
    a.ld_out = H;
    a.out_q = nullptr;
    a.out_scale_t = nullptr;
    a.n_feat_tiles = H / 256;
    a.n_kblocks = I / 128;
    a.w_rows_per_expert = H;
    a.up_row_offset = 0;
    a.gated = 0;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 2;
    if ((rc = umma_fp8_prefill_launch(a, stream))) return rc;
  } else if (fp8) {
    UmmaFp8Launch a{};
    a.tiles = tiles;
    a.num_tiles = num_tiles;
    a.tile_rows = tile_rows;
    a.dtype = c.dtype;
    a.grid = l->sm_count;
    a.row_weight = row_weight;
    a.n_routed_rows = T * k;
    a.x_scale_ld = R_cap;
    // gate_up + SiLU*mul
    a.tm_w = &l->tm_w_gu;
    a.tm_x = &l->tm_x[bi];
    a.w_scale = l->s_gate_up;
    a.x_scale_t = reinterpret_cast<const float*>(ws + L.sx);
    a.out = act;
    a.ld_out = I;
    a.n_feat_tiles = I / 128;
    a.n_kblocks = H / 128;
    a.w_rows_per_expert = 2 * I;
    a.up_row_offset = I;
    a.gated = 1;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched);
    if ((rc = umma_fp8_launch(a, stream))) return rc;
    STAGE_MARK(5);
    // re-quantise the activation rows for the down projection
    if ((rc = quant_rows_fp8(act, I, nullptr, offsets, E, T, k, I, sh_lo, sh_hi, ws + L.actq,
                             reinterpret_cast<float*>(ws + L.sact), R_cap, c.dtype, pow2, nullptr, stream)))
      return rc;
    a.tm_w = &l->tm_w_d;
    a.tm_x = &l->tm_act[bi];
    a.w_scale = l->s_down;
    a.x_scale_t = reinterpret_cast<const float*>(ws + L.sact);
    a.out = y;
    a.ld_out = H;
    a.n_feat_tiles = H / 128;
    a.n_kblocks = I / 128;
    a.w_rows_per_expert = H;
    a.up_row_offset = 0;
    a.gated = 0;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 2;
    if ((rc = umma_fp8_launch(a, stream))) return rc;
  } else if (use_umma) {
    UmmaLaunch a{};
    a.tiles = tiles;
    a.num_tiles = num_tiles;
    a.tile_rows = tile_rows;
    a.dtype = c.dtype;
    a.grid = l->sm_count;
    a.row_weight = row_weight;
    a.n_routed_rows = T * k;
    // gate_up + SiLU*mul
    a.tm_w = &l->tm_w_gu;
    a.tm_x = &l->tm_x[bi];
    // ragged tiles come through the smaller boxes over the same rows (the three tile sizes below this one)
    const int tail_lo = bi >= 3 ? bi - 3 : 0;
    const bool tails = !direct && tail_box_enabled();
    for (int i = 0; i < 3; ++i) {
      const bool on = tails && tail_lo + i < bi;
      a.tm_x_tail[i] = on ? &l->tm_x[tail_lo + i] : nullptr;
      a.tail_rows[i] = on ? kBnChoices[tail_lo + i] : 0;
    }
    a.out = act;
    a.ld_out = I;
    a.n_feat_tiles = I / 128;
    a.n_feat_valid = I;
    a.n_kblocks = H / 64;
    a.w_rows_per_expert = 2 * I;
    a.up_row_offset = I;
    a.mode = 1;
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched);
    // prefill tiles: half-size stages (K blocks of 32, 7-deep ring) keep 1.5x the bytes in flight
    const bool bk32 = tile_rows == 256 && l->bk32_ok && !direct && bk32_enabled();
    if (bk32) {
      a.bk = 32;
      a.tm_w = &l->tm_w_gu32;
      a.tm_x = &l->tm_x32;
      a.n_kblocks = H / 32;
    }
    if (direct) {
      a.tm_x = &tm_xdirect;
      a.direct.tokens = T;
      a.direct.shared_expert = c.expert_end - c.expert_begin;
      a.direct.shared_row0 = T * k;
      a.direct.sh_lo = sh_lo;
      a.direct.sh_hi = sh_hi;
      a.direct.row_token = row_token;
    }
    if ((rc = umma_launch(a, stream))) return rc;
    a.direct = DirectArgs{};
    STAGE_MARK(5);
    // down * routing weight
    a.tm_w = &l->tm_w_d;
    a.tm_x = &l->tm_act[bi];
    for (int i = 0; i < 3; ++i) a.tm_x_tail[i] = a.tm_x_tail[i] ? &l->tm_act[tail_lo + i] : nullptr;
    a.out = y;
    a.ld_out = H;
    a.n_feat_tiles = H / 128;
    a.n_feat_valid = H;
    a.n_kblocks = I / 64;
    a.w_rows_per_expert = H;
    a.up_row_offset = 0;
    a.mode = 0;
    a.bk = 0;
    if (tile_rows == 256 && H % 256 == 0 && down2_enabled()) {
      // prefill: 256 features x 256 tokens per CTA tile (two weight tiles share one token tile)
      a.mode = 3;
      a.n_feat_tiles = H / 256;
      a.up_row_offset = 128;
      if (bk32) {
        a.bk = 32;
        a.tm_w = &l->tm_w_d32;
        a.tm_x = &l->tm_act32;
        a.n_kblocks = I / 32;
      }
    }
    a.sched = reinterpret_cast<int32_t*>(ws + L.sched) + 2;
    if ((rc = umma_launch(a, stream))) return rc;
  } else {
    if ((rc = glm4moe_expert_gate_up_simt(x_rows, l->w_gate_up, tiles, num_tiles, act, H, I, tile_rows, c.dtype,
                                          stream)))
      return rc;
    STAGE_MARK(5);
    if ((rc = glm4moe_expert_down_simt(act, l->w_down, tiles, num_tiles, row_weight, y, H, I, T, k, tile_rows,
                                       c.dtype, stream)))
      return rc;
  }
  STAGE_MARK(6);
  if (!combine_done && !rows_only && (rc = glm4moe_combine(y, slot_row, out, T, k, H, sh_lo, sh_hi, c.dtype, stream))) return rc;
  STAGE_MARK(7);
#undef STAGE_MARK
  l->last_launches = g_launches - launches0;
  l->last_impl = use_umma ? GLM4MOE_IMPL_UMMA : GLM4MOE_IMPL_SIMT;
  l->last_fused = combine_done ? 1 : 0;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_layer_forward(glm4moe_layer_t* l, const void* x, void* out, int T, void* stream) {
  return forward_impl(l, x, out, T, stream, nullptr);
}

extern "C" int glm4moe_layer_forward_ex(glm4moe_layer_t* l, const void* x, void* out, int T, int shared_lo,
                                        int shared_hi, void* stream) {
  if (shared_lo < 0 || shared_hi < shared_lo) return GLM4MOE_ERR_ARG;
  return forward_impl(l, x, out, T, stream, nullptr, shared_lo, shared_hi);
}

extern "C" int glm4moe_layer_forward_timed(glm4moe_layer_t* l, const void* x, void* out, int T, void* stream,
                                           float* stage_ms) {
  if (!l || !stage_ms) return GLM4MOE_ERR_ARG;
  if (T <= 0) return GLM4MOE_ERR_ARG;
  if (!l->events_ready) {
    for (int i = 0; i < 8; ++i) GLM4MOE_CUDA_CHECK(cudaEventCreate(&l->ev[i]));
    l->events_ready = true;
  }
  int rc = forward_impl(l, x, out, T, stream, l->ev);
  if (rc) return rc;
  GLM4MOE_CUDA_CHECK(cudaEventSynchronize(l->ev[7]));
  for (int i = 0; i < 7; ++i) GLM4MOE_CUDA_CHECK(cudaEventElapsedTime(&stage_ms[i], l->ev[i], l->ev[i + 1]));
  GLM4MOE_CUDA_CHECK(cudaEventElapsedTime(&stage_ms[7], l->ev[0], l->ev[7]));
  return GLM4MOE_OK;
}

// Dense SiLU-gated MLP on the tensor-core grouped GEMM: one weight pair, all T rows ("tiles == nullptr"
// path of umma_grouped_kernel, the form the router GEMM already uses).  Stateless: tensor maps are
// encoded per call on the host, the only device state is `sched` (4 self re-arming ints).
extern "C" int glm4moe_mlp_forward(const void* x, const void* w_gate_up, const void* w_down, void* act_ws,
                                   void* out, int T, int H, int I, int dtype, int32_t* sched, void* stream) {
  if (!x || !w_gate_up || !w_down || !act_ws || !out || !sched || T < 0 || H <= 0 || I <= 0) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  // 2-byte dtypes; K of the first GEMM in whole 64-element blocks; row pitches 16-byte aligned for TMA.
  // I need not be a multiple of 128 (Air's dense layer: 10944): the ragged last feature tile is bounded
  // in the epilogue and the ragged last K block of the down GEMM is zero-filled by TMA on both operands.
  if ((dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16) || H % 128 != 0 || I % 8 != 0) return GLM4MOE_ERR_UNSUPPORTED;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // token rows per tile, per GEMM: an item streams its weight rows (256 gate+up / 128 down) plus BN token rows per
  // k-block, items = token tiles x feature tiles run in waves over the SMs: take the BN with the fewest
  // wave x (weight rows + BN) units (T = 256: one 256-row tile would leave the down GEMM with H/128 = 32 items)
  auto pick = [&](int n_feat, int w_rows) {
    int best = 16;
    long best_cost = -1;
    for (int i = 0; i < kNumBn; ++i) {
      const int bn_i = kBnChoices[i];
      const long items = static_cast<long>((T + bn_i - 1) / bn_i) * n_feat;
      const long cost = ((items + sms - 1) / sms) * (w_rows + bn_i);
      if (best_cost < 0 || cost < best_cost) {
        best_cost = cost;
        best = bn_i;
      }
      if (bn_i >= T) break;
    }
    return best;
  };
  const int bn = pick((I + 127) / 128, 256), bn_d = pick(H / 128, 128);
  alignas(64) CUtensorMap tm_wgu, tm_wd, tm_xin, tm_act;
  int rc;
  if ((rc = umma_make_tensor_map(&tm_wgu, w_gate_up, dtype, static_cast<uint64_t>(2) * I, H, 128))) return rc;
  if ((rc = umma_make_tensor_map(&tm_wd, w_down, dtype, static_cast<uint64_t>(H), I, 128))) return rc;
  if ((rc = umma_make_tensor_map(&tm_xin, x, dtype, static_cast<uint64_t>(T), H, bn))) return rc;
  if ((rc = umma_make_tensor_map(&tm_act, act_ws, dtype, static_cast<uint64_t>(T), I, bn_d))) return rc;
  UmmaLaunch a{};
  a.tiles = nullptr;
  a.num_tiles = nullptr;
  a.dense_rows = T;
  a.tile_rows = bn;
  a.dtype = dtype;
  a.grid = sms;
  a.row_weight = nullptr;
  a.n_routed_rows = 0;          // every row takes weight 1
  // gate_up + SiLU*mul (mlp.py:50-51)
  a.tm_w = &tm_wgu;
  a.tm_x = &tm_xin;
  a.out = act_ws;
  a.ld_out = I;
  a.n_feat_tiles = (I + 127) / 128;
  a.n_feat_valid = I;
  a.n_kblocks = H / 64;
  a.w_rows_per_expert = 2 * I;
  a.up_row_offset = I;
  a.mode = 1;
  a.sched = sched;
  if ((rc = umma_launch(a, stream))) return rc;
  // down (mlp.py:52)
  a.tm_w = &tm_wd;
  a.tm_x = &tm_act;
  a.out = out;
  a.ld_out = H;
  a.n_feat_tiles = H / 128;
  a.n_feat_valid = H;
  a.n_kblocks = (I + 63) / 64;
  a.w_rows_per_expert = H;
  a.up_row_offset = 0;
  a.mode = 0;
  a.tile_rows = bn_d;
  if (bn_d == 256 && H % 256 == 0 && down2_enabled()) {
    a.mode = 3;
    a.n_feat_tiles = H / 256;
    a.up_row_offset = 128;
  }
  a.sched = sched + 2;
  return umma_launch(a, stream);
}

extern "C" int glm4moe_layer_forward_host(glm4moe_layer_t* l, const void* x_host, void* out_host, int T,
                                          void* x_dev_staging, void* out_dev_staging, void* stream) {
  if (!l || !x_host || !out_host || !x_dev_staging || !out_dev_staging) return GLM4MOE_ERR_ARG;
  if (T < 0 || T > l->cfg.max_tokens) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t bytes = static_cast<size_t>(T) * l->cfg.hidden_size * dtype_size(l->cfg.dtype);
  GLM4MOE_CUDA_CHECK(cudaMemcpyAsync(x_dev_staging, x_host, bytes, cudaMemcpyHostToDevice, st));
  int rc = glm4moe_layer_forward(l, x_dev_staging, out_dev_staging, T, stream);
  if (rc) return rc;
  GLM4MOE_CUDA_CHECK(cudaMemcpyAsync(out_host, out_dev_staging, bytes, cudaMemcpyDeviceToHost, st));
  GLM4MOE_CUDA_CHECK(cudaStreamSynchronize(st));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_layer_view(glm4moe_layer_t* l, int which, void** ptr, size_t* bytes) {
  if (!l || !ptr || !bytes) return GLM4MOE_ERR_ARG;
  const glm4moe_config_t& c = l->cfg;
  const WsLayout& L = l->L;
  const size_t T = c.max_tokens, k = c.num_experts_per_tok, E = c.n_routed_experts, H = c.hidden_size,
               I = c.moe_intermediate_size, es = dtype_size(c.dtype), R = T * (k + 1);
  size_t off, n;
  switch (which) {
    case 0: off = L.idx; n = T * k * 4; break;
    case 1: off = L.w; n = T * k * 4; break;
    case 2: off = L.offsets; n = (E + 1) * 4; break;
    case 3: off = L.row_token; n = T * k * 4; break;
    case 4: off = L.slot_row; n = T * k * 4; break;
    case 5: off = L.logits; n = T * E * es; break;
    case 6: off = L.act; n = R * I * es; break;
    case 7: off = L.y; n = R * H * es; break;
    case 8: off = L.num_tiles; n = 4; break;
    case 9: off = L.tiles; n = static_cast<size_t>(L.tile_capacity) * 16; break;
    case 10: off = L.x_rows; n = R * H * es; break;
    case 11: off = L.row_weight; n = T * k * 4; break;
    default: return GLM4MOE_ERR_ARG;
  }
  *ptr = l->ws + off;
  *bytes = n;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_layer_last_launches(const glm4moe_layer_t* l) { return l ? l->last_launches : 0; }
extern "C" int glm4moe_layer_last_impl(const glm4moe_layer_t* l) { return l ? l->last_impl : 0; }
extern "C" int glm4moe_layer_last_fused(const glm4moe_layer_t* l) { return l ? l->last_fused : 0; }

// ---- what a cross-GPU wait that gives up leaves behind (ep_combine.cuh: EpErrorRecord) ------------------------
namespace {
std::mutex g_ep_watch_mu;
glm4moe::EpErrorRecord* g_ep_rec_host = nullptr;   // pinned + mapped + portable: readable after the context is lost
glm4moe::EpErrorRecord* g_ep_rec_dev = nullptr;    // the same record as the GPUs address it
long long g_ep_spin_limit = 0;                     // 0 = not decided yet (GLM4MOE_EP_TIMEOUT_MS, else the default)
constexpr long long kEpDefaultSpinLimit = 1ll << 35;   // ~18 s of SM clocks

long long cycles_of_ms(long long ms) {
  int dev = 0, khz = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev) != cudaSuccess ||
      khz <= 0)
    khz = 1965000;                                 // B200 boost clock
  return ms * static_cast<long long>(khz);
}

// host memory only: called from the set-up entry every expert-parallel user goes through (never inside a capture)
int ep_record_ensure() {
  std::lock_guard<std::mutex> lock(g_ep_watch_mu);
  if (g_ep_rec_host) return GLM4MOE_OK;
  void* h = nullptr;
  GLM4MOE_CUDA_CHECK(cudaHostAlloc(&h, sizeof(glm4moe::EpErrorRecord), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h, 0, sizeof(glm4moe::EpErrorRecord));
  void* d = nullptr;
  cudaError_t e = cudaHostGetDevicePointer(&d, h, 0);
  if (e != cudaSuccess) {
    cudaFreeHost(h);
    return static_cast<int>(e);
  }
  g_ep_rec_host = static_cast<glm4moe::EpErrorRecord*>(h);
  g_ep_rec_dev = static_cast<glm4moe::EpErrorRecord*>(d);
  return GLM4MOE_OK;
}
}  // namespace

namespace glm4moe {
EpWatch ep_watch(int rank, int site) {
  std::lock_guard<std::mutex> lock(g_ep_watch_mu);
  if (g_ep_spin_limit == 0) {
    const char* e = getenv("GLM4MOE_EP_TIMEOUT_MS");
    const long long ms = e ? atoll(e) : 0;
    g_ep_spin_limit = ms > 0 ? cycles_of_ms(ms) : kEpDefaultSpinLimit;
  }
  EpWatch w{};
  w.rec = g_ep_rec_dev;
  w.spin_limit = g_ep_spin_limit;
  w.rank = rank;
  w.site = site;
  return w;
}
}  // namespace glm4moe

extern "C" int glm4moe_ep_set_timeout_ms(int ms) {
  std::lock_guard<std::mutex> lock(g_ep_watch_mu);
  g_ep_spin_limit = ms > 0 ? cycles_of_ms(ms) : kEpDefaultSpinLimit;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_ep_last_error(int32_t* info6, int clear) {
  std::lock_guard<std::mutex> lock(g_ep_watch_mu);
  if (!g_ep_rec_host) return GLM4MOE_OK;
  volatile int32_t* r = &g_ep_rec_host->code;       // plain host reads: no CUDA call, works on a poisoned context
  const int code = r[0];
  if (info6) {
    for (int i = 0; i < 6; ++i) info6[i] = r[i];
  }
  if (clear) {
    for (int i = 0; i < 6; ++i) r[i] = 0;
  }
  if (code == kEpErrTimeout) return GLM4MOE_ERR_EP_TIMEOUT;
  if (code == kEpErrModeMismatch) return GLM4MOE_ERR_EP_MODE;
  return GLM4MOE_OK;
}

// ---- expert parallelism over peer memory (distributed.py layout "p2p") -------------------------
extern "C" int glm4moe_p2p_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
  if (!ptr || !handle64 || bytes == 0) return GLM4MOE_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  if (int rc = ep_record_ensure()) return rc;
  void* p = nullptr;
  GLM4MOE_CUDA_CHECK(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return static_cast<int>(e);
  }
  memcpy(handle64, &h, 64);
  *ptr = p;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_p2p_open(const unsigned char* handle64, void** peer_ptr) {
  if (!handle64 || !peer_ptr) return GLM4MOE_ERR_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  GLM4MOE_CUDA_CHECK(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_p2p_close(void* peer_ptr) {
  if (!peer_ptr) return GLM4MOE_ERR_ARG;
  GLM4MOE_CUDA_CHECK(cudaIpcCloseMemHandle(peer_ptr));
  return GLM4MOE_OK;
}

extern "C" int glm4moe_p2p_free(void* ptr) {
  if (!ptr) return GLM4MOE_ERR_ARG;
  GLM4MOE_CUDA_CHECK(cudaFree(ptr));
  return GLM4MOE_OK;
}

// sliced peer-memory combine: -1 = by size (default), 0 = never, 1 = whenever the layer has a staging region
static int g_ep_sliced_mode = -1;

extern "C" int glm4moe_set_ep_sliced(int mode) {
  if (mode < -1 || mode > 1) return GLM4MOE_ERR_ARG;
  g_ep_sliced_mode = mode;
  return GLM4MOE_OK;
}

extern "C" int glm4moe_layer_forward_ep_p2p(glm4moe_layer_t* l, const void* x, void* out, int T,
                                            void* const* peer_workspaces, int rank, int world, void* stream) {
  if (!l || !x || !out || !peer_workspaces) return GLM4MOE_ERR_ARG;
  const glm4moe_config_t& c = l->cfg;
  const int n_local = c.expert_end - c.expert_begin;
  if (world < 1 || world > kEpMaxRanks || rank < 0 || rank >= world || n_local <= 0 ||
      n_local * world != c.n_routed_experts || c.expert_begin != rank * n_local)
    return GLM4MOE_ERR_ARG;
  if (c.num_experts_per_tok > kEpMaxTopK || c.dtype == GLM4MOE_F32 || c.with_shared == 1 && world > 1)
    return GLM4MOE_ERR_UNSUPPORTED;
  if (peer_workspaces[rank] != l->ws) return GLM4MOE_ERR_ARG;
  if (T < 0 || T > c.max_tokens) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  const int launches0 = g_launches;
  EpCombineArgs a{};
  EpFastSignal sig{};
  sig.enabled = 1;
  sig.world = world;
  sig.rank = rank;
  sig.ctl = reinterpret_cast<int32_t*>(l->ws + l->L.ctl);
  for (int r = 0; r < world; ++r) {
    if (!peer_workspaces[r]) return GLM4MOE_ERR_ARG;
    a.ws[r] = static_cast<unsigned char*>(peer_workspaces[r]);
    sig.peer_ready[r] = reinterpret_cast<int32_t*>(a.ws[r] + l->L.ctl) + 16 + rank;
  }
  for (int i = 0; i < 2; ++i) {
    sig.y_alt[i] = l->ws + l->L.y_alt[i];
    a.off_y_alt[i] = l->L.y_alt[i];
    a.off_slot_row_alt[i] = l->L.slot_row_alt[i];
  }
  bool fast = false;
  int rc = forward_impl(l, x, nullptr, T, stream, nullptr, -1, -1, /*rows_only=*/true,
                        g_ep_fast ? &sig : nullptr, &fast);
  if (rc) return rc;
  a.fast = fast ? 1 : 0;
  a.off_y = l->L.y;
  a.off_slot_row = l->L.slot_row;
  a.off_idx = l->L.idx;
  a.off_ctl = l->L.ctl;
  a.rank = rank;
  a.world = world;
  a.tokens = T;
  a.k = c.num_experts_per_tok;
  a.H = c.hidden_size;
  a.experts_per_rank = n_local;
  a.with_shared = c.with_shared != 0;
  a.shared_per_rank = (T + world - 1) / world;
  if (c.with_shared == 1) a.shared_per_rank = T;   // world == 1: every shared row is local
  // Large batches: reduce the own token slice and store it to the peers (T/R as many NVLink reads)
  // instead of every rank pulling every row.  Same decision on every rank (depends on T only).
  a.off_stage = l->L.stage;
  const bool has_stage = world > 1 && c.with_shared == 2;
  const size_t row_bytes = static_cast<size_t>(T) * (a.k + 1) * a.H * dtype_size(c.dtype);
  a.sliced = g_ep_sliced_mode == 1 ? (has_stage || world == 1)      // world == 1: no peer, no staging needed
                                   : g_ep_sliced_mode == -1 && has_stage && row_bytes >= (16u << 20);
  if (fast) a.sliced = 0;      // the decode fast path is a pull (<= 32 tokens)
  rc = ep_combine_p2p_launch(a, out, c.dtype, stream);
  l->last_launches = g_launches - launches0;
  return rc;
}

// ---- the step in front of the block fused to it (SURVEY 8f-2 part i, 8f-4) ----------------------------------
// decode_layer.py:98-99: `hidden, residual = post_attention_layernorm(hidden, residual); hidden = mlp(hidden)` as ONE
// library call: add_rmsnorm_kernel writes the normalised tokens into the layer workspace (`xn`) and the block's
// launches follow in the same programmatic-dependent-launch chain (the norm is two launches old when the
// persistent decode kernel streams the shared expert against it before its own dependency wait).
extern "C" int glm4moe_layer_forward_prenorm(glm4moe_layer_t* l, const void* x, const void* residual,
                                             const void* norm_weight, float eps, void* residual_out, void* out,
                                             int T, void* stream) {
  if (!l || !x || !norm_weight || !out || (residual && !residual_out)) return GLM4MOE_ERR_ARG;
  const glm4moe_config_t& c = l->cfg;
  if (T < 0 || T > c.max_tokens) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  const int launches0 = g_launches;
  void* xn = l->ws + l->L.xn;
  int rc = glm4moe_add_rmsnorm(x, residual, norm_weight, eps, T, c.hidden_size, c.dtype, xn, residual_out, stream);
  if (rc) return rc;
  rc = forward_impl(l, xn, out, T, stream, nullptr);
  l->last_launches = g_launches - launches0;
  return rc;
}

extern "C" size_t glm4moe_ep_comm_bytes(int max_tokens, int H, int dtype, int world) {
  // the same limits ep_allreduce_launch applies, so that a buffer that can be sized can also be used
  if (max_tokens <= 0 || H <= 0 || H % 8 != 0 || H > GLM4MOE_EP_COMM_MAX_HIDDEN || world < 1 || world > kEpMaxRanks ||
      (dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16))
    return 0;
  return ep_comm_layout(max_tokens, H, dtype, world).total;
}

extern "C" int glm4moe_ep_allreduce_add_rmsnorm(void* const* peer_comm, int rank, int world, int max_tokens,
                                                const void* partial, const void* residual, const void* weight,
                                                float eps, int T, int H, int dtype, void* out, void* residual_out,
                                                void* stream) {
  if (!peer_comm || !partial || !out) return GLM4MOE_ERR_ARG;
  if (residual && (!weight || !residual_out)) return GLM4MOE_ERR_ARG;
  if (world < 1 || world > kEpMaxRanks || rank < 0 || rank >= world) return GLM4MOE_ERR_ARG;
  if (T < 0 || T > max_tokens || H <= 0 || H % 8 != 0) return GLM4MOE_ERR_ARG;
  if (dtype != GLM4MOE_BF16 && dtype != GLM4MOE_F16) return GLM4MOE_ERR_UNSUPPORTED;
  if (T == 0) return GLM4MOE_OK;
  EpArArgs a{};
  for (int r = 0; r < world; ++r) {
    if (!peer_comm[r]) return GLM4MOE_ERR_ARG;
    a.comm[r] = static_cast<unsigned char*>(peer_comm[r]);
  }
  a.L = ep_comm_layout(max_tokens, H, dtype, world);
  a.rank = rank;
  a.world = world;
  a.tokens = T;
  a.H = H;
  a.per = (T + world - 1) / world;
  a.eps = eps;
  a.partial = partial;
  a.residual = residual;
  a.weight = weight;
  a.out = out;
  a.residual_out = residual_out;
  // Same decision on every rank (T, H, world only).  One-shot while the pushed copies stay small: (R-1) x the
  // partial leaves every rank, against 3(R-1)/R x in the two-shot form, which pays a second flag round.
  const size_t bytes = static_cast<size_t>(T) * H * dtype_size(dtype);
  const bool fits = bytes <= a.L.os_slot && T <= kEpArOneShotRows;
  bool two = !fits || bytes * static_cast<size_t>(world - 1) > (2u << 20);
  if (g_ep_ar_mode == 0 && fits) two = false;
  if (g_ep_ar_mode == 1) two = true;
  return ep_allreduce_launch(a, dtype, two ? 1 : 0, stream);
}

extern "C" int glm4moe_layer_forward_ep_p2p_prenorm(glm4moe_layer_t* l, const void* partial, const void* residual,
                                                    const void* norm_weight, float eps, void* residual_out,
                                                    void* out, int T, void* const* peer_workspaces,
                                                    void* const* peer_comm, int comm_max_tokens, int rank,
                                                    int world, void* stream) {
  if (!l || !partial || !norm_weight || !out || !peer_workspaces || !peer_comm) return GLM4MOE_ERR_ARG;
  const glm4moe_config_t& c = l->cfg;
  if (T < 0 || T > c.max_tokens) return GLM4MOE_ERR_ARG;
  if (T == 0) return GLM4MOE_OK;
  const int launches0 = g_launches;
  void* xn = l->ws + l->L.xn;
  int rc = glm4moe_ep_allreduce_add_rmsnorm(peer_comm, rank, world, comm_max_tokens, partial, residual, norm_weight,
                                            eps, T, c.hidden_size, c.dtype, xn, residual_out, stream);
  if (rc) return rc;
  rc = glm4moe_layer_forward_ep_p2p(l, xn, out, T, peer_workspaces, rank, world, stream);
  l->last_launches = g_launches - launches0;
  return rc;
}
