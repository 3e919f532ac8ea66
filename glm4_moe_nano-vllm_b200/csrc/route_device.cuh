// Device-side pieces shared by the stand-alone router / permute kernels (router.cu, permute.cu) and the
// fused decode front-end (route_small.cu): the per-token expert selection (moe.py:37-46) and the
// single-block permutation (moe.py:135-147).
#pragma once
#include "common.cuh"

namespace glm4moe {

constexpr int kMaxSlots = 8;   // E <= 256
constexpr int kMaxTopK = 32;

// One warp selects the experts of one token.  `logit_row` may have been written earlier in the same
// kernel by other blocks (route_small.cu), hence the L2-coherent loads.
template <typename T>
__device__ __forceinline__ void select_token(const T* logit_row, const float* __restrict__ bias, int E, int k,
                                             float scale, int lane, int32_t* __restrict__ idx_row,
                                             float* __restrict__ w_row) {
  // choice (moe.py:38, fp32) is held as an order-preserving 32-bit key so a round's arg-max is two
  // warp reductions (redux.sync) instead of fifteen shuffles; slots past E get key 0 (< key(-inf)).
  uint32_t key[kMaxSlots];
  float score[kMaxSlots];
#pragma unroll
  for (int j = 0; j < kMaxSlots; ++j) {
    const int e = j * 32 + lane;
    if (e < E) {
      const float lg = Dt<T>::to_f(__ldcg(logit_row + e));
      const float s = Dt<T>::round(sigmoid_f(lg));       // moe.py:37, model dtype
      score[j] = s;
      const uint32_t u = __float_as_uint(s + bias[e]);   // moe.py:38, fp32
      key[j] = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    } else {
      score[j] = 0.f;
      key[j] = 0u;
    }
  }
  // k rounds of arg-max; ties -> lower expert id (the lane-local scan is ascending in e for a fixed
  // lane; across lanes the smallest expert id among the holders of the maximum wins).
  int my_e = -1;          // lane r keeps the r-th pick
  float my_s = 0.f;
  unsigned taken = 0;     // bit j set: my slot j already picked
  for (int r = 0; r < k; ++r) {
    uint32_t bk = 0u;
    int be = 0x7fffffff;
    float bs = 0.f;
#pragma unroll
    for (int j = 0; j < kMaxSlots; ++j) {
      const int e = j * 32 + lane;
      if (e < E && !((taken >> j) & 1u) && (key[j] > bk || be == 0x7fffffff)) {
        bk = key[j];
        be = e;
        bs = score[j];
      }
    }
    const uint32_t mk = __reduce_max_sync(0xffffffffu, bk);
    const int we = __reduce_min_sync(0xffffffffu, (bk == mk && be != 0x7fffffff) ? be : 0x7fffffff);
    const float ws = __shfl_sync(0xffffffffu, bs, we & 31);
    if ((we & 31) == lane) taken |= 1u << (we >> 5);
    if (lane == r) { my_e = we; my_s = ws; }
  }
  // renormalise: sum in fp32 over the k picks, every intermediate rounded to the model dtype
  float sum = 0.f;
  for (int r = 0; r < k; ++r) sum += __shfl_sync(0xffffffffu, my_s, r);
  const float denom = Dt<T>::round(Dt<T>::round(sum) + 1e-20f);   // moe.py:43
  float wgt = Dt<T>::round(my_s / denom);                          // moe.py:44
  wgt = Dt<T>::round(wgt * scale);                                 // moe.py:46
  // canonical slot order: ascending expert id (the order the reference accumulates in)
  int rank = 0;
  for (int r = 0; r < k; ++r) {
    const int oe = __shfl_sync(0xffffffffu, my_e, r);
    rank += (lane < k && oe < my_e) ? 1 : 0;
  }
  if (lane < k) {
    idx_row[rank] = my_e;
    w_row[rank] = wgt;
  }
}

constexpr int kTokPerBlock = 256;
constexpr int kWordsPerExpert = kTokPerBlock / 32;
constexpr int kMaxExperts = 256;

// Rows [r0, r0 + n) of stack entry `e` as ceil(n / tile_rows) tiles of near-equal size (multiples of
// 16 rows where the tile size allows: the tensor-core N granularity), written from tiles[ti].
// .w = index inside the expert | tiles of the expert << 16 (work ordering, umma_common.cuh).
__device__ __forceinline__ void emit_tiles(int4* __restrict__ tiles, int ti, int e, int r0, int n, int tile_rows) {
  const int nt = (n + tile_rows - 1) / tile_rows;
  if (nt == 0) return;
  const int gran = tile_rows < 16 ? tile_rows : 16;
  int per = ((n + nt - 1) / nt + gran - 1) / gran * gran;
  if (per > tile_rows) per = tile_rows;
  const int packed_nt = nt < 32768 ? nt : 0;     // does not fit the packing: plain per-tile order
  for (int j = 0, r = 0; j < nt; ++j, r += per)
    tiles[ti + j] = make_int4(e, r0 + r, min(per, n - r), packed_nt ? (j | (packed_nt << 16)) : 0);
}

// Decode-size batches (T <= 256): rank + scan + scatter in ONE block of 256 threads, thread t owns
// token t.  `idx` / `w` may have been written earlier in the same kernel (L2-coherent loads).
struct PermuteSmallSmem {
  uint32_t bitmap[kMaxExperts * kWordsPerExpert];
  int cnt[kMaxExperts];
  int off[kMaxExperts + 1];
  int tile_off[kMaxExperts + 1];
};

__device__ __forceinline__ void permute_small_body(
    PermuteSmallSmem& sm, const int32_t* idx, const float* w, int T, int k, int E, int e_lo, int e_hi,
    int tile_rows, int sh_lo, int sh_hi, int32_t* __restrict__ expert_offsets, int32_t* __restrict__ row_token,
    float* __restrict__ row_weight, int32_t* __restrict__ slot_row, int4* __restrict__ tiles, int tile_capacity,
    int32_t* __restrict__ num_tiles) {
  const int tid = threadIdx.x;
  const int word = tid >> 5;
  const uint32_t bit = 1u << (tid & 31);
  if (tid < T) {
    for (int s = 0; s < k; ++s) {
      const int e = __ldcg(idx + static_cast<size_t>(tid) * k + s);
      if (e >= e_lo && e < e_hi) atomicOr(&sm.bitmap[e * kWordsPerExpert + word], bit);
    }
  }
  __syncthreads();
  for (int e = tid; e < E; e += kTokPerBlock) {
    int c = 0;
#pragma unroll
    for (int wd = 0; wd < kWordsPerExpert; ++wd) c += __popc(sm.bitmap[e * kWordsPerExpert + wd]);
    sm.cnt[e] = c;
  }
  __syncthreads();
  if (tid < 32) {
    // warp scan over experts: lane handles ceil(E/32) consecutive experts
    const int per = (E + 31) / 32;
    int rows = 0, tl = 0;
    for (int i = 0; i < per; ++i) {
      const int e = tid * per + i;
      if (e < E) { rows += sm.cnt[e]; tl += (sm.cnt[e] + tile_rows - 1) / tile_rows; }
    }
    int rows_incl = rows, tl_incl = tl;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int a = __shfl_up_sync(0xffffffffu, rows_incl, o);
      const int b = __shfl_up_sync(0xffffffffu, tl_incl, o);
      if (tid >= o) { rows_incl += a; tl_incl += b; }
    }
    int roff = rows_incl - rows, toff = tl_incl - tl;
    for (int i = 0; i < per; ++i) {
      const int e = tid * per + i;
      if (e < E) {
        sm.off[e] = roff; sm.tile_off[e] = toff;
        roff += sm.cnt[e]; toff += (sm.cnt[e] + tile_rows - 1) / tile_rows;
      }
    }
    if (tid == 31) { sm.off[E] = rows_incl; sm.tile_off[E] = tl_incl; }
  }
  __syncthreads();
  const int shared_tiles = (sh_hi - sh_lo + tile_rows - 1) / tile_rows;
  const bool fits = sm.tile_off[E] + shared_tiles <= tile_capacity;
  if (tid == 0) *num_tiles = fits ? sm.tile_off[E] + shared_tiles : 0;
  for (int e = tid; e <= E; e += kTokPerBlock) expert_offsets[e] = sm.off[e];
  if (fits) {
    for (int e = tid; e < E; e += kTokPerBlock)
      emit_tiles(tiles, sm.tile_off[e], e - e_lo, sm.off[e], sm.cnt[e], tile_rows);
    if (tid == kTokPerBlock - 1)
      emit_tiles(tiles, sm.tile_off[E], e_hi - e_lo, T * k + sh_lo, sh_hi - sh_lo, tile_rows);
  }
  if (tid < T) {
    for (int s = 0; s < k; ++s) {
      const size_t i = static_cast<size_t>(tid) * k + s;
      const int e = __ldcg(idx + i);
      int row = -1;
      if (e >= e_lo && e < e_hi) {
        const uint32_t* bm = &sm.bitmap[e * kWordsPerExpert];
        int r = __popc(bm[word] & (bit - 1u));
        for (int wd = 0; wd < word; ++wd) r += __popc(bm[wd]);
        row = sm.off[e] + r;
        row_token[row] = tid;
        row_weight[row] = __ldcg(w + i);
      }
      slot_row[i] = row;
    }
  }
}

// The same bookkeeping for T * k <= 32 (token, slot) pairs by ONE warp (lane = pair), no shared memory: used
// by the router front-end merged into the persistent decode kernel (decode_fused.cu, T <= 2).  Output
// identical to permute_small_body.
__device__ __forceinline__ void permute_tiny_warp(
    const int32_t* idx, const float* w, int T, int k, int E, int e_lo, int e_hi, int tile_rows, int sh_lo, int sh_hi,
    int32_t* __restrict__ expert_offsets, int32_t* __restrict__ row_token, float* __restrict__ row_weight,
    int32_t* __restrict__ slot_row, int4* __restrict__ tiles, int tile_capacity, int32_t* __restrict__ num_tiles,
    int lane) {
  const int n = T * k;
  int e = -1;
  float wt = 0.f;
  const int t = lane / k;
  if (lane < n) {
    e = __ldcg(idx + lane);
    wt = __ldcg(w + lane);
  }
  const bool valid = lane < n && e >= e_lo && e < e_hi;
  int row = 0, cnt = 0, earlier_same = 0;
  for (int j = 0; j < n; ++j) {
    const int ej = __shfl_sync(0xffffffffu, e, j);
    const int vj = __shfl_sync(0xffffffffu, valid ? 1 : 0, j);
    const int tj = j / k;
    if (vj) {
      if (ej < e || (ej == e && tj < t)) ++row;
      if (ej == e) {
        ++cnt;
        if (tj < t) ++earlier_same;
      }
    }
  }
  const bool first = valid && earlier_same == 0;      // one lane per hit expert
  // tile index of a hit expert = hit experts below it
  int tile_i = 0, n_hit = 0;
  for (int j = 0; j < n; ++j) {
    const int ej = __shfl_sync(0xffffffffu, e, j);
    const int fj = __shfl_sync(0xffffffffu, first ? 1 : 0, j);
    if (fj) {
      ++n_hit;
      if (ej < e) ++tile_i;
    }
  }
  // rows of an expert never exceed T <= tile_rows: one tile per hit expert
  const int shared_tiles = (sh_hi - sh_lo + tile_rows - 1) / tile_rows;
  const bool fits = n_hit + shared_tiles <= tile_capacity;
  if (lane == 0) *num_tiles = fits ? n_hit + shared_tiles : 0;
  // expert_offsets[el] = valid pairs with a smaller expert id
  for (int base = 0; base <= E; base += 32) {           // warp-uniform trip count (shuffles inside)
    const int el = base + lane;
    int off = 0;
    for (int j = 0; j < n; ++j) {
      const int ej = __shfl_sync(0xffffffffu, e, j);
      const int vj = __shfl_sync(0xffffffffu, valid ? 1 : 0, j);
      if (vj && ej < el) ++off;
    }
    if (el <= E) expert_offsets[el] = off;
  }
  if (fits) {
    if (first) emit_tiles(tiles, tile_i, e - e_lo, row, cnt, tile_rows);
    if (lane == 31) emit_tiles(tiles, n_hit, e_hi - e_lo, T * k + sh_lo, sh_hi - sh_lo, tile_rows);
  }
  if (lane < n) {
    if (valid) {
      row_token[row] = t;
      row_weight[row] = wt;
    }
    slot_row[lane] = valid ? row : -1;
  }
}

}  // namespace glm4moe
