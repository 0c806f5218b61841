// K7 — TopN: one source row x many candidate rows -> intersection counts.
// Replaces the hot loop of fragment.top (fragment.go:1599-1689): Src.intersectionCount(f.row(id))
// for every ranked row, i.e. roaring.intersectionCount (roaring.go:3078-3215) over 16
// container pairs per row.
//
// One CTA owns (shard, chunk of candidate rows) and stages the WHOLE 128 KiB source row
// (16 container columns) in shared memory once. A half-warp then owns one candidate row:
// lane c handles container c, so the 16 descriptors of a row (256 B) and — because the
// loader stores a row's containers back to back — its payloads are read as contiguous
// streams, and the 16 per-container counts are combined with shuffles into one plain
// store (no atomics: a row belongs to exactly one half-warp). Small arrays / run lists
// (the bulk of a Zipf-distributed field) are probed lane-serially; containers that are
// large (bitmaps, long arrays, many runs) are handed to the whole warp afterwards.
#pragma once
#include "kernels.cuh"

namespace pb200 {

struct TopnParams {
  const FamilyDev* families;
  uint32_t family;
  const uint64_t* shards;
  uint32_t n_shards;
  const uint8_t* src_dense;            // dense source rows: (s * 16 + c) * 8192
  const uint32_t* src_n;               // cardinality per (s, c) of the source (skip empty columns)
  // candidate rows: per-shard rank tables (row indices into the fragment's row table) ...
  const uint32_t* const* rank_rowidx;  // [n_shards] device pointers (null when the shard has no fragment)
  const uint32_t* rank_n;              // [n_shards]
  const uint64_t* cand_off;            // [n_shards] offset of shard s in `counts`
  // ... or one explicit id list shared by all shards (pass 2, fragment.go:1573)
  const uint64_t* ids;
  uint32_t n_ids;
  uint32_t chunk;                      // candidates per CTA
  uint32_t chunks_per_shard;
  uint32_t* counts;                    // [sum rank_n] or [n_shards * n_ids]; every entry of the range is written
  unsigned long long* algo_bytes;
  // by-rank passes count the rank positions [range_lo, min(rank_n, range_cap, range_hi[s])) of every shard
  uint32_t range_lo, range_cap;        // range_cap == 0: no cap
  const uint32_t* range_hi;            // per shard (optional)
};

constexpr int kTopnThreads = 1024;
constexpr int kTopnSmallArray = 128; // arrays up to this many values are probed by a single lane
constexpr int kTopnSmallRuns = 8;
constexpr int kTopnIlp = 2;         // independent candidate rows per lane per iteration
constexpr int kTopnSmem = kRowContainers * kBitmapBytes;  // 128 KiB source row

__device__ __forceinline__ uint32_t popc_range_smem(const uint64_t* w, uint32_t s, uint32_t l) {  // inclusive [s, l]
  uint32_t ws = s >> 6, we = l >> 6;
  uint64_t ms = ~0ull << (s & 63), me = ~0ull >> (63 - (l & 63));
  if (ws == we) return __popcll(w[ws] & ms & me);
  uint32_t n = __popcll(w[ws] & ms) + __popcll(w[we] & me);
  for (uint32_t k = ws + 1; k < we; k++) n += __popcll(w[k]);
  return n;
}

// Quarter-warp (8 lanes) intersection count of one large container against a staged source
// column: 8 lanes x 16 B = one 128 B line per step, two steps in flight. Four containers are
// processed concurrently per warp. `ql` is the lane index inside the quarter (0..7); the
// result is valid in every lane of the quarter.
__device__ __forceinline__ uint32_t topn_big_q(const ContainerDesc& d, const uint64_t* src, int ql, bool active) {
  uint32_t cnt = 0;
  if (active) {
    const uint32_t type = desc_type(d), len = desc_len(d);
    const uint32_t* src32 = reinterpret_cast<const uint32_t*>(src);
    if (type == kArray) {  // intersectionCountArrayBitmap, roaring.go:3197 — 8 values per 128-bit load
      const uint4* a4 = reinterpret_cast<const uint4*>(d.addr);
      const uint32_t nvec = (len + 7) >> 3;  // the tail vector is padded by the loader; masked below
      for (uint32_t i = ql; i < nvec; i += 16) {
        const uint4 va = ldg128_stream(a4 + i);
        const bool two = i + 8 < nvec;
        const uint4 vb = two ? ldg128_stream(a4 + i + 8) : make_uint4(0, 0, 0, 0);
        const uint32_t wa[4] = {va.x, va.y, va.z, va.w}, wb[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
        for (int t = 0; t < 4; t++) {
          const uint32_t x = wa[t] & 0xffff, y = wa[t] >> 16, x2 = wb[t] & 0xffff, y2 = wb[t] >> 16;
          const uint32_t bx = (src32[x >> 5] >> (x & 31)) & 1u, by = (src32[y >> 5] >> (y & 31)) & 1u;
          const uint32_t bx2 = (src32[x2 >> 5] >> (x2 & 31)) & 1u, by2 = (src32[y2 >> 5] >> (y2 & 31)) & 1u;
          const uint32_t e = i * 8 + 2 * t, e2 = (i + 8) * 8 + 2 * t;
          cnt += (e < len ? bx : 0u) + (e + 1 < len ? by : 0u);
          cnt += (two && e2 < len ? bx2 : 0u) + (two && e2 + 1 < len ? by2 : 0u);
        }
      }
    } else if (type == kBitmap) {  // intersectionCountBitmapBitmap, roaring.go:3212
      const uint4* b = reinterpret_cast<const uint4*>(d.addr);
      const uint4* m4 = reinterpret_cast<const uint4*>(src);
#pragma unroll 4
      for (int j = ql; j < kBitmapWords / 2; j += 8) cnt += popc128(and128(ldg128_stream(b + j), m4[j]));
    } else {  // intersectionCountBitmapRun, roaring.go:3189 (bitmapCountRange per run)
      const ushort2* r = reinterpret_cast<const ushort2*>(d.addr);
      for (uint32_t i = ql; i < len; i += 8) { ushort2 iv = r[i]; cnt += popc_range_smem(src, iv.x, iv.y); }
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  return cnt;
}

__global__ void __launch_bounds__(kTopnThreads, 1) topn_count_kernel(TopnParams p) {
  extern __shared__ __align__(16) uint64_t s_src[];  // [16][1024]
  __shared__ uint32_t s_srcn[kRowContainers];
  const int tid = threadIdx.x, lane = tid & 31, hw = tid >> 4, sub = tid & 15;  // hw: half-warp index in the CTA
  const uint32_t n_items = p.n_shards * p.chunks_per_shard;
  unsigned long long bytes_acc = 0;

  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t s = item / p.chunks_per_shard, ch = item % p.chunks_per_shard;
    const bool by_rank = p.ids == nullptr;
    uint32_t n_cand = by_rank ? p.rank_n[s] : p.n_ids;
    if (by_rank) {
      if (p.range_cap) n_cand = min(n_cand, p.range_cap);
      if (p.range_hi) n_cand = min(n_cand, p.range_hi[s]);
    }
    const uint32_t k0 = (by_rank ? p.range_lo : 0u) + ch * p.chunk;
    if (k0 >= n_cand) continue;
    const uint32_t k1 = min(n_cand, k0 + p.chunk);
    uint32_t* out = p.counts + (by_rank ? p.cand_off[s] : uint64_t(s) * p.n_ids);
    const FamilyDev fam = p.families[p.family];
    const uint64_t shard = p.shards[s];
    FragDev f{};
    if (shard < fam.cap) f = fam.frags[shard];

    __syncthreads();  // previous item done with the staged source
    if (tid < kRowContainers) s_srcn[tid] = p.src_n[s * kRowContainers + tid];
    __syncthreads();
    uint32_t any = 0;
#pragma unroll
    for (int c = 0; c < kRowContainers; c++) any |= s_srcn[c];
    if (f.nrows == 0 || any == 0) {  // nothing can intersect: zero this chunk's counts
      for (uint32_t k = k0 + tid; k < k1; k += kTopnThreads) out[k] = 0;
      continue;
    }
    // stage the non-empty source columns
    for (int c = 0; c < kRowContainers; c++) {
      if (s_srcn[c] == 0) continue;
      const uint4* src = reinterpret_cast<const uint4*>(p.src_dense + (uint64_t(s) * kRowContainers + c) * kBitmapBytes);
      for (int j = tid; j < kBitmapWords / 2; j += kTopnThreads) reinterpret_cast<uint4*>(s_src + c * kBitmapWords)[j] = ldg128(src + j);
    }
    if (tid < kRowContainers && s_srcn[tid]) bytes_acc += kBitmapBytes;
    __syncthreads();

    const uint32_t* rowidx = by_rank ? p.rank_rowidx[s] : nullptr;
    const uint64_t* my_src = s_src + sub * kBitmapWords;
    const uint32_t* my_src32 = reinterpret_cast<const uint32_t*>(my_src);
    const bool col_live = s_srcn[sub] != 0;
    // Every warp iteration covers 2 x kTopnIlp candidate rows: one per half-warp, kTopnIlp
    // independent rows per lane so that the descriptor -> payload -> probe chains of
    // different rows overlap (the probes are shared-memory-latency bound otherwise).
    constexpr uint32_t kStride = kTopnThreads / 16;
    for (uint32_t kb = k0 + (hw & ~1u); kb < k1; kb += kStride * kTopnIlp) {
      ContainerDesc d[kTopnIlp];
      uint32_t kk[kTopnIlp], cnt[kTopnIlp];
      bool valid[kTopnIlp], small_arr[kTopnIlp], big[kTopnIlp];
      uint4 v0[kTopnIlp];
#pragma unroll
      for (int u = 0; u < kTopnIlp; u++) {
        kk[u] = kb + u * kStride + (hw & 1u);
        valid[u] = kk[u] < k1;
        uint64_t idx = ~0ull;
        if (valid[u]) {
          if (by_rank) { const uint32_t ri = rowidx[kk[u]]; if (ri != 0xffffffffu) idx = ri; }
          else {
            const uint64_t id = p.ids[kk[u]];
            if (f.dense) { if (id < f.nrows) idx = id; }
            else {
              uint32_t lo = 0, hi = f.nrows;
              while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (f.row_ids[mid] < id) lo = mid + 1; else hi = mid; }
              if (lo < f.nrows && f.row_ids[lo] == id) idx = lo;
            }
          }
        }
        d[u].addr = 0; d[u].n = 0; d[u].meta = 0;
        if (idx != ~0ull) d[u] = f.rows[idx * kRowContainers + sub];
        if (!col_live) d[u].n = 0;  // empty source column: no intersection, no loads
      }
#pragma unroll
      for (int u = 0; u < kTopnIlp; u++) {  // first 8 values of every small array: one 128-bit load each, all in flight
        const uint32_t type = desc_type(d[u]), len = desc_len(d[u]);
        small_arr[u] = d[u].n && type == kArray && len <= kTopnSmallArray;
        big[u] = d[u].n && !small_arr[u] && !(type == kRun && len <= kTopnSmallRuns);
        v0[u] = make_uint4(0, 0, 0, 0);
        if (small_arr[u]) v0[u] = ldg128_stream(reinterpret_cast<const uint4*>(d[u].addr));
        if (d[u].n) bytes_acc += desc_algo_bytes(d[u]);
        cnt[u] = 0;
      }
#pragma unroll
      for (int u = 0; u < kTopnIlp; u++) {
        const uint32_t len = small_arr[u] ? desc_len(d[u]) : 0;
        const uint32_t w[4] = {v0[u].x, v0[u].y, v0[u].z, v0[u].w};
#pragma unroll
        for (int t = 0; t < 4; t++) {  // unconditional 32-bit probes (any u16 is a valid index), masked afterwards
          const uint32_t x = w[t] & 0xffff, y = w[t] >> 16;
          const uint32_t bx = (my_src32[x >> 5] >> (x & 31)) & 1u, by = (my_src32[y >> 5] >> (y & 31)) & 1u;
          cnt[u] += (2 * t < len ? bx : 0u) + (2 * t + 1 < len ? by : 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < kTopnIlp; u++) {
        if (small_arr[u]) {
          const uint32_t len = desc_len(d[u]);
          const uint4* a4 = reinterpret_cast<const uint4*>(d[u].addr);
          for (uint32_t i = 8; i < len; i += 8) {  // payloads are 16 B aligned and padded by the loader
            const uint4 v = ldg128_stream(a4 + (i >> 3));
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int t = 0; t < 4; t++) {
              const uint32_t x = w[t] & 0xffff, y = w[t] >> 16;
              const uint32_t bx = (my_src32[x >> 5] >> (x & 31)) & 1u, by = (my_src32[y >> 5] >> (y & 31)) & 1u;
              cnt[u] += (i + 2 * t < len ? bx : 0u) + (i + 2 * t + 1 < len ? by : 0u);
            }
          }
        } else if (d[u].n && !big[u]) {  // short run list
          const ushort2* r = reinterpret_cast<const ushort2*>(d[u].addr);
          const uint32_t len = desc_len(d[u]);
          for (uint32_t i = 0; i < len; i++) { ushort2 iv = r[i]; cnt[u] += popc_range_smem(my_src, iv.x, iv.y); }
        }
      }
#pragma unroll
      for (int u = 0; u < kTopnIlp; u++) {
        // large containers: four at a time, one per quarter-warp
        uint32_t big_mask = __ballot_sync(0xffffffffu, big[u]);
        while (big_mask) {
          // the q-th lowest set bit of big_mask is served by quarter q
          uint32_t m = big_mask;
          int pick = -1;
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const int b = m ? __ffs(m) - 1 : -1;
            if (b >= 0) m &= m - 1;
            if (q == (lane >> 3)) pick = b;
          }
          big_mask = m;
          const int from = pick >= 0 ? pick : lane;
          ContainerDesc bd;
          bd.addr = __shfl_sync(0xffffffffu, d[u].addr, from);
          bd.n = __shfl_sync(0xffffffffu, d[u].n, from);
          bd.meta = __shfl_sync(0xffffffffu, d[u].meta, from);
          const uint32_t c = topn_big_q(bd, s_src + (from & 15) * kBitmapWords, lane & 7, pick >= 0);
          // route the quarter's result back to the lane that owns the container
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const int owner = __shfl_sync(0xffffffffu, pick, q * 8);
            const uint32_t val = __shfl_sync(0xffffffffu, c, q * 8);
            if (owner == lane) cnt[u] = val;
          }
        }
        // sum the 16 lanes of each half-warp
        uint32_t cc = cnt[u];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) cc += __shfl_xor_sync(0xffffffffu, cc, o);
        if (valid[u] && sub == 0) out[kk[u]] = cc;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bytes_acc += __shfl_xor_sync(0xffffffffu, bytes_acc, o);
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

// ---------------------------------------------------------------------------------------
// Pruning like fragment.top (fragment.go:1599-1689). The walk pushes the first n ranked rows whose intersection count
// is non-zero and at least MinThreshold; from then on its threshold is the smallest count pushed (later pushes are
// never smaller, so it stays), and it stops at the first row whose CACHED count is below it. Everything behind that
// row is never visited by the reference — and need not be counted here. One warp per shard replays the first part
// of the walk over the counts of the first `head` ranked rows and returns the rank position where the walk stops
// (`limit`): rows [head, limit) are counted by a second pass of topn_count_kernel, rows >= limit not at all.
//   rank_counts[s]: cached counts in rank order; counts: intersection counts (valid for positions < head)
// ---------------------------------------------------------------------------------------
__global__ void topn_limit_kernel(const uint64_t* const* __restrict__ rank_counts, const uint32_t* __restrict__ rank_n,
                                  const uint64_t* __restrict__ cand_off, const uint32_t* __restrict__ counts, uint32_t n_shards, uint32_t head,
                                  uint32_t n, uint64_t min_threshold, uint32_t* __restrict__ limit) {
  const uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (s >= n_shards) return;
  const uint32_t total = rank_n[s];
  const uint64_t* rc = rank_counts[s];
  const uint32_t* ic = counts + cand_off[s];
  const uint32_t h = min(head, total);
  // first part of the walk: which ranked rows get pushed (lanes test 32 rows at a time, in rank order)
  uint32_t pushed = 0, stop_at = 0xffffffffu;
  uint64_t thr = ~0ull;
  for (uint32_t base = 0; base < h && stop_at == 0xffffffffu; base += 32) {
    const uint32_t k = base + lane;
    bool push = false;
    uint64_t c = 0;
    if (k < h) {
      const uint64_t cached = rc[k];
      c = ic[k];
      push = cached != 0 && cached >= min_threshold && c != 0 && c >= min_threshold;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, push);
    const uint32_t need = n - pushed;
    uint32_t take = m;
    if (uint32_t(__popc(m)) >= need) {  // the n-th push happens in this group: keep only the first `need` of them
      uint32_t mm = m, kept = 0;
      for (uint32_t i = 0; i < need; i++) { const uint32_t b = mm & (0u - mm); kept |= b; mm ^= b; }
      take = kept;
      stop_at = base + 31u - uint32_t(__clz(kept));  // position of the n-th push
    }
    uint64_t mine = (take >> lane) & 1u ? c : ~0ull;
#pragma unroll
    for (int o = 16; o; o >>= 1) { const uint64_t y = __shfl_xor_sync(0xffffffffu, mine, o); mine = y < mine ? y : mine; }
    thr = mine < thr ? mine : thr;
    pushed += uint32_t(__popc(take));
  }
  uint32_t lim = total;
  if (stop_at != 0xffffffffu) {
    // n rows are in the heap: the walk continues while cached >= thr (and thr >= MinThreshold, which holds by construction)
    lim = total;
    for (uint32_t base = stop_at + 1; base < total; base += 32) {
      const uint32_t k = base + lane;
      // rows whose cached count is zero or under MinThreshold are skipped by the walk, they do not end it
      const bool below = k < total && rc[k] != 0 && rc[k] >= min_threshold && rc[k] < thr;
      const uint32_t m = __ballot_sync(0xffffffffu, below);
      if (m) { lim = base + uint32_t(__ffs(m)) - 1u; break; }
    }
  }
  if (lane == 0) limit[s] = lim;
}

// counts of the visited prefixes, packed: out[out_off[s] + k] = counts[cand_off[s] + k] for k < limit[s]
__global__ void topn_compact_kernel(const uint32_t* __restrict__ counts, const uint64_t* __restrict__ cand_off, const uint32_t* __restrict__ limit,
                                    const uint64_t* __restrict__ out_off, uint32_t n_shards, uint32_t* __restrict__ out) {
  for (uint32_t s = blockIdx.x; s < n_shards; s += gridDim.x) {
    const uint32_t* src = counts + cand_off[s];
    uint32_t* dst = out + out_off[s];
    const uint32_t n = limit[s];
    for (uint32_t k = threadIdx.x; k < n; k += blockDim.x) dst[k] = src[k];
  }
}

}  // namespace pb200
