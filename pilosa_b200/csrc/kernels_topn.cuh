// K7 — TopN: one source row x many candidate rows -> intersection counts.
// Replaces the hot loop of fragment.top (fragment.go:1599-1689): Src.intersectionCount(f.row(id))
// for every ranked row, i.e. roaring.intersectionCount (roaring.go:3078-3215) over 16
// container pairs per row. One CTA owns (shard, container column): the source column is
// staged once in shared memory as an 8 KiB bitmap, then each warp streams candidate
// containers against it (array: bit probes; bitmap: AND+popc; run: masked range popcount).
#pragma once
#include "kernels.cuh"

namespace pb200 {

struct TopnParams {
  const FamilyDev* families;
  uint32_t family;
  const uint64_t* shards;
  uint32_t n_shards;
  const uint8_t* src_dense;      // dense source rows: (s * 16 + c) * 8192 (never null: no-src TopN needs no kernel)
  // candidate rows: either per-shard rank lists (row indices into the fragment's row table)
  const uint32_t* const* rank_rowidx;  // [n_shards] device pointers (null when the shard has no fragment)
  const uint32_t* rank_n;              // [n_shards]
  const uint64_t* cand_off;            // [n_shards] offset of shard s in `counts`
  // ... or one explicit id list shared by all shards (pass 2, fragment.go:1573)
  const uint64_t* ids;
  uint32_t n_ids;
  uint32_t* counts;                    // [sum rank_n] or [n_shards * n_ids]
  unsigned long long* algo_bytes;
};

constexpr int kTopnThreads = 256;

__device__ __forceinline__ uint32_t popc_range_smem(const uint64_t* w, uint32_t s, uint32_t l) {  // inclusive [s, l]
  uint32_t ws = s >> 6, we = l >> 6;
  uint64_t ms = ~0ull << (s & 63), me = ~0ull >> (63 - (l & 63));
  if (ws == we) return __popcll(w[ws] & ms & me);
  uint32_t n = __popcll(w[ws] & ms) + __popcll(w[we] & me);
  for (uint32_t k = ws + 1; k < we; k++) n += __popcll(w[k]);
  return n;
}

__global__ void __launch_bounds__(kTopnThreads) topn_count_kernel(TopnParams p) {
  __shared__ __align__(16) uint64_t s_src[kBitmapWords];
  __shared__ uint32_t s_any;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t n_items = p.n_shards * kRowContainers;
  unsigned long long bytes_acc = 0;

  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t s = item >> 4, c = item & 15;
    const FamilyDev fam = p.families[p.family];
    const uint64_t shard = p.shards[s];
    __syncthreads();
    if (tid == 0) s_any = 0;
    __syncthreads();
    if (shard >= fam.cap) continue;
    const FragDev f = fam.frags[shard];
    if (f.nrows == 0) continue;
    // stage the source column; skip everything when it is empty (intersect with nothing)
    const uint4* src = reinterpret_cast<const uint4*>(p.src_dense + (uint64_t(s) * kRowContainers + c) * kBitmapBytes);
    uint32_t any = 0;
    for (int j = tid; j < kBitmapWords / 2; j += kTopnThreads) {
      uint4 v = ldg128(src + j);
      reinterpret_cast<uint4*>(s_src)[j] = v;
      any |= v.x | v.y | v.z | v.w;
    }
    if (any) s_any = 1;
    __syncthreads();
    if (!s_any) continue;

    const bool by_rank = p.ids == nullptr;
    const uint32_t n_cand = by_rank ? p.rank_n[s] : p.n_ids;
    const uint32_t* rowidx = by_rank ? p.rank_rowidx[s] : nullptr;
    uint32_t* out = p.counts + (by_rank ? p.cand_off[s] : uint64_t(s) * p.n_ids);

    for (uint32_t k = warp; k < n_cand; k += kTopnThreads / 32) {
      uint64_t idx;
      if (by_rank) { idx = rowidx[k]; if (idx == 0xffffffffull) continue; }
      else {
        const uint64_t id = p.ids[k];
        if (f.dense) { if (id >= f.nrows) continue; idx = id; }
        else {
          uint32_t lo = 0, hi = f.nrows;
          while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (f.row_ids[mid] < id) lo = mid + 1; else hi = mid; }
          if (lo >= f.nrows || f.row_ids[lo] != id) continue;
          idx = lo;
        }
      }
      const ContainerDesc d = f.rows[idx * kRowContainers + c];
      if (d.n == 0) continue;
      const uint32_t type = desc_type(d), len = desc_len(d);
      uint32_t cnt = 0;
      if (type == kArray) {  // intersectionCountArrayBitmap, roaring.go:3197
        const uint16_t* a = reinterpret_cast<const uint16_t*>(d.addr);
        // 8 values (16 B) per lane per step
        const uint32_t nvec = len >> 3;
        const uint4* a4 = reinterpret_cast<const uint4*>(a);
        for (uint32_t i = lane; i < nvec; i += 32) {
          uint4 v = ldg128_stream(a4 + i);
          uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int t = 0; t < 4; t++) {
            uint32_t x = w[t] & 0xffff, y = w[t] >> 16;
            cnt += uint32_t(s_src[x >> 6] >> (x & 63)) & 1u;
            cnt += uint32_t(s_src[y >> 6] >> (y & 63)) & 1u;
          }
        }
        for (uint32_t i = (nvec << 3) + lane; i < len; i += 32) { uint32_t x = a[i]; cnt += uint32_t(s_src[x >> 6] >> (x & 63)) & 1u; }
      } else if (type == kBitmap) {  // intersectionCountBitmapBitmap, roaring.go:3212
        const uint4* b = reinterpret_cast<const uint4*>(d.addr);
#pragma unroll 4
        for (int j = lane; j < kBitmapWords / 2; j += 32) {
          uint4 v = ldg128_stream(b + j);
          uint4 m = reinterpret_cast<const uint4*>(s_src)[j];
          cnt += popc128(and128(v, m));
        }
      } else {  // intersectionCountBitmapRun, roaring.go:3189 (bitmapCountRange per run)
        const ushort2* r = reinterpret_cast<const ushort2*>(d.addr);
        for (uint32_t i = lane; i < len; i += 32) { ushort2 iv = r[i]; cnt += popc_range_smem(s_src, iv.x, iv.y); }
      }
      cnt = warp_sum(cnt);
      if (lane == 0) {
        if (cnt) atomicAdd(out + k, cnt);
        bytes_acc += desc_algo_bytes(d);
      }
    }
    if (tid == 0) bytes_acc += kBitmapBytes;  // the source column
  }
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

}  // namespace pb200
