// K8/K9/K10 — bit-sliced index kernels. BSI fragments store row 0 = exists (E), row 1 = sign
// (S), row 2+i = magnitude bit i (fragment.go:90-93, sign-magnitude). Every step of
// fragment.sum / min / max / rangeOp (fragment.go:1109-1534) is a word-parallel boolean
// function of (E, S, B_i, filter/keep state), so one pass over the planes with the state
// held in registers replaces ~3 x bitDepth materialised Row operations per shard.
#pragma once
#include "kernels.cuh"
#include "tma.cuh"

namespace pb200 {

constexpr int kMaxDepth = 64;

struct BsiParams {
  const FamilyDev* families;
  uint32_t family;       // bsig_ view family
  uint32_t bit_depth;
  const uint64_t* shards;
  uint32_t n_shards;
  // optional filter row (already materialised or a plain row of another family)
  uint32_t has_filter;
  Leaf filter;
  const uint8_t* temp_base;
  // range
  uint32_t bsi_op;
  int64_t a, b;
  uint8_t* dense_out;    // range result rows: (s * 16 + c) * 8192 (null: count only)
  unsigned long long* count_out;  // optional: count_out[s * count_sstride] += |result| (Count(Row(v <op> x)) without materialising)
  uint32_t count_sstride;
  // sum
  uint32_t strict;
  unsigned long long* sum_acc;  // [0] = count, [1 + 2i] = popcount(B_i & prow), [2 + 2i] = popcount(B_i & nrow)
  // min / max
  uint32_t is_max;
  long long* mm_val;              // [n_shards]
  unsigned long long* mm_count;   // [n_shards]
  unsigned long long* algo_bytes;
};

__device__ __forceinline__ ContainerDesc bsi_row_desc(const FragDev& f, uint64_t row, uint32_t c) {
  ContainerDesc d; d.addr = 0; d.n = 0; d.meta = 0;
  if (f.nrows == 0) return d;
  uint64_t idx;
  if (f.dense) { if (row >= f.nrows) return d; idx = row; }
  else {
    uint32_t lo = 0, hi = f.nrows;
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (f.row_ids[mid] < row) lo = mid + 1; else hi = mid; }
    if (lo >= f.nrows || f.row_ids[lo] != row) return d;
    idx = lo;
  }
  return f.rows[idx * kRowContainers + c];
}

__device__ __forceinline__ ContainerDesc filter_desc(const BsiParams& p, uint32_t s, uint32_t c) {
  EvalParams ep{};
  ep.families = p.families; ep.n_families = 0xffffffffu; ep.shards = p.shards; ep.n_shards = p.n_shards; ep.temp_base = p.temp_base;
  return resolve_leaf(ep, p.filter, s, c);
}

// ---------------------------------------------------------------------------------------
// K10: range comparator -> dense result rows. State machines follow fragment.go exactly:
// rangeEQ :1288, rangeNEQ :1314, rangeLT :1331, rangeLTUnsigned :1357, rangeGT :1400,
// rangeGTUnsigned :1423, rangeBetween :1463, rangeBetweenUnsigned :1496, notNull :1456.
// ---------------------------------------------------------------------------------------
enum MachKind : int { M_NONE = 0, M_EQ, M_LTU, M_GTU, M_BETWEENU };

// `kind` is a compile-time constant in the pipe kernel (KIND != M_DYNAMIC), so the unused keep rows
// and branches disappear and the comparator fits 3 CTAs per SM; M_DYNAMIC keeps it a run-time field.
constexpr int M_DYNAMIC = -1;
template <int PER, int KIND = M_DYNAMIC>
struct Mach {
  int kind_rt = M_NONE;
  unsigned long long p1 = 0, p2 = 0;
  bool allow_eq = false, leading = true, done = false;
  uint4 f[PER], k1[PER], k2[PER];

  __device__ __forceinline__ void init(int kd, const uint4 (&filter)[PER], unsigned long long a, unsigned long long b, bool eq) {
    kind_rt = kd; p1 = a; p2 = b; allow_eq = eq; leading = true; done = false;
#pragma unroll
    for (int k = 0; k < PER; k++) { f[k] = filter[k]; k1[k] = make_uint4(0, 0, 0, 0); k2[k] = make_uint4(0, 0, 0, 0); }
  }
  // one plane, MSB -> LSB; on completion the result is in f[]
  __device__ __forceinline__ void step(int i, const uint4 (&row)[PER]) {
    const int kind = KIND == M_DYNAMIC ? kind_rt : KIND;
    if (kind == M_NONE || done) return;
    const bool bit1 = (p1 >> i) & 1ull;
    if (kind == M_EQ) {
#pragma unroll
      for (int k = 0; k < PER; k++) f[k] = bit1 ? and128(f[k], row[k]) : andnot128(f[k], row[k]);
    } else if (kind == M_LTU) {
      if (leading) {
        if (!bit1) {
#pragma unroll
          for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], row[k]);
          return;
        }
        leading = false;
      }
      if (i == 0 && !allow_eq) {
        if (!bit1) {
#pragma unroll
          for (int k = 0; k < PER; k++) f[k] = k1[k];  // return keep
        } else {
#pragma unroll
          for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(row[k], k1[k]));
        }
        done = true;
        return;
      }
      if (!bit1) {
#pragma unroll
        for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(row[k], k1[k]));
        return;
      }
      if (i > 0) {
#pragma unroll
        for (int k = 0; k < PER; k++) k1[k] = or128(k1[k], andnot128(f[k], row[k]));
      }
    } else if (kind == M_GTU) {
      if (i == 0 && !allow_eq) {
        if (bit1) {
#pragma unroll
          for (int k = 0; k < PER; k++) f[k] = k1[k];
        } else {
#pragma unroll
          for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(andnot128(f[k], row[k]), k1[k]));
        }
        done = true;
        return;
      }
      if (bit1) {
#pragma unroll
        for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(andnot128(f[k], row[k]), k1[k]));
        return;
      }
      if (i > 0) {
#pragma unroll
        for (int k = 0; k < PER; k++) k1[k] = or128(k1[k], and128(f[k], row[k]));
      }
    } else {  // M_BETWEENU: p1 = min (GTE side, keep1), p2 = max (LTE side, keep2)
      const bool bit2 = (p2 >> i) & 1ull;
      if (bit1) {
#pragma unroll
        for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(andnot128(f[k], row[k]), k1[k]));
      } else if (i > 0) {
#pragma unroll
        for (int k = 0; k < PER; k++) k1[k] = or128(k1[k], and128(f[k], row[k]));
      }
      if (!bit2) {
#pragma unroll
        for (int k = 0; k < PER; k++) f[k] = andnot128(f[k], andnot128(row[k], k2[k]));
      } else if (i > 0) {
#pragma unroll
        for (int k = 0; k < PER; k++) k2[k] = or128(k2[k], andnot128(f[k], row[k]));
      }
    }
  }
};

__device__ __forceinline__ unsigned long long uabs64(long long v) { return v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v; }

// ---------------------------------------------------------------------------------------
// K9: Min / Max (fragment.go:1145-1168 min, :1189-1210 max, :1171-1186 minUnsigned,
// :1213-1225 maxUnsigned). The reference descends the bit planes MSB -> LSB over the whole
// shard, one Row op + Count per plane. (value, count) of an extreme is a plain set function,
// so it decomposes: every SLICE of columns (what one consumer warp holds of a container
// column: 8192 columns) finds its own extreme and how many columns attain it, and slices
// combine by "better value wins, equal values add their counts". The per-plane decision is
// therefore a warp vote instead of a block-wide reduction over the shard. Both searches the
// sign split can ask for run side by side:
//   A = maxUnsigned over {negatives for Min | positives for Max}   (wins if any such column exists)
//   B = minUnsigned over {positives for Min | negatives for Max}
// bsi_minmax_kernel writes one record per (shard, container column); bsi_minmax_combine_kernel
// folds the 16 records of a shard.
// ---------------------------------------------------------------------------------------
struct MmRec { unsigned long long magA, cntA, magB, cntB; };

__device__ __forceinline__ bool mm_nz(const uint4& v) { return (v.x | v.y | v.z | v.w) != 0; }

// One thread per shard: fold the 16 column records, apply the sign rule of fragment.min / max.
__global__ void bsi_minmax_combine_kernel(BsiParams p, const MmRec* __restrict__ recs) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= p.n_shards) return;
  MmRec r{0, 0, ~0ull, 0};
  for (int c = 0; c < kRowContainers; c++) {
    const MmRec x = recs[s * kRowContainers + c];
    if (x.cntA) { if (x.magA > r.magA || !r.cntA) { r.magA = x.magA; r.cntA = x.cntA; } else if (x.magA == r.magA) r.cntA += x.cntA; }
    if (x.cntB) { if (x.magB < r.magB || !r.cntB) { r.magB = x.magB; r.cntB = x.cntB; } else if (x.magB == r.magB) r.cntB += x.cntB; }
  }
  long long val = 0;
  unsigned long long count = 0;
  if (p.bit_depth != 0) {  // with no planes the reference's loops never assign count (fragment.go:1171,1213): (0, 0)
    if (r.cntA) { val = p.is_max ? (long long)r.magA : -(long long)r.magA; count = r.cntA; }
    else if (r.cntB) { val = p.is_max ? -(long long)r.magB : (long long)r.magB; count = r.cntB; }
  }
  p.mm_val[s] = val;
  p.mm_count[s] = count;
}

// ---------------------------------------------------------------------------------------
// Plane pipe (K8 / K9 / K10). The container columns a BSI kernel consumes one after the other
// (E, S, filter, B_i ...) are 8 KiB contiguous blocks, so they are moved by the bulk copy
// engine (TMA, SASS UBLKCP) instead of by the threads that compute on them. One PRODUCER warp
// issues the cp.async.bulk copies of an item's planes into a ring of kPipeStages 8 KiB
// slots, gated per slot by an `empty` mbarrier; eight CONSUMER warps wait on the slot's
// `full` mbarrier (transaction bytes), take their 2 x 16 B, and release the slot with one
// mbarrier arrive per warp. No block-wide barrier separates planes, so the warps drift by
// up to kPipeStages planes and kPipeStages x 8 KiB per CTA stay in flight whatever the
// consumers' register footprint is. Items with an array / run plane (sparse BSI data) take
// a synchronous path among the consumers (named barrier 1).
// ---------------------------------------------------------------------------------------
constexpr int kPipeStages = 7;  // 3 CTAs / SM: 3 x (7 slots + 1 staging window) x 8 KiB + static < 227 KiB
constexpr int kPipeConsumers = 256;                 // 8 warps x 32 lanes x 2 uint4 = 8 KiB
constexpr int kPipeThreads = kPipeConsumers + 32;   // + the producer warp

__device__ __forceinline__ void consumers_sync() { asm volatile("bar.sync 1, %0;" :: "n"(kPipeConsumers) : "memory"); }

struct PlanePipe {
  uint8_t* ring;
  uint64_t* full;    // [kPipeStages], count 1 + tx bytes
  uint64_t* empty;   // [kPipeStages], count = consumer warps
  // position in the ring, advanced alike by the producer (planes issued) and the consumers (planes taken):
  // slot index, lap parity, and whether the first lap is over (the producer need not wait for `empty` before)
  uint32_t slot = 0, lap = 0;
  bool lapped = false;
  __device__ __forceinline__ void advance() {
    if (++slot == kPipeStages) { slot = 0; lap ^= 1u; lapped = true; }
  }

  __device__ __forceinline__ static bool by_tma(const ContainerDesc& d) { return d.n != 0 && desc_type(d) == kBitmap; }

  __device__ __forceinline__ void init(int tid) {
    if (tid == 0) {
      for (int i = 0; i < kPipeStages; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], kPipeConsumers / 32); }
      mbar_fence_init();
    }
  }
  // producer lane: copy plane d into the next slot once its previous contents have been released
  __device__ __forceinline__ void produce(const ContainerDesc& d) {
    if (!by_tma(d)) return;
    if (lapped) mbar_wait(&empty[slot], lap ^ 1u);  // the slot's previous use has been released by all consumer warps
    fence_proxy_async();
    mbar_arrive_expect_tx(&full[slot], kBitmapBytes);
    tma_load_1d(ring + slot * kBitmapBytes, reinterpret_cast<const void*>(d.addr), kBitmapBytes, &full[slot]);
    advance();
  }
  // consumer thread t (0..255): its two chunks of a streamable plane
  __device__ __forceinline__ void take(const ContainerDesc& d, int t, uint4 (&v)[2]) {
    if (d.n == 0) { v[0] = v[1] = make_uint4(0, 0, 0, 0); return; }
    if (desc_type(d) != kBitmap) { v[0] = v[1] = make_uint4(~0u, ~0u, ~0u, ~0u); return; }  // full run
    mbar_wait(&full[slot], lap);
    const uint4* win = reinterpret_cast<const uint4*>(ring + slot * kBitmapBytes);
    v[0] = win[t]; v[1] = win[t + kPipeConsumers];
    __syncwarp();
    if ((t & 31) == 0) mbar_arrive(&empty[slot]);
    advance();
  }
};

// any plane on the synchronous path (consumers only): array / run containers are expanded into `win` first
__device__ __forceinline__ void pipe_plane_sync(const ContainerDesc& d, int t, uint64_t* win, uint4 (&v)[2]) {
  if (d.n == 0) { v[0] = v[1] = make_uint4(0, 0, 0, 0); return; }
  if (d.n == kFullN) { v[0] = v[1] = make_uint4(~0u, ~0u, ~0u, ~0u); return; }
  if (desc_type(d) == kBitmap) {
    v[0] = ldg128_stream(reinterpret_cast<const uint4*>(d.addr) + t);
    v[1] = ldg128_stream(reinterpret_cast<const uint4*>(d.addr) + t + kPipeConsumers);
    return;
  }
  consumers_sync();  // window free
  reinterpret_cast<uint4*>(win)[t] = make_uint4(0, 0, 0, 0);
  reinterpret_cast<uint4*>(win)[t + kPipeConsumers] = make_uint4(0, 0, 0, 0);
  consumers_sync();
  stage_container(d, reinterpret_cast<uint32_t*>(win), 0, kBitmapWords, t, kPipeConsumers);
  consumers_sync();
  v[0] = reinterpret_cast<const uint4*>(win)[t]; v[1] = reinterpret_cast<const uint4*>(win)[t + kPipeConsumers];
}

// Item feed: the producer warp also resolves the descriptor table of every item (three
// dependent global loads) one item AHEAD of the consumers, double-buffered and handed over
// through mbarriers, so the pipe never drains at an item boundary and the consumer warps
// touch global memory only for results.
struct ItemFeed {
  static constexpr int kMaxPlanes = kMaxDepth + 3;
  ContainerDesc (*pd)[kMaxPlanes];  // [2][kMaxPlanes], planes in consumption order
  int* mode;                        // [2]: 0 skip, 1 streamed through the pipe, 2 synchronous path
  uint64_t* ready;                  // [2], count 32 (producer lanes)
  uint64_t* free_;                  // [2], count = consumer warps
  __device__ __forceinline__ void init(int tid) {
    if (tid == 0) {
      for (int i = 0; i < 2; i++) { mbar_init(&ready[i], 32); mbar_init(&free_[i], kPipeConsumers / 32); }
      mbar_fence_init();
    }
  }
  // producer warp: wait until buffer (it & 1) is free again
  __device__ __forceinline__ int begin_fill(uint32_t it) {
    const int buf = it & 1; const uint32_t use = it >> 1;
    if (use) mbar_wait(&free_[buf], (use - 1) & 1u);
    return buf;
  }
  __device__ __forceinline__ void publish(int buf, int m, int lane) {
    if (lane == 0) mode[buf] = m;
    __syncwarp();
    mbar_arrive(&ready[buf]);
  }
  // consumer threads
  __device__ __forceinline__ int acquire(uint32_t it) {
    const int buf = it & 1;
    mbar_wait(&ready[buf], (it >> 1) & 1u);
    return buf;
  }
  __device__ __forceinline__ void release(int buf, int lane) {
    __syncwarp();
    if (lane == 0) mbar_arrive(&free_[buf]);
  }
};

// K8: Sum. consider = E [& filter]; count = |consider|; nrow = S (unfiltered: the reference
// quirk, fragment.go:1118,1134; `strict` masks it with consider); prow = consider \ nrow;
// per plane popcount(B_i & prow), popcount(B_i & nrow), accumulated per consumer warp in shared
// memory and flushed with 2*depth+1 atomics per CTA at the end.
// Planes in consumption order: E, [filter], S, B_0 .. B_{d-1}.
__global__ void __launch_bounds__(kPipeThreads) bsi_sum_kernel(BsiParams p) {
  extern __shared__ __align__(128) uint8_t s_ring[];  // kPipeStages slots, then an 8 KiB staging window
  __shared__ __align__(8) uint64_t s_full[kPipeStages], s_empty[kPipeStages], s_ready[2], s_free[2];
  __shared__ ContainerDesc s_pd[2][ItemFeed::kMaxPlanes];
  __shared__ int s_mode[2];
  __shared__ uint32_t s_acc[kPipeConsumers / 32][2 * kMaxDepth + 1];  // per consumer warp: no atomics
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint64_t* s_win = reinterpret_cast<uint64_t*>(s_ring + kPipeStages * kBitmapBytes);
  for (int i = tid; i < (kPipeConsumers / 32) * (2 * kMaxDepth + 1); i += kPipeThreads) (&s_acc[0][0])[i] = 0;
  PlanePipe pp;
  pp.ring = s_ring; pp.full = s_full; pp.empty = s_empty;
  pp.init(tid);
  ItemFeed feed;
  feed.pd = s_pd; feed.mode = s_mode; feed.ready = s_ready; feed.free_ = s_free;
  feed.init(tid);
  __syncthreads();
  const uint32_t depth = p.bit_depth, nf = p.has_filter ? 1u : 0u, n_planes = depth + 2 + nf;
  const uint32_t n_items = p.n_shards * kRowContainers;
  uint32_t it = 0;
  if (tid >= kPipeConsumers) {  // ---- producer warp ----
    unsigned long long bytes_acc = 0;
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
      const uint32_t s = item / kRowContainers, c = item % kRowContainers;
      const FamilyDev fam = p.families[p.family];
      const uint64_t shard = p.shards[s];
      FragDev f{};
      if (shard < fam.cap) f = fam.frags[shard];
      const int buf = feed.begin_fill(it);
      int mode = f.nrows ? 1 : 0;  // no fragment: ValCount{} (executor.go:709-711)
      bool sparse = false;
      if (mode) {
        for (uint32_t j = lane; j < n_planes; j += 32) {
          ContainerDesc d;
          if (j == 0) d = bsi_row_desc(f, 0, c);
          else if (nf && j == 1) d = filter_desc(p, s, c);
          else if (j == 1 + nf) d = bsi_row_desc(f, 1, c);
          else d = bsi_row_desc(f, 2 + (j - 2 - nf), c);
          s_pd[buf][j] = d;
          bytes_acc += desc_algo_bytes(d);
          sparse |= !is_const_or_bitmap(d);
        }
        if (__any_sync(0xffffffffu, sparse)) mode = 2;
      }
      feed.publish(buf, mode, lane);
      if (mode == 1 && lane == 0)
        for (uint32_t j = 0; j < n_planes; j++) pp.produce(s_pd[buf][j]);
    }
    if (bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
  } else {  // ---- consumer warps ----
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
      const int buf = feed.acquire(it);
      const int mode = s_mode[buf];
      if (mode) {
        const ContainerDesc* pd = s_pd[buf];
        uint32_t j = 0;
        auto next_plane = [&](uint4 (&v)[2]) {
          if (mode == 1) pp.take(pd[j], tid, v); else pipe_plane_sync(pd[j], tid, s_win, v);
          j++;
        };
        uint4 cons[2], nrow[2], prow[2];
        next_plane(cons);
        if (nf) {
          uint4 fl[2];
          next_plane(fl);
          cons[0] = and128(cons[0], fl[0]); cons[1] = and128(cons[1], fl[1]);
        }
        next_plane(nrow);
        if (p.strict) { nrow[0] = and128(nrow[0], cons[0]); nrow[1] = and128(nrow[1], cons[1]); }
        prow[0] = andnot128(cons[0], nrow[0]); prow[1] = andnot128(cons[1], nrow[1]);
        const uint32_t cnt = warp_sum(popc128(cons[0]) + popc128(cons[1]));
        if (lane == 0) s_acc[wid][0] += cnt;
        for (uint32_t i = 0; i < depth; i++) {
          uint4 row[2];
          next_plane(row);
          // popcount(B_i & prow) in the low half, popcount(B_i & nrow) in the high half (<= 8192 each per warp)
          const uint32_t both = __reduce_add_sync(0xffffffffu, (popc128(and128(row[0], prow[0])) + popc128(and128(row[1], prow[1]))) |
                                                  ((popc128(and128(row[0], nrow[0])) + popc128(and128(row[1], nrow[1]))) << 16));  // one REDUX
          if (lane == 0) { s_acc[wid][1 + 2 * i] += both & 0xffffu; s_acc[wid][2 + 2 * i] += both >> 16; }
        }
      }
      feed.release(buf, lane);
    }
  }
  __syncthreads();
  for (int i = tid; i < int(2 * depth + 1); i += kPipeThreads) {
    unsigned long long tot = 0;
    for (int w = 0; w < kPipeConsumers / 32; w++) tot += s_acc[w][i];
    if (tot) atomicAdd(p.sum_acc + i, tot);
  }
}

// How a predicate maps onto the unsigned machines (fragment.go:1271-1534): machine kinds for the
// positive / negative side and how their results combine. Shared by the host (which picks the kernel
// instantiation) and the device (which picks filters and magnitudes).
//   combine 0: m0 ; 1: E \ m0 ; 2: S | m0 ; 3: pos | m0 ; 4: m0 | m1 ; 5: E (notNull)
struct RangePlan { int k0, k1, combine; };
__host__ __device__ inline RangePlan range_plan(uint32_t op, long long a, long long b) {
  switch (op) {
    case PB200_BSI_NOTNULL: return {M_NONE, M_NONE, 5};
    case PB200_BSI_EQ: return {M_EQ, M_NONE, 0};
    case PB200_BSI_NEQ: return {M_EQ, M_NONE, 1};
    case PB200_BSI_LT: case PB200_BSI_LTE: {
      const bool eq = op == PB200_BSI_LTE;
      if ((a >= 0 && eq) || (a >= -1 && !eq)) return {M_LTU, M_NONE, 2};
      return {M_GTU, M_NONE, 0};
    }
    case PB200_BSI_GT: case PB200_BSI_GTE: {
      const bool eq = op == PB200_BSI_GTE;
      if ((a >= 0 && eq) || (a >= -1 && !eq)) return {M_GTU, M_NONE, 0};
      return {M_LTU, M_NONE, 3};
    }
    default:  // BETWEEN
      if (a >= 0 || b < 0) return {M_BETWEENU, M_NONE, 0};
      return {M_LTU, M_LTU, 4};
  }
}

// K10. Planes in consumption order: E, S, B_{d-1} .. B_0. One instantiation per (machines, combine).
template <int K0, int K1, int COMBINE>
__global__ void __launch_bounds__(kPipeThreads, 3) bsi_range_kernel(BsiParams p) {
  extern __shared__ __align__(128) uint8_t s_ring[];  // kPipeStages slots, then an 8 KiB staging window
  __shared__ __align__(8) uint64_t s_full[kPipeStages], s_empty[kPipeStages], s_ready[2], s_free[2];
  __shared__ ContainerDesc s_pd[2][ItemFeed::kMaxPlanes];
  __shared__ int s_mode[2];
  const int tid = threadIdx.x, lane = tid & 31;
  uint64_t* s_win = reinterpret_cast<uint64_t*>(s_ring + kPipeStages * kBitmapBytes);
  PlanePipe pp;
  pp.ring = s_ring; pp.full = s_full; pp.empty = s_empty;
  pp.init(tid);
  ItemFeed feed;
  feed.pd = s_pd; feed.mode = s_mode; feed.ready = s_ready; feed.free_ = s_free;
  feed.init(tid);
  __syncthreads();
  const uint32_t depth = p.bit_depth;
  const uint32_t n_planes = COMBINE == 5 ? 1u : depth + 2u;
  const uint32_t n_items = p.n_shards * kRowContainers;
  uint32_t it = 0;
  if (tid >= kPipeConsumers) {  // ---- producer warp ----
    unsigned long long bytes_acc = 0;
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
      const uint32_t s = item / kRowContainers, c = item % kRowContainers;
      const FamilyDev fam = p.families[p.family];
      const uint64_t shard = p.shards[s];
      FragDev f{};
      if (shard < fam.cap) f = fam.frags[shard];
      const int buf = feed.begin_fill(it);
      bool sparse = false;
      for (uint32_t j = lane; j < n_planes; j += 32) {
        const uint64_t row = j < 2 ? uint64_t(j) : uint64_t(2 + (depth - 1 - (j - 2)));
        const ContainerDesc d = bsi_row_desc(f, row, c);  // no fragment: empty planes, the result row is zero
        s_pd[buf][j] = d;
        bytes_acc += desc_algo_bytes(d);
        sparse |= !is_const_or_bitmap(d);
      }
      const int mode = __any_sync(0xffffffffu, sparse) ? 2 : 1;
      feed.publish(buf, mode, lane);
      if (mode == 1 && lane == 0)
        for (uint32_t j = 0; j < n_planes; j++) pp.produce(s_pd[buf][j]);
    }
    if (bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
    return;
  }
  // ---- consumer warps ----
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
    const uint32_t s = item / kRowContainers;
    uint4* out = p.dense_out ? reinterpret_cast<uint4*>(p.dense_out + uint64_t(item) * kBitmapBytes) : nullptr;
    const int buf = feed.acquire(it);
    const int mode = s_mode[buf];
    const ContainerDesc* pd = s_pd[buf];
    uint32_t j = 0;
    auto next_plane = [&](uint4 (&v)[2]) {
      if (mode == 1) pp.take(pd[j], tid, v); else pipe_plane_sync(pd[j], tid, s_win, v);
      j++;
    };
    uint4 E[2], res[2];
    next_plane(E);
    if (COMBINE == 5) {
      res[0] = E[0]; res[1] = E[1];
    } else {
      uint4 S[2], pos[2], neg[2];
      next_plane(S);
#pragma unroll
      for (int k = 0; k < 2; k++) { pos[k] = andnot128(E[k], S[k]); neg[k] = and128(E[k], S[k]); }
      Mach<2, K0> m0;
      Mach<2, K1> m1;
      const long long a = p.a, b = p.b;
      const unsigned long long ua = uabs64(a), ub = uabs64(b);
      const bool eq = p.bsi_op == PB200_BSI_LTE || p.bsi_op == PB200_BSI_GTE;
      if (K0 == M_EQ) m0.init(K0, a < 0 ? neg : pos, ua, 0, false);
      else if (K1 == M_LTU) { m0.init(K0, pos, ub, 0, true); m1.init(K1, neg, ua, 0, true); }  // BETWEEN across zero
      else if (K0 == M_BETWEENU) { if (a >= 0) m0.init(K0, pos, ua, ub, true); else m0.init(K0, neg, ub, ua, true); }
      else if (COMBINE == 2) m0.init(K0, pos, ua, 0, eq);   // LT / LTE with a non-negative bound
      else if (COMBINE == 3) m0.init(K0, neg, ua, 0, eq);   // GT / GTE with a negative bound
      else if (p.bsi_op == PB200_BSI_LT || p.bsi_op == PB200_BSI_LTE) m0.init(K0, neg, ua, 0, eq);  // M_GTU over the negatives
      else m0.init(K0, pos, ua, 0, eq);                                                              // M_GTU over the positives
      for (int i = int(depth) - 1; i >= 0; i--) {  // every plane is taken, so that both sides of the pipe count alike
        uint4 row[2];
        next_plane(row);
        m0.step(i, row);
        if (K1 != M_NONE) m1.step(i, row);
      }
#pragma unroll
      for (int k = 0; k < 2; k++) {
        if (COMBINE == 0) res[k] = m0.f[k];
        else if (COMBINE == 1) res[k] = andnot128(E[k], m0.f[k]);
        else if (COMBINE == 2) res[k] = or128(S[k], m0.f[k]);
        else if (COMBINE == 3) res[k] = or128(pos[k], m0.f[k]);
        else res[k] = or128(m0.f[k], m1.f[k]);
      }
    }
    feed.release(buf, lane);
    if (out) { out[tid] = res[0]; out[tid + kPipeConsumers] = res[1]; }
    if (p.count_out) {
      const uint32_t cnt = warp_sum(popc128(res[0]) + popc128(res[1]));
      if (lane == 0 && cnt) atomicAdd(p.count_out + uint64_t(s) * p.count_sstride, (unsigned long long)cnt);
    }
  }
}

// K9 (see the slice decomposition above). Planes in consumption order: E, [filter], S, B_{d-1} .. B_0.
__global__ void __launch_bounds__(kPipeThreads, 3) bsi_minmax_kernel(BsiParams p, MmRec* __restrict__ recs) {
  extern __shared__ __align__(128) uint8_t s_ring[];  // kPipeStages slots, then an 8 KiB staging window
  __shared__ __align__(8) uint64_t s_full[kPipeStages], s_empty[kPipeStages], s_ready[2], s_free[2];
  __shared__ ContainerDesc s_pd[2][ItemFeed::kMaxPlanes];
  __shared__ int s_mode[2];
  __shared__ MmRec s_rec[kPipeConsumers / 32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint64_t* s_win = reinterpret_cast<uint64_t*>(s_ring + kPipeStages * kBitmapBytes);
  PlanePipe pp;
  pp.ring = s_ring; pp.full = s_full; pp.empty = s_empty;
  pp.init(tid);
  ItemFeed feed;
  feed.pd = s_pd; feed.mode = s_mode; feed.ready = s_ready; feed.free_ = s_free;
  feed.init(tid);
  __syncthreads();
  const uint32_t depth = p.bit_depth, nf = p.has_filter ? 1u : 0u, n_planes = depth + 2 + nf;
  const uint32_t n_items = p.n_shards * kRowContainers;
  uint32_t it = 0;
  if (tid >= kPipeConsumers) {  // ---- producer warp ----
    unsigned long long bytes_acc = 0;
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
      const uint32_t s = item / kRowContainers, c = item % kRowContainers;
      const FamilyDev fam = p.families[p.family];
      const uint64_t shard = p.shards[s];
      FragDev f{};
      if (shard < fam.cap) f = fam.frags[shard];
      const int buf = feed.begin_fill(it);
      int mode = f.nrows ? 1 : 0;
      bool sparse = false;
      if (mode) {
        for (uint32_t j = lane; j < n_planes; j += 32) {
          ContainerDesc d;
          if (j == 0) d = bsi_row_desc(f, 0, c);
          else if (nf && j == 1) d = filter_desc(p, s, c);
          else if (j == 1 + nf) d = bsi_row_desc(f, 1, c);
          else d = bsi_row_desc(f, 2 + (depth - 1 - (j - 2 - nf)), c);
          s_pd[buf][j] = d;
          bytes_acc += desc_algo_bytes(d);
          sparse |= !is_const_or_bitmap(d);
        }
        if (__any_sync(0xffffffffu, sparse)) mode = 2;
      }
      feed.publish(buf, mode, lane);
      if (mode == 1 && lane == 0)
        for (uint32_t j = 0; j < n_planes; j++) pp.produce(s_pd[buf][j]);
    }
    if (bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
    return;
  }
  // ---- consumer warps ----
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, it++) {
    const int buf = feed.acquire(it);
    const int mode = s_mode[buf];
    MmRec mine{0, 0, ~0ull, 0};
    if (mode) {
      const ContainerDesc* pd = s_pd[buf];
      uint32_t j = 0;
      auto next_plane = [&](uint4 (&v)[2]) {
        if (mode == 1) pp.take(pd[j], tid, v); else pipe_plane_sync(pd[j], tid, s_win, v);
        j++;
      };
      uint4 cons[2], sign[2];
      next_plane(cons);
      if (nf) {
        uint4 fl[2];
        next_plane(fl);
        cons[0] = and128(cons[0], fl[0]); cons[1] = and128(cons[1], fl[1]);
      }
      next_plane(sign);
      uint4 A[2], B[2];  // A: maxUnsigned candidates, B: minUnsigned candidates (see K9 above)
#pragma unroll
      for (int k = 0; k < 2; k++) {
        const uint4 neg = and128(cons[k], sign[k]), pos = andnot128(cons[k], sign[k]);
        A[k] = p.is_max ? pos : neg;
        B[k] = p.is_max ? neg : pos;
      }
      unsigned long long magA = 0, magB = 0;
      for (int i = int(depth) - 1; i >= 0; i--) {
        uint4 v[2];
        next_plane(v);
        const uint4 x0 = and128(A[0], v[0]), x1 = and128(A[1], v[1]);
        const uint4 y0 = andnot128(B[0], v[0]), y1 = andnot128(B[1], v[1]);
        if (__any_sync(0xffffffffu, mm_nz(x0) || mm_nz(x1))) { A[0] = x0; A[1] = x1; magA |= 1ull << i; }   // maxUnsigned :1216-1220
        if (__any_sync(0xffffffffu, mm_nz(y0) || mm_nz(y1))) { B[0] = y0; B[1] = y1; } else magB |= 1ull << i;  // minUnsigned :1174-1180
      }
      mine = MmRec{magA, warp_sum(popc128(A[0]) + popc128(A[1])), magB, warp_sum(popc128(B[0]) + popc128(B[1]))};
    }
    feed.release(buf, lane);
    consumers_sync();  // s_rec of the previous item has been folded
    if (lane == 0) s_rec[wid] = mine;
    consumers_sync();
    if (tid == 0) {  // fold the column's 8 slices
      MmRec r{0, 0, ~0ull, 0};
      for (int w = 0; w < kPipeConsumers / 32; w++) {
        const MmRec x = s_rec[w];
        if (x.cntA) { if (x.magA > r.magA || !r.cntA) { r.magA = x.magA; r.cntA = x.cntA; } else if (x.magA == r.magA) r.cntA += x.cntA; }
        if (x.cntB) { if (x.magB < r.magB || !r.cntB) { r.magB = x.magB; r.cntB = x.cntB; } else if (x.magB == r.magB) r.cntB += x.cntB; }
      }
      recs[item] = r;
    }
  }
}

// ---------------------------------------------------------------------------------------
// Shift by one with carry across containers (roaring.go:944-977, :4374-4388) on dense rows:
// out = (in << 1) | carry_in. The bit carried out of the shard's last column is not lost:
// Bitmap.Shift appends it as a container at key+1 (roaring.go:970-974), which inside a
// per-shard row segment is the first container of the NEXT shard's key range. That
// container is modelled as one 64-bit "overflow word" per (temp slot, shard): bit k =
// column (shard + 1) * 2^20 + k, enough for 64 nested shifts.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) shift_rows_kernel(const uint8_t* in, uint8_t* out, uint32_t n_shards) {
  const uint64_t n_words = uint64_t(n_shards) * kRowContainers * kBitmapWords;
  const uint64_t* a = reinterpret_cast<const uint64_t*>(in);
  uint64_t* o = reinterpret_cast<uint64_t*>(out);
  for (uint64_t w = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; w < n_words; w += uint64_t(gridDim.x) * blockDim.x) {
    const uint64_t in_row = w % (uint64_t(kRowContainers) * kBitmapWords);
    const uint64_t prev = in_row == 0 ? 0 : a[w - 1];
    o[w] = (a[w] << 1) | (prev >> 63);
  }
}

// overflow word of a shifted row: the operand's own overflow container shifted by one, plus
// the carry out of the operand's last word
__global__ void shift_overflow_kernel(const uint8_t* in, const unsigned long long* ovf_in, unsigned long long* ovf_out, uint32_t n_shards) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_shards) return;
  const uint64_t* a = reinterpret_cast<const uint64_t*>(in) + (uint64_t(s) + 1) * kRowContainers * kBitmapWords - 1;
  ovf_out[s] = (ovf_in[s] << 1) | (*a >> 63);
}

// The accumulator machine over overflow words only: leaves that are temp rows contribute
// their overflow word, every stored row contributes 0 (a fragment has no container in the
// next shard's key range), so Intersect drops the overflow, Union / Xor / Difference keep it —
// exactly what the roaring merge-joins do with a key that exists on one side only
// (roaring.go:593-941). One thread per (query, shard).
struct OverflowParams {
  const Prog* progs; const Instr* instrs; const Leaf* leaves;
  uint32_t n_queries, n_shards;
  const unsigned long long* ovf;   // [slot][n_shards]
  unsigned long long* out_counts;  // optional: += popcount at [q * qstride + s * sstride]
  uint32_t out_qstride, out_sstride;
  unsigned long long* ovf_out;     // optional: result word per (q, s)
};

__global__ void eval_overflow_kernel(OverflowParams p) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= uint64_t(p.n_queries) * p.n_shards) return;
  const uint32_t q = uint32_t(i / p.n_shards), s = uint32_t(i % p.n_shards);
  const Prog pg = p.progs[q];
  unsigned long long acc = 0, stack[kMaxStack];
  int sp = 0;
  bool live = false;
  for (uint32_t k = 0; k < pg.n_instr; k++) {
    const Instr in = p.instrs[pg.instr_off + k];
    unsigned long long v;
    if (in.op <= I_RANDNOT_LEAF) {
      const Leaf lf = p.leaves[pg.leaf_off + in.leaf];
      v = lf.kind == L_TEMP ? p.ovf[uint64_t(lf.family) * p.n_shards + s] : 0ull;
      if (in.op == I_LOAD) { if (live) stack[sp++] = acc; acc = v; live = true; continue; }
    } else {
      v = stack[--sp];
    }
    switch (in.op) {
      case I_AND_LEAF: case I_AND_STACK: acc &= v; break;
      case I_OR_LEAF: case I_OR_STACK: acc |= v; break;
      case I_XOR_LEAF: case I_XOR_STACK: acc ^= v; break;
      case I_ANDNOT_LEAF: acc &= ~v; break;
      default: acc = v & ~acc; break;  // I_RANDNOT_LEAF, I_ANDNOT_STACK: left \ acc
    }
  }
  if (p.ovf_out) p.ovf_out[i] = acc;
  if (p.out_counts && acc) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)__popcll(acc));
}

}  // namespace pb200
