// CUDA kernels (sm_100a) for the per-shard bitmap query hot path. No tensor cores: the
// work is HBM-bound integer/bit arithmetic (AND/OR/XOR/ANDNOT + popcount), so the design
// rules are 128-bit coalesced loads, enough bytes in flight per SM, shared-memory staging
// for array/run containers and grids sized in multiples of the SM count.
//
// Work decomposition: one ITEM = (query q, shard s, container column c) = the <= N
// containers (one per operand row) that share key&15 == c in one shard — the unit
// roaring.Bitmap.Intersect/Union/... merge-join on (roaring.go:593-941). Because every
// row of a fragment occupies exactly 16 consecutive container keys, the merge-join
// degenerates to a 16-slot gather from the row tables (layout.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "layout.h"
#include "tma.cuh"

namespace pb200 {

// Cross-GPU sum of a batch's counts without a collective library: when the last CTA of the batch's last kernel retires,
// it adds this GPU's partial counts into a mailbox slot of every participating GPU (plain system-scope reductions over
// NVLink peer memory — the peers' buffers are mapped into this process, nothing is multicast) and then bumps each
// mailbox's arrival counter; a GPU's own total is complete when its counter reaches the number of participants.
constexpr int kPeerMax = 8;
struct PushDesc {
  uint32_t n_dst;                        // 0: no push
  uint32_t n;                            // counts to push
  unsigned int* ticket;                  // local, zeroed before the launch: CTAs retired so far
  const unsigned long long* local;       // this GPU's partial counts
  unsigned long long* dst[kPeerMax];     // slot of this epoch in every participant's mailbox (own one included)
  unsigned long long* flag[kPeerMax];    // its arrival counter
};

__device__ __forceinline__ void peer_push(const PushDesc& pd, int tid, int nthreads) {
  for (uint32_t q = tid; q < pd.n; q += nthreads) {
    const unsigned long long v = __ldcg(pd.local + q);  // the counts were accumulated with L2 atomics
    if (v)
      for (uint32_t d = 0; d < pd.n_dst; d++) atomicAdd_system(pd.dst[d] + q, v);
  }
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    for (uint32_t d = 0; d < pd.n_dst; d++) atomicAdd_system(pd.flag[d], 1ull);
    *pd.ticket = 0u;  // this CTA was the ticket's last user: the slot's next batch finds it at zero
  }
}

// stand-alone form, for batches whose last kernel is not eval_fast_kernel
__global__ void peer_push_kernel(PushDesc pd) { peer_push(pd, threadIdx.x, blockDim.x); }

// Waits until `world` participants have pushed into a mailbox slot, moves the totals to `out` and re-arms the slot.
// err[0] is set when the peers do not arrive within ~2 s (a peer process died): the host reports it instead of hanging.
__global__ void peer_collect_kernel(unsigned long long* slot, unsigned long long* flag, uint32_t world, uint32_t n, unsigned long long* out,
                                    unsigned int* err) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    int ok = 1;
    while (*reinterpret_cast<volatile unsigned long long*>(flag) < world) {
      if (clock64() - t0 > 4000000000ll) { ok = 0; break; }
      __nanosleep(200);
    }
    __threadfence_system();
    s_ok = ok;
  }
  __syncthreads();
  if (!s_ok) { if (threadIdx.x == 0) atomicExch(err, 1u); return; }
  for (uint32_t q = threadIdx.x; q < n; q += blockDim.x) {
    out[q] = __ldcg(slot + q);
    slot[q] = 0ull;
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) *reinterpret_cast<volatile unsigned long long*>(flag) = 0ull;
}

struct EvalParams {
  const FamilyDev* families;
  uint32_t n_families;
  const Prog* progs;
  const Instr* instrs;
  const Leaf* leaves;
  const uint64_t* shards;
  uint32_t n_shards;
  uint32_t n_queries;
  const uint8_t* temp_base;        // dense temp rows: ((slot * n_shards + s) * 16 + c) * 8192
  unsigned long long* out_counts;  // out[q * out_qstride + s * out_sstride] += popcount (may be null)
  uint32_t out_qstride, out_sstride;
  uint8_t* dense_out;              // optional: ((q * n_shards + s) * 16 + c) * 8192 bytes of result words
  uint32_t* dense_n;               // optional: cardinality per (q, s, c)
  uint32_t* slow_list;             // scratch: items the fast kernel defers to the staged kernel
  uint32_t* slow_count;
  ContainerDesc* slow_descs;       // scratch (optional): resolved descriptors of deferred items, [slot][slow_desc_stride]
  uint32_t slow_desc_stride;
  unsigned long long* algo_bytes;  // accumulated algorithmic bytes (payload + 16 B / container)
  uint32_t fold_all;               // eval_fold_kernel walks EVERY item itself (no eval_fast_kernel pass, no deferred list)
  uint32_t* item_list;             // eval_tiny_kernel ran first: the items it left for eval_fast_kernel ...
  uint32_t* item_count;            // ... and how many (null: eval_fast_kernel walks every item)
  uint16_t* sparse_out;            // optional, with dense_out: a result container of <= kSparseListMax values is written as a sorted
                                   // list at sparse_out[item * kSparseListMax] instead of 8 KiB of words; dense_n[item] carries kSparseFlag
  PushDesc push;                   // eval_fast_kernel only: peer reduction fused into the kernel's tail (n_dst == 0: none)
};

// ---------------------------------------------------------------------------------------
// leaf resolution: (family, row) x shard x column -> ContainerDesc. Replaces
// Holder.fragment (holder.go:496) + fragment.row -> Bitmap.OffsetRange (roaring.go:535).
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ ContainerDesc resolve_leaf(const EvalParams& p, const Leaf& lf, uint32_t s, uint32_t c) {
  ContainerDesc d;
  d.addr = 0; d.n = 0; d.meta = 0;
  if (lf.kind == L_ZERO) return d;
  if (lf.kind == L_TEMP) {
    d.addr = reinterpret_cast<uint64_t>(p.temp_base) + ((uint64_t(lf.family) * p.n_shards + s) * kRowContainers + c) * kBitmapBytes;
    d.n = 1;  // unknown, non-zero
    d.meta = make_meta(kBitmap, kBitmapWords);
    return d;
  }
  if (lf.family >= p.n_families) return d;
  const FamilyDev fam = p.families[lf.family];
  const uint64_t shard = p.shards[s];
  if (shard >= fam.cap) return d;
  const FragDev f = fam.frags[shard];
  if (f.nrows == 0) return d;
  uint64_t idx;
  if (f.dense) {
    if (lf.row >= f.nrows) return d;
    idx = lf.row;
  } else {
    uint32_t lo = 0, hi = f.nrows;
    while (lo < hi) {
      uint32_t mid = (lo + hi) >> 1;
      if (f.row_ids[mid] < lf.row) lo = mid + 1; else hi = mid;
    }
    if (lo >= f.nrows || f.row_ids[lo] != lf.row) return d;
    idx = lo;
  }
  return f.rows[idx * kRowContainers + c];
}

__device__ __forceinline__ bool is_const_or_bitmap(const ContainerDesc& d) {
  return d.n == 0 || d.n == kFullN || desc_type(d) == kBitmap;
}

// Which items the fold kernel takes (the staged kernel takes the rest): LOAD first, then AND / OR / XOR / AND-NOT
// steps against leaves — Intersect, Union, Xor, Difference of rows. (Round-2 history: with atomics on the slice edges
// and a prefix-XOR run scan the XOR / AND-NOT folds were 13-19 % faster on the staged evaluator and stayed there; with
// the atomics-free slices and the E - S run fill they are 16 % faster here, PB200_FOLD_XOR=0 builds the old routing.)
// the steps a flat fold may consist of after its LOAD (the host's classification, engine.cu finish(), uses the same test)
#ifndef PB200_FOLD_XOR
#define PB200_FOLD_XOR 1
#endif
__host__ __device__ inline bool fold_op_ok(uint8_t op) {
  return op == I_AND_LEAF || op == I_OR_LEAF || (PB200_FOLD_XOR && (op == I_XOR_LEAF || op == I_ANDNOT_LEAF));
}
__device__ __forceinline__ bool fold_can_take(const Prog& pg, const Instr* ins, const ContainerDesc* descs) {
  if (pg.n_instr == 0 || ins[0].op != I_LOAD) return false;
  for (uint32_t i = 1; i < pg.n_instr; i++)
    if (!fold_op_ok(ins[i].op)) return false;
  (void)descs;
  return true;
}

__device__ __forceinline__ uint32_t desc_algo_bytes(const ContainerDesc& d) {
  if (d.n == 0) return 0;
  if (desc_densified(d)) return 16u + desc_file_bytes(d);  // algorithmic bytes follow the reference's encoding
  uint32_t t = desc_type(d), l = desc_len(d);
  return 16u + (t == kBitmap ? uint32_t(kBitmapBytes) : t == kArray ? l * 2u : l * 4u);
}

__device__ __forceinline__ uint4 ldg128(const void* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}
// streaming load: operand payloads are read once per query — do not let them evict the
// descriptor tables from L1/L2 (ld.global.nc.L1::no_allocate).
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t popc128(const uint4& v) {
  return __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
}
__device__ __forceinline__ uint4 and128(const uint4& a, const uint4& b) { return make_uint4(a.x & b.x, a.y & b.y, a.z & b.z, a.w & b.w); }
__device__ __forceinline__ uint4 or128(const uint4& a, const uint4& b) { return make_uint4(a.x | b.x, a.y | b.y, a.z | b.z, a.w | b.w); }
__device__ __forceinline__ uint4 not128(const uint4& a) { return make_uint4(~a.x, ~a.y, ~a.z, ~a.w); }
__device__ __forceinline__ uint4 xor128(const uint4& a, const uint4& b) { return make_uint4(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z, a.w ^ b.w); }
__device__ __forceinline__ uint4 andnot128(const uint4& a, const uint4& b) { return make_uint4(a.x & ~b.x, a.y & ~b.y, a.z & ~b.z, a.w & ~b.w); }

__device__ __forceinline__ uint4 apply_op(uint8_t op, const uint4& acc, const uint4& v) {
  switch (op) {
    case I_AND_LEAF: case I_AND_STACK: return and128(acc, v);
    case I_OR_LEAF: case I_OR_STACK: return or128(acc, v);
    case I_XOR_LEAF: case I_XOR_STACK: return xor128(acc, v);
    case I_ANDNOT_LEAF: return andnot128(acc, v);   // acc \ leaf
    case I_RANDNOT_LEAF: return andnot128(v, acc);  // leaf \ acc
    case I_ANDNOT_STACK: return andnot128(v, acc);  // v = popped LEFT operand: left \ acc
    default: return acc;
  }
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
  return __reduce_add_sync(0xffffffffu, v);  // one REDUX.SUM instead of a five-step shuffle tree
}
__device__ __forceinline__ unsigned long long warp_sum64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Sum of v over the lanes of `grp`, the caller's own group from __match_any_sync (all 32 lanes call). The whole warp
// and the two half warps — what a run of consecutive items gives — take shuffle trees; anything else a plain walk.
__device__ __forceinline__ unsigned long long warp_group_sum(uint32_t grp, unsigned long long v) {
  const bool whole = __all_sync(0xffffffffu, grp == 0xffffffffu);
  if (whole) return warp_sum64(v);
  const bool halves = __all_sync(0xffffffffu, grp == 0x0000ffffu || grp == 0xffff0000u);
  if (halves) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  }
  unsigned long long tot = 0;
  for (int src = 0; src < 32; src++) {
    const unsigned long long x = __shfl_sync(0xffffffffu, v, src);
    if ((grp >> src) & 1u) tot += x;
  }
  return tot;
}

// ---------------------------------------------------------------------------------------
// K2-K5 (count form), warp-level: |a AND b| when the smaller operand is a short array or a
// short run list. Lanes take the elements (runs) of the small side and test them against
// the other container where it lies: bitmap -> one word probe (intersectionCountArrayBitmap
// roaring.go:3197 / BitmapRun :3189), array -> binary search (ArrayArray :3115, ArrayRun
// :3138), runs -> binary search + interval overlap (RunRun :3156). Returns kNotSparse when
// the pair should be staged instead. The result is uniform across the warp.
// ---------------------------------------------------------------------------------------
constexpr uint32_t kNotSparse = 0xffffffffu;
constexpr uint32_t kSparseListMax = 32;        // values of a result container kept as a list (EvalParams::sparse_out)
constexpr uint32_t kSparseFlag = 0x80000000u;  // in dense_n: the container is in the list, its dense words were not written
constexpr uint32_t kSparseArrayMax = 1024;  // elements of the probing side
constexpr uint32_t kSparseRunMax = 256;     // runs of the probing side

__device__ __forceinline__ bool in_bitmap(uint64_t addr, uint32_t v) {
  return (__ldg(reinterpret_cast<const unsigned long long*>(addr) + (v >> 6)) >> (v & 63)) & 1ull;
}
__device__ __forceinline__ bool in_array(uint64_t addr, uint32_t len, uint32_t v) {
  const uint16_t* a = reinterpret_cast<const uint16_t*>(addr);
  uint32_t lo = 0, hi = len;
  while (lo < hi) { const uint32_t m = (lo + hi) >> 1; if (__ldg(a + m) < v) lo = m + 1; else hi = m; }
  return lo < len && __ldg(a + lo) == v;
}
// index of the first run whose last >= v (len if none)
__device__ __forceinline__ uint32_t run_lower(const ushort2* r, uint32_t len, uint32_t v) {
  uint32_t lo = 0, hi = len;
  while (lo < hi) { const uint32_t m = (lo + hi) >> 1; if (__ldg(r + m).y < v) lo = m + 1; else hi = m; }
  return lo;
}
__device__ __forceinline__ uint32_t popc_range_global(uint64_t addr, uint32_t s, uint32_t l) {  // inclusive [s, l]
  const unsigned long long* w = reinterpret_cast<const unsigned long long*>(addr);
  const uint32_t ws = s >> 6, we = l >> 6;
  const unsigned long long ms = ~0ull << (s & 63), me = ~0ull >> (63 - (l & 63));
  if (ws == we) return __popcll(__ldg(w + ws) & ms & me);
  uint32_t n = __popcll(__ldg(w + ws) & ms) + __popcll(__ldg(w + we) & me);
  for (uint32_t k = ws + 1; k < we; k++) n += __popcll(__ldg(w + k));
  return n;
}

__device__ __noinline__ uint32_t sparse_pair_icount(ContainerDesc a, ContainerDesc b, int lane) {
  if (a.n == 0 || b.n == 0) return 0;
  if (a.n == kFullN) return b.n;  // roaring.go:3079-3084
  if (b.n == kFullN) return a.n;
  uint32_t ta = desc_type(a), tb = desc_type(b);
  // make `a` the probing (small) side: prefer an array, then a run list; the smaller one on ties
  const bool a_probe = ta != kBitmap, b_probe = tb != kBitmap;
  if (!a_probe && !b_probe) return kNotSparse;
  if (!a_probe || (b_probe && ((tb == kArray && ta != kArray) || (tb == ta && desc_len(b) < desc_len(a))))) {
    const ContainerDesc t = a; a = b; b = t;
    ta = desc_type(a); tb = desc_type(b);
  }
  const uint32_t la = desc_len(a), lb = desc_len(b);
  uint32_t cnt = 0;
  if (ta == kArray) {
    if (la > kSparseArrayMax) return kNotSparse;
    const uint16_t* arr = reinterpret_cast<const uint16_t*>(a.addr);
    for (uint32_t i = lane; i < la; i += 32) {
      const uint32_t v = __ldg(arr + i);
      bool hit;
      if (tb == kBitmap) hit = in_bitmap(b.addr, v);
      else if (tb == kArray) hit = in_array(b.addr, lb, v);
      else { const ushort2* r = reinterpret_cast<const ushort2*>(b.addr); const uint32_t j = run_lower(r, lb, v); hit = j < lb && __ldg(r + j).x <= v; }
      cnt += hit ? 1u : 0u;
    }
  } else {  // a is a run list
    if (la > kSparseRunMax) return kNotSparse;
    if (tb == kBitmap && a.n > 16384) return kNotSparse;  // long runs over a bitmap: streaming beats probing
    const ushort2* ra = reinterpret_cast<const ushort2*>(a.addr);
    for (uint32_t i = lane; i < la; i += 32) {
      const ushort2 iv = __ldg(ra + i);
      if (tb == kBitmap) cnt += popc_range_global(b.addr, iv.x, iv.y);
      else {  // run x run: overlaps with every run of b that reaches into [start, last]
        const ushort2* rb = reinterpret_cast<const ushort2*>(b.addr);
        for (uint32_t j = run_lower(rb, lb, iv.x); j < lb; j++) {
          const ushort2 o = __ldg(rb + j);
          if (o.x > iv.y) break;
          cnt += min(uint32_t(iv.y), uint32_t(o.y)) - max(uint32_t(iv.x), uint32_t(o.x)) + 1;
        }
      }
    }
  }
  return warp_sum(cnt);
}

constexpr int kFastWarps = 8;  // warps per CTA in the fast kernel
constexpr int kUnroll = 4;     // 128-bit chunks evaluated together per lane

// ---------------------------------------------------------------------------------------
// K1 / K6-fast: one WARP per item. Handles every item whose operand containers are all
// bitmap, empty or full: the program is evaluated word-parallel straight out of global
// memory with 128-bit loads (16 chunks per lane per container), popcounted and
// warp-reduced; one atomicAdd per item. Items with array/run operands are appended to
// the slow list for eval_staged_kernel. Also covers intersectionCountBitmapBitmap
// (roaring.go:3212, :5061) and the full/empty shortcuts (roaring.go:3079-3087).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFastWarps * 32) eval_fast_kernel(EvalParams p) {
  __shared__ ContainerDesc s_desc[kFastWarps][kMaxLeaves];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint64_t n_items = uint64_t(p.n_queries) * p.n_shards * kRowContainers;
  const uint64_t warp_stride = uint64_t(gridDim.x) * kFastWarps;
  unsigned long long bytes_acc = 0;

  const uint64_t n_walk = p.item_list ? uint64_t(*p.item_count) : n_items;
  for (uint64_t k_item = uint64_t(blockIdx.x) * kFastWarps + warp; k_item < n_walk; k_item += warp_stride) {
    const uint64_t item = p.item_list ? uint64_t(p.item_list[k_item]) : k_item;
    const uint32_t c = uint32_t(item & 15);
    const uint64_t qs = item >> 4;
    const uint32_t s = uint32_t(qs % p.n_shards);
    const uint32_t q = uint32_t(qs / p.n_shards);
    const Prog pg = p.progs[q];
    const Instr* ins = p.instrs + pg.instr_off;

    // resolve leaves (lane l -> leaf l)
    bool ok = true;
    uint32_t bytes = 0;
    if (lane < pg.n_leaves) {
      ContainerDesc d = resolve_leaf(p, p.leaves[pg.leaf_off + lane], s, c);
      s_desc[warp][lane] = d;
      ok = is_const_or_bitmap(d);
      bytes = desc_algo_bytes(d);
    }
    const bool all_ok = __all_sync(0xffffffffu, ok);
    bytes = warp_sum(bytes);
    if (lane == 0) bytes_acc += bytes;
    __syncwarp();

    const bool want_dense = p.dense_out != nullptr;
    if (!all_ok) {
      // K2-K5 in-warp: a two-operand count whose smaller side is a short array / run list needs
      // no staging — membership probes (or interval overlaps) plus the cardinalities give
      // |a AND b|, and |a OR b|, |a XOR b|, |a ANDNOT b| follow from it.
      bool done = false;
      if (!want_dense && pg.n_instr == 2 && ins[0].op == I_LOAD && ins[1].op >= I_AND_LEAF && ins[1].op <= I_ANDNOT_LEAF &&
          p.leaves[pg.leaf_off + ins[0].leaf].kind != L_TEMP && p.leaves[pg.leaf_off + ins[1].leaf].kind != L_TEMP) {
        const ContainerDesc a = s_desc[warp][ins[0].leaf], b = s_desc[warp][ins[1].leaf];
        const uint32_t ic = sparse_pair_icount(a, b, lane);
        if (ic != kNotSparse) {
          uint32_t r;
          switch (ins[1].op) {
            case I_AND_LEAF: r = ic; break;
            case I_OR_LEAF: r = a.n + b.n - ic; break;
            case I_XOR_LEAF: r = a.n + b.n - 2 * ic; break;
            default: r = a.n - ic; break;  // I_ANDNOT_LEAF
          }
          if (lane == 0 && p.out_counts && r) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)r);
          done = true;
        }
      } else if (!want_dense && pg.n_instr == 1 && p.leaves[pg.leaf_off + ins[0].leaf].kind != L_TEMP) {
        // Count(Row) of an array / run container: the cardinality is in the descriptor
        const ContainerDesc d = s_desc[warp][ins[0].leaf];
        if (lane == 0 && p.out_counts && d.n) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)d.n);
        done = true;
      }
      if (!done) {
        // defer to the staged kernel; hand it the descriptors already resolved here so it does
        // not repeat the shard -> fragment -> row-table pointer chase
        uint32_t slot = 0;
        if (lane == 0) { slot = atomicAdd(p.slow_count, 1u); p.slow_list[slot] = uint32_t(item); }
        if (p.slow_descs) {
          slot = __shfl_sync(0xffffffffu, slot, 0);
          if (lane < pg.n_leaves) p.slow_descs[uint64_t(slot) * p.slow_desc_stride + lane] = s_desc[warp][lane];
        }
      }
      __syncwarp();
      continue;
    }

    uint32_t cnt = 0;
    uint8_t* dout = want_dense ? p.dense_out + item * kBitmapBytes : nullptr;

    if (!want_dense && pg.n_instr == 1) {
      // Count(Row): cardinality straight from the descriptor (roaring.go:405 Count = sum of n).
      // (L_TEMP leaves carry no cardinality, so they take the generic path.)
      const ContainerDesc d = s_desc[warp][ins[0].leaf];
      if (!(d.n == 1 && desc_type(d) == kBitmap && p.leaves[pg.leaf_off + ins[0].leaf].kind == L_TEMP)) {
        if (lane == 0 && p.out_counts && d.n) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)d.n);
        __syncwarp();
        continue;
      }
    }

    if (!want_dense && pg.n_instr == 2 && ins[1].op == I_AND_LEAF) {
      // ---- 2-operand intersect: the headline Count(Intersect(Row, Row)) path ----
      const ContainerDesc a = s_desc[warp][ins[0].leaf], b = s_desc[warp][ins[1].leaf];
      const bool a_temp = p.leaves[pg.leaf_off + ins[0].leaf].kind == L_TEMP;
      const bool b_temp = p.leaves[pg.leaf_off + ins[1].leaf].kind == L_TEMP;
      if (a.n == 0 || b.n == 0) { __syncwarp(); continue; }
      if (a.n == kFullN && !b_temp) cnt = (lane == 0) ? b.n : 0;        // roaring.go:3079
      else if (b.n == kFullN && !a_temp) cnt = (lane == 0) ? a.n : 0;   // roaring.go:3082
      else {
        const uint4* pa = reinterpret_cast<const uint4*>(a.addr) + lane;
        const uint4* pb = reinterpret_cast<const uint4*>(b.addr) + lane;
        const bool fa = a.n == kFullN, fb = b.n == kFullN;
        const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll
        for (int k0 = 0; k0 < 16; k0 += 8) {
          uint4 va[8], vb[8];
#pragma unroll
          for (int k = 0; k < 8; k++) va[k] = fa ? ones : ldg128_stream(pa + (k0 + k) * 32);
#pragma unroll
          for (int k = 0; k < 8; k++) vb[k] = fb ? ones : ldg128_stream(pb + (k0 + k) * 32);
#pragma unroll
          for (int k = 0; k < 8; k++) cnt += popc128(and128(va[k], vb[k]));
        }
      }
    } else {
      // ---- generic accumulator machine over bitmap / constant leaves ----
      for (int k0 = 0; k0 < 16; k0 += kUnroll) {
        uint4 acc[kUnroll];
        uint4 stack[kMaxStack][kUnroll];
        int sp = 0;
        bool live = false;
        for (uint32_t i = 0; i < pg.n_instr; i++) {
          const Instr in = ins[i];
          uint4 v[kUnroll];
          if (in.op <= I_RANDNOT_LEAF) {
            const ContainerDesc d = s_desc[warp][in.leaf];
            if (d.n == 0) {
#pragma unroll
              for (int k = 0; k < kUnroll; k++) v[k] = make_uint4(0, 0, 0, 0);
            } else if (d.n == kFullN && desc_type(d) != kBitmap) {
#pragma unroll
              for (int k = 0; k < kUnroll; k++) v[k] = make_uint4(~0u, ~0u, ~0u, ~0u);
            } else {
              const uint4* ptr = reinterpret_cast<const uint4*>(d.addr) + lane;
#pragma unroll
              for (int k = 0; k < kUnroll; k++) v[k] = ldg128_stream(ptr + (k0 + k) * 32);
            }
            if (in.op == I_LOAD) {
              if (live) {
#pragma unroll
                for (int k = 0; k < kUnroll; k++) stack[sp][k] = acc[k];
                sp++;
              }
#pragma unroll
              for (int k = 0; k < kUnroll; k++) acc[k] = v[k];
              live = true;
            } else {
#pragma unroll
              for (int k = 0; k < kUnroll; k++) acc[k] = apply_op(in.op, acc[k], v[k]);
            }
          } else {
            sp--;
#pragma unroll
            for (int k = 0; k < kUnroll; k++) acc[k] = apply_op(in.op, acc[k], stack[sp][k]);
          }
        }
#pragma unroll
        for (int k = 0; k < kUnroll; k++) {
          cnt += popc128(acc[k]);
          if (want_dense) reinterpret_cast<uint4*>(dout)[lane + (k0 + k) * 32] = acc[k];
        }
      }
    }
    cnt = warp_sum(cnt);
    if (lane == 0) {
      if (p.out_counts && cnt) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)cnt);
      if (p.dense_n) p.dense_n[item] = cnt;
    }
    __syncwarp();
  }
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
  if (p.push.n_dst) {  // the CTA that retires last sends this GPU's counts to the peers: compute and reduction in one kernel
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(p.push.ticket, 1u) == gridDim.x - 1 ? 1 : 0;
    __syncthreads();
    if (s_last) {
      __threadfence();
      peer_push(p.push, threadIdx.x, blockDim.x);
    }
  }
}

// ---------------------------------------------------------------------------------------
// K2 (count form) for very sparse rows: one LANE per item. A field with many rows and few bits per row (BASELINE
// config C5: p = 1e-4, 6.5 values per container) has items of two 13-byte arrays; a warp per item spends its time on
// the three dependent loads of leaf resolution with 31 lanes idle. Here every lane resolves its own item — lanes
// 0-15 / 16-31 are the 16 container columns of two consecutive shards, so the descriptor reads coalesce — and, when
// both operands are arrays of at most 16 values (or empty / full), counts the pair in registers: the values of both
// sides come in with at most two 16-byte loads each and meet in a 16 x 16 compare (intersectionCountArrayArray,
// roaring.go:3115, without the merge loop's dependent loads); Union / Xor / Difference counts follow from the
// cardinalities. Everything else — longer arrays, runs, bitmaps, programs of more than two steps — is appended to a
// list that eval_fast_kernel then walks instead of the full item space. Counts are kept per lane for as long as the
// lane's items add into the same output counter and are added once per group of lanes that share it.
// Launched only when the leaf families' containers average <= 64 B (host: tiny_of).
// ---------------------------------------------------------------------------------------
constexpr uint32_t kTinyMax = 16;

__device__ __forceinline__ void tiny_values(const ContainerDesc& d, uint32_t sentinel, uint32_t (&v)[kTinyMax]) {
  const uint32_t len = d.n == 0 ? 0u : desc_len(d);
  uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
  if (len) q0 = ldg128(reinterpret_cast<const void*>(d.addr));
  if (len > 8) q1 = ldg128(reinterpret_cast<const void*>(d.addr + 16));
  const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
  for (int k = 0; k < int(kTinyMax); k++) {
    const uint32_t x = (k & 1) ? (w[k >> 1] >> 16) : (w[k >> 1] & 0xffffu);
    v[k] = uint32_t(k) < len ? x : sentinel + uint32_t(k);  // sentinels are above 16 bits and differ per side: never equal
  }
}

__global__ void __launch_bounds__(256) eval_tiny_kernel(EvalParams p) {
  const int lane = threadIdx.x & 31;
  const uint64_t n_items = uint64_t(p.n_queries) * p.n_shards * kRowContainers;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const bool lists = p.sparse_out != nullptr;  // the rows are wanted (one query): short results go out as value lists, see EvalParams
  unsigned long long bytes_acc = 0, acc = 0;
  uint64_t acc_off = ~0ull;
  for (uint64_t base = uint64_t(blockIdx.x) * blockDim.x + (threadIdx.x & ~31u); base < n_items; base += stride) {  // warp-uniform trips
    const uint64_t item = base + lane;
    const bool valid = item < n_items;
    bool defer = false;
    uint32_t r = 0;
    uint64_t off = acc_off;
    if (valid) {
      const uint32_t c = uint32_t(item & 15);
      const uint64_t qs = item >> 4;
      const uint32_t s = uint32_t(qs % p.n_shards), q = uint32_t(qs / p.n_shards);
      const Prog pg = p.progs[q];
      const Instr* ins = p.instrs + pg.instr_off;
      off = uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride;
      if (pg.n_instr == 1 && p.leaves[pg.leaf_off + ins[0].leaf].kind != L_TEMP) {  // Count(Row): roaring.go:405
        const ContainerDesc d = resolve_leaf(p, p.leaves[pg.leaf_off + ins[0].leaf], s, c);
        r = d.n;
        if (lists) {  // the row itself is wanted: a short array IS its value list
          if (d.n == 0) p.dense_n[item] = 0;
          else if (desc_type(d) == kArray && desc_len(d) <= kSparseListMax) {
            const uint4* src = reinterpret_cast<const uint4*>(d.addr);
            uint4* dst = reinterpret_cast<uint4*>(p.sparse_out + item * kSparseListMax);
            const uint32_t len = desc_len(d);
#pragma unroll
            for (int k = 0; k < int(kSparseListMax / 8); k++)
              if (uint32_t(k) * 8 < len) dst[k] = ldg128(src + k);
            p.dense_n[item] = len | kSparseFlag;
          } else defer = true;
        }
        if (!defer) bytes_acc += desc_algo_bytes(d);
      } else if (pg.n_instr == 2 && ins[0].op == I_LOAD && ins[1].op >= I_AND_LEAF && ins[1].op <= I_ANDNOT_LEAF &&
                 p.leaves[pg.leaf_off + ins[0].leaf].kind != L_TEMP && p.leaves[pg.leaf_off + ins[1].leaf].kind != L_TEMP) {
        const ContainerDesc a = resolve_leaf(p, p.leaves[pg.leaf_off + ins[0].leaf], s, c);
        const ContainerDesc b = resolve_leaf(p, p.leaves[pg.leaf_off + ins[1].leaf], s, c);
        const bool trivial = a.n == 0 || b.n == 0 || a.n == kFullN || b.n == kFullN;  // the cardinalities decide
        const bool both_tiny = desc_type(a) == kArray && desc_len(a) <= kTinyMax && desc_type(b) == kArray && desc_len(b) <= kTinyMax;
        if (lists) {
          // the result row is wanted: an intersection or a difference of two tiny arrays is a subset of the first one, kept
          // in its order; unions, symmetric differences and anything with a full or a long operand go the general way
          const uint8_t op = ins[1].op;
          if (a.n == 0 || (op == I_AND_LEAF && b.n == 0)) p.dense_n[item] = 0;
          else if (both_tiny && (op == I_AND_LEAF || op == I_ANDNOT_LEAF)) {
            uint32_t va[kTinyMax], vb[kTinyMax];
            tiny_values(a, 0x10000u, va);
            tiny_values(b, 0x20000u, vb);
            uint16_t* lst = p.sparse_out + item * kSparseListMax;
            const uint32_t la = desc_len(a);
            uint32_t k = 0;
#pragma unroll
            for (int i = 0; i < int(kTinyMax); i++) {
              bool hit = false;
#pragma unroll
              for (int j = 0; j < int(kTinyMax); j++) hit |= va[i] == vb[j];
              if (uint32_t(i) < la && hit == (op == I_AND_LEAF)) lst[k++] = uint16_t(va[i]);
            }
            r = k;
            p.dense_n[item] = k ? (k | kSparseFlag) : 0u;
            bytes_acc += desc_algo_bytes(a) + desc_algo_bytes(b);
          } else defer = true;
        } else if (!trivial && !both_tiny) defer = true;
        else {
          uint32_t ic;
          if (a.n == 0 || b.n == 0) ic = 0;
          else if (a.n == kFullN) ic = b.n;  // roaring.go:3079-3084
          else if (b.n == kFullN) ic = a.n;
          else {
            uint32_t va[kTinyMax], vb[kTinyMax];
            tiny_values(a, 0x10000u, va);
            tiny_values(b, 0x20000u, vb);
            ic = 0;
#pragma unroll
            for (int i = 0; i < int(kTinyMax); i++)
#pragma unroll
              for (int j = 0; j < int(kTinyMax); j++) ic += va[i] == vb[j] ? 1u : 0u;
          }
          switch (ins[1].op) {
            case I_AND_LEAF: r = ic; break;
            case I_OR_LEAF: r = a.n + b.n - ic; break;
            case I_XOR_LEAF: r = a.n + b.n - 2 * ic; break;
            default: r = a.n - ic; break;  // I_ANDNOT_LEAF
          }
          bytes_acc += desc_algo_bytes(a) + desc_algo_bytes(b);
        }
      } else defer = true;
    }
    // deferred items: one list slot each, claimed with one atomic per warp
    const uint32_t dmask = __ballot_sync(0xffffffffu, defer);
    if (dmask) {
      uint32_t slot0 = 0;
      if (lane == 0) slot0 = atomicAdd(p.item_count, uint32_t(__popc(dmask)));
      slot0 = __shfl_sync(0xffffffffu, slot0, 0);
      if (defer) p.item_list[slot0 + __popc(dmask & ((1u << lane) - 1u))] = uint32_t(item);
    }
    // counts: flush when some lane moves on to another output counter
    const bool moved = valid && !defer && r && off != acc_off;
    if (__any_sync(0xffffffffu, moved && acc != 0)) {
      const uint32_t grp = __match_any_sync(0xffffffffu, acc_off);
      const unsigned long long tot = warp_group_sum(grp, acc);
      if (lane == __ffs(grp) - 1 && tot && p.out_counts) atomicAdd(p.out_counts + acc_off, tot);
      acc = 0;
    }
    if (valid && !defer && r) { acc_off = off; acc += r; }
  }
  {
    const uint32_t grp = __match_any_sync(0xffffffffu, acc_off);
    const unsigned long long tot = warp_group_sum(grp, acc);
    if (lane == __ffs(grp) - 1 && tot && p.out_counts) atomicAdd(p.out_counts + acc_off, tot);
  }
  bytes_acc = warp_sum64(bytes_acc);
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

// ---------------------------------------------------------------------------------------
// shared-memory staging of array / run containers (K2-K5 generalised): a container is
// expanded into a window [w0, w0 + nwords) of 64-bit words in shared memory, after which
// every 3x3 type pair and every op of roaring.go:3078-4957 is the same word-parallel
// boolean function. The whole CTA cooperates; the window must be pre-zeroed.
// ---------------------------------------------------------------------------------------
// XOR: the container's bits are flipped in the window instead of set (the window need not be zero; operands staged by
// different warps at the same time commute).
template <bool XOR = false>
__device__ __forceinline__ void stage_container(const ContainerDesc& d, uint32_t* win32, uint32_t w0, uint32_t nwords,
                                                int tid, int nthreads) {
  const uint32_t lo = w0 * 64, hi = (w0 + nwords) * 64;  // value range [lo, hi)
  const uint32_t type = desc_type(d), len = desc_len(d);
  if (type == kArray) {
    const uint16_t* a = reinterpret_cast<const uint16_t*>(d.addr);
    uint32_t b = 0, e = len;
    if (lo > 0) {  // lower_bound(lo)
      uint32_t l = 0, h = len;
      while (l < h) { uint32_t m = (l + h) >> 1; if (a[m] < lo) l = m + 1; else h = m; }
      b = l;
    }
    if (hi < 65536u) {  // lower_bound(hi)
      uint32_t l = b, h = len;
      while (l < h) { uint32_t m = (l + h) >> 1; if (a[m] < hi) l = m + 1; else h = m; }
      e = l;
    }
    // Eight consecutive values per thread and step (one 16-byte load when aligned); a thread's values that
    // fall into the same 32-bit word are merged in registers before the shared-memory atomicOr.
    for (uint32_t i = b + tid * 8; i < e; i += nthreads * 8) {
      uint32_t cur_w = 0xffffffffu, cur_bits = 0;
      const uint32_t m = min(8u, e - i);
      uint32_t vals[8];
      if (m == 8 && ((reinterpret_cast<uintptr_t>(a + i) & 15) == 0)) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(a + i));
        vals[0] = q.x & 0xffff; vals[1] = q.x >> 16; vals[2] = q.y & 0xffff; vals[3] = q.y >> 16;
        vals[4] = q.z & 0xffff; vals[5] = q.z >> 16; vals[6] = q.w & 0xffff; vals[7] = q.w >> 16;
      } else {
#pragma unroll
        for (int k = 0; k < 8; k++) vals[k] = k < int(m) ? uint32_t(__ldg(a + i + k)) : 0u;
      }
#pragma unroll
      for (int k = 0; k < 8; k++) {
        if (k < int(m)) {
          const uint32_t v = vals[k] - lo, w = v >> 5;
          if (w != cur_w) {
            if (cur_w != 0xffffffffu) { if (XOR) atomicXor(&win32[cur_w], cur_bits); else atomicOr(&win32[cur_w], cur_bits); }
            cur_w = w; cur_bits = 0;
          }
          cur_bits |= 1u << (v & 31);
        }
      }
      if (cur_w != 0xffffffffu) { if (XOR) atomicXor(&win32[cur_w], cur_bits); else atomicOr(&win32[cur_w], cur_bits); }
    }
  } else if (type == kRun) {
    const ushort2* r = reinterpret_cast<const ushort2*>(d.addr);
    uint32_t b = 0;
    if (lo > 0) {  // first run with last >= lo
      uint32_t l = 0, h = len;
      while (l < h) { uint32_t m = (l + h) >> 1; if (r[m].y < lo) l = m + 1; else h = m; }
      b = l;
    }
    uint32_t e = len;
    if (hi < 65536u) {  // first run with start >= hi
      uint32_t l = b, h = len;
      while (l < h) { uint32_t m = (l + h) >> 1; if (r[m].x < hi) l = m + 1; else h = m; }
      e = l;
    }
    // One run per thread (coalesced loads of the run list). The two edge words of a run can be shared with
    // a neighbouring run: shared-memory atomicOr. The words strictly inside a run belong to that run alone
    // (runs are disjoint and not adjacent): plain stores — by the owning thread when there are few, by the
    // whole warp (coalesced, run after run) when a long run has hundreds of them.
    const int lane = tid & 31;
    for (uint32_t base = b; base < e; base += nthreads) {  // warp-uniform trip count: nthreads is a multiple of 32
      const uint32_t i = base + tid;
      uint32_t ws = 0, we = 0;
      if (i < e) {
        const ushort2 iv = __ldg(r + i);
        uint32_t s = iv.x, l = iv.y;
        if (s < lo) s = lo;
        if (l >= hi) l = hi - 1;
        s -= lo; l -= lo;
        ws = s >> 5; we = l >> 5;
        const uint32_t ms = ~0u << (s & 31), me = ~0u >> (31 - (l & 31));
        if (XOR) { if (ws == we) atomicXor(&win32[ws], ms & me); else { atomicXor(&win32[ws], ms); atomicXor(&win32[we], me); } }
        else if (ws == we) atomicOr(&win32[ws], ms & me);
        else { atomicOr(&win32[ws], ms); atomicOr(&win32[we], me); }
      }
      const bool long_run = we > ws + 1 + 16;  // short interiors: by the owning thread; long ones: by the warp
      if (!long_run) for (uint32_t k = ws + 1; k < we; k++) win32[k] = XOR ? ~win32[k] : ~0u;
      uint32_t todo = __ballot_sync(0xffffffffu, long_run);
      while (todo) {
        const int src = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint32_t k0 = __shfl_sync(0xffffffffu, ws, src) + 1, k1 = __shfl_sync(0xffffffffu, we, src);
        for (uint32_t k = k0 + lane; k < k1; k += 32) win32[k] = XOR ? ~win32[k] : ~0u;
      }
    }
  }
}

constexpr int kStagedThreads = 256;
constexpr int kSlotWords = 4096;  // 32 KiB of staging windows per CTA

// ---------------------------------------------------------------------------------------
// K6-staged: one CTA per deferred item. Array/run operands are staged into shared-memory
// windows (all of them at once; the column is processed in passes when more than four
// need staging), then the accumulator machine runs word-parallel over bitmap (global),
// staged (shared) and constant operands.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kStagedThreads) eval_staged_kernel(EvalParams p) {
  __shared__ ContainerDesc s_desc[kMaxLeaves];
  __shared__ int s_slot[kMaxLeaves];
  __shared__ int s_nslots;
  __shared__ uint32_t s_red[kStagedThreads / 32];
  __shared__ unsigned long long s_base[kMaxLeaves];
  __shared__ uint32_t s_stride[kMaxLeaves];
  __shared__ __align__(16) uint4 s_const[2];
  __shared__ __align__(16) uint64_t s_win[kSlotWords];
  const int tid = threadIdx.x;
  const uint32_t n_slow = p.fold_all ? p.n_queries * p.n_shards * kRowContainers : *p.slow_count;
  if (tid == 0) { s_const[0] = make_uint4(0, 0, 0, 0); s_const[1] = make_uint4(~0u, ~0u, ~0u, ~0u); }

  for (uint32_t it = blockIdx.x; it < n_slow; it += gridDim.x) {
    const uint64_t item = p.fold_all ? uint64_t(it) : uint64_t(p.slow_list[it]);
    const uint32_t c = uint32_t(item & 15);
    const uint64_t qs = item >> 4;
    const uint32_t s = uint32_t(qs % p.n_shards);
    const uint32_t q = uint32_t(qs / p.n_shards);
    const Prog pg = p.progs[q];
    const Instr* ins = p.instrs + pg.instr_off;

    __syncthreads();  // previous item fully consumed
    if (tid < 32) {  // warp 0: descriptors + staging-slot assignment (n_leaves <= 32)
      ContainerDesc d; d.addr = 0; d.n = 0; d.meta = 0;
      if (tid < pg.n_leaves) {
        d = (p.slow_descs && !p.fold_all) ? p.slow_descs[uint64_t(it) * p.slow_desc_stride + tid] : resolve_leaf(p, p.leaves[pg.leaf_off + tid], s, c);
        s_desc[tid] = d;
      }
      const bool need = tid < pg.n_leaves && !is_const_or_bitmap(d);
      const uint32_t mask = __ballot_sync(0xffffffffu, need);
      if (tid < pg.n_leaves) s_slot[tid] = need ? int(__popc(mask & ((1u << tid) - 1u))) : -1;
      __syncwarp();
      if (tid == 0) s_nslots = fold_can_take(pg, ins, s_desc) ? -1 : __popc(mask);  // -1: eval_fold_kernel's item
    }
    __syncthreads();
    if (s_nslots < 0) continue;
    // pull the bitmap operands towards L2 while the array / run operands are being staged
    for (uint32_t l = 0; l < pg.n_leaves; l++) {
      const ContainerDesc d = s_desc[l];
      if (d.n != 0 && desc_type(d) == kBitmap && tid < kBitmapBytes / 128)
        asm volatile("prefetch.global.L2 [%0];" :: "l"(d.addr + uint64_t(tid) * 128));
    }
    const int nslots = s_nslots;
    uint32_t chunk = kBitmapWords;
    while (chunk * nslots > kSlotWords) chunk >>= 1;  // nslots <= 32 -> chunk >= 128

    uint32_t cnt = 0;
    uint8_t* dout = p.dense_out ? p.dense_out + item * kBitmapBytes : nullptr;
    for (uint32_t w0 = 0; w0 < kBitmapWords; w0 += chunk) {
      // zero + stage
      for (uint32_t i = tid; i < chunk * nslots / 2; i += kStagedThreads) reinterpret_cast<uint4*>(s_win)[i] = make_uint4(0, 0, 0, 0);
      __syncthreads();
      for (uint32_t l = 0; l < pg.n_leaves; l++)
        if (s_slot[l] >= 0) stage_container(s_desc[l], reinterpret_cast<uint32_t*>(s_win + s_slot[l] * chunk), w0, chunk, tid, kStagedThreads);
      __syncthreads();
      // operand sources for this pass, as (generic address, stride): global bitmap words, a staged
      // shared-memory window, or a 16 B constant (all zeros / all ones) read with stride 0
      if (tid < pg.n_leaves) {
        const ContainerDesc d = s_desc[tid];
        unsigned long long base; uint32_t stride = 16;
        if (d.n == 0) { base = reinterpret_cast<unsigned long long>(&s_const[0]); stride = 0; }
        else if (s_slot[tid] >= 0) base = reinterpret_cast<unsigned long long>(s_win + s_slot[tid] * chunk);
        else if (d.n == kFullN && desc_type(d) != kBitmap) { base = reinterpret_cast<unsigned long long>(&s_const[1]); stride = 0; }
        else base = d.addr + uint64_t(w0) * 8;
        s_base[tid] = base; s_stride[tid] = stride;
      }
      __syncthreads();
      // evaluate: the instruction loop is outermost so that each decode serves both of the
      // thread's 128-bit chunks (chunk/2 uint4 per operand, kStagedThreads apart)
      for (uint32_t j0 = tid; j0 < chunk / 2; j0 += 2 * kStagedThreads) {
        const uint32_t j1 = j0 + kStagedThreads;
        const bool two = j1 < chunk / 2;
        uint4 acc0 = make_uint4(0, 0, 0, 0), acc1 = acc0;
        uint4 stack0[kMaxStack], stack1[kMaxStack];
        int sp = 0;
        bool live = false;
        for (uint32_t i = 0; i < pg.n_instr; i++) {
          const Instr in = ins[i];
          if (in.op <= I_RANDNOT_LEAF) {
            const unsigned long long base = s_base[in.leaf];
            const uint32_t stride = s_stride[in.leaf];
            const uint4 v0 = *reinterpret_cast<const uint4*>(base + uint64_t(j0) * stride);
            const uint4 v1 = two ? *reinterpret_cast<const uint4*>(base + uint64_t(j1) * stride) : make_uint4(0, 0, 0, 0);
            if (in.op == I_LOAD) {
              if (live) { stack0[sp] = acc0; stack1[sp] = acc1; sp++; }
              acc0 = v0; acc1 = v1;
              live = true;
            } else {
              acc0 = apply_op(in.op, acc0, v0);
              acc1 = apply_op(in.op, acc1, v1);
            }
          } else {
            --sp;
            acc0 = apply_op(in.op, acc0, stack0[sp]);
            acc1 = apply_op(in.op, acc1, stack1[sp]);
          }
        }
        cnt += popc128(acc0);
        if (dout) reinterpret_cast<uint4*>(dout)[w0 / 2 + j0] = acc0;
        if (two) {
          cnt += popc128(acc1);
          if (dout) reinterpret_cast<uint4*>(dout)[w0 / 2 + j1] = acc1;
        }
      }
      __syncthreads();
    }
    cnt = warp_sum(cnt);
    if ((tid & 31) == 0) s_red[tid >> 5] = cnt;
    __syncthreads();
    if (tid == 0) {
      uint32_t tot = 0;
      for (int w = 0; w < kStagedThreads / 32; w++) tot += s_red[w];
      if (p.out_counts && tot) atomicAdd(p.out_counts + uint64_t(q) * p.out_qstride + uint64_t(s) * p.out_sstride, (unsigned long long)tot);
      if (p.dense_n) p.dense_n[item] = tot;
    }
  }
}

// ---------------------------------------------------------------------------------------
// K2-K6, fold form: one CTA per deferred item whose program is a flat left fold
//   acc = leaf0 ; acc = acc <op> leaf1 ; acc = acc <op> leaf2 ; ...
// (what every n-ary PQL call over rows compiles to, executor.go:1668-1733,1003-1025). The accumulator is ONE 8 KiB
// shared-memory window; each operand is applied to it IN ITS OWN ENCODING, so an operand costs its payload, not
// 65536 bits: the type pairs of roaring.go:3078-4957 become
//                bitmap (1024 words)     array (n x u16)                       run list
//   load         window = g              the array's bits, into a zero window  window = run bitmap
//   OR           w |= g                  set the values' bits                  w |= run bitmap    (union*        :3470-3839)
//   XOR          w ^= g                  flip them                             w ^= run bitmap    (xor*          :4266-4957)
//   AND-NOT      w &= ~g                 clear them                            w &= ~run bitmap   (difference*   :3920-4205)
//   AND          w &= g                  AND with the array's bitmap (in S)    w &= run bitmap    (intersect*    :3254-3418)
// Payloads are contiguous 16 B aligned streams, so the bulk copy engine moves each one global -> shared with one
// cp.async.bulk issued by one thread (local HBM only, no multicast), up to kFoldStages operands ahead of the one being
// applied; completion is an mbarrier transaction count.
//   Sorted elements -> bits without atomics: a thread takes a slice of 8 consecutive values (4 consecutive runs). A
// bitmap word belongs to the slice that holds the word's FIRST element: a thread drops its leading elements that fall
// into the word of the element in front of its slice, and reads on past the end of its slice while elements stay in
// its last word. Every word is then written by exactly one thread, once, with a plain store.
//   Run list -> bitmap without touching a word per run: starts go into one zeroed scratch bitmap S, the positions
// behind the runs' last values into another, E, and the run bitmap is the 65536-bit integer E - S: (1 << (last + 1))
// - (1 << start) is exactly the ones of one run, and disjoint runs do not interfere. A thread owns 8 consecutive
// words, subtracts them with one borrow chain (9 instructions), and gets its borrow-in — "a run is open at my first
// position" — as the parity of the bits in front of it: a warp ballot plus one shared-memory hop across the warps.
// S and E are kept zero between uses (the subtraction clears what it reads).
// The count is the popcount of the window; no result container is typed or written unless the caller wants the row.
// ---------------------------------------------------------------------------------------
constexpr int kFoldThreads = 256;                      // consumer threads (8 warps) ...
constexpr int kFoldCtaThreads = kFoldThreads + 32;     // ... plus one producer warp
#ifndef PB200_FOLD_STAGES
#define PB200_FOLD_STAGES 2
#endif
#ifndef PB200_FOLD_CTAS
#define PB200_FOLD_CTAS 5
#endif
constexpr int kFoldStages = PB200_FOLD_STAGES;          // 2: applying an operand takes longer than fetching the next one
constexpr int kFoldMeta = 2;                            // items whose descriptors are resolved ahead
constexpr int kFoldFixed = 3;                           // window, S, E
constexpr int kFoldSmem = (kFoldFixed + kFoldStages) * kBitmapBytes;  // dynamic shared memory
constexpr uint32_t kWords32 = kBitmapBytes / 4;
// barrier among the consumer warps only (the producer warp never joins it)
__device__ __forceinline__ void fold_bar() { asm volatile("bar.sync 1, %0;" :: "n"(kFoldThreads) : "memory"); }

// shared-window accesses by 32-bit shared address: the helpers below are handed addresses, not generic pointers, so no
// store pays for a generic -> shared conversion
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) { uint32_t v; asm volatile("{ .reg .u16 h; ld.shared.u16 h, [%1]; cvt.u32.u16 %0, h; }" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }

// element i of a staged payload (at shared address `base`) as a bit position
struct PosArray { uint32_t base; __device__ __forceinline__ uint32_t operator()(uint32_t i) const { return lds_u16(base + 2u * i); } };
struct PosRunStart { uint32_t base; __device__ __forceinline__ uint32_t operator()(uint32_t i) const { return lds_u16(base + 4u * i); } };
struct PosRunEnd { uint32_t base; __device__ __forceinline__ uint32_t operator()(uint32_t i) const { return lds_u16(base + 4u * i + 2u) + 1u; } };  // 65536 behind a run that ends the container

enum FoldEdit : int { E_STORE = 0, E_SET = 1, E_FLIP = 2, E_CLEAR = 3 };  // E_STORE: the target is zero, nothing to read
template <int MODE>
__device__ __forceinline__ void fold_put(uint32_t bm, uint32_t w, uint32_t bits) {
  const uint32_t a = bm + 4u * w;
  if (MODE == E_STORE) sts_u32(a, bits);
  else {
    const uint32_t old = lds_u32(a);
    sts_u32(a, MODE == E_SET ? (old | bits) : MODE == E_FLIP ? (old ^ bits) : (old & ~bits));
  }
}

// One slice: elements [first, first + cnt) are in v[0 .. cnt), cnt <= N; `pos` reads any element of the sequence.
// MODE: how a word of the bitmap at shared address bm is edited (E_STORE for the zeroed S / E, else the window).
// GUARD: the last element may be position 65536, which has no word.
template <int N, bool FULL, int MODE, bool GUARD, class POS>
__device__ __forceinline__ void fold_set_slice(uint32_t bm, const uint32_t (&v)[N], uint32_t cnt, uint32_t first, uint32_t len, const POS& pos) {
  // the word in front of the slice belongs to an earlier slice: bits gathered for it are dropped. Words ascend, so
  // "cur_w is mine" is "cur_w != foreign" at every flush — no flag is carried through the loop.
  const uint32_t foreign = first ? (pos(first - 1) >> 5) : 0xffffffffu;
  uint32_t cur_w = foreign, cur_b = 0;
#pragma unroll
  for (int k = 0; k < N; k++) {
    if (FULL || uint32_t(k) < cnt) {
      const uint32_t w = v[k] >> 5;
      const uint32_t bit = __funnelshift_l(0u, 1u, v[k]);  // 1 << (v & 31)
      const bool chg = w != cur_w;
      if (chg && cur_w != foreign) fold_put<MODE>(bm, cur_w, cur_b);
      cur_b = chg ? bit : (cur_b | bit);
      cur_w = w;
    }
  }
  if (cur_w == foreign) return;  // the whole slice lies in a word of an earlier slice
  for (uint32_t i = first + cnt; i < len; i++) {  // the rest of the last word: at most 31 elements
    const uint32_t x = pos(i);
    if ((x >> 5) != cur_w) break;
    cur_b |= __funnelshift_l(0u, 1u, x);
  }
  if (!GUARD || cur_w < kWords32) fold_put<MODE>(bm, cur_w, cur_b);
}

// The bits of a sorted array container set / flipped / cleared in the window, or stored into the zeroed scratch S.
template <int MODE>
__device__ __noinline__ void fold_array_set(uint32_t bm, uint32_t st, uint32_t len, int tid) {
  const PosArray pos{st};
  for (uint32_t first = uint32_t(tid) * 8u; first < len; first += kFoldThreads * 8u) {
    uint4 q;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(st + 2u * first) : "memory");
    const uint32_t v[8] = {q.x & 0xffffu, q.x >> 16, q.y & 0xffffu, q.y >> 16, q.z & 0xffffu, q.z >> 16, q.w & 0xffffu, q.w >> 16};
    const uint32_t cnt = len - first;
    if (cnt >= 8u) fold_set_slice<8, true, MODE, false>(bm, v, 8u, first, len, pos);
    else fold_set_slice<8, false, MODE, false>(bm, v, cnt, first, len, pos);
  }
}

// r = e - s - borrow over 8 words, least significant first
__device__ __forceinline__ void fold_sub8(uint32_t (&r)[8], const uint32_t (&e)[8], const uint32_t (&s)[8], uint32_t borrow) {
  asm("{\n\t"
      ".reg .u32 t;\n\t"
      "sub.cc.u32 t, 0, %24;\n\t"
      "subc.cc.u32 %0, %8, %16;\n\t"
      "subc.cc.u32 %1, %9, %17;\n\t"
      "subc.cc.u32 %2, %10, %18;\n\t"
      "subc.cc.u32 %3, %11, %19;\n\t"
      "subc.cc.u32 %4, %12, %20;\n\t"
      "subc.cc.u32 %5, %13, %21;\n\t"
      "subc.cc.u32 %6, %14, %22;\n\t"
      "subc.u32 %7, %15, %23;\n\t"
      "}"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
      : "r"(e[0]), "r"(e[1]), "r"(e[2]), "r"(e[3]), "r"(e[4]), "r"(e[5]), "r"(e[6]), "r"(e[7]),
        "r"(s[0]), "r"(s[1]), "r"(s[2]), "r"(s[3]), "r"(s[4]), "r"(s[5]), "r"(s[6]), "r"(s[7]), "r"(borrow));
}

// A run list applied to the window (OP: I_LOAD / I_OR_LEAF / I_AND_LEAF). S, E: the zeroed scratch bitmaps; left zeroed.
__device__ __noinline__ void fold_runs(const int OP, uint32_t* win32, uint32_t* S, uint32_t* E, uint32_t st, uint32_t len, uint32_t* s_par, int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  const PosRunStart ps{st};
  const PosRunEnd pe{st};
  const uint32_t S_a = smem_u32(S), E_a = smem_u32(E);
  for (uint32_t first = uint32_t(tid) * 4u; first < len; first += kFoldThreads * 4u) {
    uint4 q;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(st + 4u * first) : "memory");
    const uint32_t vs[4] = {q.x & 0xffffu, q.y & 0xffffu, q.z & 0xffffu, q.w & 0xffffu};
    const uint32_t ve[4] = {(q.x >> 16) + 1u, (q.y >> 16) + 1u, (q.z >> 16) + 1u, (q.w >> 16) + 1u};
    const uint32_t cnt = len - first;
    if (cnt >= 4u) {
      fold_set_slice<4, true, E_STORE, false>(S_a, vs, 4u, first, len, ps);
      fold_set_slice<4, true, E_STORE, true>(E_a, ve, 4u, first, len, pe);
    } else {
      fold_set_slice<4, false, E_STORE, false>(S_a, vs, cnt, first, len, ps);
      fold_set_slice<4, false, E_STORE, true>(E_a, ve, cnt, first, len, pe);
    }
  }
  fold_bar();
  // words 8 tid .. 8 tid + 7 of E - S
  uint4* S4 = reinterpret_cast<uint4*>(S);
  uint4* E4 = reinterpret_cast<uint4*>(E);
  const uint4 s0 = S4[2 * tid], s1 = S4[2 * tid + 1], e0 = E4[2 * tid], e1 = E4[2 * tid + 1];
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  S4[2 * tid] = zero4; S4[2 * tid + 1] = zero4; E4[2 * tid] = zero4; E4[2 * tid + 1] = zero4;
  const uint32_t s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  const uint32_t e[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
  uint32_t par = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) par ^= s[k] ^ e[k];
  const uint32_t votes = __ballot_sync(0xffffffffu, (__popc(par) & 1u) != 0);
  if (lane == 0) s_par[warp] = __popc(votes) & 1u;
  fold_bar();
  const uint32_t wvotes = __ballot_sync(0xffffffffu, lane < kFoldThreads / 32 && s_par[lane & (kFoldThreads / 32 - 1)] != 0);
  // starts minus ends in front of this thread's words: 0 or 1 — a run is open there
  const uint32_t borrow = (__popc(votes & ((1u << lane) - 1u)) + __popc(wvotes & ((1u << warp) - 1u))) & 1u;
  uint32_t r[8];
  fold_sub8(r, e, s, borrow);
  uint4* W4 = reinterpret_cast<uint4*>(win32);
  if (OP == I_LOAD) {  // the run bitmap is the window: nothing to read, nothing to clear beforehand
    W4[2 * tid] = make_uint4(r[0], r[1], r[2], r[3]);
    W4[2 * tid + 1] = make_uint4(r[4], r[5], r[6], r[7]);
    return;
  }
  const uint4 a0 = W4[2 * tid], a1 = W4[2 * tid + 1];
  const uint32_t a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
  uint32_t o[8];
  if (OP == I_OR_LEAF) {
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = a[k] | r[k];
  } else if (OP == I_AND_LEAF) {
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = a[k] & r[k];
  } else if (OP == I_XOR_LEAF) {
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = a[k] ^ r[k];
  } else {
#pragma unroll
    for (int k = 0; k < 8; k++) o[k] = a[k] & ~r[k];
  }
  W4[2 * tid] = make_uint4(o[0], o[1], o[2], o[3]);
  W4[2 * tid + 1] = make_uint4(o[4], o[5], o[6], o[7]);
}

// Payload bytes the copy engine moves for an operand (a multiple of 16; 0 = constant operand, nothing to move).
__device__ __forceinline__ uint32_t fold_bytes(const ContainerDesc& d) {
  if (d.n == 0) return 0;
  const uint32_t type = desc_type(d), len = desc_len(d);
  if (type == kBitmap) return kBitmapBytes;
  if (d.n == kFullN) return 0;
  return type == kArray ? ((len * 2 + 15) & ~15u) : ((len * 4 + 15) & ~15u);
}

struct FoldMeta {  // one deferred item, resolved by the producer warp
  uint64_t item;
  uint32_t q, s, n_instr, take;
  Instr ins[kMaxInstr];
  ContainerDesc desc[kMaxLeaves];
};

// The producer warp of the fold evaluators (see eval_fold_kernel): walks the CTA's items one ahead of the consumers,
// resolves their descriptors into the meta ring and keeps the copy engine busy.
__device__ __forceinline__ void fold_producer(const EvalParams& p, uint32_t n_slow, int lane, uint8_t* s_dyn, FoldMeta* s_meta,
                                              uint64_t* s_full, uint64_t* s_empty, uint64_t* s_mfull, uint64_t* s_mempty) {
    uint32_t g = 0;  // payload loads issued so far (lane 0)
    uint32_t j = 0;  // items handled so far
    unsigned long long bytes_acc = 0;
    for (uint32_t it = blockIdx.x; it < n_slow; it += gridDim.x, j++) {
      const uint32_t ms = j % kFoldMeta;
      if (j >= uint32_t(kFoldMeta)) mbar_wait(&s_mempty[ms], ((j / kFoldMeta) - 1u) & 1u);
      FoldMeta& m = s_meta[ms];
      const uint64_t item = p.fold_all ? uint64_t(it) : uint64_t(p.slow_list[it]);
      const uint32_t c = uint32_t(item & 15);
      const uint64_t qs = item >> 4;
      const uint32_t sidx = uint32_t(qs % p.n_shards), q = uint32_t(qs / p.n_shards);
      const Prog pg = p.progs[q];
      const Instr* ins = p.instrs + pg.instr_off;
      uint32_t bytes = 0;
      if (uint32_t(lane) < pg.n_leaves) {
        const ContainerDesc d = (p.slow_descs && !p.fold_all) ? p.slow_descs[uint64_t(it) * p.slow_desc_stride + lane]
                                                              : resolve_leaf(p, p.leaves[pg.leaf_off + lane], sidx, c);
        m.desc[lane] = d;
        bytes = desc_algo_bytes(d);
        // the staging ring is two operands deep; the payloads of the whole item are pulled towards L2 now, an item ahead
        // of the consumers, so that the ring's loads find them there
        const uint32_t fb = fold_bytes(d);
        if (fb && pg.n_instr > 1) l2_prefetch_bulk(reinterpret_cast<const void*>(d.addr), fb);
      }
      if (p.fold_all) {  // no eval_fast_kernel pass did the byte accounting
        bytes = warp_sum(bytes);
        if (lane == 0) bytes_acc += bytes;
      }
      for (uint32_t k = lane; k < pg.n_instr; k += 32) m.ins[k] = ins[k];
      __syncwarp();
      if (lane == 0) {
        bool take = fold_can_take(pg, m.ins, m.desc);
        m.item = item; m.q = q; m.s = sidx; m.n_instr = pg.n_instr; m.take = take ? 1u : 0u;
        // Count(Row): the cardinality is in the descriptor (roaring.go:405 Count = sum of n); temps carry none
        if (take && pg.n_instr == 1 && !p.dense_out && p.leaves[pg.leaf_off + m.ins[0].leaf].kind != L_TEMP) { m.take = 2u; take = false; }
        // A pure intersection with an empty operand is empty (roaring.go:3079-3087 per pair; Row.Intersect over no
        // container): when only counts are wanted nothing of the item has to be loaded. take = 3: the consumers skip it.
        if (take && pg.n_instr > 1 && !p.dense_out && !p.dense_n) {
          bool pure_and = true, any_nil = m.desc[m.ins[0].leaf].n == 0;
          for (uint32_t k = 1; k < pg.n_instr; k++) { pure_and &= m.ins[k].op == I_AND_LEAF; any_nil |= m.desc[m.ins[k].leaf].n == 0; }
          if (pure_and && any_nil) { m.take = 3u; take = false; }
        }
        mbar_arrive(&s_mfull[ms]);  // (release) the item's descriptors are published
        if (take)
          for (uint32_t k = 0; k < pg.n_instr; k++) {
            const ContainerDesc d = m.desc[m.ins[k].leaf];
            const uint32_t bytes = fold_bytes(d);
            if (!bytes) continue;
            const uint32_t slot = g % kFoldStages;
            if (g >= uint32_t(kFoldStages)) mbar_wait(&s_empty[slot], ((g / kFoldStages) - 1u) & 1u);
            mbar_arrive_expect_tx(&s_full[slot], bytes);
            tma_load_1d(s_dyn + size_t(slot + kFoldFixed) * kBitmapBytes, reinterpret_cast<const void*>(d.addr), bytes, &s_full[slot]);
            g++;
          }
      }
      __syncwarp();
    }
    if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

// Warp-specialised: warp 8 is the PRODUCER — it walks the CTA's items one ahead of the consumers, resolves their
// descriptors into a two-deep ring (meta_full / meta_empty mbarriers) and keeps the copy engine busy: every operand
// payload is one cp.async.bulk into the next free slot of a kFoldStages-deep staging ring (full / empty mbarriers),
// issued as soon as the slot is released, whatever item it belongs to. Warps 0-7 are the CONSUMERS: they never wait
// for global memory except on a cold pipeline, only for `full`, and synchronise among themselves with a named barrier.
__global__ void __launch_bounds__(kFoldCtaThreads, PB200_FOLD_CTAS) eval_fold_kernel(EvalParams p) {
  extern __shared__ __align__(128) uint8_t s_dyn[];  // [window][S][E][kFoldStages x staging], 8 KiB each
  __shared__ FoldMeta s_meta[kFoldMeta];
  __shared__ uint32_t s_red[kFoldThreads / 32], s_red2[kFoldThreads / 32];
  __shared__ __align__(8) uint64_t s_full[kFoldStages], s_empty[kFoldStages], s_mfull[kFoldMeta], s_mempty[kFoldMeta];
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t n_slow = p.fold_all ? p.n_queries * p.n_shards * kRowContainers : *p.slow_count;
  if (tid == 0) {
    for (int i = 0; i < kFoldStages; i++) { mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], 1); }
    for (int i = 0; i < kFoldMeta; i++) { mbar_init(&s_mfull[i], 1); mbar_init(&s_mempty[i], 1); }
    mbar_fence_init();
  }
  __syncthreads();

  if (tid >= kFoldThreads) {
    fold_producer(p, n_slow, lane, s_dyn, s_meta, s_full, s_empty, s_mfull, s_mempty);  // ---------------- producer warp ----------------
    return;
  }

  // ---------------- consumer warps ----------------
  // regions 0 and 1 are the window and S; they trade places when an array is LOADed (it is set into the zeroed S with
  // plain stores, the old window is cleared to become the next S — no clear-then-barrier-then-edit of the window)
  uint32_t wsel = 0;
  uint32_t* s_E = reinterpret_cast<uint32_t*>(s_dyn + 2 * kBitmapBytes);
  const uint32_t dyn_a = smem_u32(s_dyn);
  const uint4 zero4 = make_uint4(0, 0, 0, 0), ones4 = make_uint4(~0u, ~0u, ~0u, ~0u);
  for (int i = tid; i < 3 * kBitmapBytes / 16; i += kFoldThreads) reinterpret_cast<uint4*>(s_dyn)[i] = zero4;  // S and E start out zero (and stay so between uses)
  fold_bar();
  uint32_t g = 0, j = 0;
  // Counts: every warp keeps a running sum of its part of the windows for as long as the items add into the same
  // output counter (for Count that is the whole launch: all shards of a query share one), and adds it once — one
  // atomic per warp and counter instead of one per item on what is often a single address.
  uint64_t acc_off = ~0ull;
  unsigned long long acc = 0;
  auto add_count = [&](uint64_t off, uint32_t v) {
    if (off != acc_off) {
      if (lane == 0 && acc) atomicAdd(p.out_counts + acc_off, acc);
      acc_off = off; acc = 0;
    }
    acc += v;
  };
  for (uint32_t it = blockIdx.x; it < n_slow; it += gridDim.x, j++) {
    const uint32_t ms = j % kFoldMeta;
    mbar_wait(&s_mfull[ms], (j / kFoldMeta) & 1u);
    const FoldMeta& m = s_meta[ms];
    const uint32_t n_instr = m.n_instr;
    if (m.take != 1u) {  // 0: not a flat fold, eval_staged_kernel handles it; 2: Count(Row) straight from the descriptor; 3: empty intersection
      if (tid < 32 && m.take == 2u && p.out_counts)
        add_count(uint64_t(m.q) * p.out_qstride + uint64_t(m.s) * p.out_sstride, m.desc[m.ins[0].leaf].n);
      fold_bar();
      if (tid == 0) mbar_arrive(&s_mempty[ms]);
      continue;
    }
#pragma unroll 1
    for (uint32_t i = 0; i < n_instr; i++) {
      const Instr in = m.ins[i];
      const ContainerDesc d = m.desc[in.leaf];
      const uint32_t type = desc_type(d), len = desc_len(d);
      const bool nil = d.n == 0, full = d.n == kFullN && type != kBitmap;
      const bool staged = !nil && !full;
      const uint32_t slot = g % kFoldStages;
      const uint4* st4 = reinterpret_cast<const uint4*>(s_dyn + size_t(slot + kFoldFixed) * kBitmapBytes);
      const uint32_t st_a = dyn_a + (slot + kFoldFixed) * kBitmapBytes;
      if (staged) mbar_wait(&s_full[slot], (g / kFoldStages) & 1u);  // this operand's payload has landed
      uint32_t* s_win = reinterpret_cast<uint32_t*>(s_dyn + wsel * kBitmapBytes);
      uint32_t* s_S = reinterpret_cast<uint32_t*>(s_dyn + (wsel ^ 1u) * kBitmapBytes);
      uint4* win4 = reinterpret_cast<uint4*>(s_win);
      uint4* S4 = reinterpret_cast<uint4*>(s_S);
      const uint32_t win_a = dyn_a + wsel * kBitmapBytes, S_a = dyn_a + (wsel ^ 1u) * kBitmapBytes;
      const int op = in.op;
      // (fold_can_take admits LOAD first, then OR and AND steps only; the other operations' code is not instantiated here)
      if (op == I_LOAD) {
        if (!staged) {
          win4[tid] = full ? ones4 : zero4;
          win4[tid + kFoldThreads] = full ? ones4 : zero4;
        } else if (type == kBitmap) {
          win4[tid] = st4[tid];
          win4[tid + kFoldThreads] = st4[tid + kFoldThreads];
        } else if (type == kArray) {
          fold_array_set<E_STORE>(S_a, st_a, len, tid);
          win4[tid] = zero4; win4[tid + kFoldThreads] = zero4;
          wsel ^= 1u;
        } else {
          fold_runs(I_LOAD, s_win, s_S, s_E, st_a, len, s_red, tid);
        }
      } else if (nil) {
        if (op == I_AND_LEAF) { win4[tid] = zero4; win4[tid + kFoldThreads] = zero4; }
      } else if (full) {
        if (op == I_OR_LEAF) { win4[tid] = ones4; win4[tid + kFoldThreads] = ones4; }
        else if (op == I_ANDNOT_LEAF) { win4[tid] = zero4; win4[tid + kFoldThreads] = zero4; }
        else if (op == I_XOR_LEAF) { win4[tid] = not128(win4[tid]); win4[tid + kFoldThreads] = not128(win4[tid + kFoldThreads]); }
      } else if (type == kBitmap) {
        const uint4 g0 = st4[tid], g1 = st4[tid + kFoldThreads];
        const uint4 w0 = win4[tid], w1 = win4[tid + kFoldThreads];
        win4[tid] = op == I_OR_LEAF ? or128(w0, g0) : op == I_AND_LEAF ? and128(w0, g0) : op == I_XOR_LEAF ? xor128(w0, g0) : andnot128(w0, g0);
        win4[tid + kFoldThreads] = op == I_OR_LEAF ? or128(w1, g1) : op == I_AND_LEAF ? and128(w1, g1) : op == I_XOR_LEAF ? xor128(w1, g1) : andnot128(w1, g1);
      } else if (type == kArray) {
        if (op == I_OR_LEAF) fold_array_set<E_SET>(win_a, st_a, len, tid);
#if PB200_FOLD_XOR
        else if (op == I_XOR_LEAF) fold_array_set<E_FLIP>(win_a, st_a, len, tid);
        else if (op == I_ANDNOT_LEAF) fold_array_set<E_CLEAR>(win_a, st_a, len, tid);
#endif
        else {  // AND: the array as a bitmap in S, then word by word (intersectArrayBitmap roaring.go:3404)
          fold_array_set<E_STORE>(S_a, st_a, len, tid);
          fold_bar();
          const uint4 g0 = S4[tid], g1 = S4[tid + kFoldThreads];
          S4[tid] = zero4; S4[tid + kFoldThreads] = zero4;
          win4[tid] = and128(win4[tid], g0);
          win4[tid + kFoldThreads] = and128(win4[tid + kFoldThreads], g1);
        }
      } else {  // run list
        fold_runs(op, s_win, s_S, s_E, st_a, len, s_red, tid);
      }
      fold_bar();  // the operand is applied by every consumer
      if (staged) {
        if (tid == 0) {  // its staging slot goes back to the producer (the consumers' reads are ordered before the copy engine's next write)
          fence_proxy_async();
          mbar_arrive(&s_empty[slot]);
        }
        g++;
      }
    }

    const uint4* win4 = reinterpret_cast<const uint4*>(s_dyn + wsel * kBitmapBytes);
    const uint4 a = win4[tid], b = win4[tid + kFoldThreads];
    const uint32_t ca = popc128(a), cb = popc128(b);
    if (!p.dense_n) {  // no per-container cardinalities wanted (a count, or a temp row): every warp adds its part, one barrier less per item
      if (p.dense_out) {
        uint4* dout = reinterpret_cast<uint4*>(p.dense_out + m.item * kBitmapBytes);
        dout[tid] = a; dout[tid + kFoldThreads] = b;
      }
      const uint32_t cnt = warp_sum(ca + cb);
      if (p.out_counts) add_count(uint64_t(m.q) * p.out_qstride + uint64_t(m.s) * p.out_sstride, cnt);
      fold_bar();  // the window and the item's descriptors are no longer needed
      if (tid == 0) mbar_arrive(&s_mempty[ms]);
      continue;
    }
    // the row is wanted: thread t holds words 4t .. 4t+3 (a) and 1024 + 4t .. (b) of the container; inclusive scans of
    // both halves give the total and, for a short result, every thread's place in the value list
    uint32_t xa = ca, xb = cb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t ya = __shfl_up_sync(0xffffffffu, xa, o), yb = __shfl_up_sync(0xffffffffu, xb, o);
      if (lane >= o) { xa += ya; xb += yb; }
    }
    if (lane == 31) { s_red[tid >> 5] = xa; s_red2[tid >> 5] = xb; }
    fold_bar();
    uint32_t tot_a = 0, tot_b = 0, pre_a = 0, pre_b = 0;
#pragma unroll
    for (int w = 0; w < kFoldThreads / 32; w++) {
      const uint32_t ra = s_red[w], rb = s_red2[w];
      if (w < (tid >> 5)) { pre_a += ra; pre_b += rb; }
      tot_a += ra; tot_b += rb;
    }
    const uint32_t tot = tot_a + tot_b;
    const bool as_list = p.sparse_out && tot <= kSparseListMax;
    if (as_list) {
      uint16_t* lst = p.sparse_out + m.item * kSparseListMax;
      uint32_t ia = pre_a + xa - ca, ib = tot_a + pre_b + xb - cb;
      const uint32_t wa[4] = {a.x, a.y, a.z, a.w}, wb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        for (uint32_t v = wa[k]; v; v &= v - 1) lst[ia++] = uint16_t((4 * tid + k) * 32 + __ffs(v) - 1);
        for (uint32_t v = wb[k]; v; v &= v - 1) lst[ib++] = uint16_t((4 * (tid + kFoldThreads) + k) * 32 + __ffs(v) - 1);
      }
    } else if (p.dense_out) {
      uint4* dout = reinterpret_cast<uint4*>(p.dense_out + m.item * kBitmapBytes);
      dout[tid] = a; dout[tid + kFoldThreads] = b;
    }
    if (tid == 0) {
      if (p.out_counts && tot) atomicAdd(p.out_counts + uint64_t(m.q) * p.out_qstride + uint64_t(m.s) * p.out_sstride, (unsigned long long)tot);
      p.dense_n[m.item] = as_list ? (tot | kSparseFlag) : tot;
      mbar_arrive(&s_mempty[ms]);  // the item's descriptors are no longer needed
    }
    fold_bar();  // s_red and the window are free for the next item
  }
  if (lane == 0 && acc) atomicAdd(p.out_counts + acc_off, acc);
}

// ---------------------------------------------------------------------------------------
// K13 — Union of many plain rows into one dense temp row (the chunk sub-programs of a wide
// n-ary Union, and of the subtrahends of a wide Difference): executeUnionShard's left fold
// (executor.go:1693-1713) -> Row.Union -> roaring union (roaring.go:3420-3839) for any number
// of operands at once. OR needs no operand kept apart, so every operand of a column is
// accumulated into ONE 8 KiB shared-memory window — arrays and runs through the same
// expansion the staged kernel uses, bitmaps word by word — and the window is written once.
// One CTA per (shard, container column) item, persistent grid.
// ---------------------------------------------------------------------------------------
constexpr int kUnionThreads = 256;

template <bool XOR>
__global__ void __launch_bounds__(kUnionThreads) union_rows_kernel(EvalParams p) {
  __shared__ __align__(16) uint32_t s_win[kBitmapWords * 2];
  __shared__ ContainerDesc s_desc[kUnionThreads];  // a batch of operands: one descriptor lookup per thread
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const Prog pg = p.progs[0];
  const uint32_t n_items = p.n_shards * kRowContainers;
  unsigned long long bytes_acc = 0;
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t s = item / kRowContainers, c = item % kRowContainers;
    __syncthreads();  // the previous item's window is no longer read
    uint4* win4 = reinterpret_cast<uint4*>(s_win);
    for (int j = tid; j < kBitmapWords / 2; j += kUnionThreads) win4[j] = make_uint4(0, 0, 0, 0);
    for (uint32_t l0 = 0; l0 < pg.n_leaves; l0 += kUnionThreads) {  // any number of operands, a batch of descriptors at a time
      const uint32_t nl = min(uint32_t(kUnionThreads), pg.n_leaves - l0);
      if (tid < int(nl)) {
        const ContainerDesc d = resolve_leaf(p, p.leaves[pg.leaf_off + l0 + tid], s, c);
        s_desc[tid] = d;
        bytes_acc += desc_algo_bytes(d);
      }
      __syncthreads();
      // arrays (the bulk of a sparse field): a warp per operand, so that eight payload reads are in flight per CTA
      // instead of one; their expansion uses shared-memory atomicOr only, so operands may overlap freely
      for (uint32_t l = warp; l < nl; l += kUnionThreads / 32) {
        const ContainerDesc d = s_desc[l];
        if (d.n && desc_type(d) == kArray) stage_container<XOR>(d, s_win, 0, kBitmapWords, lane, 32);
      }
      __syncthreads();
      for (uint32_t l = 0; l < nl; l++) {  // bitmaps and runs: the whole CTA on one operand
        const ContainerDesc d = s_desc[l];
        if (d.n == 0 || desc_type(d) == kArray) continue;  // uniform over the CTA
        if (desc_type(d) == kBitmap) {  // every thread owns the same words for every bitmap operand
          const uint4* src = reinterpret_cast<const uint4*>(d.addr);
          for (int j = tid; j < kBitmapWords / 2; j += kUnionThreads) win4[j] = XOR ? xor128(win4[j], ldg128_stream(src + j)) : or128(win4[j], ldg128_stream(src + j));
        } else {
          stage_container<XOR>(d, s_win, 0, kBitmapWords, tid, kUnionThreads);
        }
        __syncthreads();  // plain stores of this operand before the atomics / read-modify-writes of the next one
      }
      __syncthreads();  // s_desc is rewritten by the next batch
    }
    __syncthreads();
    uint4* dst = reinterpret_cast<uint4*>(p.dense_out + size_t(item) * kBitmapBytes);
    for (int j = tid; j < kBitmapWords / 2; j += kUnionThreads) dst[j] = win4[j];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bytes_acc += __shfl_xor_sync(0xffffffffu, bytes_acc, o);
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

}  // namespace pb200
