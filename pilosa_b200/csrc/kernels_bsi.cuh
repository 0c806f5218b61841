// K8/K9/K10 — bit-sliced index kernels. BSI fragments store row 0 = exists (E), row 1 = sign
// (S), row 2+i = magnitude bit i (fragment.go:90-93, sign-magnitude). Every step of
// fragment.sum / min / max / rangeOp (fragment.go:1109-1534) is a word-parallel boolean
// function of (E, S, B_i, filter/keep state), so one pass over the planes with the state
// held in registers replaces ~3 x bitDepth materialised Row operations per shard.
#pragma once
#include "kernels.cuh"
#include "tma.cuh"

namespace pb200 {

// ---------------------------------------------------------------------------------------
// Plane ring: the container columns a BSI kernel consumes one after the other (E, S, filter,
// B_i ...) are streamed through a ring of kRingDepth 8 KiB shared-memory slots by the bulk
// copy engine. One thread issues cp.async.bulk for plane q + kRingDepth as soon as plane q
// has been consumed, so kRingDepth x 8 KiB per CTA are always in flight however many
// registers the comparator state needs. Bitmap containers take the TMA path; array / run
// containers are staged into their slot by the CTA; empty / full containers are constants.
// ---------------------------------------------------------------------------------------
constexpr int kRingDepth = 4;
constexpr int kRingBytes = kRingDepth * kBitmapBytes;

struct PlaneRing {
  uint8_t* ring;     // kRingDepth x 8 KiB, 128 B aligned
  uint64_t* full;    // kRingDepth mbarriers (count 1 + transaction bytes)
  uint32_t phase;    // expected parity per slot (bit s), tracked identically by every thread

  __device__ __forceinline__ static bool by_tma(const ContainerDesc& d) { return d.n != 0 && desc_type(d) == kBitmap; }

  // thread 0 only: start the copy of plane sequence number q
  __device__ __forceinline__ void issue(const ContainerDesc& d, uint32_t q) {
    if (!by_tma(d)) return;
    const uint32_t slot = q % kRingDepth;
    fence_proxy_async();  // earlier generic-proxy traffic on this slot is ordered before the async write
    mbar_arrive_expect_tx(&full[slot], kBitmapBytes);
    tma_load_1d(ring + slot * kBitmapBytes, reinterpret_cast<const void*>(d.addr), kBitmapBytes, &full[slot]);
  }

  // whole CTA: this thread's PER chunks of plane q. The caller must __syncthreads() before
  // thread 0 re-issues into the same slot.
  template <int NT, int PER>
  __device__ __forceinline__ void take(const ContainerDesc& d, uint32_t q, int tid, uint4 (&v)[PER]) {
    const uint32_t slot = q % kRingDepth;
    uint4* win = reinterpret_cast<uint4*>(ring + slot * kBitmapBytes);
    if (d.n == 0) {
#pragma unroll
      for (int k = 0; k < PER; k++) v[k] = make_uint4(0, 0, 0, 0);
    } else if (desc_type(d) == kBitmap) {
      mbar_wait(&full[slot], (phase >> slot) & 1u);
      phase ^= 1u << slot;
#pragma unroll
      for (int k = 0; k < PER; k++) v[k] = win[tid + k * NT];
    } else if (d.n == kFullN) {
#pragma unroll
      for (int k = 0; k < PER; k++) v[k] = make_uint4(~0u, ~0u, ~0u, ~0u);
    } else {
      for (int j = tid; j < kBitmapWords / 2; j += NT) win[j] = make_uint4(0, 0, 0, 0);
      __syncthreads();
      stage_container(d, reinterpret_cast<uint32_t*>(win), 0, kBitmapWords, tid, NT);
      __syncthreads();
#pragma unroll
      for (int k = 0; k < PER; k++) v[k] = win[tid + k * NT];
    }
  }
};

constexpr int kBsiThreads = 256;  // one CTA per (shard, container column): 2 x uint4 per thread
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

// Block-cooperative load of a whole container column as words: thread t receives uint4 #t
// and #(t + nthreads) ... of the 512 uint4. Array/run containers are staged through the
// shared window `win` (8 KiB). All threads of the CTA must call with the same descriptor.
template <int NT, int PER>
__device__ __forceinline__ void load_column(const ContainerDesc& d, uint64_t* win, int tid, uint4 (&v)[PER]) {
  static_assert(NT * PER == kBitmapWords / 2, "column = 512 uint4");
  if (d.n == 0) {
#pragma unroll
    for (int k = 0; k < PER; k++) v[k] = make_uint4(0, 0, 0, 0);
  } else if (desc_type(d) == kBitmap) {
    const uint4* ptr = reinterpret_cast<const uint4*>(d.addr) + tid;
#pragma unroll
    for (int k = 0; k < PER; k++) v[k] = ldg128_stream(ptr + k * NT);
  } else if (d.n == kFullN) {
#pragma unroll
    for (int k = 0; k < PER; k++) v[k] = make_uint4(~0u, ~0u, ~0u, ~0u);
  } else {
    __syncthreads();  // window free
    for (int j = tid; j < kBitmapWords / 2; j += NT) reinterpret_cast<uint4*>(win)[j] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    stage_container(d, reinterpret_cast<uint32_t*>(win), 0, kBitmapWords, tid, NT);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PER; k++) v[k] = reinterpret_cast<const uint4*>(win)[tid + k * NT];
  }
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

template <int PER>
struct Mach {
  int kind = M_NONE;
  unsigned long long p1 = 0, p2 = 0;
  bool allow_eq = false, leading = true, done = false;
  uint4 f[PER], k1[PER], k2[PER];

  __device__ __forceinline__ void init(int kd, const uint4 (&filter)[PER], unsigned long long a, unsigned long long b, bool eq) {
    kind = kd; p1 = a; p2 = b; allow_eq = eq; leading = true; done = false;
#pragma unroll
    for (int k = 0; k < PER; k++) { f[k] = filter[k]; k1[k] = make_uint4(0, 0, 0, 0); k2[k] = make_uint4(0, 0, 0, 0); }
  }
  // one plane, MSB -> LSB; on completion the result is in f[]
  __device__ __forceinline__ void step(int i, const uint4 (&row)[PER]) {
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

__global__ void __launch_bounds__(kBsiThreads) bsi_range_kernel(BsiParams p) {
  constexpr int PER = 2;
  extern __shared__ __align__(128) uint8_t s_ring[];
  __shared__ __align__(8) uint64_t s_full[kRingDepth];
  __shared__ ContainerDesc s_pd[kMaxDepth + 2];  // planes in consumption order: E, S, B_{d-1} .. B_0
  const int tid = threadIdx.x;
  const uint32_t n_items = p.n_shards * kRowContainers;
  unsigned long long bytes_acc = 0;
  PlaneRing pr;
  pr.ring = s_ring; pr.full = s_full; pr.phase = 0;
  if (tid == 0) {
    for (int i = 0; i < kRingDepth; i++) mbar_init(&s_full[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  uint32_t q0 = 0;  // planes consumed so far by this CTA (ring sequence numbers run across items)
  const uint32_t n_planes = p.bsi_op == PB200_BSI_NOTNULL ? 1u : p.bit_depth + 2u;

  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t s = item >> 4, c = item & 15;
    uint4* out = p.dense_out ? reinterpret_cast<uint4*>(p.dense_out + uint64_t(item) * kBitmapBytes) : nullptr;
    uint32_t cnt = 0;
    const FamilyDev fam = p.families[p.family];
    const uint64_t shard = p.shards[s];
    FragDev f{};
    if (shard < fam.cap) f = fam.frags[shard];
    if (tid < int(n_planes)) {
      const uint64_t row = tid < 2 ? uint64_t(tid) : uint64_t(2 + (p.bit_depth - 1 - (tid - 2)));
      const ContainerDesc d = bsi_row_desc(f, row, c);
      s_pd[tid] = d;
      bytes_acc += desc_algo_bytes(d);
    }
    __syncthreads();
    if (tid == 0)
      for (uint32_t j = 0; j < n_planes && j < uint32_t(kRingDepth); j++) pr.issue(s_pd[j], q0 + j);

    uint4 E[PER], S[PER], res[PER];
    pr.take<kBsiThreads, PER>(s_pd[0], q0 + 0, tid, E);
    __syncthreads();
    if (tid == 0 && uint32_t(kRingDepth) < n_planes) pr.issue(s_pd[kRingDepth], q0 + kRingDepth);
    if (p.bsi_op == PB200_BSI_NOTNULL) {
#pragma unroll
      for (int k = 0; k < PER; k++) { if (out) out[tid + k * kBsiThreads] = E[k]; cnt += popc128(E[k]); }
      if (p.count_out) {
        cnt = warp_sum(cnt);
        if ((tid & 31) == 0 && cnt) atomicAdd(p.count_out + uint64_t(s) * p.count_sstride, (unsigned long long)cnt);
      }
      q0 += n_planes;
      continue;
    }
    pr.take<kBsiThreads, PER>(s_pd[1], q0 + 1, tid, S);
    __syncthreads();
    if (tid == 0 && 1u + kRingDepth < n_planes) pr.issue(s_pd[1 + kRingDepth], q0 + 1 + kRingDepth);
    uint4 pos[PER], neg[PER];
#pragma unroll
    for (int k = 0; k < PER; k++) { pos[k] = andnot128(E[k], S[k]); neg[k] = and128(E[k], S[k]); }

    Mach<PER> m0, m1;
    int combine = 0;  // 0: m0 ; 1: E \ m0 ; 2: S | m0 ; 3: pos | m0 ; 4: m0 | m1
    const long long a = p.a, b = p.b;
    const unsigned long long ua = uabs64(a), ub = uabs64(b);
    switch (p.bsi_op) {
      case PB200_BSI_EQ: m0.init(M_EQ, a < 0 ? neg : pos, ua, 0, false); break;
      case PB200_BSI_NEQ: m0.init(M_EQ, a < 0 ? neg : pos, ua, 0, false); combine = 1; break;
      case PB200_BSI_LT: case PB200_BSI_LTE: {
        const bool eq = p.bsi_op == PB200_BSI_LTE;
        if ((a >= 0 && eq) || (a >= -1 && !eq)) { m0.init(M_LTU, pos, ua, 0, eq); combine = 2; }
        else m0.init(M_GTU, neg, ua, 0, eq);
        break;
      }
      case PB200_BSI_GT: case PB200_BSI_GTE: {
        const bool eq = p.bsi_op == PB200_BSI_GTE;
        if ((a >= 0 && eq) || (a >= -1 && !eq)) m0.init(M_GTU, pos, ua, 0, eq);
        else { m0.init(M_LTU, neg, ua, 0, eq); combine = 3; }
        break;
      }
      default:  // BETWEEN
        if (a >= 0) m0.init(M_BETWEENU, pos, ua, ub, true);
        else if (b < 0) m0.init(M_BETWEENU, neg, ub, ua, true);
        else { m0.init(M_LTU, pos, ub, 0, true); m1.init(M_LTU, neg, ua, 0, true); combine = 4; }
        break;
    }
    // magnitude planes MSB -> LSB through the ring
    for (int i = int(p.bit_depth) - 1; i >= 0; i--) {
      const uint32_t j = 2u + uint32_t(int(p.bit_depth) - 1 - i);
      uint4 row[PER];
      pr.take<kBsiThreads, PER>(s_pd[j], q0 + j, tid, row);
      __syncthreads();  // every thread has read slot (q0 + j) % depth: it may be refilled
      if (tid == 0 && j + kRingDepth < n_planes) pr.issue(s_pd[j + kRingDepth], q0 + j + kRingDepth);
      m0.step(i, row);
      m1.step(i, row);
    }
    q0 += n_planes;
#pragma unroll
    for (int k = 0; k < PER; k++) {
      switch (combine) {
        case 0: res[k] = m0.f[k]; break;
        case 1: res[k] = andnot128(E[k], m0.f[k]); break;
        case 2: res[k] = or128(S[k], m0.f[k]); break;
        case 3: res[k] = or128(pos[k], m0.f[k]); break;
        default: res[k] = or128(m0.f[k], m1.f[k]); break;
      }
      if (out) out[tid + k * kBsiThreads] = res[k];
      cnt += popc128(res[k]);
    }
    if (p.count_out) {
      cnt = warp_sum(cnt);
      if ((tid & 31) == 0 && cnt) atomicAdd(p.count_out + uint64_t(s) * p.count_sstride, (unsigned long long)cnt);
    }
    __syncthreads();  // s_pd is rewritten by the next item
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bytes_acc += __shfl_xor_sync(0xffffffffu, bytes_acc, o);
  if ((tid & 31) == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

// ---------------------------------------------------------------------------------------
// K8: Sum. consider = E [& filter]; count = |consider|; nrow = S (unfiltered: the reference
// quirk, fragment.go:1118,1134; `strict` masks it with consider); prow = consider \ nrow;
// per plane popcount(B_i & prow), popcount(B_i & nrow). Persistent CTAs accumulate in
// shared memory and flush 2*depth+1 atomics once.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBsiThreads) bsi_sum_kernel(BsiParams p) {
  constexpr int PER = 2;
  extern __shared__ __align__(128) uint8_t s_ring[];
  __shared__ __align__(8) uint64_t s_full[kRingDepth];
  __shared__ ContainerDesc s_pd[kMaxDepth + 3];  // planes in consumption order: E, [filter], S, B_0 .. B_{d-1}
  __shared__ unsigned long long s_acc[2 * kMaxDepth + 1];
  const int tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 2 * kMaxDepth + 1; i += kBsiThreads) s_acc[i] = 0;
  PlaneRing pr;
  pr.ring = s_ring; pr.full = s_full; pr.phase = 0;
  if (tid == 0) {
    for (int i = 0; i < kRingDepth; i++) mbar_init(&s_full[i], 1);
    mbar_fence_init();
  }
  __syncthreads();
  const uint32_t n_items = p.n_shards * kRowContainers;
  const uint32_t nf = p.has_filter ? 1u : 0u;
  const uint32_t n_planes = p.bit_depth + 2u + nf;
  uint32_t q0 = 0;
  unsigned long long bytes_acc = 0;
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t s = item >> 4, c = item & 15;
    const FamilyDev fam = p.families[p.family];
    const uint64_t shard = p.shards[s];
    FragDev f{};
    if (shard < fam.cap) f = fam.frags[shard];
    if (f.nrows == 0) continue;  // no fragment: ValCount{} (executor.go:709-711)
    __syncthreads();             // s_pd of the previous item fully consumed
    if (tid < int(n_planes)) {
      ContainerDesc d;
      if (tid == 0) d = bsi_row_desc(f, 0, c);
      else if (nf && tid == 1) d = filter_desc(p, s, c);
      else if (tid == int(1 + nf)) d = bsi_row_desc(f, 1, c);
      else d = bsi_row_desc(f, 2 + (tid - 2 - nf), c);
      s_pd[tid] = d;
      bytes_acc += desc_algo_bytes(d);
    }
    __syncthreads();
    if (tid == 0)
      for (uint32_t j = 0; j < n_planes && j < uint32_t(kRingDepth); j++) pr.issue(s_pd[j], q0 + j);
    uint32_t j = 0;
    auto next_plane = [&](uint4 (&v)[PER]) {
      pr.take<kBsiThreads, PER>(s_pd[j], q0 + j, tid, v);
      __syncthreads();  // slot consumed by every thread: refill it
      if (tid == 0 && j + kRingDepth < n_planes) pr.issue(s_pd[j + kRingDepth], q0 + j + kRingDepth);
      j++;
    };
    uint4 cons[PER], nrow[PER], prow[PER];
    next_plane(cons);
    if (nf) {
      uint4 fl[PER];
      next_plane(fl);
#pragma unroll
      for (int k = 0; k < PER; k++) cons[k] = and128(cons[k], fl[k]);
    }
    next_plane(nrow);
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < PER; k++) {
      if (p.strict) nrow[k] = and128(nrow[k], cons[k]);
      prow[k] = andnot128(cons[k], nrow[k]);
      cnt += popc128(cons[k]);
    }
    cnt = warp_sum(cnt);
    if (lane == 0 && cnt) atomicAdd(&s_acc[0], (unsigned long long)cnt);
    for (uint32_t i = 0; i < p.bit_depth; i++) {
      uint4 row[PER];
      next_plane(row);
      uint32_t pc = 0, nc = 0;
#pragma unroll
      for (int k = 0; k < PER; k++) { pc += popc128(and128(row[k], prow[k])); nc += popc128(and128(row[k], nrow[k])); }
      pc = warp_sum(pc); nc = warp_sum(nc);
      if (lane == 0) {
        if (pc) atomicAdd(&s_acc[1 + 2 * i], (unsigned long long)pc);
        if (nc) atomicAdd(&s_acc[2 + 2 * i], (unsigned long long)nc);
      }
    }
    q0 += n_planes;
  }
  __syncthreads();
  for (uint32_t i = tid; i < 2 * p.bit_depth + 1; i += kBsiThreads)
    if (s_acc[i]) atomicAdd(p.sum_acc + i, s_acc[i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) bytes_acc += __shfl_xor_sync(0xffffffffu, bytes_acc, o);
  if (lane == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
}

// ---------------------------------------------------------------------------------------
// K9: Min / Max. One CTA per shard (1024 threads x 8 uint4 = the whole 128 KiB row state in
// registers); per plane one block-wide popcount decides whether the candidate set shrinks
// (fragment.go:1171-1186 minUnsigned, :1213-1225 maxUnsigned, :1145-1168 min, :1189-1210 max).
// ---------------------------------------------------------------------------------------
constexpr int kMmThreads = 1024;

__device__ __forceinline__ unsigned long long block_sum_1024(uint32_t v, unsigned long long* s_red, int tid) {
  v = warp_sum(v);
  __syncthreads();
  if ((tid & 31) == 0) s_red[tid >> 5] = v;
  __syncthreads();
  unsigned long long tot = 0;
#pragma unroll
  for (int w = 0; w < kMmThreads / 32; w++) tot += s_red[w];
  return tot;
}

// Column c of the row is handled by threads [64c, 64c+64): 8 uint4 each.
__device__ __forceinline__ void load_row_1024(const FragDev& f, uint64_t row, uint64_t* win16, int tid, uint4 (&v)[8],
                                              unsigned long long& bytes_acc) {
  const int c = tid >> 6, t = tid & 63;
  const ContainerDesc d = bsi_row_desc(f, row, c);
  if (t == 0) bytes_acc += desc_algo_bytes(d);
  // staging (rare for BSI planes): every column has its own 8 KiB window
  const bool need_stage = d.n != 0 && d.n != kFullN && desc_type(d) != kBitmap;
  const int any_stage = __syncthreads_or(need_stage);
  if (any_stage) {
    uint64_t* win = win16 + size_t(c) * kBitmapWords;
    for (int j = t; j < kBitmapWords / 2; j += 64) reinterpret_cast<uint4*>(win)[j] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    if (need_stage) stage_container(d, reinterpret_cast<uint32_t*>(win), 0, kBitmapWords, t, 64);
    __syncthreads();
  }
  if (d.n == 0) {
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = make_uint4(0, 0, 0, 0);
  } else if (desc_type(d) == kBitmap) {
    const uint4* ptr = reinterpret_cast<const uint4*>(d.addr) + t;
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = ldg128_stream(ptr + k * 64);
  } else if (d.n == kFullN) {
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = make_uint4(~0u, ~0u, ~0u, ~0u);
  } else {
    const uint4* win = reinterpret_cast<const uint4*>(win16 + size_t(c) * kBitmapWords);
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = win[t + k * 64];
  }
  if (any_stage) __syncthreads();
}

__global__ void __launch_bounds__(kMmThreads) bsi_minmax_kernel(BsiParams p) {
  extern __shared__ __align__(16) uint64_t s_win16[];  // 16 x 8 KiB staging windows
  __shared__ unsigned long long s_red[kMmThreads / 32];
  const int tid = threadIdx.x;
  unsigned long long bytes_acc = 0;
  for (uint32_t s = blockIdx.x; s < p.n_shards; s += gridDim.x) {
    const FamilyDev fam = p.families[p.family];
    const uint64_t shard = p.shards[s];
    FragDev f{};
    if (shard < fam.cap) f = fam.frags[shard];
    long long val = 0;
    unsigned long long count = 0;
    if (f.nrows != 0) {
      uint4 cons[8], tmp[8];
      load_row_1024(f, 0, s_win16, tid, cons, bytes_acc);
      if (p.has_filter) {
        // the filter row: resolve per column like any leaf
        const int c = tid >> 6, t = tid & 63;
        const ContainerDesc dF = filter_desc(p, s, c);
        if (t == 0) bytes_acc += desc_algo_bytes(dF);
        const bool need_stage = dF.n != 0 && dF.n != kFullN && desc_type(dF) != kBitmap;
        const int any_stage = __syncthreads_or(need_stage);
        uint64_t* win = s_win16 + size_t(c) * kBitmapWords;
        if (any_stage) {
          for (int j = t; j < kBitmapWords / 2; j += 64) reinterpret_cast<uint4*>(win)[j] = make_uint4(0, 0, 0, 0);
          __syncthreads();
          if (need_stage) stage_container(dF, reinterpret_cast<uint32_t*>(win), 0, kBitmapWords, t, 64);
          __syncthreads();
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
          uint4 fv;
          if (dF.n == 0) fv = make_uint4(0, 0, 0, 0);
          else if (desc_type(dF) == kBitmap) fv = ldg128_stream(reinterpret_cast<const uint4*>(dF.addr) + t + k * 64);
          else if (dF.n == kFullN) fv = make_uint4(~0u, ~0u, ~0u, ~0u);
          else fv = reinterpret_cast<const uint4*>(win)[t + k * 64];
          cons[k] = and128(cons[k], fv);
        }
        if (any_stage) __syncthreads();
      }
      uint32_t c0 = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) c0 += popc128(cons[k]);
      const unsigned long long n_cons = block_sum_1024(c0, s_red, tid);
      if (n_cons != 0) {
        load_row_1024(f, 1, s_win16, tid, tmp, bytes_acc);  // sign row
        // which unsigned search, over which candidate set?
        uint32_t cs = 0;
        if (!p.is_max) {
#pragma unroll
          for (int k = 0; k < 8; k++) cs += popc128(and128(tmp[k], cons[k]));  // negatives present?
        } else {
#pragma unroll
          for (int k = 0; k < 8; k++) cs += popc128(andnot128(cons[k], tmp[k]));  // positives present?
        }
        const unsigned long long n_side = block_sum_1024(cs, s_red, tid);
        bool use_max_unsigned, negate;
        uint4 filt[8];
        if (!p.is_max) {
          if (n_side) { use_max_unsigned = true; negate = true;
#pragma unroll
            for (int k = 0; k < 8; k++) filt[k] = and128(tmp[k], cons[k]);
          } else { use_max_unsigned = false; negate = false;
#pragma unroll
            for (int k = 0; k < 8; k++) filt[k] = cons[k];
          }
        } else {
          if (!n_side) { use_max_unsigned = false; negate = true;
#pragma unroll
            for (int k = 0; k < 8; k++) filt[k] = cons[k];
          } else { use_max_unsigned = true; negate = false;
#pragma unroll
            for (int k = 0; k < 8; k++) filt[k] = andnot128(cons[k], tmp[k]);
          }
        }
        unsigned long long mag = 0;
        for (int i = int(p.bit_depth) - 1; i >= 0; i--) {
          load_row_1024(f, 2 + i, s_win16, tid, tmp, bytes_acc);
          uint32_t pc = 0;
#pragma unroll
          for (int k = 0; k < 8; k++) {
            tmp[k] = use_max_unsigned ? and128(tmp[k], filt[k]) : andnot128(filt[k], tmp[k]);
            pc += popc128(tmp[k]);
          }
          count = block_sum_1024(pc, s_red, tid);
          if (count > 0) {
            if (use_max_unsigned) mag += 1ull << i;
#pragma unroll
            for (int k = 0; k < 8; k++) filt[k] = tmp[k];
          } else {
            if (!use_max_unsigned) mag += 1ull << i;
            if (i == 0) {
              uint32_t fc = 0;
#pragma unroll
              for (int k = 0; k < 8; k++) fc += popc128(filt[k]);
              count = block_sum_1024(fc, s_red, tid);
            }
          }
        }
        val = negate ? -(long long)mag : (long long)mag;
      }
    }
    if (tid == 0) { p.mm_val[s] = val; p.mm_count[s] = count; }
  }
  if ((tid & 63) == 0 && bytes_acc && p.algo_bytes) atomicAdd(p.algo_bytes, bytes_acc);
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
