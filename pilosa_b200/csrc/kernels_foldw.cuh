// K2-K6, fold form without a barrier among the consumers (count-only launches). NOT BUILT BY DEFAULT: parity-green, but
// measured slower than eval_fold_kernel on the mixed-container config (0.34 vs 0.26 ms: no barrier stall, 72 % issue
// utilisation, but twice the instructions — every warp pays the border search and the partly filled chunks of every
// operand). Kept as the measured alternative; engine.cu includes it only with -DPB200_FOLDW=1.
//
// eval_fold_kernel's eight consumer warps apply every operand together and meet at a named barrier after it (three
// times for a run list): ncu attributes 29 % of its stall samples to that barrier and the issue slots are busy 53 % of
// the time. Here each consumer warp owns one eighth of the window for good — positions [8192 w, 8192 (w + 1)), 256
// words — and of the scratch bitmaps S and E, and applies to it the part of every operand that falls into it:
//   bitmap     its 1 KiB of the staged words;
//   array      the values in its range, found by a 16-ary search of both range borders at once (half a warp per
//              border, three rounds for 4096 values), taken in aligned chunks of 8 per lane with the slice-ownership
//              rule of eval_fold_kernel (a word belongs to the chunk holding its first element): plain stores;
//   run list   the runs that reach into its range (same search on the runs' last / start fields), starts and ends as
//              bits of its parts of S and E, run bitmap = E - S over the part's 256 words: a lane owns 8 words, the
//              borrow-in of a lane is the parity of the bits in front of it (one ballot — no shared-memory hop, the
//              part is exactly one warp wide), the borrow-in of the part is "a run straddles my lower border".
// Nothing a warp writes is read by another warp, so the only synchronisation is __syncwarp and the mbarriers of the
// staging ring (a slot is released when all eight warps have arrived). Warps drift apart by up to the depth of the
// ring and of the descriptor ring; the count of an item is eight partial popcounts that go to the per-warp running sums.
// Launches that want the rows (dense_out / dense_n) keep eval_fold_kernel: the cardinality per container needs all
// eight parts at once.
#pragma once
#include "kernels.cuh"

namespace pb200 {

constexpr uint32_t kPartWords = kWords32 / (kFoldThreads / 32);  // 256 32-bit words = 8192 positions per consumer warp
constexpr uint32_t kPartBits = kPartWords * 32;

// Both borders of a range in one go: lanes 0-15 look for the first element that is not below key_lo, lanes 16-31 for
// key_hi; `below(i, key)` says whether element i (< len) is below the key. 16-ary: three rounds for 4096 elements.
template <class BELOW>
__device__ __forceinline__ void warp_two_bounds(const BELOW& below, uint32_t len, uint32_t key_lo, uint32_t key_hi, int lane,
                                                uint32_t* out_lo, uint32_t* out_hi) {
  const int half = lane >> 4, l = lane & 15;
  const uint32_t key = half ? key_hi : key_lo;
  uint32_t lo = 0, hi = len;  // the answer lies in [lo, hi]
  while (__any_sync(0xffffffffu, hi > lo)) {
    const uint32_t span = hi - lo, step = (span + 15) >> 4;
    const uint32_t idx = lo + uint32_t(l) * step;
    const bool pred = hi > lo && idx < hi && below(idx, key);
    const uint32_t votes = (__ballot_sync(0xffffffffu, pred) >> (half * 16)) & 0xffffu;
    const uint32_t c = __popc(votes);  // elements are sorted: the lanes that say "below" are the first c
    if (hi > lo) {
      const uint32_t nlo = c ? lo + (c - 1) * step + 1 : lo;
      const uint32_t nhi = min(hi, lo + c * step);
      lo = nlo; hi = nhi;
    }
  }
  *out_lo = __shfl_sync(0xffffffffu, lo, 0);
  *out_hi = __shfl_sync(0xffffffffu, lo, 16);
}

// One lane's chunk of a sorted sequence: element j of the chunk is v[j]; only j in [kb, ke) belong to the part.
// `prev_w` is the word of the element in front of v[kb] (kNoWord if there is none in this part); `next(i)` reads the
// elements behind the chunk (i counts from 0 at v[ke]; returns kNoPos when the part's elements are exhausted).
// Positions are relative to the part. MODE as in fold_put.
constexpr uint32_t kNoWord = 0xffffffffu, kNoPos = 0xffffffffu;
template <int N, int MODE, class NEXT>
__device__ __forceinline__ void foldw_chunk(uint32_t bm, const uint32_t (&v)[N], uint32_t kb, uint32_t ke, uint32_t prev_w, const NEXT& next) {
  uint32_t cur_w = prev_w, cur_b = 0;
  const uint32_t foreign = prev_w;
#pragma unroll
  for (int k = 0; k < N; k++) {
    if (uint32_t(k) >= kb && uint32_t(k) < ke) {
      const uint32_t w = v[k] >> 5;
      const uint32_t bit = __funnelshift_l(0u, 1u, v[k]);
      const bool chg = w != cur_w;
      if (chg && cur_w != foreign) fold_put<MODE>(bm, cur_w, cur_b);
      cur_b = chg ? bit : (cur_b | bit);
      cur_w = w;
    }
  }
  if (cur_w == foreign) return;  // nothing of this chunk starts a word
  for (uint32_t i = 0;; i++) {   // the rest of the last word: at most 31 elements
    const uint32_t x = next(i);
    if (x == kNoPos || (x >> 5) != cur_w) break;
    cur_b |= __funnelshift_l(0u, 1u, x);
  }
  if (cur_w < kPartWords) fold_put<MODE>(bm, cur_w, cur_b);  // (a run's end may lie behind the part)
}

__device__ __forceinline__ uint4 lds_v4(uint32_t a) {
  uint4 q;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(a) : "memory");
  return q;
}
__device__ __forceinline__ void sts_v4(uint32_t a, const uint4& q) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a), "r"(q.x), "r"(q.y), "r"(q.z), "r"(q.w) : "memory");
}

// The part of a sorted array container (staged at shared address st, len values) that falls into [base, base + 8192),
// set / flipped / cleared in the warp's part of the window, or stored into its part of the zero scratch S (bm).
template <int MODE>
__device__ __noinline__ void foldw_array(uint32_t bm, uint32_t st, uint32_t len, uint32_t base, int lane) {
  uint32_t lo, hi;
  warp_two_bounds([&](uint32_t i, uint32_t key) { return lds_u16(st + 2u * i) < key; }, len, base, base + kPartBits, lane, &lo, &hi);
  if (lo >= hi) return;
  for (uint32_t c = (lo >> 3) + uint32_t(lane); c * 8u < hi; c += 32u) {
    const uint32_t e0 = c * 8u;
    const uint4 q = lds_v4(st + 2u * e0);
    const uint32_t v[8] = {(q.x & 0xffffu) - base, (q.x >> 16) - base, (q.y & 0xffffu) - base, (q.y >> 16) - base,
                           (q.z & 0xffffu) - base, (q.z >> 16) - base, (q.w & 0xffffu) - base, (q.w >> 16) - base};
    const uint32_t kb = lo > e0 ? lo - e0 : 0u, ke = min(hi - e0, 8u);
    const uint32_t prev_w = (kb == 0 && e0 > lo) ? ((lds_u16(st + 2u * (e0 - 1)) - base) >> 5) : kNoWord;
    const uint32_t after = e0 + ke;
    foldw_chunk<8, MODE>(bm, v, kb, ke, prev_w, [&](uint32_t i) { return after + i < hi ? lds_u16(st + 2u * (after + i)) - base : kNoPos; });
  }
}

// r = e - s - borrow over 8 words (fold_sub8 of eval_fold_kernel), then the operation on the warp's part of the window.
// The run list is staged at shared address st (len runs: start | last << 16).
__device__ __noinline__ void foldw_runs(const int OP, uint32_t win, uint32_t S, uint32_t E, uint32_t st, uint32_t len, uint32_t base, int lane) {
  uint32_t lo, hi;
  // lower border: first run whose last value reaches the part; upper border: first run that starts behind the part
  warp_two_bounds([&](uint32_t i, uint32_t key) { return (lane < 16 ? lds_u16(st + 4u * i + 2u) : lds_u16(st + 4u * i)) < key; },
                  len, base, base + kPartBits, lane, &lo, &hi);
  uint32_t open = 0;  // a run straddles the lower border: ones from the part's first position on, no start bit
  if (lo < hi) open = lds_u16(st + 4u * lo) < base ? 1u : 0u;
  if (lo < hi) {
    const uint32_t lo_s = lo + open;
    for (uint32_t c = (lo >> 2) + uint32_t(lane); c * 4u < hi; c += 32u) {
      const uint32_t e0 = c * 4u;
      const uint4 q = lds_v4(st + 4u * e0);
      const uint32_t vs[4] = {(q.x & 0xffffu) - base, (q.y & 0xffffu) - base, (q.z & 0xffffu) - base, (q.w & 0xffffu) - base};
      const uint32_t ve[4] = {(q.x >> 16) + 1u - base, (q.y >> 16) + 1u - base, (q.z >> 16) + 1u - base, (q.w >> 16) + 1u - base};
      const uint32_t ke = min(hi - e0, 4u), after = e0 + ke;
      const uint32_t kbs = lo_s > e0 ? lo_s - e0 : 0u, kbe = lo > e0 ? lo - e0 : 0u;
      const uint32_t prev_s = (kbs == 0 && e0 > lo_s) ? ((lds_u16(st + 4u * (e0 - 1)) - base) >> 5) : kNoWord;
      const uint32_t prev_e = (kbe == 0 && e0 > lo) ? ((lds_u16(st + 4u * (e0 - 1) + 2u) + 1u - base) >> 5) : kNoWord;
      foldw_chunk<4, E_STORE>(S, vs, kbs, ke, prev_s, [&](uint32_t i) { return after + i < hi ? lds_u16(st + 4u * (after + i)) - base : kNoPos; });
      foldw_chunk<4, E_STORE>(E, ve, kbe, ke, prev_e, [&](uint32_t i) { return after + i < hi ? lds_u16(st + 4u * (after + i) + 2u) + 1u - base : kNoPos; });
    }
  }
  __syncwarp();
  // words 8 lane .. 8 lane + 7 of the part
  const uint32_t o = 32u * uint32_t(lane);
  const uint4 s0 = lds_v4(S + o), s1 = lds_v4(S + o + 16), e0 = lds_v4(E + o), e1 = lds_v4(E + o + 16);
  const uint4 zero4 = make_uint4(0, 0, 0, 0);
  sts_v4(S + o, zero4); sts_v4(S + o + 16, zero4); sts_v4(E + o, zero4); sts_v4(E + o + 16, zero4);
  const uint32_t s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  const uint32_t e[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
  uint32_t par = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) par ^= s[k] ^ e[k];
  const uint32_t votes = __ballot_sync(0xffffffffu, (__popc(par) & 1u) != 0);
  const uint32_t borrow = (__popc(votes & ((1u << lane) - 1u)) + open) & 1u;
  uint32_t r[8];
  fold_sub8(r, e, s, borrow);
  if (OP == I_LOAD) {
    sts_v4(win + o, make_uint4(r[0], r[1], r[2], r[3]));
    sts_v4(win + o + 16, make_uint4(r[4], r[5], r[6], r[7]));
    return;
  }
  const uint4 a0 = lds_v4(win + o), a1 = lds_v4(win + o + 16);
  const uint32_t a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
  uint32_t w[8];
#pragma unroll
  for (int k = 0; k < 8; k++)
    w[k] = OP == I_OR_LEAF ? (a[k] | r[k]) : OP == I_AND_LEAF ? (a[k] & r[k]) : OP == I_XOR_LEAF ? (a[k] ^ r[k]) : (a[k] & ~r[k]);
  sts_v4(win + o, make_uint4(w[0], w[1], w[2], w[3]));
  sts_v4(win + o + 16, make_uint4(w[4], w[5], w[6], w[7]));
}

__global__ void __launch_bounds__(kFoldCtaThreads, PB200_FOLD_CTAS) eval_foldw_kernel(EvalParams p) {
  extern __shared__ __align__(128) uint8_t s_dyn[];  // [window][S][E][kFoldStages x staging], 8 KiB each
  __shared__ FoldMeta s_meta[kFoldMeta];
  __shared__ __align__(8) uint64_t s_full[kFoldStages], s_empty[kFoldStages], s_mfull[kFoldMeta], s_mempty[kFoldMeta];
  constexpr uint32_t kWarps = kFoldThreads / 32;
  const int tid = threadIdx.x, lane = tid & 31;
  const uint32_t n_slow = p.fold_all ? p.n_queries * p.n_shards * kRowContainers : *p.slow_count;
  if (tid == 0) {
    for (int i = 0; i < kFoldStages; i++) { mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], kWarps); }  // a slot is free when every warp has let go
    for (int i = 0; i < kFoldMeta; i++) { mbar_init(&s_mfull[i], 1); mbar_init(&s_mempty[i], kWarps); }
    mbar_fence_init();
  }
  __syncthreads();
  if (tid >= kFoldThreads) {
    fold_producer(p, n_slow, lane, s_dyn, s_meta, s_full, s_empty, s_mfull, s_mempty);
    return;
  }

  // ---------------- consumer warps: warp w owns words [256 w, 256 w + 256) of the window, of S and of E ----------------
  const uint32_t warp = uint32_t(tid) >> 5, base = warp * kPartBits;
  const uint32_t dyn_a = smem_u32(s_dyn);
  const uint32_t part = warp * kPartWords * 4u;                 // byte offset of the part inside a region
  const uint32_t E_a = dyn_a + 2u * kBitmapBytes + part;
  const uint32_t lo16 = part + 16u * uint32_t(lane);            // this lane's two 16 B pieces of a part: lo16, lo16 + 512
  const uint4 zero4 = make_uint4(0, 0, 0, 0), ones4 = make_uint4(~0u, ~0u, ~0u, ~0u);
  uint32_t wsel = 0;  // which of regions 0 / 1 is this warp's window (the other is its S); see eval_fold_kernel
  for (uint32_t r = 0; r < 3; r++) { sts_v4(dyn_a + r * kBitmapBytes + lo16, zero4); sts_v4(dyn_a + r * kBitmapBytes + lo16 + 512u, zero4); }
  __syncwarp();
  uint32_t g = 0, j = 0;
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
    const uint32_t n_instr = m.n_instr, take = m.take;
    const uint64_t off = uint64_t(m.q) * p.out_qstride + uint64_t(m.s) * p.out_sstride;
    if (take != 1u) {  // 0: eval_staged_kernel's item; 2: Count(Row) straight from the descriptor (warp 0 adds it)
      if (warp == 0 && take == 2u && p.out_counts) add_count(off, m.desc[m.ins[0].leaf].n);
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_mempty[ms]);
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
      const uint32_t st_a = dyn_a + (slot + kFoldFixed) * kBitmapBytes;
      if (staged) mbar_wait(&s_full[slot], (g / kFoldStages) & 1u);  // this operand's payload has landed
      const uint32_t win_r = dyn_a + wsel * kBitmapBytes, S_r = dyn_a + (wsel ^ 1u) * kBitmapBytes;  // regions
      const uint32_t win_a = win_r + part, S_a = S_r + part;                                           // this warp's parts
      const int op = in.op;
      if (op == I_LOAD) {
        if (!staged) {
          sts_v4(win_r + lo16, full ? ones4 : zero4); sts_v4(win_r + lo16 + 512u, full ? ones4 : zero4);
        } else if (type == kBitmap) {
          sts_v4(win_r + lo16, lds_v4(st_a + lo16)); sts_v4(win_r + lo16 + 512u, lds_v4(st_a + lo16 + 512u));
        } else if (type == kArray) {
          foldw_array<E_STORE>(S_a, st_a, len, base, lane);
          sts_v4(win_r + lo16, zero4); sts_v4(win_r + lo16 + 512u, zero4);
          wsel ^= 1u;
        } else {
          foldw_runs(I_LOAD, win_a, S_a, E_a, st_a, len, base, lane);
        }
      } else if (nil) {
        if (op == I_AND_LEAF) { sts_v4(win_r + lo16, zero4); sts_v4(win_r + lo16 + 512u, zero4); }
      } else if (full) {
        if (op == I_OR_LEAF) { sts_v4(win_r + lo16, ones4); sts_v4(win_r + lo16 + 512u, ones4); }
        else if (op == I_ANDNOT_LEAF) { sts_v4(win_r + lo16, zero4); sts_v4(win_r + lo16 + 512u, zero4); }
        else if (op == I_XOR_LEAF) { sts_v4(win_r + lo16, not128(lds_v4(win_r + lo16))); sts_v4(win_r + lo16 + 512u, not128(lds_v4(win_r + lo16 + 512u))); }
      } else if (type == kBitmap) {
        const uint4 g0 = lds_v4(st_a + lo16), g1 = lds_v4(st_a + lo16 + 512u);
        const uint4 w0 = lds_v4(win_r + lo16), w1 = lds_v4(win_r + lo16 + 512u);
        sts_v4(win_r + lo16, op == I_OR_LEAF ? or128(w0, g0) : op == I_AND_LEAF ? and128(w0, g0) : op == I_XOR_LEAF ? xor128(w0, g0) : andnot128(w0, g0));
        sts_v4(win_r + lo16 + 512u, op == I_OR_LEAF ? or128(w1, g1) : op == I_AND_LEAF ? and128(w1, g1) : op == I_XOR_LEAF ? xor128(w1, g1) : andnot128(w1, g1));
      } else if (type == kArray) {
        if (op == I_OR_LEAF) foldw_array<E_SET>(win_a, st_a, len, base, lane);
        else if (op == I_XOR_LEAF) foldw_array<E_FLIP>(win_a, st_a, len, base, lane);
        else if (op == I_ANDNOT_LEAF) foldw_array<E_CLEAR>(win_a, st_a, len, base, lane);
        else {  // AND: the array's bits in the part of S, then word by word
          foldw_array<E_STORE>(S_a, st_a, len, base, lane);
          __syncwarp();
          const uint4 g0 = lds_v4(S_r + lo16), g1 = lds_v4(S_r + lo16 + 512u);
          sts_v4(S_r + lo16, zero4); sts_v4(S_r + lo16 + 512u, zero4);
          sts_v4(win_r + lo16, and128(lds_v4(win_r + lo16), g0));
          sts_v4(win_r + lo16 + 512u, and128(lds_v4(win_r + lo16 + 512u), g1));
        }
      } else {
        foldw_runs(op, win_a, S_a, E_a, st_a, len, base, lane);
      }
      __syncwarp();  // the operand is applied to this warp's part (and read out of the staging slot) by all of its lanes
      if (staged) {
        if (lane == 0) mbar_arrive(&s_empty[slot]);
        g++;
      }
    }
    const uint32_t win_r = dyn_a + wsel * kBitmapBytes;
    uint32_t cnt = popc128(lds_v4(win_r + lo16)) + popc128(lds_v4(win_r + lo16 + 512u));
    cnt = warp_sum(cnt);
    if (p.out_counts) add_count(off, cnt);
    __syncwarp();
    if (lane == 0) mbar_arrive(&s_mempty[ms]);
  }
  if (lane == 0 && acc) atomicAdd(p.out_counts + acc_off, acc);
}

}  // namespace pb200
