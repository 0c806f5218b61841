// K11: bitmap-returning calls. The dense result rows of eval_dense
// ([shard][16 containers][1024 words] + cardinalities) are turned ON THE DEVICE into what the
// reference hands back per shard — a roaring bitmap whose containers carry the canonical type
// of Container.optimize (roaring.go:2289-2338) — already laid out as the Pilosa roaring byte
// stream (roaring.go:1044-1122: cookie, keyN, {key u64, type u16, n-1 u16}*, offset u32*,
// payloads; run payload = u16 count + {start,last}*), so that the host copies compact bytes
// instead of 128 KiB of words per shard and never re-scans them.
//
//   rows_classify_kernel  warp per container: count runs (bitmapCountRuns roaring.go:2249), pick type
//   rows_layout_kernel    one CTA: stream sizes, exclusive scan over shards, headers, offsets
//   rows_emit_kernel      CTA per container: array / run / bitmap payload
#pragma once
#include "layout.h"

namespace pb200 {

constexpr uint32_t kRowsEmitThreads = 256;

__device__ __forceinline__ uint32_t rows_payload_bytes(uint32_t meta) {
  const uint32_t type = meta & 3u, len = meta >> 2;
  return type == kArray ? len * 2 : type == kRun ? 2 + len * 4 : type == kBitmap ? kBitmapBytes : 0;
}

// Container.optimize's choice from (n, runs): run if runs <= 2048 && runs <= n/2; else array if n < 4096;
// else bitmap; n == 0 -> nil. Returned as type | len << 2.
__device__ __forceinline__ uint32_t rows_pick_type(uint32_t n, uint32_t runs) {
  if (n == 0) return kNil;
  if (runs <= 2048 && runs <= n / 2) return kRun | (runs << 2);
  if (n < 4096) return kArray | (n << 2);
  return kBitmap | (uint32_t(kBitmapWords) << 2);
}

__global__ void __launch_bounds__(256) rows_classify_kernel(const uint64_t* __restrict__ dense, const uint32_t* __restrict__ dense_n,
                                                            uint32_t n_items, uint32_t* __restrict__ meta) {
  const uint32_t lane = threadIdx.x & 31, warps = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; item < n_items; item += warps) {
    const uint32_t n = dense_n[item];
    if (n == 0 || n == 65536) {  // nil / full (one run)
      if (lane == 0) meta[item] = rows_pick_type(n, n ? 1 : 0);
      continue;
    }
    const uint4* src = reinterpret_cast<const uint4*>(dense + uint64_t(item) * kBitmapWords);
    uint32_t starts = 0, carry = 0;  // carry: top bit of the word before this iteration's first word
    #pragma unroll 4
    for (int it = 0; it < 16; it++) {
      const uint4 v = src[it * 32 + lane];
      const uint64_t w0 = uint64_t(v.x) | (uint64_t(v.y) << 32), w1 = uint64_t(v.z) | (uint64_t(v.w) << 32);
      const uint32_t top1 = uint32_t(w1 >> 63);
      uint32_t prev = __shfl_up_sync(0xffffffffu, top1, 1);
      if (lane == 0) prev = carry;
      starts += __popcll(w0 & ~((w0 << 1) | prev)) + __popcll(w1 & ~((w1 << 1) | (w0 >> 63)));
      carry = __shfl_sync(0xffffffffu, top1, 31);
    }
    #pragma unroll
    for (int o = 16; o; o >>= 1) starts += __shfl_xor_sync(0xffffffffu, starts, o);
    if (lane == 0) meta[item] = rows_pick_type(n, starts);
  }
}

__device__ __forceinline__ void rows_put16(uint8_t* p, uint32_t v) { *reinterpret_cast<uint16_t*>(p) = uint16_t(v); }
__device__ __forceinline__ void rows_put32(uint8_t* p, uint32_t v) { rows_put16(p, v); rows_put16(p + 2, v >> 16); }
__device__ __forceinline__ void rows_put64(uint8_t* p, uint64_t v) { rows_put32(p, uint32_t(v)); rows_put32(p + 4, uint32_t(v >> 32)); }

// One CTA of 1024 threads, a thread per shard (chunks of 1024 with a running total): stream
// size = 8 + 16 * keyN + payloads; streams are packed back to back (all sizes are even, so
// every field stays 2-byte aligned). `ovf` (optional) is the Shift overflow word per shard:
// a 17th container under key (shard + 1) * 16, small enough to be written right here.
__global__ void __launch_bounds__(1024) rows_layout_kernel(const uint32_t* __restrict__ meta, const uint32_t* __restrict__ dense_n,
                                                           const uint64_t* __restrict__ shards, const unsigned long long* __restrict__ ovf,
                                                           uint32_t n_shards, uint8_t* __restrict__ out, unsigned long long* __restrict__ shard_off,
                                                           unsigned long long* __restrict__ item_off) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long running;
  const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n_shards; base += 1024) {
    const uint32_t s = base + threadIdx.x;
    uint32_t keyn = 0, pay = 0, ometa = 0;
    unsigned long long ow = 0;
    if (s < n_shards) {
      for (int c = 0; c < kRowContainers; c++) {
        const uint32_t m = meta[s * kRowContainers + c];
        if (m & 3u) { keyn++; pay += rows_payload_bytes(m); }
      }
      ow = ovf ? ovf[s] : 0ull;
      if (ow) {
        ometa = rows_pick_type(__popcll(ow), __popcll(ow & ~(ow << 1)));
        keyn++; pay += rows_payload_bytes(ometa);
      }
    }
    const unsigned long long size = s < n_shards ? 8ull + 16ull * keyn + pay : 0ull;
    unsigned long long x = size;  // inclusive scan over the chunk
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    if (wid == 0) {
      unsigned long long t = warp_tot[lane], z = t;
      #pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, z, o); if (lane >= o) z += y; }
      warp_tot[lane] = z - t;  // exclusive
    }
    __syncthreads();
    const unsigned long long start = running + warp_tot[wid] + x - size;
    if (s < n_shards) {
      shard_off[s] = start;
      uint8_t* p = out + start;
      rows_put32(p, 12348u);  // magic, storage version 0, no flags (roaring.go:1056)
      rows_put32(p + 4, keyn);
      uint8_t* hdr = p + 8;
      uint8_t* offs = p + 8 + 12ull * keyn;
      uint32_t off = 8 + 16 * keyn;
      const uint64_t shard = shards[s];
      for (int c = 0; c < kRowContainers; c++) {
        const uint32_t item = s * kRowContainers + c, m = meta[item];
        if (!(m & 3u)) continue;
        rows_put64(hdr, shard * kRowContainers + c);
        rows_put16(hdr + 8, m & 3u);
        rows_put16(hdr + 10, dense_n[item] - 1);
        hdr += 12;
        rows_put32(offs, off); offs += 4;
        item_off[item] = start + off;
        off += rows_payload_bytes(m);
      }
      if (ow) {
        rows_put64(hdr, (shard + 1) * kRowContainers);
        rows_put16(hdr + 8, ometa & 3u);
        rows_put16(hdr + 10, __popcll(ow) - 1);
        rows_put32(offs, off);
        uint8_t* q = p + off;
        if ((ometa & 3u) == kArray) {
          for (unsigned long long v = ow; v; v &= v - 1) { rows_put16(q, __ffsll((long long)v) - 1); q += 2; }
        } else {  // run (a bitmap needs n >= 4096)
          rows_put16(q, ometa >> 2); q += 2;
          unsigned long long st = ow & ~(ow << 1), en = ow & ~(ow >> 1);
          for (; st; st &= st - 1, en &= en - 1) { rows_put16(q, __ffsll((long long)st) - 1); rows_put16(q + 2, __ffsll((long long)en) - 1); q += 4; }
        }
      }
    }
    __syncthreads();
    if (threadIdx.x == 1023) running += warp_tot[31] + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) shard_off[n_shards] = running;
}

// CTA per non-empty container; thread t owns words 4t .. 4t+3.
__global__ void __launch_bounds__(kRowsEmitThreads) rows_emit_kernel(const uint64_t* __restrict__ dense, const uint32_t* __restrict__ meta,
                                                                     const unsigned long long* __restrict__ item_off, uint32_t n_items,
                                                                     uint8_t* __restrict__ out) {
  __shared__ uint32_t warp_tot[kRowsEmitThreads / 32];
  const uint32_t t = threadIdx.x, lane = t & 31, wid = t >> 5;
  for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
    const uint32_t m = meta[item], type = m & 3u;
    if (type == kNil) continue;
    const uint64_t* w = dense + uint64_t(item) * kBitmapWords;
    uint8_t* dst = out + item_off[item];
    if (type == kBitmap) {
      const uint64_t a = reinterpret_cast<uint64_t>(dst);
      if ((a & 15) == 0) {
        const uint4* s4 = reinterpret_cast<const uint4*>(w); uint4* d4 = reinterpret_cast<uint4*>(dst);
        for (uint32_t i = t; i < kBitmapBytes / 16; i += kRowsEmitThreads) d4[i] = s4[i];
      } else if ((a & 7) == 0) {
        uint64_t* d8 = reinterpret_cast<uint64_t*>(dst);
        for (uint32_t i = t; i < kBitmapWords; i += kRowsEmitThreads) d8[i] = w[i];
      } else if ((a & 3) == 0) {
        const uint32_t* s2 = reinterpret_cast<const uint32_t*>(w); uint32_t* d2 = reinterpret_cast<uint32_t*>(dst);
        for (uint32_t i = t; i < kBitmapBytes / 4; i += kRowsEmitThreads) d2[i] = s2[i];
      } else {
        const uint16_t* s1 = reinterpret_cast<const uint16_t*>(w); uint16_t* d1 = reinterpret_cast<uint16_t*>(dst);
        for (uint32_t i = t; i < kBitmapBytes / 2; i += kRowsEmitThreads) d1[i] = s1[i];
      }
      continue;
    }
    const ulonglong2 lo = reinterpret_cast<const ulonglong2*>(w)[2 * t], hi = reinterpret_cast<const ulonglong2*>(w)[2 * t + 1];
    const uint64_t v[4] = {lo.x, lo.y, hi.x, hi.y};
    uint64_t sm[4], em[4];  // array: sm = the words; run: run-start and run-end masks
    uint32_t cnt = 0;       // array: bits; run: starts | ends << 16
    if (type == kArray) {
      #pragma unroll
      for (int i = 0; i < 4; i++) { sm[i] = v[i]; cnt += __popcll(v[i]); }
    } else {
      uint64_t prev_top = t ? w[4 * t - 1] >> 63 : 0ull;
      const uint64_t next_low = t + 1 < kRowsEmitThreads ? w[4 * t + 4] & 1ull : 0ull;
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        const uint64_t nl = i < 3 ? v[i + 1] & 1ull : next_low;
        sm[i] = v[i] & ~((v[i] << 1) | prev_top);
        em[i] = v[i] & ~((v[i] >> 1) | (nl << 63));
        prev_top = v[i] >> 63;
        cnt += __popcll(sm[i]) + (uint32_t(__popcll(em[i])) << 16);
      }
    }
    uint32_t x = cnt;  // block exclusive scan (both halves at once: at most 4096 / 2048 per half)
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();  // warp_tot of the previous item is no longer read
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    uint32_t pre = x - cnt;
    for (uint32_t i = 0; i < wid; i++) pre += warp_tot[i];
    if (type == kArray) {
      uint16_t* d = reinterpret_cast<uint16_t*>(dst) + pre;
      #pragma unroll
      for (int i = 0; i < 4; i++)
        for (uint64_t b = sm[i]; b; b &= b - 1) *d++ = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1);
    } else {
      if (t == 0) *reinterpret_cast<uint16_t*>(dst) = uint16_t(m >> 2);
      uint16_t* ds = reinterpret_cast<uint16_t*>(dst + 2) + 2 * (pre & 0xffffu);
      uint16_t* de = reinterpret_cast<uint16_t*>(dst + 2) + 2 * (pre >> 16) + 1;
      #pragma unroll
      for (int i = 0; i < 4; i++) {
        for (uint64_t b = sm[i]; b; b &= b - 1) { *ds = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1); ds += 2; }
        for (uint64_t b = em[i]; b; b &= b - 1) { *de = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1); de += 2; }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// K12 — protobuf emission of a result row: the `Columns` field of internal.Row as the reference's
// generated marshaller writes it (internal/public.pb.go:745-766, from encodeRow
// encoding/proto/proto.go:408-457): tag 0x0A, varint(payload length), then every column id of
// Row.Columns() (row.go:302-327) as a base-128 varint. Three launches over the dense rows the
// evaluator left in HBM: sizes per (shard, container) slot, one scan, emission.
// A slot is one of the 16 containers of a shard's row or, as the 17th, its Shift overflow word.
// The varint length of a column id only changes at 2^7 and 2^14 (inside the first container of
// shard 0, on 64-bit word boundaries) and at multiples of 2^20 beyond, so it is constant per word.
// ---------------------------------------------------------------------------------------
constexpr int kProtoSlots = kRowContainers + 1;
constexpr int kProtoThreads = 256;

__device__ __forceinline__ uint32_t varint_len(uint64_t x) { return x < 128 ? 1u : uint32_t(70 - __clzll((long long)x)) / 7u; }
__device__ __forceinline__ uint8_t* put_varint(uint8_t* p, uint64_t x) {
  while (x >= 0x80) { *p++ = uint8_t(x) | 0x80; x >>= 7; }
  *p++ = uint8_t(x);
  return p;
}

// A warp per slot: bytes its columns take as varints.
__global__ void __launch_bounds__(kProtoThreads) proto_size_kernel(const uint64_t* __restrict__ dense, const uint32_t* __restrict__ dense_n,
                                                                   const uint16_t* __restrict__ sparse, const uint64_t* __restrict__ shards,
                                                                   const unsigned long long* __restrict__ ovf, uint32_t n_shards, uint32_t* __restrict__ sizes) {
  const uint32_t lane = threadIdx.x & 31, n_slots = n_shards * kProtoSlots;
  const uint32_t warps = gridDim.x * (kProtoThreads / 32);
  for (uint32_t slot = blockIdx.x * (kProtoThreads / 32) + (threadIdx.x >> 5); slot < n_slots; slot += warps) {
    const uint32_t s = slot / kProtoSlots, c = slot % kProtoSlots;
    uint32_t bytes = 0;
    if (c == kRowContainers) {
      if (lane == 0 && ovf)
        for (unsigned long long v = ovf[s]; v; v &= v - 1) bytes += varint_len(((shards[s] + 1) << 20) + uint32_t(__ffsll((long long)v) - 1));
    } else {
      const uint32_t item = s * kRowContainers + c;
      const uint32_t dn = dense_n[item];
      const uint64_t base = (shards[s] << 20) + (uint64_t(c) << 16);
      if (dn & kSparseFlag) {  // a short result kept as a value list: one value per lane
        if (lane < (dn & ~kSparseFlag)) bytes = varint_len(base + sparse[size_t(item) * kSparseListMax + lane]);
      } else if (dn) {
        const uint64_t* w = dense + size_t(item) * kBitmapWords;
        for (uint32_t i = lane; i < kBitmapWords; i += 32) bytes += uint32_t(__popcll(w[i])) * varint_len(base + i * 64);
      }
    }
    bytes = warp_sum(bytes);
    if (lane == 0) sizes[slot] = bytes;
  }
}

// One CTA: exclusive scan of the slot sizes (payload-relative offsets); totals[0] = message length
// (0 for an empty row: the field is omitted), totals[1] = length of tag + length prefix.
// Every thread sums a contiguous chunk of slots, the 1024 chunk sums are scanned once, and a second walk over the
// chunk writes the offsets: two barriers whatever the number of slots (139 264 for 8192 shards).
__global__ void __launch_bounds__(1024) proto_layout_kernel(const uint32_t* __restrict__ sizes, uint32_t n_slots,
                                                            unsigned long long* __restrict__ slot_off, unsigned long long* __restrict__ totals) {
  __shared__ unsigned long long warp_tot[32];
  const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const uint32_t per = (n_slots + 1023) / 1024;
  const uint32_t b = min(n_slots, threadIdx.x * per), e = min(n_slots, b + per);
  unsigned long long mine = 0;
  for (uint32_t i = b; i < e; i++) mine += sizes[i];
  unsigned long long x = mine;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) warp_tot[wid] = x;
  __syncthreads();
  if (wid == 0) {
    unsigned long long t = warp_tot[lane], z = t;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long y = __shfl_up_sync(0xffffffffu, z, o); if (lane >= o) z += y; }
    warp_tot[lane] = z - t;
    if (lane == 31) {
      const unsigned long long payload = z;
      const unsigned long long hdr = payload ? 1ull + varint_len(payload) : 0ull;
      totals[0] = hdr + payload;
      totals[1] = hdr;
    }
  }
  __syncthreads();
  unsigned long long run = warp_tot[wid] + x - mine;
  for (uint32_t i = b; i < e; i++) { slot_off[i] = run; run += sizes[i]; }
}

// A CTA per slot: thread t owns words 4t .. 4t+3 of the container, a block scan places its bytes.
__global__ void __launch_bounds__(kProtoThreads) proto_emit_kernel(const uint64_t* __restrict__ dense, const uint32_t* __restrict__ dense_n,
                                                                   const uint16_t* __restrict__ sparse, const uint32_t* __restrict__ sizes,
                                                                   const unsigned long long* __restrict__ slot_off,
                                                                   const unsigned long long* __restrict__ totals, const uint64_t* __restrict__ shards,
                                                                   const unsigned long long* __restrict__ ovf, uint32_t n_shards,
                                                                   uint8_t* __restrict__ out) {
  __shared__ uint32_t warp_tot[kProtoThreads / 32];
  const uint32_t t = threadIdx.x, lane = t & 31, wid = t >> 5, n_slots = n_shards * kProtoSlots;
  const unsigned long long hdr = totals[1];
  if (blockIdx.x == 0 && t == 0 && hdr) { out[0] = 0x0a; put_varint(out + 1, totals[0] - hdr); }
  // value lists and Shift overflow words: a warp per slot (a sparse row over thousands of shards is nothing but these)
  for (uint32_t slot = blockIdx.x * (kProtoThreads / 32) + wid; slot < n_slots; slot += gridDim.x * (kProtoThreads / 32)) {
    if (sizes[slot] == 0) continue;
    const uint32_t s = slot / kProtoSlots, c = slot % kProtoSlots;
    uint8_t* dst = out + hdr + slot_off[slot];
    if (c == kRowContainers) {
      if (lane == 0)
        for (unsigned long long v = ovf[s]; v; v &= v - 1) dst = put_varint(dst, ((shards[s] + 1) << 20) + uint32_t(__ffsll((long long)v) - 1));
      continue;
    }
    const uint32_t dn = dense_n[s * kRowContainers + c];
    if (!(dn & kSparseFlag)) continue;
    const uint64_t col = (shards[s] << 20) + (uint64_t(c) << 16) + (lane < (dn & ~kSparseFlag) ? sparse[(size_t(s) * kRowContainers + c) * kSparseListMax + lane] : 0u);
    const uint32_t len = lane < (dn & ~kSparseFlag) ? varint_len(col) : 0u;
    uint32_t x = len;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (len) put_varint(dst + x - len, col);
  }
  // containers kept as words: a CTA per slot
  for (uint32_t slot = blockIdx.x; slot < n_slots; slot += gridDim.x) {
    const uint32_t s = slot / kProtoSlots, c = slot % kProtoSlots;
    if (c == kRowContainers || sizes[slot] == 0 || (dense_n[s * kRowContainers + c] & kSparseFlag)) continue;  // uniform over the CTA
    uint8_t* dst = out + hdr + slot_off[slot];
    const uint64_t* w = dense + (size_t(s) * kRowContainers + c) * kBitmapWords + 4 * t;
    const uint64_t base = (shards[s] << 20) + (uint64_t(c) << 16) + uint64_t(4 * t) * 64;
    uint64_t word[4];
    uint32_t bytes = 0;
    #pragma unroll
    for (int i = 0; i < 4; i++) { word[i] = w[i]; bytes += uint32_t(__popcll(word[i])) * varint_len(base + i * 64); }
    uint32_t x = bytes;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();  // warp_tot of the previous slot is no longer read
    if (lane == 31) warp_tot[wid] = x;
    __syncthreads();
    uint32_t pre = x - bytes;
    for (uint32_t i = 0; i < wid; i++) pre += warp_tot[i];
    uint8_t* p = dst + pre;
    #pragma unroll
    for (int i = 0; i < 4; i++)
      for (uint64_t b = word[i]; b; b &= b - 1) p = put_varint(p, base + i * 64 + uint32_t(__ffsll((long long)b) - 1));
  }
}

}  // namespace pb200
