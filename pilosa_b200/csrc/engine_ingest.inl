// Included by engine.cu: writes applied to the RESIDENT copy of a fragment.
//
// fragment.setBit / clearBit (fragment.go:680-779), importPositions (:2051-2098) and importRoaring (:2243-2280 ->
// Bitmap.ImportRoaringBits roaring.go:1463-1558) touch the containers a write names and nothing else. Here a write
// is one small upload (the positions, or the import's own container payloads) and one kernel: a CTA per touched
// container decodes it into a shared-memory window, applies the sets / clears / the operand container, re-types the
// result by Container.optimize's rule (roaring.go:2289-2338), writes the new payload into the context's write
// arena and patches the 16-byte descriptor in the fragment's row table. Nothing of the fragment travels over PCIe
// and no cudaMalloc happens per write (the arena grows by large chunks). Rows the table does not hold yet are added
// by rebuilding the (small) row table on the device. The space of replaced payloads is reclaimed when the fragments
// that wrote into the arena have all been reloaded or dropped (the reference, too, compacts at the next snapshot).

namespace {

struct IngestTask {            // 64 bytes
  ContainerDesc* target;       // descriptor to rewrite, inside the fragment's row table
  uint64_t pad0;
  ContainerDesc other;         // operand container of an import (payload in the upload buffer); n == 0: none
  uint32_t set_off, set_n;     // values to set: offset (in u16 units) into the upload buffer, count
  uint32_t clr_off, clr_n;     // values to clear, applied after the sets
  uint32_t other_clear;        // the import subtracts `other` instead of adding it
  uint32_t pad1[3];
};
static_assert(sizeof(IngestTask) == 64, "IngestTask must be 64 B");

struct IngestResult {  // 16 bytes
  int32_t delta;       // new cardinality - old cardinality
  uint32_t n;          // new cardinality
  uint32_t meta;       // new type | len << 2
  uint32_t old_bytes;  // payload bytes of the container that was replaced
};

constexpr int kIngestThreads = 256;

__global__ void __launch_bounds__(kIngestThreads) ingest_apply_kernel(const IngestTask* __restrict__ tasks, uint32_t n_tasks,
                                                                       const uint16_t* __restrict__ vals, uint8_t* arena,
                                                                       unsigned long long* arena_top, IngestResult* __restrict__ results) {
  __shared__ __align__(16) uint64_t s_w[kBitmapWords];  // the container being rewritten
  __shared__ __align__(16) uint64_t s_o[kBitmapWords];  // the operand of a subtracting import
  __shared__ uint32_t s_tot[kIngestThreads / 32];
  __shared__ unsigned long long s_off;
  const uint32_t t = threadIdx.x, lane = t & 31, wid = t >> 5;
  uint32_t* w32 = reinterpret_cast<uint32_t*>(s_w);
  uint4* w4 = reinterpret_cast<uint4*>(s_w);
  uint4* o4 = reinterpret_cast<uint4*>(s_o);
  for (uint32_t task = blockIdx.x; task < n_tasks; task += gridDim.x) {
    const IngestTask tk = tasks[task];
    const ContainerDesc d = *tk.target;
    __syncthreads();
    // 1. the container as it stands
    if (d.n && desc_type(d) == kBitmap) {
      const uint4* g = reinterpret_cast<const uint4*>(d.addr);
      w4[t] = g[t]; w4[t + kIngestThreads] = g[t + kIngestThreads];
    } else {
      w4[t] = make_uint4(0, 0, 0, 0); w4[t + kIngestThreads] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    if (d.n && desc_type(d) != kBitmap) stage_container(d, w32, 0, kBitmapWords, int(t), kIngestThreads);
    __syncthreads();
    // 2. the operand container of an import: OR in, or subtract
    if (tk.other.n) {
      const bool ob = desc_type(tk.other) == kBitmap;
      if (!tk.other_clear) {
        if (ob) {
          const uint4* g = reinterpret_cast<const uint4*>(tk.other.addr);
          w4[t] = or128(w4[t], g[t]); w4[t + kIngestThreads] = or128(w4[t + kIngestThreads], g[t + kIngestThreads]);
        } else {
          stage_container(tk.other, w32, 0, kBitmapWords, int(t), kIngestThreads);
        }
      } else {
        if (ob) {
          const uint4* g = reinterpret_cast<const uint4*>(tk.other.addr);
          o4[t] = g[t]; o4[t + kIngestThreads] = g[t + kIngestThreads];
        } else {
          o4[t] = make_uint4(0, 0, 0, 0); o4[t + kIngestThreads] = make_uint4(0, 0, 0, 0);
          __syncthreads();
          stage_container(tk.other, reinterpret_cast<uint32_t*>(s_o), 0, kBitmapWords, int(t), kIngestThreads);
        }
        __syncthreads();
        w4[t] = andnot128(w4[t], o4[t]); w4[t + kIngestThreads] = andnot128(w4[t + kIngestThreads], o4[t + kIngestThreads]);
      }
      __syncthreads();
    }
    // 3. single positions: sets, then clears
    for (uint32_t i = t; i < tk.set_n; i += kIngestThreads) { const uint32_t v = vals[tk.set_off + i]; atomicOr(&w32[v >> 5], 1u << (v & 31)); }
    __syncthreads();
    for (uint32_t i = t; i < tk.clr_n; i += kIngestThreads) { const uint32_t v = vals[tk.clr_off + i]; atomicAnd(&w32[v >> 5], ~(1u << (v & 31))); }
    __syncthreads();
    // 4. cardinality, runs (bitmapCountRuns roaring.go:2249), type; thread t owns 64-bit words 4t .. 4t+3
    const ulonglong2 lo = reinterpret_cast<const ulonglong2*>(s_w)[2 * t], hi = reinterpret_cast<const ulonglong2*>(s_w)[2 * t + 1];
    const uint64_t v[4] = {lo.x, lo.y, hi.x, hi.y};
    uint64_t sm[4], em[4];
    uint32_t bits = 0, starts = 0;
    {
      uint64_t prev_top = t ? s_w[4 * t - 1] >> 63 : 0ull;
      const uint64_t next_low = t + 1 < uint32_t(kIngestThreads) ? s_w[4 * t + 4] & 1ull : 0ull;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const uint64_t nl = i < 3 ? v[i + 1] & 1ull : next_low;
        sm[i] = v[i] & ~((v[i] << 1) | prev_top);
        em[i] = v[i] & ~((v[i] >> 1) | (nl << 63));
        prev_top = v[i] >> 63;
        bits += __popcll(v[i]);
        starts += __popcll(sm[i]);
      }
    }
    // block-wide inclusive scans of (bits, starts): the totals pick the type, the prefixes place the output
    uint32_t xb = bits, xs = starts;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t yb = __shfl_up_sync(0xffffffffu, xb, o), ys = __shfl_up_sync(0xffffffffu, xs, o);
      if (lane >= uint32_t(o)) { xb += yb; xs += ys; }
    }
    if (lane == 31) s_tot[wid] = xb | (xs << 16);  // per warp: at most 8192 bits (14 bits), 4096 starts
    __syncthreads();
    uint32_t pre_b = xb - bits, pre_s = xs - starts, n_new = 0, runs = 0;
    for (uint32_t i = 0; i < uint32_t(kIngestThreads / 32); i++) {
      const uint32_t tot = s_tot[i];
      if (i < wid) { pre_b += tot & 0xffffu; pre_s += tot >> 16; }
      n_new += tot & 0xffffu; runs += tot >> 16;
    }
    const uint32_t meta = rows_pick_type(n_new, runs);
    const uint32_t type = meta & 3u, len = meta >> 2;
    const uint32_t bytes = type == kArray ? len * 2 : type == kRun ? len * 4 : type == kBitmap ? uint32_t(kBitmapBytes) : 0u;
    if (t == 0) s_off = bytes ? atomicAdd(arena_top, (unsigned long long)((bytes + 127u) & ~127u)) : 0ull;
    __syncthreads();
    uint8_t* dst = arena + s_off;
    // 5. the new payload, in the layout of layout.h (array: n x u16; run: runs x {start, last}; bitmap: 1024 x u64)
    if (type == kBitmap) {
      uint4* d4 = reinterpret_cast<uint4*>(dst);
      d4[t] = w4[t]; d4[t + kIngestThreads] = w4[t + kIngestThreads];
    } else if (type == kArray) {
      uint16_t* o = reinterpret_cast<uint16_t*>(dst) + pre_b;
#pragma unroll
      for (int i = 0; i < 4; i++)
        for (uint64_t b = v[i]; b; b &= b - 1) *o++ = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1);
    } else if (type == kRun) {
      // run k starts at the k-th start bit and ends at the k-th end bit: both sequences have the same prefix counts
      uint16_t* os = reinterpret_cast<uint16_t*>(dst) + 2 * pre_s;
      uint32_t ends_before = pre_s;  // ends in front of this thread = starts in front of it, minus a run that is still open
      if (t && (s_w[4 * t - 1] >> 63) && (v[0] & 1ull)) ends_before -= 1;
      uint16_t* oe = reinterpret_cast<uint16_t*>(dst) + 2 * ends_before + 1;
#pragma unroll
      for (int i = 0; i < 4; i++) {
        for (uint64_t b = sm[i]; b; b &= b - 1) { *os = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1); os += 2; }
        for (uint64_t b = em[i]; b; b &= b - 1) { *oe = uint16_t((4 * t + i) * 64 + __ffsll((long long)b) - 1); oe += 2; }
      }
    }
    if (t == 0) {
      ContainerDesc nd;
      nd.addr = bytes ? reinterpret_cast<uint64_t>(dst) : 0ull;
      nd.n = n_new;
      nd.meta = n_new ? make_meta(type, len) : 0u;
      *tk.target = nd;
      IngestResult r;
      r.delta = int32_t(n_new) - int32_t(d.n);
      r.n = n_new; r.meta = nd.meta;
      r.old_bytes = d.n == 0 ? 0u : (desc_type(d) == kBitmap ? uint32_t(kBitmapBytes) : desc_type(d) == kArray ? desc_len(d) * 2u : desc_len(d) * 4u);
      results[task] = r;
    }
  }
}

// Rebuilds a fragment's row table for a new (ascending) set of rows: every new table row copies the 16 descriptors of the
// old row with the same id, or starts empty. Works between dense (index = row id) and sparse (index = rank of the id) tables.
__global__ void table_rebuild_kernel(FragDev old, const uint64_t* __restrict__ new_ids, uint32_t new_rows, uint32_t new_dense,
                                     ContainerDesc* __restrict__ new_table) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per descriptor
  if (i >= new_rows * kRowContainers) return;
  const uint32_t r = i / kRowContainers, c = i % kRowContainers;
  const uint64_t id = new_dense ? uint64_t(r) : new_ids[r];
  ContainerDesc d; d.addr = 0; d.n = 0; d.meta = 0;
  if (old.nrows) {
    if (old.dense) { if (id < old.nrows) d = old.rows[id * kRowContainers + c]; }
    else {
      uint32_t lo = 0, hi = old.nrows;
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (old.row_ids[mid] < id) lo = mid + 1; else hi = mid; }
      if (lo < old.nrows && old.row_ids[lo] == id) d = old.rows[uint64_t(lo) * kRowContainers + c];
    }
  }
  new_table[i] = d;
}

constexpr size_t kArenaChunk = 256ull << 20;

// Room for `bytes` more in the write arena (host-side bump of the device-visible top; allocations are 128 B aligned).
// A chunk that is full is retired — descriptors still point into it — and a new one started.
int arena_reserve(pb200_ctx* ctx, size_t bytes) {
  WriteArena& a = ctx->arena;
  bytes += 256;
  if (a.base && a.top + bytes <= a.cap) return PB200_OK;
  const size_t cap = std::max<size_t>(kArenaChunk, (bytes + (size_t(64) << 20)) & ~((size_t(1) << 20) - 1));
  if (ctx->budget && ctx->used + cap > ctx->budget) return fail(ctx, PB200_ERR_NOMEM, "HBM budget exhausted (write arena)");
  void* p = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&p, cap));
  if (!a.d_top) CUDA_TRY(ctx, cudaMalloc(&a.d_top, 8));
  if (a.base) a.retired.push_back({a.base, a.cap});
  a.base = static_cast<uint8_t*>(p); a.cap = cap; a.top = 0;
  ctx->used += cap;
  return PB200_OK;
}

// the fragment starts referencing the arena
void arena_attach(pb200_ctx* ctx, Fragment& f) {
  if (!f.in_arena) { f.in_arena = true; ctx->arena.live_frags++; }
}

// host part of a fragment's present-row list: insert ids (ascending, unique) that are not listed yet, with count 0
void host_add_rows(Fragment& f, const std::vector<uint64_t>& ids) {
  std::vector<uint64_t> nid; std::vector<uint64_t> ncnt;
  nid.reserve(f.row_ids.size() + ids.size()); ncnt.reserve(nid.capacity());
  size_t i = 0, j = 0;
  while (i < f.row_ids.size() || j < ids.size()) {
    if (j == ids.size() || (i < f.row_ids.size() && f.row_ids[i] < ids[j])) { nid.push_back(f.row_ids[i]); ncnt.push_back(f.row_counts[i]); i++; }
    else if (i < f.row_ids.size() && f.row_ids[i] == ids[j]) { nid.push_back(f.row_ids[i]); ncnt.push_back(f.row_counts[i]); i++; j++; }
    else { nid.push_back(ids[j]); ncnt.push_back(0); j++; }
  }
  f.row_ids.swap(nid); f.row_counts.swap(ncnt);
}

// Makes sure the rows `need` (ascending, unique) have a row in the device table of `f`; rebuilds the table in the arena
// when some do not. Returns with f.dev / the family table / the peers' tables current.
int ensure_rows(pb200_ctx* ctx, Family& fam, uint32_t family, uint64_t shard, Fragment& f, const std::vector<uint64_t>& need) {
  std::vector<uint64_t> missing;
  for (uint64_t id : need) {
    const bool have = f.dev.dense ? id < f.dev.nrows : std::binary_search(f.row_ids.begin(), f.row_ids.end(), id);
    if (!have) missing.push_back(id);
  }
  // dense tables: a row inside the table may be unknown to the host list (it had no container so far)
  if (f.dev.dense) {
    std::vector<uint64_t> add;
    for (uint64_t id : need) if (!std::binary_search(f.row_ids.begin(), f.row_ids.end(), id)) add.push_back(id);
    if (missing.empty()) { if (!add.empty()) { host_add_rows(f, add); f.rank_pos_valid = false; } return PB200_OK; }
  }
  if (missing.empty()) return PB200_OK;
  const bool was_dense = f.dev.dense != 0;
  host_add_rows(f, need);
  const uint64_t max_row = f.row_ids.back();
  bool dense = was_dense && max_row < 2 * f.row_ids.size() + 64;  // the rule build_image applies at load
  if (!was_dense) dense = false;
  if (!dense && f.row_ids.size() > 0xffffffffull) return fail(ctx, PB200_ERR_RANGE, "too many rows");
  // dense tables grow with slack so that consecutive new row ids do not rebuild every time
  const uint32_t new_rows = dense ? uint32_t(std::min<uint64_t>(std::max<uint64_t>(max_row + 1, uint64_t(f.dev.nrows) * 3 / 2), 2 * f.row_ids.size() + 64))
                                  : uint32_t(f.row_ids.size());
  const size_t ids_bytes = dense ? 0 : align_up(size_t(new_rows) * 8, 128);
  const size_t tab_bytes = align_up(size_t(new_rows) * kRowContainers * sizeof(ContainerDesc), 128);
  TRY(arena_reserve(ctx, ids_bytes + tab_bytes));
  WriteArena& a = ctx->arena;
  uint8_t* d_ids = a.base + a.top;
  ContainerDesc* d_tab = reinterpret_cast<ContainerDesc*>(a.base + a.top + ids_bytes);
  a.top += ids_bytes + tab_bytes;
  if (!dense) CUDA_TRY(ctx, cudaMemcpyAsync(d_ids, f.row_ids.data(), size_t(new_rows) * 8, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t n_desc = new_rows * kRowContainers;
  table_rebuild_kernel<<<(n_desc + 255) / 256, 256, 0, ctx->stream>>>(f.dev, reinterpret_cast<const uint64_t*>(d_ids), new_rows, dense ? 1u : 0u, d_tab);
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // f.row_ids is pageable; queries issued so far have also drained
  f.dev.row_ids = reinterpret_cast<const uint64_t*>(d_ids);
  f.dev.rows = d_tab;
  f.dev.nrows = new_rows;
  f.dev.dense = dense ? 1u : 0u;
  arena_attach(ctx, f);
  fam.h_frags[shard] = f.dev;
  CUDA_TRY(ctx, cudaMemcpy(fam.d_frags + shard, &f.dev, sizeof(FragDev), cudaMemcpyHostToDevice));
  if (ctx->publish) TRY(ctx->publish(family, shard, f.dev, f.all_bitmap, true));
  f.d_rank_valid = false; f.rank_pos_valid = false;
  ctx->generation++;
  return PB200_OK;
}

struct IngestWork {  // one touched container, host side
  uint64_t key;      // row * 16 + c
  std::vector<uint16_t> sets, clears;
  const HostContainer* other = nullptr;
};

// Runs the tasks of one write against fragment `f`; `results` comes back in task order.
int run_ingest(pb200_ctx* ctx, Family& fam, Fragment& f, std::vector<IngestWork>& work, bool other_clear, std::vector<IngestResult>& results) {
  results.clear();
  const size_t n = work.size();
  if (n == 0) return PB200_OK;
  // upload buffer: [tasks][u16 values][operand payloads]
  size_t n_vals = 0, pay_bytes = 0;
  for (auto& w : work) {
    n_vals += w.sets.size() + w.clears.size();
    if (w.other) pay_bytes += align_up(payload_bytes(w.other->type, w.other->len), 128) + 128;
  }
  const size_t o_vals = n * sizeof(IngestTask);
  const size_t o_pay = align_up(o_vals + n_vals * 2, 128);
  const size_t total = o_pay + pay_bytes;
  TRY(ensure_pin(ctx, ctx->h_ing, total));
  TRY(ensure_dev(ctx, ctx->d_ing, total));
  TRY(ensure_dev(ctx, ctx->d_ing_res, n * sizeof(IngestResult) + 8));
  TRY(ensure_pin(ctx, ctx->h_ing_res, n * sizeof(IngestResult) + 8));
  TRY(arena_reserve(ctx, n * (kBitmapBytes + 128)));
  uint8_t* h = static_cast<uint8_t*>(ctx->h_ing.p);
  uint8_t* dv = static_cast<uint8_t*>(ctx->d_ing.p);
  auto* tasks = reinterpret_cast<IngestTask*>(h);
  uint16_t* hv = reinterpret_cast<uint16_t*>(h + o_vals);
  size_t vo = 0, po = o_pay;
  for (size_t i = 0; i < n; i++) {
    IngestWork& w = work[i];
    IngestTask tk{};
    const uint64_t row = w.key >> 4;
    uint64_t idx;
    if (f.dev.dense) idx = row;
    else idx = uint64_t(std::lower_bound(f.row_ids.begin(), f.row_ids.end(), row) - f.row_ids.begin());
    tk.target = const_cast<ContainerDesc*>(f.dev.rows) + idx * kRowContainers + (w.key & 15);
    tk.set_off = uint32_t(vo); tk.set_n = uint32_t(w.sets.size());
    std::copy(w.sets.begin(), w.sets.end(), hv + vo); vo += w.sets.size();
    tk.clr_off = uint32_t(vo); tk.clr_n = uint32_t(w.clears.size());
    std::copy(w.clears.begin(), w.clears.end(), hv + vo); vo += w.clears.size();
    tk.other_clear = other_clear ? 1u : 0u;
    if (w.other && w.other->n) {
      const HostContainer& oc = *w.other;
      const uint32_t pb = payload_bytes(oc.type, oc.len);
      std::memcpy(h + po, oc.data, pb);
      std::memset(h + po + pb, 0, align_up(pb, 128) + 128 - pb);
      tk.other.addr = reinterpret_cast<uint64_t>(dv + po);
      tk.other.n = oc.n;
      tk.other.meta = make_meta(oc.type, oc.len);
      po += align_up(pb, 128) + 128;
    }
    tasks[i] = tk;
  }
  WriteArena& a = ctx->arena;
  unsigned long long top = a.top;
  CUDA_TRY(ctx, cudaMemcpyAsync(a.d_top, &top, 8, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dv, h, total, cudaMemcpyHostToDevice, ctx->stream));
  const uint32_t grid = uint32_t(std::min<size_t>(n, size_t(ctx->sm_count) * 4));
  ingest_apply_kernel<<<grid, kIngestThreads, 0, ctx->stream>>>(reinterpret_cast<const IngestTask*>(dv), uint32_t(n),
                                                               reinterpret_cast<const uint16_t*>(dv + o_vals), a.base, a.d_top,
                                                               static_cast<IngestResult*>(ctx->d_ing_res.p));
  CUDA_TRY(ctx, cudaGetLastError());
  uint8_t* hr = static_cast<uint8_t*>(ctx->h_ing_res.p);
  CUDA_TRY(ctx, cudaMemcpyAsync(hr, ctx->d_ing_res.p, n * sizeof(IngestResult), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(hr + n * sizeof(IngestResult), a.d_top, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  results.assign(reinterpret_cast<IngestResult*>(hr), reinterpret_cast<IngestResult*>(hr) + n);
  std::memcpy(&top, hr + n * sizeof(IngestResult), 8);
  a.top = size_t(top);
  ctx->timing.h2d_bytes = total + 8;
  ctx->timing.d2h_bytes = n * sizeof(IngestResult) + 8;
  // host bookkeeping: row counts, statistics, container-type flags
  arena_attach(ctx, f);
  bool sparse_now = false;
  for (size_t i = 0; i < n; i++) {
    const IngestResult& r = results[i];
    const uint64_t row = work[i].key >> 4;
    const size_t j = size_t(std::lower_bound(f.row_ids.begin(), f.row_ids.end(), row) - f.row_ids.begin());
    f.row_counts[j] = uint64_t(int64_t(f.row_counts[j]) + r.delta);
    if (f.row_counts[j] == 0) f.maybe_empty_rows = true;
    const uint32_t type = r.meta & 3u, len = (r.meta >> 2) & 0x1fffu;
    const uint32_t nb = r.n == 0 ? 0u : type == kBitmap ? uint32_t(kBitmapBytes) : type == kArray ? len * 2u : len * 4u;
    const bool had = r.old_bytes != 0, has = r.n != 0;
    f.n_containers += (has ? 1 : 0) - (had ? 1 : 0);
    fam.containers_total += (has ? 1 : 0) - (had ? 1 : 0);
    f.payload_bytes += nb; f.payload_bytes -= r.old_bytes;
    fam.payload_total += nb; fam.payload_total -= r.old_bytes;
    if (has && type != kBitmap && r.n != kFullN) sparse_now = true;
  }
  if (sparse_now && f.all_bitmap) { f.all_bitmap = false; fam.n_not_all_bitmap++; }
  f.d_rank_valid = false; f.rank_pos_valid = false;
  ctx->generation++;
  return PB200_OK;
}

}  // namespace

namespace {

bool row_in_table(const Fragment& f, uint64_t row) {
  return f.dev.dense ? row < f.dev.nrows : std::binary_search(f.row_ids.begin(), f.row_ids.end(), row);
}

}  // namespace

extern "C" int pb200_frag_apply(pb200_ctx* ctx, pb200_family family, uint64_t shard, const uint64_t* set_pos, size_t n_set,
                                const uint64_t* clear_pos, size_t n_clear) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if ((n_set && !set_pos) || (n_clear && !clear_pos)) return PB200_ERR_INVALID;
  Family& fam = ctx->families[family];
  auto it = fam.frags.find(shard);
  if (it == fam.frags.end()) return frag_apply_host(ctx, family, shard, set_pos, n_set, clear_pos, n_clear);
  Fragment& f = it->second;
  for (size_t i = 0; i < n_set; i++)
    if ((set_pos[i] >> 20) >= (1ull << 44)) return fail(ctx, PB200_ERR_RANGE, "position out of range");
  std::map<uint64_t, IngestWork> by_key;  // container key -> its values
  std::set<uint64_t> set_rows, named_rows;
  for (size_t i = 0; i < n_set; i++) { by_key[set_pos[i] >> 16].sets.push_back(uint16_t(set_pos[i])); set_rows.insert(set_pos[i] >> 20); named_rows.insert(set_pos[i] >> 20); }
  for (size_t i = 0; i < n_clear; i++) { by_key[clear_pos[i] >> 16].clears.push_back(uint16_t(clear_pos[i])); named_rows.insert(clear_pos[i] >> 20); }
  rank_state_build(f, ctx->rank_cache_size);  // the cache as it stands before the write (rewrite_fragment does the same)
  TRY(ensure_rows(ctx, fam, family, shard, f, std::vector<uint64_t>(set_rows.begin(), set_rows.end())));
  std::vector<IngestWork> work;
  for (auto& kv : by_key) {
    if (!row_in_table(f, kv.first >> 4)) continue;  // clears on a row the fragment does not have
    if (!std::binary_search(f.row_ids.begin(), f.row_ids.end(), kv.first >> 4)) continue;  // dense table slot that never held data: nothing to clear
    kv.second.key = kv.first;
    work.push_back(std::move(kv.second));
  }
  std::vector<IngestResult> res;
  TRY(run_ingest(ctx, fam, f, work, false, res));
  // importPositions' recount (fragment.go:2080-2093): the cached count of every row a position names is re-read
  // from storage (BulkAdd), then Recalculate
  for (uint64_t row : named_rows) {
    const auto r = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), row);
    rank_bulk_add(f, row, (r != f.row_ids.end() && *r == row) ? f.row_counts[size_t(r - f.row_ids.begin())] : 0);
  }
  rank_recalculate(f, ctx->rank_cache_size);
  return PB200_OK;
}

extern "C" int pb200_frag_import_roaring(pb200_ctx* ctx, pb200_family family, uint64_t shard, const void* data, size_t len, int clear,
                                         int64_t* changed) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (!data) return fail(ctx, PB200_ERR_INVALID, "no roaring bitmap provided");  // roaring.go:1464
  std::vector<HostContainer> other;
  std::string err;
  if (!parse_any(static_cast<const uint8_t*>(data), len, other, err)) return fail(ctx, PB200_ERR_FORMAT, err);
  for (auto& c : other)
    if ((c.key >> 4) >= (1ull << 44)) return fail(ctx, PB200_ERR_FORMAT, "container key out of range");
  Family& fam = ctx->families[family];
  auto it = fam.frags.find(shard);
  if (it == fam.frags.end()) return frag_import_host(ctx, family, shard, other, clear, changed);
  Fragment& f = it->second;
  // operands beyond the descriptor's length field become the bitmaps Container.optimize would make of them (as at load)
  for (auto& c : other) {
    if (!((c.type == kArray && c.len > 4096) || (c.type == kRun && c.len > 2048))) continue;
    std::vector<uint8_t> words(kBitmapBytes);
    to_words(c, reinterpret_cast<uint64_t*>(words.data()));
    c.owned.swap(words); c.data = c.owned.data(); c.type = kBitmap; c.len = kBitmapWords;
  }
  std::set<uint64_t> need;
  if (!clear) for (auto& c : other) if (c.n) need.insert(c.key >> 4);
  rank_state_build(f, ctx->rank_cache_size);
  TRY(ensure_rows(ctx, fam, family, shard, f, std::vector<uint64_t>(need.begin(), need.end())));
  std::vector<IngestWork> work;
  for (auto& c : other) {
    if (c.n == 0) continue;
    if (!row_in_table(f, c.key >> 4) || !std::binary_search(f.row_ids.begin(), f.row_ids.end(), c.key >> 4)) continue;  // subtracting from a row that is not there
    IngestWork w; w.key = c.key; w.other = &c;
    work.push_back(std::move(w));
  }
  std::vector<IngestResult> res;
  TRY(run_ingest(ctx, fam, f, work, clear != 0, res));
  int64_t n = 0;
  std::map<uint64_t, int64_t> row_set;  // ImportRoaringBits' rowSet: signed change per row
  for (size_t i = 0; i < res.size(); i++) {
    n += clear ? -int64_t(res[i].delta) : int64_t(res[i].delta);
    if (res[i].delta) row_set[work[i].key >> 4] += res[i].delta;
  }
  if (changed) *changed = n;
  if (n == 0) return PB200_OK;  // no row changed: the cache stays as it is
  // fragment.importRoaring (fragment.go:2258-2276): BulkAdd(row, Get(row) + changes) for every changed row
  // (uint64 arithmetic), then Recalculate
  for (auto& kv : row_set) {
    if (kv.second == 0) continue;
    const auto e = f.rank_state.entries.find(kv.first);
    rank_bulk_add(f, kv.first, (e == f.rank_state.entries.end() ? 0 : e->second) + uint64_t(kv.second));
  }
  rank_recalculate(f, ctx->rank_cache_size);
  return PB200_OK;
}
