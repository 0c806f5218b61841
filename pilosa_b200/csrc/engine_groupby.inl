// Included by engine.cu: Rows and GroupBy.
//
// GroupBy in the reference is a nested iterator per shard (groupByIterator executor.go:3060-3230)
// whose only arithmetic is card(filter & r0 & r1 & ... & r_last) for the combinations it visits.
// Here ALL combinations are counted on the device in one pass per chunk — the prefix rows
// filter & r0 & ... & r_{k-2} are evaluated densely for every (prefix tuple, shard) and the TopN
// count kernel runs the last field's rows against each of them — and the iterator's bookkeeping
// (ascending order, `previous` seek / wrap rules, per-shard limit, mergeGroupCounts) is replayed on
// the host over that count tensor, one thread per shard.

namespace {

// a row is present in a shard when the fragment's row table lists it (non-empty containers only)
std::vector<uint64_t> live_row_ids(const Fragment& f) {  // rows a write has emptied stay listed on the host with count 0
  if (!f.maybe_empty_rows) return f.row_ids;
  std::vector<uint64_t> out;
  for (size_t i = 0; i < f.row_ids.size(); i++) if (f.row_counts[i]) out.push_back(f.row_ids[i]);
  return out;
}

std::vector<uint64_t> present_rows(const Fragment& f, const uint64_t* only, uint64_t n_only) {
  if (!only) return live_row_ids(f);
  const std::vector<uint64_t> live = live_row_ids(f);
  std::vector<uint64_t> out;
  std::set_intersection(live.begin(), live.end(), only, only + n_only, std::back_inserter(out));
  return out;
}

struct GbTensor {  // counts for every (prefix tuple q, shard s, last-field row j), as the device produced them
  uint32_t k = 0, n_shards = 0;
  std::vector<std::vector<uint64_t>> G;       // candidate row ids per field (ascending, union over shards)
  std::vector<uint32_t> final_counts;         // [(q * n_shards + s) * |G_last| + j]  (k == 1: q = 0)
  std::vector<uint64_t> prefix_counts;        // [q * n_shards + s]: card(filter & r0 & ... & r_{k-2}); empty when k == 1
  uint64_t n_last() const { return G[k - 1].size(); }
};

// The iterator of one shard, over indices into G_i instead of rows (groupByIterator :3060-3230,
// rowIterator fragment.go:2759-2791).
struct GbWalk {
  const GbTensor& t;
  uint32_t s;
  std::vector<std::vector<uint32_t>> lists;  // per field: indices (into G_i) of the rows this shard has, ascending
  std::vector<size_t> cur;
  std::vector<uint32_t> ids;                 // current index per field
  bool done = false;

  GbWalk(const GbTensor& tt, uint32_t shard_pos) : t(tt), s(shard_pos), lists(tt.k), cur(tt.k, 0), ids(tt.k, 0) {}

  void seek(size_t i, uint64_t row_id) {  // rowIterator.Seek
    const auto& g = t.G[i];
    cur[i] = size_t(std::lower_bound(lists[i].begin(), lists[i].end(), row_id, [&](uint32_t idx, uint64_t v) { return g[idx] < v; }) - lists[i].begin());
  }
  bool next(size_t i, uint32_t* idx, bool* wrapped) {  // rowIterator.Next; wrap = i != 0
    *wrapped = false;
    if (cur[i] >= lists[i].size()) {
      if (i == 0 || lists[i].empty()) { *wrapped = true; return false; }
      cur[i] = 0;
      *wrapped = true;
    }
    *idx = lists[i][cur[i]++];
    return true;
  }
  uint64_t prefix_index() const {  // q of (ids[0] .. ids[k-2])
    uint64_t q = 0;
    for (uint32_t i = 0; i + 1 < t.k; i++) q = q * t.G[i].size() + ids[i];
    return q;
  }
  uint64_t final_count() const {
    const uint64_t q = t.k == 1 ? 0 : prefix_index();
    return t.final_counts[(q * t.n_shards + s) * t.n_last() + ids[t.k - 1]];
  }
  // rows[i].row.IsEmpty() after nextAtIdx stored it. Only the k == 1 row and the level k-2 prefix can be
  // empty in a way this walk needs to know: the last field's row is raw (never empty), and skipping empty
  // rows at lower levels is, in the reference's own words (:3168), an optimisation that can be removed.
  bool level_empty(size_t i) const {
    if (t.k == 1) return final_count() == 0;
    if (i + 2 == t.k) return t.prefix_counts[prefix_index() * t.n_shards + s] == 0;
    return false;
  }
  void next_at(size_t i) {  // :3167-3191
    for (;;) {
      uint32_t idx; bool wrapped;
      if (!next(i, &idx, &wrapped)) { done = true; return; }
      if (wrapped && i != 0) next_at(i - 1);
      if (done) return;  // the reference would keep wrapping for ever here (see pilosa_b200.h)
      ids[i] = idx;
      if (!level_empty(i)) break;
    }
  }
  bool next_group(pb200_group* out, uint64_t* dense_index = nullptr) {  // :3193-3230
    uint64_t count;
    for (;;) {
      if (done) return false;
      count = final_count();
      if (count == 0) { next_at(t.k - 1); continue; }
      break;
    }
    *out = pb200_group{};
    for (uint32_t i = 0; i < t.k; i++) out->rows[i] = t.G[i][ids[i]];
    out->count = count;
    if (dense_index) *dense_index = (t.k == 1 ? 0 : prefix_index()) * t.n_last() + ids[t.k - 1];  // groups in key order = index order
    next_at(t.k - 1);
    return true;
  }
  // newGroupByIterator :3086-3165
  void init(const pb200_groupby_child* ch) {
    bool ignore_prev = false;
    for (size_t i = 0; i < t.k; i++) {
      uint64_t prev = ch[i].previous;
      if (ch[i].has_previous && !ignore_prev) {
        if (i == t.k - 1) prev++;
        seek(i, prev);
      }
      uint32_t idx; bool wrapped;
      if (!next(i, &idx, &wrapped)) { done = true; return; }
      ids[i] = idx;
      if (ch[i].has_previous && t.G[i][idx] != prev) ignore_prev = true;
      if (wrapped) {
        for (size_t j = i; j-- > 0;) {
          uint32_t idx2; bool w2;
          if (!next(j, &idx2, &w2)) { done = true; return; }
          ids[j] = idx2;
          if (!w2) break;
        }
      }
    }
  }
};

int group_compare(const pb200_group& a, const pb200_group& b, uint32_t k) {  // GroupCount.Compare :1203-1214
  for (uint32_t i = 0; i < k; i++) {
    if (a.rows[i] < b.rows[i]) return -1;
    if (a.rows[i] > b.rows[i]) return 1;
  }
  return 0;
}

std::vector<pb200_group> merge_group_counts(std::vector<pb200_group> a, const std::vector<pb200_group>& b, uint64_t limit, uint32_t k) {  // :1166-1200
  limit = std::min<uint64_t>(limit, a.size() + b.size());
  std::vector<pb200_group> ret;
  size_t i = 0, j = 0;
  while (i < a.size() && j < b.size() && ret.size() < limit) {
    const int c = group_compare(a[i], b[j], k);
    if (c < 0) ret.push_back(a[i++]);
    else if (c == 0) { a[i].count += b[j].count; ret.push_back(a[i]); i++; j++; }
    else ret.push_back(b[j++]);
  }
  for (; i < a.size() && ret.size() < limit; i++) ret.push_back(a[i]);
  for (; j < b.size() && ret.size() < limit; j++) ret.push_back(b[j]);
  return ret;
}

// sums[q * n_ids + j] = sum over the S shards of counts[(q * S + s) * n_ids + j]: the reduce of a GroupBy without a
// limit or a `previous` bound, where every shard's walk emits exactly its non-zero counts (one thread per group)
__global__ void gb_sum_kernel(const uint32_t* __restrict__ counts, uint32_t Q, uint32_t S, uint32_t n_ids, unsigned long long* __restrict__ sums) {
  const uint64_t n = uint64_t(Q) * n_ids;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += uint64_t(gridDim.x) * blockDim.x) {
    const uint64_t q = i / n_ids, j = i % n_ids;
    const uint32_t* c = counts + q * S * n_ids + j;
    unsigned long long t = 0;
    for (uint32_t s = 0; s < S; s++) t += c[uint64_t(s) * n_ids];
    sums[i] = t;
  }
}

// Runs the TopN count kernel for `n_virtual` (prefix, shard) source rows resident in d_dense against the
// explicit id list `ids`; counts land in h_counts[v * n_ids + j] — or, with d_sums, stay on the device and are summed
// over the S shards of each of the n_virtual / S prefix tuples into d_sums[q * n_ids + j].
int gb_count_against(pb200_ctx* ctx, pb200_family family, const uint64_t* virtual_shards, uint32_t n_virtual,
                     const std::vector<uint64_t>& ids, uint32_t* h_counts, unsigned long long* d_sums = nullptr, uint32_t S = 0) {
  const size_t n_ids = ids.size();
  const size_t o_sh = 0, o_ids = size_t(n_virtual) * 8, total_bytes = o_ids + n_ids * 8;
  TRY(ensure_dev(ctx, ctx->d_topn, total_bytes));
  TRY(ensure_pin(ctx, ctx->h_topn, total_bytes));
  uint8_t* h = static_cast<uint8_t*>(ctx->h_topn.p);
  uint8_t* d = static_cast<uint8_t*>(ctx->d_topn.p);
  std::memcpy(h + o_sh, virtual_shards, size_t(n_virtual) * 8);
  std::memcpy(h + o_ids, ids.data(), n_ids * 8);
  CUDA_TRY(ctx, cudaMemcpyAsync(d, h, total_bytes, cudaMemcpyHostToDevice, ctx->stream));
  ctx->timing.h2d_bytes += total_bytes;
  const size_t n_counts = size_t(n_virtual) * n_ids;
  TRY(ensure_dev(ctx, ctx->d_bsi, n_counts * 4));
  TopnParams tp{};
  tp.families = ctx->d_families; tp.family = family;
  tp.shards = reinterpret_cast<const uint64_t*>(d + o_sh); tp.n_shards = n_virtual;
  tp.src_dense = static_cast<const uint8_t*>(ctx->d_dense.p);
  tp.src_n = static_cast<const uint32_t*>(ctx->d_dense_n.p);
  tp.ids = reinterpret_cast<const uint64_t*>(d + o_ids);
  tp.n_ids = uint32_t(n_ids);
  tp.chunk = 2048;
  tp.chunks_per_shard = std::max<uint32_t>(1, uint32_t((n_ids + tp.chunk - 1) / tp.chunk));
  tp.counts = static_cast<uint32_t*>(ctx->d_bsi.p);
  tp.algo_bytes = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_misc.p) + 8);
  if (!ctx->topn_attr_set) {
    CUDA_TRY(ctx, cudaFuncSetAttribute(topn_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopnSmem));
    ctx->topn_attr_set = true;
  }
  const uint64_t items = uint64_t(n_virtual) * tp.chunks_per_shard;
  const uint32_t grid = uint32_t(std::min<uint64_t>(items, resident_ctas(ctx, OCC_TOPN, topn_count_kernel, kTopnThreads, kTopnSmem)));
  topn_count_kernel<<<grid, kTopnThreads, kTopnSmem, ctx->stream>>>(tp);
  ctx->timing.launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  if (d_sums) {
    const uint32_t Q = n_virtual / S;
    const uint64_t n = uint64_t(Q) * n_ids;
    gb_sum_kernel<<<uint32_t(std::min<uint64_t>((n + 255) / 256, uint64_t(ctx->sm_count) * 8)), 256, 0, ctx->stream>>>(tp.counts, Q, S, uint32_t(n_ids), d_sums);
    ctx->timing.launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    return PB200_OK;
  }
  TRY(ensure_pin(ctx, ctx->h_out, n_counts * 4));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_out.p, tp.counts, n_counts * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes += n_counts * 4;
  std::memcpy(h_counts, ctx->h_out.p, n_counts * 4);
  return PB200_OK;
}

// the 16 per-container cardinalities of every dense row in d_dense -> one count per row
int gb_dense_row_counts(pb200_ctx* ctx, uint32_t n_rows, uint64_t* out) {
  std::vector<uint32_t> dn(size_t(n_rows) * kRowContainers);
  CUDA_TRY(ctx, cudaMemcpyAsync(dn.data(), ctx->d_dense_n.p, dn.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes += dn.size() * 4;
  for (uint32_t r = 0; r < n_rows; r++) {
    uint64_t c = 0;
    for (int j = 0; j < kRowContainers; j++) c += dn[size_t(r) * kRowContainers + j];
    out[r] = c;
  }
  return PB200_OK;
}

}  // namespace

// which table rows (id >= start) of one fragment have bit `lo` set in container column c (filterColumn fragment.go:2622)
__global__ void rows_with_column_kernel(FragDev f, uint64_t start, uint32_t c, uint32_t lo, uint8_t* __restrict__ flags) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= f.nrows) return;
  const uint64_t id = f.dense ? uint64_t(t) : f.row_ids[t];
  uint8_t hit = 0;
  if (id >= start) {
    const ContainerDesc d = f.rows[uint64_t(t) * kRowContainers + c];
    if (d.n == kFullN) hit = 1;
    else if (d.n != 0) {
      const uint32_t type = desc_type(d), len = desc_len(d);
      if (type == kBitmap) hit = in_bitmap(d.addr, lo);
      else if (type == kArray) hit = in_array(d.addr, len, lo);
      else {
        const ushort2* r = reinterpret_cast<const ushort2*>(d.addr);
        const uint32_t j = run_lower(r, len, lo);
        hit = j < len && __ldg(r + j).x <= lo;
      }
    }
  }
  flags[t] = hit;
}

extern "C" int pb200_rows(pb200_ctx* ctx, pb200_family family, uint64_t start, uint32_t flags, uint64_t column, uint64_t limit,
                          const uint64_t* shards, uint32_t n_shards, uint64_t* out_rows, size_t cap, size_t* n_out) {
  ENTER(ctx);
  if (!n_out) return PB200_ERR_INVALID;
  *n_out = 0;
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  TRY(validate_shards(ctx, shards, n_shards));
  Family& fam = ctx->families[family];
  const bool has_column = (flags & PB200_ROWS_COLUMN) != 0;
  std::vector<uint64_t> acc;
  for (uint32_t si = 0; si < n_shards; si++) {
    auto it = fam.frags.find(shards[si]);
    if (it == fam.frags.end()) continue;
    if (has_column && column / PB200_SHARD_WIDTH != shards[si]) continue;  // executor.go:1405-1409
    const Fragment& f = it->second;
    std::vector<uint64_t> mine;
    if (!has_column) {
      const std::vector<uint64_t> live = live_row_ids(f);
      mine.assign(std::lower_bound(live.begin(), live.end(), start), live.end());
    } else if (f.dev.nrows) {
      TRY(sync_families(ctx));
      TRY(ensure_dev(ctx, ctx->d_slow, f.dev.nrows));
      const uint64_t col = column % PB200_SHARD_WIDTH;
      rows_with_column_kernel<<<(f.dev.nrows + 255) / 256, 256, 0, ctx->stream>>>(f.dev, start, uint32_t(col >> 16), uint32_t(col & 0xffff),
                                                                               static_cast<uint8_t*>(ctx->d_slow.p));
      CUDA_TRY(ctx, cudaGetLastError());
      std::vector<uint8_t> hit(f.dev.nrows);
      CUDA_TRY(ctx, cudaMemcpyAsync(hit.data(), ctx->d_slow.p, hit.size(), cudaMemcpyDeviceToHost, ctx->stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
      for (uint32_t t = 0; t < f.dev.nrows; t++)
        if (hit[t]) mine.push_back(f.dev.dense ? uint64_t(t) : f.row_ids[t]);
    }
    if (mine.size() > limit) mine.resize(size_t(limit));  // filterWithLimit (fragment.go:2612)
    // RowIDs.merge (executor.go:1043-1069)
    std::vector<uint64_t> merged;
    size_t i = 0, j = 0;
    while (i < acc.size() && j < mine.size() && merged.size() < limit) {
      if (acc[i] < mine[j]) merged.push_back(acc[i++]);
      else if (acc[i] > mine[j]) merged.push_back(mine[j++]);
      else { merged.push_back(mine[j]); i++; j++; }
    }
    while (i < acc.size() && merged.size() < limit) merged.push_back(acc[i++]);
    while (j < mine.size() && merged.size() < limit) merged.push_back(mine[j++]);
    acc.swap(merged);
  }
  *n_out = acc.size();
  if (acc.size() > cap) return fail(ctx, PB200_ERR_RANGE, "row buffer too small");
  if (!acc.empty() && !out_rows) return PB200_ERR_INVALID;
  std::copy(acc.begin(), acc.end(), out_rows);
  return PB200_OK;
}

extern "C" int pb200_groupby(pb200_ctx* ctx, const pb200_groupby_child* children, uint32_t n_children, const pb200_node* filter_prog,
                             uint32_t n_filter_nodes, uint64_t limit, const uint64_t* shards, uint32_t n_shards, pb200_group* out,
                             size_t cap, size_t* n_out) {
  ENTER(ctx);
  if (!n_out) return PB200_ERR_INVALID;
  *n_out = 0;
  if (n_children == 0) return fail(ctx, PB200_ERR_INVALID, "need at least one child call");  // executor.go:1073
  if (!children || n_children > PB200_GROUPBY_MAX_FIELDS) return fail(ctx, PB200_ERR_UNSUPPORTED, "GroupBy takes 1 to 4 Rows() children");
  TRY(validate_shards(ctx, shards, n_shards));
  const uint32_t k = n_children;
  for (uint32_t i = 0; i < k; i++) {
    if (children[i].family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
    if (children[i].n_rows && !children[i].rows) return PB200_ERR_INVALID;
  }
  const bool has_filter = filter_prog && n_filter_nodes;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));

  // shards that take part (every field needs a fragment there, :3110-3113) and their row lists
  std::vector<uint32_t> live;                            // positions in `shards`
  std::vector<std::vector<std::vector<uint64_t>>> rows;  // [live][field] -> row ids
  for (uint32_t si = 0; si < n_shards; si++) {
    std::vector<std::vector<uint64_t>> r(k);
    bool ok = true;
    for (uint32_t i = 0; i < k && ok; i++) {
      auto it = ctx->families[children[i].family].frags.find(shards[si]);
      if (it == ctx->families[children[i].family].frags.end()) { ok = false; break; }
      r[i] = present_rows(it->second, children[i].n_rows ? children[i].rows : nullptr, children[i].n_rows);
    }
    if (ok) { live.push_back(si); rows.push_back(std::move(r)); }
  }
  GbTensor t;
  t.k = k; t.n_shards = uint32_t(live.size());
  t.G.resize(k);
  for (uint32_t i = 0; i < k; i++) {
    for (auto& r : rows) t.G[i].insert(t.G[i].end(), r[i].begin(), r[i].end());
    std::sort(t.G[i].begin(), t.G[i].end());
    t.G[i].erase(std::unique(t.G[i].begin(), t.G[i].end()), t.G[i].end());
  }
  bool nothing = live.empty();
  for (uint32_t i = 0; i < k; i++) nothing = nothing || t.G[i].empty();
  std::vector<pb200_group> acc;
  if (!nothing) {
    const uint32_t S = t.n_shards;
    std::vector<uint64_t> live_shards(S);
    for (uint32_t v = 0; v < S; v++) live_shards[v] = shards[live[v]];
    uint64_t n_prefix = 1;
    for (uint32_t i = 0; i + 1 < k; i++) {
      n_prefix *= t.G[i].size();
      if (n_prefix > (1ull << 27)) return fail(ctx, PB200_ERR_RANGE, "GroupBy cross product too large (rows x shards > 2^27 counts)");
    }
    const uint64_t n_last = t.G[k - 1].size();
    if (n_prefix * S * n_last > (1ull << 27)) return fail(ctx, PB200_ERR_RANGE, "GroupBy cross product too large (rows x shards > 2^27 counts)");
    // Without a limit that can cut a shard's list and without a `previous` bound, the walk of a shard emits exactly its
    // non-zero counts, and the reduce over the shards is a plain sum per group: that sum is taken on the device and only
    // the group totals come back (no per-shard counts over PCIe, no host walk). (k == 1 without a filter has its counts
    // on the host already.)
    uint64_t n_groups_all = n_prefix * n_last;
    bool any_prev = false;
    for (uint32_t i = 0; i < k; i++) any_prev |= children[i].has_previous != 0;
    const bool device_reduce = !any_prev && limit >= n_groups_all && (k >= 2 || has_filter) && n_groups_all <= (1ull << 24);
    unsigned long long* d_sums = nullptr;
    if (device_reduce) {
      TRY(ensure_dev(ctx, ctx->d_gb, size_t(n_groups_all) * 8));
      d_sums = static_cast<unsigned long long*>(ctx->d_gb.p);
    } else {
      t.final_counts.assign(size_t(n_prefix * S * n_last), 0);
    }
    TRY(sync_families(ctx));
    const size_t row_bytes = size_t(kRowContainers) * kBitmapBytes;

    if (has_filter) {  // the filter row of every shard, evaluated once: kept as dense temp slot 0
      BatchDev fbd{};
      TRY(eval_dense(ctx, filter_prog, n_filter_nodes, live_shards.data(), S, fbd));
      if (k >= 2) {
        TRY(ensure_dev(ctx, ctx->d_temp, size_t(S) * row_bytes));  // eval_dense is done with its own temps
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_temp.p, ctx->d_dense.p, size_t(S) * row_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
      }
    }
    if (k == 1) {
      if (has_filter) {  // counts of filter & row: the filter rows are the sources
        TRY(gb_count_against(ctx, children[0].family, live_shards.data(), S, t.G[0], device_reduce ? nullptr : t.final_counts.data(), d_sums, S));
      } else {           // Row.Count() of every present row: the loader already holds the exact cardinalities
        for (uint32_t v = 0; v < S; v++) {
          const Fragment& f = ctx->families[children[0].family].frags.find(live_shards[v])->second;
          for (size_t r = 0; r < f.row_ids.size(); r++) {
            const auto g = std::lower_bound(t.G[0].begin(), t.G[0].end(), f.row_ids[r]);
            if (g != t.G[0].end() && *g == f.row_ids[r]) t.final_counts[size_t(v) * n_last + size_t(g - t.G[0].begin())] = uint32_t(f.row_counts[r]);
          }
        }
      }
    } else {
      if (!device_reduce) t.prefix_counts.assign(size_t(n_prefix * S), 0);
      // prefix tuples in chunks of at most 8192 dense rows (1 GiB)
      const uint64_t q_per_chunk = std::max<uint64_t>(1, 8192 / S);
      std::vector<pb200_node> nodes;
      std::vector<uint32_t> prog_off;
      std::vector<uint64_t> vshards;
      for (uint64_t q0 = 0; q0 < n_prefix; q0 += q_per_chunk) {
        const uint32_t Q = uint32_t(std::min<uint64_t>(q_per_chunk, n_prefix - q0));
        nodes.clear(); prog_off.clear();
        for (uint32_t qi = 0; qi < Q; qi++) {
          prog_off.push_back(uint32_t(nodes.size()));
          uint64_t rem = q0 + qi;
          uint32_t idx[PB200_GROUPBY_MAX_FIELDS] = {0, 0, 0, 0};
          for (int i = int(k) - 2; i >= 0; i--) { idx[i] = uint32_t(rem % t.G[i].size()); rem /= t.G[i].size(); }
          uint32_t n_leaves = 0;
          if (has_filter) { pb200_node nd{}; nd.op = kOpInternalTemp; nd.row = 0; nodes.push_back(nd); n_leaves++; }
          for (uint32_t i = 0; i + 1 < k; i++) {
            pb200_node nd{}; nd.op = PB200_OP_ROW; nd.family = children[i].family; nd.row = t.G[i][idx[i]];
            nodes.push_back(nd); n_leaves++;
          }
          if (n_leaves > 1) { pb200_node nd{}; nd.op = PB200_OP_INTERSECT; nd.nargs = n_leaves; nodes.push_back(nd); }
        }
        prog_off.push_back(uint32_t(nodes.size()));
        Compiled cp;
        TRY(compile_batch(ctx, nodes.data(), prog_off.data(), Q, cp, true));
        BatchDev bd{};
        TRY(upload_batch(ctx, cp, live_shards.data(), S, bd));
        const size_t n_dense = size_t(Q) * S;
        TRY(ensure_dev(ctx, ctx->d_dense, n_dense * row_bytes));
        TRY(ensure_dev(ctx, ctx->d_dense_n, n_dense * kRowContainers * 4));
        TRY(launch_eval(ctx, bd, bd.progs, Q, S, true, nullptr, 0, 0, static_cast<uint8_t*>(ctx->d_dense.p),
                        static_cast<uint32_t*>(ctx->d_dense_n.p), max_leaves_of(cp)));
        if (!device_reduce) TRY(gb_dense_row_counts(ctx, uint32_t(n_dense), t.prefix_counts.data() + size_t(q0) * S));
        vshards.resize(n_dense);
        for (uint32_t qi = 0; qi < Q; qi++)
          for (uint32_t v = 0; v < S; v++) vshards[size_t(qi) * S + v] = live_shards[v];
        TRY(gb_count_against(ctx, children[k - 1].family, vshards.data(), uint32_t(n_dense), t.G[k - 1],
                             device_reduce ? nullptr : t.final_counts.data() + size_t(q0) * S * n_last, d_sums ? d_sums + size_t(q0) * n_last : nullptr, S));
      }
    }
    ctx->dense_shards_dev = nullptr;  // d_dense no longer holds a TopN source
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    std::vector<unsigned long long> h_sums;
    if (device_reduce) {
      h_sums.resize(size_t(n_groups_all));
      CUDA_TRY(ctx, cudaMemcpyAsync(h_sums.data(), d_sums, h_sums.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
      ctx->timing.d2h_bytes += h_sums.size() * 8;
    }
    TRY(finish_call(ctx));
    if (device_reduce) {
      for (uint64_t gi = 0; gi < n_groups_all; gi++) {  // key order = index order
        if (!h_sums[size_t(gi)]) continue;
        pb200_group g{};
        uint64_t rem = gi;
        for (int i = int(k) - 1; i >= 0; i--) { g.rows[i] = t.G[i][size_t(rem % t.G[i].size())]; rem /= t.G[i].size(); }
        g.count = h_sums[size_t(gi)];
        acc.push_back(g);
      }
    } else {

    // per-shard walks, then the reduce. mergeGroupCounts (:1166-1200) is a sorted union with summed counts, cut to `limit`:
    // what survives the reference's left fold over the shards are the `limit` smallest keys with the sums of everything
    // the shards' walks emitted for them (a surviving key is never cut on the way). The keys are positions in the
    // cross product of the sorted candidate lists, so the sums go into one dense array in index order; the pairwise
    // merges remain for cross products too large for that.
    uint64_t n_groups = 1;
    for (uint32_t i = 0; i < k; i++) n_groups *= t.G[i].size();
    const bool dense_reduce = n_groups <= (1ull << 22);
    std::vector<std::vector<pb200_group>> per(dense_reduce ? 0 : S);
    std::vector<std::vector<std::pair<uint64_t, uint64_t>>> per_idx(dense_reduce ? S : 0);  // (group index, count)
    parallel_shards(S, [&](uint32_t v) {
      GbWalk w(t, v);
      for (uint32_t i = 0; i < k; i++) {
        const auto& r = rows[v][i];
        w.lists[i].reserve(r.size());
        for (uint64_t id : r) w.lists[i].push_back(uint32_t(std::lower_bound(t.G[i].begin(), t.G[i].end(), id) - t.G[i].begin()));
      }
      w.init(children);
      pb200_group g;
      uint64_t gi = 0;
      if (dense_reduce) { while (per_idx[v].size() < limit && w.next_group(&g, &gi)) per_idx[v].push_back({gi, g.count}); }
      else { while (per[v].size() < limit && w.next_group(&g)) per[v].push_back(g); }  // executeGroupByShard :1262-1274
    });
    if (dense_reduce) {
      std::vector<uint64_t> sum(size_t(n_groups), 0);
      std::vector<uint8_t> seen(size_t(n_groups), 0);
      for (uint32_t v = 0; v < S; v++)
        for (auto& e : per_idx[v]) { sum[size_t(e.first)] += e.second; seen[size_t(e.first)] = 1; }
      for (uint64_t gi = 0; gi < n_groups && acc.size() < limit; gi++) {
        if (!seen[size_t(gi)]) continue;
        pb200_group g{};
        uint64_t rem = gi;
        for (int i = int(k) - 1; i >= 0; i--) { g.rows[i] = t.G[i][size_t(rem % t.G[i].size())]; rem /= t.G[i].size(); }
        g.count = sum[size_t(gi)];
        acc.push_back(g);
      }
    } else {
      for (uint32_t v = 0; v < S; v++) acc = merge_group_counts(std::move(acc), per[v], limit, k);
    }
    }
  } else {
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    TRY(finish_call(ctx));
  }
  *n_out = acc.size();
  if (acc.size() > cap) return fail(ctx, PB200_ERR_RANGE, "group buffer too small");
  if (!acc.empty() && !out) return PB200_ERR_INVALID;
  std::copy(acc.begin(), acc.end(), out);
  return PB200_OK;
}
