// Included by engine.cu: bitmap-returning calls, TopN, BSI aggregates.

namespace {

uint8_t* temp_slot(pb200_ctx* ctx, uint32_t slot, uint32_t n_shards) {
  return static_cast<uint8_t*>(ctx->d_temp.p) + size_t(slot) * n_shards * kRowContainers * kBitmapBytes;
}

template <class K>
int launch_range_pipe(pb200_ctx* ctx, int slot, K kernel, const BsiParams& p, uint32_t items) {
  constexpr int smem = (kPipeStages + 1) * kBitmapBytes;
  if (!(ctx->range_attr_set & (1u << slot))) {
    CUDA_TRY(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    ctx->range_attr_set |= 1u << slot;
  }
  if (ctx->occ_range[slot] == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kPipeThreads, smem) != cudaSuccess || n < 1) n = 1;
    ctx->occ_range[slot] = n;
  }
  const uint32_t grid = std::min<uint32_t>(items, uint32_t(ctx->sm_count) * uint32_t(ctx->occ_range[slot]));
  kernel<<<grid, kPipeThreads, smem, ctx->stream>>>(p);
  return PB200_OK;
}

int launch_bsi_range(pb200_ctx* ctx, const BatchDev& bd, const pb200_node& nd, uint32_t n_shards, uint8_t* dense_out,
                     unsigned long long* count_out = nullptr, uint32_t count_sstride = 0) {
  BsiParams p{};
  p.count_out = count_out; p.count_sstride = count_sstride;
  p.families = ctx->d_families; p.family = nd.family; p.bit_depth = nd.bit_depth;
  p.shards = bd.shards; p.n_shards = n_shards;
  p.bsi_op = nd.bsi_op; p.a = nd.a; p.b = nd.b;
  p.dense_out = dense_out;
  p.algo_bytes = algo_ptr(ctx);
  const uint32_t items = n_shards * kRowContainers;
  {
    const RangePlan plan = range_plan(p.bsi_op, p.a, p.b);
    int rc = PB200_ERR_INVALID;
#define PB200_RANGE_CASE(slot, K0, K1, CB) \
    if (plan.k0 == K0 && plan.k1 == K1 && plan.combine == CB) rc = launch_range_pipe(ctx, slot, bsi_range_kernel<K0, K1, CB>, p, items);
    PB200_RANGE_CASE(0, M_EQ, M_NONE, 0)
    PB200_RANGE_CASE(1, M_EQ, M_NONE, 1)
    PB200_RANGE_CASE(2, M_LTU, M_NONE, 2)
    PB200_RANGE_CASE(3, M_LTU, M_NONE, 3)
    PB200_RANGE_CASE(4, M_GTU, M_NONE, 0)
    PB200_RANGE_CASE(5, M_BETWEENU, M_NONE, 0)
    PB200_RANGE_CASE(6, M_LTU, M_LTU, 4)
    PB200_RANGE_CASE(7, M_NONE, M_NONE, 5)
#undef PB200_RANGE_CASE
    if (rc != PB200_OK) return rc == PB200_ERR_INVALID ? fail(ctx, rc, "unknown BSI range op") : rc;
  }
  ctx->timing.launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

bool has_shift(const Compiled& cp) {
  for (auto& pt : cp.pre) if (pt.kind == 1) return true;
  return false;
}

unsigned long long* ovf_slot(pb200_ctx* ctx, uint32_t slot, uint32_t n_shards) {
  return static_cast<unsigned long long*>(ctx->d_ovf.p) + size_t(slot) * n_shards;
}

// Overflow words (see shift_overflow_kernel) of `n_queries` programs: added to the counts and / or stored.
int launch_overflow(pb200_ctx* ctx, const BatchDev& bd, const Prog* progs, uint32_t n_queries, uint32_t n_shards,
                    unsigned long long* out_counts, uint32_t qstride, uint32_t sstride, unsigned long long* ovf_out) {
  OverflowParams op{};
  op.progs = progs; op.instrs = bd.instrs; op.leaves = bd.leaves; op.n_queries = n_queries; op.n_shards = n_shards;
  op.ovf = static_cast<const unsigned long long*>(ctx->d_ovf.p);
  op.out_counts = out_counts; op.out_qstride = qstride; op.out_sstride = sstride; op.ovf_out = ovf_out;
  const uint64_t n = uint64_t(n_queries) * n_shards;
  eval_overflow_kernel<<<uint32_t((n + 127) / 128), 128, 0, ctx->stream>>>(op);
  ctx->timing.launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

// K13: one chunk of a wide Union (plain Row leaves only) into a dense temp row.
int launch_union_rows(pb200_ctx* ctx, const BatchDev& bd, const Prog* prog, uint32_t n_shards, uint8_t* dense_out, bool xor_rows = false) {
  const uint64_t n_items = uint64_t(n_shards) * kRowContainers;
  if (n_items == 0) return PB200_OK;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  EvalParams p{};
  p.families = ctx->d_families;
  p.n_families = uint32_t(ctx->families.size());
  p.progs = prog; p.instrs = bd.instrs; p.leaves = bd.leaves; p.shards = bd.shards;
  p.n_shards = n_shards; p.n_queries = 1;
  p.temp_base = static_cast<const uint8_t*>(ctx->d_temp.p);
  p.dense_out = dense_out;
  p.algo_bytes = algo_ptr(ctx);
  const uint32_t grid = uint32_t(std::min<uint64_t>(n_items, uint64_t(ctx->sm_count) * 8));
  if (xor_rows) union_rows_kernel<true><<<grid, kUnionThreads, 0, ctx->stream>>>(p);
  else union_rows_kernel<false><<<grid, kUnionThreads, 0, ctx->stream>>>(p);
  ctx->timing.launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

int run_pretasks(pb200_ctx* ctx, const Compiled& cp, const BatchDev& bd, uint32_t n_shards) {
  if (has_shift(cp)) {  // overflow words of every temp slot (+1 for a result), zero unless a shift sets them
    const size_t bytes = size_t(cp.n_temp_slots + 1) * n_shards * 8;
    TRY(ensure_dev(ctx, ctx->d_ovf, bytes));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_ovf.p, 0, bytes, ctx->stream));
  }
  for (size_t i = 0; i < cp.pre.size(); i++) {
    const auto& pt = cp.pre[i];
    if (pt.kind == 0) {
      if (pt.count_only)  // Count(Row(v <op> x)): popcount in the comparator kernel, nothing materialised
        TRY(launch_bsi_range(ctx, bd, pt.node, n_shards, nullptr, pt.count_out, pt.count_sstride));
      else
        TRY(launch_bsi_range(ctx, bd, pt.node, n_shards, temp_slot(ctx, pt.slot, n_shards)));
    } else if (pt.kind == 2) {
      if (pt.union_rows) TRY(launch_union_rows(ctx, bd, bd.progs + cp.progs.size() + i, n_shards, temp_slot(ctx, pt.slot, n_shards), pt.xor_rows));
      else TRY(launch_eval(ctx, bd, bd.progs + cp.progs.size() + i, 1, n_shards, true, nullptr, 0, 0, temp_slot(ctx, pt.slot, n_shards), nullptr));
      if (has_shift(cp) && !pt.union_rows)  // plain rows carry no overflow: the slot's word stays zero
        TRY(launch_overflow(ctx, bd, bd.progs + cp.progs.size() + i, 1, n_shards, nullptr, 0, 0, ovf_slot(ctx, pt.slot, n_shards)));
    } else {
      const uint64_t n_words = uint64_t(n_shards) * kRowContainers * kBitmapWords;
      const uint32_t grid = uint32_t(std::min<uint64_t>((n_words + 255) / 256, uint64_t(ctx->sm_count) * 16));
      shift_rows_kernel<<<grid, 256, 0, ctx->stream>>>(temp_slot(ctx, pt.src_slot, n_shards), temp_slot(ctx, pt.slot, n_shards), n_shards);
      shift_overflow_kernel<<<(n_shards + 127) / 128, 128, 0, ctx->stream>>>(temp_slot(ctx, pt.src_slot, n_shards), ovf_slot(ctx, pt.src_slot, n_shards),
                                                                            ovf_slot(ctx, pt.slot, n_shards), n_shards);
      ctx->timing.launches += 2;
      CUDA_TRY(ctx, cudaGetLastError());
    }
  }
  return PB200_OK;
}

// Compile one program and evaluate it into dense rows [n_shards][16][1024 words] at
// ctx->d_dense with per-container cardinalities at ctx->d_dense_n. Leaves timing open:
// the caller records ev1.
int eval_dense(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, BatchDev& bd, bool allow_lists = false) {
  Compiled cp;
  Compiler cc(ctx, cp);
  if (!cc.compile(prog, n_nodes)) return fail(ctx, PB200_ERR_INVALID, cc.err);
  TRY(sync_families(ctx));
  TRY(upload_batch(ctx, cp, shards, n_shards, bd));
  const size_t items = size_t(n_shards) * kRowContainers;
  TRY(ensure_dev(ctx, ctx->d_dense, items * kBitmapBytes));
  TRY(ensure_dev(ctx, ctx->d_dense_n, items * 4));
  if (cp.n_temp_slots) TRY(ensure_dev(ctx, ctx->d_temp, size_t(cp.n_temp_slots) * items * kBitmapBytes));
  TRY(run_pretasks(ctx, cp, bd, n_shards));
  ctx->dense_has_ovf = has_shift(cp);
  if (ctx->dense_has_ovf) {
    ctx->dense_ovf_dev = ovf_slot(ctx, cp.n_temp_slots, n_shards);
    TRY(launch_overflow(ctx, bd, bd.progs, 1, n_shards, nullptr, 0, 0, ctx->dense_ovf_dev));
  }
  // K12 reads short result containers as value lists (64 B instead of 8 KiB each way); every other consumer of the dense rows gets words
  ctx->dense_lists = allow_lists && !ctx->dense_has_ovf;
  if (ctx->dense_lists) TRY(ensure_dev(ctx, ctx->d_sparse, items * kSparseListMax * 2));
  return launch_eval(ctx, bd, bd.progs, 1, n_shards, true, nullptr, 0, 0, static_cast<uint8_t*>(ctx->d_dense.p), static_cast<uint32_t*>(ctx->d_dense_n.p),
                     max_leaves_of(cp), cp.any_bushy, false, nullptr, cp.any_flat, tiny_of(cp), ctx->dense_lists ? static_cast<uint16_t*>(ctx->d_sparse.p) : nullptr);
}

int eval_one_shard(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, std::vector<uint64_t>& words, std::vector<uint32_t>& ns,
                   uint64_t* overflow) {
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  BatchDev bd{};
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  TRY(eval_dense(ctx, prog, n_nodes, &shard, 1, bd));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  words.resize(size_t(kRowContainers) * kBitmapWords);
  ns.resize(kRowContainers);
  CUDA_TRY(ctx, cudaMemcpyAsync(words.data(), ctx->d_dense.p, words.size() * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ns.data(), ctx->d_dense_n.p, ns.size() * 4, cudaMemcpyDeviceToHost, ctx->stream));
  *overflow = 0;
  if (ctx->dense_has_ovf) CUDA_TRY(ctx, cudaMemcpyAsync(overflow, ctx->dense_ovf_dev, 8, cudaMemcpyDeviceToHost, ctx->stream));
  TRY(finish_call(ctx));
  ctx->timing.d2h_bytes += words.size() * 8 + ns.size() * 4;
  return PB200_OK;
}

}  // namespace

namespace {

// K11: evaluate one program over the shards and serialise every shard's row on the device
// (kernels_rows.cuh). Leaves ctx->d_rows (packed streams) and ctx->d_rows_off
// ([n_shards + 1] u64 stream offsets) filled, in stream order.
// Serialises the dense rows currently in ctx->d_dense / d_dense_n (n_shards rows, shard numbers at `shards_dev`).
int emit_rows_from_dense(pb200_ctx* ctx, const uint64_t* shards_dev, uint32_t n_shards) {
  const uint32_t items = n_shards * kRowContainers;
  TRY(ensure_dev(ctx, ctx->d_rows_meta, size_t(items) * 4));
  TRY(ensure_dev(ctx, ctx->d_rows_item_off, size_t(items) * 8));
  TRY(ensure_dev(ctx, ctx->d_rows_off, size_t(n_shards + 1) * 8));
  // worst case: every container a bitmap, plus headers and a small overflow container per shard
  TRY(ensure_dev(ctx, ctx->d_rows, size_t(items) * kBitmapBytes + size_t(n_shards) * (8 + 17 * 16 + 256)));
  const uint64_t* dense = static_cast<const uint64_t*>(ctx->d_dense.p);
  const uint32_t* dense_n = static_cast<const uint32_t*>(ctx->d_dense_n.p);
  uint32_t* meta = static_cast<uint32_t*>(ctx->d_rows_meta.p);
  unsigned long long* item_off = static_cast<unsigned long long*>(ctx->d_rows_item_off.p);
  rows_classify_kernel<<<std::min<uint32_t>((items + 7) / 8, uint32_t(ctx->sm_count) * 8), 256, 0, ctx->stream>>>(dense, dense_n, items, meta);
  rows_layout_kernel<<<1, 1024, 0, ctx->stream>>>(meta, dense_n, shards_dev, ctx->dense_has_ovf ? ctx->dense_ovf_dev : nullptr, n_shards,
                                                  static_cast<uint8_t*>(ctx->d_rows.p), static_cast<unsigned long long*>(ctx->d_rows_off.p), item_off);
  rows_emit_kernel<<<std::min<uint32_t>(items, uint32_t(ctx->sm_count) * 8), kRowsEmitThreads, 0, ctx->stream>>>(dense, meta, item_off, items,
                                                                                                            static_cast<uint8_t*>(ctx->d_rows.p));
  ctx->timing.launches += 3;
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

int eval_rows_device(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards) {
  BatchDev bd{};
  TRY(eval_dense(ctx, prog, n_nodes, shards, n_shards, bd));
  return emit_rows_from_dense(ctx, bd.shards, n_shards);
}

// Second half of pb200_eval_shards: offsets, then the packed streams, device -> host.
int fetch_rows(pb200_ctx* ctx, uint32_t n_shards, void* out, size_t cap, uint64_t* out_off, size_t* out_len) {
  std::vector<uint64_t> local_off;
  uint64_t* off = out_off;
  if (!off) { local_off.resize(size_t(n_shards) + 1); off = local_off.data(); }
  CUDA_TRY(ctx, cudaMemcpyAsync(off, ctx->d_rows_off.p, size_t(n_shards + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const size_t total = size_t(off[n_shards]);
  *out_len = total;
  ctx->timing.d2h_bytes += size_t(n_shards + 1) * 8;
  if (out) {
    if (cap < total) { finish_call(ctx); return fail(ctx, PB200_ERR_RANGE, "output buffer too small"); }
    CUDA_TRY(ctx, cudaMemcpyAsync(out, ctx->d_rows.p, total, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->timing.d2h_bytes += total;
  }
  return finish_call(ctx);
}

int eval_shards_impl(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards,
                     void* out, size_t cap, uint64_t* out_off, size_t* out_len) {
  if (!out_len) return PB200_ERR_INVALID;
  TRY(validate_shards(ctx, shards, n_shards));
  *out_len = 0;
  if (n_shards == 0) { if (out_off) out_off[0] = 0; return PB200_OK; }
  if (n_shards > (1u << 24)) return fail(ctx, PB200_ERR_RANGE, "too many shards in one call");
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  TRY(eval_rows_device(ctx, prog, n_nodes, shards, n_shards));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  return fetch_rows(ctx, n_shards, out, cap, out_off, out_len);
}

}  // namespace

extern "C" int pb200_eval(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, void* out, size_t cap, size_t* out_len) {
  ENTER(ctx);
  return eval_shards_impl(ctx, prog, n_nodes, &shard, 1, out, cap, nullptr, out_len);
}

extern "C" int pb200_eval_shards(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards,
                                 void* out, size_t cap, uint64_t* out_off, size_t* out_len) {
  ENTER(ctx);
  return eval_shards_impl(ctx, prog, n_nodes, shards, n_shards, out, cap, out_off, out_len);
}

namespace {

// K12 over the dense rows currently in ctx->d_dense / d_dense_n: size, lay out and (when `out` is given) emit the
// Columns field; the caller has recorded ev0.
int emit_proto_from_dense(pb200_ctx* ctx, const uint64_t* shards_dev, uint32_t n_shards, void* out, size_t cap, size_t* out_len) {
  const uint32_t n_slots = n_shards * kProtoSlots;
  TRY(ensure_dev(ctx, ctx->d_rows_meta, size_t(n_slots) * 4));
  TRY(ensure_dev(ctx, ctx->d_rows_item_off, size_t(n_slots) * 8));
  const uint64_t* dense = static_cast<const uint64_t*>(ctx->d_dense.p);
  const uint32_t* dense_n = static_cast<const uint32_t*>(ctx->d_dense_n.p);
  const unsigned long long* ovf = ctx->dense_has_ovf ? ctx->dense_ovf_dev : nullptr;
  uint32_t* sizes = static_cast<uint32_t*>(ctx->d_rows_meta.p);
  unsigned long long* slot_off = static_cast<unsigned long long*>(ctx->d_rows_item_off.p);
  unsigned long long* totals = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_misc.p) + 16);
  const uint16_t* sparse = ctx->dense_lists ? static_cast<const uint16_t*>(ctx->d_sparse.p) : nullptr;  // (no list flags are set without it)
  proto_size_kernel<<<std::min<uint32_t>((n_slots + 7) / 8, uint32_t(ctx->sm_count) * 8), kProtoThreads, 0, ctx->stream>>>(dense, dense_n, sparse, shards_dev, ovf,
                                                                                                                   n_shards, sizes);
  proto_layout_kernel<<<1, 1024, 0, ctx->stream>>>(sizes, n_slots, slot_off, totals);
  ctx->timing.launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  unsigned long long h_tot[2] = {0, 0};
  CUDA_TRY(ctx, cudaMemcpyAsync(h_tot, totals, 16, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->timing.d2h_bytes += 16;
  const size_t total = size_t(h_tot[0]);
  *out_len = total;
  if (!out || total == 0) return finish_call(ctx);
  if (cap < total) { finish_call(ctx); return fail(ctx, PB200_ERR_RANGE, "output buffer too small"); }
  TRY(ensure_dev(ctx, ctx->d_rows, total));
  proto_emit_kernel<<<std::min<uint32_t>(n_slots, uint32_t(ctx->sm_count) * 8), kProtoThreads, 0, ctx->stream>>>(
      dense, dense_n, sparse, sizes, slot_off, totals, shards_dev, ovf, n_shards, static_cast<uint8_t*>(ctx->d_rows.p));
  ctx->timing.launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(out, ctx->d_rows.p, total, cudaMemcpyDeviceToHost, ctx->stream));
  ctx->timing.d2h_bytes += total;
  return finish_call(ctx);
}

}  // namespace

extern "C" int pb200_eval_proto_columns(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards,
                                        void* out, size_t cap, size_t* out_len) {
  ENTER(ctx);
  if (!out_len) return PB200_ERR_INVALID;
  TRY(validate_shards(ctx, shards, n_shards));
  *out_len = 0;
  if (n_shards == 0) return PB200_OK;
  if (n_shards > (1u << 24)) return fail(ctx, PB200_ERR_RANGE, "too many shards in one call");
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  BatchDev bd{};
  TRY(eval_dense(ctx, prog, n_nodes, shards, n_shards, bd, true));
  return emit_proto_from_dense(ctx, bd.shards, n_shards, out, cap, out_len);
}

extern "C" int pb200_eval_columns(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, uint64_t* out_cols, size_t cap, size_t* out_n) {
  ENTER(ctx);
  if (!out_n) return PB200_ERR_INVALID;
  std::vector<uint64_t> words;
  std::vector<uint32_t> ns;
  uint64_t overflow = 0;
  TRY(eval_one_shard(ctx, prog, n_nodes, shard, words, ns, &overflow));
  size_t total = size_t(__builtin_popcountll(overflow));
  for (uint32_t n : ns) total += n;
  *out_n = total;
  if (!out_cols) return PB200_OK;
  if (cap < total) return fail(ctx, PB200_ERR_RANGE, "output buffer too small");
  size_t k = 0;
  const uint64_t base = shard * PB200_SHARD_WIDTH;
  for (size_t w = 0; w < words.size(); w++) {
    uint64_t v = words[w];
    while (v) { out_cols[k++] = base + w * 64 + __builtin_ctzll(v); v &= v - 1; }
  }
  for (uint64_t v = overflow; v; v &= v - 1) out_cols[k++] = base + PB200_SHARD_WIDTH + __builtin_ctzll(v);
  return PB200_OK;
}

// ---------------------------------------------------------------------------------------
// TopN
// ---------------------------------------------------------------------------------------
namespace {

struct HPair { uint64_t id, count; };

// Go's container/heap over pairHeap (cache.go:331-353: min-heap on Count). The order in
// which fragment.top pops its results depends on these exact sift rules.
struct PairHeap {
  std::vector<HPair> a;
  bool less(int i, int j) const { return a[i].count < a[j].count; }
  void up(int j) { for (;;) { int i = (j - 1) / 2; if (i == j || !less(j, i)) break; std::swap(a[i], a[j]); j = i; } }
  void down(int i0, int n) {
    int i = i0;
    for (;;) {
      int j1 = 2 * i + 1;
      if (j1 >= n || j1 < 0) break;
      int j = j1;
      if (int j2 = j1 + 1; j2 < n && less(j2, j1)) j = j2;
      if (!less(j, i)) break;
      std::swap(a[i], a[j]);
      i = j;
    }
  }
  void push(HPair p) { a.push_back(p); up(int(a.size()) - 1); }
  HPair pop() { int n = int(a.size()) - 1; std::swap(a[0], a[n]); down(0, n); HPair p = a.back(); a.pop_back(); return p; }
};

void sort_pairs(std::vector<HPair>& p) {  // count desc; ties by ascending id (deterministic refinement of sort.Sort)
  std::stable_sort(p.begin(), p.end(), [](const HPair& x, const HPair& y) { return x.count != y.count ? x.count > y.count : x.id < y.id; });
}

// Host-side shard loop on a few threads (the reference runs its per-shard jobs on a worker
// pool of NumCPU goroutines, executor.go:83-107; the selection walks are independent per shard).
// persistent helper threads of the host-side shard loops (started once per process; a TopN call makes three of these
// loops and thread start-up was a visible part of it)
std::mutex& host_pool_mutex() { static std::mutex mu; return mu; }
std::vector<std::unique_ptr<DevWorker>>& host_pool() { static std::vector<std::unique_ptr<DevWorker>> pool; return pool; }

template <class F>
void parallel_shards(uint32_t n, F&& fn) {
  std::mutex& pool_mu = host_pool_mutex();
  std::vector<std::unique_ptr<DevWorker>>& pool = host_pool();
  unsigned hw = std::thread::hardware_concurrency();
  const unsigned nt = std::min<unsigned>(hw ? hw : 4, 12);
  if (n < 256 || nt < 2) { for (uint32_t s = 0; s < n; s++) fn(s); return; }
  std::unique_lock<std::mutex> lk(pool_mu);  // concurrent big loops (the devices of a pb200_multi) take turns on the helpers

  if (pool.empty()) for (unsigned t = 0; t + 1 < nt; t++) pool.emplace_back(new DevWorker());
  std::atomic<uint32_t> next{0};
  auto body = [&]() -> int {
    for (;;) {
      const uint32_t b = next.fetch_add(16);
      if (b >= n) break;
      for (uint32_t s = b; s < std::min(n, b + 16); s++) fn(s);
    }
    return 0;
  };
  for (auto& w : pool) w->post(body);
  body();
  for (auto& w : pool) w->wait();
}

// fragment.top (fragment.go:1568-1700) over candidates (ids[k], cached counts[k]) whose
// src-intersection counts are already known (`isect[k]`, ignored when !has_src).
std::vector<HPair> top_walk(const uint64_t* ids, const uint64_t* counts, size_t n_pairs, const uint32_t* isect, bool has_src, int n,
                            uint64_t min_threshold, uint64_t tanimoto, uint64_t src_count) {
  double min_tan = 0, max_tan = 0;
  if (tanimoto > 0 && has_src) {
    min_tan = double(src_count * tanimoto) / 100;
    max_tan = double(src_count * 100) / double(tanimoto);
  } else {
    tanimoto = 0;
  }
  PairHeap results;
  for (size_t k = 0; k < n_pairs; k++) {
    const uint64_t row_id = ids[k], cnt = counts[k];
    if (cnt == 0) continue;
    if (tanimoto > 0) { if (double(cnt) <= min_tan || double(cnt) >= max_tan) continue; }
    else if (cnt < min_threshold) continue;
    if (n == 0 || int(results.a.size()) < n) {
      uint64_t count = has_src ? isect[k] : cnt;
      if (count == 0) continue;
      if (tanimoto > 0) {
        double t = std::ceil(double(count * 100) / double(cnt + src_count - count));
        if (t <= double(tanimoto)) continue;
      } else if (count < min_threshold) continue;
      results.push({row_id, count});
      if (n > 0 && int(results.a.size()) == n && !has_src) break;
      continue;
    }
    const uint64_t threshold = results.a[0].count;
    if (threshold < min_threshold || cnt < threshold) break;
    const uint64_t count = isect[k];
    if (count < threshold) continue;
    results.push({row_id, count});
  }
  std::vector<HPair> r(results.a.size());
  size_t x = r.size(), i = 1;
  while (!results.a.empty()) { r[x - i] = results.pop(); i++; }
  return r;
}

uint64_t exact_row_count(const Fragment& f, uint64_t id) {
  auto it = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), id);
  if (it == f.row_ids.end() || *it != id) return 0;
  return f.row_counts[size_t(it - f.row_ids.begin())];
}

// Device copy of a fragment's rank order as row-table indices; built when the rank list
// changes (load / pb200_rank_set), reused by every TopN call.
int ensure_rank_table(pb200_ctx* ctx, Fragment& f) {
  if (f.d_rank_valid) return PB200_OK;
  if (f.d_rank) { CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream)); CUDA_TRY(ctx, cudaFree(f.d_rank)); f.d_rank = nullptr; f.d_rank_counts = nullptr; }
  const size_t n = f.rank_ids.size();
  if (n) {
    const size_t o_cnt = align_up(n * 4, 16);
    std::vector<uint8_t> h(o_cnt + n * 8);
    uint32_t* idx = reinterpret_cast<uint32_t*>(h.data());
    for (size_t k = 0; k < n; k++) {
      const uint64_t id = f.rank_ids[k];
      if (f.dev.dense) idx[k] = id < f.dev.nrows ? uint32_t(id) : 0xffffffffu;
      else {
        auto it = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), id);
        idx[k] = (it != f.row_ids.end() && *it == id) ? uint32_t(it - f.row_ids.begin()) : 0xffffffffu;
      }
    }
    std::memcpy(h.data() + o_cnt, f.rank_counts.data(), n * 8);
    void* d = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&d, h.size()));
    CUDA_TRY(ctx, cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice));
    f.d_rank = static_cast<uint32_t*>(d);
    f.d_rank_counts = reinterpret_cast<uint64_t*>(static_cast<uint8_t*>(d) + o_cnt);
  }
  f.d_rank_valid = true;
  return PB200_OK;
}

// One TopN pass over `shards` (executor.go:905-930 executeTopNShards without the final sort):
// per-shard fragment.top results.
// What pass 1 of a two-pass TopN leaves behind for pass 2 of the same call: the exact intersection count of
// EVERY ranked row of every shard is already on the host, so pass 2 (the same counts for the merged candidate
// ids, executor.go:889-902) is a lookup instead of a second trip over the head rows.
struct TopnCarry {
  bool valid = false;
  std::vector<uint64_t> off;         // shard s owns isect[off[s] .. ) in rank order ...
  std::vector<uint32_t> limit;       // ... for its first limit[s] rank positions (pruned pass 1 counts no further)
  const uint32_t* isect = nullptr;   // pinned host buffer of pass 1 (ctx->h_out)
  std::vector<uint64_t> src_counts;  // |source row| per shard (Tanimoto)
};

void ensure_rank_pos(Fragment& f) {
  if (f.rank_pos_valid) return;
  f.rank_pos_of_row.assign(f.row_ids.size(), 0xffffffffu);
  for (size_t p = 0; p < f.rank_ids.size(); p++) {
    auto it = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), f.rank_ids[p]);
    if (it != f.row_ids.end() && *it == f.rank_ids[p]) f.rank_pos_of_row[size_t(it - f.row_ids.begin())] = uint32_t(p);
  }
  f.rank_pos_valid = true;
}

// Pass 2 from pass 1's counts. Returns false (nothing written) when some candidate row exists in a shard
// but is not in that shard's rank table: the regular device pass handles that case.
bool topn_pass2_from_carry(pb200_ctx* ctx, pb200_family family, const uint64_t* ids, size_t n_ids, uint64_t min_threshold,
                           uint64_t tanimoto, const uint64_t* shards, uint32_t n_shards, const TopnCarry& carry,
                           std::vector<std::vector<HPair>>& per_shard) {
  Family& fam = ctx->families[family];
  std::vector<Fragment*> frs(n_shards, nullptr), todo;
  for (uint32_t s = 0; s < n_shards; s++) {
    auto it = fam.frags.find(shards[s]);
    if (it == fam.frags.end()) continue;
    frs[s] = &it->second;
    if (!it->second.rank_pos_valid) todo.push_back(&it->second);
  }
  std::sort(todo.begin(), todo.end());
  todo.erase(std::unique(todo.begin(), todo.end()), todo.end());
  parallel_shards(uint32_t(todo.size()), [&](uint32_t i) { ensure_rank_pos(*todo[i]); });
  std::atomic<bool> miss{false};
  std::vector<std::vector<HPair>> out(n_shards);
  parallel_shards(n_shards, [&](uint32_t s) {
    if (!frs[s] || miss.load(std::memory_order_relaxed)) return;
    const Fragment& f = *frs[s];
    struct Cand { uint64_t id, count; uint32_t isect; };
    std::vector<Cand> cd;
    for (size_t k = 0; k < n_ids; k++) {
      const auto it = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), ids[k]);
      if (it == f.row_ids.end() || *it != ids[k]) continue;  // the shard does not have this row
      const size_t j = size_t(it - f.row_ids.begin());
      const uint32_t pos = f.rank_pos_of_row[j];
      if (pos == 0xffffffffu || pos >= carry.limit[s]) { miss.store(true, std::memory_order_relaxed); return; }
      uint64_t c = f.rank_from_host ? f.rank_counts[pos] : f.row_counts[j];
      if (c == 0) c = f.row_counts[j];
      if (c > 0) cd.push_back({ids[k], c, carry.isect[carry.off[s] + pos]});
    }
    std::stable_sort(cd.begin(), cd.end(), [](const Cand& x, const Cand& y) { return x.count != y.count ? x.count > y.count : x.id < y.id; });
    std::vector<uint64_t> cid(cd.size()), cc(cd.size());
    std::vector<uint32_t> ci(cd.size());
    for (size_t k = 0; k < cd.size(); k++) { cid[k] = cd[k].id; cc[k] = cd[k].count; ci[k] = cd[k].isect; }
    out[s] = top_walk(cid.data(), cc.data(), cid.size(), ci.data(), true, 0, min_threshold, tanimoto, carry.src_counts[s]);
  });
  if (miss.load()) return false;
  per_shard.swap(out);
  return true;
}

int topn_pass(pb200_ctx* ctx, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes, uint32_t n,
              const uint64_t* ids, size_t n_ids, uint64_t min_threshold, uint64_t tanimoto, const uint64_t* shards,
              uint32_t n_shards, std::vector<std::vector<HPair>>& per_shard, bool reuse_src = false, TopnCarry* carry = nullptr) {
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (tanimoto > 100) return fail(ctx, PB200_ERR_INVALID, "Tanimoto Threshold is from 1 to 100 only");  // executor.go:988
  if (min_threshold == 0) min_threshold = 1;  // defaultMinThreshold, executor.go:37,984
  Family& fam = ctx->families[family];
  const bool has_src = src_prog && n_src_nodes;
  per_shard.assign(n_shards, {});
  if (n_shards == 0) return PB200_OK;
  static const bool trace = getenv("PB200_TRACE") != nullptr;
  auto t_start = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!trace) return;
    auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[pb200 topn] %-18s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_start).count());
    t_start = now;
  };

  std::vector<Fragment*> frs(n_shards, nullptr);
  for (uint32_t s = 0; s < n_shards; s++) {
    auto it = fam.frags.find(shards[s]);
    if (it != fam.frags.end()) frs[s] = &it->second;
  }
  // explicit ids: candidate pairs per shard (topBitmapPairs, fragment.go:1702-1738), sorted by count
  std::vector<std::vector<uint64_t>> id_ids, id_counts;
  std::vector<std::vector<uint32_t>> id_pos;  // position of each sorted candidate in the caller's `ids`
  if (n_ids) {
    id_ids.resize(n_shards); id_counts.resize(n_shards); id_pos.resize(n_shards);
    parallel_shards(n_shards, [&](uint32_t s) {
      if (!frs[s]) return;
      const Fragment& f = *frs[s];
      std::unordered_map<uint64_t, uint64_t> cache;
      if (f.rank_from_host) for (size_t k = 0; k < f.rank_ids.size(); k++) cache[f.rank_ids[k]] = f.rank_counts[k];
      std::vector<std::pair<HPair, uint32_t>> cd;
      for (size_t k = 0; k < n_ids; k++) {
        uint64_t c = 0;
        if (f.rank_from_host) { auto ci = cache.find(ids[k]); if (ci != cache.end()) c = ci->second; }
        if (c == 0) c = exact_row_count(f, ids[k]);
        if (c > 0) cd.push_back({{ids[k], c}, uint32_t(k)});
      }
      std::stable_sort(cd.begin(), cd.end(), [](const auto& x, const auto& y) {
        return x.first.count != y.first.count ? x.first.count > y.first.count : x.first.id < y.first.id;
      });
      for (auto& e : cd) { id_ids[s].push_back(e.first.id); id_counts[s].push_back(e.first.count); id_pos[s].push_back(e.second); }
    });
  }

  lap("candidates");
  std::vector<uint64_t> off(n_shards + 1, 0);
  std::vector<uint32_t> visible;  // with a source: how many rank positions of each shard have a count on the host
  std::vector<uint64_t> src_counts(n_shards, 0);
  const uint32_t* isect = nullptr;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  if (has_src) {
    for (uint32_t s = 0; s < n_shards; s++) {
      if (frs[s] && !n_ids) TRY(ensure_rank_table(ctx, *frs[s]));
      off[s + 1] = off[s] + ((frs[s] && !n_ids) ? frs[s]->rank_ids.size() : 0);
    }
    BatchDev bd{};
    CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if (reuse_src && ctx->dense_shards_dev) {
      bd.shards = ctx->dense_shards_dev;  // pass 2: the source rows of pass 1 are still resident in d_dense
    } else {
      TRY(eval_dense(ctx, src_prog, n_src_nodes, shards, n_shards, bd));
      ctx->dense_shards_dev = bd.shards;
    }
    if (trace) { cudaStreamSynchronize(ctx->stream); lap("  rank tables + src"); }
    const uint64_t total = off[n_shards];
    const size_t n_counts = n_ids ? size_t(n_shards) * n_ids : size_t(total);
    // Pruned pass (the reference's own rule, fragment.go:1673-1688): count the head of every shard's ranking, let the
    // device replay the start of the walk to find where it stops, count only up to there, fetch only those counts.
    const bool prune = ctx->topn_prune && !n_ids && n > 0 && tanimoto == 0;
    // d_topn: [cand_off u64 x ns][out_off u64 x ns][rank_n u32 x ns][limit u32 x ns][idx ptrs x ns][count ptrs x ns][ids u64 x n_ids] ; counts in d_bsi
    size_t o_off = 0;
    size_t o_out = o_off + size_t(n_shards) * 8;
    size_t o_n = o_out + size_t(n_shards) * 8;
    size_t o_lim = o_n + size_t(n_shards) * 4;
    size_t o_ptr = align_up(o_lim + size_t(n_shards) * 4, 16);
    size_t o_cptr = o_ptr + size_t(n_shards) * 8;
    size_t o_ids = o_cptr + size_t(n_shards) * 8;
    size_t total_bytes = o_ids + n_ids * 8;
    TRY(ensure_dev(ctx, ctx->d_topn, total_bytes));
    TRY(ensure_pin(ctx, ctx->h_topn, total_bytes));  // not h_stage: the program upload issued above may still be reading it
    TRY(ensure_dev(ctx, ctx->d_bsi, std::max<size_t>(n_counts, 1) * 4));
    uint8_t* h = static_cast<uint8_t*>(ctx->h_topn.p);
    uint8_t* d = static_cast<uint8_t*>(ctx->d_topn.p);
    std::memcpy(h + o_off, off.data(), size_t(n_shards) * 8);
    auto* hn = reinterpret_cast<uint32_t*>(h + o_n);
    auto* hp = reinterpret_cast<const uint32_t**>(h + o_ptr);
    auto* hcp = reinterpret_cast<const uint64_t**>(h + o_cptr);
    uint32_t max_cand = uint32_t(n_ids);
    for (uint32_t s = 0; s < n_shards; s++) {
      const bool use = frs[s] && !n_ids;
      hn[s] = use ? uint32_t(frs[s]->rank_ids.size()) : 0;
      hp[s] = use ? frs[s]->d_rank : nullptr;
      hcp[s] = use ? frs[s]->d_rank_counts : nullptr;
      max_cand = std::max(max_cand, hn[s]);
    }
    if (n_ids) std::memcpy(h + o_ids, ids, n_ids * 8);
    CUDA_TRY(ctx, cudaMemcpyAsync(d, h, o_out, cudaMemcpyHostToDevice, ctx->stream));  // cand_off
    CUDA_TRY(ctx, cudaMemcpyAsync(d + o_n, h + o_n, total_bytes - o_n, cudaMemcpyHostToDevice, ctx->stream));
    ctx->timing.h2d_bytes += total_bytes;
    uint32_t* d_counts = static_cast<uint32_t*>(ctx->d_bsi.p);
    uint32_t* d_limit = reinterpret_cast<uint32_t*>(d + o_lim);
    TopnParams tp{};
    tp.families = ctx->d_families; tp.family = family; tp.shards = bd.shards; tp.n_shards = n_shards;
    tp.src_dense = static_cast<const uint8_t*>(ctx->d_dense.p);
    tp.src_n = static_cast<const uint32_t*>(ctx->d_dense_n.p);
    tp.rank_rowidx = reinterpret_cast<const uint32_t* const*>(d + o_ptr);
    tp.rank_n = reinterpret_cast<const uint32_t*>(d + o_n);
    tp.cand_off = reinterpret_cast<const uint64_t*>(d + o_off);
    tp.ids = n_ids ? reinterpret_cast<const uint64_t*>(d + o_ids) : nullptr;
    tp.n_ids = uint32_t(n_ids);
    tp.chunk = 2048;
    tp.counts = d_counts;
    tp.algo_bytes = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_misc.p) + 8);
    auto launch_counts = [&](uint32_t span) -> int {  // one pass of the count kernel over `span` candidates per shard at most
      if (!ctx->topn_attr_set) {
        CUDA_TRY(ctx, cudaFuncSetAttribute(topn_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTopnSmem));
        ctx->topn_attr_set = true;
      }
      tp.chunks_per_shard = std::max<uint32_t>(1, (span + tp.chunk - 1) / tp.chunk);
      const uint64_t items = uint64_t(n_shards) * tp.chunks_per_shard;
      const uint32_t grid = uint32_t(std::min<uint64_t>(items, resident_ctas(ctx, OCC_TOPN, topn_count_kernel, kTopnThreads, kTopnSmem)));
      topn_count_kernel<<<grid, kTopnThreads, kTopnSmem, ctx->stream>>>(tp);
      ctx->timing.launches++;
      CUDA_TRY(ctx, cudaGetLastError());
      return PB200_OK;
    };
    std::vector<uint32_t> limit(n_shards, 0);   // rank positions of shard s whose counts reach the host: [0, limit[s])
    std::vector<uint64_t> host_off = off;       // where they start in the host buffer
    size_t n_fetch = n_counts;
    const uint32_t* d_fetch = d_counts;
    if (n_counts && prune) {
      const uint32_t head = std::min<uint32_t>(max_cand, std::max<uint32_t>(4 * n, 64));
      tp.range_lo = 0; tp.range_cap = head; tp.range_hi = nullptr;
      TRY(launch_counts(head));
      topn_limit_kernel<<<(n_shards * 32 + 255) / 256, 256, 0, ctx->stream>>>(reinterpret_cast<const uint64_t* const*>(d + o_cptr), tp.rank_n, tp.cand_off, d_counts,
                                                                           n_shards, head, n, min_threshold, d_limit);
      ctx->timing.launches++;
      CUDA_TRY(ctx, cudaGetLastError());
      if (max_cand > head) {
        tp.range_lo = head; tp.range_cap = 0; tp.range_hi = d_limit;
        TRY(launch_counts(max_cand - head));
      }
      CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
      CUDA_TRY(ctx, cudaMemcpyAsync(limit.data(), d_limit, size_t(n_shards) * 4, cudaMemcpyDeviceToHost, ctx->stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
      uint64_t acc = 0;
      for (uint32_t s = 0; s < n_shards; s++) { host_off[s] = acc; acc += limit[s]; }
      host_off[n_shards] = acc;
      n_fetch = size_t(acc);
      std::memcpy(h + o_out, host_off.data(), size_t(n_shards) * 8);
      CUDA_TRY(ctx, cudaMemcpyAsync(d + o_out, h + o_out, size_t(n_shards) * 8, cudaMemcpyHostToDevice, ctx->stream));
      TRY(ensure_dev(ctx, ctx->d_topn_compact, std::max<size_t>(n_fetch, 1) * 4));
      topn_compact_kernel<<<std::min<uint32_t>(n_shards, uint32_t(ctx->sm_count) * 8), 128, 0, ctx->stream>>>(
          d_counts, tp.cand_off, d_limit, reinterpret_cast<const uint64_t*>(d + o_out), n_shards, static_cast<uint32_t*>(ctx->d_topn_compact.p));
      ctx->timing.launches++;
      CUDA_TRY(ctx, cudaGetLastError());
      d_fetch = static_cast<const uint32_t*>(ctx->d_topn_compact.p);
      ctx->timing.d2h_bytes += size_t(n_shards) * 4;
    } else {
      if (n_counts) { tp.range_lo = 0; tp.range_cap = 0; tp.range_hi = nullptr; TRY(launch_counts(max_cand)); }
      CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
      for (uint32_t s = 0; s < n_shards; s++) limit[s] = uint32_t(off[s + 1] - off[s]);
    }
    if (trace) { cudaStreamSynchronize(ctx->stream); lap("  count kernels"); }
    TRY(ensure_pin(ctx, ctx->h_out, std::max<size_t>(n_fetch, 1) * 4 + size_t(n_shards) * kRowContainers * 4));
    uint32_t* h_counts = static_cast<uint32_t*>(ctx->h_out.p);
    uint32_t* h_dn = h_counts + std::max<size_t>(n_fetch, 1);
    if (n_fetch) CUDA_TRY(ctx, cudaMemcpyAsync(h_counts, d_fetch, n_fetch * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(h_dn, ctx->d_dense_n.p, size_t(n_shards) * kRowContainers * 4, cudaMemcpyDeviceToHost, ctx->stream));
    TRY(finish_call(ctx));
    ctx->timing.d2h_bytes += n_fetch * 4 + size_t(n_shards) * kRowContainers * 4;
    for (uint32_t s = 0; s < n_shards; s++)
      for (int c = 0; c < kRowContainers; c++) src_counts[s] += h_dn[size_t(s) * kRowContainers + c];
    isect = h_counts;
    off = host_off;
    visible = limit;
    lap("device + copies");
    if (carry && !n_ids) { carry->valid = true; carry->off = off; carry->limit = limit; carry->isect = h_counts; carry->src_counts = src_counts; }
  }

  // per-shard selection walk (host: sequential, data-dependent, a few thousand steps per shard)
  parallel_shards(n_shards, [&](uint32_t s) {
    if (!frs[s]) return;
    const Fragment& f = *frs[s];
    if (n_ids) {
      std::vector<uint32_t> tmp;
      const uint32_t* is = nullptr;
      if (has_src) {  // device counts are laid out in `ids` order; candidates were re-sorted by count
        tmp.resize(id_ids[s].size());
        for (size_t k = 0; k < tmp.size(); k++) tmp[k] = isect[size_t(s) * n_ids + id_pos[s][k]];
        is = tmp.data();
      }
      per_shard[s] = top_walk(id_ids[s].data(), id_counts[s].data(), id_ids[s].size(), is, has_src, 0, min_threshold, tanimoto, src_counts[s]);
    } else {
      // a pruned pass holds counts for the prefix the walk visits; the walk ends there by the same rule the device applied
      per_shard[s] = top_walk(f.rank_ids.data(), f.rank_counts.data(), has_src ? size_t(visible[s]) : f.rank_ids.size(), has_src ? isect + off[s] : nullptr, has_src,
                              int(n), min_threshold, tanimoto, src_counts[s]);
    }
  });
  lap("selection walk");
  return PB200_OK;
}

std::vector<HPair> merge_pairs(const std::vector<std::vector<HPair>>& per_shard) {  // Pairs.Add (cache.go:356-374) + sort
  // the shards mostly report the same few ids: sum them in a hash map, then sort only the distinct ones
  std::unordered_map<uint64_t, uint64_t> sum;
  sum.reserve(1024);
  for (auto& v : per_shard)
    for (auto& p : v) sum[p.id] += p.count;
  std::vector<HPair> out;
  out.reserve(sum.size());
  for (auto& kv : sum) out.push_back({kv.first, kv.second});
  sort_pairs(out);
  return out;
}

int write_pairs(pb200_ctx* ctx, const std::vector<HPair>& v, pb200_pair* out, size_t cap, size_t* n_out) {
  if (!n_out) return PB200_ERR_INVALID;
  *n_out = v.size();
  if (v.size() > cap) return fail(ctx, PB200_ERR_RANGE, "pair buffer too small");
  for (size_t i = 0; i < v.size(); i++) { out[i].id = v[i].id; out[i].count = v[i].count; }
  return PB200_OK;
}

}  // namespace

extern "C" int pb200_topn_shard(pb200_ctx* ctx, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes, uint32_t n,
                                const uint64_t* ids, size_t n_ids, uint64_t min_threshold, uint64_t tanimoto_threshold,
                                uint64_t shard, pb200_pair* out, size_t cap, size_t* n_out) {
  ENTER(ctx);
  if (n_ids && !ids) return PB200_ERR_INVALID;
  std::vector<std::vector<HPair>> per;
  TRY(topn_pass(ctx, family, src_prog, n_src_nodes, n, ids, n_ids, min_threshold, tanimoto_threshold, &shard, 1, per));
  return write_pairs(ctx, per[0], out, cap, n_out);
}

extern "C" int pb200_topn(pb200_ctx* ctx, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes, uint32_t n,
                          const uint64_t* ids, size_t n_ids, uint64_t min_threshold, uint64_t tanimoto_threshold, uint32_t flags,
                          const uint64_t* shards, uint32_t n_shards, pb200_pair* out, size_t cap, size_t* n_out) {
  ENTER(ctx);
  if (n_ids && !ids) return PB200_ERR_INVALID;
  TRY(validate_shards(ctx, shards, n_shards));
  std::vector<std::vector<HPair>> per;
  TopnCarry carry;
  TRY(topn_pass(ctx, family, src_prog, n_src_nodes, n, ids, n_ids, min_threshold, tanimoto_threshold, shards, n_shards, per, false, &carry));
  std::vector<HPair> pairs = merge_pairs(per);
  if (pairs.empty() || n_ids > 0 || (flags & PB200_TOPN_REMOTE)) return write_pairs(ctx, pairs, out, cap, n_out);  // executor.go:884-887
  // second pass with the candidate ids of pass 1 (executor.go:889-902)
  pb200_timing t1 = ctx->timing;
  std::vector<uint64_t> ids2;
  for (auto& p : pairs) ids2.push_back(p.id);
  std::sort(ids2.begin(), ids2.end());
  const uint64_t thr = min_threshold ? min_threshold : 1;  // defaultMinThreshold, as topn_pass applies it
  if (!(carry.valid && topn_pass2_from_carry(ctx, family, ids2.data(), ids2.size(), thr, tanimoto_threshold, shards, n_shards, carry, per))) {
    TRY(topn_pass(ctx, family, src_prog, n_src_nodes, n, ids2.data(), ids2.size(), min_threshold, tanimoto_threshold, shards, n_shards, per,
                  /*reuse_src=*/true));
    ctx->timing.kernel_ms += t1.kernel_ms; ctx->timing.algo_bytes += t1.algo_bytes; ctx->timing.launches += t1.launches;
    ctx->timing.h2d_bytes += t1.h2d_bytes; ctx->timing.d2h_bytes += t1.d2h_bytes;
  }
  ctx->dense_shards_dev = nullptr;
  std::vector<HPair> trimmed = merge_pairs(per);
  if (n != 0 && n < trimmed.size()) trimmed.resize(n);
  return write_pairs(ctx, trimmed, out, cap, n_out);
}

// ---------------------------------------------------------------------------------------
// BSI aggregates
// ---------------------------------------------------------------------------------------
namespace {

// Prepares the optional filter of a BSI aggregate: a plain Row() leaf is used in place,
// anything else is evaluated into temp slot 0.
int bsi_prepare(pb200_ctx* ctx, pb200_family fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, const uint64_t* shards,
                uint32_t n_shards, BsiParams& p, BatchDev& bd) {
  if (fam >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (bit_depth > 63) return fail(ctx, PB200_ERR_INVALID, "bit depth > 63");
  TRY(validate_shards(ctx, shards, n_shards));
  Compiled cp;
  Compiler cc(ctx, cp);
  const bool has_filter = filt && nf;
  bool direct = false;
  if (has_filter) {
    if (!cc.compile(filt, nf)) return fail(ctx, PB200_ERR_INVALID, cc.err);
    direct = cp.pre.empty() && cp.progs[0].n_instr == 1 && cp.leaves[cp.progs[0].leaf_off + cp.instrs[cp.progs[0].instr_off].leaf].kind != L_TEMP;
  }
  TRY(sync_families(ctx));
  TRY(upload_batch(ctx, cp, shards, n_shards, bd));
  p = BsiParams{};
  p.families = ctx->d_families; p.family = fam; p.bit_depth = bit_depth;
  p.shards = bd.shards; p.n_shards = n_shards;
  p.algo_bytes = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_misc.p) + 8);
  if (has_filter) {
    p.has_filter = 1;
    if (direct) {
      p.filter = cp.leaves[cp.progs[0].leaf_off + cp.instrs[cp.progs[0].instr_off].leaf];
    } else {
      // evaluate into one extra temp slot placed after the program's own temps
      const uint32_t slot = cp.n_temp_slots;
      TRY(ensure_dev(ctx, ctx->d_temp, size_t(slot + 1) * n_shards * kRowContainers * kBitmapBytes));
      TRY(run_pretasks(ctx, cp, bd, n_shards));
      TRY(launch_eval(ctx, bd, bd.progs, 1, n_shards, true, nullptr, 0, 0, temp_slot(ctx, slot, n_shards), nullptr));
      p.filter.kind = L_TEMP; p.filter.family = slot; p.filter.row = 0;
    }
    p.temp_base = static_cast<const uint8_t*>(ctx->d_temp.p);
  }
  return PB200_OK;
}

int bsi_minmax_impl(pb200_ctx* ctx, pb200_family fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, int is_max,
                    const uint64_t* shards, uint32_t n_shards, std::vector<int64_t>& vals, std::vector<uint64_t>& counts) {
  vals.assign(n_shards, 0);
  counts.assign(n_shards, 0);
  if (n_shards == 0) return PB200_OK;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  BsiParams p; BatchDev bd{};
  TRY(bsi_prepare(ctx, fam, bit_depth, filt, nf, shards, n_shards, p, bd));
  const uint32_t items = n_shards * kRowContainers;
  TRY(ensure_dev(ctx, ctx->d_bsi, size_t(n_shards) * 16 + size_t(items) * sizeof(MmRec)));
  p.is_max = is_max ? 1 : 0;
  p.mm_val = static_cast<long long*>(ctx->d_bsi.p);
  p.mm_count = reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_bsi.p) + size_t(n_shards) * 8);
  MmRec* recs = reinterpret_cast<MmRec*>(static_cast<uint8_t*>(ctx->d_bsi.p) + size_t(n_shards) * 16);
  {
    constexpr int smem = (kPipeStages + 1) * kBitmapBytes;
    if (!ctx->mm_attr_set) { CUDA_TRY(ctx, cudaFuncSetAttribute(bsi_minmax_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); ctx->mm_attr_set = true; }
    const uint32_t grid = std::min<uint32_t>(items, resident_ctas(ctx, OCC_MM, bsi_minmax_kernel, kPipeThreads, smem));
    bsi_minmax_kernel<<<grid, kPipeThreads, smem, ctx->stream>>>(p, recs);
  }
  bsi_minmax_combine_kernel<<<(n_shards + 127) / 128, 128, 0, ctx->stream>>>(p, recs);
  ctx->timing.launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  TRY(ensure_pin(ctx, ctx->h_out, size_t(n_shards) * 16));  // values then counts, contiguous on the device too
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_out.p, p.mm_val, size_t(n_shards) * 16, cudaMemcpyDeviceToHost, ctx->stream));
  TRY(finish_call(ctx));
  std::memcpy(vals.data(), ctx->h_out.p, size_t(n_shards) * 8);
  std::memcpy(counts.data(), static_cast<const uint8_t*>(ctx->h_out.p) + size_t(n_shards) * 8, size_t(n_shards) * 8);
  ctx->timing.d2h_bytes += size_t(n_shards) * 16;
  return PB200_OK;
}

int bsi_minmax_reduce(pb200_ctx* ctx, pb200_family fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, int is_max,
                      const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  if (!out_val || !out_count) return PB200_ERR_INVALID;
  std::vector<int64_t> vals; std::vector<uint64_t> counts;
  TRY(bsi_minmax_impl(ctx, fam, bit_depth, filt, nf, is_max, shards, n_shards, vals, counts));
  // ValCount.smaller / larger (executor.go:3010-3029), folded in shard-list order
  int64_t v = 0; uint64_t c = 0;
  for (uint32_t s = 0; s < n_shards; s++) {
    const bool take = c == 0 || ((is_max ? vals[s] > v : vals[s] < v) && counts[s] > 0);
    if (take) { v = vals[s]; c = counts[s]; }
  }
  if (c == 0) v = 0;  // executor.go:470-472,506-508
  *out_val = v; *out_count = c;
  return PB200_OK;
}

}  // namespace

extern "C" int pb200_bsi_sum(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                             uint32_t flags, const uint64_t* shards, uint32_t n_shards, int64_t* out_sum, uint64_t* out_count) {
  ENTER(ctx);
  if (!out_sum || !out_count) return PB200_ERR_INVALID;
  *out_sum = 0; *out_count = 0;
  if (n_shards == 0) return PB200_OK;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  timing_begin(ctx);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
  BsiParams p; BatchDev bd{};
  TRY(bsi_prepare(ctx, bsi_family, bit_depth, filter_prog, n_filter_nodes, shards, n_shards, p, bd));
  const size_t n_acc = 2 * size_t(kMaxDepth) + 1;
  TRY(ensure_dev(ctx, ctx->d_bsi, n_acc * 8));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_bsi.p, 0, n_acc * 8, ctx->stream));
  p.strict = (flags & PB200_SUM_STRICT) ? 1 : 0;
  p.sum_acc = static_cast<unsigned long long*>(ctx->d_bsi.p);
  const uint32_t items = n_shards * kRowContainers;
  {
    constexpr int smem = (kPipeStages + 1) * kBitmapBytes;
    if (!ctx->sum_attr_set) { CUDA_TRY(ctx, cudaFuncSetAttribute(bsi_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); ctx->sum_attr_set = true; }
    const uint32_t grid = std::min<uint32_t>(items, resident_ctas(ctx, OCC_SUM, bsi_sum_kernel, kPipeThreads, smem));
    bsi_sum_kernel<<<grid, kPipeThreads, smem, ctx->stream>>>(p);
  }
  ctx->timing.launches++;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
  std::vector<unsigned long long> acc(n_acc);
  CUDA_TRY(ctx, cudaMemcpyAsync(acc.data(), ctx->d_bsi.p, n_acc * 8, cudaMemcpyDeviceToHost, ctx->stream));
  TRY(finish_call(ctx));
  ctx->timing.d2h_bytes += n_acc * 8;
  // sum = SUM_i 2^i * (|B_i & prow| - |B_i & nrow|) in wrapping 64-bit arithmetic (fragment.go:1130-1138)
  uint64_t sum = 0;
  for (uint32_t i = 0; i < bit_depth; i++) sum += (uint64_t(1) << i) * acc[1 + 2 * i] - (uint64_t(1) << i) * acc[2 + 2 * i];
  *out_count = acc[0];
  *out_sum = acc[0] ? int64_t(sum) : 0;  // executor.go:438-440: {0,0} when count is 0
  return PB200_OK;
}

extern "C" int pb200_bsi_min(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                             const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  ENTER(ctx);
  return bsi_minmax_reduce(ctx, bsi_family, bit_depth, filter_prog, n_filter_nodes, 0, shards, n_shards, out_val, out_count);
}

extern "C" int pb200_bsi_max(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                             const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  ENTER(ctx);
  return bsi_minmax_reduce(ctx, bsi_family, bit_depth, filter_prog, n_filter_nodes, 1, shards, n_shards, out_val, out_count);
}

extern "C" int pb200_bsi_minmax_shards(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                                       int is_max, const uint64_t* shards, uint32_t n_shards, int64_t* out_vals, uint64_t* out_counts) {
  ENTER(ctx);
  if (n_shards && (!out_vals || !out_counts)) return PB200_ERR_INVALID;
  std::vector<int64_t> vals; std::vector<uint64_t> counts;
  TRY(bsi_minmax_impl(ctx, bsi_family, bit_depth, filter_prog, n_filter_nodes, is_max, shards, n_shards, vals, counts));
  std::copy(vals.begin(), vals.end(), out_vals);
  std::copy(counts.begin(), counts.end(), out_counts);
  return PB200_OK;
}
