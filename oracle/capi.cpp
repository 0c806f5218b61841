// ORACLE — TEST INFRASTRUCTURE ONLY (see roaring.hpp header).
// extern "C" surface of the CPU oracle for the ctypes test driver (oracle/__init__.py).
#include <chrono>
#include <cstdio>

#include "pilosa.hpp"

using namespace oracle;

namespace {
thread_local std::string g_err;

struct Orc {
  Holder holder;
  Executor exec;
  Orc() { exec.holder = &holder; }
};

CPtr build_container(int type, const uint16_t* vals, int n) {
  // vals: array -> values; run -> (start,last) pairs, n = number of pairs; bitmap -> values
  if (type == kNil) return nullptr;
  if (type == kArray) return Container::make_array(std::vector<uint16_t>(vals, vals + n));
  if (type == kRun) {
    std::vector<Interval> r;
    for (int i = 0; i < n; i++) r.push_back({vals[2 * i], vals[2 * i + 1]});
    return Container::make_run(std::move(r));
  }
  auto b = Container::make_bitmap();
  for (int i = 0; i < n; i++) b->bitmap[vals[i] >> 6] |= uint64_t(1) << (vals[i] & 63);
  int32_t c = 0;
  for (uint64_t w : b->bitmap) c += popc(w);
  b->n = c;
  return b;
}

template <class F> int guard(F&& f) {
  try { f(); return 0; }
  catch (const std::exception& e) { g_err = e.what(); return -1; }
}
}  // namespace

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

// ---- container level -------------------------------------------------------
// op: 0 intersect, 1 union, 2 difference, 3 xor, 4 flip(a), 5 optimize(a), 6 convert(a -> type tb), 7 shift(a)
// Output: *out_type, *out_n (cardinality), out_vals (sorted values, cap 65536), *out_repr_len
// (array: n, run: number of runs, bitmap: 1024), *carry for shift.
int orc_container_op(int op, int ta, const uint16_t* va, int na, int tb, const uint16_t* vb, int nb,
                     int* out_type, int* out_n, uint16_t* out_vals, int* out_repr_len, int* carry) {
  return guard([&] {
    CPtr a = build_container(ta, va, na);
    CPtr b = op <= 3 ? build_container(tb, vb, nb) : nullptr;
    CPtr r;
    bool cy = false;
    switch (op) {
      case 0: r = intersect(a, b); break;
      case 1: r = union_(a, b); break;
      case 2: r = difference(a, b); break;
      case 3: r = xor_(a, b); break;
      case 4: r = flip(a); break;
      case 5: r = optimize(a); break;
      case 6: r = convert_to(a, CType(tb)); break;
      case 7: r = shift(a, &cy); break;
      default: throw std::runtime_error("bad op");
    }
    if (carry) *carry = cy;
    *out_type = r ? r->type : kNil;
    *out_n = N(r);
    auto vals = values_of(r);
    if ((int)vals.size() != N(r)) throw std::runtime_error("container n mismatch: n=" + std::to_string(N(r)) + " values=" + std::to_string(vals.size()));
    std::copy(vals.begin(), vals.end(), out_vals);
    *out_repr_len = !r ? 0 : r->type == kArray ? (int)r->array.size() : r->type == kRun ? (int)r->runs.size() : kBitmapN;
  });
}

int orc_container_icount(int ta, const uint16_t* va, int na, int tb, const uint16_t* vb, int nb, int* out) {
  return guard([&] { *out = intersection_count(build_container(ta, va, na), build_container(tb, vb, nb)); });
}

// misc container helpers: 0 countRuns, 1 bitmapCountRange(start,end), 2 count
int orc_container_stat(int what, int ta, const uint16_t* va, int na, int start, int end, int* out) {
  return guard([&] {
    CPtr a = build_container(ta, va, na);
    if (what == 0) *out = count_runs(*a);
    else if (what == 1) *out = bitmap_count_range(*a, start, end);
    else *out = N(a);
  });
}

// ---- bitmap level ------------------------------------------------------------
void* orc_bitmap_new(const uint64_t* vals, size_t n) {
  auto* b = new Bitmap();
  for (size_t i = 0; i < n; i++) b->direct_add(vals[i]);
  return b;
}
void orc_bitmap_free(void* b) { delete (Bitmap*)b; }
int orc_bitmap_add(void* b, uint64_t v) { return ((Bitmap*)b)->direct_add(v); }
int orc_bitmap_remove(void* b, uint64_t v) { return ((Bitmap*)b)->direct_remove(v); }
void orc_bitmap_optimize(void* b) { ((Bitmap*)b)->optimize_all(); }
uint64_t orc_bitmap_count(void* b) { return ((Bitmap*)b)->count(); }
// op: 0 intersect 1 union 2 difference 3 xor 4 shift
void* orc_bitmap_op(int op, void* a, void* b) {
  auto* A = (Bitmap*)a; auto* B = (Bitmap*)b;
  auto* o = new Bitmap();
  switch (op) {
    case 0: *o = intersect(*A, *B); break;
    case 1: *o = union_(*A, *B); break;
    case 2: *o = difference(*A, *B); break;
    case 3: *o = xor_(*A, *B); break;
    case 4: *o = shift1(*A); break;
  }
  return o;
}
void* orc_bitmap_union_many(void* a, void** others, int n) {
  std::vector<const Bitmap*> v;
  for (int i = 0; i < n; i++) v.push_back((Bitmap*)others[i]);
  auto* o = new Bitmap();
  *o = union_many(*(Bitmap*)a, v);
  return o;
}
uint64_t orc_bitmap_intersection_count(void* a, void* b) { return intersection_count(*(Bitmap*)a, *(Bitmap*)b); }
size_t orc_bitmap_slice(void* b, uint64_t* out, size_t cap) {
  auto v = ((Bitmap*)b)->slice();
  if (out) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), out);
  return v.size();
}
// container types in key order: out_keys/out_types/out_ns (cap entries) -> returns number of containers
size_t orc_bitmap_containers(void* b, uint64_t* keys, int* types, int* ns, size_t cap) {
  size_t i = 0;
  for (auto& kv : ((Bitmap*)b)->cs) {
    if (!kv.second) continue;
    if (i < cap) { keys[i] = kv.first; types[i] = kv.second->type; ns[i] = kv.second->n; }
    i++;
  }
  return i;
}
size_t orc_bitmap_serialize(void* b, int do_optimize, uint8_t* out, size_t cap) {
  auto v = serialize(*(Bitmap*)b, do_optimize != 0);
  if (out && cap >= v.size()) std::copy(v.begin(), v.end(), out);
  return v.size();
}
void* orc_bitmap_deserialize(const uint8_t* data, size_t len) {  // Bitmap.UnmarshalBinary: either format
  try { return new Bitmap(unmarshal_binary(data, len)); }
  catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
// append an op record (typ 0..3) to a byte buffer held by the caller: returns encoded bytes
size_t orc_encode_op(int typ, const uint64_t* values, size_t n, uint8_t* out, size_t cap) {
  std::vector<uint8_t> v;
  append_op(v, (uint8_t)typ, std::vector<uint64_t>(values, values + n));
  if (out && cap >= v.size()) std::copy(v.begin(), v.end(), out);
  return v.size();
}

// ---- row level (row_test.go) -----------------------------------------------------
void* orc_row_new(const uint64_t* cols, size_t n) { return new Row(Row::from_columns(std::vector<uint64_t>(cols, cols + n))); }
void orc_row_free(void* r) { delete (Row*)r; }
// op: 0 intersect 1 union 2 difference 3 xor 4 shift 5 merge(a <- b, returns copy of a)
void* orc_row_op(int op, void* a, void* b) {
  auto* A = (Row*)a; auto* B = (Row*)b;
  switch (op) {
    case 0: return new Row(A->intersect(*B));
    case 1: return new Row(A->union_(*B));
    case 2: return new Row(A->difference(*B));
    case 3: return new Row(A->xor_(*B));
    case 4: return new Row(A->shift());
    case 5: A->merge(*B); return new Row(*A);
  }
  return nullptr;
}
uint64_t orc_row_count(void* r) { return ((Row*)r)->count(); }
uint64_t orc_row_intersection_count(void* a, void* b) { return ((Row*)a)->intersection_count(*(Row*)b); }
size_t orc_row_columns(void* r, uint64_t* out, size_t cap) {
  auto v = ((Row*)r)->columns();
  if (out) std::copy(v.begin(), v.begin() + std::min(cap, v.size()), out);
  return v.size();
}

// ---- holder / fragments ----------------------------------------------------------
void* orc_new(int workers) { auto* o = new Orc(); o->exec.workers = workers > 0 ? workers : (int)std::thread::hardware_concurrency(); return o; }
void orc_free(void* h) { delete (Orc*)h; }
void orc_set_workers(void* h, int w) { ((Orc*)h)->exec.workers = w; }
uint32_t orc_family_create(void* h) { return ((Orc*)h)->holder.family_create(); }

int orc_frag_load(void* h, uint32_t fam, uint64_t shard, const uint8_t* data, size_t len) {
  return guard([&] {
    Fragment* f = ((Orc*)h)->holder.fragment(fam, shard, true);
    f->storage = deserialize(data, len);
    f->rebuild_cache_from_storage();
  });
}
int orc_frag_set_bit(void* h, uint32_t fam, uint64_t shard, uint64_t row, uint64_t col) {
  return ((Orc*)h)->holder.fragment(fam, shard, true)->set_bit(row, col);
}
int orc_frag_clear_bit(void* h, uint32_t fam, uint64_t shard, uint64_t row, uint64_t col) {
  return ((Orc*)h)->holder.fragment(fam, shard, true)->clear_bit(row, col);
}
int orc_frag_set_value(void* h, uint32_t fam, uint64_t shard, uint64_t col, uint32_t bit_depth, int64_t v) {
  return ((Orc*)h)->holder.fragment(fam, shard, true)->set_value(col, bit_depth, v);
}
int orc_frag_clear_value(void* h, uint32_t fam, uint64_t shard, uint64_t col, uint32_t bit_depth, int64_t v) {
  return ((Orc*)h)->holder.fragment(fam, shard, true)->clear_value(col, bit_depth, v);
}
int orc_frag_value(void* h, uint32_t fam, uint64_t shard, uint64_t col, uint32_t bit_depth, int64_t* out) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  return f ? f->value(col, bit_depth, out) : 0;
}
void orc_frag_cache_config(void* h, uint32_t fam, uint64_t shard, int cache_none, uint32_t max_entries) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard, true);
  f->cache_none = cache_none != 0;
  f->cache.max_entries = max_entries;
}
void orc_frag_cache_add(void* h, uint32_t fam, uint64_t shard, uint64_t id, uint64_t n) {  // rankCache.Add cache.go:168-181
  ((Orc*)h)->holder.fragment(fam, shard, true)->cache.add(id, n);
}
size_t orc_frag_cache_len(void* h, uint32_t fam, uint64_t shard) {  // rankCache.Len cache.go:202-206
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  return f ? f->cache.entries.size() : 0;
}
void orc_frag_recalculate_cache(void* h, uint32_t fam, uint64_t shard) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  if (f) f->recalculate_cache();
}
void orc_frag_rebuild_cache(void* h, uint32_t fam, uint64_t shard) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  if (f) f->rebuild_cache_from_storage();
}
size_t orc_frag_rankings(void* h, uint32_t fam, uint64_t shard, uint64_t* ids, uint64_t* counts, size_t cap) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  if (!f) return 0;
  size_t n = f->cache.rankings.size();
  for (size_t i = 0; i < n && i < cap; i++) { ids[i] = f->cache.rankings[i].id; counts[i] = f->cache.rankings[i].count; }
  return n;
}
size_t orc_frag_serialize(void* h, uint32_t fam, uint64_t shard, uint8_t* out, size_t cap) {
  Fragment* f = ((Orc*)h)->holder.fragment(fam, shard);
  if (!f) return 0;
  auto v = serialize(f->storage, true);
  if (out && cap >= v.size()) std::copy(v.begin(), v.end(), out);
  return v.size();
}

// ---- queries -----------------------------------------------------------------------
static std::vector<uint64_t> vec(const uint64_t* p, size_t n) { return p ? std::vector<uint64_t>(p, p + n) : std::vector<uint64_t>(); }

int orc_count(void* h, const pb200_node* prog, uint32_t n, const uint64_t* shards, uint32_t ns, int fused, uint64_t* out) {
  return guard([&] { *out = ((Orc*)h)->exec.count(prog, n, vec(shards, ns), fused != 0); });
}
int orc_count_shards(void* h, const pb200_node* prog, uint32_t n, const uint64_t* shards, uint32_t ns, int fused, uint64_t* out) {
  return guard([&] {
    auto v = ((Orc*)h)->exec.count_shards(prog, n, vec(shards, ns), fused != 0);
    std::copy(v.begin(), v.end(), out);
  });
}
// n_queries Count calls in flight together on the shared worker pool (Executor::count_batch).
int orc_count_batch(void* h, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries, const uint64_t* shards, uint32_t ns,
                    int fused, uint64_t* out) {
  return guard([&] {
    auto v = ((Orc*)h)->exec.count_batch(progs, prog_off, n_queries, vec(shards, ns), fused != 0);
    std::copy(v.begin(), v.end(), out);
  });
}
int orc_eval_columns(void* h, const pb200_node* prog, uint32_t n, uint64_t shard, uint64_t* out, size_t cap, size_t* out_n) {
  return guard([&] {
    auto cols = ((Orc*)h)->exec.eval_shard(prog, n, shard).columns();
    *out_n = cols.size();
    if (out) std::copy(cols.begin(), cols.begin() + std::min(cap, cols.size()), out);
  });
}
// The `Columns` field of internal.Row as the generated marshaller writes it (internal/public.pb.go:745-766,
// reached from encodeRow, encoding/proto/proto.go:408-457): tag 0x0A, varint(payload length), the columns as
// base-128 varints; nothing at all for an empty row. Columns are Row.Columns() (row.go:302-327) of the call's
// result, segment by segment in shard-list order (Row.Merge keeps the segments sorted by shard when the list is).
int orc_eval_proto_columns(void* h, const pb200_node* prog, uint32_t n, const uint64_t* shards, uint32_t ns, uint8_t* out, size_t cap,
                           size_t* out_len) {
  return guard([&] {
    std::vector<uint8_t> payload;
    for (uint32_t i = 0; i < ns; i++) {
      for (uint64_t num : ((Orc*)h)->exec.eval_shard(prog, n, shards[i]).columns()) {
        while (num >= 0x80) { payload.push_back(uint8_t((num & 0x7f) | 0x80)); num >>= 7; }
        payload.push_back(uint8_t(num));
      }
    }
    std::vector<uint8_t> msg;
    if (!payload.empty()) {
      msg.push_back(0x0a);
      uint64_t l = payload.size();
      while (l >= 0x80) { msg.push_back(uint8_t((l & 0x7f) | 0x80)); l >>= 7; }
      msg.push_back(uint8_t(l));
      msg.insert(msg.end(), payload.begin(), payload.end());
    }
    *out_len = msg.size();
    if (out && cap >= msg.size()) std::copy(msg.begin(), msg.end(), out);
  });
}
// optimized Pilosa-roaring bytes of the shard's result segment
int orc_eval_roaring(void* h, const pb200_node* prog, uint32_t n, uint64_t shard, uint8_t* out, size_t cap, size_t* out_len) {
  return guard([&] {
    Row r = ((Orc*)h)->exec.eval_shard(prog, n, shard);
    Bitmap b;
    if (!r.segments.empty()) b = r.segments.begin()->second;
    auto v = serialize(b, true);
    *out_len = v.size();
    if (out && cap >= v.size()) std::copy(v.begin(), v.end(), out);
  });
}
int orc_topn(void* h, uint32_t fam, const pb200_node* src, uint32_t nsrc, uint32_t n, const uint64_t* ids, size_t nids,
             uint64_t min_threshold, uint64_t tanimoto, uint32_t flags, const uint64_t* shards, uint32_t ns, pb200_pair* out, size_t cap, size_t* n_out) {
  return guard([&] {
    auto r = ((Orc*)h)->exec.topn(fam, src, nsrc, n, vec(ids, nids), min_threshold, tanimoto, vec(shards, ns), (flags & PB200_TOPN_REMOTE) != 0);
    *n_out = r.size();
    for (size_t i = 0; i < r.size() && i < cap; i++) out[i] = {r[i].id, r[i].count};
  });
}
int orc_topn_shard(void* h, uint32_t fam, const pb200_node* src, uint32_t nsrc, uint32_t n, const uint64_t* ids, size_t nids,
                   uint64_t min_threshold, uint64_t tanimoto, uint64_t shard, pb200_pair* out, size_t cap, size_t* n_out) {
  return guard([&] {
    auto r = ((Orc*)h)->exec.topn_shard(fam, src, nsrc, n, vec(ids, nids), min_threshold, tanimoto, shard);
    *n_out = r.size();
    for (size_t i = 0; i < r.size() && i < cap; i++) out[i] = {r[i].id, r[i].count};
  });
}
// fragment.importRoaring (fragment.go:2243-2330): ImportRoaringBits into storage, then the rank cache
// entries of the touched rows are refreshed from exact counts (:2080-2095)
int orc_frag_drop(void* h, uint32_t fam, uint64_t shard) {
  return guard([&] {
    auto& fams = ((Orc*)h)->holder.families;
    if (fam >= fams.size()) throw std::runtime_error("unknown family");
    fams[fam].erase(shard);
  });
}
int orc_frag_import_positions(void* h, uint32_t fam, uint64_t shard, const uint64_t* set, size_t n_set, const uint64_t* clear, size_t n_clear,
                              int64_t* changed) {
  return guard([&] {
    Fragment* f = ((Orc*)h)->holder.fragment(fam, shard, true);
    const int64_t n = f->import_positions(set, n_set, clear, n_clear);
    if (changed) *changed = n;
  });
}
int orc_frag_import_roaring(void* h, uint32_t fam, uint64_t shard, const uint8_t* data, size_t len, int clear, int64_t* changed) {
  return guard([&] {
    Fragment* f = ((Orc*)h)->holder.fragment(fam, shard, true);
    Bitmap other = unmarshal_binary(data, len);
    std::map<uint64_t, int64_t> row_set;
    const int64_t n = import_roaring_bits(f->storage, data, len, clear != 0, &row_set);
    if (changed) *changed = n;
    for (auto& kv : other.cs)
      if ((kv.first >> 4) > f->max_row_id) f->max_row_id = kv.first >> 4;
    // fragment.importRoaring's cache update (fragment.go:2258-2276): the cached count of every row that
    // changed moves by the signed change (uint64 arithmetic), then the ranking is rebuilt at once
    bool any = false;
    for (auto& kv : row_set) {
      if (kv.second == 0 || f->cache_none) continue;
      any = true;
      f->cache.bulk_add(kv.first, f->cache.get(kv.first) + uint64_t(kv.second));
    }
    if (any) f->cache.recalculate();
  });
}
int orc_rows(void* h, uint32_t fam, uint64_t start, uint32_t flags, uint64_t column, uint64_t limit, const uint64_t* shards, uint32_t ns,
             uint64_t* out, size_t cap, size_t* n_out) {
  return guard([&] {
    auto r = ((Orc*)h)->exec.rows(fam, start, (flags & PB200_ROWS_COLUMN) != 0, column, limit != ~0ull, limit, vec(shards, ns));
    *n_out = r.size();
    for (size_t i = 0; i < r.size() && i < cap; i++) out[i] = r[i];
  });
}
int orc_groupby(void* h, const pb200_groupby_child* ch, uint32_t nch, const pb200_node* filt, uint32_t nf, uint64_t limit,
                const uint64_t* shards, uint32_t ns, pb200_group* out, size_t cap, size_t* n_out) {
  return guard([&] {
    if (nch > PB200_GROUPBY_MAX_FIELDS) throw std::runtime_error("too many GroupBy fields");
    std::vector<Executor::GroupByChild> c(nch);
    for (uint32_t i = 0; i < nch; i++) {
      c[i].fam = ch[i].family; c[i].has_prev = ch[i].has_previous != 0; c[i].prev = ch[i].previous;
      if (ch[i].rows) c[i].child_rows.assign(ch[i].rows, ch[i].rows + ch[i].n_rows);
    }
    auto r = ((Orc*)h)->exec.group_by(c, filt, nf, limit, vec(shards, ns));
    *n_out = r.size();
    for (size_t i = 0; i < r.size() && i < cap; i++) {
      pb200_group g{};
      for (size_t j = 0; j < r[i].rows.size(); j++) g.rows[j] = r[i].rows[j];
      g.count = r[i].count;
      out[i] = g;
    }
  });
}
int orc_bsi_sum(void* h, uint32_t fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, uint32_t flags,
                const uint64_t* shards, uint32_t ns, int64_t* out_sum, uint64_t* out_count) {
  return guard([&] {
    ValCount v = ((Orc*)h)->exec.sum(fam, bit_depth, filt, nf, vec(shards, ns), (flags & PB200_SUM_STRICT) != 0);
    *out_sum = v.val; *out_count = (uint64_t)v.count;
  });
}
int orc_bsi_minmax(void* h, uint32_t fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, int is_max,
                   const uint64_t* shards, uint32_t ns, int64_t* out_val, uint64_t* out_count) {
  return guard([&] {
    ValCount v = ((Orc*)h)->exec.minmax(fam, bit_depth, filt, nf, is_max != 0, vec(shards, ns));
    *out_val = v.val; *out_count = (uint64_t)v.count;
  });
}
int orc_bsi_minmax_shards(void* h, uint32_t fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, int is_max,
                          const uint64_t* shards, uint32_t ns, int64_t* out_vals, uint64_t* out_counts) {
  return guard([&] {
    auto v = ((Orc*)h)->exec.minmax_shards(fam, bit_depth, filt, nf, is_max != 0, vec(shards, ns));
    for (size_t i = 0; i < v.size(); i++) { out_vals[i] = v[i].val; out_counts[i] = (uint64_t)v[i].count; }
  });
}
// bsiGroup helpers (field.go:1526-1621)
int64_t orc_base_value(int64_t base, uint32_t bit_depth, uint32_t op, int64_t value, int* out_of_range) {
  BsiGroup g; g.base = base; g.bit_depth = bit_depth;
  bool oor; int64_t v = g.base_value(op, value, &oor); *out_of_range = oor; return v;
}
int orc_base_value_between(int64_t base, uint32_t bit_depth, int64_t lo, int64_t hi, int64_t* olo, int64_t* ohi) {
  BsiGroup g; g.base = base; g.bit_depth = bit_depth;
  return g.base_value_between(lo, hi, olo, ohi);
}
uint32_t orc_bit_depth(uint64_t v) { return bit_depth_u64(v); }
uint32_t orc_bit_depth_i64(int64_t v) { return bit_depth_i64(v); }

}  // extern "C"
