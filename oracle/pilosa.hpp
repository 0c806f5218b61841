// ORACLE — TEST INFRASTRUCTURE ONLY (see roaring.hpp header).
//
// CPU restatement of the reference's storage objects and executor for the hot path:
// Row (row.go), fragment row/BSI/top (fragment.go), rank cache + Pairs (cache.go),
// bsiGroup (field.go) and the executor's shard map-reduce (executor.go).
#pragma once
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <set>
#include <thread>
#include <unordered_map>

#include "../include/pilosa_b200.h"  // pb200_node / pb200_pair layouts only (interface, no code)
#include "roaring.hpp"

namespace oracle {

constexpr uint64_t kShardWidth = uint64_t(1) << 20;  // fragment.go:53
constexpr uint64_t kBsiExistsBit = 0, kBsiSignBit = 1, kBsiOffsetBit = 2;  // fragment.go:90-93
constexpr uint64_t kDefaultMinThreshold = 1;  // executor.go:37

// ---------------------------------------------------------------------------
// Row: one segment per shard (row.go:27-35, 332-344).
// ---------------------------------------------------------------------------
struct Row {
  std::map<uint64_t, Bitmap> segments;  // shard -> data

  static Row from_columns(const std::vector<uint64_t>& cols) {  // row.go NewRow(columns...)
    Row r;
    for (uint64_t c : cols) r.segments[c / kShardWidth].direct_add(c);
    return r;
  }
  uint64_t count() const {  // row.go:294-300
    uint64_t n = 0;
    for (auto& s : segments) n += s.second.count();
    return n;
  }
  bool any() const {  // row.go:123-130
    for (auto& s : segments) if (s.second.any()) return true;
    return false;
  }
  std::vector<uint64_t> columns() const {  // row.go:320-327
    std::vector<uint64_t> out;
    for (auto& s : segments) { auto v = s.second.slice(); out.insert(out.end(), v.begin(), v.end()); }
    return out;
  }
  // row.go:107-120
  Row intersect(const Row& o) const {
    Row r;
    for (auto& s : segments) {
      auto it = o.segments.find(s.first);
      if (it != o.segments.end()) r.segments[s.first] = oracle::intersect(s.second, it->second);
    }
    return r;
  }
  // row.go:153-195 (pairwise form; the n-way form is union_many below)
  Row union_(const Row& o) const {
    Row r = *this;
    for (auto& s : o.segments) {
      auto it = r.segments.find(s.first);
      if (it == r.segments.end()) r.segments[s.first] = s.second;
      else it->second = oracle::union_(it->second, s.second);
    }
    return r;
  }
  Row union_many(const std::vector<const Row*>& others) const {
    if (others.size() == 1) return union_(*others[0]);
    Row r;
    std::map<uint64_t, std::vector<const Bitmap*>> by_shard;
    for (auto& s : segments) by_shard[s.first].push_back(&s.second);
    for (auto* o : others) for (auto& s : o->segments) by_shard[s.first].push_back(&s.second);
    for (auto& [shard, list] : by_shard) {
      if (list.size() == 1) r.segments[shard] = *list[0];
      else if (list.size() == 2) r.segments[shard] = oracle::union_(*list[0], *list[1]);
      else r.segments[shard] = oracle::union_many(*list[0], std::vector<const Bitmap*>(list.begin() + 1, list.end()));
    }
    return r;
  }
  // row.go:198-213
  Row difference(const Row& o) const {
    Row r;
    for (auto& s : segments) {
      auto it = o.segments.find(s.first);
      if (it == o.segments.end()) r.segments[s.first] = s.second;
      else r.segments[s.first] = oracle::difference(s.second, it->second);
    }
    return r;
  }
  // row.go:133-150
  Row xor_(const Row& o) const {
    Row r = *this;
    for (auto& s : o.segments) {
      auto it = r.segments.find(s.first);
      if (it == r.segments.end()) r.segments[s.first] = s.second;
      else it->second = oracle::xor_(it->second, s.second);
    }
    return r;
  }
  // row.go:217-235 (n must be 1; roaring.go:944)
  Row shift() const {
    Row r;
    for (auto& s : segments) r.segments[s.first] = shift1(s.second);
    return r;
  }
  // row.go:91-104
  uint64_t intersection_count(const Row& o) const {
    uint64_t n = 0;
    for (auto& s : segments) {
      auto it = o.segments.find(s.first);
      if (it != o.segments.end()) n += oracle::intersection_count(s.second, it->second);
    }
    return n;
  }
  // row.go:67-88, 352-359: bit-by-bit merge of other into this.
  void merge(const Row& o) {
    for (auto& s : o.segments) {
      auto& dst = segments[s.first];
      for (uint64_t v : s.second.slice()) dst.direct_add(v);
    }
  }
};

// ---------------------------------------------------------------------------
// Pairs (cache.go:304-374) and Go's container/heap (stdlib; the pop order of
// fragment.top's result depends on its exact sift rules).
// ---------------------------------------------------------------------------
struct Pair { uint64_t id, count; };

struct PairHeap {  // cache.go:331-353; min-heap on Count
  std::vector<Pair> a;
  bool less(int i, int j) const { return a[i].count < a[j].count; }
  void up(int j) {
    for (;;) {
      int i = (j - 1) / 2;
      if (i == j || !less(j, i)) break;
      std::swap(a[i], a[j]);
      j = i;
    }
  }
  bool down(int i0, int n) {
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
    return i > i0;
  }
  void push(Pair p) { a.push_back(p); up((int)a.size() - 1); }
  Pair pop() {
    int n = (int)a.size() - 1;
    std::swap(a[0], a[n]);
    down(0, n);
    Pair p = a.back();
    a.pop_back();
    return p;
  }
};

// cache.go:356-374 Pairs.Add; the Go map order is unspecified so the merged
// list is returned sorted by id (callers sort by count afterwards).
inline std::vector<Pair> pairs_add(const std::vector<Pair>& p, const std::vector<Pair>& other) {
  std::map<uint64_t, uint64_t> m;
  for (auto& x : p) m[x.id] = x.count;
  for (auto& x : other) m[x.id] += x.count;
  std::vector<Pair> out;
  for (auto& kv : m) out.push_back({kv.first, kv.second});
  return out;
}

// sort.Sort(Pairs) (cache.go:343-347) is count-desc and unstable in Go; the
// deterministic refinement used here and by the engine: ties by ascending id.
inline void sort_pairs(std::vector<Pair>& p) {
  std::stable_sort(p.begin(), p.end(), [](const Pair& x, const Pair& y) {
    return x.count != y.count ? x.count > y.count : x.id < y.id;
  });
}

// ---------------------------------------------------------------------------
// rankCache (cache.go:136-288). Time-based invalidation debouncing (:237-244) is
// not modelled: recalculate() is explicit, as the reference tests do
// (fragment_internal_test.go calls f.RecalculateCache()).
// ---------------------------------------------------------------------------
struct RankCache {
  uint32_t max_entries = 50000;  // field.go:47 DefaultCacheSize
  std::map<uint64_t, uint64_t> entries;
  std::vector<Pair> rankings;
  uint64_t threshold_value = 0;

  void add(uint64_t id, uint64_t n) {  // :168-181
    if (n < threshold_value && n > 0) return;
    entries[id] = n;
  }
  void bulk_add(uint64_t id, uint64_t n) {  // :184-192 (no n > 0 exception: a count of 0 never replaces an entry once the threshold is 1)
    if (n < threshold_value) return;
    entries[id] = n;
  }
  uint64_t get(uint64_t id) const { auto it = entries.find(id); return it == entries.end() ? 0 : it->second; }
  void recalculate() {  // :246-282
    std::vector<Pair> r;
    for (auto& kv : entries) r.push_back({kv.first, kv.second});
    sort_pairs(r);
    std::vector<Pair> removed;
    if (r.size() > max_entries) {
      threshold_value = r[max_entries].count;
      removed.assign(r.begin() + max_entries, r.end());
      r.resize(max_entries);
    } else {
      threshold_value = 1;
    }
    rankings = std::move(r);
    int threshold_buffer = int(1.1 * double(max_entries));  // cache.go:31,161
    if ((int)entries.size() > threshold_buffer)
      for (auto& p : removed) entries.erase(p.id);
  }
};

// ---------------------------------------------------------------------------
// fragment (fragment.go): one roaring Bitmap holding every row of one shard.
// ---------------------------------------------------------------------------
struct TopOptions {  // fragment.go:1741-1756 (attr filters stay on the host)
  int n = 0;
  const Row* src = nullptr;
  std::vector<uint64_t> row_ids;
  uint64_t min_threshold = 0;
  uint64_t tanimoto_threshold = 0;
};

struct Fragment {
  uint64_t shard = 0;
  Bitmap storage;
  RankCache cache;
  bool cache_none = false;
  uint64_t max_row_id = 0;

  uint64_t pos(uint64_t row, uint64_t col) const { return row * kShardWidth + (col % kShardWidth); }  // :3078

  // fragment.go:621-641 rowFromStorage -> Bitmap.OffsetRange (roaring.go:535)
  Row row(uint64_t row_id) const {
    Row r;
    r.segments[shard] = offset_range(storage, shard * kShardWidth, row_id * kShardWidth, (row_id + 1) * kShardWidth);
    return r;
  }
  uint64_t row_count(uint64_t row_id) const {
    uint64_t n = 0;
    for (auto it = storage.cs.lower_bound(row_id * 16); it != storage.cs.end() && it->first < (row_id + 1) * 16; ++it)
      n += N(it->second);
    return n;
  }
  // fragment.rows (fragment.go:2668-2712) with the filters of :2606-2657 applied in the order the
  // callers build them: filterColumn, filterWithRows, filterWithLimit (last, :2608-2611).
  struct RowsOpts {
    uint64_t start = 0;
    bool has_column = false; uint64_t column = 0;
    const std::vector<uint64_t>* in_rows = nullptr;  // filterWithRows (ascending)
    bool has_limit = false; uint64_t limit = 0;
  };
  std::vector<uint64_t> rows(const RowsOpts& o) const {
    std::vector<uint64_t> out;
    uint64_t last_row = ~0ull, limit = o.limit;
    size_t loc = 0;
    for (auto it = storage.cs.lower_bound(o.start * 16); it != storage.cs.end(); ++it) {
      if (!it->second || N(it->second) == 0) continue;  // nil containers are skipped by the iterator (containers_slice.go:239-254)
      const uint64_t key = it->first, vrow = key >> 4;
      if (vrow == last_row) continue;
      bool add = true, done = false;
      if (o.has_column) {  // filterColumn :2622
        const uint64_t col_id = o.column % kShardWidth;
        const uint64_t col_key = (vrow * kShardWidth + col_id) >> 16;
        add = col_key == key && contains(it->second, uint16_t(col_id & 0xffff));
      }
      if (add && o.in_rows) {  // filterWithRows :2636
        const auto& r = *o.in_rows;
        if (loc >= r.size()) { add = false; done = true; }
        else {
          loc = size_t(std::lower_bound(r.begin() + loc, r.end(), vrow) - r.begin());
          if (loc >= r.size()) { add = false; done = true; }
          else if (r[loc] == vrow) { if (loc == r.size() - 1) done = true; }
          else add = false;
        }
      }
      if (add && o.has_limit) {  // filterWithLimit :2612
        if (limit > 0) limit--; else { add = false; done = true; }
      }
      if (add) { last_row = vrow; out.push_back(vrow); }
      if (done) return out;
    }
    return out;
  }
  bool set_bit(uint64_t row_id, uint64_t col) {  // :680-724
    bool changed = storage.direct_add(pos(row_id, col));
    if (!changed) return false;
    if (!cache_none) cache.add(row_id, row_count(row_id));
    if (row_id > max_row_id) max_row_id = row_id;
    return true;
  }
  bool clear_bit(uint64_t row_id, uint64_t col) {  // :742-779
    bool changed = storage.direct_remove(pos(row_id, col));
    if (!changed) return false;
    if (!cache_none) cache.add(row_id, row_count(row_id));
    return true;
  }
  // importPositions (fragment.go:2051-2098): storage.AddN(set), storage.RemoveN(clear), then the cache count of
  // every row named by a position is re-read from storage (BulkAdd, :2080-2090) and the ranking rebuilt (:2093).
  // Positions are fragment positions (row * ShardWidth + column % ShardWidth). Returns the bits changed.
  int64_t import_positions(const uint64_t* set, size_t n_set, const uint64_t* clear, size_t n_clear) {
    int64_t changed = 0;
    std::set<uint64_t> row_set;
    for (size_t i = 0; i < n_set; i++) { changed += storage.direct_add(set[i]); row_set.insert(set[i] / kShardWidth); }
    for (size_t i = 0; i < n_clear; i++) { changed += storage.direct_remove(clear[i]); row_set.insert(clear[i] / kShardWidth); }
    for (uint64_t row : row_set) {
      if (!cache_none) cache.bulk_add(row, row_count(row));
      if (row > max_row_id && row_count(row)) max_row_id = row;
    }
    if (!cache_none) cache.recalculate();
    return changed;
  }
  // fragment.go:975-1040 setValueBase (sign-magnitude, exists row last)
  bool set_value(uint64_t col, unsigned bit_depth, int64_t value) {
    bool changed = false;
    uint64_t uvalue = value < 0 ? uint64_t(-value) : uint64_t(value);
    for (unsigned i = 0; i < bit_depth; i++) {
      if (uvalue & (uint64_t(1) << i)) changed |= set_bit(kBsiOffsetBit + i, col);
      else changed |= clear_bit(kBsiOffsetBit + i, col);
    }
    changed |= set_bit(kBsiExistsBit, col);
    if (value < 0) changed |= set_bit(kBsiSignBit, col);
    else changed |= clear_bit(kBsiSignBit, col);
    return changed;
  }
  // fragment.go:926-929 clearValue -> setValueBase(clear = true): magnitude bits follow
  // `value`, then the exists and sign bits are cleared (:1008-1030)
  bool clear_value(uint64_t col, unsigned bit_depth, int64_t value) {
    bool changed = false;
    uint64_t uvalue = value < 0 ? uint64_t(-value) : uint64_t(value);
    for (unsigned i = 0; i < bit_depth; i++) {
      if (uvalue & (uint64_t(1) << i)) changed |= set_bit(kBsiOffsetBit + i, col);
      else changed |= clear_bit(kBsiOffsetBit + i, col);
    }
    changed |= clear_bit(kBsiExistsBit, col);
    changed |= clear_bit(kBsiSignBit, col);
    return changed;
  }
  // fragment.go:893-924
  bool value(uint64_t col, unsigned bit_depth, int64_t* out) const {
    if (!storage.contains_value(pos(kBsiExistsBit, col))) return false;
    int64_t v = 0;
    for (unsigned i = 0; i < bit_depth; i++)
      if (storage.contains_value(pos(kBsiOffsetBit + i, col))) v |= int64_t(1) << i;
    if (storage.contains_value(pos(kBsiSignBit, col))) v = -v;
    *out = v;
    return true;
  }
  void recalculate_cache() {  // rebuild from storage then rank (fragment.go RecalculateCache)
    cache.recalculate();
  }
  void rebuild_cache_from_storage() {  // what openCache + import recount amount to (:2080-2095)
    cache.entries.clear();
    uint64_t cur = ~uint64_t(0), n = 0;
    for (auto& kv : storage.cs) {
      uint64_t r = kv.first / 16;
      if (r != cur) { if (cur != ~uint64_t(0) && n) cache.entries[cur] = n; cur = r; n = 0; }
      n += N(kv.second);
    }
    if (cur != ~uint64_t(0) && n) cache.entries[cur] = n;
    if (!storage.cs.empty()) max_row_id = storage.cs.rbegin()->first / 16;
    cache.recalculate();
  }

  // ---- BSI (fragment.go:1109-1534) ----
  void sum(const Row* filter, unsigned bit_depth, int64_t* out_sum, uint64_t* out_count, bool strict = false) const {  // :1109-1141
    Row consider = row(kBsiExistsBit);
    if (filter) consider = consider.intersect(*filter);
    *out_count = consider.count();
    Row nrow = row(kBsiSignBit);
    if (strict) nrow = nrow.intersect(consider);  // NOT the reference; opt-in
    Row prow = consider.difference(nrow);
    int64_t s = 0;
    for (unsigned i = 0; i < bit_depth; i++) {
      Row r = row(kBsiOffsetBit + i);
      int64_t psum = int64_t((uint64_t(1) << i) * r.intersection_count(prow));
      int64_t nsum = int64_t((uint64_t(1) << i) * r.intersection_count(nrow));
      s = int64_t(uint64_t(s) + uint64_t(psum) - uint64_t(nsum));
    }
    *out_sum = s;
  }
  void min_unsigned(Row filter, unsigned bit_depth, int64_t* out, uint64_t* count) const {  // :1171-1186
    int64_t m = 0; uint64_t c = 0;
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = filter.difference(row(kBsiOffsetBit + i));
      c = r.count();
      if (c > 0) filter = r;
      else { m += int64_t(1) << i; if (i == 0) c = filter.count(); }
    }
    *out = m; *count = c;
  }
  void max_unsigned(Row filter, unsigned bit_depth, int64_t* out, uint64_t* count) const {  // :1213-1225
    int64_t m = 0; uint64_t c = 0;
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = row(kBsiOffsetBit + i).intersect(filter);
      c = r.count();
      if (c > 0) { m += int64_t(1) << i; filter = r; }
      else if (i == 0) c = filter.count();
    }
    *out = m; *count = c;
  }
  void min(const Row* filter, unsigned bit_depth, int64_t* out, uint64_t* count) const {  // :1145-1168
    Row consider = row(kBsiExistsBit);
    if (filter) consider = consider.intersect(*filter);
    *out = 0; *count = 0;
    if (consider.count() == 0) return;
    Row neg = row(kBsiSignBit).intersect(consider);
    if (neg.any()) { max_unsigned(neg, bit_depth, out, count); *out = -*out; return; }
    min_unsigned(consider, bit_depth, out, count);
  }
  void max(const Row* filter, unsigned bit_depth, int64_t* out, uint64_t* count) const {  // :1189-1210
    Row consider = row(kBsiExistsBit);
    if (filter) consider = consider.intersect(*filter);
    *out = 0; *count = 0;
    if (!consider.any()) return;
    Row p = consider.difference(row(kBsiSignBit));
    if (!p.any()) { min_unsigned(consider, bit_depth, out, count); *out = -*out; return; }
    max_unsigned(p, bit_depth, out, count);
  }

  Row range_eq(unsigned bit_depth, int64_t predicate) const {  // :1288-1312
    Row b = row(kBsiExistsBit);
    uint64_t up = uint64_t(predicate);
    if (predicate < 0) { up = uint64_t(-predicate); b = b.intersect(row(kBsiSignBit)); }
    else b = b.difference(row(kBsiSignBit));
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = row(kBsiOffsetBit + i);
      if ((up >> i) & 1) b = b.intersect(r); else b = b.difference(r);
    }
    return b;
  }
  Row range_neq(unsigned bit_depth, int64_t predicate) const {  // :1314-1329
    return row(kBsiExistsBit).difference(range_eq(bit_depth, predicate));
  }
  Row range_lt_unsigned(Row filter, unsigned bit_depth, uint64_t predicate, bool allow_eq) const {  // :1357-1398
    Row keep;
    bool leading_zeros = true;
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = row(kBsiOffsetBit + i);
      uint64_t bit = (predicate >> i) & 1;
      if (leading_zeros) {
        if (bit == 0) { filter = filter.difference(r); continue; }
        leading_zeros = false;
      }
      if (i == 0 && !allow_eq) {
        if (bit == 0) return keep;
        return filter.difference(r.difference(keep));
      }
      if (bit == 0) { filter = filter.difference(r.difference(keep)); continue; }
      if (i > 0) keep = keep.union_(filter.difference(r));
    }
    return filter;
  }
  Row range_gt_unsigned(Row filter, unsigned bit_depth, uint64_t predicate, bool allow_eq) const {  // :1423-1453
    Row keep;
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = row(kBsiOffsetBit + i);
      uint64_t bit = (predicate >> i) & 1;
      if (i == 0 && !allow_eq) {
        if (bit == 1) return keep;
        return filter.difference(filter.difference(r).difference(keep));
      }
      if (bit == 1) { filter = filter.difference(filter.difference(r).difference(keep)); continue; }
      if (i > 0) keep = keep.union_(filter.intersect(r));
    }
    return filter;
  }
  Row range_lt(unsigned bit_depth, int64_t predicate, bool allow_eq) const {  // :1331-1354
    Row b = row(kBsiExistsBit);
    uint64_t up = predicate < 0 ? uint64_t(-predicate) : uint64_t(predicate);
    if ((predicate >= 0 && allow_eq) || (predicate >= -1 && !allow_eq)) {
      Row p = range_lt_unsigned(b.difference(row(kBsiSignBit)), bit_depth, up, allow_eq);
      return row(kBsiSignBit).union_(p);
    }
    return range_gt_unsigned(b.intersect(row(kBsiSignBit)), bit_depth, up, allow_eq);
  }
  Row range_gt(unsigned bit_depth, int64_t predicate, bool allow_eq) const {  // :1400-1421
    Row b = row(kBsiExistsBit);
    uint64_t up = predicate < 0 ? uint64_t(-predicate) : uint64_t(predicate);
    if ((predicate >= 0 && allow_eq) || (predicate >= -1 && !allow_eq))
      return range_gt_unsigned(b.difference(row(kBsiSignBit)), bit_depth, up, allow_eq);
    Row neg = range_lt_unsigned(b.intersect(row(kBsiSignBit)), bit_depth, up, allow_eq);
    return b.difference(row(kBsiSignBit)).union_(neg);
  }
  Row range_between_unsigned(Row filter, unsigned bit_depth, uint64_t pmin, uint64_t pmax) const {  // :1496-1534
    Row keep1, keep2;
    for (int i = int(bit_depth) - 1; i >= 0; i--) {
      Row r = row(kBsiOffsetBit + i);
      uint64_t bit1 = (pmin >> i) & 1, bit2 = (pmax >> i) & 1;
      if (bit1 == 1) filter = filter.difference(filter.difference(r).difference(keep1));
      else if (i > 0) keep1 = keep1.union_(filter.intersect(r));
      if (bit2 == 0) filter = filter.difference(r.difference(keep2));
      else if (i > 0) keep2 = keep2.union_(filter.difference(r));
    }
    return filter;
  }
  Row range_between(unsigned bit_depth, int64_t pmin, int64_t pmax) const {  // :1463-1493
    Row b = row(kBsiExistsBit);
    uint64_t umin = pmin < 0 ? uint64_t(-pmin) : uint64_t(pmin);
    uint64_t umax = pmax < 0 ? uint64_t(-pmax) : uint64_t(pmax);
    if (pmin >= 0) return range_between_unsigned(b.difference(row(kBsiSignBit)), bit_depth, umin, umax);
    if (pmax < 0) return range_between_unsigned(b.intersect(row(kBsiSignBit)), bit_depth, umax, umin);
    Row p = range_lt_unsigned(b.difference(row(kBsiSignBit)), bit_depth, umax, true);
    Row n = range_lt_unsigned(b.intersect(row(kBsiSignBit)), bit_depth, umin, true);
    return p.union_(n);
  }
  Row range_op(uint32_t op, unsigned bit_depth, int64_t a, int64_t b2) const {  // :1271-1286
    switch (op) {
      case PB200_BSI_EQ: return range_eq(bit_depth, a);
      case PB200_BSI_NEQ: return range_neq(bit_depth, a);
      case PB200_BSI_LT: return range_lt(bit_depth, a, false);
      case PB200_BSI_LTE: return range_lt(bit_depth, a, true);
      case PB200_BSI_GT: return range_gt(bit_depth, a, false);
      case PB200_BSI_GTE: return range_gt(bit_depth, a, true);
      case PB200_BSI_BETWEEN: return range_between(bit_depth, a, b2);
      case PB200_BSI_NOTNULL: return row(kBsiExistsBit);  // :1456-1460
      default: throw std::runtime_error("invalid range operation");
    }
  }

  // ---- TopN (fragment.go:1568-1738) ----
  std::vector<Pair> top_bitmap_pairs(const std::vector<uint64_t>& row_ids) const {  // :1702-1738
    if (cache_none || row_ids.empty()) return cache.rankings;
    std::vector<Pair> pairs;
    for (uint64_t id : row_ids) {
      if (uint64_t n = cache.get(id); n > 0) { pairs.push_back({id, n}); continue; }
      uint64_t c = row_count(id);
      if (c > 0) pairs.push_back({id, c});
    }
    sort_pairs(pairs);  // sort.Sort(bitmapPairs): count desc; ties by id here
    return pairs;
  }
  std::vector<Pair> top(TopOptions opt) const {  // :1568-1700
    std::vector<Pair> pairs = top_bitmap_pairs(opt.row_ids);
    if (!opt.row_ids.empty()) opt.n = 0;
    uint64_t tanimoto = 0, src_count = 0;
    double min_tan = 0, max_tan = 0;
    if (opt.tanimoto_threshold > 0 && opt.src) {
      tanimoto = opt.tanimoto_threshold;
      src_count = opt.src->count();
      min_tan = double(src_count * tanimoto) / 100;
      max_tan = double(src_count * 100) / double(tanimoto);
    }
    PairHeap results;
    for (auto& pr : pairs) {
      uint64_t row_id = pr.id, cnt = pr.count;
      if (cnt == 0) continue;
      if (tanimoto > 0) {
        if (double(cnt) <= min_tan || double(cnt) >= max_tan) continue;
      } else if (cnt < opt.min_threshold) {
        continue;
      }
      if (opt.n == 0 || (int)results.a.size() < opt.n) {
        uint64_t count = cnt;
        if (opt.src) count = opt.src->intersection_count(row(row_id));
        if (count == 0) continue;
        if (tanimoto > 0) {
          double t = std::ceil(double(count * 100) / double(cnt + src_count - count));
          if (t <= double(tanimoto)) continue;
        } else if (count < opt.min_threshold) {
          continue;
        }
        results.push({row_id, count});
        if (opt.n > 0 && (int)results.a.size() == opt.n && !opt.src) break;
        continue;
      }
      uint64_t threshold = results.a[0].count;
      if (threshold < opt.min_threshold || cnt < threshold) break;
      uint64_t count = opt.src->intersection_count(row(row_id));
      if (count < threshold) continue;
      results.push({row_id, count});
    }
    std::vector<Pair> r(results.a.size());
    size_t x = r.size(), i = 1;
    while (!results.a.empty()) { r[x - i] = results.pop(); i++; }
    return r;
  }
};

// ---------------------------------------------------------------------------
// bsiGroup (field.go:1505-1586)
// ---------------------------------------------------------------------------
struct BsiGroup {
  int64_t min = 0, max = 0, base = 0;
  unsigned bit_depth = 0;
  int64_t bit_depth_min() const { return base - (int64_t(1) << bit_depth) + 1; }
  int64_t bit_depth_max() const { return base + (int64_t(1) << bit_depth) - 1; }
  // field.go:1526-1552; op uses pb200_bsi_op codes
  int64_t base_value(uint32_t op, int64_t value, bool* out_of_range) const {
    int64_t mn = bit_depth_min(), mx = bit_depth_max(), bv = 0;
    *out_of_range = false;
    if (op == PB200_BSI_GT || op == PB200_BSI_GTE) {
      if (value > mx) { *out_of_range = true; return 0; }
      else if (value > mn) bv = value - base;
    } else if (op == PB200_BSI_LT || op == PB200_BSI_LTE) {
      if (value < mn) { *out_of_range = true; return 0; }
      else if (value > mx) bv = mx - base;
      else bv = value - base;
    } else if (op == PB200_BSI_EQ || op == PB200_BSI_NEQ) {
      if (value < mn || value > mx) { *out_of_range = true; return 0; }
      bv = value - base;
    }
    return bv;
  }
  // field.go:1555-1569
  bool base_value_between(int64_t lo, int64_t hi, int64_t* olo, int64_t* ohi) const {
    int64_t mn = bit_depth_min(), mx = bit_depth_max();
    *olo = *ohi = 0;
    if (hi < mn || lo > mx) return true;
    if (lo < mn) lo = mn;
    if (hi > mx) hi = mx;
    *olo = lo - base; *ohi = hi - base;
    return false;
  }
};

// field.go:1606-1621
inline unsigned bit_depth_u64(uint64_t v) {
  for (unsigned i = 0; i < 63; i++) if (v < (uint64_t(1) << i)) return i;
  return 63;
}
inline unsigned bit_depth_i64(int64_t v) { return v < 0 ? bit_depth_u64(uint64_t(-v)) : bit_depth_u64(uint64_t(v)); }

// ---------------------------------------------------------------------------
// ValCount (executor.go:2997-3029)
// ---------------------------------------------------------------------------
struct ValCount {
  int64_t val = 0, count = 0;
  ValCount add(const ValCount& o) const { return {int64_t(uint64_t(val) + uint64_t(o.val)), count + o.count}; }
  ValCount smaller(const ValCount& o) const { return (count == 0 || (o.val < val && o.count > 0)) ? o : *this; }
  ValCount larger(const ValCount& o) const { return (count == 0 || (o.val > val && o.count > 0)) ? o : *this; }
};

// ---------------------------------------------------------------------------
// Holder + executor (holder.go:496-502; executor.go). Families stand for
// (index, field, view) triples; programs are the flattened post-order call
// trees of include/pilosa_b200.h.
// ---------------------------------------------------------------------------
struct Holder {
  std::vector<std::map<uint64_t, Fragment>> families;

  uint32_t family_create() { families.emplace_back(); return (uint32_t)families.size() - 1; }
  Fragment* fragment(uint32_t fam, uint64_t shard, bool create = false) {
    if (fam >= families.size()) throw std::runtime_error("unknown family");
    auto& m = families[fam];
    auto it = m.find(shard);
    if (it != m.end()) return &it->second;
    if (!create) return nullptr;
    Fragment& f = m[shard];
    f.shard = shard;
    return &f;
  }
};

// The executor's worker pool (executor.go:83-107): NewExecutor starts workerPoolSize goroutines ONCE; each runs
// worker() (:2604-2611), taking jobs off the shared `work` channel for as long as the executor lives. mapperLocal
// (:2576-2601) pushes one job per shard onto that channel and then collects one response per job; jobs of different
// requests interleave on the same workers. Restated with threads created once, a queue of job groups (one group =
// the jobs one caller pushed) and a completion wait per group.
class WorkerPool {
 public:
  explicit WorkerPool(int n) {
    for (int i = 0; i < std::max(1, n); i++) threads_.emplace_back([this] { worker(); });
  }
  ~WorkerPool() {
    { std::lock_guard<std::mutex> g(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  int size() const { return (int)threads_.size(); }

  // Pushes jobs fn(0) .. fn(n-1) and blocks until every one of them has reported back.
  void run(size_t n, const std::function<void(size_t)>& fn) {
    if (n == 0) return;
    auto g = std::make_shared<Group>();
    g->fn = &fn; g->n = n;
    { std::lock_guard<std::mutex> lk(mu_); groups_.push_back(g); }
    cv_.notify_all();
    std::unique_lock<std::mutex> lk(g->mu);
    g->cv.wait(lk, [&] { return g->done.load(std::memory_order_acquire) == n; });
    if (g->err) std::rethrow_exception(g->err);
  }

 private:
  struct Group {
    const std::function<void(size_t)>* fn = nullptr;
    size_t n = 0;
    std::atomic<size_t> next{0}, done{0};
    std::exception_ptr err;
    std::mutex mu;
    std::condition_variable cv;
  };
  void worker() {
    std::shared_ptr<Group> g;
    for (;;) {
      if (!g) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || !groups_.empty(); });
        if (stop_) return;
        g = groups_.front();
      }
      const size_t i = g->next.fetch_add(1, std::memory_order_relaxed);
      if (i >= g->n) {  // this group has no job left to hand out: retire it from the queue
        std::lock_guard<std::mutex> lk(mu_);
        if (!groups_.empty() && groups_.front() == g) groups_.pop_front();
        g.reset();
        continue;
      }
      try { (*g->fn)(i); }
      catch (...) { std::lock_guard<std::mutex> lk(g->mu); g->err = std::current_exception(); }
      if (g->done.fetch_add(1, std::memory_order_acq_rel) + 1 == g->n) {
        std::lock_guard<std::mutex> lk(g->mu);
        g->cv.notify_all();
      }
    }
  }
  std::vector<std::thread> threads_;
  std::deque<std::shared_ptr<Group>> groups_;
  std::mutex mu_;
  std::condition_variable cv_;
  bool stop_ = false;
};

struct Executor {
  Holder* holder;
  int workers = 1;  // executor.go:83-107 worker pool; server/config.go:172 NumCPU
  mutable std::unique_ptr<WorkerPool> pool_;
  mutable std::mutex pool_mu_;
  WorkerPool& pool() const {  // started on first use, restarted when `workers` changes
    std::lock_guard<std::mutex> g(pool_mu_);
    if (!pool_ || pool_->size() != std::max(1, workers)) pool_.reset(new WorkerPool(workers));
    return *pool_;
  }

  // executor.go:654-680 executeBitmapCallShard on a flattened program.
  Row eval_shard(const pb200_node* prog, uint32_t n, uint64_t shard) const {
    std::vector<Row> st;
    for (uint32_t k = 0; k < n; k++) {
      const pb200_node& nd = prog[k];
      switch (nd.op) {
        case PB200_OP_ROW: {  // executor.go:1494-1499
          Fragment* f = holder->fragment(nd.family, shard);
          st.push_back(f ? f->row(nd.row) : Row());
          break;
        }
        case PB200_OP_EMPTY: st.push_back(Row()); break;
        case PB200_OP_INTERSECT: case PB200_OP_UNION: case PB200_OP_DIFFERENCE: case PB200_OP_XOR: {
          uint32_t na = nd.nargs;
          if (na == 0) {
            if (nd.op == PB200_OP_UNION || nd.op == PB200_OP_XOR) { st.push_back(Row()); break; }
            throw std::runtime_error("empty Intersect/Difference query is currently not supported");  // :1673,1008
          }
          if (st.size() < na) throw std::runtime_error("malformed program");
          size_t base = st.size() - na;
          Row acc = std::move(st[base]);
          for (uint32_t j = 1; j < na; j++) {
            const Row& r = st[base + j];
            if (nd.op == PB200_OP_INTERSECT) acc = acc.intersect(r);
            else if (nd.op == PB200_OP_UNION) acc = acc.union_(r);
            else if (nd.op == PB200_OP_DIFFERENCE) acc = acc.difference(r);
            else acc = acc.xor_(r);
          }
          st.resize(base);
          st.push_back(std::move(acc));
          break;
        }
        case PB200_OP_NOT: {  // :1737-1768
          if (st.empty()) throw std::runtime_error("Not() requires an input row");
          Fragment* f = holder->fragment(nd.family, shard);
          Row ex = f ? f->row(0) : Row();
          Row r = ex.difference(st.back());
          st.back() = std::move(r);
          break;
        }
        case PB200_OP_SHIFT: {  // :1771-1790
          if (st.empty()) throw std::runtime_error("Shift() requires an input row");
          st.back() = st.back().shift();
          break;
        }
        case PB200_OP_BSI_RANGE: {  // :1536-1665 (host already re-based the predicate)
          Fragment* f = holder->fragment(nd.family, shard);
          st.push_back(f ? f->range_op(nd.bsi_op, nd.bit_depth, nd.a, nd.b) : Row());
          break;
        }
        default: throw std::runtime_error("unknown op");
      }
    }
    if (st.size() != 1) throw std::runtime_error("malformed program");
    return std::move(st[0]);
  }

  // executor.go:2576-2611 mapperLocal: one job per shard pushed to the (persistent) worker pool, one response each.
  template <class T>
  std::vector<T> map_shards(const std::vector<uint64_t>& shards, const std::function<T(uint64_t)>& fn) const {
    std::vector<T> out(shards.size());
    if (workers <= 1 || shards.size() <= 1) { for (size_t i = 0; i < shards.size(); i++) out[i] = fn(shards[i]); return out; }
    pool().run(shards.size(), [&](size_t i) { out[i] = fn(shards[i]); });
    return out;
  }

  // Many Count calls in flight at once — what concurrent HTTP requests give the reference: every request runs its
  // own mapperLocal, and all of them push their per-shard jobs onto the ONE shared pool (executor.go:99-105), so
  // the workers see queries x shards jobs. Each job is executeCount's map function for one shard (:1804-1810), the
  // per-query reduce is `uint64 +` (:1813). out[q] = Count of program q over `shards`.
  std::vector<uint64_t> count_batch(const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                                    const std::vector<uint64_t>& shards, bool fused = false) const {
    const size_t ns = shards.size();
    std::vector<uint64_t> per(size_t(n_queries) * ns, 0), out(n_queries, 0);
    auto job = [&](size_t j) {
      const uint32_t q = uint32_t(j / ns);
      per[j] = count_one_shard(progs + prog_off[q], prog_off[q + 1] - prog_off[q], shards[j % ns], fused);
    };
    if (workers <= 1) for (size_t j = 0; j < per.size(); j++) job(j);
    else pool().run(per.size(), job);
    for (uint32_t q = 0; q < n_queries; q++)
      for (size_t s = 0; s < ns; s++) out[q] += per[size_t(q) * ns + s];
    return out;
  }

  // The map function of executeCount for one shard. fused=false: executeBitmapCallShard materialises the row, then
  // Row.Count (what the reference does). fused=true: a top-level two-leaf Intersect goes through intersectionCount
  // without materialising (BASELINE.md mode B "best-effort CPU").
  uint64_t count_one_shard(const pb200_node* prog, uint32_t n, uint64_t shard, bool fused) const {
    if (fused && n == 3 && prog[0].op == PB200_OP_ROW && prog[1].op == PB200_OP_ROW && prog[2].op == PB200_OP_INTERSECT && prog[2].nargs == 2) {
      Fragment* fa = holder->fragment(prog[0].family, shard);
      Fragment* fb = holder->fragment(prog[1].family, shard);
      if (!fa || !fb) return 0;
      return fa->row(prog[0].row).intersection_count(fb->row(prog[1].row));
    }
    return eval_shard(prog, n, shard).count();
  }

  // executor.go:1793-1825. fused=false materialises then counts (what the
  // reference does); fused=true uses intersectionCount for a top-level
  // two-leaf Intersect (BASELINE.md mode B "best-effort CPU").
  uint64_t count(const pb200_node* prog, uint32_t n, const std::vector<uint64_t>& shards, bool fused = false) const {
    auto per = count_shards(prog, n, shards, fused);
    uint64_t tot = 0;
    for (uint64_t v : per) tot += v;
    return tot;
  }
  std::vector<uint64_t> count_shards(const pb200_node* prog, uint32_t n, const std::vector<uint64_t>& shards, bool fused = false) const {
    return map_shards<uint64_t>(shards, [&](uint64_t shard) -> uint64_t { return count_one_shard(prog, n, shard, fused); });
  }

  // executor.go:933-1000 executeTopNShard
  std::vector<Pair> topn_shard(uint32_t fam, const pb200_node* src, uint32_t nsrc, uint32_t n,
                               const std::vector<uint64_t>& ids, uint64_t min_threshold, uint64_t tanimoto,
                               uint64_t shard) const {
    Fragment* f = holder->fragment(fam, shard);
    if (!f) return {};
    Row src_row;
    TopOptions opt;
    if (src && nsrc) { src_row = eval_shard(src, nsrc, shard); opt.src = &src_row; }
    if (min_threshold == 0) min_threshold = kDefaultMinThreshold;
    if (tanimoto > 100) throw std::runtime_error("Tanimoto Threshold is from 1 to 100 only");
    opt.n = (int)n; opt.row_ids = ids; opt.min_threshold = min_threshold; opt.tanimoto_threshold = tanimoto;
    return f->top(opt);
  }
  // executor.go:905-930 executeTopNShards
  std::vector<Pair> topn_shards(uint32_t fam, const pb200_node* src, uint32_t nsrc, uint32_t n,
                                const std::vector<uint64_t>& ids, uint64_t min_threshold, uint64_t tanimoto,
                                const std::vector<uint64_t>& shards) const {
    auto per = map_shards<std::vector<Pair>>(shards, [&](uint64_t shard) {
      return topn_shard(fam, src, nsrc, n, ids, min_threshold, tanimoto, shard);
    });
    std::vector<Pair> acc;
    for (auto& p : per) acc = pairs_add(acc, p);
    sort_pairs(acc);
    return acc;
  }
  // executor.go:863-903 executeTopN (two passes)
  std::vector<Pair> topn(uint32_t fam, const pb200_node* src, uint32_t nsrc, uint32_t n,
                         const std::vector<uint64_t>& ids, uint64_t min_threshold, uint64_t tanimoto,
                         const std::vector<uint64_t>& shards, bool remote = false) const {
    auto pairs = topn_shards(fam, src, nsrc, n, ids, min_threshold, tanimoto, shards);
    if (pairs.empty() || !ids.empty() || remote) return pairs;  // executor.go:884-887
    std::vector<uint64_t> ids2;
    for (auto& p : pairs) ids2.push_back(p.id);
    std::sort(ids2.begin(), ids2.end());
    auto trimmed = topn_shards(fam, src, nsrc, n, ids2, min_threshold, tanimoto, shards);
    if (n != 0 && n < trimmed.size()) trimmed.resize(n);
    return trimmed;
  }

  // RowIDs.merge (executor.go:1043-1069)
  static std::vector<uint64_t> rowids_merge(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b, uint64_t limit) {
    std::vector<uint64_t> r;
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size() && r.size() < limit) {
      if (a[i] < b[j]) r.push_back(a[i++]);
      else if (a[i] > b[j]) r.push_back(b[j++]);
      else { r.push_back(b[j]); i++; j++; }
    }
    while (i < a.size() && r.size() < limit) r.push_back(a[i++]);
    while (j < b.size() && r.size() < limit) r.push_back(b[j++]);
    return r;
  }
  // executeRows / executeRowsShard (executor.go:1277-1441), standard view only (time views stay on the host)
  std::vector<uint64_t> rows(uint32_t fam, uint64_t start, bool has_column, uint64_t column, bool has_limit, uint64_t limit,
                             const std::vector<uint64_t>& shards) const {
    const uint64_t lim = has_limit ? limit : ~0ull;
    auto per = map_shards<std::vector<uint64_t>>(shards, [&](uint64_t shard) -> std::vector<uint64_t> {
      Fragment* f = holder->fragment(fam, shard);
      if (!f) return {};
      if (has_column && column / kShardWidth != shard) return {};  // :1405-1409
      Fragment::RowsOpts o;
      o.start = start; o.has_column = has_column; o.column = column; o.has_limit = has_limit; o.limit = limit;
      return f->rows(o);
    });
    std::vector<uint64_t> acc;
    for (auto& v : per) acc = rowids_merge(acc, v, lim);
    return acc;
  }

  // ---- GroupBy (executor.go:1071-1275, groupByIterator :3060-3230, rowIterator fragment.go:2759-2791) ----
  struct GroupCount { std::vector<uint64_t> rows; uint64_t count = 0; };
  struct GroupByChild {
    uint32_t fam = 0;
    bool has_prev = false; uint64_t prev = 0;
    std::vector<uint64_t> child_rows;  // pre-executed Rows(limit= / column=); empty = no filterWithRows (:3118-3120)
  };
  struct RowIter {  // fragment.go:2759-2791
    const Fragment* f = nullptr;
    std::vector<uint64_t> ids;
    size_t cur = 0;
    bool wrap = false;
    void seek(uint64_t row_id) { cur = size_t(std::lower_bound(ids.begin(), ids.end(), row_id) - ids.begin()); }
    bool next(uint64_t* id, bool* wrapped) {
      *wrapped = false;
      if (cur >= ids.size()) {
        if (!wrap || ids.empty()) { *wrapped = true; return false; }
        seek(0);
        *wrapped = true;
      }
      *id = ids[cur++];
      return true;
    }
  };
  struct GroupByIter {
    std::vector<RowIter> its;
    std::vector<Row> rows;
    std::vector<uint64_t> ids;
    const Row* filter = nullptr;
    bool done = false;
    void set_row(size_t i, uint64_t id) {  // the three cases of nextAtIdx :3180-3186
      Row nr = its[i].f->row(id);
      if (i == 0 && filter) rows[i] = nr.intersect(*filter);
      else if (i == 0 || i == rows.size() - 1) rows[i] = nr;
      else rows[i] = nr.intersect(rows[i - 1]);
      ids[i] = id;
    }
    void next_at(size_t i) {  // :3167-3191
      for (;;) {
        uint64_t id; bool wrapped;
        if (!its[i].next(&id, &wrapped)) { done = true; return; }
        if (wrapped && i != 0) next_at(i - 1);
        // The reference keeps looping here after the recursion has set done; with >= 3 fields and no
        // non-empty row left at this level that loop never ends (the iterator wraps for ever). Stop instead.
        if (done) return;
        set_row(i, id);
        if (rows[i].count() != 0) break;
      }
    }
    bool next(GroupCount* out) {  // :3193-3230
      uint64_t count;
      for (;;) {
        if (done) return false;
        const size_t k = rows.size();
        count = k == 1 ? rows[0].count() : rows[k - 1].intersection_count(rows[k - 2]);
        if (count == 0) { next_at(k - 1); continue; }
        break;
      }
      out->rows = ids;
      out->count = count;
      next_at(rows.size() - 1);
      return true;
    }
  };
  std::vector<GroupCount> group_by_shard(const std::vector<GroupByChild>& ch, const pb200_node* filt, uint32_t nf, uint64_t limit,
                                         uint64_t shard) const {
    Row filter_row;
    GroupByIter g;
    if (filt && nf) { filter_row = eval_shard(filt, nf, shard); g.filter = &filter_row; }
    const size_t k = ch.size();
    g.its.resize(k); g.rows.resize(k); g.ids.assign(k, 0);
    bool ignore_prev = false;
    for (size_t i = 0; i < k; i++) {  // newGroupByIterator :3086-3165
      Fragment* f = holder->fragment(ch[i].fam, shard);
      if (!f) return {};  // "this whole shard doesn't have all it needs to continue"
      Fragment::RowsOpts o;
      if (!ch[i].child_rows.empty()) o.in_rows = &ch[i].child_rows;
      g.its[i].f = f; g.its[i].ids = f->rows(o); g.its[i].wrap = i != 0;
      uint64_t prev = ch[i].prev;
      if (ch[i].has_prev && !ignore_prev) {
        if (i == k - 1) prev++;
        g.its[i].seek(prev);
      }
      uint64_t id; bool wrapped;
      if (!g.its[i].next(&id, &wrapped)) { g.done = true; break; }
      g.rows[i] = f->row(id); g.ids[i] = id;
      if (ch[i].has_prev && id != prev) ignore_prev = true;
      if (wrapped) {
        bool dead = false;
        for (size_t j = i; j-- > 0;) {
          uint64_t id2; bool w2;
          if (!g.its[j].next(&id2, &w2)) { dead = true; break; }
          g.rows[j] = g.its[j].f->row(id2); g.ids[j] = id2;
          if (!w2) break;
        }
        if (dead) { g.done = true; break; }
      }
    }
    if (!g.done) {
      if (g.filter && k > 0) g.rows[0] = g.rows[0].intersect(*g.filter);
      for (size_t i = 1; i + 1 < k; i++) g.rows[i] = g.rows[i].intersect(g.rows[i - 1]);
    }
    std::vector<GroupCount> results;  // executeGroupByShard :1262-1274
    GroupCount gc;
    while (results.size() < limit && g.next(&gc))
      if (gc.count > 0) results.push_back(gc);
    return results;
  }
  static int group_compare(const GroupCount& a, const GroupCount& b) {  // :1203-1214
    for (size_t i = 0; i < a.rows.size(); i++) {
      if (a.rows[i] < b.rows[i]) return -1;
      if (a.rows[i] > b.rows[i]) return 1;
    }
    return 0;
  }
  static std::vector<GroupCount> merge_group_counts(std::vector<GroupCount> a, const std::vector<GroupCount>& b, uint64_t limit) {  // :1166-1200
    limit = std::min<uint64_t>(limit, a.size() + b.size());
    std::vector<GroupCount> ret;
    size_t i = 0, j = 0;
    while (i < a.size() && j < b.size() && ret.size() < limit) {
      const int c = group_compare(a[i], b[j]);
      if (c < 0) ret.push_back(a[i++]);
      else if (c == 0) { a[i].count += b[j].count; ret.push_back(a[i]); i++; j++; }
      else ret.push_back(b[j++]);
    }
    for (; i < a.size() && ret.size() < limit; i++) ret.push_back(a[i]);
    for (; j < b.size() && ret.size() < limit; j++) ret.push_back(b[j]);
    return ret;
  }
  // executeGroupBy's map-reduce (:1122-1137); offset / final limit are applied by the caller (:1139-1157)
  std::vector<GroupCount> group_by(const std::vector<GroupByChild>& ch, const pb200_node* filt, uint32_t nf, uint64_t limit,
                                   const std::vector<uint64_t>& shards) const {
    if (ch.empty()) throw std::runtime_error("need at least one child call");
    auto per = map_shards<std::vector<GroupCount>>(shards, [&](uint64_t shard) { return group_by_shard(ch, filt, nf, limit, shard); });
    std::vector<GroupCount> acc;
    for (auto& v : per) acc = merge_group_counts(acc, v, limit);
    return acc;
  }

  // executor.go:409-442 + 683-721 (Base is added by the caller)
  ValCount sum(uint32_t fam, unsigned bit_depth, const pb200_node* filt, uint32_t nf,
               const std::vector<uint64_t>& shards, bool strict = false) const {
    auto per = map_shards<ValCount>(shards, [&](uint64_t shard) -> ValCount {
      Fragment* f = holder->fragment(fam, shard);
      if (!f) return {};
      Row fr; const Row* fp = nullptr;
      if (filt && nf) { fr = eval_shard(filt, nf, shard); fp = &fr; }
      int64_t s; uint64_t c;
      f->sum(fp, bit_depth, &s, &c, strict);
      return {s, (int64_t)c};
    });
    ValCount acc;
    for (auto& v : per) acc = acc.add(v);
    if (acc.count == 0) return {};
    return acc;
  }
  std::vector<ValCount> minmax_shards(uint32_t fam, unsigned bit_depth, const pb200_node* filt, uint32_t nf,
                                      bool is_max, const std::vector<uint64_t>& shards) const {
    return map_shards<ValCount>(shards, [&](uint64_t shard) -> ValCount {
      Fragment* f = holder->fragment(fam, shard);
      if (!f) return {};
      Row fr; const Row* fp = nullptr;
      if (filt && nf) { fr = eval_shard(filt, nf, shard); fp = &fr; }
      int64_t v; uint64_t c;
      if (is_max) f->max(fp, bit_depth, &v, &c); else f->min(fp, bit_depth, &v, &c);
      return {v, (int64_t)c};
    });
  }
  // executor.go:445-514: reduce in ascending shard-list order.
  ValCount minmax(uint32_t fam, unsigned bit_depth, const pb200_node* filt, uint32_t nf, bool is_max,
                  const std::vector<uint64_t>& shards) const {
    auto per = minmax_shards(fam, bit_depth, filt, nf, is_max, shards);
    ValCount acc;
    for (auto& v : per) acc = is_max ? acc.larger(v) : acc.smaller(v);
    if (acc.count == 0) return {};
    return acc;
  }
};

}  // namespace oracle
