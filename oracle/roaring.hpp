// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement (C++17, no dependencies) of the roaring container algebra of
// the reference (travisturner/pilosa, roaring/roaring.go). It exists to CHECK the
// CUDA engine: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may call into it. The product library (libpilosa_b200)
// never links or loads anything under oracle/.
//
// Parity pin: the reference is Go and no Go toolchain exists in this image, so the
// reference itself cannot be executed here. The oracle is pinned instead against
// the reference's own golden vectors and known-answer tests, restated in
// tests/test_oracle_*.py (SURVEY.md §8c lists them).
//
// Every function cites the reference file:line it follows. The code is a
// restatement of the published algorithms, not a transliteration: containers are
// immutable values behind shared_ptr (the reference uses frozen/COW pointers).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace oracle {

// roaring/roaring.go:55-69, 1984-1992
constexpr int kBitmapN = 1024;
constexpr int32_t kMaxContainerVal = 0xffff;
constexpr int32_t kFullN = 65536;
constexpr int32_t kArrayMaxSize = 4096;
constexpr int32_t kRunMaxSize = 2048;
constexpr uint64_t kMaxContainerKey = (uint64_t(1) << 48) - 1;
constexpr uint32_t kMagicNumber = 12348;

enum CType : uint8_t { kNil = 0, kArray = 1, kBitmap = 2, kRun = 3 };

struct Interval {  // roaring.go:1989-1992 (inclusive [start,last])
  uint16_t start, last;
  bool operator==(const Interval& o) const { return start == o.start && last == o.last; }
};

inline int popc(uint64_t x) { return __builtin_popcountll(x); }

struct Container;
using CPtr = std::shared_ptr<const Container>;
using MPtr = std::shared_ptr<Container>;

// roaring/container_stash.go:39-46 — tagged union of array | bitmap | run.
struct Container {
  CType type = kArray;
  int32_t n = 0;
  std::vector<uint16_t> array;
  std::vector<uint64_t> bitmap;  // kBitmapN words when type == kBitmap
  std::vector<Interval> runs;

  static MPtr make_array(std::vector<uint16_t> v) {
    auto c = std::make_shared<Container>();
    c->type = kArray;
    c->n = (int32_t)v.size();
    c->array = std::move(v);
    return c;
  }
  static MPtr make_bitmap() {
    auto c = std::make_shared<Container>();
    c->type = kBitmap;
    c->bitmap.assign(kBitmapN, 0);
    return c;
  }
  static MPtr make_bitmap(std::vector<uint64_t> w, int32_t n) {
    auto c = std::make_shared<Container>();
    c->type = kBitmap;
    w.resize(kBitmapN, 0);
    c->bitmap = std::move(w);
    c->n = n;
    return c;
  }
  // NewContainerRun computes n from the runs (roaring.go NewContainerRun).
  static MPtr make_run(std::vector<Interval> r) {
    auto c = std::make_shared<Container>();
    c->type = kRun;
    int32_t n = 0;
    for (auto& iv : r) n += 1 + int32_t(iv.last) - int32_t(iv.start);
    c->n = n;
    c->runs = std::move(r);
    return c;
  }
  bool contains_bit(uint16_t v) const { return (bitmap[v >> 6] >> (v & 63)) & 1; }
};

inline int32_t N(const CPtr& c) { return c ? c->n : 0; }  // container_stash.go:210-215
inline int32_t N(const Container* c) { return c ? c->n : 0; }

inline const CPtr& full_container() {  // roaring.go:78
  static CPtr f = Container::make_run({Interval{0, 0xffff}});
  return f;
}

// ---------------------------------------------------------------------------
// counting helpers
// ---------------------------------------------------------------------------

// roaring.go:2040-2074
inline int32_t bitmap_count_range(const Container& c, int32_t start, int32_t end) {
  uint64_t n = 0;
  int32_t i = start / 64, j = end / 64;
  const auto& b = c.bitmap;
  if (i == j) {
    unsigned offi = unsigned(start % 64), offj = unsigned(64 - end % 64);
    // Go shifts >= 64 yield 0; emulate.
    uint64_t v = b[i] >> offi;
    unsigned sh = offj + offi;
    v = sh >= 64 ? 0 : (v << sh);
    return popc(v);
  }
  if (unsigned off = unsigned(start) % 64; off != 0) {
    n += popc(b[i] >> off);
    i++;
  }
  for (; i < j; i++) n += popc(b[i]);
  if (j < (int32_t)b.size()) {
    unsigned off = 64 - (unsigned(end) % 64);
    n += off >= 64 ? 0 : popc(b[j] << off);
  }
  return (int32_t)n;
}

// roaring.go:2249-2257
inline int32_t bitmap_count_runs(const std::vector<uint64_t>& b) {
  int32_t r = 0;
  for (int i = 0; i < kBitmapN - 1; i++) {
    uint64_t v = b[i], v1 = b[i + 1];
    r += popc((v << 1) & ~v) + int32_t((v >> 63) & ~v1);
  }
  uint64_t vl = b[kBitmapN - 1];
  r += popc((vl << 1) & ~vl) + int32_t(vl >> 63);
  return r;
}

// roaring.go:2259-2268
inline int32_t array_count_runs(const std::vector<uint16_t>& a) {
  int32_t r = 0, prev = -2;
  for (uint16_t v : a) {
    if (prev + 1 != int32_t(v)) r++;
    prev = v;
  }
  return r;
}

// roaring.go:2274-2286
inline int32_t count_runs(const Container& c) {
  switch (c.type) {
    case kArray: return array_count_runs(c.array);
    case kBitmap: return bitmap_count_runs(c.bitmap);
    case kRun: return (int32_t)c.runs.size();
    default: return 0;
  }
}

// ---------------------------------------------------------------------------
// conversions (roaring.go:2556-2834)
// ---------------------------------------------------------------------------

inline MPtr bitmap_to_array(const Container& c) {  // :2556-2603
  std::vector<uint16_t> out;
  out.reserve(c.n > 0 ? c.n : 0);
  for (int i = 0; i < kBitmapN; i++) {
    uint64_t w = c.bitmap[i];
    while (w) {
      uint64_t t = w & (~w + 1);
      out.push_back(uint16_t(i * 64 + popc(t - 1)));
      w ^= t;
    }
  }
  return Container::make_array(std::move(out));
}

inline MPtr array_to_bitmap(const Container& c) {  // :2606-2637
  auto o = Container::make_bitmap();
  for (uint16_t v : c.array) o->bitmap[v >> 6] |= uint64_t(1) << (v & 63);
  o->n = c.n;
  return o;
}

inline MPtr run_to_bitmap(const Container& c) {  // :2640-2675
  auto o = Container::make_bitmap();
  for (auto& r : c.runs)
    for (int v = r.start; v <= int(r.last); v++) o->bitmap[v >> 6] |= uint64_t(1) << (v & 63);
  o->n = c.n;
  return o;
}

inline MPtr bitmap_to_run(const Container& c) {  // :2678-2746
  std::vector<Interval> runs;
  if (c.n == 0) return Container::make_run({});
  const auto& b = c.bitmap;
  uint64_t cur = b[0];
  uint32_t i = 0;
  for (;;) {
    while (cur == 0 && i < kBitmapN - 1) cur = b[++i];
    if (cur == 0) break;
    uint16_t start = uint16_t(64 * i + __builtin_ctzll(cur));
    cur |= cur - 1;
    while (cur == ~uint64_t(0) && i < kBitmapN - 1) cur = b[++i];
    if (cur == ~uint64_t(0)) {
      runs.push_back({start, 0xffff});
      break;
    }
    uint16_t last = uint16_t(64 * i + __builtin_ctzll(~cur));
    runs.push_back({start, uint16_t(last - 1)});
    cur &= cur + 1;
  }
  auto o = Container::make_run(std::move(runs));
  o->n = c.n;
  return o;
}

inline MPtr array_to_run(const Container& c) {  // :2749-2791
  if (c.n == 0) return Container::make_run({});
  std::vector<Interval> runs;
  const auto& a = c.array;
  uint16_t start = a[0];
  for (size_t i = 1; i < a.size(); i++) {
    if (uint16_t(a[i] - a[i - 1]) > 1) {
      runs.push_back({start, a[i - 1]});
      start = a[i];
    }
  }
  runs.push_back({start, a.back()});
  auto o = Container::make_run(std::move(runs));
  o->n = c.n;
  return o;
}

inline MPtr run_to_array(const Container& c) {  // :2794-2834
  std::vector<uint16_t> out;
  out.reserve(c.n > 0 ? c.n : 0);
  for (auto& r : c.runs)
    for (int v = r.start; v <= int(r.last); v++) out.push_back(uint16_t(v));
  return Container::make_array(std::move(out));
}

// roaring.go:2289-2338 — canonical type: run if runs<=2048 && runs<=n/2; else
// array if n<4096; else bitmap; empty -> nil.
inline CPtr optimize(const CPtr& c) {
  if (N(c) == 0) return nullptr;
  int32_t runs = count_runs(*c);
  CType nt;
  if (runs <= kRunMaxSize && runs <= c->n / 2) nt = kRun;
  else if (c->n < kArrayMaxSize) nt = kArray;
  else nt = kBitmap;
  if (nt == c->type) return c;
  switch (c->type) {
    case kArray: return nt == kBitmap ? CPtr(array_to_bitmap(*c)) : CPtr(array_to_run(*c));
    case kBitmap: return nt == kArray ? CPtr(bitmap_to_array(*c)) : CPtr(bitmap_to_run(*c));
    case kRun: return nt == kBitmap ? CPtr(run_to_bitmap(*c)) : CPtr(run_to_array(*c));
    default: return c;
  }
}

// Force a container into a given representation (test helper; mirrors the
// conversions used by roaring_helpers_test.go fixtures).
inline CPtr convert_to(const CPtr& c, CType t) {
  if (!c || c->type == t) return c;
  if (c->type == kArray) return t == kBitmap ? CPtr(array_to_bitmap(*c)) : CPtr(array_to_run(*c));
  if (c->type == kBitmap) return t == kArray ? CPtr(bitmap_to_array(*c)) : CPtr(bitmap_to_run(*c));
  return t == kArray ? CPtr(run_to_array(*c)) : CPtr(run_to_bitmap(*c));
}

// Decode any container to the sorted list of its values.
inline std::vector<uint16_t> values_of(const CPtr& c) {
  if (!c) return {};
  if (c->type == kArray) return c->array;
  if (c->type == kBitmap) return bitmap_to_array(*c)->array;
  return run_to_array(*c)->array;
}

// ---------------------------------------------------------------------------
// search helpers
// ---------------------------------------------------------------------------

// roaring.go:4612-4649 (search32): index, or -(insert+1) when absent.
inline int32_t search32(const std::vector<uint16_t>& a, uint16_t value) {
  int32_t n = (int32_t)a.size();
  if (n == 0) return -1;
  if (a[n - 1] == value) return n - 1;
  int32_t lo = 0, hi = n - 1;
  while (lo + 16 <= hi) {
    int32_t i = int32_t(uint32_t(lo + hi) >> 1);
    uint16_t v = a[i];
    if (v < value) lo = i + 1;
    else if (v > value) hi = i - 1;
    else return i;
  }
  for (; lo <= hi; lo++) {
    uint16_t v = a[lo];
    if (v == value) return lo;
    if (v > value) break;
  }
  return -(lo + 1);
}

// roaring.go:2409-2417
inline std::pair<int32_t, bool> bin_search_runs(uint16_t v, const std::vector<Interval>& a) {
  int32_t lo = 0, hi = (int32_t)a.size();
  while (lo < hi) {
    int32_t m = (lo + hi) / 2;
    if (a[m].last >= v) hi = m; else lo = m + 1;
  }
  if (lo < (int32_t)a.size()) return {lo, v >= a[lo].start && v <= a[lo].last};
  return {lo, false};
}

inline bool contains(const CPtr& c, uint16_t v) {  // roaring.go:2232-2243
  if (!c) return false;
  if (c->type == kArray) return search32(c->array, v) >= 0;
  if (c->type == kRun) return bin_search_runs(v, c->runs).second;
  return c->contains_bit(v);
}

// ---------------------------------------------------------------------------
// single-value add / remove (roaring.go:2118-2230, 2426-2506). Value semantics:
// returns a new container (the reference mutates in place when not frozen).
// ---------------------------------------------------------------------------

inline CPtr bitmap_add(const CPtr& c, uint16_t v, bool* added) {  // :2178-2192
  if (c->contains_bit(v)) { *added = false; return c; }
  auto o = std::make_shared<Container>(*c);
  o->bitmap[v >> 6] |= uint64_t(1) << (v & 63);
  o->n++;
  *added = true;
  return o;
}

inline CPtr array_add(const CPtr& c, uint16_t v, bool* added) {  // :2144-2176
  if (c->n > 0 && c->n < kArrayMaxSize && c->array.back() < v) {
    auto o = std::make_shared<Container>(*c);
    o->array.push_back(v);
    o->n++;
    *added = true;
    return o;
  }
  int32_t i = search32(c->array, v);
  if (i >= 0) { *added = false; return c; }
  if (c->n >= kArrayMaxSize) return bitmap_add(array_to_bitmap(*c), v, added);
  auto o = std::make_shared<Container>(*c);
  i = -i - 1;
  o->array.insert(o->array.begin() + i, v);
  o->n++;
  *added = true;
  return o;
}

inline CPtr run_add(const CPtr& c, uint16_t v, bool* added) {  // :2194-2246
  if (c->runs.empty()) {
    *added = true;
    return Container::make_run({Interval{v, v}});
  }
  const auto& runs0 = c->runs;
  int32_t i = bin_search_runs(v, runs0).first;
  if (i == (int32_t)runs0.size()) i--;
  Interval iv = runs0[i];
  if (v >= iv.start && iv.last >= v) { *added = false; return c; }
  auto o = std::make_shared<Container>(*c);
  auto& runs = o->runs;
  if (iv.last < v) {
    if (iv.last == uint16_t(v - 1)) runs[i].last++;
    else runs.push_back({v, v});
  } else if (uint16_t(v + 1) == iv.start) {
    if (i > 0 && runs[i - 1].last == uint16_t(v - 1)) {
      runs[i - 1].last = iv.last;
      runs.erase(runs.begin() + i);
    } else {
      runs[i].start--;
    }
  } else if (i > 0 && uint16_t(v - 1) == runs[i - 1].last) {
    runs[i - 1].last++;
  } else {
    runs.insert(runs.begin() + i, Interval{v, v});
  }
  o->n++;
  *added = true;
  return o;
}

inline CPtr add(const CPtr& c, uint16_t v, bool* added) {  // :2118-2130
  if (!c) { *added = true; return Container::make_array({v}); }
  if (c->type == kArray) return array_add(c, v, added);
  if (c->type == kRun) return run_add(c, v, added);
  return bitmap_add(c, v, added);
}

inline CPtr remove(const CPtr& c, uint16_t v, bool* removed) {  // :2426-2506
  *removed = false;
  if (!c) return nullptr;
  if (c->type == kArray) {
    int32_t i = search32(c->array, v);
    if (i < 0) return c;
    *removed = true;
    if (c->n == 1) return nullptr;
    auto o = std::make_shared<Container>(*c);
    o->array.erase(o->array.begin() + i);
    o->n--;
    return o;
  }
  if (c->type == kBitmap) {
    if (!c->contains_bit(v)) return c;
    *removed = true;
    if (c->n == 1) return nullptr;
    auto o = std::make_shared<Container>(*c);
    o->bitmap[v >> 6] &= ~(uint64_t(1) << (v & 63));
    o->n--;
    if (o->n == kArrayMaxSize) return bitmap_to_array(*o);  // :2470-2473 (sic: at ==4096)
    return o;
  }
  auto [i, found] = bin_search_runs(v, c->runs);
  if (!found) return c;
  *removed = true;
  if (c->n == 1) return nullptr;
  auto o = std::make_shared<Container>(*c);
  auto& runs = o->runs;
  if (v == runs[i].last && v == runs[i].start) runs.erase(runs.begin() + i);
  else if (v == runs[i].last) runs[i].last--;
  else if (v == runs[i].start) runs[i].start++;
  else if (v > runs[i].start) {
    uint16_t last = runs[i].last;
    runs[i].last = uint16_t(v - 1);
    runs.insert(runs.begin() + i + 1, Interval{uint16_t(v + 1), last});
  }
  o->n--;
  return o;
}

// ---------------------------------------------------------------------------
// intersectionCount (roaring.go:3078-3215)
// ---------------------------------------------------------------------------

inline int32_t icount_array_array(const Container& a, const Container& b) {  // :3115
  const auto* ca = &a.array; const auto* cb = &b.array;
  if (ca->size() > cb->size()) std::swap(ca, cb);
  size_t nb = cb->size(), j = 0;
  int32_t n = 0;
  for (uint16_t va : *ca) {
    while ((*cb)[j] < va) { if (++j >= nb) return n; }
    if ((*cb)[j] == va) n++;
  }
  return n;
}

inline int32_t icount_array_run(const Container& a, const Container& b) {  // :3138
  const auto& arr = a.array; const auto& runs = b.runs;
  size_t na = arr.size(), nb = runs.size();
  int32_t n = 0;
  for (size_t i = 0, j = 0; i < na && j < nb;) {
    uint16_t va = arr[i]; Interval vb = runs[j];
    if (va < vb.start) i++;
    else if (va <= vb.last) { i++; n++; }
    else j++;
  }
  return n;
}

inline int32_t icount_run_run(const Container& a, const Container& b) {  // :3156
  const auto& ra = a.runs; const auto& rb = b.runs;
  size_t na = ra.size(), nb = rb.size();
  int32_t n = 0;
  for (size_t i = 0, j = 0; i < na && j < nb;) {
    Interval va = ra[i], vb = rb[j];
    if (va.last < vb.start) i++;
    else if (va.start > vb.last) j++;
    else if (va.last > vb.last && va.start >= vb.start) { n += 1 + int32_t(vb.last - va.start); j++; }
    else if (va.last > vb.last && va.start < vb.start) { n += 1 + int32_t(vb.last - vb.start); j++; }
    else if (va.last <= vb.last && va.start >= vb.start) { n += 1 + int32_t(va.last - va.start); i++; }
    else if (va.last <= vb.last && va.start < vb.start) { n += 1 + int32_t(va.last - vb.start); i++; }
  }
  return n;
}

inline int32_t icount_bitmap_run(const Container& a, const Container& b) {  // :3189
  int32_t n = 0;
  for (auto& iv : b.runs) n += bitmap_count_range(a, iv.start, int32_t(iv.last) + 1);
  return n;
}

inline int32_t icount_array_bitmap(const Container& a, const Container& b) {  // :3197
  int32_t n = 0;
  for (uint16_t v : a.array) n += int32_t((b.bitmap[v >> 6] >> (v & 63)) & 1);
  return n;
}

inline int32_t icount_bitmap_bitmap(const Container& a, const Container& b) {  // :3212, :5061
  uint64_t n = 0;
  for (int i = 0; i < kBitmapN; i++) n += popc(a.bitmap[i] & b.bitmap[i]);
  return (int32_t)n;
}

inline int32_t intersection_count(const CPtr& a, const CPtr& b) {  // :3078-3113
  if (N(a) == kFullN) return N(b);
  if (N(b) == kFullN) return N(a);
  if (N(a) == 0 || N(b) == 0) return 0;
  if (a->type == kArray) {
    if (b->type == kArray) return icount_array_array(*a, *b);
    if (b->type == kRun) return icount_array_run(*a, *b);
    return icount_array_bitmap(*a, *b);
  } else if (a->type == kRun) {
    if (b->type == kArray) return icount_array_run(*b, *a);
    if (b->type == kRun) return icount_run_run(*a, *b);
    return icount_bitmap_run(*b, *a);
  } else {
    if (b->type == kArray) return icount_array_bitmap(*b, *a);
    if (b->type == kRun) return icount_bitmap_run(*a, *b);
    return icount_bitmap_bitmap(*a, *b);
  }
}

// ---------------------------------------------------------------------------
// run builder: roaring.go:3597-3619 (runAppendInterval)
// ---------------------------------------------------------------------------
inline int32_t run_append(std::vector<Interval>& runs, Interval v) {
  if (runs.empty()) {
    runs.push_back(v);
    return int32_t(v.last - v.start) + 1;
  }
  Interval last = runs.back();
  if (last.last == kMaxContainerVal) return 0;
  if (int32_t(last.last) + 1 >= int32_t(v.start) && v.last > last.last) {
    runs.back().last = v.last;
    return int32_t(v.last - last.last);
  } else if (int32_t(last.last) + 1 < int32_t(v.start)) {
    runs.push_back(v);
    return int32_t(v.last - v.start) + 1;
  }
  return 0;
}

inline MPtr run_with_n(std::vector<Interval> runs, int32_t n) {
  auto c = Container::make_run(std::move(runs));
  c->n = n;
  return c;
}

// ---------------------------------------------------------------------------
// intersect (roaring.go:3217-3437)
// ---------------------------------------------------------------------------

inline CPtr intersect_array_array(const Container& a, const Container& b) {  // :3254
  std::vector<uint16_t> out;
  const auto& aa = a.array; const auto& ab = b.array;
  for (size_t i = 0, j = 0; i < aa.size() && j < ab.size();) {
    if (aa[i] < ab[j]) i++;
    else if (aa[i] > ab[j]) j++;
    else { out.push_back(aa[i]); i++; j++; }
  }
  return Container::make_array(std::move(out));
}

inline CPtr intersect_array_run(const Container& a, const Container& b) {  // :3276
  std::vector<uint16_t> out;
  const auto& aa = a.array; const auto& rb = b.runs;
  for (size_t i = 0, j = 0; i < aa.size() && j < rb.size();) {
    if (aa[i] < rb[j].start) i++;
    else if (aa[i] > rb[j].last) j++;
    else { out.push_back(aa[i]); i++; }
  }
  return Container::make_array(std::move(out));
}

inline CPtr intersect_run_run(const Container& a, const Container& b) {  // :3296
  std::vector<Interval> out;
  const auto& ra = a.runs; const auto& rb = b.runs;
  int32_t n = 0;
  for (size_t i = 0, j = 0; i < ra.size() && j < rb.size();) {
    Interval va = ra[i], vb = rb[j];
    if (va.last < vb.start) i++;
    else if (vb.last < va.start) j++;
    else if (va.last > vb.last && va.start >= vb.start) { n += run_append(out, {va.start, vb.last}); j++; }
    else if (va.last > vb.last && va.start < vb.start) { n += run_append(out, vb); j++; }
    else if (va.last <= vb.last && va.start >= vb.start) { n += run_append(out, va); i++; }
    else if (va.last <= vb.last && va.start < vb.start) { n += run_append(out, {vb.start, va.last}); i++; }
  }
  size_t nr = out.size();
  auto o = run_with_n(std::move(out), n);
  if (n < kArrayMaxSize && int32_t(nr) > n / 2) return run_to_array(*o);   // :3330
  if ((int32_t)nr > kRunMaxSize) return run_to_bitmap(*o);                  // :3332
  return o;
}

inline CPtr intersect_bitmap_run(const Container& a, const Container& b) {  // :3340
  const auto& runs = b.runs;
  if (b.n <= kArrayMaxSize || a.n <= kArrayMaxSize) {
    std::vector<uint16_t> out;
    for (auto& iv : runs)
      for (int v = iv.start; v <= int(iv.last); v++)
        if (a.contains_bit(uint16_t(v))) out.push_back(uint16_t(v));
    return Container::make_array(std::move(out));
  }
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (auto& vb : runs) {  // word-masked copy; same set as the reference's :3360-3399 loop
    int ws = vb.start >> 6, we = vb.last >> 6;
    for (int w = ws; w <= we; w++) {
      uint64_t mask = ~uint64_t(0);
      if (w == ws) mask &= ~uint64_t(0) << (vb.start & 63);
      if (w == we) mask &= ~uint64_t(0) >> (63 - (vb.last & 63));
      uint64_t bits = a.bitmap[w] & mask;
      o->bitmap[w] |= bits;
      n += popc(bits);
    }
  }
  o->n = n;
  return o;
}

inline CPtr intersect_array_bitmap(const Container& a, const Container& b) {  // :3404
  std::vector<uint16_t> out;
  for (uint16_t v : a.array) if (b.contains_bit(v)) out.push_back(v);
  return Container::make_array(std::move(out));
}

inline CPtr intersect_bitmap_bitmap(const Container& a, const Container& b) {  // :3420 (always bitmap)
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (int i = 0; i < kBitmapN; i++) {
    o->bitmap[i] = a.bitmap[i] & b.bitmap[i];
    n += popc(o->bitmap[i]);
  }
  o->n = n;
  return o;
}

inline CPtr intersect(const CPtr& a, const CPtr& b) {  // :3217-3252
  if (N(a) == kFullN) return b;
  if (N(b) == kFullN) return a;
  if (N(a) == 0 || N(b) == 0) return nullptr;
  if (a->type == kArray) {
    if (b->type == kArray) return intersect_array_array(*a, *b);
    if (b->type == kRun) return intersect_array_run(*a, *b);
    return intersect_array_bitmap(*a, *b);
  } else if (a->type == kRun) {
    if (b->type == kArray) return intersect_array_run(*b, *a);
    if (b->type == kRun) return intersect_run_run(*a, *b);
    return intersect_bitmap_run(*b, *a);
  } else {
    if (b->type == kArray) return intersect_array_bitmap(*b, *a);
    if (b->type == kRun) return intersect_bitmap_run(*a, *b);
    return intersect_bitmap_bitmap(*a, *b);
  }
}

// ---------------------------------------------------------------------------
// bitmap range helpers (roaring.go:3675-3768)
// ---------------------------------------------------------------------------
inline void bitmap_set_range(Container& c, uint64_t i, uint64_t j) {  // :3675 — [i,j)
  auto& b = c.bitmap;
  uint64_t x = i >> 6, y = (j - 1) >> 6;
  uint64_t X = ~uint64_t(0) << (i % 64), Y = ~uint64_t(0) >> (63 - ((j - 1) % 64));
  int32_t n = c.n;
  if (x == y) {
    n += int32_t((j - i) - popc(b[x] & (X & Y)));
    b[x] |= X & Y;
  } else {
    n += popc(X) - popc(b[x] & X);
    b[x] |= X;
    for (uint64_t k = x + 1; k < y; k++) { n += 64 - popc(b[k]); b[k] = ~uint64_t(0); }
    n += popc(Y) - popc(b[y] & Y);
    b[y] |= Y;
  }
  c.n = n;
}

inline void bitmap_xor_range(Container& c, uint64_t i, uint64_t j) {  // :3719
  auto& b = c.bitmap;
  uint64_t x = i >> 6, y = (j - 1) >> 6;
  uint64_t X = ~uint64_t(0) << (i % 64), Y = ~uint64_t(0) >> (63 - ((j - 1) % 64));
  int32_t n = c.n;
  auto flip = [&](uint64_t k, uint64_t m) { int c0 = popc(b[k]); b[k] ^= m; n += popc(b[k]) - c0; };
  if (x == y) flip(x, X & Y);
  else {
    flip(x, X);
    for (uint64_t k = x + 1; k < y; k++) flip(k, ~uint64_t(0));
    flip(y, Y);
  }
  c.n = n;
}

inline void bitmap_zero_range(Container& c, uint64_t i, uint64_t j) {  // :3747
  auto& b = c.bitmap;
  uint64_t x = i >> 6, y = (j - 1) >> 6;
  uint64_t X = ~uint64_t(0) << (i % 64), Y = ~uint64_t(0) >> (63 - ((j - 1) % 64));
  int32_t n = c.n;
  if (x == y) { n -= popc(b[x] & (X & Y)); b[x] &= ~(X & Y); }
  else {
    n -= popc(b[x] & X); b[x] &= ~X;
    for (uint64_t k = x + 1; k < y; k++) { n -= popc(b[k]); b[k] = 0; }
    n -= popc(b[y] & Y); b[y] &= ~Y;
  }
  c.n = n;
}

// ---------------------------------------------------------------------------
// union (roaring.go:3439-3883)
// ---------------------------------------------------------------------------

inline CPtr union_array_array(const Container& a, const Container& b) {  // :3470 (no conversion, :3506)
  std::vector<uint16_t> out(a.array.size() + b.array.size());
  auto e = std::set_union(a.array.begin(), a.array.end(), b.array.begin(), b.array.end(), out.begin());
  out.resize(e - out.begin());
  return Container::make_array(std::move(out));
}

inline CPtr union_array_run(const Container& a, const Container& b) {  // :3559
  std::vector<Interval> out;
  const auto& aa = a.array; const auto& rb = b.runs;
  size_t na = aa.size(), nb = rb.size();
  int32_t n = 0;
  for (size_t i = 0, j = 0; i < na || j < nb;) {
    if (i < na && (j >= nb || aa[i] < rb[j].start)) { n += run_append(out, {aa[i], aa[i]}); i++; }
    else { n += run_append(out, rb[j]); j++; }
  }
  size_t nr = out.size();
  auto o = run_with_n(std::move(out), n);
  if (n < kArrayMaxSize) return run_to_array(*o);            // :3583
  if ((int32_t)nr > kRunMaxSize) return run_to_bitmap(*o);    // :3585
  return o;
}

inline CPtr union_run_run(const Container& a, const Container& b) {  // :3621
  std::vector<Interval> out;
  const auto& ra = a.runs; const auto& rb = b.runs;
  size_t na = ra.size(), nb = rb.size();
  int32_t n = 0;
  for (size_t i = 0, j = 0; i < na || j < nb;) {
    if (i < na && (j >= nb || ra[i].start < rb[j].start)) { n += run_append(out, ra[i]); i++; }
    else { n += run_append(out, rb[j]); j++; }
  }
  size_t nr = out.size();
  auto o = run_with_n(std::move(out), n);
  if ((int32_t)nr > kRunMaxSize) return run_to_bitmap(*o);  // :3644
  return o;
}

inline CPtr union_bitmap_run(const Container& a, const Container& b) {  // :3650
  auto o = std::make_shared<Container>(a);
  for (auto& r : b.runs) bitmap_set_range(*o, r.start, uint64_t(r.last) + 1);
  return o;
}

inline CPtr union_array_bitmap(const Container& a, const Container& b) {  // :3816
  auto o = std::make_shared<Container>(b);
  for (uint16_t v : a.array)
    if (!o->contains_bit(v)) { o->bitmap[v >> 6] |= uint64_t(1) << (v & 63); o->n++; }
  return o;
}

inline CPtr union_bitmap_bitmap(const Container& a, const Container& b) {  // :3841
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (int i = 0; i < kBitmapN; i++) { o->bitmap[i] = a.bitmap[i] | b.bitmap[i]; n += popc(o->bitmap[i]); }
  o->n = n;
  return o;
}

// The Bitmap-level merge only calls union() with two non-nil containers
// (roaring.go:638-660); nil operands are handled here for the test harness the
// way the Bitmap level would (the other side, frozen).
inline CPtr union_(const CPtr& a, const CPtr& b) {  // :3439-3468
  if (N(a) == kFullN || N(b) == kFullN) return full_container();
  if (!a) return b;
  if (!b) return a;
  if (a->type == kArray) {
    if (b->type == kArray) return union_array_array(*a, *b);
    if (b->type == kRun) return union_array_run(*a, *b);
    return union_array_bitmap(*a, *b);
  } else if (a->type == kRun) {
    if (b->type == kArray) return union_array_run(*b, *a);
    if (b->type == kRun) return union_run_run(*a, *b);
    return union_bitmap_run(*b, *a);
  } else {
    if (b->type == kArray) return union_array_bitmap(*b, *a);
    if (b->type == kRun) return union_bitmap_run(*a, *b);
    return union_bitmap_bitmap(*a, *b);
  }
}

// ---------------------------------------------------------------------------
// difference (roaring.go:3885-4230)
// ---------------------------------------------------------------------------

inline CPtr flip_bitmap(const Container& b) {  // :3057
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (int i = 0; i < kBitmapN; i++) { o->bitmap[i] = ~b.bitmap[i]; n += popc(o->bitmap[i]); }
  o->n = n;
  return o;
}

inline CPtr difference_array_array(const Container& a, const Container& b) {  // :3920
  // output.add() appends in ascending order; the result can never exceed
  // len(a) <= ... the reference's arrayAdd converts to bitmap once n reaches
  // 4096 (roaring.go:2163) — reproduce that type rule.
  std::vector<uint16_t> out;
  std::set_difference(a.array.begin(), a.array.end(), b.array.begin(), b.array.end(), std::back_inserter(out));
  auto o = Container::make_array(std::move(out));
  if (o->n > kArrayMaxSize) return array_to_bitmap(*o);
  return o;
}

inline CPtr difference_array_run(const Container& a, const Container& b) {  // :3947
  std::vector<uint16_t> out;
  const auto& aa = a.array; const auto& rb = b.runs;
  size_t i = 0, j = 0;
  while (i < aa.size()) {
    if (aa[i] < rb[j].start) { out.push_back(aa[i]); i++; continue; }
    if (aa[i] <= rb[j].last) { i++; continue; }
    j++;
    if (j == rb.size()) break;
  }
  for (; i < aa.size(); i++) out.push_back(aa[i]);
  return Container::make_array(std::move(out));
}

inline CPtr difference_bitmap_run(const Container& a, const Container& b) {  // :3992 (stays bitmap)
  auto o = std::make_shared<Container>(a);
  for (auto& r : b.runs) bitmap_zero_range(*o, r.start, uint64_t(r.last) + 1);
  return o;
}

inline CPtr difference_run_array(const Container& a, const Container& b) {  // :4003
  std::vector<Interval> runs;
  const auto& ab = b.array;
  size_t bidx = 0;
  uint16_t vb = ab[0];
  bool stop = false;
  for (size_t r = 0; r < a.runs.size() && !stop; r++) {
    Interval run = a.runs[r];
    uint16_t start = run.start;
    while (vb < run.start) {
      if (++bidx >= ab.size()) break;
      vb = ab[bidx];
    }
    while (vb >= run.start && vb <= run.last) {
      if (vb == start) {
        if (vb == 65535) { stop = true; break; }
        start++;
        if (++bidx >= ab.size()) break;
        vb = ab[bidx];
        continue;
      }
      runs.push_back({start, uint16_t(vb - 1)});
      if (vb == 65535) { stop = true; break; }
      start = uint16_t(vb + 1);
      if (++bidx >= ab.size()) break;
      vb = ab[bidx];
    }
    if (stop) break;  // `break RUNLOOP` skips the trailing append (:4022-4036)
    if (start <= run.last) runs.push_back({start, run.last});
  }
  return optimize(Container::make_run(std::move(runs)));  // :4050-4052
}

inline CPtr difference_run_bitmap(const Container& a, const Container& b) {  // :4056
  const auto& ra = a.runs;
  if (!ra.empty() && ra[0].start == 0 && ra[0].last == 65535) return flip_bitmap(b);
  std::vector<Interval> runs;
  for (auto& in : ra) {
    Interval run = in;
    bool add = true;
    for (int bit = in.start; bit <= int(in.last); bit++) {
      if (b.contains_bit(uint16_t(bit))) {
        if (run.start == bit) {
          if (bit == 65535) add = false;
          run.start++;
        } else if (bit == run.last) {
          run.last--;
        } else {
          run.last = uint16_t(bit - 1);
          if (run.last >= run.start) runs.push_back(run);
          run.start = uint16_t(bit + 1);
          run.last = in.last;
        }
        if (run.start > run.last) break;
      }
    }
    if (run.start <= run.last && add) runs.push_back(run);
  }
  size_t nr = runs.size();
  auto o = Container::make_run(std::move(runs));
  if (o->n < kArrayMaxSize && int32_t(nr) > o->n / 2) return run_to_array(*o);  // :4102
  if ((int32_t)nr > kRunMaxSize) return run_to_bitmap(*o);                       // :4104
  return o;
}

inline CPtr difference_run_run(const Container& a, const Container& b) {  // :4111 (no conversion)
  const auto& ra = a.runs; const auto& rb = b.runs;
  size_t apos = 0, bpos = 0, alen = ra.size(), blen = rb.size();
  int32_t astart = ra[0].start, alast = ra[0].last, bstart = rb[0].start, blast = rb[0].last;
  std::vector<Interval> runs;
  while (apos < alen && bpos < blen) {
    if (alast < bstart) {
      runs.push_back({uint16_t(astart), uint16_t(alast)});
      if (++apos < alen) { astart = ra[apos].start; alast = ra[apos].last; }
    } else if (blast < astart) {
      if (++bpos < blen) { bstart = rb[bpos].start; blast = rb[bpos].last; }
    } else {
      if (astart < bstart) runs.push_back({uint16_t(astart), uint16_t(bstart - 1)});
      if (alast > blast) astart = blast + 1;
      else if (++apos < alen) { astart = ra[apos].start; alast = ra[apos].last; }
    }
  }
  if (apos < alen) {
    runs.push_back({uint16_t(astart), uint16_t(alast)});
    for (apos++; apos < alen; apos++) runs.push_back(ra[apos]);
  }
  return Container::make_run(std::move(runs));
}

inline CPtr difference_array_bitmap(const Container& a, const Container& b) {  // :4171
  std::vector<uint16_t> out;
  for (uint16_t v : a.array) if (!b.contains_bit(v)) out.push_back(v);
  return Container::make_array(std::move(out));
}

inline CPtr difference_bitmap_array(const Container& a, const Container& b) {  // :4188
  auto o = std::make_shared<Container>(a);
  for (uint16_t v : b.array)
    if (o->contains_bit(v)) { o->bitmap[v >> 6] &= ~(uint64_t(1) << (v & 63)); o->n--; }
  if (o->n < kArrayMaxSize) return bitmap_to_array(*o);  // :4201
  return o;
}

inline CPtr difference_bitmap_bitmap(const Container& a, const Container& b) {  // :4207
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (int i = 0; i < kBitmapN; i++) { o->bitmap[i] = a.bitmap[i] & ~b.bitmap[i]; n += popc(o->bitmap[i]); }
  o->n = n;
  if (n < kArrayMaxSize) return bitmap_to_array(*o);  // :4226
  return o;
}

inline CPtr difference(const CPtr& a, const CPtr& b) {  // :3885-3918
  if (N(a) == 0 || N(b) == kFullN) return nullptr;
  if (N(b) == 0) return a;
  if (a->type == kArray) {
    if (b->type == kArray) return difference_array_array(*a, *b);
    if (b->type == kRun) return difference_array_run(*a, *b);
    return difference_array_bitmap(*a, *b);
  } else if (a->type == kRun) {
    if (b->type == kArray) return difference_run_array(*a, *b);
    if (b->type == kRun) return difference_run_run(*a, *b);
    return difference_run_bitmap(*a, *b);
  } else {
    if (b->type == kArray) return difference_bitmap_array(*a, *b);
    if (b->type == kRun) return difference_bitmap_run(*a, *b);
    return difference_bitmap_bitmap(*a, *b);
  }
}

// ---------------------------------------------------------------------------
// xor (roaring.go:4232-4340, 4739-4957)
// ---------------------------------------------------------------------------

inline CPtr xor_array_array(const Container& a, const Container& b) {  // :4266 (no conversion)
  std::vector<uint16_t> out;
  std::set_symmetric_difference(a.array.begin(), a.array.end(), b.array.begin(), b.array.end(),
                                std::back_inserter(out));
  return Container::make_array(std::move(out));
}

inline CPtr xor_array_bitmap(const Container& a, const Container& b) {  // :4297
  // Per-element add/remove on a clone of b; the container type can flip at the
  // 4096 boundary in both directions (bitmapRemove :2470, arrayAdd :2163).
  CPtr o = std::make_shared<Container>(b);
  bool ch;
  for (uint16_t v : a.array) {
    if (b.contains_bit(v)) o = remove(o, v, &ch);
    else o = add(o, v, &ch);
  }
  if (o && o->type == kBitmap) {
    int32_t cnt = 0;
    for (uint64_t w : o->bitmap) cnt += popc(w);
    if (cnt < kArrayMaxSize) return bitmap_to_array(*o);  // :4310
  }
  return o;
}

inline CPtr xor_bitmap_bitmap(const Container& a, const Container& b) {  // :4317
  auto o = Container::make_bitmap();
  int32_t n = 0;
  for (int i = 0; i < kBitmapN; i++) { o->bitmap[i] = a.bitmap[i] ^ b.bitmap[i]; n += popc(o->bitmap[i]); }
  o->n = n;
  if (n < kArrayMaxSize) return bitmap_to_array(*o);  // :4336
  return o;
}

inline CPtr xor_array_run(const Container& a, const Container& b) {  // :4739
  std::vector<Interval> out;
  const auto& aa = a.array; const auto& rb = b.runs;
  int na = (int)aa.size(), nb = (int)rb.size();
  Interval vb{0, 0};
  uint16_t va = 0;
  int lastI = -1, lastJ = -1;
  int32_t n = 0;
  for (int i = 0, j = 0; i < na || j < nb;) {
    if (i < na && i != lastI) va = aa[i];
    if (j < nb && j != lastJ) vb = rb[j];
    lastI = i; lastJ = j;
    if (i < na && (j >= nb || va < vb.start)) { n += run_append(out, {va, va}); i++; }
    else if (j < nb && (i >= na || va > vb.last)) { n += run_append(out, vb); j++; }
    else if (va > vb.start) {
      if (va < vb.last) {
        n += run_append(out, {vb.start, uint16_t(va - 1)});
        i++;
        vb.start = uint16_t(va + 1);
        if (vb.start > vb.last) j++;
      } else if (va > vb.last) {
        n += run_append(out, vb); j++;
      } else {
        vb.last--;
        if (vb.start <= vb.last) n += run_append(out, vb);
        j++; i++;
      }
    } else {  // va == vb.start
      if (vb.start == kMaxContainerVal) j++;
      else { vb.start++; if (vb.start > vb.last) j++; }
      i++;
    }
  }
  size_t nr = out.size();
  auto o = run_with_n(std::move(out), n);
  if (n < kArrayMaxSize) return run_to_array(*o);            // :4798
  if ((int32_t)nr > kRunMaxSize) return run_to_bitmap(*o);    // :4800
  return o;
}

struct XorState { bool va_valid = false, vb_valid = false; Interval va{0, 0}, vb{0, 0}; };  // :4895

inline bool xor_compare(XorState& x, Interval& r1) {  // :4807-4892
  if (!x.va_valid || !x.vb_valid) {
    if (x.vb_valid) { x.vb_valid = false; r1 = x.vb; return true; }
    if (x.va_valid) { x.va_valid = false; r1 = x.va; return true; }
    return false;
  }
  bool has = false;
  if (x.va.last < x.vb.start) { x.va_valid = false; r1 = x.va; has = true; }
  else if (x.vb.last < x.va.start) { x.vb_valid = false; r1 = x.vb; has = true; }
  else if (x.va.start == x.vb.start && x.va.last == x.vb.last) { x.va_valid = x.vb_valid = false; }
  else if (x.va.start <= x.vb.start && x.va.last >= x.vb.last) {  // vb inside
    x.vb_valid = false;
    if (x.va.start != x.vb.start) { r1 = {x.va.start, uint16_t(x.vb.start - 1)}; has = true; }
    if (x.vb.last == kMaxContainerVal) x.va_valid = false;
    else { x.va.start = uint16_t(x.vb.last + 1); if (x.va.start > x.va.last) x.va_valid = false; }
  } else if (x.vb.start <= x.va.start && x.vb.last >= x.va.last) {  // va inside
    x.va_valid = false;
    if (x.vb.start != x.va.start) { r1 = {x.vb.start, uint16_t(x.va.start - 1)}; has = true; }
    if (x.va.last == kMaxContainerVal) x.vb_valid = false;
    else { x.vb.start = uint16_t(x.va.last + 1); if (x.vb.start > x.vb.last) x.vb_valid = false; }
  } else if (x.va.start < x.vb.start && x.va.last <= x.vb.last) {  // va first overlap
    x.va_valid = false;
    r1 = {x.va.start, uint16_t(x.vb.start - 1)}; has = true;
    if (x.va.last == kMaxContainerVal) x.vb_valid = false;
    else { x.vb.start = uint16_t(x.va.last + 1); if (x.vb.start > x.vb.last) x.vb_valid = false; }
  } else if (x.vb.start < x.va.start && x.vb.last <= x.va.last) {  // vb first overlap
    x.vb_valid = false;
    r1 = {x.vb.start, uint16_t(x.va.start - 1)}; has = true;
    if (x.vb.last == kMaxContainerVal) x.va_valid = false;
    else { x.va.start = uint16_t(x.vb.last + 1); if (x.va.start > x.va.last) x.va_valid = false; }
  }
  return has;
}

inline CPtr xor_run_run(const Container& a, const Container& b) {  // :4901
  const auto& ra = a.runs; const auto& rb = b.runs;
  int na = (int)ra.size(), nb = (int)rb.size();
  std::vector<Interval> out;
  int lastI = -1, lastJ = -1;
  XorState st;
  int32_t n = 0;
  for (int i = 0, j = 0; i < na || j < nb;) {
    if (i < na && lastI != i) { st.va = ra[i]; st.va_valid = true; }
    if (j < nb && lastJ != j) { st.vb = rb[j]; st.vb_valid = true; }
    lastI = i; lastJ = j;
    Interval r1{0, 0};
    if (xor_compare(st, r1)) n += run_append(out, r1);
    if (!st.va_valid) i++;
    if (!st.vb_valid) j++;
  }
  size_t l = out.size();
  auto o = run_with_n(std::move(out), n);
  if (n < kArrayMaxSize && int32_t(l) > n / 2) return run_to_array(*o);
  if ((int32_t)l > kRunMaxSize) return run_to_bitmap(*o);
  return o;
}

inline CPtr xor_bitmap_run(const Container& a, const Container& b) {  // :4948
  auto o = std::make_shared<Container>(a);
  for (auto& r : b.runs) bitmap_xor_range(*o, r.start, uint64_t(r.last) + 1);
  return o;
}

inline CPtr xor_(const CPtr& a, const CPtr& b) {  // :4232-4264
  if (N(a) == 0) return b;
  if (N(b) == 0) return a;
  if (a->type == kArray) {
    if (b->type == kArray) return xor_array_array(*a, *b);
    if (b->type == kRun) return xor_array_run(*a, *b);
    return xor_array_bitmap(*a, *b);
  } else if (a->type == kRun) {
    if (b->type == kArray) return xor_array_run(*b, *a);
    if (b->type == kRun) return xor_run_run(*a, *b);
    return xor_bitmap_run(*b, *a);
  } else {
    if (b->type == kArray) return xor_array_bitmap(*b, *a);
    if (b->type == kRun) return xor_bitmap_run(*a, *b);
    return xor_bitmap_bitmap(*a, *b);
  }
}

// flip (roaring.go:3039-3076): used by TestContainerCombinations.
inline CPtr flip(const CPtr& a) {
  if (a->type == kArray) return flip_bitmap(*array_to_bitmap(*a));
  if (a->type == kRun) return flip_bitmap(*run_to_bitmap(*a));
  return flip_bitmap(*a);
}

// ---------------------------------------------------------------------------
// shift by one (roaring.go:4345-4413)
// ---------------------------------------------------------------------------
inline CPtr shift(const CPtr& c, bool* carry) {
  *carry = false;
  if (N(c) == 0) return nullptr;
  if (c->type == kArray) {  // :4358
    std::vector<uint16_t> out;
    for (uint16_t v : c->array) {
      if (uint16_t(v + 1) == 0) *carry = true; else out.push_back(uint16_t(v + 1));
    }
    return Container::make_array(std::move(out));
  }
  if (c->type == kRun) {  // :4391
    std::vector<Interval> out;
    for (Interval v : c->runs) {
      if (uint16_t(v.start + 1) == 0) { *carry = true; break; }
      else if (uint16_t(v.last + 1) == 0) { v.start++; *carry = true; }
      else { v.start++; v.last++; *carry = false; }
      out.push_back(v);
    }
    return Container::make_run(std::move(out));
  }
  auto o = Container::make_bitmap();  // :4374
  uint64_t last_carry = 0, cy = 0;
  for (int i = 0; i < kBitmapN; i++) {
    uint64_t v = c->bitmap[i];
    cy = v >> 63;
    o->bitmap[i] = (v << 1) | last_carry;
    last_carry = cy;
  }
  o->n = c->n - int32_t(cy);
  *carry = cy != 0;
  return o;
}

// ---------------------------------------------------------------------------
// Bitmap: ordered key -> container map (roaring.go:145-160; Containers iface
// :80-137). Iteration skips nil values like containers_slice.go:239-254.
// ---------------------------------------------------------------------------
struct Bitmap {
  std::map<uint64_t, CPtr> cs;
  uint8_t flags = 0;
  int ops = 0, op_n = 0;

  void put(uint64_t key, CPtr c) { cs[key] = std::move(c); }
  CPtr get(uint64_t key) const { auto it = cs.find(key); return it == cs.end() ? nullptr : it->second; }

  uint64_t count() const {  // :405
    uint64_t n = 0;
    for (auto& kv : cs) n += N(kv.second);
    return n;
  }
  bool any() const {  // :410
    for (auto& kv : cs) if (N(kv.second) > 0) return true;
    return false;
  }
  bool direct_add(uint64_t v) {  // roaring.go DirectAdd
    bool added;
    uint64_t k = v >> 16;
    CPtr c = get(k);
    CPtr o = add(c, uint16_t(v & 0xffff), &added);
    if (added) cs[k] = o;
    return added;
  }
  bool direct_remove(uint64_t v) {
    bool removed;
    uint64_t k = v >> 16;
    CPtr c = get(k);
    if (!c) return false;
    CPtr o = remove(c, uint16_t(v & 0xffff), &removed);
    if (removed) { if (o) cs[k] = o; else cs.erase(k); }
    return removed;
  }
  bool contains_value(uint64_t v) const { return contains(get(v >> 16), uint16_t(v & 0xffff)); }

  std::vector<uint64_t> slice() const {
    std::vector<uint64_t> out;
    for (auto& kv : cs)
      for (uint16_t v : values_of(kv.second)) out.push_back((kv.first << 16) | v);
    return out;
  }
  void optimize_all() {  // :1009-1013; nil results stay stored and are skipped by iteration
    for (auto it = cs.begin(); it != cs.end();) {
      it->second = optimize(it->second);
      if (!it->second) it = cs.erase(it); else ++it;
    }
  }
};

// roaring.go:535-558 — extract [start,end) re-keyed at offset (all multiples of 2^16).
inline Bitmap offset_range(const Bitmap& b, uint64_t offset, uint64_t start, uint64_t end) {
  uint64_t off = offset >> 16, hi0 = start >> 16, hi1 = end >> 16;
  Bitmap o;
  for (auto it = b.cs.lower_bound(hi0); it != b.cs.end() && it->first < hi1; ++it)
    if (it->second) o.cs.emplace_hint(o.cs.end(), off + (it->first - hi0), it->second);
  return o;
}

// roaring.go:568-590
inline uint64_t intersection_count(const Bitmap& a, const Bitmap& b) {
  uint64_t n = 0;
  auto i = a.cs.begin(), j = b.cs.begin();
  while (i != a.cs.end() && j != b.cs.end()) {
    if (i->first < j->first) ++i;
    else if (i->first > j->first) ++j;
    else { n += (uint64_t)intersection_count(i->second, j->second); ++i; ++j; }
  }
  return n;
}

// roaring.go:593-615
inline Bitmap intersect(const Bitmap& a, const Bitmap& b) {
  Bitmap o;
  auto i = a.cs.begin(), j = b.cs.begin();
  while (i != a.cs.end() && j != b.cs.end()) {
    if (i->first < j->first) ++i;
    else if (i->first > j->first) ++j;
    else { o.cs.emplace_hint(o.cs.end(), i->first, intersect(i->second, j->second)); ++i; ++j; }
  }
  return o;
}

// roaring.go:618-660 (single-operand path: unionIntoTargetSingle)
inline Bitmap union_(const Bitmap& a, const Bitmap& b) {
  Bitmap o;
  auto i = a.cs.begin(), j = b.cs.begin();
  while (i != a.cs.end() || j != b.cs.end()) {
    bool hi = i != a.cs.end(), hj = j != b.cs.end();
    if (hi && (!hj || i->first < j->first)) { o.cs.emplace_hint(o.cs.end(), i->first, i->second); ++i; }
    else if (hj && (!hi || i->first > j->first)) { o.cs.emplace_hint(o.cs.end(), j->first, j->second); ++j; }
    else { o.cs.emplace_hint(o.cs.end(), i->first, union_(i->second, j->second)); ++i; ++j; }
  }
  return o;
}

// roaring.go:737-886 n-way union: per key, union every operand's container into
// one target; a full operand short-circuits to the shared full run container;
// targets expected to hold >= 512 bits are widened to a bitmap first
// (:810-842); cardinalities are repaired once at the end (:885). The result is
// set-identical to folding union_() pairwise.
inline Bitmap union_many(const Bitmap& a, const std::vector<const Bitmap*>& others) {
  Bitmap o = a;
  std::map<uint64_t, std::vector<CPtr>> by_key;
  for (auto* b : others)
    for (auto& kv : b->cs) if (kv.second) by_key[kv.first].push_back(kv.second);
  for (auto& [key, list] : by_key) {
    CPtr t = o.get(key);
    if (N(t) == kFullN) continue;
    bool full = false;
    int64_t expected = N(t);
    for (auto& c : list) { if (c->n == kFullN) full = true; expected += c->n; }
    if (full) { o.cs[key] = full_container(); continue; }
    size_t first = 0;
    if (!t) {
      if (list.size() == 1) { o.cs[key] = list[0]; continue; }
      if (expected >= 512 && list[0]->type != kBitmap) t = Container::make_bitmap();
      else { t = list[0]; first = 1; }
    } else if (expected >= 512 && t->type != kBitmap) {
      t = t->type == kArray ? CPtr(array_to_bitmap(*t)) : CPtr(run_to_bitmap(*t));
    }
    for (size_t k = first; k < list.size(); k++) {  // Container.unionInPlace :2347-2397
      const CPtr& c = list[k];
      if (t->type == kArray && c->type == kArray) { t = optimize(union_array_array(*t, *c)); continue; }  // :3512-3556
      MPtr m = t->type == kBitmap ? std::make_shared<Container>(*t)
             : t->type == kArray ? array_to_bitmap(*t) : run_to_bitmap(*t);
      if (c->type == kBitmap) for (int w = 0; w < kBitmapN; w++) m->bitmap[w] |= c->bitmap[w];
      else if (c->type == kArray) for (uint16_t v : c->array) m->bitmap[v >> 6] |= uint64_t(1) << (v & 63);
      else for (auto& r : c->runs) { m->n = 0; bitmap_set_range(*m, r.start, uint64_t(r.last) + 1); }
      int32_t n = 0;
      for (uint64_t w : m->bitmap) n += popc(w);  // Repair :3002-3026
      m->n = n;
      t = m;
    }
    o.cs[key] = t;
  }
  return o;
}

// roaring.go:889-913
inline Bitmap difference(const Bitmap& a, const Bitmap& b) {
  Bitmap o;
  auto i = a.cs.begin(), j = b.cs.begin();
  while (i != a.cs.end() || j != b.cs.end()) {
    bool hi = i != a.cs.end(), hj = j != b.cs.end();
    if (hi && (!hj || i->first < j->first)) { o.cs.emplace_hint(o.cs.end(), i->first, i->second); ++i; }
    else if (hj && (!hi || i->first > j->first)) { ++j; }
    else { o.cs.emplace_hint(o.cs.end(), i->first, difference(i->second, j->second)); ++i; ++j; }
  }
  return o;
}

// roaring.go:916-941
inline Bitmap xor_(const Bitmap& a, const Bitmap& b) {
  Bitmap o;
  auto i = a.cs.begin(), j = b.cs.begin();
  while (i != a.cs.end() || j != b.cs.end()) {
    bool hi = i != a.cs.end(), hj = j != b.cs.end();
    if (hi && (!hj || i->first < j->first)) { o.cs.emplace_hint(o.cs.end(), i->first, i->second); ++i; }
    else if (hj && (!hi || i->first > j->first)) { o.cs.emplace_hint(o.cs.end(), j->first, j->second); ++j; }
    else { o.cs.emplace_hint(o.cs.end(), i->first, xor_(i->second, j->second)); ++i; ++j; }
  }
  return o;
}

// roaring.go:944-977 — shift by one with carry into the next container key.
inline Bitmap shift1(const Bitmap& b) {
  Bitmap o;
  bool last_carry = false;
  uint64_t last_key = 0;
  bool ch;
  for (auto& kv : b.cs) {
    if (!kv.second) continue;
    uint64_t ki = kv.first;
    if (last_carry && ki > last_key + 1) {
      o.put(last_key + 1, Container::make_array({0}));
      last_carry = false;
    }
    bool carry;
    CPtr s = shift(kv.second, &carry);
    if (last_carry) s = add(s, 0, &ch);
    if (N(s) > 0) o.put(ki, s);
    last_carry = carry;
    last_key = ki;
  }
  if (last_carry && last_key != kMaxContainerKey) o.put(last_key + 1, Container::make_array({0}));
  return o;
}

// ---------------------------------------------------------------------------
// Pilosa roaring file format (roaring.go:1044-1122 writer, :1562-1653 reader,
// container payloads :2867-2931, ops log :4459-4559). Little endian.
// ---------------------------------------------------------------------------
inline uint32_t fnv32a(const uint8_t* p, size_t n, uint32_t h = 2166136261u) {
  for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; }
  return h;
}

inline size_t container_size(const Container& c) {  // :2917-2925
  if (c.type == kArray) return c.array.size() * 2;
  if (c.type == kRun) return c.runs.size() * 4 + 2;
  return kBitmapN * 8;
}

template <class T> inline void put_le(std::vector<uint8_t>& out, T v) {
  for (size_t i = 0; i < sizeof(T); i++) out.push_back(uint8_t(v >> (8 * i)));
}

// writeToUnoptimized (:1052-1122); set do_optimize for WriteTo (:1044-1047).
inline std::vector<uint8_t> serialize(Bitmap b, bool do_optimize = true) {
  if (do_optimize) b.optimize_all();
  std::vector<uint8_t> out;
  uint32_t count = 0;
  for (auto& kv : b.cs) if (N(kv.second) > 0) count++;
  put_le<uint32_t>(out, kMagicNumber | (uint32_t(b.flags) << 24));
  put_le<uint32_t>(out, count);
  for (auto& kv : b.cs) {
    if (N(kv.second) <= 0) continue;
    put_le<uint64_t>(out, kv.first);
    put_le<uint16_t>(out, uint16_t(kv.second->type));
    put_le<uint16_t>(out, uint16_t(kv.second->n - 1));
  }
  uint32_t offset = 8 + count * 16;
  for (auto& kv : b.cs) {
    if (N(kv.second) <= 0) continue;
    put_le<uint32_t>(out, offset);
    offset += (uint32_t)container_size(*kv.second);
  }
  for (auto& kv : b.cs) {
    if (N(kv.second) <= 0) continue;
    const Container& c = *kv.second;
    if (c.type == kArray) for (uint16_t v : c.array) put_le<uint16_t>(out, v);
    else if (c.type == kBitmap) for (uint64_t w : c.bitmap) put_le<uint64_t>(out, w);
    else {
      put_le<uint16_t>(out, uint16_t(c.runs.size()));
      for (auto& r : c.runs) { put_le<uint16_t>(out, r.start); put_le<uint16_t>(out, r.last); }
    }
  }
  return out;
}

// op.WriteTo (:4459-4503) for add (0) / remove (1) / addBatch (2) / removeBatch (3).
inline void append_op(std::vector<uint8_t>& out, uint8_t typ, const std::vector<uint64_t>& values) {
  std::vector<uint8_t> buf;
  buf.push_back(typ);
  if (typ <= 1) put_le<uint64_t>(buf, values.at(0));
  else put_le<uint64_t>(buf, (uint64_t)values.size());
  put_le<uint32_t>(buf, 0);  // checksum placeholder
  if (typ >= 2) for (uint64_t v : values) put_le<uint64_t>(buf, v);
  uint32_t h = fnv32a(buf.data(), 9);
  h = fnv32a(buf.data() + 13, buf.size() - 13, h);
  for (int i = 0; i < 4; i++) buf[9 + i] = uint8_t(h >> (8 * i));
  out.insert(out.end(), buf.begin(), buf.end());
}

template <class T> inline T get_le(const uint8_t* p) {
  T v = 0;
  for (size_t i = 0; i < sizeof(T); i++) v |= T(p[i]) << (8 * i);
  return v;
}

Bitmap deserialize(const uint8_t* data, size_t len);

Bitmap unmarshal_official(const uint8_t* data, size_t len);

// Bitmap.UnmarshalBinary (roaring.go:5133-5179): Pilosa format by its magic number, else the
// official roaring format.
inline Bitmap unmarshal_binary(const uint8_t* data, size_t len) {
  if (len >= 2 && get_le<uint16_t>(data) == kMagicNumber) return deserialize(data, len);
  return unmarshal_official(data, len);
}

// ImportRoaringBits (roaring.go:1463-1558): set or clear every bit of a serialized bitmap (either
// format); returns the number of bits that changed. `row_set`, when given, receives the signed
// change per row of `row_size` containers (the rowSet result, :1484, :1498, :1514, :1524).
inline int64_t import_roaring_bits(Bitmap& b, const uint8_t* data, size_t len, bool clear,
                                   std::map<uint64_t, int64_t>* row_set = nullptr, uint64_t row_size = 16) {
  Bitmap other = unmarshal_binary(data, len);
  int64_t changed = 0;
  for (auto& kv : other.cs) {
    if (!kv.second) continue;
    CPtr cur = b.get(kv.first);
    const int64_t before = N(cur);
    if (clear && before == 0) continue;  // :1493-1495
    CPtr res = clear ? difference(cur, kv.second) : (cur ? union_(cur, kv.second) : kv.second);
    const int64_t after = N(res);
    changed += clear ? before - after : after - before;
    if (row_set && after != before) (*row_set)[kv.first / row_size] += after - before;
    if (after > 0) b.cs[kv.first] = res; else b.cs.erase(kv.first);
  }
  return changed;
}

// readOfficialHeader + readOffsets / readWithRuns (roaring.go:5075-5245): the standard 32-bit roaring
// serialization (16-bit container keys). Error texts are the reference's; reads that would panic in Go
// (slice bounds) are reported as "... truncated".
inline Bitmap unmarshal_official(const uint8_t* data, size_t len) {
  constexpr uint32_t kCookieNoRuns = 12346, kCookieRuns = 12347;
  Bitmap b;
  if (len < 8) throw std::runtime_error("reading roaring header: buffer too small, expecting at least 8 bytes, was " + std::to_string(len));
  const uint32_t cookie = get_le<uint32_t>(data);
  size_t pos = 4;
  uint32_t size = 0;
  bool have_runs = false;
  const uint8_t* is_run = nullptr;
  if (cookie == kCookieNoRuns) {
    size = get_le<uint32_t>(data + pos);
    pos += 4;
  } else if ((cookie & 0xffff) == kCookieRuns) {
    have_runs = true;
    size = uint32_t(uint16_t(cookie >> 16)) + 1;
    const size_t nb = (size + 7) / 8;
    if (pos + nb > len) throw std::runtime_error("reading roaring header: malformed bitmap, is-run bitmap overruns buffer at " + std::to_string(pos + nb));
    is_run = data + pos;
    pos += nb;
  } else {
    throw std::runtime_error("reading roaring header: did not find expected serialCookie in header");
  }
  const size_t header = pos;
  if (size > (1u << 16)) throw std::runtime_error("reading roaring header: it is logically impossible to have more than (1<<16) containers");
  if (pos + 4 * size_t(size) >= len)
    throw std::runtime_error("reading roaring header: malformed bitmap, key-cardinality slice overruns buffer at " + std::to_string(pos + 4 * size_t(size)));
  pos += 4 * size_t(size);
  struct Hdr { uint64_t key; int type; int32_t card; };
  std::vector<Hdr> hdr(size);
  for (uint32_t i = 0; i < size; i++) {
    const uint8_t* p = data + header + size_t(i) * 4;
    const int32_t card = int32_t(get_le<uint16_t>(p + 2)) + 1;
    int type = card < 4096 ? kArray : kBitmap;  // ArrayMaxSize, :5091-5097
    if (is_run && (is_run[i / 8] & (1u << (i % 8)))) type = kRun;
    hdr[i] = {get_le<uint16_t>(p), type, card};
  }
  auto put_array = [&](const Hdr& h, size_t off) {
    if (off + size_t(h.card) * 2 > len) throw std::runtime_error("array payload truncated");
    auto c = std::make_shared<Container>();
    c->type = kArray; c->n = h.card; c->array.resize(h.card);
    for (int32_t k = 0; k < h.card; k++) c->array[k] = get_le<uint16_t>(data + off + size_t(k) * 2);
    b.cs[h.key] = c;
  };
  auto put_bitmap = [&](const Hdr& h, size_t off) {
    if (off + kBitmapN * 8 > len) throw std::runtime_error("bitmap payload truncated");
    auto c = std::make_shared<Container>();
    c->type = kBitmap; c->n = h.card; c->bitmap.resize(kBitmapN);
    std::memcpy(c->bitmap.data(), data + off, kBitmapN * 8);
    b.cs[h.key] = c;
  };
  if (!have_runs) {  // readOffsets :5181-5211
    for (uint32_t i = 0; i < size; i++) {
      if (pos + size_t(i) * 4 + 4 > len) throw std::runtime_error("reading offsets from official roaring format: offset incomplete: len=" + std::to_string(len - pos - size_t(i) * 4));
      const uint32_t off = get_le<uint32_t>(data + pos + size_t(i) * 4);
      if (off >= len) throw std::runtime_error("reading offsets from official roaring format: offset out of bounds: off=" + std::to_string(off) + ", len=" + std::to_string(len));
      if (hdr[i].type == kArray) put_array(hdr[i], off); else put_bitmap(hdr[i], off);
    }
  } else {           // readWithRuns :5213-5245: payloads back to back, no offset section
    if (len < pos + 2) throw std::runtime_error("reading offsets from official roaring format: offset incomplete: len=" + std::to_string(len));
    for (uint32_t i = 0; i < size; i++) {
      const Hdr& h = hdr[i];
      if (h.type == kRun) {
        if (pos + 2 > len) throw std::runtime_error("run header truncated");
        const uint16_t rc = get_le<uint16_t>(data + pos);
        if (pos + 2 + size_t(rc) * 4 > len) throw std::runtime_error("run payload truncated");
        auto c = std::make_shared<Container>();
        c->type = kRun; c->n = h.card; c->runs.resize(rc);
        for (uint16_t k = 0; k < rc; k++) {  // start:length -> start:last
          const uint16_t st = get_le<uint16_t>(data + pos + 2 + size_t(k) * 4), ln = get_le<uint16_t>(data + pos + 4 + size_t(k) * 4);
          c->runs[k] = {st, uint16_t(st + ln)};
        }
        b.cs[h.key] = c;
        pos += 2 + size_t(rc) * 4;
      } else if (h.type == kArray) {
        put_array(h, pos);
        pos += size_t(h.card) * 2;
      } else {
        put_bitmap(h, pos);
        pos += kBitmapN * 8;
      }
    }
  }
  return b;
}

// unmarshalPilosaRoaring (:1562-1653) — throws std::runtime_error with the
// reference's error text on malformed input.
inline Bitmap deserialize(const uint8_t* data, size_t len) {
  Bitmap b;
  if (len < 8) throw std::runtime_error("data too small");
  uint32_t magic = get_le<uint16_t>(data), version = data[2];
  b.flags = data[3];
  if (magic != kMagicNumber) throw std::runtime_error("invalid roaring file, magic number " + std::to_string(magic) + " is incorrect");
  if (version != 0) throw std::runtime_error("wrong roaring version, file is v" + std::to_string(version) + ", server requires v0");
  uint32_t key_n = get_le<uint32_t>(data + 4);
  if (len < 8 + uint64_t(key_n) * 12) throw std::runtime_error("malformed bitmap, key-cardinality not provided");
  struct Hdr { uint64_t key; uint16_t typ; int32_t n; };
  std::vector<Hdr> hdr(key_n);
  for (uint32_t i = 0; i < key_n; i++) {
    const uint8_t* p = data + 8 + size_t(i) * 12;
    hdr[i] = {get_le<uint64_t>(p), get_le<uint16_t>(p + 8), int32_t(get_le<uint16_t>(p + 10)) + 1};
  }
  size_t ops_off = 8 + size_t(key_n) * 12;
  if (len < ops_off + size_t(key_n) * 4) throw std::runtime_error("offset section truncated");
  const uint8_t* offs = data + ops_off;
  for (uint32_t i = 0; i < key_n; i++) {
    uint32_t off = get_le<uint32_t>(offs + size_t(i) * 4);
    if (off >= len) throw std::runtime_error("offset out of bounds: off=" + std::to_string(off) + ", len=" + std::to_string(len));
    const Hdr& h = hdr[i];
    auto c = std::make_shared<Container>();
    c->n = h.n;
    if (h.typ == kRun) {
      if (off + 2 > len) throw std::runtime_error("run header truncated");
      uint16_t rc = get_le<uint16_t>(data + off);
      if (off + 2 + size_t(rc) * 4 > len) throw std::runtime_error("run payload truncated");
      c->type = kRun;
      c->runs.resize(rc);
      for (uint16_t k = 0; k < rc; k++)
        c->runs[k] = {get_le<uint16_t>(data + off + 2 + k * 4), get_le<uint16_t>(data + off + 4 + k * 4)};
      ops_off = off + 2 + size_t(rc) * 4;
    } else if (h.typ == kArray) {
      if (off + size_t(h.n) * 2 > len) throw std::runtime_error("array payload truncated");
      c->type = kArray;
      c->array.resize(h.n);
      for (int32_t k = 0; k < h.n; k++) c->array[k] = get_le<uint16_t>(data + off + k * 2);
      ops_off = off + size_t(h.n) * 2;
    } else if (h.typ == kBitmap) {
      if (off + kBitmapN * 8 > len) throw std::runtime_error("bitmap payload truncated");
      c->type = kBitmap;
      c->bitmap.resize(kBitmapN);
      std::memcpy(c->bitmap.data(), data + off, kBitmapN * 8);
      ops_off = off + kBitmapN * 8;
    } else {
      throw std::runtime_error("unknown container type " + std::to_string(h.typ));
    }
    b.cs[h.key] = c;
  }
  if (key_n == 0) ops_off = 8;
  // ops log
  const uint8_t* buf = data + ops_off;
  size_t rem = len - ops_off;
  while (rem > 0) {
    if (rem < 13) throw std::runtime_error("op data out of bounds: len=" + std::to_string(rem));
    uint8_t typ = buf[0];
    uint64_t value = get_le<uint64_t>(buf + 1);
    uint32_t chk = get_le<uint32_t>(buf + 9);
    uint32_t h = fnv32a(buf, 9);
    size_t size = 13;
    if (typ <= 1) {
      // single add / remove
    } else if (typ <= 3) {
      if (value > (uint64_t(1) << 59)) throw std::runtime_error("maximum operation size exceeded");
      if (rem < 13 + value * 8) throw std::runtime_error("op data truncated");
      h = fnv32a(buf + 13, value * 8, h);
      size = 13 + value * 8;
    } else if (typ <= 5) {
      if (rem < 13 + 4 + value) throw std::runtime_error("op data truncated");
      h = fnv32a(buf + 13, 4 + value, h);
      size = 17 + value;
    } else {
      throw std::runtime_error("unknown op type: " + std::to_string(typ));
    }
    if (chk != h) throw std::runtime_error("checksum mismatch");
    int cnt = 0;
    switch (typ) {
      case 0: b.direct_add(value); cnt = 1; break;
      case 1: b.direct_remove(value); cnt = 1; break;
      case 2: for (uint64_t k = 0; k < value; k++) b.direct_add(get_le<uint64_t>(buf + 13 + k * 8)); cnt = (int)value; break;
      case 3: for (uint64_t k = 0; k < value; k++) b.direct_remove(get_le<uint64_t>(buf + 13 + k * 8)); cnt = (int)value; break;
      case 4: import_roaring_bits(b, buf + 17, value, false); cnt = (int)get_le<uint32_t>(buf + 13); break;
      case 5: import_roaring_bits(b, buf + 17, value, true); cnt = (int)get_le<uint32_t>(buf + 13); break;
    }
    b.ops++;
    b.op_n += cnt;
    buf += size;
    rem -= size;
  }
  return b;
}

// ---------------------------------------------------------------------------
// Second, independent oracle: sorted-slice set algebra (roaring/naive.go:54-176).
// ---------------------------------------------------------------------------
namespace naive {
inline std::vector<uint64_t> intersect(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
  std::vector<uint64_t> o; std::set_intersection(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(o)); return o;
}
inline std::vector<uint64_t> union_(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
  std::vector<uint64_t> o; std::set_union(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(o)); return o;
}
inline std::vector<uint64_t> difference(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
  std::vector<uint64_t> o; std::set_difference(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(o)); return o;
}
inline std::vector<uint64_t> xor_(const std::vector<uint64_t>& a, const std::vector<uint64_t>& b) {
  std::vector<uint64_t> o; std::set_symmetric_difference(a.begin(), a.end(), b.begin(), b.end(), std::back_inserter(o)); return o;
}
}  // namespace naive

}  // namespace oracle
