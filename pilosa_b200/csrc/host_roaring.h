// Host side of the fragment loader: parses the bytes of a Pilosa fragment file
// (roaring snapshot + ops log) and lays the containers out as the device image described
// in layout.h. Pure C++ (no CUDA calls) so it can be exercised without a GPU.
//
// Format (little endian), following roaring/roaring.go:1044-1122 (writer) and
// :1562-1653 (reader):  u32 cookie = 12348 | version<<16 | flags<<24 ; u32 keyN ;
// keyN x {u64 key, u16 type, u16 n-1} ; keyN x u32 offset ; payloads (array: n x u16,
// bitmap: 1024 x u64, run: u16 count + count x {u16 start, u16 last}) ; ops log
// (roaring.go:4459-4559: u8 type, u64 value/len, u32 fnv32a checksum, payload).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "layout.h"

namespace pb200 {

struct HostContainer {
  uint64_t key = 0;
  uint32_t type = kNil;
  uint32_t n = 0;
  uint32_t len = 0;              // array: n ; run: runs ; bitmap: 1024
  const uint8_t* data = nullptr; // payload (run: after the 2 B count); may point into the caller's bytes
  std::vector<uint8_t> owned;    // backing store when re-encoded
};

inline uint32_t payload_bytes(uint32_t type, uint32_t len) {
  return type == kArray ? len * 2 : type == kRun ? len * 4 : type == kBitmap ? kBitmapBytes : 0;
}

template <class T> inline T rd(const uint8_t* p) { T v; std::memcpy(&v, p, sizeof(T)); return v; }

inline uint32_t fnv32a(const uint8_t* p, size_t n, uint32_t h = 2166136261u) {
  for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; }
  return h;
}

// ---- container codec (only used for ops-log replay and pb200_frag_apply) ----------
inline void to_words(const HostContainer& c, uint64_t* w) {
  std::memset(w, 0, kBitmapBytes);
  if (c.type == kBitmap) std::memcpy(w, c.data, kBitmapBytes);
  else if (c.type == kArray) {
    for (uint32_t i = 0; i < c.len; i++) { uint16_t v = rd<uint16_t>(c.data + 2 * i); w[v >> 6] |= 1ull << (v & 63); }
  } else if (c.type == kRun) {
    for (uint32_t i = 0; i < c.len; i++) {
      uint32_t s = rd<uint16_t>(c.data + 4 * i), l = rd<uint16_t>(c.data + 4 * i + 2);
      for (uint32_t ws = s >> 6, we = l >> 6, k = ws; k <= we && l >= s; k++) {
        uint64_t m = ~0ull;
        if (k == ws) m &= ~0ull << (s & 63);
        if (k == we) m &= ~0ull >> (63 - (l & 63));
        w[k] |= m;
      }
    }
  }
}

// Canonical representation of a 65536-bit word array, by the rule of
// Container.optimize (roaring.go:2289-2338): run if runs <= 2048 && runs <= n/2; else
// array if n < 4096; else bitmap; empty -> nil.
inline void from_words(const uint64_t* w, HostContainer& c) {
  uint32_t n = 0, runs = 0;
  for (int i = 0; i < kBitmapWords; i++) {
    uint64_t v = w[i], nx = i + 1 < kBitmapWords ? w[i + 1] : 0;
    n += __builtin_popcountll(v);
    runs += __builtin_popcountll((v << 1) & ~v) + uint32_t((v >> 63) & ~nx);  // run ends (roaring.go:2249)
  }
  c.n = n;
  c.owned.clear();
  if (n == 0) { c.type = kNil; c.len = 0; c.data = nullptr; return; }
  if (runs <= 2048 && runs <= n / 2) {
    c.type = kRun; c.len = runs; c.owned.resize(size_t(runs) * 4);
    uint32_t k = 0; int start = -1;
    for (int i = 0; i < kBitmapWords; i++) {
      uint64_t v = w[i];
      if (start < 0 && v == 0) continue;
      if (start >= 0 && v == ~0ull) continue;
      for (int b = 0; b < 64; b++) {
        bool set = (v >> b) & 1;
        if (set && start < 0) start = i * 64 + b;
        else if (!set && start >= 0) {
          uint16_t s = uint16_t(start), l = uint16_t(i * 64 + b - 1);
          std::memcpy(&c.owned[k * 4], &s, 2); std::memcpy(&c.owned[k * 4 + 2], &l, 2); k++; start = -1;
        }
      }
    }
    if (start >= 0) { uint16_t s = uint16_t(start), l = 0xffff; std::memcpy(&c.owned[k * 4], &s, 2); std::memcpy(&c.owned[k * 4 + 2], &l, 2); k++; }
  } else if (n < 4096) {
    c.type = kArray; c.len = n; c.owned.resize(size_t(n) * 2);
    uint32_t k = 0;
    for (int i = 0; i < kBitmapWords; i++) {
      uint64_t v = w[i];
      while (v) { uint16_t x = uint16_t(i * 64 + __builtin_ctzll(v)); std::memcpy(&c.owned[2 * k++], &x, 2); v &= v - 1; }
    }
  } else {
    c.type = kBitmap; c.len = kBitmapWords; c.owned.resize(kBitmapBytes);
    std::memcpy(c.owned.data(), w, kBitmapBytes);
  }
  c.data = c.owned.data();
}

inline void apply_bit(std::map<uint64_t, HostContainer>& m, uint64_t pos, bool set, std::vector<uint64_t>& scratch) {
  uint64_t key = pos >> 16;
  auto it = m.find(key);
  if (it == m.end()) {
    if (!set) return;
    it = m.emplace(key, HostContainer{}).first;
    it->second.key = key;
  }
  scratch.resize(kBitmapWords);
  to_words(it->second, scratch.data());
  uint32_t v = pos & 0xffff;
  if (set) scratch[v >> 6] |= 1ull << (v & 63); else scratch[v >> 6] &= ~(1ull << (v & 63));
  from_words(scratch.data(), it->second);
  if (it->second.n == 0) m.erase(it);
}

// Batch form: group positions by container so each touched container is decoded once.
inline void apply_bits(std::map<uint64_t, HostContainer>& m, const uint64_t* pos, size_t n, bool set) {
  std::vector<uint64_t> sorted(pos, pos + n);
  std::sort(sorted.begin(), sorted.end());
  std::vector<uint64_t> w(kBitmapWords);
  size_t i = 0;
  while (i < sorted.size()) {
    uint64_t key = sorted[i] >> 16;
    auto it = m.find(key);
    if (it == m.end() && !set) { while (i < sorted.size() && (sorted[i] >> 16) == key) i++; continue; }
    if (it == m.end()) { it = m.emplace(key, HostContainer{}).first; it->second.key = key; }
    to_words(it->second, w.data());
    for (; i < sorted.size() && (sorted[i] >> 16) == key; i++) {
      uint32_t v = sorted[i] & 0xffff;
      if (set) w[v >> 6] |= 1ull << (v & 63); else w[v >> 6] &= ~(1ull << (v & 63));
    }
    from_words(w.data(), it->second);
    if (it->second.n == 0) m.erase(it);
  }
}

// ---- parser ---------------------------------------------------------------------------
// Parses the snapshot section into `out` (ascending key order, zero-copy payload pointers
// into `data`) and reports where the ops log starts. Returns false with `err` set on
// malformed input (never reads out of bounds).
inline bool parse_snapshot(const uint8_t* data, size_t len, std::vector<HostContainer>& out, size_t* ops_off,
                           uint8_t* flags, std::string& err) {
  if (len < 8) { err = "data too small"; return false; }
  uint32_t magic = rd<uint16_t>(data), version = data[2];
  if (flags) *flags = data[3];
  if (magic != 12348) { err = "invalid roaring file, magic number " + std::to_string(magic) + " is incorrect"; return false; }
  if (version != 0) { err = "wrong roaring version, file is v" + std::to_string(version) + ", server requires v0"; return false; }
  uint64_t key_n = rd<uint32_t>(data + 4);
  if (len < 8 + key_n * 16) { err = "malformed bitmap, key-cardinality/offset sections truncated for " + std::to_string(key_n) + " containers"; return false; }
  out.clear();
  out.resize(key_n);
  const uint8_t* hdr = data + 8;
  const uint8_t* offs = data + 8 + key_n * 12;
  size_t end = 8 + key_n * 16;
  for (uint64_t i = 0; i < key_n; i++) {
    HostContainer& c = out[i];
    c.key = rd<uint64_t>(hdr + i * 12);
    c.type = rd<uint16_t>(hdr + i * 12 + 8);
    c.n = uint32_t(rd<uint16_t>(hdr + i * 12 + 10)) + 1;
    uint64_t off = rd<uint32_t>(offs + i * 4);
    if (off >= len) { err = "offset out of bounds: off=" + std::to_string(off) + ", len=" + std::to_string(len); return false; }
    if (c.type == kArray) {
      c.len = c.n;
      if (off + uint64_t(c.len) * 2 > len) { err = "array payload truncated"; return false; }
      c.data = data + off;
      end = off + size_t(c.len) * 2;
    } else if (c.type == kBitmap) {
      c.len = kBitmapWords;
      if (off + kBitmapBytes > len) { err = "bitmap payload truncated"; return false; }
      c.data = data + off;
      end = off + kBitmapBytes;
    } else if (c.type == kRun) {
      if (off + 2 > len) { err = "run header truncated"; return false; }
      c.len = rd<uint16_t>(data + off);
      if (off + 2 + uint64_t(c.len) * 4 > len) { err = "run payload truncated"; return false; }
      c.data = data + off + 2;
      end = off + 2 + size_t(c.len) * 4;
    } else {
      err = "unknown container type " + std::to_string(c.type); return false;
    }
    if (i > 0 && out[i - 1].key >= c.key) { err = "container keys not strictly ascending"; return false; }
  }
  *ops_off = end;
  return true;
}

// The official (32-bit, 16-bit container keys) roaring serialization that Bitmap.UnmarshalBinary also
// accepts (readOfficialHeader / readOffsets / readWithRuns, roaring.go:5075-5245): cookie 12346 = arrays
// and bitmaps with an offset section; cookie 12347 | (n-1) << 16 = is-run bitmap, payloads back to back,
// runs stored as start:length. Error texts are the reference's; out-of-range reads become "... truncated".
inline bool parse_official(const uint8_t* data, size_t len, std::vector<HostContainer>& out, std::string& err) {
  out.clear();
  if (len < 8) { err = "reading roaring header: buffer too small, expecting at least 8 bytes, was " + std::to_string(len); return false; }
  const uint32_t cookie = rd<uint32_t>(data);
  size_t pos = 4;
  uint64_t size = 0;
  const uint8_t* is_run = nullptr;
  if (cookie == 12346) {
    size = rd<uint32_t>(data + pos);
    pos += 4;
  } else if ((cookie & 0xffff) == 12347) {
    size = uint64_t(uint16_t(cookie >> 16)) + 1;
    const size_t nb = size_t(size + 7) / 8;
    if (pos + nb > len) { err = "reading roaring header: malformed bitmap, is-run bitmap overruns buffer at " + std::to_string(pos + nb); return false; }
    is_run = data + pos;
    pos += nb;
  } else {
    err = "reading roaring header: did not find expected serialCookie in header"; return false;
  }
  const size_t header = pos;
  if (size > (1u << 16)) { err = "reading roaring header: it is logically impossible to have more than (1<<16) containers"; return false; }
  if (pos + 4 * size >= len) { err = "reading roaring header: malformed bitmap, key-cardinality slice overruns buffer at " + std::to_string(pos + 4 * size); return false; }
  pos += size_t(4 * size);
  out.resize(size_t(size));
  for (uint64_t i = 0; i < size; i++) {
    HostContainer& c = out[i];
    c.key = rd<uint16_t>(data + header + i * 4);
    c.n = uint32_t(rd<uint16_t>(data + header + i * 4 + 2)) + 1;
    c.type = c.n < 4096 ? kArray : kBitmap;
    if (is_run && (is_run[i / 8] & (1u << (i % 8)))) c.type = kRun;
    if (i > 0 && out[i - 1].key >= c.key) { err = "container keys not strictly ascending"; return false; }
  }
  for (uint64_t i = 0; i < size; i++) {
    HostContainer& c = out[i];
    size_t off = pos;
    if (!is_run) {
      if (pos + i * 4 + 4 > len) { err = "reading offsets from official roaring format: offset incomplete: len=" + std::to_string(len - pos - i * 4); return false; }
      off = rd<uint32_t>(data + pos + i * 4);
      if (off >= len) { err = "reading offsets from official roaring format: offset out of bounds: off=" + std::to_string(off) + ", len=" + std::to_string(len); return false; }
    }
    if (c.type == kRun) {
      if (off + 2 > len) { err = "run header truncated"; return false; }
      c.len = rd<uint16_t>(data + off);
      if (off + 2 + size_t(c.len) * 4 > len) { err = "run payload truncated"; return false; }
      c.owned.resize(size_t(c.len) * 4);
      for (uint32_t k = 0; k < c.len; k++) {  // start:length -> start:last
        const uint16_t st = rd<uint16_t>(data + off + 2 + k * 4), last = uint16_t(st + rd<uint16_t>(data + off + 4 + k * 4));
        std::memcpy(&c.owned[k * 4], &st, 2); std::memcpy(&c.owned[k * 4 + 2], &last, 2);
      }
      c.data = c.owned.data();
      pos = off + 2 + size_t(c.len) * 4;
    } else if (c.type == kArray) {
      c.len = c.n;
      if (off + size_t(c.len) * 2 > len) { err = "array payload truncated"; return false; }
      c.data = data + off;
      if (is_run) pos = off + size_t(c.len) * 2;
    } else {
      c.len = kBitmapWords;
      if (off + kBitmapBytes > len) { err = "bitmap payload truncated"; return false; }
      c.data = data + off;
      if (is_run) pos = off + kBitmapBytes;
    }
  }
  return true;
}

bool parse_fragment(const uint8_t* data, size_t len, std::vector<HostContainer>& out, std::string& err);

// Bitmap.UnmarshalBinary (roaring.go:5133-5179): Pilosa format (with its ops log) by the magic number, else official.
inline bool parse_any(const uint8_t* data, size_t len, std::vector<HostContainer>& out, std::string& err) {
  if (len >= 2 && rd<uint16_t>(data) == 12348) return parse_fragment(data, len, out, err);
  return parse_official(data, len, out, err);
}

// ImportRoaringBits (roaring.go:1463-1558) on the host copy: OR (or AND-NOT) every container of `other`
// into the map; returns the number of bits that changed and, in `row_set` when given, the signed change per
// row (its rowSet result: key / 16 containers per row).
inline int64_t merge_containers(std::map<uint64_t, HostContainer>& m, const std::vector<HostContainer>& other, bool clear,
                                std::map<uint64_t, int64_t>* row_set = nullptr) {
  std::vector<uint64_t> w(kBitmapWords), w2(kBitmapWords);
  int64_t changed = 0;
  for (auto& oc : other) {
    if (oc.n == 0) continue;
    auto it = m.find(oc.key);
    if (it == m.end()) {
      if (clear) continue;
      it = m.emplace(oc.key, HostContainer{}).first; it->second.key = oc.key;
    }
    const int64_t before = it->second.n;
    to_words(it->second, w.data());
    to_words(oc, w2.data());
    for (int k = 0; k < kBitmapWords; k++) w[k] = clear ? (w[k] & ~w2[k]) : (w[k] | w2[k]);
    from_words(w.data(), it->second);
    changed += clear ? before - int64_t(it->second.n) : int64_t(it->second.n) - before;
    if (row_set && int64_t(it->second.n) != before) (*row_set)[oc.key >> 4] += int64_t(it->second.n) - before;
    if (it->second.n == 0) m.erase(it);
  }
  return changed;
}


// Replays the ops log (roaring.go:1627-1650) on top of the snapshot containers.
inline bool replay_ops(const uint8_t* buf, size_t rem, std::map<uint64_t, HostContainer>& m, std::string& err) {
  std::vector<uint64_t> scratch;
  while (rem > 0) {
    if (rem < 13) { err = "op data out of bounds: len=" + std::to_string(rem); return false; }
    uint8_t typ = buf[0];
    uint64_t value = rd<uint64_t>(buf + 1);
    uint32_t chk = rd<uint32_t>(buf + 9);
    uint32_t h = fnv32a(buf, 9);
    size_t size = 13;
    if (typ <= 1) {
    } else if (typ <= 3) {
      if (value > (1ull << 59)) { err = "maximum operation size exceeded"; return false; }
      if (rem < 13 + value * 8) { err = "op data truncated"; return false; }
      h = fnv32a(buf + 13, value * 8, h);
      size = 13 + value * 8;
    } else if (typ <= 5) {
      if (value > rem || rem < 17 + value) { err = "op data truncated"; return false; }
      h = fnv32a(buf + 13, 4 + value, h);
      size = 17 + value;
    } else { err = "unknown op type: " + std::to_string(typ); return false; }
    if (chk != h) { err = "checksum mismatch"; return false; }
    switch (typ) {
      case 0: apply_bit(m, value, true, scratch); break;
      case 1: apply_bit(m, value, false, scratch); break;
      case 2: case 3: {
        std::vector<uint64_t> vals(value);
        for (uint64_t k = 0; k < value; k++) vals[k] = rd<uint64_t>(buf + 13 + k * 8);
        apply_bits(m, vals.data(), vals.size(), typ == 2);
        break;
      }
      case 4: case 5: {  // embedded roaring bitmap: set / clear all of its bits (roaring.go:1463-1558)
        std::vector<HostContainer> other;
        if (!parse_any(buf + 17, value, other, err)) return false;
        merge_containers(m, other, typ == 5);
        break;
      }
    }
    buf += size; rem -= size;
  }
  return true;
}

// Full parse: snapshot + ops log. Containers whose payload was rewritten own their bytes.
inline bool parse_fragment(const uint8_t* data, size_t len, std::vector<HostContainer>& out, std::string& err) {
  size_t ops_off = 0;
  if (!parse_snapshot(data, len, out, &ops_off, nullptr, err)) return false;
  if (ops_off >= len) return true;
  std::map<uint64_t, HostContainer> m;
  for (auto& c : out) m.emplace(c.key, std::move(c));
  if (!replay_ops(data + ops_off, len - ops_off, m, err)) return false;
  out.clear();
  for (auto& kv : m) {
    out.push_back(std::move(kv.second));
    if (!out.back().owned.empty()) out.back().data = out.back().owned.data();
  }
  return true;
}

// Writes containers (ascending keys, already canonical) in the Pilosa roaring format
// (roaring.go:1052-1122 writeToUnoptimized).
inline void serialize_containers(const std::vector<HostContainer>& cs, std::vector<uint8_t>& out) {
  auto put = [&](uint64_t v, int nbytes) { for (int i = 0; i < nbytes; i++) out.push_back(uint8_t(v >> (8 * i))); };
  out.clear();
  uint32_t count = 0;
  for (auto& c : cs) if (c.n) count++;
  put(12348, 4);
  put(count, 4);
  for (auto& c : cs) { if (!c.n) continue; put(c.key, 8); put(c.type, 2); put(c.n - 1, 2); }
  uint32_t off = 8 + count * 16;
  for (auto& c : cs) { if (!c.n) continue; put(off, 4); off += payload_bytes(c.type, c.len) + (c.type == kRun ? 2 : 0); }
  for (auto& c : cs) {
    if (!c.n) continue;
    if (c.type == kRun) put(c.len, 2);
    out.insert(out.end(), c.data, c.data + payload_bytes(c.type, c.len));
  }
}

// ---- device image --------------------------------------------------------------------
struct FragImage {
  std::vector<uint8_t> blob;  // [row_ids (sparse only) | descs | payloads]; desc.addr = offset in blob
  size_t row_ids_off = 0, descs_off = 0, payload_off = 0;
  uint32_t nrows = 0;         // rows in the table (dense: max_row_id + 1)
  bool dense = true;
  uint64_t n_containers = 0, payload_bytes = 0, n_real_rows = 0;
  bool all_bitmap = true;     // every container is a bitmap (or a full run)
  std::vector<uint64_t> row_ids;  // host copy (ascending) for TopN bookkeeping
  std::vector<uint64_t> row_counts;  // cardinality per real row (same order as row_ids)
};

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// densify_min_bytes > 0: array / run containers whose payload is at least that large are stored as 8 KiB
// bitmaps (HBM capacity traded for the streaming path of the evaluator); 0 keeps every container as loaded.
inline bool build_image(const std::vector<HostContainer>& cs, FragImage& img, std::string& err, uint32_t densify_min_bytes = 0) {
  auto densify = [&](const HostContainer& c) {
    return densify_min_bytes && c.n && c.n != kFullN && (c.type == kArray || c.type == kRun) && payload_bytes(c.type, c.len) >= densify_min_bytes;
  };
  // An array above ArrayMaxSize or a run list above runMaxSize is legal input (the reference's readers accept it,
  // roaring.go:1562-1653, :5081-5245) but does not fit the descriptor's 13-bit length: such a container is stored as
  // the bitmap Container.optimize would turn it into (roaring.go:2289-2338).
  auto oversize = [](const HostContainer& c) { return (c.type == kArray && c.len > 4096) || (c.type == kRun && c.len > 2048); };
  // rows present
  img.row_ids.clear(); img.row_counts.clear();
  for (auto& c : cs) {
    uint64_t row = c.key >> 4;
    if (img.row_ids.empty() || img.row_ids.back() != row) { img.row_ids.push_back(row); img.row_counts.push_back(0); }
    img.row_counts.back() += c.n;
  }
  img.n_real_rows = img.row_ids.size();
  uint64_t max_row = img.row_ids.empty() ? 0 : img.row_ids.back();
  img.dense = img.row_ids.empty() || max_row < 2 * img.row_ids.size() + 64;
  if (!img.dense && img.row_ids.size() > 0xffffffffull) { err = "too many rows"; return false; }
  img.nrows = img.row_ids.empty() ? 0 : uint32_t(img.dense ? max_row + 1 : img.row_ids.size());
  size_t off = 0;
  img.row_ids_off = off;
  if (!img.dense) off += img.row_ids.size() * 8;
  off = align_up(off, 128);
  img.descs_off = off;
  off += size_t(img.nrows) * kRowContainers * sizeof(ContainerDesc);
  off = align_up(off, 128);
  img.payload_off = off;
  size_t psize = 0;
  for (auto& c : cs) {
    const bool dz = densify(c) || oversize(c);
    psize = align_up(psize, c.type == kBitmap || dz ? 128 : 16);
    psize += dz ? kBitmapBytes : payload_bytes(c.type, c.len);
  }
  img.blob.assign(off + align_up(psize, 128) + 128, 0);  // +128: kernels may over-read a 16 B tail
  if (!img.dense) std::memcpy(img.blob.data() + img.row_ids_off, img.row_ids.data(), img.row_ids.size() * 8);
  auto* descs = reinterpret_cast<ContainerDesc*>(img.blob.data() + img.descs_off);
  size_t p = img.payload_off;
  size_t ri = 0;
  img.n_containers = 0; img.payload_bytes = 0; img.all_bitmap = true;
  for (auto& c : cs) {
    if (c.n == 0 || c.type == kNil) continue;
    uint64_t row = c.key >> 4;
    while (img.row_ids[ri] != row) ri++;
    size_t slot = (img.dense ? size_t(row) : ri) * kRowContainers + (c.key & 15);
    const bool ov = oversize(c);
    const bool dz = !ov && densify(c);
    p = align_up(p, c.type == kBitmap || dz || ov ? 128 : 16);
    uint32_t pb = payload_bytes(c.type, c.len);
    descs[slot].addr = p;
    descs[slot].n = c.n;
    if (ov) {
      to_words(c, reinterpret_cast<uint64_t*>(img.blob.data() + p));
      descs[slot].meta = make_meta(kBitmap, kBitmapWords);
      pb = kBitmapBytes;
    } else if (dz) {
      to_words(c, reinterpret_cast<uint64_t*>(img.blob.data() + p));  // 128 B aligned inside the blob
      descs[slot].meta = make_meta_densified(pb);
      pb = kBitmapBytes;
    } else {
      std::memcpy(img.blob.data() + p, c.data, pb);
      descs[slot].meta = make_meta(c.type, c.len);
    }
    p += pb;
    img.n_containers++;
    img.payload_bytes += pb;
    if (!(c.type == kBitmap || dz || ov || c.n == kFullN)) img.all_bitmap = false;
  }
  return true;
}

}  // namespace pb200
