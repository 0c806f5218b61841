// Synthetic index generator: seeded splitmix64 data emitted as
// the exact bytes of Pilosa fragment files (roaring.go:1044-1122), every container in the
// canonical type Container.optimize (roaring.go:2289-2338) would give it — i.e. what a
// snapshotted Pilosa fragment holds. Host-only helper library (libpb200_synth.so) used by
// the tests and bench.py to feed BOTH the engine (pb200_frag_load) and the CPU oracle with
// identical bytes. Not part of the query path.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

constexpr int kWords = 1024;

struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next() {  // splitmix64
    uint64_t z = (s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
  }
  double uniform() { return double(next() >> 11) * (1.0 / 9007199254740992.0); }
  uint64_t below(uint64_t n) { return n ? next() % n : 0; }
};

inline uint64_t mix_seed(uint64_t seed, uint64_t a, uint64_t b, uint64_t c) {
  Rng r(seed ^ (a * 0x9e3779b97f4a7c15ull) ^ (b * 0xc2b2ae3d27d4eb4full) ^ (c * 0x165667b19e3779f9ull));
  r.next();
  return r.next();
}

// 64 independent Bernoulli(p16 / 65536) bits: fold the binary expansion of p from the
// least significant bit with AND (0) / OR (1) of fresh random words.
inline uint64_t bernoulli_word(Rng& r, uint32_t p16) {
  if (p16 == 0) return 0;
  if (p16 >= 65536) return ~0ull;
  uint64_t acc = 0;
  int tz = __builtin_ctz(p16);
  for (int b = tz; b < 16; b++) {
    uint64_t w = r.next();
    acc = ((p16 >> b) & 1) ? (acc | w) : (acc & w);
  }
  return acc;
}

struct Out {
  std::vector<uint8_t> bytes;
};

struct Cont {
  uint64_t key;
  uint16_t type;  // 1 array 2 bitmap 3 run
  uint32_t n;
  std::vector<uint8_t> payload;  // run: includes the u16 count header
};

// canonical container from 1024 words (optimize rule)
bool encode_words(const uint64_t* w, uint64_t key, Cont& c) {
  uint32_t n = 0, runs = 0;
  for (int i = 0; i < kWords; i++) {
    uint64_t v = w[i], nx = i + 1 < kWords ? w[i + 1] : 0;
    n += __builtin_popcountll(v);
    runs += __builtin_popcountll((v << 1) & ~v) + uint32_t((v >> 63) & ~nx);
  }
  if (n == 0) return false;
  c.key = key; c.n = n; c.payload.clear();
  if (runs <= 2048 && runs <= n / 2) {
    c.type = 3;
    c.payload.resize(2 + size_t(runs) * 4);
    uint16_t rc = uint16_t(runs);
    std::memcpy(c.payload.data(), &rc, 2);
    size_t k = 0; int start = -1;
    auto emit = [&](int s, int l) { uint16_t a = uint16_t(s), b = uint16_t(l); std::memcpy(&c.payload[2 + k * 4], &a, 2); std::memcpy(&c.payload[4 + k * 4], &b, 2); k++; };
    for (int i = 0; i < kWords; i++) {
      uint64_t v = w[i];
      if (start < 0 && v == 0) continue;
      if (start >= 0 && v == ~0ull) continue;
      for (int b = 0; b < 64; b++) {
        bool set = (v >> b) & 1;
        if (set && start < 0) start = i * 64 + b;
        else if (!set && start >= 0) { emit(start, i * 64 + b - 1); start = -1; }
      }
    }
    if (start >= 0) emit(start, 65535);
  } else if (n < 4096) {
    c.type = 1;
    c.payload.resize(size_t(n) * 2);
    size_t k = 0;
    for (int i = 0; i < kWords; i++) {
      uint64_t v = w[i];
      while (v) { uint16_t x = uint16_t(i * 64 + __builtin_ctzll(v)); std::memcpy(&c.payload[2 * k++], &x, 2); v &= v - 1; }
    }
  } else {
    c.type = 2;
    c.payload.resize(8192);
    std::memcpy(c.payload.data(), w, 8192);
  }
  return true;
}

void serialize(const std::vector<Cont>& cs, std::vector<uint8_t>& out) {
  size_t total = 8 + cs.size() * 16;
  for (auto& c : cs) total += c.payload.size();
  out.resize(total);
  uint8_t* p = out.data();
  uint32_t cookie = 12348, cnt = uint32_t(cs.size());
  std::memcpy(p, &cookie, 4); std::memcpy(p + 4, &cnt, 4);
  uint8_t* h = p + 8;
  for (auto& c : cs) { uint16_t n1 = uint16_t(c.n - 1); std::memcpy(h, &c.key, 8); std::memcpy(h + 8, &c.type, 2); std::memcpy(h + 10, &n1, 2); h += 12; }
  uint64_t off = 8 + cs.size() * 16;
  for (auto& c : cs) { uint32_t o = uint32_t(off); std::memcpy(h, &o, 4); h += 4; off += c.payload.size(); }
  for (auto& c : cs) { std::memcpy(h, c.payload.data(), c.payload.size()); h += c.payload.size(); }
}

// Streaming writer for big all-bitmap fragments: avoids holding Cont copies.
struct StreamFrag {
  std::vector<Cont> small;  // only metadata is kept when payload_direct
};

}  // namespace

extern "C" {

void* pbs_alloc() { return new Out(); }
void pbs_free(void* h) { delete static_cast<Out*>(h); }
const uint8_t* pbs_data(void* h) { return static_cast<Out*>(h)->bytes.data(); }
size_t pbs_size(void* h) { return static_cast<Out*>(h)->bytes.size(); }

// Rows [row_lo, row_lo + n_rows) of one shard, every column iid Bernoulli(p)
// (C1: p = 0.10 -> bitmap containers; C5: p = 1e-4 -> arrays / short runs).
// The stream of row r, container c depends only on (seed, shard, r, c).
void* pbs_gen_bernoulli_rows(uint64_t seed, uint64_t shard, const uint64_t* row_ids, uint64_t n_rows, double p);

void* pbs_gen_bernoulli(uint64_t seed, uint64_t shard, uint64_t row_lo, uint64_t n_rows, double p) {
  std::vector<uint64_t> ids(n_rows);
  for (uint64_t i = 0; i < n_rows; i++) ids[i] = row_lo + i;
  return pbs_gen_bernoulli_rows(seed, shard, ids.data(), n_rows, p);
}

// Same streams, explicit (ascending) row ids: lets the CPU baseline hold only the rows a
// bounded sample of queries touches while the GPU holds the full index.
void* pbs_gen_bernoulli_rows(uint64_t seed, uint64_t shard, const uint64_t* row_ids, uint64_t n_rows, double p) {
  auto* o = new Out();
  std::vector<Cont> cs;
  cs.reserve(n_rows * 16);
  std::vector<uint64_t> w(kWords);
  const bool sparse = p < 1.0 / 256;
  const uint32_t p16 = uint32_t(std::lround(p * 65536.0));
  for (uint64_t ri = 0; ri < n_rows; ri++) {
    const uint64_t r = row_ids[ri];
    for (uint64_t c = 0; c < 16; c++) {
      Rng rng(mix_seed(seed, shard, r, c));
      if (sparse) {
        std::fill(w.begin(), w.end(), 0);
        if (p > 0) {
          const double lq = std::log1p(-p);
          double pos = -1;
          for (;;) {
            double u = rng.uniform();
            if (u <= 0) u = 1e-300;
            pos += 1.0 + std::floor(std::log(u) / lq);
            if (pos >= 65536.0) break;
            uint32_t v = uint32_t(pos);
            w[v >> 6] |= 1ull << (v & 63);
          }
        }
      } else {
        for (int i = 0; i < kWords; i++) w[i] = bernoulli_word(rng, p16);
      }
      Cont ct;
      if (encode_words(w.data(), r * 16 + c, ct)) cs.push_back(std::move(ct));
    }
  }
  serialize(cs, o->bytes);
  return o;
}

// C2: per (row, container) one of three classes by RNG: 1/3 array (n ~ U[1,4095], uniform
// positions), 1/3 bitmap (p ~ U[0.07, 0.9]), 1/3 run (k ~ U[1,2047] runs, geometric
// lengths); then canonicalised like a snapshot would.
void* pbs_gen_mixed(uint64_t seed, uint64_t shard, uint64_t row_lo, uint64_t n_rows) {
  auto* o = new Out();
  std::vector<Cont> cs;
  std::vector<uint64_t> w(kWords);
  for (uint64_t r = row_lo; r < row_lo + n_rows; r++) {
    for (uint64_t c = 0; c < 16; c++) {
      Rng rng(mix_seed(seed, shard, r, c));
      std::fill(w.begin(), w.end(), 0);
      uint64_t cls = rng.below(3);
      if (cls == 0) {
        uint32_t n = 1 + uint32_t(rng.below(4095));
        for (uint32_t i = 0; i < n; i++) { uint32_t v = uint32_t(rng.below(65536)); w[v >> 6] |= 1ull << (v & 63); }
      } else if (cls == 1) {
        double p = 0.07 + 0.83 * rng.uniform();
        uint32_t p16 = uint32_t(std::lround(p * 65536.0));
        for (int i = 0; i < kWords; i++) w[i] = bernoulli_word(rng, p16);
      } else {
        uint32_t k = 1 + uint32_t(rng.below(2047));
        // k runs with geometric lengths and geometric gaps filling the container on average
        double mean_len = 65536.0 / (2.0 * k) + 1.0;
        double pos = 0;
        for (uint32_t i = 0; i < k && pos < 65536.0; i++) {
          double gap = std::floor(-std::log(std::max(rng.uniform(), 1e-300)) * mean_len);
          double len = 2.0 + std::floor(-std::log(std::max(rng.uniform(), 1e-300)) * mean_len);
          uint32_t s = uint32_t(std::min(pos + 1 + gap, 65535.0));
          uint32_t l = uint32_t(std::min(double(s) + len - 1, 65535.0));
          for (uint32_t wd = s >> 6; wd <= (l >> 6); wd++) {
            uint64_t m = ~0ull;
            if (wd == (s >> 6)) m &= ~0ull << (s & 63);
            if (wd == (l >> 6)) m &= ~0ull >> (63 - (l & 63));
            w[wd] |= m;
          }
          pos = double(l) + 1;
        }
      }
      Cont ct;
      if (encode_words(w.data(), r * 16 + c, ct)) cs.push_back(std::move(ct));
    }
  }
  serialize(cs, o->bytes);
  return o;
}

// C4: BSI fragment (view bsig_<field>): rows 0 = exists, 1 = sign, 2+i = magnitude bit i
// (fragment.go:90-93). Each column is set with probability `density`; values are uniform in
// [vmin, vmax] stored sign-magnitude (fragment.go:936-973). bit_depth rows of magnitude.
void* pbs_gen_bsi(uint64_t seed, uint64_t shard, uint32_t bit_depth, int64_t vmin, int64_t vmax, double density) {
  auto* o = new Out();
  const size_t row_words = 16 * kWords;
  std::vector<uint64_t> planes((size_t(bit_depth) + 2) * row_words, 0);
  Rng rng(mix_seed(seed, shard, 0xb51, 0));
  const uint64_t span = uint64_t(vmax - vmin) + 1;  // 0 when the range is the full 2^64
  for (uint64_t col = 0; col < (1ull << 20); col++) {
    if (density < 1.0 && rng.uniform() >= density) continue;
    int64_t v = vmin + int64_t(span ? rng.next() % span : rng.next());
    uint64_t uv = v < 0 ? 0ull - uint64_t(v) : uint64_t(v);
    const size_t wi = col >> 6; const uint64_t bit = 1ull << (col & 63);
    planes[wi] |= bit;
    if (v < 0) planes[row_words + wi] |= bit;
    for (uint32_t i = 0; i < bit_depth; i++)
      if ((uv >> i) & 1) planes[(2 + size_t(i)) * row_words + wi] |= bit;
  }
  std::vector<Cont> cs;
  for (uint64_t r = 0; r < uint64_t(bit_depth) + 2; r++)
    for (uint64_t c = 0; c < 16; c++) {
      Cont ct;
      if (encode_words(planes.data() + r * row_words + c * kWords, r * 16 + c, ct)) cs.push_back(std::move(ct));
    }
  serialize(cs, o->bytes);
  return o;
}

// C3: a ranked set field with n_rows rows; each column draws k rows from
// Zipf(s, v, imax = n_rows - 1): P(r) ~ (v + r)^-s, like the reference's
// getZipfRowsSliceRoaring test generator (fragment_internal_test.go:2377-2399). Sampling
// uses an alias table (exact distribution, not Go's rejection-inversion stream).
void* pbs_gen_zipf(uint64_t seed, uint64_t shard, uint32_t n_rows, uint32_t k, double s, double v) {
  auto* o = new Out();
  // alias table
  std::vector<double> prob(n_rows);
  double tot = 0;
  for (uint32_t r = 0; r < n_rows; r++) { prob[r] = std::pow(v + double(r), -s); tot += prob[r]; }
  std::vector<double> q(n_rows);
  std::vector<uint32_t> alias(n_rows, 0), small, large;
  for (uint32_t r = 0; r < n_rows; r++) { q[r] = prob[r] / tot * n_rows; (q[r] < 1.0 ? small : large).push_back(r); }
  while (!small.empty() && !large.empty()) {
    uint32_t a = small.back(), b = large.back(); small.pop_back();
    alias[a] = b; q[b] = q[b] + q[a] - 1.0;
    if (q[b] < 1.0) { large.pop_back(); small.push_back(b); }
  }
  for (uint32_t a : small) q[a] = 1.0;
  for (uint32_t b : large) q[b] = 1.0;
  // draw
  Rng rng(mix_seed(seed, shard, 0x21bf, 0));
  std::vector<uint32_t> counts(n_rows + 1, 0);
  std::vector<uint32_t> rows(size_t(k) << 20);
  for (size_t i = 0; i < rows.size(); i++) {
    uint64_t x = rng.next();
    uint32_t r = uint32_t((x >> 32) % n_rows);
    double u = double(x & 0xffffffffu) * (1.0 / 4294967296.0);
    if (u >= q[r]) r = alias[r];
    rows[i] = r;
    counts[r + 1]++;
  }
  for (uint32_t r = 0; r < n_rows; r++) counts[r + 1] += counts[r];
  std::vector<uint32_t> cols(rows.size());
  {
    std::vector<uint32_t> cur(counts.begin(), counts.end() - 1);
    for (size_t i = 0; i < rows.size(); i++) cols[cur[rows[i]]++] = uint32_t(i / k);  // ascending column order per row
  }
  std::vector<Cont> cs;
  std::vector<uint64_t> w(16 * kWords);
  for (uint32_t r = 0; r < n_rows; r++) {
    if (counts[r] == counts[r + 1]) continue;
    std::fill(w.begin(), w.end(), 0);
    for (uint32_t i = counts[r]; i < counts[r + 1]; i++) { uint32_t c = cols[i]; w[c >> 6] |= 1ull << (c & 63); }
    for (uint64_t c = 0; c < 16; c++) {
      Cont ct;
      if (encode_words(w.data() + c * kWords, uint64_t(r) * 16 + c, ct)) cs.push_back(std::move(ct));
    }
  }
  serialize(cs, o->bytes);
  return o;
}

// Arbitrary fragment from explicit storage positions (pos = row * 2^20 + col % 2^20),
// canonicalised: the small literal fixtures of the reference's tests.
void* pbs_from_positions(const uint64_t* pos, size_t n) {
  auto* o = new Out();
  std::vector<uint64_t> p(pos, pos + n);
  std::sort(p.begin(), p.end());
  p.erase(std::unique(p.begin(), p.end()), p.end());
  std::vector<Cont> cs;
  std::vector<uint64_t> w(kWords);
  size_t i = 0;
  while (i < p.size()) {
    uint64_t key = p[i] >> 16;
    std::fill(w.begin(), w.end(), 0);
    for (; i < p.size() && (p[i] >> 16) == key; i++) { uint32_t v = p[i] & 0xffff; w[v >> 6] |= 1ull << (v & 63); }
    Cont ct;
    if (encode_words(w.data(), key, ct)) cs.push_back(std::move(ct));
  }
  serialize(cs, o->bytes);
  return o;
}

}  // extern "C"
