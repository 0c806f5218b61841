// libpilosa_b200: context, fragment loader, program compiler and the C ABI entry points.
// Single translation unit with the kernels (kernels*.cuh) so nvcc -lineinfo maps the ncu
// source page onto this tree.
#include "engine.h"

#include <algorithm>
#include <chrono>
#include <set>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <condition_variable>
#include <memory>
#include <thread>

#include "kernels.cuh"
// A measured alternative to eval_fold_kernel for count-only launches (consumer warps that share no barrier): fewer stalls,
// twice the instructions, 0.34 vs 0.26 ms on the mixed-container config. Built only with -DPB200_FOLDW=1.
#ifndef PB200_FOLDW
#define PB200_FOLDW 0
#endif
#if PB200_FOLDW
#include "kernels_foldw.cuh"
#endif
#include "kernels_bsi.cuh"
#include "kernels_topn.cuh"
#include "kernels_rows.cuh"

using namespace pb200;

// ---------------------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------------------
#define CUDA_TRY(ctx, call)                                                                   \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                       \
      return PB200_ERR_CUDA;                                                                  \
    }                                                                                         \
  } while (0)

#define TRY(expr)             \
  do {                        \
    int rc__ = (expr);        \
    if (rc__ != PB200_OK) return rc__; \
  } while (0)

namespace { void peer_teardown(pb200_ctx* ctx); }  // engine_peer.inl

static int fail(pb200_ctx* ctx, int code, const std::string& msg) {
  ctx->err = msg;
  return code;
}

static int ensure_dev(pb200_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap) return PB200_OK;
  size_t ncap = std::max(bytes, b.cap * 2);
  ncap = (ncap + 255) & ~size_t(255);
  if (b.p) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, cudaFree(b.p));
    b.p = nullptr; b.cap = 0;
  }
  CUDA_TRY(ctx, cudaMalloc(&b.p, ncap));
  b.cap = ncap;
  return PB200_OK;
}

static int ensure_pin(pb200_ctx* ctx, PinBuf& b, size_t bytes) {
  if (bytes <= b.cap) return PB200_OK;
  size_t ncap = std::max(bytes, b.cap * 2);
  if (b.p) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, cudaFreeHost(b.p));
    b.p = nullptr; b.cap = 0;
  }
  CUDA_TRY(ctx, cudaMallocHost(&b.p, ncap));
  b.cap = ncap;
  return PB200_OK;
}

struct DeviceGuard {  // make ctx->device current for the duration of a call
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

#define ENTER(ctx)                                   \
  if (!(ctx)) return PB200_ERR_INVALID;              \
  std::lock_guard<std::mutex> lock__((ctx)->mu);     \
  DeviceGuard guard__((ctx)->device);                \
  (ctx)->err.clear()

namespace {

// A persistent host thread: issues work for one device of a pb200_multi, or runs a slice of a host-side shard loop. Hand-off is a spin on an atomic for a short while after each job
// (a burst of queries finds the worker awake: ~1 us) and a condition variable after that.
class DevWorker {
 public:
  DevWorker() : th_([this] { loop(); }) {}
  ~DevWorker() {
    { std::lock_guard<std::mutex> g(mu_); stop_ = true; }
    cv_.notify_all();
    th_.join();
  }
  void post(std::function<int()> fn) {
    fn_ = std::move(fn);
    state_.store(1, std::memory_order_release);
    { std::lock_guard<std::mutex> g(mu_); }
    cv_.notify_one();
  }
  int wait() {
    for (uint32_t spins = 0; state_.load(std::memory_order_acquire) != 2; spins++)
      if (spins > 2000) std::this_thread::yield();
    state_.store(0, std::memory_order_relaxed);
    return rc_;
  }

 private:
  void loop() {
    for (;;) {
      bool have = false;
      const auto t0 = std::chrono::steady_clock::now();
      for (uint32_t spins = 0;; spins++) {
        if (state_.load(std::memory_order_acquire) == 1) { have = true; break; }
        if ((spins & 255) == 255 && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(200)) break;
      }
      if (!have) {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return stop_ || state_.load(std::memory_order_acquire) == 1; });
        if (stop_) return;
      }
      rc_ = fn_();
      state_.store(2, std::memory_order_release);
    }
  }
  std::function<int()> fn_;
  int rc_ = 0;
  std::atomic<int> state_{0};  // 0 idle, 1 job posted, 2 job done
  std::mutex mu_;
  std::condition_variable cv_;
  bool stop_ = false;
  std::thread th_;
};

}  // namespace

// ---------------------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------------------
extern "C" int pb200_abi_version(void) { return PB200_ABI_VERSION; }

extern "C" int pb200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" const char* pb200_strerror(int status) {
  switch (status) {
    case PB200_OK: return "ok";
    case PB200_ERR_INVALID: return "invalid argument";
    case PB200_ERR_CUDA: return "CUDA error";
    case PB200_ERR_FORMAT: return "malformed roaring data";
    case PB200_ERR_NOMEM: return "HBM budget exhausted";
    case PB200_ERR_NOTFOUND: return "not found";
    case PB200_ERR_RANGE: return "buffer too small or size limit exceeded";
    case PB200_ERR_UNSUPPORTED: return "unsupported";
    default: return "unknown status";
  }
}

extern "C" const char* pb200_last_error(pb200_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int pb200_init(int device, size_t hbm_budget_bytes, pb200_ctx** out) {
  if (!out) return PB200_ERR_INVALID;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0 || device < 0 || device >= ndev) return PB200_ERR_CUDA;  // no CPU fallback
  auto* ctx = new pb200_ctx();
  ctx->device = device;
  ctx->budget = hbm_budget_bytes;
  DeviceGuard g(device);
  cudaDeviceProp prop{};
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete ctx; return PB200_ERR_CUDA; }
  ctx->sm_count = prop.multiProcessorCount;
  // the fold evaluator's window + scratch + staging ring may exceed the 48 KiB a kernel gets without asking
  if (cudaFuncSetAttribute(eval_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFoldSmem) != cudaSuccess) { delete ctx; return PB200_ERR_CUDA; }
#if PB200_FOLDW
  if (cudaFuncSetAttribute(eval_foldw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFoldSmem) != cudaSuccess) { delete ctx; return PB200_ERR_CUDA; }
#endif
  if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess) {
    delete ctx;
    return PB200_ERR_CUDA;
  }
  ctx->stream = ctx->own_stream;
  *out = ctx;
  return PB200_OK;
}

static void arena_detach(pb200_ctx* ctx, Fragment& f) {  // the fragment stops referencing the write arena
  if (!f.in_arena) return;
  f.in_arena = false;
  WriteArena& a = ctx->arena;
  if (a.live_frags && --a.live_frags == 0) {  // nothing points into it any more: reclaim (retired chunks freed, current one reused)
    for (auto& r : a.retired) { cudaFree(r.first); ctx->used -= std::min(ctx->used, r.second); }
    a.retired.clear();
    a.top = 0;
  }
}

static void free_fragment(Fragment& f) {
  if (f.blob) cudaFree(f.blob);
  if (f.d_rank) cudaFree(f.d_rank);
  f.blob = nullptr;
  f.d_rank = nullptr;
  f.d_rank_counts = nullptr;
  f.d_rank_valid = false;
  f.rank_pos_valid = false;
}

extern "C" int pb200_shutdown(pb200_ctx* ctx) {
  if (!ctx) return PB200_ERR_INVALID;
  {
    std::lock_guard<std::mutex> lock(ctx->mu);
    DeviceGuard g(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto& fam : ctx->families) {
      for (auto& kv : fam.frags) free_fragment(kv.second);
      if (fam.d_frags) cudaFree(fam.d_frags);
    }
    if (ctx->d_families) cudaFree(ctx->d_families);
    for (DevBuf* b : {&ctx->d_rows, &ctx->d_rows_off, &ctx->d_rows_meta, &ctx->d_rows_item_off, &ctx->d_ovf, &ctx->d_slow_descs, &ctx->d_rest, &ctx->d_sparse, &ctx->d_gb, &ctx->d_batch, &ctx->d_out, &ctx->d_slow, &ctx->d_temp, &ctx->d_dense, &ctx->d_dense_n, &ctx->d_misc, &ctx->d_topn, &ctx->d_bsi, &ctx->d_ing, &ctx->d_ing_res, &ctx->d_topn_compact})
      if (b->p) cudaFree(b->p);
    for (auto& r : ctx->arena.retired) cudaFree(r.first);
    if (ctx->arena.base) cudaFree(ctx->arena.base);
    if (ctx->arena.d_top) cudaFree(ctx->arena.d_top);
    for (PinBuf* b : {&ctx->h_stage, &ctx->h_out, &ctx->h_algo, &ctx->h_topn, &ctx->h_ing, &ctx->h_ing_res})
      if (b->p) cudaFreeHost(b->p);
    for (auto& sl : ctx->slots) {
      if (sl.d_batch.p) cudaFree(sl.d_batch.p);
      if (sl.d_out.p) cudaFree(sl.d_out.p);
      if (sl.h_stage.p) cudaFreeHost(sl.h_stage.p);
      if (sl.h_out.p) cudaFreeHost(sl.h_out.p);
      if (sl.k0) cudaEventDestroy(sl.k0);
      if (sl.k1) cudaEventDestroy(sl.k1);
      if (sl.done) cudaEventDestroy(sl.done);
    }
    peer_teardown(ctx);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  }
  delete ctx;
  return PB200_OK;
}

extern "C" int pb200_set_stream(pb200_ctx* ctx, void* cuda_stream) {
  ENTER(ctx);
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->stream = cuda_stream ? static_cast<cudaStream_t>(cuda_stream) : ctx->own_stream;
  return PB200_OK;
}

extern "C" int pb200_last_timing(pb200_ctx* ctx, pb200_timing* out) {
  if (!ctx || !out) return PB200_ERR_INVALID;
  std::lock_guard<std::mutex> lock(ctx->mu);
  *out = ctx->timing;
  return PB200_OK;
}

// ---------------------------------------------------------------------------------------
// families and fragments
// ---------------------------------------------------------------------------------------
extern "C" int pb200_family_create(pb200_ctx* ctx, pb200_family* out) {
  ENTER(ctx);
  if (!out) return PB200_ERR_INVALID;
  ctx->families.emplace_back();
  ctx->families_dirty = true;
  ctx->generation++;
  *out = uint32_t(ctx->families.size() - 1);
  return PB200_OK;
}

static int family_grow(pb200_ctx* ctx, Family& fam, uint64_t shard) {
  if (shard < fam.cap) return PB200_OK;
  if (shard >= (1ull << 26)) return fail(ctx, PB200_ERR_RANGE, "shard number too large");
  uint32_t ncap = std::max<uint32_t>(64, fam.cap);
  while (ncap <= shard) ncap *= 2;
  FragDev* nd = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&nd, size_t(ncap) * sizeof(FragDev)));
  fam.h_frags.resize(ncap, FragDev{});
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaMemcpy(nd, fam.h_frags.data(), size_t(ncap) * sizeof(FragDev), cudaMemcpyHostToDevice));
  if (fam.d_frags) CUDA_TRY(ctx, cudaFree(fam.d_frags));
  fam.d_frags = nd;
  fam.cap = ncap;
  ctx->families_dirty = true;
  ctx->generation++;
  return PB200_OK;
}

static int sync_families(pb200_ctx* ctx) {
  if (!ctx->families_dirty) return PB200_OK;
  uint32_t n = uint32_t(ctx->families.size());
  if (n > ctx->d_families_cap) {
    uint32_t ncap = std::max<uint32_t>(16, ctx->d_families_cap);
    while (ncap < n) ncap *= 2;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->d_families) CUDA_TRY(ctx, cudaFree(ctx->d_families));
    CUDA_TRY(ctx, cudaMalloc(&ctx->d_families, size_t(ncap) * sizeof(FamilyDev)));
    ctx->d_families_cap = ncap;
  }
  std::vector<FamilyDev> h(n);
  for (uint32_t i = 0; i < n; i++) { h[i].frags = ctx->families[i].d_frags; h[i].cap = ctx->families[i].cap; h[i].pad = 0; }
  if (n) {
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    CUDA_TRY(ctx, cudaMemcpy(ctx->d_families, h.data(), size_t(n) * sizeof(FamilyDev), cudaMemcpyHostToDevice));
  }
  ctx->families_dirty = false;
  return PB200_OK;
}

static void default_rank(Fragment& f, uint32_t max_entries) {
  // Exact row counts, count desc / id asc, trimmed to the cache size (DefaultCacheSize = 50000, field.go:47);
  // rows with count 0 never enter the cache.
  size_t n = f.row_ids.size();
  std::vector<uint32_t> order(n);
  for (size_t i = 0; i < n; i++) order[i] = uint32_t(i);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return f.row_counts[a] > f.row_counts[b]; });
  if (n > max_entries) order.resize(max_entries);
  f.rank_ids.clear(); f.rank_counts.clear();
  for (uint32_t i : order) { f.rank_ids.push_back(f.row_ids[i]); f.rank_counts.push_back(f.row_counts[i]); }
  f.rank_from_host = false;
}

// rankCache.recalculate (cache.go:245-281) over the engine-held entries: rank by count (descending; ties by
// ascending id, where the reference's sort.Sort leaves them unspecified), cut at the cache size, remember the
// count at the cut as the admission threshold, and drop the cut entries once the map outgrows 1.1 x the size.
// The ranking becomes the fragment's TopN candidate order.
static void rank_recalculate(Fragment& f, uint32_t max_entries) {
  auto& rs = f.rank_state;
  std::vector<std::pair<uint64_t, uint64_t>> r(rs.entries.begin(), rs.entries.end());  // (id, count)
  std::sort(r.begin(), r.end(), [](const std::pair<uint64_t, uint64_t>& a, const std::pair<uint64_t, uint64_t>& b) {
    return a.second != b.second ? a.second > b.second : a.first < b.first;
  });
  size_t keep = r.size();
  if (r.size() > max_entries) { rs.threshold_value = r[max_entries].second; keep = max_entries; }
  else rs.threshold_value = 1;
  if (rs.entries.size() > size_t(int(1.1 * double(max_entries))))  // thresholdFactor cache.go:31
    for (size_t i = keep; i < r.size(); i++) rs.entries.erase(r[i].first);
  f.rank_ids.resize(keep); f.rank_counts.resize(keep);
  for (size_t i = 0; i < keep; i++) { f.rank_ids[i] = r[i].first; f.rank_counts[i] = r[i].second; }
  f.rank_from_host = true;  // the cached counts are what TopN reports, as with a list handed over by the host
  f.d_rank_valid = false;
  f.rank_pos_valid = false;
}

// The cache as it stands before the first import recount: the list the host handed over (pb200_rank_set), else
// every present row with its exact count, ranked once — what openCache + Recalculate leave (fragment.go:404-463).
static void rank_state_build(Fragment& f, uint32_t max_entries) {
  auto& rs = f.rank_state;
  if (rs.live) return;
  rs.entries.clear();
  if (f.rank_from_host) for (size_t k = 0; k < f.rank_ids.size(); k++) rs.entries[f.rank_ids[k]] = f.rank_counts[k];
  else for (size_t j = 0; j < f.row_ids.size(); j++) if (f.row_counts[j]) rs.entries[f.row_ids[j]] = f.row_counts[j];
  rank_recalculate(f, max_entries);
  rs.live = true;
}

static int upload_fragment(pb200_ctx* ctx, Family& fam, uint64_t shard, const std::vector<HostContainer>& cs) {
  FragImage img;
  std::string err;
  if (!build_image(cs, img, err, ctx->densify_min_bytes)) return fail(ctx, PB200_ERR_FORMAT, err);
  if (ctx->budget && ctx->used + img.blob.size() > ctx->budget) return fail(ctx, PB200_ERR_NOMEM, "HBM budget exhausted");
  TRY(family_grow(ctx, fam, shard));
  void* blob = nullptr;
  CUDA_TRY(ctx, cudaMalloc(&blob, img.blob.size()));
  struct BlobGuard {  // the new copy is released on every failure path below
    void* p;
    ~BlobGuard() { if (p) cudaFree(p); }
  } guard{blob};
  // fix up payload addresses now that the base is known
  auto* descs = reinterpret_cast<ContainerDesc*>(img.blob.data() + img.descs_off);
  const uint64_t base = reinterpret_cast<uint64_t>(blob);
  for (size_t i = 0; i < size_t(img.nrows) * kRowContainers; i++)
    if (descs[i].n) descs[i].addr += base;
  CUDA_TRY(ctx, cudaMemcpy(blob, img.blob.data(), img.blob.size(), cudaMemcpyHostToDevice));

  Fragment nf;
  nf.blob = blob;
  nf.blob_bytes = img.blob.size();
  nf.row_ids_off = img.row_ids_off; nf.descs_off = img.descs_off; nf.payload_off = img.payload_off;
  nf.dev.row_ids = reinterpret_cast<const uint64_t*>(base + img.row_ids_off);
  nf.dev.rows = reinterpret_cast<const ContainerDesc*>(base + img.descs_off);
  nf.dev.nrows = img.nrows;
  nf.dev.dense = img.dense ? 1u : 0u;
  nf.n_containers = img.n_containers;
  nf.payload_bytes = img.payload_bytes;
  nf.all_bitmap = img.all_bitmap;
  nf.row_ids = std::move(img.row_ids);
  nf.row_counts = std::move(img.row_counts);
  default_rank(nf, ctx->rank_cache_size);

  // Order: wait for the queries that may still read the old copy, point the device table at the new copy, and only
  // then let go of the old one — a failure before that leaves the old fragment fully in place.
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaMemcpy(fam.d_frags + shard, &nf.dev, sizeof(FragDev), cudaMemcpyHostToDevice));
  if (ctx->publish) TRY(ctx->publish(uint32_t(&fam - ctx->families.data()), shard, nf.dev, nf.all_bitmap, true));
  guard.p = nullptr;
  auto it = fam.frags.find(shard);
  if (it != fam.frags.end()) {
    if (!it->second.all_bitmap) fam.n_not_all_bitmap--;
    fam.payload_total -= it->second.payload_bytes; fam.containers_total -= it->second.n_containers;
    ctx->used -= it->second.blob_bytes;
    arena_detach(ctx, it->second);
    free_fragment(it->second);
    fam.frags.erase(it);
  }
  ctx->used += nf.blob_bytes;
  fam.payload_total += nf.payload_bytes; fam.containers_total += nf.n_containers;
  if (!nf.all_bitmap) fam.n_not_all_bitmap++;
  fam.h_frags[shard] = nf.dev;
  fam.frags.emplace(shard, std::move(nf));
  ctx->generation++;
  return PB200_OK;
}

extern "C" int pb200_host_check_roaring(const void* roaring_bytes, size_t len, uint64_t* n_containers, uint64_t* n_bits, char* err,
                                        size_t err_cap) {
  std::vector<HostContainer> cs;
  std::string msg;
  if (!roaring_bytes && len) return PB200_ERR_INVALID;
  const bool ok = parse_any(static_cast<const uint8_t*>(roaring_bytes), len, cs, msg);
  if (err && err_cap) { std::snprintf(err, err_cap, "%s", ok ? "" : msg.c_str()); }
  if (!ok) return PB200_ERR_FORMAT;
  FragImage img;
  if (!build_image(cs, img, msg)) { if (err && err_cap) std::snprintf(err, err_cap, "%s", msg.c_str()); return PB200_ERR_FORMAT; }
  uint64_t bits = 0;
  for (auto& c : cs) bits += c.n;
  if (n_containers) *n_containers = img.n_containers;
  if (n_bits) *n_bits = bits;
  return PB200_OK;
}

extern "C" int pb200_frag_load(pb200_ctx* ctx, pb200_family family, uint64_t shard, const void* roaring_bytes, size_t len) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (!roaring_bytes && len) return PB200_ERR_INVALID;
  std::vector<HostContainer> cs;
  std::string err;
  if (!parse_fragment(static_cast<const uint8_t*>(roaring_bytes), len, cs, err)) return fail(ctx, PB200_ERR_FORMAT, err);
  return upload_fragment(ctx, ctx->families[family], shard, cs);
}

// Rebuild the host view of a resident fragment from its device blob (no host shadow is kept).
static int download_containers(pb200_ctx* ctx, const Fragment& f, std::vector<uint8_t>& blob, std::vector<HostContainer>& cs) {
  blob.resize(f.blob_bytes);
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaMemcpy(blob.data(), f.blob, f.blob_bytes, cudaMemcpyDeviceToHost));
  const auto* descs = reinterpret_cast<const ContainerDesc*>(blob.data() + f.descs_off);
  const uint64_t base = reinterpret_cast<uint64_t>(f.blob);
  cs.clear();
  for (uint32_t r = 0; r < f.dev.nrows; r++) {
    uint64_t row = f.dev.dense ? r : f.row_ids[r];
    for (int c = 0; c < kRowContainers; c++) {
      const ContainerDesc& d = descs[size_t(r) * kRowContainers + c];
      if (!d.n) continue;
      HostContainer hc;
      hc.key = row * kRowContainers + c;
      hc.type = desc_type(d); hc.n = d.n; hc.len = desc_len(d);
      hc.data = blob.data() + (d.addr - base);
      cs.push_back(std::move(hc));
    }
  }
  return PB200_OK;
}

// Replace the device copy of a fragment by `cs` and bring its rank cache along: the rankCache state (built on first
// use) is carried over the rewrite, `update` applies the write's BulkAdd calls to it, then Recalculate.
template <class Update>
static int rewrite_fragment(pb200_ctx* ctx, Family& fam, uint64_t shard, const std::vector<HostContainer>& cs, Update&& update) {
  Fragment::RankState st;
  st.live = true;  // a fragment created by this write starts from an empty cache with threshold 0 (NewRankCache cache.go:158)
  auto it = fam.frags.find(shard);
  if (it != fam.frags.end()) {
    rank_state_build(it->second, ctx->rank_cache_size);
    st = std::move(it->second.rank_state);
    it->second.rank_state = Fragment::RankState{};
  }
  const int rc = upload_fragment(ctx, fam, shard, cs);
  it = fam.frags.find(shard);
  if (it == fam.frags.end()) return rc;
  Fragment& f = it->second;
  f.rank_state = std::move(st);
  if (rc != PB200_OK) return rc;  // the old copy is still in place, with its cache
  update(f);
  rank_recalculate(f, ctx->rank_cache_size);
  return PB200_OK;
}

static void rank_bulk_add(Fragment& f, uint64_t id, uint64_t n) {  // rankCache.BulkAdd cache.go:184-192
  if (n < f.rank_state.threshold_value) return;
  f.rank_state.entries[id] = n;
}

// Host-merge form of a position write: only used to CREATE a fragment (the shard had none); writes to a resident
// fragment run on the device (engine_ingest.inl).
static int frag_apply_host(pb200_ctx* ctx, pb200_family family, uint64_t shard, const uint64_t* set_pos, size_t n_set,
                           const uint64_t* clear_pos, size_t n_clear) {
  Family& fam = ctx->families[family];
  std::vector<uint8_t> blob;
  std::vector<HostContainer> cs;
  auto it = fam.frags.find(shard);
  if (it != fam.frags.end()) TRY(download_containers(ctx, it->second, blob, cs));
  std::map<uint64_t, HostContainer> m;
  for (auto& c : cs) m.emplace(c.key, std::move(c));
  if (n_set) apply_bits(m, set_pos, n_set, true);
  if (n_clear) apply_bits(m, clear_pos, n_clear, false);
  cs.clear();
  for (auto& kv : m) {
    cs.push_back(std::move(kv.second));
    if (!cs.back().owned.empty()) cs.back().data = cs.back().owned.data();
  }
  // importPositions' recount (fragment.go:2080-2093): the cached count of every row a position names is re-read
  // from storage (BulkAdd), then Recalculate
  std::set<uint64_t> row_set;
  for (size_t i = 0; i < n_set; i++) row_set.insert(set_pos[i] >> 20);
  for (size_t i = 0; i < n_clear; i++) row_set.insert(clear_pos[i] >> 20);
  return rewrite_fragment(ctx, fam, shard, cs, [&](Fragment& f) {
    for (uint64_t row : row_set) {
      const auto r = std::lower_bound(f.row_ids.begin(), f.row_ids.end(), row);
      rank_bulk_add(f, row, (r != f.row_ids.end() && *r == row) ? f.row_counts[size_t(r - f.row_ids.begin())] : 0);
    }
  });
}

extern "C" int pb200_set_option(pb200_ctx* ctx, uint32_t option, uint64_t value) {
  ENTER(ctx);
  switch (option) {
    case PB200_OPT_DENSIFY_MIN_BYTES:
      if (value > kBitmapBytes) return fail(ctx, PB200_ERR_INVALID, "densify threshold above 8192 bytes");
      ctx->densify_min_bytes = uint32_t(value);
      return PB200_OK;
    case PB200_OPT_PEER_ALLREDUCE:
      if (value && !ctx->peer.on) return fail(ctx, PB200_ERR_INVALID, "no peers connected (pb200_peer_connect)");
      ctx->peer_allreduce = value != 0;
      return PB200_OK;
    case PB200_OPT_TOPN_PRUNE:
      ctx->topn_prune = value != 0;
      return PB200_OK;
    case PB200_OPT_RANK_CACHE_SIZE:
      if (value == 0 || value > 0xffffffffull) return fail(ctx, PB200_ERR_INVALID, "rank cache size out of range");
      ctx->rank_cache_size = uint32_t(value);
      return PB200_OK;
    default: return fail(ctx, PB200_ERR_INVALID, "unknown option");
  }
}

extern "C" int pb200_host_register(pb200_ctx* ctx, void* ptr, size_t len) {
  ENTER(ctx);
  if (!ptr || !len) return PB200_ERR_INVALID;
  CUDA_TRY(ctx, cudaHostRegister(ptr, len, cudaHostRegisterDefault));
  return PB200_OK;
}

extern "C" int pb200_host_unregister(pb200_ctx* ctx, void* ptr) {
  ENTER(ctx);
  if (!ptr) return PB200_ERR_INVALID;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaHostUnregister(ptr));
  return PB200_OK;
}

// Host-merge form of an import (creates the fragment); see frag_apply_host.
static int frag_import_host(pb200_ctx* ctx, pb200_family family, uint64_t shard, const std::vector<HostContainer>& other, int clear, int64_t* changed) {
  Family& fam = ctx->families[family];
  std::vector<uint8_t> blob;
  std::vector<HostContainer> cs;
  auto it = fam.frags.find(shard);
  if (it != fam.frags.end()) TRY(download_containers(ctx, it->second, blob, cs));
  std::map<uint64_t, HostContainer> m;
  for (auto& c : cs) m.emplace(c.key, std::move(c));
  std::map<uint64_t, int64_t> row_set;
  const int64_t n = merge_containers(m, other, clear != 0, &row_set);
  if (changed) *changed = n;
  if (n == 0 && it != fam.frags.end()) return PB200_OK;  // nothing to rewrite, no row changed
  cs.clear();
  for (auto& kv : m) {
    cs.push_back(std::move(kv.second));
    if (!cs.back().owned.empty()) cs.back().data = cs.back().owned.data();
  }
  // fragment.importRoaring (fragment.go:2258-2276): BulkAdd(row, Get(row) + changes) for every changed row
  // (uint64 arithmetic), then Recalculate
  return rewrite_fragment(ctx, fam, shard, cs, [&](Fragment& f) {
    for (auto& kv : row_set) {
      if (kv.second == 0) continue;
      const auto e = f.rank_state.entries.find(kv.first);
      rank_bulk_add(f, kv.first, (e == f.rank_state.entries.end() ? 0 : e->second) + uint64_t(kv.second));
    }
  });
}

extern "C" int pb200_rank_get(pb200_ctx* ctx, pb200_family family, uint64_t shard, uint64_t* row_ids, uint64_t* counts, size_t cap, size_t* n) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (!n) return PB200_ERR_INVALID;
  auto it = ctx->families[family].frags.find(shard);
  if (it == ctx->families[family].frags.end()) return fail(ctx, PB200_ERR_NOTFOUND, "no such fragment");
  const Fragment& f = it->second;
  *n = f.rank_ids.size();
  if (!row_ids && !counts) return PB200_OK;
  if (cap < f.rank_ids.size()) return fail(ctx, PB200_ERR_RANGE, "rank list larger than the buffer");
  for (size_t k = 0; k < f.rank_ids.size(); k++) {
    if (row_ids) row_ids[k] = f.rank_ids[k];
    if (counts) counts[k] = f.rank_counts[k];
  }
  return PB200_OK;
}

extern "C" int pb200_frag_drop(pb200_ctx* ctx, pb200_family family, uint64_t shard) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  Family& fam = ctx->families[family];
  auto it = fam.frags.find(shard);
  if (it == fam.frags.end()) return fail(ctx, PB200_ERR_NOTFOUND, "no such fragment");
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  FragDev empty{};
  fam.h_frags[shard] = empty;
  CUDA_TRY(ctx, cudaMemcpy(fam.d_frags + shard, &empty, sizeof(FragDev), cudaMemcpyHostToDevice));
  if (ctx->publish) TRY(ctx->publish(family, shard, empty, true, false));
  if (!it->second.all_bitmap) fam.n_not_all_bitmap--;
  fam.payload_total -= it->second.payload_bytes; fam.containers_total -= it->second.n_containers;
  ctx->used -= it->second.blob_bytes;
  arena_detach(ctx, it->second);
  free_fragment(it->second);
  fam.frags.erase(it);
  ctx->generation++;
  return PB200_OK;
}

extern "C" int pb200_rank_set(pb200_ctx* ctx, pb200_family family, uint64_t shard, const uint64_t* row_ids, const uint64_t* counts, size_t n) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (n && (!row_ids || !counts)) return PB200_ERR_INVALID;
  auto it = ctx->families[family].frags.find(shard);
  if (it == ctx->families[family].frags.end()) return fail(ctx, PB200_ERR_NOTFOUND, "no such fragment");
  it->second.rank_ids.assign(row_ids, row_ids + n);
  it->second.rank_counts.assign(counts, counts + n);
  it->second.rank_from_host = true;
  it->second.rank_state = Fragment::RankState{};  // the host's list replaces whatever the import recount kept
  it->second.d_rank_valid = false;  // rebuilt on the next TopN
  it->second.rank_pos_valid = false;
  return PB200_OK;
}

extern "C" int pb200_frag_stat(pb200_ctx* ctx, pb200_family family, uint64_t shard, uint64_t* n_rows, uint64_t* n_containers, uint64_t* payload_bytes) {
  ENTER(ctx);
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  auto it = ctx->families[family].frags.find(shard);
  if (it == ctx->families[family].frags.end()) return fail(ctx, PB200_ERR_NOTFOUND, "no such fragment");
  if (n_rows) *n_rows = it->second.row_ids.size();
  if (n_containers) *n_containers = it->second.n_containers;
  if (payload_bytes) *payload_bytes = it->second.payload_bytes;
  return PB200_OK;
}

// ---------------------------------------------------------------------------------------
// program compiler: post-order pb200_node tree -> accumulator machine (layout.h)
// ---------------------------------------------------------------------------------------
namespace {

struct TNode {
  uint32_t op = 0;  // pb200_op of an n-ary node, or 0 for a leaf, or PB200_OP_NOT
  int leaf = -1;    // leaf index (leaf nodes; NOT: the existence leaf)
  uint32_t kid0 = 0, n_kids = 0;  // children: Compiler::kid_pool[kid0 .. kid0 + n_kids) — no allocation per node
};

constexpr uint32_t kOpInternalTemp = 0x100;  // leaf = dense temp row slot `row`; engine-internal (GroupBy), refused from the C ABI

struct Compiler {
  pb200_ctx* ctx;
  Compiled& out;
  bool internal_ok = false;
  uint32_t leaf_base = 0;
  std::vector<TNode> tree;
  std::vector<int> kid_pool;      // children of the n-ary nodes of the call being compiled
  std::vector<int> st;            // operand stack of compile_once ...
  std::vector<Instr> ins_buf;     // ... and the instruction list of finish: reused from call to call of a batch
  std::string err;

  Compiler(pb200_ctx* c, Compiled& o) : ctx(c), out(o) {}
  int kid(const TNode& t, size_t k) const { return kid_pool[t.kid0 + k]; }
  void set_kids(TNode& t, const int* b, const int* e2) { t.kid0 = uint32_t(kid_pool.size()); t.n_kids = uint32_t(e2 - b); kid_pool.insert(kid_pool.end(), b, e2); }

  int add_leaf(uint32_t kind, uint32_t family, uint64_t row) {
    Leaf lf; lf.kind = kind; lf.family = family; lf.row = row;
    out.leaves.push_back(lf);
    if (kind == L_TEMP) { out.may_need_staging = true; out.leaf_payload += kBitmapBytes; out.leaf_containers += 1; }
    else if (kind == L_ROW && family < ctx->families.size()) {
      const Family& fam = ctx->families[family];
      if (fam.n_not_all_bitmap) out.may_need_staging = true;
      out.leaf_payload += fam.payload_total; out.leaf_containers += fam.containers_total;
    }
    return int(out.leaves.size() - 1 - leaf_base);
  }
  int leaf_node(int leaf) { TNode t; t.leaf = leaf; tree.push_back(t); return int(tree.size() - 1); }

  void emit(int n, std::vector<Instr>& ins) {
    const TNode& t = tree[n];
    if (t.op == 0) { ins.push_back({I_LOAD, uint8_t(t.leaf)}); return; }
    if (t.op == PB200_OP_NOT) { emit(kid(t, 0), ins); ins.push_back({I_RANDNOT_LEAF, uint8_t(t.leaf)}); return; }
    uint8_t lop, sop;
    switch (t.op) {
      case PB200_OP_INTERSECT: lop = I_AND_LEAF; sop = I_AND_STACK; break;
      case PB200_OP_UNION: lop = I_OR_LEAF; sop = I_OR_STACK; break;
      case PB200_OP_XOR: lop = I_XOR_LEAF; sop = I_XOR_STACK; break;
      default: lop = I_ANDNOT_LEAF; sop = I_ANDNOT_STACK; break;
    }
    emit(kid(t, 0), ins);
    for (size_t k = 1; k < t.n_kids; k++) {
      const TNode& ch = tree[kid(t, k)];
      if (ch.op == 0) ins.push_back({lop, uint8_t(ch.leaf)});
      else { emit(kid(t, k), ins); ins.push_back({sop, 0}); }
    }
  }

  // Emits the program of subtree `root` as a new Prog; returns false on limits.
  bool finish(int root, Prog& pg) {
    std::vector<Instr>& ins = ins_buf;
    ins.clear();
    emit(root, ins);
    int sp = 0, maxsp = 0; bool live = false;
    for (auto& i : ins) {
      if (i.op == I_LOAD) { if (live) { sp++; maxsp = std::max(maxsp, sp); } live = true; }
      else if (i.op >= I_AND_STACK) sp--;
    }
    if (maxsp > kMaxStack) { err = "expression too deeply nested (stack depth > 4)"; return false; }
    if (ins.size() > size_t(kMaxInstr)) { err = "expression too long (> 64 instructions)"; return false; }
    bool flat = !ins.empty() && ins[0].op == I_LOAD;
    for (size_t k = 1; k < ins.size(); k++)
      if (!fold_op_ok(ins[k].op)) flat = false;  // what fold_can_take refuses
    (flat ? out.any_flat : out.any_bushy) = true;
    pg.instr_off = uint32_t(out.instrs.size());
    pg.n_instr = uint32_t(ins.size());
    pg.leaf_off = leaf_base;
    pg.n_leaves = uint32_t(out.leaves.size() - leaf_base);
    if (pg.n_leaves > uint32_t(kMaxLeaves)) { err = "too many leaves in one expression (> 32)"; return false; }
    out.instrs.insert(out.instrs.end(), ins.begin(), ins.end());
    return true;
  }

  // Compiles one post-order program; appends a Prog to out.progs. A call whose leaves do not fit one program
  // (kMaxLeaves) is compiled a second time with its wide n-ary operations over plain rows evaluated chunk-wise
  // into dense temp rows first — the reference has no limit on the operands of Union / Intersect / Xor /
  // Difference (executor.go:1667-1735), and a time-range Row is such a Union over many views (:1501-1531).
  bool compile(const pb200_node* nodes, uint32_t n) {
    const size_t s_leaves = out.leaves.size(), s_instrs = out.instrs.size(), s_pre = out.pre.size(), s_progs = out.progs.size();
    const uint32_t s_slots = out.n_temp_slots;
    split_min = 0;
    if (compile_once(nodes, n)) return true;
    if (err.rfind("too many leaves", 0) != 0 && err.rfind("expression too long", 0) != 0) return false;
    out.leaves.resize(s_leaves); out.instrs.resize(s_instrs); out.pre.resize(s_pre); out.progs.resize(s_progs);
    out.n_temp_slots = s_slots;
    split_min = 3;
    return compile_once(nodes, n);
  }

  uint32_t split_min = 0;  // 0: never; else n-ary operations over at least this many plain Row leaves go through temps
  size_t pre_base = 0;     // first pre-task of the call being compiled

  // The trailing operands that are plain Row leaves (and the tail of the leaf table) get leaf windows of their own
  // below the call's window and are evaluated into temp slots that take their place as operands: a Union (also the
  // subtrahends of a Difference) as ONE sub-program of any width (union_rows_kernel), Intersect / Xor as one
  // sub-program of the same operation per chunk of kMaxLeaves.
  bool wide_split(const pb200_node& nd, std::vector<int>& st) {
    const bool diff = nd.op == PB200_OP_DIFFERENCE;
    const size_t n_rel = out.leaves.size() - leaf_base;
    uint32_t m = 0;  // the longest run of trailing operands that are plain Row leaves at the tail of the leaf table
    while (m < nd.nargs - (diff ? 1u : 0u) && m < n_rel) {
      const TNode& tn = tree[st[st.size() - 1 - m]];
      if (tn.op != 0 || tn.leaf != int(n_rel - 1 - m) || out.leaves[leaf_base + tn.leaf].kind != L_ROW) break;
      m++;
    }
    if (m < 2) return false;
    const uint32_t lead = nd.nargs - m;  // operands before the run stay operands (Difference: its minuend among them)
    // [leaves of the call so far | m wide leaves] -> [m wide leaves | leaves of the call so far]: the call's window moves up by m
    std::rotate(out.leaves.begin() + leaf_base, out.leaves.end() - m, out.leaves.end());
    for (size_t p = pre_base; p < out.pre.size(); p++)
      if (out.pre[p].kind == 2 && out.pre[p].prog.leaf_off == leaf_base) out.pre[p].prog.leaf_off += m;  // Shift operands compiled earlier
    const uint32_t wide_base = leaf_base;
    leaf_base += m;
    st.resize(st.size() - m);
    const uint8_t lop = nd.op == PB200_OP_INTERSECT ? I_AND_LEAF : nd.op == PB200_OP_XOR ? I_XOR_LEAF : I_OR_LEAF;
    std::vector<int> temps;
    const bool one_pass = lop == I_OR_LEAF || lop == I_XOR_LEAF;  // OR and XOR commute and need no operand kept apart
    if (one_pass) {  // a Union (Xor) of rows is one accumulating pass whatever its width (union_rows_kernel)
      Compiled::PreTask ev; ev.kind = 2; ev.slot = out.n_temp_slots++;
      ev.prog.instr_off = uint32_t(out.instrs.size()); ev.prog.n_instr = 0;
      ev.prog.leaf_off = wide_base; ev.prog.n_leaves = m;
      ev.union_rows = true;
      ev.xor_rows = lop == I_XOR_LEAF;
      out.pre.push_back(ev);
      temps.push_back(leaf_node(add_leaf(L_TEMP, ev.slot, 0)));
    }
    for (uint32_t b = 0; !one_pass && b < m; b += uint32_t(kMaxLeaves)) {
      const uint32_t cnt = std::min<uint32_t>(uint32_t(kMaxLeaves), m - b);
      Compiled::PreTask ev; ev.kind = 2; ev.slot = out.n_temp_slots++;
      ev.prog.instr_off = uint32_t(out.instrs.size()); ev.prog.n_instr = cnt;
      ev.prog.leaf_off = wide_base + b; ev.prog.n_leaves = cnt;
      out.instrs.push_back({I_LOAD, 0});
      for (uint32_t j = 1; j < cnt; j++) out.instrs.push_back({lop, uint8_t(j)});
      out.pre.push_back(ev);
      temps.push_back(leaf_node(add_leaf(L_TEMP, ev.slot, 0)));
    }
    if (lead == 0 && temps.size() == 1) { st.push_back(temps[0]); return true; }
    TNode t; t.op = nd.op;
    std::vector<int> ks(st.end() - lead, st.end());
    st.resize(st.size() - lead);
    ks.insert(ks.end(), temps.begin(), temps.end());
    set_kids(t, ks.data(), ks.data() + ks.size());
    tree.push_back(t);
    st.push_back(int(tree.size() - 1));
    return true;
  }

  bool compile_once(const pb200_node* nodes, uint32_t n) {
    if (!nodes || n == 0) { err = "empty program"; return false; }
    leaf_base = uint32_t(out.leaves.size());
    pre_base = out.pre.size();
    tree.clear();
    kid_pool.clear();
    st.clear();
    uint32_t n_shifts = 0;
    for (uint32_t k = 0; k < n; k++) {
      const pb200_node& nd = nodes[k];
      switch (nd.op) {
        case PB200_OP_ROW:
          if (nd.family >= ctx->families.size()) { err = "unknown family in Row()"; return false; }
          st.push_back(leaf_node(add_leaf(L_ROW, nd.family, nd.row)));
          break;
        case PB200_OP_EMPTY: st.push_back(leaf_node(add_leaf(L_ZERO, 0, 0))); break;
        case kOpInternalTemp:
          if (!internal_ok) { err = "unknown op " + std::to_string(nd.op); return false; }
          st.push_back(leaf_node(add_leaf(L_TEMP, uint32_t(nd.row), 0)));
          break;
        case PB200_OP_BSI_RANGE: {
          if (nd.family >= ctx->families.size()) { err = "unknown family in BSI range"; return false; }
          if (nd.bsi_op < PB200_BSI_EQ || nd.bsi_op > PB200_BSI_NOTNULL) { err = "invalid range operation"; return false; }
          if (nd.bit_depth > 63) { err = "bit depth > 63"; return false; }
          Compiled::PreTask pt; pt.kind = 0; pt.node = nd; pt.slot = out.n_temp_slots++;
          out.pre.push_back(pt);
          st.push_back(leaf_node(add_leaf(L_TEMP, pt.slot, 0)));
          break;
        }
        case PB200_OP_INTERSECT: case PB200_OP_UNION: case PB200_OP_DIFFERENCE: case PB200_OP_XOR: {
          if (nd.nargs == 0) {
            if (nd.op == PB200_OP_UNION || nd.op == PB200_OP_XOR) { st.push_back(leaf_node(add_leaf(L_ZERO, 0, 0))); break; }
            err = nd.op == PB200_OP_INTERSECT ? "empty Intersect query is currently not supported"   // executor.go:1673
                                              : "empty Difference query is currently not supported";  // executor.go:1008
            return false;
          }
          if (st.size() < nd.nargs) { err = "malformed program: operand stack underflow"; return false; }
          if (nd.nargs == 1) break;  // identity
          if (split_min && nd.nargs >= split_min && wide_split(nd, st)) break;
          TNode t; t.op = nd.op;
          set_kids(t, st.data() + (st.size() - nd.nargs), st.data() + st.size());
          st.resize(st.size() - nd.nargs);
          tree.push_back(t);
          st.push_back(int(tree.size() - 1));
          break;
        }
        case PB200_OP_NOT: {
          if (st.empty()) { err = "Not() requires an input row"; return false; }  // executor.go:1741
          if (nd.family >= ctx->families.size()) { err = "index does not support existence tracking"; return false; }
          TNode t; t.op = PB200_OP_NOT; t.leaf = add_leaf(L_ROW, nd.family, 0);
          set_kids(t, &st.back(), &st.back() + 1);
          st.pop_back();
          tree.push_back(t);
          st.push_back(int(tree.size() - 1));
          break;
        }
        case PB200_OP_SHIFT: {
          if (st.empty()) { err = "Shift() requires an input row"; return false; }  // executor.go:1777
          if (++n_shifts > 64) { err = "too many Shift steps in one call (limit 64: the carried bits live in one 64-bit word per shard)"; return false; }
          // materialise the operand, then shift it by one with carry across containers
          Compiled::PreTask ev; ev.kind = 2; ev.slot = out.n_temp_slots++;
          // The operand's leaves sit in [leaf_base, now): compile it as its own program over the same table.
          if (!finish(st.back(), ev.prog)) return false;
          st.pop_back();
          out.pre.push_back(ev);
          Compiled::PreTask sh; sh.kind = 1; sh.src_slot = ev.slot; sh.slot = out.n_temp_slots++;
          out.pre.push_back(sh);
          st.push_back(leaf_node(add_leaf(L_TEMP, sh.slot, 0)));
          break;
        }
        default: err = "unknown op " + std::to_string(nd.op); return false;
      }
    }
    if (st.size() != 1) { err = "malformed program: " + std::to_string(st.size()) + " values left on the stack"; return false; }
    Prog pg;
    if (!finish(st[0], pg)) return false;
    out.progs.push_back(pg);
    return true;
  }
};

}  // namespace

// ---------------------------------------------------------------------------------------
// batch upload + evaluation
// ---------------------------------------------------------------------------------------
namespace {

struct BatchDev {  // device pointers of an uploaded batch
  const Prog* progs; const Instr* instrs; const Leaf* leaves; const uint64_t* shards;
};

int upload_batch(pb200_ctx* ctx, const Compiled& cp, const uint64_t* shards, uint32_t n_shards, BatchDev& bd) {
  size_t o_prog = 0;
  size_t o_leaf = align_up(o_prog + cp.progs.size() * sizeof(Prog) + cp.pre.size() * sizeof(Prog), 16);
  size_t o_ins = align_up(o_leaf + cp.leaves.size() * sizeof(Leaf), 16);
  size_t o_sh = align_up(o_ins + cp.instrs.size() * sizeof(Instr), 16);
  size_t total = o_sh + size_t(n_shards) * 8;
  TRY(ensure_pin(ctx, ctx->h_stage, total));
  TRY(ensure_dev(ctx, ctx->d_batch, total));
  uint8_t* h = static_cast<uint8_t*>(ctx->h_stage.p);
  std::memcpy(h + o_prog, cp.progs.data(), cp.progs.size() * sizeof(Prog));
  Prog* pre_progs = reinterpret_cast<Prog*>(h + o_prog + cp.progs.size() * sizeof(Prog));
  for (size_t i = 0; i < cp.pre.size(); i++) pre_progs[i] = cp.pre[i].prog;
  std::memcpy(h + o_leaf, cp.leaves.data(), cp.leaves.size() * sizeof(Leaf));
  std::memcpy(h + o_ins, cp.instrs.data(), cp.instrs.size() * sizeof(Instr));
  std::memcpy(h + o_sh, shards, size_t(n_shards) * 8);
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_batch.p, h, total, cudaMemcpyHostToDevice, ctx->stream));
  uint8_t* d = static_cast<uint8_t*>(ctx->d_batch.p);
  bd.progs = reinterpret_cast<const Prog*>(d + o_prog);
  bd.leaves = reinterpret_cast<const Leaf*>(d + o_leaf);
  bd.instrs = reinterpret_cast<const Instr*>(d + o_ins);
  bd.shards = reinterpret_cast<const uint64_t*>(d + o_sh);
  ctx->timing.h2d_bytes += total;
  return PB200_OK;
}

// Where the kernels of the call being enqueued add their algorithmic bytes: a Count call keeps the counter (and the
// deferred-item counters) behind its counts, so that one memset clears and one copy fetches everything; other calls
// use the context's word.
unsigned long long* algo_ptr(pb200_ctx* ctx) {
  return ctx->algo_dev ? ctx->algo_dev : reinterpret_cast<unsigned long long*>(static_cast<uint8_t*>(ctx->d_misc.p) + 8);
}

// Persistent grids are sized to exactly the number of CTAs that can be resident (SMs x CTAs
// per SM from the occupancy calculator): a grid larger than that runs a second, thinner wave.
template <class K>
uint32_t resident_ctas(pb200_ctx* ctx, int slot, K kernel, int block, size_t dyn_smem) {
  if (ctx->occ[slot] == 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, block, dyn_smem) != cudaSuccess || n < 1) n = 1;
    ctx->occ[slot] = n;
  }
  return uint32_t(ctx->sm_count) * uint32_t(ctx->occ[slot]);
}
enum { OCC_FAST = 0, OCC_STAGED, OCC_SUM, OCC_MM, OCC_TOPN, OCC_FOLD, OCC_TINY, OCC_FOLDW };

// Launches the fast + staged kernels for `n_queries` programs starting at `progs`.
int launch_eval(pb200_ctx* ctx, const BatchDev& bd, const Prog* progs, uint32_t n_queries, uint32_t n_shards, bool staging,
                unsigned long long* out_counts, uint32_t qstride, uint32_t sstride, uint8_t* dense_out, uint32_t* dense_n,
                uint32_t max_leaves = kMaxLeaves, bool bushy = true, bool fold_all = false, const PushDesc* push = nullptr, bool flat = true,
                bool tiny = false, uint16_t* sparse_out = nullptr) {
  const uint64_t n_items = uint64_t(n_queries) * n_shards * kRowContainers;
  if (n_items == 0) return PB200_OK;
  if (n_items >= (1ull << 32)) return fail(ctx, PB200_ERR_RANGE, "batch too large: queries x shards x 16 must be < 2^32");
  // Containers of a kilobyte and more on average (mixed array / bitmap / run data): nearly every item would be deferred,
  // so the fold evaluator walks all items itself and the first pass with its deferred list is skipped.
  fold_all = fold_all && staging;
  // scratch: slow list + counter + algo-bytes counter
  size_t slow_bytes = staging && !fold_all ? size_t(n_items) * 4 : 4;
  TRY(ensure_dev(ctx, ctx->d_slow, slow_bytes));
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  EvalParams p{};
  p.families = ctx->d_families;
  p.n_families = uint32_t(ctx->families.size());
  p.progs = progs; p.instrs = bd.instrs; p.leaves = bd.leaves; p.shards = bd.shards;
  p.n_shards = n_shards; p.n_queries = n_queries;
  p.temp_base = static_cast<const uint8_t*>(ctx->d_temp.p);
  p.out_counts = out_counts; p.out_qstride = qstride; p.out_sstride = sstride;
  p.dense_out = dense_out; p.dense_n = dense_n; p.sparse_out = sparse_out;
  p.slow_list = static_cast<uint32_t*>(ctx->d_slow.p);
  p.slow_count = ctx->slow_dev ? ctx->slow_dev : static_cast<uint32_t*>(ctx->d_misc.p);
  p.algo_bytes = algo_ptr(ctx);
  // descriptors of deferred items travel from the fast to the staged kernel when they fit a modest scratch budget
  const uint64_t desc_bytes = n_items * uint64_t(max_leaves) * sizeof(ContainerDesc);
  p.fold_all = fold_all ? 1u : 0u;
  if (staging && !fold_all && desc_bytes <= (256ull << 20)) {
    TRY(ensure_dev(ctx, ctx->d_slow_descs, size_t(desc_bytes)));
    p.slow_descs = static_cast<ContainerDesc*>(ctx->d_slow_descs.p);
    p.slow_desc_stride = max_leaves;
  }
  const bool push_fused = push && !staging;  // eval_fast_kernel is the batch's only kernel: the peer reduction rides in its tail
  if (push_fused) p.push = *push;
  // Very sparse rows (containers of a few values): a first pass with one lane per item counts the pairs of tiny arrays
  // in registers and leaves eval_fast_kernel a list of the rest.
  // (only when the items outnumber the warps the device holds at once: below that a warp per item is one wave anyway)
  tiny = tiny && staging && !fold_all && (dense_out ? (sparse_out && dense_n && n_queries == 1) : out_counts != nullptr) && n_items > uint64_t(ctx->sm_count) * 64;
  if (tiny) TRY(ensure_dev(ctx, ctx->d_rest, size_t(n_items) * 4));
  if (!fold_all) {
    // slow_count and, behind it, the tiny pass's item count (a Count call's first launch finds them cleared with its counts)
    if (ctx->slow_zeroed) ctx->slow_zeroed = false;
    else CUDA_TRY(ctx, cudaMemsetAsync(p.slow_count, 0, 8, ctx->stream));
    if (tiny) {
      p.item_list = static_cast<uint32_t*>(ctx->d_rest.p);
      p.item_count = p.slow_count + 1;
      const uint32_t tgrid = uint32_t(std::min<uint64_t>((n_items + 255) / 256, resident_ctas(ctx, OCC_TINY, eval_tiny_kernel, 256, 0)));
      eval_tiny_kernel<<<tgrid, 256, 0, ctx->stream>>>(p);
      ctx->timing.launches++;
    }
    const uint64_t want = (n_items + kFastWarps - 1) / kFastWarps;
    uint32_t resident = resident_ctas(ctx, OCC_FAST, eval_fast_kernel, kFastWarps * 32, 0);
    if (push && resident > 1) resident -= 1;  // one CTA slot stays free: the one-CTA collect kernel of an older batch can run beside this grid
    const uint32_t grid = uint32_t(std::min<uint64_t>(want, resident));
    eval_fast_kernel<<<grid, kFastWarps * 32, 0, ctx->stream>>>(p);
    ctx->timing.launches++;
  }
  if (staging) {  // deferred items: flat left folds go to the fold kernel, anything else to the general staged evaluator
    if (flat) {
#if PB200_FOLDW
      if (!dense_out && !dense_n) {  // counts only: the consumers need no barrier among themselves (kernels_foldw.cuh)
        const uint32_t fgrid = uint32_t(std::min<uint64_t>(n_items, resident_ctas(ctx, OCC_FOLDW, eval_foldw_kernel, kFoldCtaThreads, kFoldSmem)));
        eval_foldw_kernel<<<fgrid, kFoldCtaThreads, kFoldSmem, ctx->stream>>>(p);
      } else
#endif
      {
        const uint32_t fgrid = uint32_t(std::min<uint64_t>(n_items, resident_ctas(ctx, OCC_FOLD, eval_fold_kernel, kFoldCtaThreads, kFoldSmem)));
        eval_fold_kernel<<<fgrid, kFoldCtaThreads, kFoldSmem, ctx->stream>>>(p);
      }
      ctx->timing.launches++;
    }
    if (bushy) {
      const uint32_t sgrid = uint32_t(std::min<uint64_t>(n_items, resident_ctas(ctx, OCC_STAGED, eval_staged_kernel, kStagedThreads, 0)));
      eval_staged_kernel<<<sgrid, kStagedThreads, 0, ctx->stream>>>(p);
      ctx->timing.launches++;
    }
  }
  if (push && !push_fused) {
    peer_push_kernel<<<1, 256, 0, ctx->stream>>>(*push);
    ctx->timing.launches++;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

int validate_shards(pb200_ctx* ctx, const uint64_t* shards, uint32_t n_shards) {
  if (n_shards && !shards) return fail(ctx, PB200_ERR_INVALID, "null shard list");
  return PB200_OK;
}

void timing_begin(pb200_ctx* ctx) {
  ctx->timing = pb200_timing{};
  cudaMemsetAsync(static_cast<uint8_t*>(ctx->d_misc.p) + 8, 0, 8, ctx->stream);
}

// Runs the pre-tasks (BSI range / shift / sub-expression temps) of a compiled batch.
int run_pretasks(pb200_ctx* ctx, const Compiled& cp, const BatchDev& bd, uint32_t n_shards);
bool has_shift(const Compiled& cp);
int launch_overflow(pb200_ctx* ctx, const BatchDev& bd, const Prog* progs, uint32_t n_queries, uint32_t n_shards,
                    unsigned long long* out_counts, uint32_t qstride, uint32_t sstride, unsigned long long* ovf_out);

// engine_peer.inl
int peer_begin_epoch(pb200_ctx* ctx, unsigned long long* target, uint32_t n, PushDesc& out, int* slot_out);
int peer_collect(pb200_ctx* ctx, int slot);
int peer_end_epoch(pb200_ctx* ctx, int slot);
int peer_check_error(pb200_ctx* ctx);
int pb200_peer_flush_locked(pb200_ctx* ctx);

// Ends a query call: fetches the algorithmic-byte counter with the other results, waits for
// the stream once, and reads the kernel time between the ev0/ev1 events.
int finish_call(pb200_ctx* ctx) {
  TRY(ensure_pin(ctx, ctx->h_algo, 64));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_algo.p, static_cast<uint8_t*>(ctx->d_misc.p) + 8, 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  float ms = 0;
  if (cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess) ctx->timing.kernel_ms = ms;
  ctx->timing.algo_bytes = *static_cast<unsigned long long*>(ctx->h_algo.p);
  return PB200_OK;
}

// mean payload of the containers the batch's Row leaves come from; at a kilobyte and more the fold evaluator runs alone
bool fold_all_of(const Compiled& cp) {  // (a batch with programs the fold kernel refuses keeps the first pass: all-bitmap items are cheapest there)
  return cp.may_need_staging && !cp.any_bushy && cp.leaf_containers && cp.leaf_payload / cp.leaf_containers >= 1024;
}

// containers of at most a few dozen bytes on average: the lane-per-item first pass pays (eval_tiny_kernel)
bool tiny_of(const Compiled& cp) {
  return cp.may_need_staging && cp.leaf_containers && cp.leaf_payload / cp.leaf_containers <= 64;
}

uint32_t max_leaves_of(const Compiled& cp) {
  uint32_t m = 1;
  for (auto& pg : cp.progs) m = std::max(m, pg.n_leaves);
  for (auto& pt : cp.pre) if (!pt.union_rows) m = std::max(m, pt.prog.n_leaves);
  return m;
}

// Compiles a batch of post-order programs.
int compile_batch(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries, Compiled& cp, bool internal_ok = false) {
  Compiler cc(ctx, cp);
  cc.internal_ok = internal_ok;
  for (uint32_t q = 0; q < n_queries; q++) {
    const uint32_t b = prog_off[q], e = prog_off[q + 1];
    if (e < b) return fail(ctx, PB200_ERR_INVALID, "prog_off not ascending");
    if (!cc.compile(progs + b, e - b)) return fail(ctx, PB200_ERR_INVALID, cc.err);
  }
  return PB200_OK;
}

// First half of a Count call: compile, upload, launch (asynchronous on the stream). The
// counts stay in the slot's device buffer until count_end fetches them. Up to two batches
// may be in flight: the host prepares batch i+1 while the kernels of batch i run.
int count_begin(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries, const uint64_t* shards,
                uint32_t n_shards, bool per_shard) {
  TRY(validate_shards(ctx, shards, n_shards));
  if (ctx->n_pending >= 2) return fail(ctx, PB200_ERR_INVALID, "two Count batches already in flight: call pb200_count_batch_end first");
  // The older batch's device counts were handed out: whatever the caller queued on the compute stream since
  // (an in-place NCCL all-reduce) belongs to that batch, so its `done` moves behind it before new work is queued.
  if (ctx->n_pending == 1 && ctx->slots[ctx->head].external) CUDA_TRY(ctx, cudaEventRecord(ctx->slots[ctx->head].done, ctx->stream));
  const int si = (ctx->head + ctx->n_pending) & 1;
  CountSlot& sl = ctx->slots[si];
  sl.external = false;
  sl.peer_slot = -1;
  if (!sl.k0) {
    CUDA_TRY(ctx, cudaEventCreate(&sl.k0)); CUDA_TRY(ctx, cudaEventCreate(&sl.k1));
    CUDA_TRY(ctx, cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
  }
  if (!ctx->copy_stream) CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  const size_t n_out = per_shard ? size_t(n_queries) * n_shards : n_queries;
  sl.n_out = n_out;
  sl.launched = false;
  sl.seq = ++ctx->begin_seq;
  std::swap(ctx->d_out, sl.d_out);  // ensure_* work on the context's named buffers
  int rc = ensure_dev(ctx, ctx->d_out, (std::max<size_t>(n_out, 1) + 2) * 8);  // [counts][algorithmic bytes][deferred-item counters]
  std::swap(ctx->d_out, sl.d_out);
  TRY(rc);
  std::swap(ctx->h_out, sl.h_out);
  rc = ensure_pin(ctx, ctx->h_out, std::max<size_t>(n_out, 1) * 8 + 8);
  std::swap(ctx->h_out, sl.h_out);
  TRY(rc);
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  ctx->timing = pb200_timing{};
  ctx->n_pending++;
  unsigned long long* d_out = static_cast<unsigned long long*>(sl.d_out.p);
  CUDA_TRY(ctx, cudaEventRecord(sl.k0, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(d_out, 0, (std::max<size_t>(n_out, 1) + 2) * 8, ctx->stream));
  struct Counters {  // the slot's own counters are the context's for the duration of this call
    pb200_ctx* c;
    Counters(pb200_ctx* cc, unsigned long long* base, size_t n) : c(cc) {
      c->algo_dev = base + n; c->slow_dev = reinterpret_cast<uint32_t*>(base + n + 1); c->slow_zeroed = true;
    }
    ~Counters() { c->algo_dev = nullptr; c->slow_dev = nullptr; c->slow_zeroed = false; }
  } counters(ctx, d_out, std::max<size_t>(n_out, 1));
  if (n_queries != 0 && n_shards == 0) {
    if (ctx->peer_allreduce && !per_shard) {  // a rank without local shards still takes part in the sum (its zeros, its arrival)
      PushDesc pd{};
      const int rc0 = peer_begin_epoch(ctx, d_out, n_queries, pd, &sl.peer_slot);
      if (rc0 != PB200_OK) { ctx->n_pending--; return rc0; }
      peer_push_kernel<<<1, 256, 0, ctx->stream>>>(pd);
      CUDA_TRY(ctx, cudaGetLastError());
      TRY(peer_end_epoch(ctx, sl.peer_slot));
    }
  }
  if (n_queries != 0 && n_shards != 0) {
    Compiled cp;
    rc = compile_batch(ctx, progs, prog_off, n_queries, cp);
    if (rc != PB200_OK) { ctx->n_pending--; return rc; }
    // Count(Row(v <op> x)) with nothing around it: let the comparator kernel count, skip the dense row
    for (uint32_t q = 0; q < n_queries; q++) {
      const Prog& pg = cp.progs[q];
      if (pg.n_instr != 1) continue;
      Leaf& lf = cp.leaves[pg.leaf_off + cp.instrs[pg.instr_off].leaf];
      if (lf.kind != L_TEMP) continue;
      for (auto& pt : cp.pre)
        if (pt.kind == 0 && pt.slot == lf.family) {
          pt.count_only = true;
          pt.count_out = d_out + (per_shard ? size_t(q) * n_shards : q);
          pt.count_sstride = per_shard ? 1 : 0;
          lf.kind = L_ZERO;  // the evaluator adds nothing for this query
          break;
        }
    }
    rc = sync_families(ctx);
    BatchDev bd{};
    if (rc == PB200_OK) {
      std::swap(ctx->d_batch, sl.d_batch); std::swap(ctx->h_stage, sl.h_stage);  // this slot's own upload buffers
      rc = upload_batch(ctx, cp, shards, n_shards, bd);
      std::swap(ctx->d_batch, sl.d_batch); std::swap(ctx->h_stage, sl.h_stage);
    }
    bool need_temp = false;
    for (auto& pt : cp.pre) need_temp |= !(pt.kind == 0 && pt.count_only);
    if (rc == PB200_OK && need_temp) rc = ensure_dev(ctx, ctx->d_temp, size_t(cp.n_temp_slots) * n_shards * kRowContainers * kBitmapBytes);
    if (rc == PB200_OK) rc = run_pretasks(ctx, cp, bd, n_shards);
    PushDesc pd{};
    const bool reduce = ctx->peer_allreduce && !per_shard;
    if (rc == PB200_OK && reduce) rc = peer_begin_epoch(ctx, d_out, n_queries, pd, &sl.peer_slot);
    if (rc == PB200_OK)
      rc = launch_eval(ctx, bd, bd.progs, n_queries, n_shards, cp.may_need_staging, d_out, per_shard ? n_shards : 1, per_shard ? 1 : 0,
                       nullptr, nullptr, max_leaves_of(cp), cp.any_bushy, fold_all_of(cp), reduce && !has_shift(cp) ? &pd : nullptr, cp.any_flat, tiny_of(cp));
    if (rc == PB200_OK && reduce && has_shift(cp)) {  // the overflow pass below adds to the counts: push after it
      rc = launch_overflow(ctx, bd, bd.progs, n_queries, n_shards, d_out, 1, 0, nullptr);
      if (rc == PB200_OK) { peer_push_kernel<<<1, 256, 0, ctx->stream>>>(pd); ctx->timing.launches++; }
    } else
    if (rc == PB200_OK && has_shift(cp))  // bits shifted past a shard's last column count too (roaring.go:970-974)
      rc = launch_overflow(ctx, bd, bd.progs, n_queries, n_shards, d_out, per_shard ? n_shards : 1, per_shard ? 1 : 0, nullptr);
    if (rc == PB200_OK && reduce) rc = peer_end_epoch(ctx, sl.peer_slot);
    if (rc != PB200_OK) { ctx->n_pending--; return rc; }
  }
  CUDA_TRY(ctx, cudaEventRecord(sl.k1, ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(sl.done, ctx->stream));
  sl.launched = true;
  return PB200_OK;
}

// Device pointer of the counts of the most recently begun batch (for in-place collectives).
void* count_device_ptr(pb200_ctx* ctx) {
  if (ctx->n_pending == 0) return nullptr;
  return ctx->slots[(ctx->head + ctx->n_pending - 1) & 1].d_out.p;
}

// Second half, oldest batch first: device -> host copy of the counts on the copy stream, so
// that a younger batch already queued on the compute stream keeps the GPU busy meanwhile.
int count_end(pb200_ctx* ctx, uint64_t* out) {
  if (ctx->n_pending == 0) return fail(ctx, PB200_ERR_INVALID, "no Count batch in flight");
  CountSlot& sl = ctx->slots[ctx->head];
  const size_t n_out = sl.n_out;
  if (n_out && !out) return PB200_ERR_INVALID;
  uint8_t* h = static_cast<uint8_t*>(sl.h_out.p);
  const size_t fetch = (std::max<size_t>(n_out, 1) + 1) * 8;  // the counts and, behind them, the algorithmic bytes: one copy
  if (ctx->n_pending == 1 && sl.peer_slot < 0) {
    // Nothing younger is queued and nothing arrives from peers: the compute stream itself carries the copy (whatever the
    // host queued on it after _begin, an in-place NCCL all-reduce, runs before it) — no event, no second stream.
    CUDA_TRY(ctx, cudaMemcpyAsync(h, sl.d_out.p, fetch, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    // Re-record `done` when nothing younger is queued, so that work the host enqueued on the compute stream after
    // _begin is waited for too; the copy runs on the copy stream, beside a younger batch's kernels.
    if (ctx->n_pending == 1) CUDA_TRY(ctx, cudaEventRecord(sl.done, ctx->stream));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copy_stream, sl.done, 0));
    if (sl.peer_slot >= 0) {  // the sum over the ranks: wait for the peers' counts beside whatever the compute stream is running
      TRY(peer_collect(ctx, sl.peer_slot));
      sl.peer_slot = -1;
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(h, sl.d_out.p, fetch, cudaMemcpyDeviceToHost, ctx->copy_stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->copy_stream));
  }
  if (ctx->peer_allreduce) TRY(peer_check_error(ctx));
  float ms = 0;
  if (cudaEventElapsedTime(&ms, sl.k0, sl.k1) == cudaSuccess) ctx->timing.kernel_ms = ms;
  std::memcpy(&ctx->timing.algo_bytes, h + std::max<size_t>(n_out, 1) * 8, 8);
  if (n_out) std::memcpy(out, h, n_out * 8);
  ctx->timing.d2h_bytes = n_out * 8 + 8;
  sl.launched = false;
  ctx->head ^= 1;
  ctx->n_pending--;
  return PB200_OK;
}

int count_impl(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries, const uint64_t* shards,
               uint32_t n_shards, bool per_shard, uint64_t* out) {
  if (!out) return PB200_ERR_INVALID;
  TRY(count_begin(ctx, progs, prog_off, n_queries, shards, n_shards, per_shard));
  return count_end(ctx, out);
}

}  // namespace

#include "engine_queries.inl"
#include "engine_groupby.inl"

// ---- pb200_count from many host threads: a combining queue -------------------------------------------------------
// The reference runs workerPoolSize goroutines and many HTTP requests side by side (executor.go:83-107), each of which
// would issue its own Count here. One launch per call serialises them at the per-call cost (~45 us); instead a caller
// files its request and, if nobody is combining, becomes the combiner: it takes every request that has arrived with
// the same shard list, runs them as ONE batch (the batched kernels are what reaches the HBM roofline) and hands the
// results out; requests that arrive meanwhile are taken by the next round. A lone caller pays one extra mutex.
namespace {

constexpr size_t kCombineMax = 1024;
#ifndef PB200_GATHER_US
#define PB200_GATHER_US 12
#endif
constexpr int kGatherMaxUs = PB200_GATHER_US;  // longest wait for the callers of the last round

void run_count_group(pb200_ctx* ctx, std::vector<CountReq*>& grp) {
  std::lock_guard<std::mutex> lock(ctx->mu);
  DeviceGuard guard(ctx->device);
  ctx->err.clear();
  if (grp.size() == 1) {
    CountReq* r = grp[0];
    const uint32_t off[2] = {0, r->n_nodes};
    r->rc = count_impl(ctx, r->prog, off, 1, r->shards, r->n_shards, false, r->out);
    if (r->rc != PB200_OK) r->err = ctx->err;
    return;
  }
  std::vector<pb200_node> flat;
  std::vector<uint32_t> off(grp.size() + 1, 0);
  for (size_t i = 0; i < grp.size(); i++) {
    flat.insert(flat.end(), grp[i]->prog, grp[i]->prog + grp[i]->n_nodes);
    off[i + 1] = uint32_t(flat.size());
  }
  std::vector<uint64_t> out(grp.size(), 0);
  const int rc = count_impl(ctx, flat.data(), off.data(), uint32_t(grp.size()), grp[0]->shards, grp[0]->n_shards, false, out.data());
  if (rc == PB200_OK) {
    for (size_t i = 0; i < grp.size(); i++) { *grp[i]->out = out[i]; grp[i]->rc = PB200_OK; }
    ctx->cq_batches++; ctx->cq_queries += grp.size();
    return;
  }
  // some program of the group is at fault: each call gets its own verdict
  for (CountReq* r : grp) {
    const uint32_t o1[2] = {0, r->n_nodes};
    ctx->err.clear();
    r->rc = count_impl(ctx, r->prog, o1, 1, r->shards, r->n_shards, false, r->out);
    if (r->rc != PB200_OK) r->err = ctx->err;
  }
}

int combine_counts(pb200_ctx* ctx, CountReq& me) {
  // file the request: one compare-and-swap, no lock (callers re-entering together would queue up on a mutex)
  ctx->cq_n.fetch_add(1, std::memory_order_relaxed);  // (before the push: the count never runs behind the list)
  me.next = ctx->cq_head.load(std::memory_order_relaxed);
  while (!ctx->cq_head.compare_exchange_weak(me.next, &me, std::memory_order_release, std::memory_order_relaxed)) {}
  while (!me.done.load(std::memory_order_acquire)) {
    if (!ctx->cq_busy.load(std::memory_order_relaxed) && !ctx->cq_busy.exchange(true, std::memory_order_acquire)) {
      // this thread combines until its own request has been served; a waiting thread takes over then
      while (!me.done.load(std::memory_order_relaxed)) {
        // The callers of the last round are on their way back with their next request: when that round was a batch, give
        // them up to 12 us to arrive so that this launch is as wide — until as many requests wait as callers were seen
        // in and during the last round (the ones it served plus the ones that arrived meanwhile: counting only the
        // former lets the callers settle into two cohorts that alternate, each launch half as wide). A lone caller
        // never waits: last round = 1.
        if (ctx->cq_last > 1 && ctx->cq_n.load(std::memory_order_relaxed) < ctx->cq_target) {
          const auto t0 = std::chrono::steady_clock::now();
          for (int i = 0; i < 4000; i++) {
            if ((i & 7) == 7 && (ctx->cq_n.load(std::memory_order_relaxed) >= ctx->cq_target ||
                                 std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(kGatherMaxUs))) break;
            __builtin_ia32_pause();
          }
        }
        // take the arrivals (newest first) and put them behind what earlier rounds left over, oldest first
        CountReq* lst = ctx->cq_head.exchange(nullptr, std::memory_order_acquire);
        const size_t old = ctx->cq.size();
        for (CountReq* r = lst; r; r = r->next) ctx->cq.push_back(r);
        std::reverse(ctx->cq.begin() + old, ctx->cq.end());
        if (ctx->cq.empty()) break;  // (cannot happen while this thread's own request is unserved)
        std::vector<CountReq*> grp, rest;
        CountReq* head = ctx->cq.front();
        for (CountReq* r : ctx->cq) {
          const bool same = grp.size() < kCombineMax && r->n_shards == head->n_shards && r->n_nodes && r->prog && r->out &&
                            (r->n_shards == 0 || (r->shards && head->shards && std::memcmp(r->shards, head->shards, size_t(r->n_shards) * 8) == 0));
          (r == head || same ? grp : rest).push_back(r);
        }
        ctx->cq.swap(rest);
        ctx->cq_last = grp.size();
        ctx->cq_n.fetch_sub(grp.size(), std::memory_order_relaxed);  // cq_n: requests filed and not yet launched
        run_count_group(ctx, grp);
        ctx->cq_target = grp.size() + ctx->cq_n.load(std::memory_order_relaxed);
        for (CountReq* r : grp) r->done.store(true, std::memory_order_release);  // last access to r: its owner may return now
        if (ctx->cq_sleepers.load(std::memory_order_acquire)) { std::lock_guard<std::mutex> lk(ctx->cq_mu); ctx->cq_cv.notify_all(); }
      }
      ctx->cq_busy.store(false, std::memory_order_release);
      // a caller that sleeps may have been served, or may have to take over
      if (ctx->cq_sleepers.load(std::memory_order_acquire)) { std::lock_guard<std::mutex> lk(ctx->cq_mu); ctx->cq_cv.notify_all(); }
      continue;
    }
    // Somebody else is combining. A launch takes tens of microseconds: spin that long before sleeping — a futex wake-up
    // costs as much as the launch and would leave the GPU idle between rounds.
    bool progress = false;
    const auto s0 = std::chrono::steady_clock::now();
    for (int i = 0; !progress; i++) {
      progress = me.done.load(std::memory_order_acquire) || !ctx->cq_busy.load(std::memory_order_acquire);
      if (progress || ((i & 63) == 63 && std::chrono::steady_clock::now() - s0 > std::chrono::microseconds(150))) break;
      __builtin_ia32_pause();
    }
    if (!progress) {
      std::unique_lock<std::mutex> lk(ctx->cq_mu);
      ctx->cq_sleepers.fetch_add(1, std::memory_order_acq_rel);
      if (!me.done.load(std::memory_order_acquire) && ctx->cq_busy.load(std::memory_order_acquire))
        ctx->cq_cv.wait_for(lk, std::chrono::microseconds(200));
      ctx->cq_sleepers.fetch_sub(1, std::memory_order_acq_rel);
    }
  }
  if (me.rc != PB200_OK) { std::lock_guard<std::mutex> lock(ctx->mu); ctx->err = me.err; }
  return me.rc;
}

}  // namespace

extern "C" int pb200_combine_stats(pb200_ctx* ctx, uint64_t* launches, uint64_t* queries) {
  ENTER(ctx);
  if (launches) *launches = ctx->cq_batches;
  if (queries) *queries = ctx->cq_queries;
  return PB200_OK;
}

extern "C" int pb200_count(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, uint64_t* out_count) {
  if (!ctx) return PB200_ERR_INVALID;
  if (!out_count) return PB200_ERR_INVALID;
  CountReq me;
  me.prog = prog; me.n_nodes = n_nodes; me.shards = shards; me.n_shards = n_shards; me.out = out_count;
  return combine_counts(ctx, me);
}

extern "C" int pb200_count_batch(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                                 const uint64_t* shards, uint32_t n_shards, uint64_t* out_counts) {
  ENTER(ctx);
  if (n_queries && (!progs || !prog_off)) return PB200_ERR_INVALID;
  return count_impl(ctx, progs, prog_off, n_queries, shards, n_shards, false, out_counts);
}

extern "C" int pb200_count_shards(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, uint64_t* out_per_shard) {
  ENTER(ctx);
  const uint32_t off[2] = {0, n_nodes};
  return count_impl(ctx, prog, off, 1, shards, n_shards, true, out_per_shard);
}

// ---- split Count batch: launch now, fetch later (multi-GPU hosts reduce the device-resident
// counts with NCCL between the two calls) ----
extern "C" int pb200_count_batch_begin(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                                       const uint64_t* shards, uint32_t n_shards) {
  ENTER(ctx);
  if (n_queries && (!progs || !prog_off)) return PB200_ERR_INVALID;
  return count_begin(ctx, progs, prog_off, n_queries, shards, n_shards, false);
}

extern "C" int pb200_count_batch_device(pb200_ctx* ctx, void** device_counts) {
  ENTER(ctx);
  if (!device_counts) return PB200_ERR_INVALID;
  *device_counts = count_device_ptr(ctx);
  if (!*device_counts) return fail(ctx, PB200_ERR_INVALID, "no Count batch in flight");
  ctx->slots[(ctx->head + ctx->n_pending - 1) & 1].external = true;
  return PB200_OK;
}

extern "C" int pb200_count_batch_end(pb200_ctx* ctx, uint64_t* out_counts) {
  ENTER(ctx);
  return count_end(ctx, out_counts);
}

// ---- prepared batches: compile + upload once, launch many times, results stay in HBM ----
struct pb200_batch {
  DevBuf d_batch, d_counts;
  Compiled cp;
  BatchDev bd{};
  uint32_t n_queries = 0, n_shards = 0;
  uint64_t generation = 0;  // ctx->generation the batch was last validated against
};

// A prepared batch outlives loads, writes and drops of the fragments it names. Its programs hold only (family, row)
// leaves, so all that can go stale is (a) the device family table the launch passes to the kernels and (b) whether
// some leaf family now holds array / run containers, i.e. whether the staged kernel and its scratch are needed.
static int batch_revalidate(pb200_ctx* ctx, pb200_batch* b) {
  TRY(sync_families(ctx));
  if (b->generation == ctx->generation) return PB200_OK;
  bool staging = false;
  for (const Leaf& lf : b->cp.leaves) {
    if (lf.kind == L_TEMP) staging = true;
    else if (lf.kind == L_ROW) {
      if (lf.family >= ctx->families.size()) return fail(ctx, PB200_ERR_INVALID, "prepared batch names a family that no longer exists");
      if (ctx->families[lf.family].n_not_all_bitmap) staging = true;
    }
  }
  b->cp.may_need_staging = staging;
  b->generation = ctx->generation;
  return PB200_OK;
}

extern "C" int pb200_batch_prepare(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                                   const uint64_t* shards, uint32_t n_shards, pb200_batch** out) {
  ENTER(ctx);
  if (!out || !n_queries || !progs || !prog_off || !n_shards || !shards) return PB200_ERR_INVALID;
  auto* b = new pb200_batch();
  int rc = compile_batch(ctx, progs, prog_off, n_queries, b->cp);
  if (rc == PB200_OK && (b->cp.n_temp_slots || !b->cp.pre.empty())) rc = fail(ctx, PB200_ERR_UNSUPPORTED, "prepared batches do not take BSI range / Shift leaves");
  if (rc == PB200_OK) rc = sync_families(ctx);
  if (rc == PB200_OK) {
    std::swap(ctx->d_batch, b->d_batch);  // upload into the batch's own buffer
    rc = upload_batch(ctx, b->cp, shards, n_shards, b->bd);
    std::swap(ctx->d_batch, b->d_batch);
  }
  if (rc == PB200_OK) rc = ensure_dev(ctx, b->d_counts, size_t(n_queries) * 8);
  if (rc == PB200_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) rc = fail(ctx, PB200_ERR_CUDA, "sync after batch upload failed");
  if (rc != PB200_OK) {
    if (b->d_batch.p) cudaFree(b->d_batch.p);
    if (b->d_counts.p) cudaFree(b->d_counts.p);
    delete b;
    return rc;
  }
  b->n_queries = n_queries; b->n_shards = n_shards;
  b->generation = ctx->generation;
  *out = b;
  return PB200_OK;
}

extern "C" int pb200_batch_launch(pb200_ctx* ctx, pb200_batch* b) {
  ENTER(ctx);
  if (!b) return PB200_ERR_INVALID;
  TRY(batch_revalidate(ctx, b));
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  ctx->timing.launches = 0;
  CUDA_TRY(ctx, cudaMemsetAsync(b->d_counts.p, 0, size_t(b->n_queries) * 8, ctx->stream));
  return launch_eval(ctx, b->bd, b->bd.progs, b->n_queries, b->n_shards, b->cp.may_need_staging,
                     static_cast<unsigned long long*>(b->d_counts.p), 1, 0, nullptr, nullptr, max_leaves_of(b->cp), b->cp.any_bushy, fold_all_of(b->cp), nullptr, b->cp.any_flat, tiny_of(b->cp));
}

extern "C" int pb200_batch_device_counts(pb200_ctx* ctx, pb200_batch* b, void** device_counts) {
  if (!ctx || !b || !device_counts) return PB200_ERR_INVALID;
  *device_counts = b->d_counts.p;
  return PB200_OK;
}

extern "C" int pb200_batch_fetch(pb200_ctx* ctx, pb200_batch* b, uint64_t* out_counts) {
  ENTER(ctx);
  if (!b || !out_counts) return PB200_ERR_INVALID;
  for (int k = 0; k < 4 && ctx->peer.on; k++)  // an all-reduced launch of this batch still in flight: pick its total up first
    if (ctx->peer.pend[k].open && ctx->peer.pend[k].target == b->d_counts.p) TRY(pb200_peer_flush_locked(ctx));
  TRY(ensure_pin(ctx, ctx->h_out, size_t(b->n_queries) * 8));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_out.p, b->d_counts.p, size_t(b->n_queries) * 8, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  std::memcpy(out_counts, ctx->h_out.p, size_t(b->n_queries) * 8);
  return peer_check_error(ctx);
}

extern "C" int pb200_batch_free(pb200_ctx* ctx, pb200_batch* b) {
  ENTER(ctx);
  if (!b) return PB200_OK;
  cudaStreamSynchronize(ctx->stream);
  if (b->d_batch.p) cudaFree(b->d_batch.p);
  if (b->d_counts.p) cudaFree(b->d_counts.p);
  delete b;
  return PB200_OK;
}

// =======================================================================================
// Allocation-free Count(Intersect(Row a, Row b)) for callers that must not allocate.
//
// Who needs this: code that runs inside a process where somebody else owns the device heap
// and a stray cudaMalloc is off limits — a host embedded in a PyTorch program (the caching
// allocator has usually reserved the free HBM, so a raw cudaMalloc can fail even though
// "free" memory exists, and cudaMalloc/cudaFree stall every stream of the device), a
// deployment that handed the whole HBM budget to resident fragments, or a latency-critical
// query thread whose policy forbids anything that can block on the driver's allocator. Such
// callers carve the small input / output buffers out of their own pool and pass them in;
// everything the kernels need beyond that (the compiled two-leaf programs and the list of
// container columns deferred to the staged kernel) lives in the fixed array below.
//
// The array is part of the library image: every GPU that loads libpilosa_b200.so holds
// PB200_NOALLOC_SCRATCH_BYTES of it until the process exits, used or not. All calls on a
// device share it and it keeps whatever the previous call left behind, so the entry point
// re-zeroes the words that must start at zero, and two calls whose kernels overlap on the
// GPU (different contexts / streams on one device) race on it — keep them on one stream.
// =======================================================================================
static __device__ __align__(256) unsigned char g_noalloc_scratch[PB200_NOALLOC_SCRATCH_BYTES];

namespace {

// byte offsets inside g_noalloc_scratch (all 256 B aligned)
constexpr size_t kNaCounters = 0;                                                     // u32 slow_count @0, u64 algo_bytes @8
constexpr size_t kNaProgs = 256;                                                      // Prog[MAX_PAIRS]
constexpr size_t kNaLeaves = kNaProgs + size_t(PB200_NOALLOC_MAX_PAIRS) * sizeof(Prog);      // Leaf[2 * MAX_PAIRS]
constexpr size_t kNaInstrs = kNaLeaves + size_t(PB200_NOALLOC_MAX_PAIRS) * 2 * sizeof(Leaf); // Instr[2 * MAX_PAIRS]
constexpr size_t kNaSlow = (kNaInstrs + size_t(PB200_NOALLOC_MAX_PAIRS) * 2 * sizeof(Instr) + 255) / 256 * 256;  // u32[MAX_ITEMS]
constexpr size_t kNaEnd = kNaSlow + size_t(PB200_NOALLOC_MAX_ITEMS) * 4;
static_assert(kNaEnd <= PB200_NOALLOC_SCRATCH_BYTES, "scratch layout exceeds PB200_NOALLOC_SCRATCH_BYTES");

// Writes the program of pair q — LOAD row a ; AND row b — into the scratch array. Device
// code may name the array directly; only host code needs its address.
__global__ void noalloc_expand_kernel(const uint64_t* pairs, uint32_t n_pairs, uint32_t family) {
  const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n_pairs) return;
  Prog* progs = reinterpret_cast<Prog*>(g_noalloc_scratch + kNaProgs);
  Leaf* leaves = reinterpret_cast<Leaf*>(g_noalloc_scratch + kNaLeaves);
  Instr* instrs = reinterpret_cast<Instr*>(g_noalloc_scratch + kNaInstrs);
  Prog pg; pg.instr_off = 2 * q; pg.n_instr = 2; pg.leaf_off = 2 * q; pg.n_leaves = 2;
  progs[q] = pg;
  Leaf a; a.kind = L_ROW; a.family = family; a.row = pairs[2 * q];
  Leaf b; b.kind = L_ROW; b.family = family; b.row = pairs[2 * q + 1];
  leaves[2 * q] = a; leaves[2 * q + 1] = b;
  instrs[2 * q] = Instr{I_LOAD, 0};
  instrs[2 * q + 1] = Instr{I_AND_LEAF, 1};
}

std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {};  // device address of the array, per device ordinal

// One-time set-up per device, with that device current: in host code the array's name is
// only a stand-in, not a device address, so the address has to be asked for.
int noalloc_setup(pb200_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= 64) return fail(ctx, PB200_ERR_RANGE, "device ordinal out of range for the no-alloc path");
  std::lock_guard<std::mutex> g(g_noalloc_mu);
  if (!g_noalloc_ptr[ctx->device]) {
    void* p = nullptr;
    CUDA_TRY(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
    if (!p) return fail(ctx, PB200_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[ctx->device] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_ptr[ctx->device];
  return PB200_OK;
}

}  // namespace

extern "C" int pb200_count_pairs_noalloc(pb200_ctx* ctx, pb200_family family, const uint64_t* d_pairs, uint32_t n_pairs,
                                         const uint64_t* d_shards, uint32_t n_shards, uint64_t* d_counts) {
  ENTER(ctx);
  if (n_pairs == 0) return PB200_OK;
  if (!d_pairs || !d_counts || (n_shards && !d_shards)) return fail(ctx, PB200_ERR_INVALID, "null device buffer");
  if (family >= ctx->families.size()) return fail(ctx, PB200_ERR_NOTFOUND, "unknown family");
  if (n_pairs > PB200_NOALLOC_MAX_PAIRS)
    return fail(ctx, PB200_ERR_RANGE, "pb200_count_pairs_noalloc: n_pairs " + std::to_string(n_pairs) + " exceeds PB200_NOALLOC_MAX_PAIRS (" +
                                          std::to_string(PB200_NOALLOC_MAX_PAIRS) + "); split the batch");
  const uint64_t n_items = uint64_t(n_pairs) * n_shards * kRowContainers;
  if (n_items > PB200_NOALLOC_MAX_ITEMS)
    return fail(ctx, PB200_ERR_RANGE, "pb200_count_pairs_noalloc: n_pairs x n_shards x 16 = " + std::to_string(n_items) +
                                          " exceeds PB200_NOALLOC_MAX_ITEMS (" + std::to_string(PB200_NOALLOC_MAX_ITEMS) + "); split the batch");
  if (ctx->families_dirty || !ctx->d_families)
    return fail(ctx, PB200_ERR_INVALID, "device fragment tables are stale: run one ordinary query (e.g. pb200_count) after loading fragments "
                                        "before using the no-alloc entry point (refreshing them allocates)");
  unsigned char* scratch = nullptr;
  TRY(noalloc_setup(ctx, &scratch));
  // the array still holds the previous call's counters; the caller's counts are accumulated with atomics
  CUDA_TRY(ctx, cudaMemsetAsync(scratch + kNaCounters, 0, 16, ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(d_counts, 0, size_t(n_pairs) * 8, ctx->stream));
  if (n_shards == 0) return PB200_OK;
  noalloc_expand_kernel<<<(n_pairs + 255) / 256, 256, 0, ctx->stream>>>(d_pairs, n_pairs, family);
  EvalParams p{};
  p.families = ctx->d_families;
  p.n_families = uint32_t(ctx->families.size());
  p.progs = reinterpret_cast<const Prog*>(scratch + kNaProgs);
  p.leaves = reinterpret_cast<const Leaf*>(scratch + kNaLeaves);
  p.instrs = reinterpret_cast<const Instr*>(scratch + kNaInstrs);
  p.shards = d_shards;
  p.n_shards = n_shards; p.n_queries = n_pairs;
  p.out_counts = reinterpret_cast<unsigned long long*>(d_counts);
  p.out_qstride = 1; p.out_sstride = 0;
  p.slow_list = reinterpret_cast<uint32_t*>(scratch + kNaSlow);
  p.slow_count = reinterpret_cast<uint32_t*>(scratch + kNaCounters);
  p.algo_bytes = reinterpret_cast<unsigned long long*>(scratch + kNaCounters + 8);
  const uint32_t grid = uint32_t(std::min<uint64_t>((n_items + kFastWarps - 1) / kFastWarps, uint64_t(ctx->sm_count) * 8));
  eval_fast_kernel<<<grid, kFastWarps * 32, 0, ctx->stream>>>(p);
  const uint32_t sgrid = uint32_t(std::min<uint64_t>(n_items, uint64_t(ctx->sm_count) * PB200_FOLD_CTAS));
  eval_fold_kernel<<<sgrid, kFoldCtaThreads, kFoldSmem, ctx->stream>>>(p);  // a two-leaf Intersect is a flat fold
  CUDA_TRY(ctx, cudaGetLastError());
  return PB200_OK;
}

#include "engine_ingest.inl"
#include "engine_peer.inl"
#include "engine_multi.inl"
