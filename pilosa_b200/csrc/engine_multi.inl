// Included by engine.cu: several GPUs of one box behind ONE handle (pb200_multi).
//
// The reference does its whole fan-out and reduce inside one call: executor.mapReduce / mapper / mapperLocal
// (executor.go:2458-2611) hand per-shard jobs to the workers and fold the partial results with the call's reduce
// function (uint64 + :1813, ValCount.add/smaller/larger :3002-3029, Pairs.Add cache.go:356-374, Row.Merge
// executor.go:595-602). pb200_multi gives a Go host the same shape over the GPUs of a box: one context per device,
// shard s lives on device s % ndev (every family of a shard on the same device), a call takes the
// caller's shard list, runs on the owning devices concurrently and returns the reduced answer.
//
// NVLink is used in two ways, both with plain peer loads / stores (all-to-all peer access is enabled at init):
//  * every device's fragment table lists ALL shards; the entries of remote shards point into the owner's HBM. A
//    call whose operands are small (a sparse Count over a few hundred shards) is therefore run by ONE device over
//    the peers' memory — one launch, no fan-out latency — while anything bandwidth-bound goes to the owners;
//  * bitmap-returning calls are evaluated by the owners and their dense result rows are pulled into the first
//    device by a gather kernel reading peer memory, so that typing, serialisation (K11) and protobuf emission (K12)
//    run once over the whole shard list and ONE device -> host copy carries the answer.

namespace {

struct FamStat {  // what the router knows about a family: bytes and rows resident over all devices
  std::unordered_map<uint64_t, std::pair<uint64_t, uint64_t>> per_shard;  // shard -> (rows, payload bytes)
  uint64_t rows = 0, bytes = 0;
};

}  // namespace

struct pb200_multi {
  std::vector<pb200_ctx*> ctx;
  std::vector<std::unique_ptr<DevWorker>> workers;  // workers[d] issues for ctx[d], d >= 1; the calling thread issues for ctx[0]
  std::mutex mu;                                    // one multi call at a time
  std::string err;
  bool peer = false;                                // all-to-all peer access is on
  uint32_t rr = 0;                                  // next device for a single-device call
  uint64_t small_call_bytes = 24ull << 20;          // estimated operand bytes below which one device runs the whole call ...
  uint32_t small_call_shards = 1024;                // ... provided queries x shards is no more than this
  std::vector<FamStat> fam;
  pb200_timing timing{};
  DevBuf g_dense, g_dense_n, g_sparse, g_ovf, g_shards, g_pos;  // on ctx[0]'s device: gathered result rows of a bitmap-returning call
  std::vector<cudaEvent_t> ev;                      // per device: "dense rows ready"
};

namespace {

int mfail(pb200_multi* m, int code, const std::string& msg) {
  m->err = msg;
  return code;
}

struct ShardSplit {
  std::vector<std::vector<uint64_t>> shards;  // per device, in the caller's order
  std::vector<std::vector<uint32_t>> pos;     // position of each in the caller's list
};

ShardSplit split_shards(const pb200_multi* m, const uint64_t* shards, uint32_t n, bool want_pos = true) {
  const size_t nd = m->ctx.size();
  ShardSplit sp;
  sp.shards.resize(nd); sp.pos.resize(nd);
  for (size_t d = 0; d < nd; d++) {  // (a call over thousands of shards should not pay for vector growth: this is on every query's path)
    sp.shards[d].reserve(n / nd + 8);
    if (want_pos) sp.pos[d].reserve(n / nd + 8);
  }
  for (uint32_t i = 0; i < n; i++) {
    const size_t d = size_t(shards[i] % nd);
    sp.shards[d].push_back(shards[i]);
    if (want_pos) sp.pos[d].push_back(i);
  }
  return sp;
}

// Runs fn(d) for every device with work, concurrently: device 0 on the calling thread, the others on their workers.
template <class F>
int fan_out(pb200_multi* m, const std::vector<uint8_t>& active, F&& fn) {
  const size_t nd = m->ctx.size();
  std::vector<int> rc(nd, PB200_OK);
  for (size_t d = 1; d < nd; d++)
    if (active[d]) m->workers[d]->post([&fn, d] { return fn(d); });
  if (active[0]) rc[0] = fn(0);
  for (size_t d = 1; d < nd; d++)
    if (active[d]) rc[d] = m->workers[d]->wait();
  for (size_t d = 0; d < nd; d++)
    if (rc[d] != PB200_OK) {
      m->err = "device " + std::to_string(m->ctx[d]->device) + ": " + m->ctx[d]->err;
      return rc[d];
    }
  return PB200_OK;
}

std::vector<uint8_t> active_of(const ShardSplit& sp) {
  std::vector<uint8_t> a(sp.shards.size());
  for (size_t d = 0; d < a.size(); d++) a[d] = !sp.shards[d].empty();
  return a;
}

void timing_merge(pb200_multi* m, const std::vector<uint8_t>& active) {
  pb200_timing t{};
  for (size_t d = 0; d < m->ctx.size(); d++) {
    if (!active[d]) continue;
    const pb200_timing& s = m->ctx[d]->timing;
    t.kernel_ms = std::max(t.kernel_ms, s.kernel_ms);  // the devices run side by side
    t.algo_bytes += s.algo_bytes; t.h2d_bytes += s.h2d_bytes; t.d2h_bytes += s.d2h_bytes; t.launches += s.launches;
  }
  m->timing = t;
}

// Publishes one fragment-table entry of device `owner` into every other device's table (called by the owner's
// upload / drop, before the previous copy is freed): queries any device may still be running over the old copy are
// waited for, then the entry is replaced.
int publish_entry(pb200_multi* m, size_t owner, uint32_t family, uint64_t shard, const FragDev& dev, bool all_bitmap, bool present) {
  for (size_t j = 0; j < m->ctx.size(); j++) {
    if (j == owner) continue;
    pb200_ctx* c = m->ctx[j];
    std::lock_guard<std::mutex> lock(c->mu);
    DeviceGuard g(c->device);
    if (family >= c->families.size()) return fail(c, PB200_ERR_NOTFOUND, "family missing on a peer device");
    Family& fam = c->families[family];
    TRY(family_grow(c, fam, shard));
    CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    if (fam.remote_flags.size() <= shard) fam.remote_flags.resize(fam.cap, 0);
    if (fam.remote_flags[shard] & 2) fam.n_not_all_bitmap--;
    fam.remote_flags[shard] = present ? uint8_t(1 | (all_bitmap ? 0 : 2)) : 0;
    if (fam.remote_flags[shard] & 2) fam.n_not_all_bitmap++;
    fam.h_frags[shard] = present ? dev : FragDev{};
    CUDA_TRY(c, cudaMemcpy(fam.d_frags + shard, &fam.h_frags[shard], sizeof(FragDev), cudaMemcpyHostToDevice));
    c->generation++;
  }
  return PB200_OK;
}

void stat_update(pb200_multi* m, pb200_family family, uint64_t shard) {
  pb200_ctx* c = m->ctx[shard % m->ctx.size()];
  FamStat& fs = m->fam[family];
  auto old = fs.per_shard.find(shard);
  if (old != fs.per_shard.end()) { fs.rows -= old->second.first; fs.bytes -= old->second.second; fs.per_shard.erase(old); }
  auto it = c->families[family].frags.find(shard);
  if (it == c->families[family].frags.end()) return;
  const uint64_t rows = it->second.row_ids.size(), bytes = it->second.payload_bytes + it->second.n_containers * 16;
  fs.per_shard[shard] = {rows, bytes};
  fs.rows += rows; fs.bytes += bytes;
}

// Operand bytes a batch of programs will read, from the per-family averages (bytes per row per fragment).
uint64_t estimate_bytes(const pb200_multi* m, const pb200_node* progs, uint32_t n_nodes, uint32_t n_shards) {
  double per_shard = 0;
  for (uint32_t k = 0; k < n_nodes; k++) {
    const pb200_node& nd = progs[k];
    if (nd.family >= m->fam.size()) continue;
    const FamStat& fs = m->fam[nd.family];
    const double row_bytes = fs.rows ? double(fs.bytes) / double(fs.rows) : 0;
    if (nd.op == PB200_OP_ROW || nd.op == PB200_OP_NOT) per_shard += row_bytes;
    else if (nd.op == PB200_OP_BSI_RANGE) per_shard += row_bytes * (nd.bit_depth + 2);
  }
  return uint64_t(per_shard * n_shards);
}

// Where a Count call runs: -1 = on the owners of its shards, else the one device that runs all of it over peer memory.
int route_count(pb200_multi* m, const pb200_node* progs, uint32_t n_nodes, uint32_t n_queries, uint32_t n_shards) {
  if (m->ctx.size() == 1) return 0;
  if (!m->peer) return -1;
  // Remote reads pay NVLink latency per dependent load (shard -> fragment -> row table -> payload), so what one device
  // can cover over its peers is bounded by the NUMBER of (query, shard) pairs as well as by the bytes: measured, a
  // sparse two-row Count over 512 shards costs the same either way, over 8192 shards the owners are twice as fast, and
  // a batch of 256 such queries over 64 shards is 2x slower on one device than on the owners.
  if (uint64_t(n_queries) * n_shards > m->small_call_shards || estimate_bytes(m, progs, n_nodes, n_shards) >= m->small_call_bytes) return -1;
  return int(m->rr++ % m->ctx.size());
}

#define MENTER(m)                                  \
  if (!(m)) return PB200_ERR_INVALID;              \
  std::lock_guard<std::mutex> mlock__((m)->mu);    \
  (m)->err.clear()

#define MOWNER(m, shard) ((m)->ctx[size_t((shard) % (m)->ctx.size())])
#define MFORWARD(m, c, call)                 \
  do {                                       \
    const int rc__ = (call);                 \
    if (rc__ != PB200_OK) (m)->err = (c)->err; \
    return rc__;                             \
  } while (0)

int mcheck_family(pb200_multi* m, pb200_family f) {
  return f < m->fam.size() ? PB200_OK : mfail(m, PB200_ERR_NOTFOUND, "unknown family");
}

int count_batch_multi(pb200_multi* m, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries, const uint64_t* shards,
                      uint32_t n_shards, uint64_t* out) {
  if (n_queries && (!progs || !prog_off || !out)) return PB200_ERR_INVALID;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  const int one = route_count(m, progs, n_queries ? prog_off[n_queries] : 0, n_queries, n_shards);
  if (one >= 0) {
    const int rc = pb200_count_batch(m->ctx[one], progs, prog_off, n_queries, shards, n_shards, out);
    if (rc != PB200_OK) m->err = m->ctx[one]->err;
    m->timing = m->ctx[one]->timing;
    return rc;
  }
  const ShardSplit sp = split_shards(m, shards, n_shards, false);
  const std::vector<uint8_t> active = active_of(sp);
  std::vector<std::vector<uint64_t>> part(m->ctx.size());
  TRY(fan_out(m, active, [&](size_t d) {
    part[d].assign(n_queries, 0);
    return pb200_count_batch(m->ctx[d], progs, prog_off, n_queries, sp.shards[d].data(), uint32_t(sp.shards[d].size()), part[d].data());
  }));
  for (uint32_t q = 0; q < n_queries; q++) out[q] = 0;
  bool any = false;
  for (size_t d = 0; d < part.size(); d++) {
    if (!active[d]) continue;
    any = true;
    for (uint32_t q = 0; q < n_queries; q++) out[q] += part[d][q];  // reduceFn of executeCount: uint64 + (executor.go:1813)
  }
  if (!any && n_queries) {  // no shard at all: the programs are still validated, as on one device
    const int rc = pb200_count_batch(m->ctx[0], progs, prog_off, n_queries, shards, 0, out);
    if (rc != PB200_OK) { m->err = m->ctx[0]->err; return rc; }
  }
  timing_merge(m, active);
  return PB200_OK;
}

// Copy of the result rows of an owner (peer memory, plain loads) into the gathered buffers of device 0, a warp per
// container: 8 KiB of words, or — when the owner kept a short container as a value list (src_sparse given) — its 64
// bytes; with lists on, empty containers move nothing.
__global__ void gather_rows_kernel(const uint4* __restrict__ src, const uint32_t* __restrict__ src_n, const uint4* __restrict__ src_sparse,
                                   const unsigned long long* __restrict__ src_ovf, const uint32_t* __restrict__ pos, uint32_t n_local,
                                   uint4* __restrict__ dst, uint32_t* __restrict__ dst_n, uint4* __restrict__ dst_sparse,
                                   unsigned long long* __restrict__ dst_ovf) {
  constexpr uint32_t kVec = kBitmapBytes / 16;              // uint4 per container
  constexpr uint32_t kListVec = kSparseListMax * 2 / 16;    // uint4 per value list
  const uint32_t lane = threadIdx.x & 31, warps = gridDim.x * (blockDim.x >> 5);
  const uint32_t n_items = n_local * kRowContainers;
  for (uint32_t item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); item < n_items; item += warps) {
    const uint32_t dn = src_n[item];
    const uint32_t to = pos[item / kRowContainers] * kRowContainers + item % kRowContainers;
    if (src_sparse && (dn & kSparseFlag)) {
      if (lane < kListVec) dst_sparse[size_t(to) * kListVec + lane] = src_sparse[size_t(item) * kListVec + lane];
    } else if (dn || !src_sparse) {
      const uint4* f = src + size_t(item) * kVec;
      uint4* t = dst + size_t(to) * kVec;
#pragma unroll 4
      for (uint32_t k = lane; k < kVec; k += 32) t[k] = f[k];
    }
    if (lane == 0) dst_n[to] = dn;
  }
  if (dst_ovf)
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_local; i += gridDim.x * blockDim.x) dst_ovf[pos[i]] = src_ovf ? src_ovf[i] : 0ull;
}

// Owners evaluate the call into dense rows; device 0 pulls them together in the caller's shard order and is left
// with ctx[0]->d_dense / d_dense_n / overflow words describing ALL n_shards rows, shard numbers at *shards_dev.
// On return ctx[0]'s mutex is NOT held; the caller re-enters ctx[0] for the emission.
int gather_dense(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, const uint64_t** shards_dev,
                 bool allow_lists = false) {
  if (!m->peer && m->ctx.size() > 1) return mfail(m, PB200_ERR_UNSUPPORTED, "bitmap-returning multi-GPU calls need peer access between the devices");
  const ShardSplit sp = split_shards(m, shards, n_shards);
  const std::vector<uint8_t> active = active_of(sp);
  pb200_ctx* c0 = m->ctx[0];
  // every owner: evaluate its shards (asynchronously), leave an event behind
  std::vector<uint8_t> has_ovf(m->ctx.size(), 0);
  TRY(fan_out(m, active, [&](size_t d) {
    pb200_ctx* c = m->ctx[d];
    std::lock_guard<std::mutex> lock(c->mu);
    DeviceGuard g(c->device);
    c->err.clear();
    TRY(ensure_dev(c, c->d_misc, 64));
    timing_begin(c);
    CUDA_TRY(c, cudaEventRecord(c->ev0, c->stream));
    BatchDev bd{};
    TRY(eval_dense(c, prog, n_nodes, sp.shards[d].data(), uint32_t(sp.shards[d].size()), bd, allow_lists));
    CUDA_TRY(c, cudaEventRecord(c->ev1, c->stream));
    has_ovf[d] = c->dense_has_ovf ? 1 : 0;
    CUDA_TRY(c, cudaEventRecord(m->ev[d], c->stream));
    return int(PB200_OK);
  }));
  bool any_ovf = false;
  for (uint8_t v : has_ovf) any_ovf |= v != 0;
  // device 0: room for all rows, the shard list and the positions of every owner's rows
  std::lock_guard<std::mutex> lock(c0->mu);
  DeviceGuard g(c0->device);
  c0->err.clear();
  auto fail0 = [&](int rc) { m->err = c0->err; return rc; };
  const size_t items = size_t(n_shards) * kRowContainers;
  auto ensure0 = [&](DevBuf& b, size_t bytes) {
    std::swap(c0->d_out, b);
    const int rc = ensure_dev(c0, c0->d_out, bytes);
    std::swap(c0->d_out, b);
    return rc;
  };
  if (!active[0]) {  // device 0 owns none of the shards: its timing window opens here
    if (ensure_dev(c0, c0->d_misc, 64) != PB200_OK) return fail0(PB200_ERR_CUDA);
    timing_begin(c0);
    if (cudaEventRecord(c0->ev0, c0->stream) != cudaSuccess) return fail0(fail(c0, PB200_ERR_CUDA, "cudaEventRecord failed"));
  }
  int rc = ensure0(m->g_dense, items * kBitmapBytes);
  if (rc == PB200_OK) rc = ensure0(m->g_dense_n, items * 4);
  const bool lists = allow_lists && !any_ovf;  // (a program property: every owner decided the same)
  if (rc == PB200_OK && lists) rc = ensure0(m->g_sparse, items * kSparseListMax * 2);
  if (rc == PB200_OK) rc = ensure0(m->g_ovf, size_t(n_shards) * 8);
  if (rc == PB200_OK) rc = ensure0(m->g_shards, size_t(n_shards) * 8);
  if (rc == PB200_OK) rc = ensure0(m->g_pos, size_t(n_shards) * 4);
  if (rc != PB200_OK) return fail0(rc);
  if (cudaMemcpyAsync(m->g_shards.p, shards, size_t(n_shards) * 8, cudaMemcpyHostToDevice, c0->stream) != cudaSuccess)
    return fail0(fail(c0, PB200_ERR_CUDA, "upload of the shard list failed"));
  std::vector<uint32_t> pos_all;
  std::vector<size_t> pos_off(m->ctx.size(), 0);
  for (size_t d = 0; d < m->ctx.size(); d++) { pos_off[d] = pos_all.size(); pos_all.insert(pos_all.end(), sp.pos[d].begin(), sp.pos[d].end()); }
  if (cudaMemcpyAsync(m->g_pos.p, pos_all.data(), pos_all.size() * 4, cudaMemcpyHostToDevice, c0->stream) != cudaSuccess)
    return fail0(fail(c0, PB200_ERR_CUDA, "upload of the row positions failed"));
  if (cudaStreamSynchronize(c0->stream) != cudaSuccess) return fail0(fail(c0, PB200_ERR_CUDA, "sync failed"));  // pos_all is pageable
  for (size_t d = 0; d < m->ctx.size(); d++) {
    if (!active[d]) continue;
    pb200_ctx* c = m->ctx[d];
    const uint32_t nl = uint32_t(sp.shards[d].size());
    if (cudaStreamWaitEvent(c0->stream, m->ev[d], 0) != cudaSuccess) return fail0(fail(c0, PB200_ERR_CUDA, "cudaStreamWaitEvent failed"));
    const uint64_t warps = uint64_t(nl) * kRowContainers;
    const uint32_t grid = uint32_t(std::min<uint64_t>((warps + 7) / 8, uint64_t(c0->sm_count) * 8));
    gather_rows_kernel<<<grid, 256, 0, c0->stream>>>(static_cast<const uint4*>(c->d_dense.p), static_cast<const uint32_t*>(c->d_dense_n.p),
                                                     lists ? static_cast<const uint4*>(c->d_sparse.p) : nullptr,
                                                     has_ovf[d] ? c->dense_ovf_dev : nullptr, static_cast<const uint32_t*>(m->g_pos.p) + pos_off[d], nl,
                                                     static_cast<uint4*>(m->g_dense.p), static_cast<uint32_t*>(m->g_dense_n.p),
                                                     lists ? static_cast<uint4*>(m->g_sparse.p) : nullptr,
                                                     any_ovf ? static_cast<unsigned long long*>(m->g_ovf.p) : nullptr);
  }
  if (cudaGetLastError() != cudaSuccess) return fail0(fail(c0, PB200_ERR_CUDA, "gather launch failed"));
  *shards_dev = static_cast<const uint64_t*>(m->g_shards.p);
  c0->dense_has_ovf = any_ovf;
  c0->dense_lists = lists;
  c0->dense_ovf_dev = static_cast<unsigned long long*>(m->g_ovf.p);
  timing_merge(m, active);
  return PB200_OK;
}

// The gathered rows take the place of ctx[0]'s own dense buffers for the duration of the emission.
struct GatherSwap {
  pb200_multi* m;
  explicit GatherSwap(pb200_multi* mm) : m(mm) { swap(); }
  ~GatherSwap() { swap(); }
  void swap() { std::swap(m->ctx[0]->d_dense, m->g_dense); std::swap(m->ctx[0]->d_dense_n, m->g_dense_n); std::swap(m->ctx[0]->d_sparse, m->g_sparse); }
};

}  // namespace

// ---------------------------------------------------------------------------------------
// lifecycle
// ---------------------------------------------------------------------------------------
extern "C" int pb200_init_multi(const int* devices, int ndev, size_t hbm_budget_bytes_per_device, pb200_multi** out) {
  if (!out || !devices || ndev < 1 || ndev > 64) return PB200_ERR_INVALID;
  *out = nullptr;
  for (int i = 0; i < ndev; i++)
    for (int j = 0; j < i; j++)
      if (devices[i] == devices[j]) return PB200_ERR_INVALID;
  auto* m = new pb200_multi();
  for (int i = 0; i < ndev; i++) {
    pb200_ctx* c = nullptr;
    const int rc = pb200_init(devices[i], hbm_budget_bytes_per_device, &c);
    if (rc != PB200_OK) {
      for (pb200_ctx* p : m->ctx) pb200_shutdown(p);
      delete m;
      return rc;
    }
    m->ctx.push_back(c);
  }
  // all-to-all peer access: the single-device path and the result gather read the owners' HBM over NVLink
  m->peer = ndev > 1;
  for (int i = 0; i < ndev && m->peer; i++)
    for (int j = 0; j < ndev && m->peer; j++) {
      if (i == j) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) != cudaSuccess || !can) m->peer = false;
    }
  for (int i = 0; i < ndev && m->peer; i++) {
    DeviceGuard g(devices[i]);
    for (int j = 0; j < ndev; j++) {
      if (i == j) continue;
      const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (e != cudaSuccess) { m->peer = false; cudaGetLastError(); }
    }
  }
  m->ev.resize(size_t(ndev), nullptr);
  for (int i = 0; i < ndev; i++) {
    DeviceGuard g(devices[i]);
    if (cudaEventCreateWithFlags(&m->ev[size_t(i)], cudaEventDisableTiming) != cudaSuccess) {
      for (pb200_ctx* p : m->ctx) pb200_shutdown(p);
      delete m;
      return PB200_ERR_CUDA;
    }
  }
  m->workers.resize(size_t(ndev));
  for (int i = 1; i < ndev; i++) m->workers[size_t(i)].reset(new DevWorker());
  for (size_t i = 0; i < m->ctx.size(); i++) {
    pb200_ctx* c = m->ctx[i];
    c->publish = [m, i](uint32_t family, uint64_t shard, const FragDev& dev, bool all_bitmap, bool present) {
      return m->ctx.size() > 1 && m->peer ? publish_entry(m, i, family, shard, dev, all_bitmap, present) : int(PB200_OK);
    };
  }
  *out = m;
  return PB200_OK;
}

extern "C" int pb200_multi_shutdown(pb200_multi* m) {
  if (!m) return PB200_ERR_INVALID;
  {
    std::lock_guard<std::mutex> lock(m->mu);
    m->workers.clear();
    for (pb200_ctx* c : m->ctx) { DeviceGuard g(c->device); cudaStreamSynchronize(c->stream); }
    {
      DeviceGuard g(m->ctx[0]->device);
      for (DevBuf* b : {&m->g_dense, &m->g_dense_n, &m->g_sparse, &m->g_ovf, &m->g_shards, &m->g_pos})
        if (b->p) cudaFree(b->p);
    }
    for (size_t i = 0; i < m->ev.size(); i++)
      if (m->ev[i]) { DeviceGuard g(m->ctx[i]->device); cudaEventDestroy(m->ev[i]); }
    for (pb200_ctx* c : m->ctx) { c->publish = nullptr; pb200_shutdown(c); }
  }
  delete m;
  return PB200_OK;
}

extern "C" const char* pb200_multi_last_error(pb200_multi* m) { return m ? m->err.c_str() : "null handle"; }
extern "C" int pb200_multi_ndev(pb200_multi* m) { return m ? int(m->ctx.size()) : 0; }
extern "C" int pb200_multi_peer_access(pb200_multi* m) { return m && m->peer ? 1 : 0; }

extern "C" int pb200_multi_ctx(pb200_multi* m, int i, pb200_ctx** out) {
  if (!m || !out || i < 0 || size_t(i) >= m->ctx.size()) return PB200_ERR_INVALID;
  *out = m->ctx[size_t(i)];
  return PB200_OK;
}

extern "C" int pb200_multi_last_timing(pb200_multi* m, pb200_timing* out) {
  if (!m || !out) return PB200_ERR_INVALID;
  std::lock_guard<std::mutex> lock(m->mu);
  *out = m->timing;
  return PB200_OK;
}

extern "C" int pb200_multi_host_register(pb200_multi* m, void* ptr, size_t len) {
  MENTER(m);
  MFORWARD(m, m->ctx[0], pb200_host_register(m->ctx[0], ptr, len));
}
extern "C" int pb200_multi_host_unregister(pb200_multi* m, void* ptr) {
  MENTER(m);
  MFORWARD(m, m->ctx[0], pb200_host_unregister(m->ctx[0], ptr));
}

extern "C" int pb200_multi_set_option(pb200_multi* m, uint32_t option, uint64_t value) {
  MENTER(m);
  if (option == PB200_OPT_SMALL_CALL_BYTES) { m->small_call_bytes = value; return PB200_OK; }
  for (pb200_ctx* c : m->ctx) {
    const int rc = pb200_set_option(c, option, value);
    if (rc != PB200_OK) { m->err = c->err; return rc; }
  }
  return PB200_OK;
}

// ---------------------------------------------------------------------------------------
// families and fragments: a family handle is the same number on every device; a fragment lives on shard % ndev
// ---------------------------------------------------------------------------------------
extern "C" int pb200_multi_family_create(pb200_multi* m, pb200_family* out) {
  MENTER(m);
  if (!out) return PB200_ERR_INVALID;
  pb200_family f0 = 0;
  for (size_t d = 0; d < m->ctx.size(); d++) {
    pb200_family f = 0;
    const int rc = pb200_family_create(m->ctx[d], &f);
    if (rc != PB200_OK) { m->err = m->ctx[d]->err; return rc; }
    if (d == 0) f0 = f;
    else if (f != f0) return mfail(m, PB200_ERR_INVALID, "family tables of the devices diverged (families must be created through the multi handle)");
  }
  m->fam.resize(size_t(f0) + 1);
  *out = f0;
  return PB200_OK;
}

extern "C" int pb200_multi_frag_load(pb200_multi* m, pb200_family family, uint64_t shard, const void* roaring_bytes, size_t len) {
  MENTER(m);
  TRY(mcheck_family(m, family));
  pb200_ctx* c = MOWNER(m, shard);
  const int rc = pb200_frag_load(c, family, shard, roaring_bytes, len);
  if (rc != PB200_OK) { m->err = c->err; return rc; }
  stat_update(m, family, shard);
  return PB200_OK;
}

extern "C" int pb200_multi_frag_apply(pb200_multi* m, pb200_family family, uint64_t shard, const uint64_t* set_pos, size_t n_set,
                                      const uint64_t* clear_pos, size_t n_clear) {
  MENTER(m);
  TRY(mcheck_family(m, family));
  pb200_ctx* c = MOWNER(m, shard);
  const int rc = pb200_frag_apply(c, family, shard, set_pos, n_set, clear_pos, n_clear);
  if (rc != PB200_OK) { m->err = c->err; return rc; }
  stat_update(m, family, shard);
  return PB200_OK;
}

extern "C" int pb200_multi_frag_import_roaring(pb200_multi* m, pb200_family family, uint64_t shard, const void* data, size_t len, int clear,
                                               int64_t* changed) {
  MENTER(m);
  TRY(mcheck_family(m, family));
  pb200_ctx* c = MOWNER(m, shard);
  const int rc = pb200_frag_import_roaring(c, family, shard, data, len, clear, changed);
  if (rc != PB200_OK) { m->err = c->err; return rc; }
  stat_update(m, family, shard);
  return PB200_OK;
}

extern "C" int pb200_multi_frag_drop(pb200_multi* m, pb200_family family, uint64_t shard) {
  MENTER(m);
  TRY(mcheck_family(m, family));
  pb200_ctx* c = MOWNER(m, shard);
  const int rc = pb200_frag_drop(c, family, shard);
  if (rc != PB200_OK) { m->err = c->err; return rc; }
  stat_update(m, family, shard);
  return PB200_OK;
}

extern "C" int pb200_multi_rank_set(pb200_multi* m, pb200_family family, uint64_t shard, const uint64_t* row_ids, const uint64_t* counts, size_t n) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_rank_set(c, family, shard, row_ids, counts, n));
}

extern "C" int pb200_multi_rank_get(pb200_multi* m, pb200_family family, uint64_t shard, uint64_t* row_ids, uint64_t* counts, size_t cap, size_t* n) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_rank_get(c, family, shard, row_ids, counts, cap, n));
}

extern "C" int pb200_multi_frag_stat(pb200_multi* m, pb200_family family, uint64_t shard, uint64_t* n_rows, uint64_t* n_containers, uint64_t* payload_bytes) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_frag_stat(c, family, shard, n_rows, n_containers, payload_bytes));
}

// ---------------------------------------------------------------------------------------
// queries
// ---------------------------------------------------------------------------------------
extern "C" int pb200_multi_count(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, uint64_t* out_count) {
  MENTER(m);
  if (!out_count) return PB200_ERR_INVALID;
  const uint32_t off[2] = {0, n_nodes};
  return count_batch_multi(m, prog, off, 1, shards, n_shards, out_count);
}

extern "C" int pb200_multi_count_batch(pb200_multi* m, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                                       const uint64_t* shards, uint32_t n_shards, uint64_t* out_counts) {
  MENTER(m);
  return count_batch_multi(m, progs, prog_off, n_queries, shards, n_shards, out_counts);
}

extern "C" int pb200_multi_count_shards(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards,
                                        uint64_t* out_per_shard) {
  MENTER(m);
  if (n_shards && (!shards || !out_per_shard)) return PB200_ERR_INVALID;
  const ShardSplit sp = split_shards(m, shards, n_shards);
  const std::vector<uint8_t> active = active_of(sp);
  std::vector<std::vector<uint64_t>> part(m->ctx.size());
  TRY(fan_out(m, active, [&](size_t d) {
    part[d].assign(sp.shards[d].size(), 0);
    return pb200_count_shards(m->ctx[d], prog, n_nodes, sp.shards[d].data(), uint32_t(sp.shards[d].size()), part[d].data());
  }));
  for (size_t d = 0; d < part.size(); d++)
    for (size_t i = 0; i < part[d].size(); i++) out_per_shard[sp.pos[d][i]] = part[d][i];
  timing_merge(m, active);
  return PB200_OK;
}

extern "C" int pb200_multi_bsi_sum(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                                   uint32_t flags, const uint64_t* shards, uint32_t n_shards, int64_t* out_sum, uint64_t* out_count) {
  MENTER(m);
  if (!out_sum || !out_count) return PB200_ERR_INVALID;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  *out_sum = 0; *out_count = 0;
  const ShardSplit sp = split_shards(m, shards, n_shards);
  const std::vector<uint8_t> active = active_of(sp);
  std::vector<int64_t> sums(m->ctx.size(), 0);
  std::vector<uint64_t> counts(m->ctx.size(), 0);
  TRY(fan_out(m, active, [&](size_t d) {
    return pb200_bsi_sum(m->ctx[d], bsi_family, bit_depth, filter_prog, n_filter_nodes, flags, sp.shards[d].data(), uint32_t(sp.shards[d].size()),
                         &sums[d], &counts[d]);
  }));
  uint64_t s = 0, c = 0;  // ValCount.add (executor.go:3002-3008), wrapping like the reference's int64 +
  for (size_t d = 0; d < sums.size(); d++) { s += uint64_t(sums[d]); c += counts[d]; }
  *out_count = c;
  *out_sum = c ? int64_t(s) : 0;
  timing_merge(m, active);
  return PB200_OK;
}

extern "C" int pb200_multi_bsi_minmax_shards(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog,
                                             uint32_t n_filter_nodes, int is_max, const uint64_t* shards, uint32_t n_shards, int64_t* out_vals,
                                             uint64_t* out_counts) {
  MENTER(m);
  if (n_shards && (!shards || !out_vals || !out_counts)) return PB200_ERR_INVALID;
  const ShardSplit sp = split_shards(m, shards, n_shards);
  const std::vector<uint8_t> active = active_of(sp);
  std::vector<std::vector<int64_t>> v(m->ctx.size());
  std::vector<std::vector<uint64_t>> c(m->ctx.size());
  TRY(fan_out(m, active, [&](size_t d) {
    v[d].assign(sp.shards[d].size(), 0); c[d].assign(sp.shards[d].size(), 0);
    return pb200_bsi_minmax_shards(m->ctx[d], bsi_family, bit_depth, filter_prog, n_filter_nodes, is_max, sp.shards[d].data(),
                                   uint32_t(sp.shards[d].size()), v[d].data(), c[d].data());
  }));
  for (size_t d = 0; d < v.size(); d++)
    for (size_t i = 0; i < v[d].size(); i++) { out_vals[sp.pos[d][i]] = v[d][i]; out_counts[sp.pos[d][i]] = c[d][i]; }
  timing_merge(m, active);
  return PB200_OK;
}

namespace {
int minmax_multi(pb200_multi* m, pb200_family fam, uint32_t bit_depth, const pb200_node* filt, uint32_t nf, int is_max, const uint64_t* shards,
                 uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  if (!out_val || !out_count) return PB200_ERR_INVALID;
  std::vector<int64_t> vals(n_shards, 0);
  std::vector<uint64_t> counts(n_shards, 0);
  TRY(pb200_multi_bsi_minmax_shards(m, fam, bit_depth, filt, nf, is_max, shards, n_shards, vals.data(), counts.data()));
  // ValCount.smaller / larger (executor.go:3010-3029), folded in the caller's shard-list order
  int64_t v = 0; uint64_t c = 0;
  for (uint32_t s = 0; s < n_shards; s++) {
    const bool take = c == 0 || ((is_max ? vals[s] > v : vals[s] < v) && counts[s] > 0);
    if (take) { v = vals[s]; c = counts[s]; }
  }
  if (c == 0) v = 0;
  *out_val = v; *out_count = c;
  return PB200_OK;
}
}  // namespace

extern "C" int pb200_multi_bsi_min(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                                   const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  return minmax_multi(m, bsi_family, bit_depth, filter_prog, n_filter_nodes, 0, shards, n_shards, out_val, out_count);
}
extern "C" int pb200_multi_bsi_max(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog, uint32_t n_filter_nodes,
                                   const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count) {
  return minmax_multi(m, bsi_family, bit_depth, filter_prog, n_filter_nodes, 1, shards, n_shards, out_val, out_count);
}

// executeTopN (executor.go:863-903) over the devices: pass 1 on every owner (fragment.top per shard), Pairs.Add over
// all of them, then pass 2 for the merged candidate ids — from the counts pass 1 left on each owner's host side
// where possible — and the final sort + trim.
extern "C" int pb200_multi_topn(pb200_multi* m, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes, uint32_t n, const uint64_t* ids,
                                size_t n_ids, uint64_t min_threshold, uint64_t tanimoto_threshold, uint32_t flags, const uint64_t* shards,
                                uint32_t n_shards, pb200_pair* out, size_t cap, size_t* n_out) {
  MENTER(m);
  if (n_ids && !ids) return PB200_ERR_INVALID;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  TRY(mcheck_family(m, family));
  const ShardSplit sp = split_shards(m, shards, n_shards);
  std::vector<uint8_t> active = active_of(sp);
  if (n_shards == 0) active[0] = 1;  // argument checks still run
  const size_t nd = m->ctx.size();
  std::vector<std::vector<std::vector<HPair>>> per(nd);
  std::vector<TopnCarry> carry(nd);
  std::vector<pb200_timing> t1(nd);
  auto finish = [&](const std::vector<HPair>& v) {
    for (size_t d = 0; d < nd; d++) m->ctx[d]->dense_shards_dev = nullptr;
    timing_merge(m, active);
    if (!n_out) return int(PB200_ERR_INVALID);
    *n_out = v.size();
    if (v.size() > cap) return mfail(m, PB200_ERR_RANGE, "pair buffer too small");
    for (size_t i = 0; i < v.size(); i++) { out[i].id = v[i].id; out[i].count = v[i].count; }
    return int(PB200_OK);
  };
  TRY(fan_out(m, active, [&](size_t d) {
    pb200_ctx* c = m->ctx[d];
    std::lock_guard<std::mutex> lock(c->mu);
    DeviceGuard g(c->device);
    c->err.clear();
    return topn_pass(c, family, src_prog, n_src_nodes, n, ids, n_ids, min_threshold, tanimoto_threshold, sp.shards[d].data(),
                     uint32_t(sp.shards[d].size()), per[d], false, &carry[d]);
  }));
  std::vector<std::vector<HPair>> all;
  for (size_t d = 0; d < nd; d++) for (auto& v : per[d]) all.push_back(std::move(v));
  std::vector<HPair> pairs = merge_pairs(all);
  if (pairs.empty() || n_ids > 0 || (flags & PB200_TOPN_REMOTE)) return finish(pairs);  // executor.go:884-887
  std::vector<uint64_t> ids2;
  for (auto& p : pairs) ids2.push_back(p.id);
  std::sort(ids2.begin(), ids2.end());
  const uint64_t thr = min_threshold ? min_threshold : 1;
  TRY(fan_out(m, active, [&](size_t d) {
    pb200_ctx* c = m->ctx[d];
    std::lock_guard<std::mutex> lock(c->mu);
    DeviceGuard g(c->device);
    c->err.clear();
    const uint64_t* sh = sp.shards[d].data();
    const uint32_t ns = uint32_t(sp.shards[d].size());
    if (carry[d].valid && topn_pass2_from_carry(c, family, ids2.data(), ids2.size(), thr, tanimoto_threshold, sh, ns, carry[d], per[d])) return int(PB200_OK);
    const pb200_timing first = c->timing;
    TRY(topn_pass(c, family, src_prog, n_src_nodes, n, ids2.data(), ids2.size(), min_threshold, tanimoto_threshold, sh, ns, per[d], /*reuse_src=*/true));
    c->timing.kernel_ms += first.kernel_ms; c->timing.algo_bytes += first.algo_bytes; c->timing.launches += first.launches;
    c->timing.h2d_bytes += first.h2d_bytes; c->timing.d2h_bytes += first.d2h_bytes;
    return int(PB200_OK);
  }));
  all.clear();
  for (size_t d = 0; d < nd; d++) for (auto& v : per[d]) all.push_back(std::move(v));
  std::vector<HPair> trimmed = merge_pairs(all);
  if (n != 0 && n < trimmed.size()) trimmed.resize(n);
  return finish(trimmed);
}

extern "C" int pb200_multi_topn_shard(pb200_multi* m, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes, uint32_t n,
                                      const uint64_t* ids, size_t n_ids, uint64_t min_threshold, uint64_t tanimoto_threshold, uint64_t shard,
                                      pb200_pair* out, size_t cap, size_t* n_out) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_topn_shard(c, family, src_prog, n_src_nodes, n, ids, n_ids, min_threshold, tanimoto_threshold, shard, out, cap, n_out));
}

extern "C" int pb200_multi_rows(pb200_multi* m, pb200_family family, uint64_t start, uint32_t flags, uint64_t column, uint64_t limit,
                                const uint64_t* shards, uint32_t n_shards, uint64_t* out_rows, size_t cap, size_t* n_out) {
  MENTER(m);
  if (!n_out) return PB200_ERR_INVALID;
  *n_out = 0;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  TRY(mcheck_family(m, family));
  const ShardSplit sp = split_shards(m, shards, n_shards);
  const std::vector<uint8_t> active = active_of(sp);
  std::vector<std::vector<uint64_t>> part(m->ctx.size());
  TRY(fan_out(m, active, [&](size_t d) {
    pb200_ctx* c = m->ctx[d];
    size_t n = 0;
    const uint64_t* sh = sp.shards[d].data();
    const uint32_t ns = uint32_t(sp.shards[d].size());
    int rc = pb200_rows(c, family, start, flags, column, limit, sh, ns, nullptr, 0, &n);
    if (rc != PB200_OK && rc != PB200_ERR_RANGE) return rc;
    part[d].resize(n);
    if (n == 0) return int(PB200_OK);
    return pb200_rows(c, family, start, flags, column, limit, sh, ns, part[d].data(), n, &n);
  }));
  // RowIDs.merge (executor.go:1043-1069): ascending union cut at `limit` — associative, so per-device merges fold the same way
  std::vector<uint64_t> acc;
  for (size_t d = 0; d < part.size(); d++) {
    std::vector<uint64_t> merged;
    std::set_union(acc.begin(), acc.end(), part[d].begin(), part[d].end(), std::back_inserter(merged));
    if (merged.size() > limit) merged.resize(size_t(limit));
    acc.swap(merged);
  }
  *n_out = acc.size();
  if (acc.size() > cap) return mfail(m, PB200_ERR_RANGE, "row buffer too small");
  if (!acc.empty() && !out_rows) return PB200_ERR_INVALID;
  std::copy(acc.begin(), acc.end(), out_rows);
  return PB200_OK;
}

extern "C" int pb200_multi_groupby(pb200_multi* m, const pb200_groupby_child* children, uint32_t n_children, const pb200_node* filter_prog,
                                   uint32_t n_filter_nodes, uint64_t limit, const uint64_t* shards, uint32_t n_shards, pb200_group* out, size_t cap,
                                   size_t* n_out) {
  MENTER(m);
  if (!n_out) return PB200_ERR_INVALID;
  *n_out = 0;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  const ShardSplit sp = split_shards(m, shards, n_shards);
  std::vector<uint8_t> active = active_of(sp);
  if (n_shards == 0) active[0] = 1;  // argument checks still run
  std::vector<std::vector<pb200_group>> part(m->ctx.size());
  TRY(fan_out(m, active, [&](size_t d) {
    pb200_ctx* c = m->ctx[d];
    const uint64_t* sh = sp.shards[d].data();
    const uint32_t ns = uint32_t(sp.shards[d].size());
    size_t n = 0;
    size_t room = 1024;
    for (;;) {
      part[d].resize(room);
      const int rc = pb200_groupby(c, children, n_children, filter_prog, n_filter_nodes, limit, sh, ns, part[d].data(), room, &n);
      if (rc == PB200_ERR_RANGE && n > room) { room = n; continue; }
      if (rc != PB200_OK) return rc;
      part[d].resize(n);
      return int(PB200_OK);
    }
  }));
  // mergeGroupCounts (executor.go:1166-1200): the groups a device kept are the first `limit` of its shards' merge, so
  // folding the devices gives what folding the shards one by one gives
  std::vector<pb200_group> acc;
  for (size_t d = 0; d < part.size(); d++)
    if (active[d]) acc = merge_group_counts(std::move(acc), part[d], limit, n_children);
  *n_out = acc.size();
  if (acc.size() > cap) return mfail(m, PB200_ERR_RANGE, "group buffer too small");
  if (!acc.empty() && !out) return PB200_ERR_INVALID;
  std::copy(acc.begin(), acc.end(), out);
  timing_merge(m, active);
  return PB200_OK;
}

extern "C" int pb200_multi_eval(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, void* out, size_t cap, size_t* out_len) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_eval(c, prog, n_nodes, shard, out, cap, out_len));
}

extern "C" int pb200_multi_eval_columns(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, uint64_t* out_cols, size_t cap,
                                        size_t* out_n) {
  MENTER(m);
  pb200_ctx* c = MOWNER(m, shard);
  MFORWARD(m, c, pb200_eval_columns(c, prog, n_nodes, shard, out_cols, cap, out_n));
}

extern "C" int pb200_multi_eval_shards(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards, void* out,
                                       size_t cap, uint64_t* out_off, size_t* out_len) {
  MENTER(m);
  if (!out_len) return PB200_ERR_INVALID;
  *out_len = 0;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  if (n_shards == 0 || m->ctx.size() == 1) MFORWARD(m, m->ctx[0], pb200_eval_shards(m->ctx[0], prog, n_nodes, shards, n_shards, out, cap, out_off, out_len));
  if (n_shards > (1u << 24)) return mfail(m, PB200_ERR_RANGE, "too many shards in one call");
  const uint64_t* shards_dev = nullptr;
  TRY(gather_dense(m, prog, n_nodes, shards, n_shards, &shards_dev));
  pb200_ctx* c0 = m->ctx[0];
  std::lock_guard<std::mutex> lock(c0->mu);
  DeviceGuard g(c0->device);
  GatherSwap swap(m);
  int rc = emit_rows_from_dense(c0, shards_dev, n_shards);
  if (rc == PB200_OK) rc = fetch_rows(c0, n_shards, out, cap, out_off, out_len);
  if (rc != PB200_OK) m->err = c0->err;
  m->timing.d2h_bytes += c0->timing.d2h_bytes;
  return rc;
}

extern "C" int pb200_multi_eval_proto_columns(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards, uint32_t n_shards,
                                              void* out, size_t cap, size_t* out_len) {
  MENTER(m);
  if (!out_len) return PB200_ERR_INVALID;
  *out_len = 0;
  if (n_shards && !shards) return mfail(m, PB200_ERR_INVALID, "null shard list");
  if (n_shards == 0 || m->ctx.size() == 1) MFORWARD(m, m->ctx[0], pb200_eval_proto_columns(m->ctx[0], prog, n_nodes, shards, n_shards, out, cap, out_len));
  if (n_shards > (1u << 24)) return mfail(m, PB200_ERR_RANGE, "too many shards in one call");
  const uint64_t* shards_dev = nullptr;
  TRY(gather_dense(m, prog, n_nodes, shards, n_shards, &shards_dev, true));  // short containers travel as value lists
  pb200_ctx* c0 = m->ctx[0];
  std::lock_guard<std::mutex> lock(c0->mu);
  DeviceGuard g(c0->device);
  GatherSwap swap(m);
  const int rc = emit_proto_from_dense(c0, shards_dev, n_shards, out, cap, out_len);
  if (rc != PB200_OK) m->err = c0->err;
  m->timing.d2h_bytes += c0->timing.d2h_bytes;
  return rc;
}
