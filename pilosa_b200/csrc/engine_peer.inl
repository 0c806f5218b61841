// Included by engine.cu: summing a Count batch over the GPUs of a box when every GPU has its own PROCESS (the layout
// `torchrun` gives: one rank per device). The reduce function of executeCount is `uint64 +` (executor.go:1813); here it
// is the tail of the batch's own kernel: the last CTA to retire adds this GPU's partial counts into a mailbox slot of
// every rank with system-scope reductions over NVLink (the peers' mailboxes are mapped into this process through CUDA
// IPC; plain peer stores, nothing is multicast) and bumps the slot's arrival counter. No collective library runs
// between two batches, so no rank ever waits inside the launch sequence for another one: a slot's total is picked
// up two batches later (or when the host asks for it) by a one-CTA kernel that waits for the counter.
//
// Epochs: batch e uses slot e & 3 of every mailbox. Before batch e is launched, slot (e - 2) & 3 is collected in
// stream order (if the host has not fetched it already), which also re-arms it; a peer can therefore be at most two
// batches ahead, and never writes into a slot that still holds an older batch.

namespace {

constexpr size_t kPeerSlotBytes = size_t(PB200_PEER_MAX_COUNTS) * 8;
constexpr size_t kPeerFlagOff = 4 * kPeerSlotBytes;          // 4 arrival counters, 128 B apart
constexpr size_t kPeerTicketOff = kPeerFlagOff + 4 * 128;    // 4 CTA tickets (local use), 128 B apart
constexpr size_t kPeerErrOff = kPeerTicketOff + 4 * 128;
constexpr size_t kPeerBoxBytes = kPeerErrOff + 128;

// Picks up the total of the batch in `slot`: a one-CTA kernel on the COPY stream (it waits for the peers' counts beside
// whatever the compute stream is running — launches with a peer reduction leave one CTA slot free for it).
int peer_collect(pb200_ctx* ctx, int slot) {
  PeerState& ps = ctx->peer;
  PeerState::Pending& pd = ps.pend[slot];
  if (!pd.open) return PB200_OK;
  if (!ctx->copy_stream) CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copy_stream, ps.launched[slot], 0));
  peer_collect_kernel<<<1, 256, 0, ctx->copy_stream>>>(reinterpret_cast<unsigned long long*>(ps.box + size_t(slot) * kPeerSlotBytes),
                                                      reinterpret_cast<unsigned long long*>(ps.box + kPeerFlagOff + size_t(slot) * 128), uint32_t(ps.world),
                                                      pd.n, pd.target, reinterpret_cast<unsigned int*>(ps.box + kPeerErrOff));
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaEventRecord(ps.collected[slot], ctx->copy_stream));
  ps.collected_valid[slot] = true;
  pd.open = false;
  ctx->timing.launches++;
  return PB200_OK;
}

// Opens the next epoch for a batch whose partial counts are accumulated at `target`: flow control, descriptor.
int peer_begin_epoch(pb200_ctx* ctx, unsigned long long* target, uint32_t n, PushDesc& out, int* slot_out) {
  PeerState& ps = ctx->peer;
  if (!ps.on) return fail(ctx, PB200_ERR_INVALID, "no peers connected (pb200_peer_connect)");
  if (n > PB200_PEER_MAX_COUNTS) return fail(ctx, PB200_ERR_RANGE, "batch larger than PB200_PEER_MAX_COUNTS queries");
  const uint64_t e = ps.epoch++;
  const int slot = int(e & 3);
  if (e >= 2) {  // batch e-2 is picked up (which re-arms its slot) before this batch's kernel may push: peers stay within two batches
    const int old = int((e - 2) & 3);
    TRY(peer_collect(ctx, old));
    if (ps.collected_valid[old]) CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ps.collected[old], 0));
  }
  if (ps.pend[slot].open) {  // cannot happen with the rule above; kept as a guard
    TRY(peer_collect(ctx, slot));
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ps.collected[slot], 0));
  }
  out = PushDesc{};
  out.n_dst = uint32_t(ps.world);
  out.n = n;
  out.ticket = reinterpret_cast<unsigned int*>(ps.box + kPeerTicketOff + size_t(slot) * 128);  // left at zero by its last user
  out.local = target;
  for (int d = 0; d < ps.world; d++) {
    out.dst[d] = reinterpret_cast<unsigned long long*>(ps.peer_box[d] + size_t(slot) * kPeerSlotBytes);
    out.flag[d] = reinterpret_cast<unsigned long long*>(ps.peer_box[d] + kPeerFlagOff + size_t(slot) * 128);
  }
  ps.pend[slot].target = target; ps.pend[slot].n = n; ps.pend[slot].open = true;
  *slot_out = slot;
  return PB200_OK;
}

// After the batch's kernels (and its push) are queued on the compute stream.
int peer_end_epoch(pb200_ctx* ctx, int slot) {
  CUDA_TRY(ctx, cudaEventRecord(ctx->peer.launched[slot], ctx->stream));
  return PB200_OK;
}

int peer_check_error(pb200_ctx* ctx) {  // after a synchronising copy: did a collect give up waiting?
  PeerState& ps = ctx->peer;
  if (!ps.on) return PB200_OK;
  unsigned int e = 0;
  CUDA_TRY(ctx, cudaMemcpy(&e, ps.box + kPeerErrOff, 4, cudaMemcpyDeviceToHost));
  return e ? fail(ctx, PB200_ERR_CUDA, "peer reduction timed out: a participating GPU did not deliver its counts") : PB200_OK;
}

void peer_teardown(pb200_ctx* ctx) {
  PeerState& ps = ctx->peer;
  for (int d = 0; d < ps.world; d++)
    if (ps.on && d != ps.rank && ps.peer_box[d]) cudaIpcCloseMemHandle(ps.peer_box[d]);
  if (ps.box) cudaFree(ps.box);
  for (int k = 0; k < 4; k++) { if (ps.launched[k]) cudaEventDestroy(ps.launched[k]); if (ps.collected[k]) cudaEventDestroy(ps.collected[k]); }
  ps = PeerState{};
}

}  // namespace

extern "C" int pb200_peer_export(pb200_ctx* ctx, void* handle_out) {
  ENTER(ctx);
  if (!handle_out) return PB200_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == PB200_PEER_HANDLE_BYTES, "IPC handle size");
  PeerState& ps = ctx->peer;
  if (!ps.box) {
    CUDA_TRY(ctx, cudaMalloc(reinterpret_cast<void**>(&ps.box), kPeerBoxBytes));
    CUDA_TRY(ctx, cudaMemset(ps.box, 0, kPeerBoxBytes));
    for (int k = 0; k < 4; k++) {
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ps.launched[k], cudaEventDisableTiming));
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&ps.collected[k], cudaEventDisableTiming));
    }
  }
  cudaIpcMemHandle_t h;
  CUDA_TRY(ctx, cudaIpcGetMemHandle(&h, ps.box));
  std::memcpy(handle_out, &h, sizeof h);
  return PB200_OK;
}

extern "C" int pb200_peer_connect(pb200_ctx* ctx, int rank, int world, const void* handles) {
  ENTER(ctx);
  PeerState& ps = ctx->peer;
  if (!handles || world < 1 || world > kPeerMax || rank < 0 || rank >= world) return fail(ctx, PB200_ERR_INVALID, "bad rank / world (at most 8 participants)");
  if (!ps.box) return fail(ctx, PB200_ERR_INVALID, "call pb200_peer_export first");
  if (ps.on) return fail(ctx, PB200_ERR_INVALID, "peers already connected");
  for (int d = 0; d < world; d++) {
    if (d == rank) { ps.peer_box[d] = ps.box; continue; }
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const uint8_t*>(handles) + size_t(d) * sizeof h, sizeof h);
    void* p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int k = 0; k < d; k++) if (k != rank && ps.peer_box[k]) { cudaIpcCloseMemHandle(ps.peer_box[k]); ps.peer_box[k] = nullptr; }
      return fail(ctx, PB200_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
    }
    ps.peer_box[d] = static_cast<uint8_t*>(p);
  }
  ps.rank = rank; ps.world = world; ps.on = true; ps.epoch = 0;
  return PB200_OK;
}

namespace {
int pb200_peer_flush_locked(pb200_ctx* ctx) {
  if (!ctx->peer.on) return PB200_OK;
  const uint64_t e = ctx->peer.epoch;
  for (uint64_t k = e >= 4 ? e - 4 : 0; k < e; k++) {  // oldest first
    const int slot = int(k & 3);
    const bool was_open = ctx->peer.pend[slot].open;
    TRY(peer_collect(ctx, slot));
    if (was_open) CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->peer.collected[slot], 0));  // later work on the compute stream sees the totals
  }
  return PB200_OK;
}
}  // namespace

extern "C" int pb200_peer_flush(pb200_ctx* ctx) {
  ENTER(ctx);
  return pb200_peer_flush_locked(ctx);
}

extern "C" int pb200_batch_launch_allreduce(pb200_ctx* ctx, pb200_batch* b) {
  ENTER(ctx);
  if (!b) return PB200_ERR_INVALID;
  TRY(batch_revalidate(ctx, b));
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  ctx->timing.launches = 0;
  PushDesc pd;
  int slot = 0;
  TRY(peer_begin_epoch(ctx, static_cast<unsigned long long*>(b->d_counts.p), b->n_queries, pd, &slot));
  CUDA_TRY(ctx, cudaMemsetAsync(b->d_counts.p, 0, size_t(b->n_queries) * 8, ctx->stream));
  TRY(launch_eval(ctx, b->bd, b->bd.progs, b->n_queries, b->n_shards, b->cp.may_need_staging, static_cast<unsigned long long*>(b->d_counts.p), 1, 0,
                  nullptr, nullptr, max_leaves_of(b->cp), b->cp.any_bushy, fold_all_of(b->cp), &pd, b->cp.any_flat, tiny_of(b->cp)));
  return peer_end_epoch(ctx, slot);
}

// ---------------------------------------------------------------------------------------
// Scratch reservation: the query path grows its device / pinned scratch on demand (cudaMalloc, and a cudaFree of the
// smaller buffer, which synchronises the whole device). A serving host sizes it once, up front, for the largest batch
// it will issue; afterwards no Count / batch call of that size or smaller allocates.
// ---------------------------------------------------------------------------------------
extern "C" int pb200_reserve(pb200_ctx* ctx, uint32_t max_queries, uint32_t max_shards, uint32_t max_nodes_per_query) {
  ENTER(ctx);
  if (!max_queries || !max_shards) return PB200_ERR_INVALID;
  const uint64_t n_items = uint64_t(max_queries) * max_shards * kRowContainers;
  if (n_items >= (1ull << 32)) return fail(ctx, PB200_ERR_RANGE, "queries x shards x 16 must be < 2^32");
  const uint32_t nodes = std::max<uint32_t>(3, max_nodes_per_query);
  const uint32_t leaves = std::min<uint32_t>(nodes, uint32_t(kMaxLeaves));
  // upload image of a batch: programs, leaves, instructions, shard list (upload_batch)
  const size_t batch_bytes = 2 * size_t(max_queries) * (sizeof(Prog) + size_t(nodes) * (sizeof(Leaf) + sizeof(Instr))) + size_t(max_shards) * 8 + 256;
  TRY(ensure_dev(ctx, ctx->d_misc, 64));
  TRY(ensure_dev(ctx, ctx->d_slow, size_t(n_items) * 4));
  TRY(ensure_dev(ctx, ctx->d_rest, size_t(n_items) * 4));  // the lane-per-item pass's hand-over list (very sparse rows)
  const uint64_t desc_bytes = n_items * uint64_t(leaves) * sizeof(ContainerDesc);
  if (desc_bytes <= (256ull << 20)) TRY(ensure_dev(ctx, ctx->d_slow_descs, size_t(desc_bytes)));
  TRY(ensure_dev(ctx, ctx->d_batch, batch_bytes));
  TRY(ensure_pin(ctx, ctx->h_stage, batch_bytes));
  TRY(ensure_dev(ctx, ctx->d_out, (size_t(max_queries) + 2) * 8));
  TRY(ensure_pin(ctx, ctx->h_out, size_t(max_queries) * 8 + 8));
  TRY(ensure_pin(ctx, ctx->h_algo, 64));
  if (!ctx->copy_stream) CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  for (CountSlot& sl : ctx->slots) {  // the two slots of pb200_count / pb200_count_batch(_begin)
    if (!sl.k0) {
      CUDA_TRY(ctx, cudaEventCreate(&sl.k0)); CUDA_TRY(ctx, cudaEventCreate(&sl.k1));
      CUDA_TRY(ctx, cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    std::swap(ctx->d_out, sl.d_out); std::swap(ctx->d_batch, sl.d_batch); std::swap(ctx->h_stage, sl.h_stage); std::swap(ctx->h_out, sl.h_out);
    int rc = ensure_dev(ctx, ctx->d_out, (size_t(max_queries) + 2) * 8);
    if (rc == PB200_OK) rc = ensure_dev(ctx, ctx->d_batch, batch_bytes);
    if (rc == PB200_OK) rc = ensure_pin(ctx, ctx->h_stage, batch_bytes);
    if (rc == PB200_OK) rc = ensure_pin(ctx, ctx->h_out, size_t(max_queries) * 8 + 8);
    std::swap(ctx->d_out, sl.d_out); std::swap(ctx->d_batch, sl.d_batch); std::swap(ctx->h_stage, sl.h_stage); std::swap(ctx->h_out, sl.h_out);
    TRY(rc);
  }
  TRY(sync_families(ctx));
  return PB200_OK;
}
