// Host-side engine state behind the C ABI (include/pilosa_b200.h).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/pilosa_b200.h"
#include "host_roaring.h"
#include "layout.h"

namespace pb200 {

struct DevBuf {  // growable cudaMalloc-backed scratch
  void* p = nullptr;
  size_t cap = 0;
};
struct PinBuf {  // growable pinned host staging
  void* p = nullptr;
  size_t cap = 0;
};

// Where rewritten containers and rebuilt row tables go (engine_ingest.inl): large chunks, bump-allocated, 128 B aligned.
struct WriteArena {
  uint8_t* base = nullptr;             // the chunk being filled
  size_t cap = 0, top = 0;             // top: host copy of the bump pointer (authoritative between launches)
  unsigned long long* d_top = nullptr; // device copy the ingest kernel bumps
  std::vector<std::pair<void*, size_t>> retired;  // full chunks (pointer, bytes) that descriptors still point into
  uint32_t live_frags = 0;             // fragments referencing the arena; at 0 everything is reclaimed
};

struct Fragment {
  bool in_arena = false;  // some descriptor or the row table points into the write arena
  bool maybe_empty_rows = false;  // a write may have emptied a listed row (row_counts == 0): Rows / GroupBy skip those
  void* blob = nullptr;  // one cudaMalloc: [row_ids | descs | payloads]
  size_t blob_bytes = 0;
  FragDev dev{};
  size_t row_ids_off = 0, descs_off = 0, payload_off = 0;
  uint64_t n_containers = 0, payload_bytes = 0;
  bool all_bitmap = true;
  std::vector<uint64_t> row_ids;     // ascending ids of the rows present
  std::vector<uint64_t> row_counts;  // exact cardinality per present row
  // TopN candidate order (rank cache, cache.go:136-288). Either handed over by the host
  // (pb200_rank_set) or derived from exact row counts (count desc, id asc, 50000 entries).
  std::vector<uint64_t> rank_ids, rank_counts;
  bool rank_from_host = false;
  uint32_t* d_rank = nullptr;  // device copy of the rank order as row-table indices (TopN) ...
  uint64_t* d_rank_counts = nullptr;  // ... and of the cached counts in rank order (same allocation)
  bool d_rank_valid = false;
  std::vector<uint32_t> rank_pos_of_row;  // rank position of row_ids[j] (0xffffffff: not ranked); host index for TopN pass 2
  bool rank_pos_valid = false;
  // rankCache state (cache.go:136-288) for the import recount of fragment.importRoaring
  // (fragment.go:2258-2276): id -> cached count and the admission threshold. Built on the first
  // pb200_frag_import_roaring of the fragment, carried over every rewrite of its device copy.
  struct RankState {
    bool live = false;
    uint64_t threshold_value = 0;
    std::unordered_map<uint64_t, uint64_t> entries;
  } rank_state;
};

struct Family {
  std::unordered_map<uint64_t, Fragment> frags;
  std::vector<FragDev> h_frags;  // indexed by shard
  FragDev* d_frags = nullptr;
  uint32_t cap = 0;
  uint64_t n_not_all_bitmap = 0;  // fragments holding array/run containers (local ones and, in a pb200_multi, the peers')
  uint64_t payload_total = 0, containers_total = 0;  // over the local fragments: mean container size steers the evaluator choice
  std::vector<uint8_t> remote_flags;  // pb200_multi: per shard, bit 0 = the entry points into a peer's HBM, bit 1 = not all bitmap
};

// A compiled batch of programs, ready for upload.
struct Compiled {
  std::vector<Prog> progs;
  std::vector<Instr> instrs;
  std::vector<Leaf> leaves;
  struct PreTask {
    int kind;             // 0 = BSI range -> temp slot, 1 = shift(temp a) -> temp b, 2 = eval sub-program -> temp slot
    pb200_node node{};    // kind 0
    uint32_t slot = 0, src_slot = 0;
    Prog prog{};          // kind 2: sub-program (offsets into instrs / leaves)
    bool union_rows = false;              // kind 2 that is a plain Union of Row leaves: one accumulating pass (K13)
    bool xor_rows = false;                // ... or a plain Xor of Row leaves: the same pass, flipping instead of setting
    bool count_only = false;              // kind 0 feeding a bare Count(): count in the comparator kernel
    unsigned long long* count_out = nullptr;
    uint32_t count_sstride = 0;
  };
  std::vector<PreTask> pre;
  uint32_t n_temp_slots = 0;
  bool may_need_staging = false;  // some leaf family holds array/run containers (or temps)
  uint64_t leaf_payload = 0, leaf_containers = 0;  // summed over the Row leaves' families (see Family)
  bool any_flat = false;          // some program is a fold eval_fold_kernel takes (LOAD, then AND / OR steps against leaves)
  bool any_bushy = false;         // some program is not a flat left fold (stack ops, Not): eval_staged_kernel is needed beside eval_fold_kernel
};

// One Count batch in flight (pb200_count_batch_begin .. _end). Two slots let a host compile and
// upload batch i+1 while the kernels of batch i run; results leave on a separate copy stream.
struct CountSlot {
  DevBuf d_batch, d_out;
  PinBuf h_stage, h_out;
  cudaEvent_t k0 = nullptr, k1 = nullptr, done = nullptr;
  size_t n_out = 0;
  bool launched = false;
  int peer_slot = -1;     // mailbox slot of the batch's peer reduction (PB200_OPT_PEER_ALLREDUCE), -1: none
  bool external = false;  // the caller took the device counts (pb200_count_batch_device) and may have queued work on them
  uint64_t seq = 0;  // begin order
};

}  // namespace pb200

namespace pb200 {
// One single-query Count call waiting to be combined with the calls of other host threads (engine.cu: combine_counts).
struct CountReq {
  const pb200_node* prog; uint32_t n_nodes;
  const uint64_t* shards; uint32_t n_shards;
  uint64_t* out;
  int rc = 0;
  CountReq* next = nullptr;  // arrival list (newest first), linked with one compare-and-swap per caller
  std::atomic<bool> done{false};
  std::string err;
};
// Peer reduction between processes (engine_peer.inl): this context's mailbox and the peers' as mapped here.
struct PeerState {
  bool on = false;
  int rank = 0, world = 1;
  uint8_t* box = nullptr;
  uint8_t* peer_box[8] = {};
  uint64_t epoch = 0;
  struct Pending { unsigned long long* target = nullptr; uint32_t n = 0; bool open = false; } pend[4];
  cudaEvent_t launched[4] = {}, collected[4] = {};  // per slot: the batch's kernels are queued / its total has been picked up
  bool collected_valid[4] = {};
};
}  // namespace pb200

struct pb200_ctx {
  pb200::PeerState peer;
  bool peer_allreduce = false;  // PB200_OPT_PEER_ALLREDUCE: the split Count batches (_begin / _end) return counts summed over the peers
  int device = 0;
  // combining queue of pb200_count: threads that arrive while a launch is in flight are folded into the next batch
  std::mutex cq_mu;
  std::condition_variable cq_cv;
  std::atomic<pb200::CountReq*> cq_head{nullptr};  // arrivals: pushed lock-free by the callers, taken whole by the combiner
  std::vector<pb200::CountReq*> cq;                // taken, not yet served; owned by whoever holds cq_busy (cq_mu: only for callers that sleep)
  std::atomic<bool> cq_busy{false};
  size_t cq_last = 0;  // requests in the last round
  size_t cq_target = 0;               // ... plus the ones that arrived during it: the callers the next round waits for
  std::atomic<size_t> cq_n{0};        // requests filed and not yet launched
  std::atomic<int> cq_sleepers{0};    // callers that gave up spinning and wait on cq_cv
  uint64_t cq_batches = 0, cq_queries = 0;  // statistics: launches made for how many single-query calls
  int sm_count = 148;
  size_t budget = 0, used = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::mutex mu;
  std::string err;
  std::vector<pb200::Family> families;
  pb200::FamilyDev* d_families = nullptr;
  uint32_t d_families_cap = 0;
  bool families_dirty = true;
  // pb200_multi: called with every new / dropped fragment-table entry of this device, before the previous copy of the
  // fragment is freed, so that the peers' tables follow and no peer still reads the old copy
  std::function<int(uint32_t family, uint64_t shard, const pb200::FragDev& dev, bool all_bitmap, bool present)> publish;
  uint64_t generation = 0;  // bumped by every load / write / drop / table growth: prepared batches re-validate against it
  pb200::DevBuf d_gb;            // GroupBy: group totals summed over the shards on the device
  pb200::DevBuf d_sparse;        // value lists of short result containers (K12 path, EvalParams::sparse_out)
  bool dense_lists = false;      // d_dense_n carries kSparseFlag entries backed by d_sparse
  pb200::DevBuf d_slow_descs, d_ovf, d_rest;  // d_rest: items eval_tiny_kernel leaves to eval_fast_kernel
  pb200::DevBuf d_rows, d_rows_off, d_rows_meta, d_rows_item_off;  // K11: serialised result rows (kernels_rows.cuh)
  bool dense_has_ovf = false;                      // the rows in d_dense carry Shift overflow words
  unsigned long long* dense_ovf_dev = nullptr;
  pb200::DevBuf d_batch, d_out, d_slow, d_temp, d_dense, d_dense_n, d_misc, d_topn, d_bsi;
  pb200::PinBuf h_stage, h_out, h_algo, h_topn, h_ing, h_ing_res;
  pb200::DevBuf d_ing, d_ing_res, d_topn_compact;
  bool topn_prune = true;  // PB200_OPT_TOPN_PRUNE
  pb200::WriteArena arena;
  uint32_t densify_min_bytes = 0;  // PB200_OPT_DENSIFY_MIN_BYTES (0 = containers stay as loaded)
  uint32_t rank_cache_size = 50000;  // PB200_OPT_RANK_CACHE_SIZE (field.go:47 DefaultCacheSize)
  bool topn_attr_set = false, sum_attr_set = false, mm_attr_set = false;
  uint32_t range_attr_set = 0;  // per instantiation of bsi_range_pipe_kernel
  int occ_range[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int occ[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // resident CTAs per SM of each persistent kernel (occupancy calculator)
  const uint64_t* dense_shards_dev = nullptr;  // shard list of the source rows currently in d_dense (TopN pass 1 -> 2)
  pb200::CountSlot slots[2];     // split Count batches in flight (count_begin / count_end), oldest first
  int n_pending = 0, head = 0;   // ring of pending slots
  uint64_t begin_seq = 0;
  cudaStream_t copy_stream = nullptr;
  unsigned long long* algo_dev = nullptr;  // set while a Count call is enqueued: its own algorithmic-bytes counter (behind its counts) ...
  uint32_t* slow_dev = nullptr;            // ... and deferred-item counters; null: the words in d_misc
  bool slow_zeroed = false;                // the call's memset has cleared slow_dev already (consumed by its first launch_eval)
  pb200_timing timing{};
};
