/*
 * pilosa_b200.h — C ABI of libpilosa_b200.so, the B200 (sm_100a) execution engine for
 * Pilosa's per-shard bitmap query hot path.
 *
 * The reference (travisturner/pilosa) is pure Go and has no FFI for this path; the
 * boundary below is what a cgo shim placed behind the reference's own call surface
 * binds (see INTEGRATION.md for the shim). Each entry point cites the reference
 * function(s) it replaces, as file:line in the reference tree.
 *
 * Conventions
 *  - every function returns an int status: 0 = OK, <0 = error; pb200_strerror(status)
 *    gives a static message and pb200_last_error(ctx) the detailed text of the last
 *    failure on that context (Go side: turn into `error`; never a crash/longjmp).
 *  - all integers are fixed width, little endian; all buffers are caller allocated.
 *  - the library never retains a host pointer after a call returns (cgo pointer rule;
 *    reference containers may alias an mmap that is unmapped at the next snapshot,
 *    fragment.go:394-431).
 *  - a context is bound to ONE CUDA device. Calls on the same context are serialised
 *    by an internal mutex (the reference serialises per fragment, fragment.go:601);
 *    use one context per GPU and batch entry points for throughput. pb200_multi (below)
 *    puts the GPUs of a box behind one handle: partition, fan-out and reduce inside the call.
 *  - there is NO CPU fallback: if no CUDA device is usable pb200_init fails with
 *    PB200_ERR_CUDA and nothing else can be called.
 */
#ifndef PILOSA_B200_H
#define PILOSA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PB200_ABI_VERSION 1

/* ShardWidth = 2^20 columns (shardwidth/20.go, fragment.go:53): 16 containers per row. */
#define PB200_SHARD_WIDTH_EXP 20
#define PB200_SHARD_WIDTH (1ull << PB200_SHARD_WIDTH_EXP)
#define PB200_ROW_CONTAINERS 16

/* status codes */
#define PB200_OK 0
#define PB200_ERR_INVALID (-1)   /* bad argument / malformed expression            */
#define PB200_ERR_CUDA (-2)      /* CUDA runtime error (text in pb200_last_error)   */
#define PB200_ERR_FORMAT (-3)    /* malformed Pilosa roaring bytes                  */
#define PB200_ERR_NOMEM (-4)     /* HBM budget exhausted                            */
#define PB200_ERR_NOTFOUND (-5)  /* unknown family / fragment                       */
#define PB200_ERR_RANGE (-6)     /* caller buffer too small / size limit exceeded   */
#define PB200_ERR_UNSUPPORTED (-7)

typedef struct pb200_ctx pb200_ctx;

/* A "family" is one (index, field, view) triple: the set of fragments, one per shard,
 * that holder.go:496-502 Holder.fragment(index, field, view, shard) would return. */
typedef uint32_t pb200_family;

/* ---- expression programs ------------------------------------------------------------
 * A bitmap call tree (executor.go:654-680 executeBitmapCallShard) is passed flattened in
 * post order. N-ary nodes consume their `nargs` operands from the evaluation stack and
 * combine them as a LEFT FOLD in PQL order, exactly like executor.go:1668-1733,1003-1025.
 * As in the reference, an n-ary node may have any number of operands: one device program holds
 * 32 leaves, and a call with more is compiled with its wide operations over plain Row operands
 * evaluated into dense temp rows first (same result; a Union of rows in one extra pass whatever
 * its width, Intersect / Xor in passes of 32 rows).
 */
enum pb200_op {
  PB200_OP_ROW = 1,        /* Row(f=row): fragment.row, fragment.go:600-641; executor.go:1444-1499 */
  PB200_OP_INTERSECT = 2,  /* executeIntersectShard executor.go:1668; Row.Intersect row.go:107     */
  PB200_OP_UNION = 3,      /* executeUnionShard executor.go:1693; Row.Union row.go:153             */
  PB200_OP_DIFFERENCE = 4, /* executeDifferenceShard executor.go:1003; Row.Difference row.go:198   */
  PB200_OP_XOR = 5,        /* executeXorShard executor.go:1715; Row.Xor row.go:133                 */
  PB200_OP_NOT = 6,        /* executeNotShard executor.go:1737-1768: family = existence field,
                              row = 0, one operand: exists.row(0) \ operand                         */
  PB200_OP_SHIFT = 7,      /* executeShiftShard executor.go:1771; Bitmap.Shift roaring.go:944 (one step; nest for n>1;
                              each step is a leaf of the call, so at most 31 deep). The bit leaving the shard's last column is kept under container key
                              (shard+1)*16 of the same segment, as roaring.go:970-974 does. */
  PB200_OP_BSI_RANGE = 8,  /* fragment.rangeOp / rangeBetween / notNull fragment.go:1271-1534      */
  PB200_OP_EMPTY = 9       /* NewRow(): used by hosts for out-of-range BSI predicates
                              (executor.go:1611,1641) and Union()/Xor() with no children           */
};

enum pb200_bsi_op {        /* pql/token.go:22-31 */
  PB200_BSI_EQ = 1, PB200_BSI_NEQ = 2, PB200_BSI_LT = 3, PB200_BSI_LTE = 4,
  PB200_BSI_GT = 5, PB200_BSI_GTE = 6, PB200_BSI_BETWEEN = 7, PB200_BSI_NOTNULL = 8
};

typedef struct pb200_node {
  uint32_t op;         /* enum pb200_op                                                    */
  uint32_t family;     /* ROW / NOT / BSI_RANGE: family handle                             */
  uint64_t row;        /* ROW: row id                                                      */
  uint32_t nargs;      /* INTERSECT/UNION/DIFFERENCE/XOR: operand count (>=1; UNION/XOR >=0)*/
  uint32_t bsi_op;     /* BSI_RANGE: enum pb200_bsi_op                                     */
  uint32_t bit_depth;  /* BSI_RANGE: bsiGroup.BitDepth (field.go:1505-1512)                */
  uint32_t reserved;
  int64_t a;           /* BSI_RANGE: predicate (already re-based by the host,
                          field.go:1526-1567 baseValue / baseValueBetween lo)              */
  int64_t b;           /* BSI_RANGE BETWEEN: hi predicate                                  */
} pb200_node;

typedef struct pb200_pair {  /* cache.go:337-341 Pair{ID, Count} */
  uint64_t id;
  uint64_t count;
} pb200_pair;

typedef struct pb200_timing {
  double kernel_ms;        /* device time of the kernels of the last query call (CUDA events) */
  double h2d_ms, d2h_ms;   /* copies issued by the last call                                   */
  uint64_t algo_bytes;     /* algorithmic bytes of the last call: sum of operand container
                              payloads read once + 16 B descriptor each        */
  uint64_t h2d_bytes, d2h_bytes;
  uint32_t launches;       /* kernels launched by the last call                               */
  uint32_t reserved;
} pb200_timing;

/* ---- lifecycle ------------------------------------------------------------------------ */

/* Bind a context to CUDA device `device`. hbm_budget_bytes == 0 means "no explicit budget".
 * Replaces nothing in the reference (there is no device today); called where
 * server.Open wires the holder (server.go). */
int pb200_init(int device, size_t hbm_budget_bytes, pb200_ctx** out);
int pb200_shutdown(pb200_ctx* ctx);
const char* pb200_strerror(int status);
const char* pb200_last_error(pb200_ctx* ctx);
int pb200_abi_version(void);
int pb200_device_count(void);   /* usable CUDA devices (0 when there is none: nothing else can then be called) */

/* Run the engine's kernels on a caller-owned CUDA stream (a cudaStream_t passed as void*,
 * e.g. torch.cuda.current_stream().cuda_stream). NULL restores the context's own stream. */
int pb200_set_stream(pb200_ctx* ctx, void* cuda_stream);

/* ---- fragments (storage objects resident in HBM) --------------------------------------- */

/* holder.go:496-502: create the (index, field, view) handle fragments are filed under. */
int pb200_family_create(pb200_ctx* ctx, pb200_family* out);

/* Load / replace the fragment of `family` for `shard` from the exact bytes of a Pilosa
 * fragment file: roaring snapshot + appended ops log, which is replayed
 * (roaring.go:1562-1653 unmarshalPilosaRoaring; format roaring.go:1044-1122; ops
 * roaring.go:4459-4559). Go passes its mmap slice (fragment.go:330-401 openStorage).
 * Container keys must lie in rows of this shard: key = row*16 + (col>>16)
 * (fragment.go:3078-3081 pos). The bytes are copied; nothing is retained. */
int pb200_frag_load(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                    const void* roaring_bytes, size_t len);

/* Write-path coherence: apply set/clear of storage positions (pos = row*2^20 + col%2^20)
 * to the device copy; hook for fragment.go:680-779 (set/clearBit), :2051 importPositions,
 * :2243 importRoaring. Clears are applied after sets. The rank cache follows importPositions
 * (:2080-2093): the cached count of every row a position names is re-read from the rewritten
 * fragment (BulkAdd: dropped when below the admission threshold, so an emptied row keeps its old
 * count as in the reference), then the ranking is rebuilt; see pb200_frag_import_roaring for the
 * cache state the engine keeps. */
int pb200_frag_apply(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                     const uint64_t* set_pos, size_t n_set,
                     const uint64_t* clear_pos, size_t n_clear);

/* Union (clear = 0) or subtract (clear = 1) a serialized roaring bitmap of fragment positions
 * (row * 2^20 + column % 2^20) into the resident fragment and report how many bits changed:
 * fragment.importRoaring (fragment.go:2243) -> Bitmap.ImportRoaringBits (roaring.go:1463-1558).
 * `data` is either Pilosa roaring (with ops log) or the OFFICIAL roaring format with 16-bit
 * keys that /import-roaring receives (readOfficialHeader .. readWithRuns roaring.go:5075-5245;
 * run payloads are start:length there). The fragment is created if the shard had none.
 * The rank-cache update of importRoaring (fragment.go:2258-2276) is done here as well: the engine
 * keeps the fragment's rankCache entries and admission threshold (cache.go:136-288; seeded from the
 * list pb200_rank_set handed over, else from the exact row counts, cut at
 * PB200_OPT_RANK_CACHE_SIZE), moves the cached count of every changed row by its signed change
 * (BulkAdd(row, Get(row) + changes), uint64 arithmetic, values below the threshold dropped) and
 * re-ranks at once (Recalculate), so the next TopN walks the order the reference would. Ties rank
 * by ascending id (the reference's sort leaves them unspecified). pb200_rank_get reads the
 * ranking back for the Go cache; pb200_rank_set replaces it with the host's list and
 * pb200_frag_load with a fresh ranking of exact counts. */
int pb200_frag_import_roaring(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                              const void* data, size_t len, int clear, int64_t* changed);

/* fragment.go:574 closeStorage / drop of a fragment. */
int pb200_frag_drop(pb200_ctx* ctx, pb200_family family, uint64_t shard);

/* The Go rank cache stays authoritative for TopN candidate order (cache.go:136-288
 * rankCache.Top(), fragment.go:1702-1738 topBitmapPairs): hand the engine the ranked
 * (row id, cached count) list of one fragment, in rank order. */
int pb200_rank_set(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                   const uint64_t* row_ids, const uint64_t* counts, size_t n);

/* The fragment's current TopN candidate order, rank by rank: what rankCache.Top() (cache.go:288)
 * holds after the engine's own import recount, the list of pb200_rank_set, or the default ranking
 * of exact counts. *n receives the length; with row_ids == counts == NULL nothing else is written;
 * PB200_ERR_RANGE when cap is too small. */
int pb200_rank_get(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                   uint64_t* row_ids, uint64_t* counts, size_t cap, size_t* n);

/* Host-only validation of fragment-file bytes with the loader's own parser (snapshot + ops
 * log replay), without a context or a GPU: the checks pb200_frag_load performs before it
 * touches the device. Returns PB200_OK or PB200_ERR_FORMAT; the reference's counterpart is
 * the error path of roaring.go:1562-1653 unmarshalPilosaRoaring. Optional outputs: number of
 * non-empty containers and total cardinality; `err`/`err_cap` receive the message. */
int pb200_host_check_roaring(const void* roaring_bytes, size_t len, uint64_t* n_containers,
                             uint64_t* n_bits, char* err, size_t err_cap);

/* Number of rows / containers / payload bytes resident for a fragment (diagnostics). */
int pb200_frag_stat(pb200_ctx* ctx, pb200_family family, uint64_t shard,
                    uint64_t* n_rows, uint64_t* n_containers, uint64_t* payload_bytes);

/* ---- queries --------------------------------------------------------------------------- */

/* Count(<bitmap call>) over the given shards: executor.go:1793-1825 executeCount with
 * mapFn = executeBitmapCallShard(...).Count() and reduceFn = uint64 add, i.e. the whole
 * executor.go:2576-2611 mapperLocal pass in one call. */
int pb200_count(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes,
                const uint64_t* shards, uint32_t n_shards, uint64_t* out_count);

/* Batch of independent Count queries sharing one shard list. Programs are concatenated
 * in `progs`; query q occupies nodes [prog_off[q], prog_off[q+1]). out_counts[n_queries]. */
int pb200_count_batch(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off,
                      uint32_t n_queries, const uint64_t* shards, uint32_t n_shards,
                      uint64_t* out_counts);

/* Same as pb200_count but also returns per-shard counts (the mapFn results before the
 * reduce), out_per_shard[n_shards]; used by hosts that merge across GPUs themselves. */
int pb200_count_shards(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes,
                       const uint64_t* shards, uint32_t n_shards, uint64_t* out_per_shard);

/* Split form of pb200_count_batch: _begin compiles, uploads and launches asynchronously;
 * _end copies the counts of the OLDEST batch in flight to the host (on a separate copy
 * stream) and waits for them. Two uses:
 *  - pipelining: up to two batches may be in flight per context, so a host calls
 *    _begin(i+1) before _end(i) and its compile/upload work overlaps the kernels of batch i;
 *  - merging across GPUs: the per-query partial counts of the most recently begun batch are
 *    resident in HBM at *device_counts (uint64[n_queries]) where an NCCL all-reduce over
 *    NVLink, enqueued on the context's stream right after _device, sums them in place (the
 *    reduceFn of executor.go:1813 across devices). Work queued on the context's stream
 *    between _device and the next _begin / _end call belongs to that batch: its _end waits
 *    for it too, so the two uses combine (_begin, _device + all-reduce, _begin, _device +
 *    all-reduce, _end, ...). */
int pb200_count_batch_begin(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off,
                            uint32_t n_queries, const uint64_t* shards, uint32_t n_shards);
int pb200_count_batch_device(pb200_ctx* ctx, void** device_counts);
int pb200_count_batch_end(pb200_ctx* ctx, uint64_t* out_counts);

/* Prepared batches: compile + upload a batch of Count programs once (Row / Intersect /
 * Union / Difference / Xor / Not trees), then launch it any number of times with no
 * host<->device traffic; counts stay in HBM until fetched. The serving loop of a host that
 * re-issues the same dashboard queries; also what bench.py times as the device-resident
 * throughput. Launches are asynchronous on the context's stream. */
typedef struct pb200_batch pb200_batch;
int pb200_batch_prepare(pb200_ctx* ctx, const pb200_node* progs, const uint32_t* prog_off,
                        uint32_t n_queries, const uint64_t* shards, uint32_t n_shards,
                        pb200_batch** out);
int pb200_batch_launch(pb200_ctx* ctx, pb200_batch* batch);
int pb200_batch_device_counts(pb200_ctx* ctx, pb200_batch* batch, void** device_counts);
int pb200_batch_fetch(pb200_ctx* ctx, pb200_batch* batch, uint64_t* out_counts);
int pb200_batch_free(pb200_ctx* ctx, pb200_batch* batch);

/* Count(Intersect(Row(family=a), Row(family=b))) for n_pairs row pairs WITHOUT allocating.
 * This entry point has no counterpart in the reference; it exists for hosts that are not
 * allowed to allocate device memory while serving: a host living inside a PyTorch (or RMM)
 * process whose pooling allocator already owns the free HBM — where a raw cudaMalloc can
 * fail or stall every stream — a deployment that gave its whole HBM budget to resident
 * fragments, or a latency-critical query thread that must never block in the driver's
 * allocator. It calls no cudaMalloc / cudaFree / pooled or library allocator at all.
 *
 *  - d_pairs  : DEVICE buffer, uint64[2 * n_pairs] = {a0, b0, a1, b1, ...} row ids
 *  - d_shards : DEVICE buffer, uint64[n_shards] shard numbers
 *  - d_counts : DEVICE buffer, uint64[n_pairs], overwritten with the counts
 *  The caller owns all three (cut from its own pool). Everything else the kernels need —
 *  the expanded programs and the list of array/run container columns handed to the staged
 *  kernel — lives in one fixed array inside the library image, PB200_NOALLOC_SCRATCH_BYTES
 *  per GPU that loads the library, held until process exit.
 *  Limits (PB200_ERR_RANGE beyond them; split the batch): n_pairs <= PB200_NOALLOC_MAX_PAIRS
 *  and n_pairs * n_shards * 16 <= PB200_NOALLOC_MAX_ITEMS.
 *  The call is asynchronous on the context's stream (pb200_set_stream): synchronise that
 *  stream before reading d_counts. The first call on a device looks up the array's device
 *  address (one-time set-up); make that first call before starting any stream capture.
 *  All calls on a device share the array and it keeps the previous call's contents: the
 *  entry point re-zeroes what must start at zero, and calls that overlap on the GPU
 *  (two contexts or streams on the same device) race — issue them on one stream.
 *  Fragment tables must be current: after loading fragments run one ordinary query first
 *  (refreshing the tables allocates, so this entry point refuses to do it). */
#define PB200_NOALLOC_MAX_PAIRS 4096u
#define PB200_NOALLOC_MAX_ITEMS (1u << 20)
#define PB200_NOALLOC_SCRATCH_BYTES ((4u << 20) + (256u << 10))
int pb200_count_pairs_noalloc(pb200_ctx* ctx, pb200_family family, const uint64_t* d_pairs,
                              uint32_t n_pairs, const uint64_t* d_shards, uint32_t n_shards,
                              uint64_t* d_counts);

/* Evaluate a bitmap call on ONE shard and return the resulting row segment as Pilosa
 * roaring bytes (containers optimize()d, keys = shard*16 + i as fragment.go:629
 * OffsetRange produces). Two-call protocol: with out == NULL only *out_len is written.
 * Replaces executor.go:654-680 executeBitmapCallShard for bitmap-returning calls. */
int pb200_eval(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, uint64_t shard,
               void* out, size_t cap, size_t* out_len);

/* The same for ALL local shards of a bitmap-returning call in one launch sequence: what
 * executor.mapperLocal (executor.go:2576) + executeBitmapCallShard (:654) produce before
 * Row.Merge (:595-602), one row segment per shard. The segments are typed (Container.optimize,
 * roaring.go:2289-2338) and serialised ON THE DEVICE; `out` receives the per-shard Pilosa
 * roaring streams packed back to back in shard-list order, stream i at
 * out + out_off[i] with length out_off[i+1] - out_off[i] (out_off: n_shards + 1 entries, may be
 * NULL). A shard without data yields the 8-byte empty stream. With out == NULL only *out_len
 * (and out_off) are written; if cap is too small PB200_ERR_RANGE is returned with the needed
 * size in *out_len. Upper bound for sizing: n_shards * (128 KiB + 536) bytes. */
int pb200_eval_shards(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes,
                      const uint64_t* shards, uint32_t n_shards, void* out, size_t cap,
                      uint64_t* out_off, size_t* out_len);

/* The same result as the protobuf the HTTP / internal client answer carries: the `Columns` field
 * of internal.Row (internal/public.proto: repeated uint64 Columns = 1) exactly as the generated
 * marshaller writes it (internal/public.pb.go:745-766, reached from encodeRow / encodeQueryResponse,
 * encoding/proto/proto.go:408-457): tag 0x0A, varint(payload length), every column id of
 * Row.Columns() (row.go:302-327) as a varint — segment by segment in shard-list order (pass the
 * shards ascending, as Row.Merge keeps them). An empty row yields 0 bytes (the field is omitted).
 * Sized and written ON THE DEVICE over all the local shards of the call; Go appends its Attrs /
 * Keys fields (tags 0x12 / 0x1A) behind these bytes to complete the message. With out == NULL only
 * *out_len is written; PB200_ERR_RANGE with the needed size in *out_len when cap is too small. */
int pb200_eval_proto_columns(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes,
                             const uint64_t* shards, uint32_t n_shards, void* out, size_t cap,
                             size_t* out_len);

/* Context options. No reference counterpart.
 * PB200_OPT_DENSIFY_MIN_BYTES (default 0 = off): fragments loaded or rewritten AFTER the call keep
 * array / run containers whose file payload is at least `value` bytes (<= 8192) as 8 KiB bitmaps
 * in HBM. Results are unchanged; set-op calls over such containers take the streaming bitmap
 * kernel instead of the shared-memory expansion (about 2x on the mixed-container config at
 * value = 1024) at the price of HBM capacity (at most 8192 / value times the payload).
 * pb200_last_timing().algo_bytes keeps counting the reference's encoding. */
#define PB200_OPT_DENSIFY_MIN_BYTES 1u
/* PB200_OPT_RANK_CACHE_SIZE (default 50000 = DefaultCacheSize, field.go:47): entries the rank
 * cache of a fragment keeps (rankCache.maxEntries); applies to rankings derived after the call. */
#define PB200_OPT_RANK_CACHE_SIZE 2u
/* PB200_OPT_TOPN_PRUNE (default 1): TopN with a source row counts, per shard, only the ranked rows fragment.top's walk
 * visits (fragment.go:1673-1688: it stops at the first row whose cached count is below the smallest count among the
 * first n pushed) — the head of the ranking first, the stop position derived on the device, then the rest up to it.
 * 0 counts every ranked row (same results; for measuring both byte figures of SURVEY.md 8d, A_full and A_pruned). */
#define PB200_OPT_TOPN_PRUNE 3u
int pb200_set_option(pb200_ctx* ctx, uint32_t option, uint64_t value);

/* Page-lock a caller-owned host buffer (cudaHostRegister) so that result copies into it — the
 * packed streams of pb200_eval_shards are 128 KiB per dense shard — run as direct DMA at PCIe
 * speed instead of through the driver's pageable staging. Optional: every entry point also works
 * with ordinary memory. A Go host registers its reusable result arena once (the memory must not
 * move: C.malloc or an mmap, not a GC-managed slice). No reference counterpart. */
int pb200_host_register(pb200_ctx* ctx, void* ptr, size_t len);
int pb200_host_unregister(pb200_ctx* ctx, void* ptr);

/* ---- Rows / GroupBy --------------------------------------------- */

/* Rows(field[, previous=][, column=][, limit=]) over the local shards: executeRows /
 * executeRowsShard (executor.go:1277-1441) = fragment.rows (fragment.go:2668-2712) with
 * filterColumn / filterWithLimit (:2606-2629) per shard, merged by RowIDs.merge (:1043-1069).
 * `start` is previous + 1 (0 when absent); `limit` is UINT64_MAX when absent. A row exists in a
 * shard when it has a non-empty container there (what a snapshotted fragment holds). With
 * PB200_ROWS_COLUMN only the shard holding `column` can contribute (:1405-1409). Time-quantum
 * views are merged by the host (one call per view). */
#define PB200_ROWS_COLUMN 1u
int pb200_rows(pb200_ctx* ctx, pb200_family family, uint64_t start, uint32_t flags, uint64_t column,
               uint64_t limit, const uint64_t* shards, uint32_t n_shards, uint64_t* out_rows,
               size_t cap, size_t* n_out);

/* GroupBy(Rows(f0), Rows(f1), ..., filter=, limit=): executeGroupBy's map-reduce
 * (executor.go:1071-1137), executeGroupByShard (:1236-1275), groupByIterator (:3060-3230)
 * and mergeGroupCounts (:1166-1200) for the local shards. Every shard walks the cross
 * product of its rows in ascending (f0, f1, ...) order and keeps its first `limit` groups
 * with a non-zero count; shards are merged in shard-list order (equal groups add up, at
 * most `limit` kept). `offset` and the final limit are the caller's (:1139-1157). The counts
 * |filter & r0 & r1 & ...| of ALL combinations come from the device in one pass (the same kernel
 * as TopN: a prefix row per shard staged in shared memory against the last field's rows); only
 * the iterator's bookkeeping (`previous` seek / wrap rules :3121-3153) runs on the host.
 * child.rows (optional, ascending): the result of a Rows() child that carried limit= or
 * column= and was therefore executed first (:1105-1116) -> filterWithRows (fragment.go:2636).
 * Limits: 1..4 fields; (product of candidate rows) x shards <= 2^27 counts per call.
 * One deviation: with >= 3 fields the reference's nextAtIdx can spin for ever once the first
 * field is exhausted (:3167-3191); this implementation ends the iteration there. */
#define PB200_GROUPBY_MAX_FIELDS 4
typedef struct pb200_groupby_child {
  pb200_family family;   /* standard view of the field                                   */
  uint32_t has_previous; /* Rows(previous=) present                                      */
  uint64_t previous;
  const uint64_t* rows;  /* pre-executed Rows() result, or NULL                          */
  uint64_t n_rows;
} pb200_groupby_child;   /* 32 bytes */
typedef struct pb200_group {
  uint64_t rows[PB200_GROUPBY_MAX_FIELDS]; /* row id per field, unused entries 0         */
  uint64_t count;
} pb200_group;           /* 40 bytes */
int pb200_groupby(pb200_ctx* ctx, const pb200_groupby_child* children, uint32_t n_children,
                  const pb200_node* filter_prog, uint32_t n_filter_nodes, uint64_t limit,
                  const uint64_t* shards, uint32_t n_shards, pb200_group* out, size_t cap,
                  size_t* n_out);

/* Evaluate a bitmap call on one shard into the sorted list of absolute column ids
 * (row.go:302-327 Row.Columns). Two-call protocol as above (cap in elements). */
int pb200_eval_columns(pb200_ctx* ctx, const pb200_node* prog, uint32_t n_nodes, uint64_t shard,
                       uint64_t* out_cols, size_t cap, size_t* out_n);

/* TopN over the ranked rows of `family`: executor.go:863-1000 executeTopN (both passes and
 * the Pairs.Add merge, cache.go:356-374) with fragment.top semantics (fragment.go:1568-1700)
 * per shard. src_prog may be NULL (no Src). ids non-NULL selects the explicit-ids path
 * (N forced 0, fragment.go:1573). min_threshold 0 means defaultMinThreshold=1
 * (executor.go:37,984). Attribute filters stay on the host. On return *n_out pairs are
 * written (sorted by count desc, ties by ascending id); cap is the capacity of `out`.
 * flags & PB200_TOPN_REMOTE: behave like a remote node of a distributed query
 * (execOptions.Remote, executor.go:884-887): run the first pass only and return its merged
 * pairs untrimmed, so that a coordinator (another GPU's host) can merge and issue pass 2. */
#define PB200_TOPN_REMOTE 1u
int pb200_topn(pb200_ctx* ctx, pb200_family family, const pb200_node* src_prog,
               uint32_t n_src_nodes, uint32_t n, const uint64_t* ids, size_t n_ids,
               uint64_t min_threshold, uint64_t tanimoto_threshold, uint32_t flags,
               const uint64_t* shards, uint32_t n_shards, pb200_pair* out, size_t cap,
               size_t* n_out);

/* One shard of TopN, exactly fragment.top (fragment.go:1568-1700): pass-1 style when
 * ids == NULL, explicit-ids style otherwise. Output order = the reference's heap-pop order. */
int pb200_topn_shard(pb200_ctx* ctx, pb200_family family, const pb200_node* src_prog,
                     uint32_t n_src_nodes, uint32_t n, const uint64_t* ids, size_t n_ids,
                     uint64_t min_threshold, uint64_t tanimoto_threshold, uint64_t shard,
                     pb200_pair* out, size_t cap, size_t* n_out);

/* BSI aggregates over a bsig_ view family (view.go:37). filter_prog may be NULL.
 * Sum: fragment.sum (fragment.go:1109-1141) per shard + ValCount.add (executor.go:3002);
 * the caller adds count*Base (executor.go:718). Reproduces the reference quirk that
 * magnitudes of ALL negative columns are subtracted even outside the filter
 * (fragment.go:1118,1134) unless PB200_SUM_STRICT is set in flags. */
#define PB200_SUM_STRICT 1u
int pb200_bsi_sum(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth,
                  const pb200_node* filter_prog, uint32_t n_filter_nodes, uint32_t flags,
                  const uint64_t* shards, uint32_t n_shards, int64_t* out_sum, uint64_t* out_count);

/* Min / Max: fragment.min/max (fragment.go:1145-1225) per shard, reduced with
 * ValCount.smaller / larger (executor.go:3010-3029) in ascending shard-list order
 * ("lowest shard wins ties; counts are not summed" — one legal reference outcome).
 * Values are un-based: the caller adds Base (executor.go:759,797). */
int pb200_bsi_min(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth,
                  const pb200_node* filter_prog, uint32_t n_filter_nodes,
                  const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count);
int pb200_bsi_max(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth,
                  const pb200_node* filter_prog, uint32_t n_filter_nodes,
                  const uint64_t* shards, uint32_t n_shards, int64_t* out_val, uint64_t* out_count);

/* Per-shard (value, count) pairs of Min (is_max=0) or Max (is_max=1) before the reduce. */
int pb200_bsi_minmax_shards(pb200_ctx* ctx, pb200_family bsi_family, uint32_t bit_depth,
                            const pb200_node* filter_prog, uint32_t n_filter_nodes, int is_max,
                            const uint64_t* shards, uint32_t n_shards,
                            int64_t* out_vals, uint64_t* out_counts);

/* ---- several GPUs of one box behind one handle --------------------------------------------
 * The reference does its fan-out and reduce inside ONE call: executor.mapReduce / mapper / mapperLocal
 * (executor.go:2458-2611) push per-shard jobs to the worker pool (:83-107) and fold the partial results
 * with the call's reduce function (uint64 + :1813; ValCount.add / smaller / larger :3002-3029; Pairs.Add
 * cache.go:356-374; Row.Merge executor.go:595-602). pb200_multi is that call over the GPUs of a box:
 * one context per device inside ONE process, shard s resident on devices[s % ndev] (every family of a
 * shard on the same device), and every entry point below takes the caller's WHOLE shard list, runs on the
 * owning devices concurrently and returns the reduced answer — a Go host binds these and needs neither
 * NCCL nor a partition of its own. Semantics are those of the single-device entry point of the same name.
 *   - all-to-all peer access (NVLink) is enabled at init. Every device's fragment table lists all shards,
 *     remote entries pointing into the owner's HBM, so a Count whose operands are small (estimated below
 *     PB200_OPT_SMALL_CALL_BYTES, default 24 MiB, and at most 1024 (query, shard) pairs) is run by ONE device over its peers' memory — one
 *     launch, no fan-out latency; anything larger runs on the owners.
 *   - bitmap-returning calls (pb200_multi_eval_shards / _eval_proto_columns) are evaluated by the owners;
 *     the first device gathers their dense result rows over NVLink, types / serialises them once for the
 *     whole shard list and ONE device -> host copy carries the bytes (identical to one device's output).
 *     They need peer access (PB200_ERR_UNSUPPORTED otherwise; pb200_multi_peer_access tells).
 *   - TopN runs pass 1 on every owner, merges (Pairs.Add), runs pass 2 for the merged candidate ids from
 *     the counts pass 1 left behind, then sorts and trims: executeTopN (executor.go:863-903) as a whole.
 *   - writes (load / apply / import / drop / rank_set) go through the handle too: they wait for queries on
 *     EVERY device before a fragment's previous copy is released and keep the peers' tables current. Do
 *     not call the single-device write entry points on the contexts pb200_multi_ctx returns.
 *   - multi calls on one handle are serialised; the handle may be used from any thread. */
typedef struct pb200_multi pb200_multi;
int pb200_init_multi(const int* devices, int ndev, size_t hbm_budget_bytes_per_device, pb200_multi** out);
int pb200_multi_shutdown(pb200_multi* m);
const char* pb200_multi_last_error(pb200_multi* m);
int pb200_multi_ndev(pb200_multi* m);
int pb200_multi_peer_access(pb200_multi* m);                      /* 1 when all-to-all peer access is on */
int pb200_multi_ctx(pb200_multi* m, int i, pb200_ctx** out);      /* the i-th device's context (statistics, pb200_last_timing) */
int pb200_multi_last_timing(pb200_multi* m, pb200_timing* out);   /* kernel_ms = max over devices, bytes / launches summed */
/* pb200_set_option on every device; PB200_OPT_SMALL_CALL_BYTES is the routing threshold above. */
#define PB200_OPT_SMALL_CALL_BYTES 100u
int pb200_multi_set_option(pb200_multi* m, uint32_t option, uint64_t value);
int pb200_multi_host_register(pb200_multi* m, void* ptr, size_t len);   /* result arena of the first device's copies */
int pb200_multi_host_unregister(pb200_multi* m, void* ptr);
int pb200_multi_family_create(pb200_multi* m, pb200_family* out);  /* the same handle on every device */
int pb200_multi_frag_load(pb200_multi* m, pb200_family family, uint64_t shard, const void* roaring_bytes, size_t len);
int pb200_multi_frag_apply(pb200_multi* m, pb200_family family, uint64_t shard, const uint64_t* set_pos, size_t n_set,
                           const uint64_t* clear_pos, size_t n_clear);
int pb200_multi_frag_import_roaring(pb200_multi* m, pb200_family family, uint64_t shard, const void* data, size_t len,
                                    int clear, int64_t* changed);
int pb200_multi_frag_drop(pb200_multi* m, pb200_family family, uint64_t shard);
int pb200_multi_rank_set(pb200_multi* m, pb200_family family, uint64_t shard, const uint64_t* row_ids,
                         const uint64_t* counts, size_t n);
int pb200_multi_rank_get(pb200_multi* m, pb200_family family, uint64_t shard, uint64_t* row_ids, uint64_t* counts,
                         size_t cap, size_t* n);
int pb200_multi_frag_stat(pb200_multi* m, pb200_family family, uint64_t shard, uint64_t* n_rows, uint64_t* n_containers,
                          uint64_t* payload_bytes);
int pb200_multi_count(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards,
                      uint32_t n_shards, uint64_t* out_count);
int pb200_multi_count_batch(pb200_multi* m, const pb200_node* progs, const uint32_t* prog_off, uint32_t n_queries,
                            const uint64_t* shards, uint32_t n_shards, uint64_t* out_counts);
int pb200_multi_count_shards(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards,
                             uint32_t n_shards, uint64_t* out_per_shard);
int pb200_multi_eval(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, uint64_t shard, void* out, size_t cap,
                     size_t* out_len);
int pb200_multi_eval_columns(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, uint64_t shard,
                             uint64_t* out_cols, size_t cap, size_t* out_n);
int pb200_multi_eval_shards(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards,
                            uint32_t n_shards, void* out, size_t cap, uint64_t* out_off, size_t* out_len);
int pb200_multi_eval_proto_columns(pb200_multi* m, const pb200_node* prog, uint32_t n_nodes, const uint64_t* shards,
                                   uint32_t n_shards, void* out, size_t cap, size_t* out_len);
int pb200_multi_rows(pb200_multi* m, pb200_family family, uint64_t start, uint32_t flags, uint64_t column,
                     uint64_t limit, const uint64_t* shards, uint32_t n_shards, uint64_t* out_rows, size_t cap,
                     size_t* n_out);
int pb200_multi_groupby(pb200_multi* m, const pb200_groupby_child* children, uint32_t n_children,
                        const pb200_node* filter_prog, uint32_t n_filter_nodes, uint64_t limit,
                        const uint64_t* shards, uint32_t n_shards, pb200_group* out, size_t cap, size_t* n_out);
int pb200_multi_topn(pb200_multi* m, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes,
                     uint32_t n, const uint64_t* ids, size_t n_ids, uint64_t min_threshold,
                     uint64_t tanimoto_threshold, uint32_t flags, const uint64_t* shards, uint32_t n_shards,
                     pb200_pair* out, size_t cap, size_t* n_out);
int pb200_multi_topn_shard(pb200_multi* m, pb200_family family, const pb200_node* src_prog, uint32_t n_src_nodes,
                           uint32_t n, const uint64_t* ids, size_t n_ids, uint64_t min_threshold,
                           uint64_t tanimoto_threshold, uint64_t shard, pb200_pair* out, size_t cap, size_t* n_out);
int pb200_multi_bsi_sum(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog,
                        uint32_t n_filter_nodes, uint32_t flags, const uint64_t* shards, uint32_t n_shards,
                        int64_t* out_sum, uint64_t* out_count);
int pb200_multi_bsi_min(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog,
                        uint32_t n_filter_nodes, const uint64_t* shards, uint32_t n_shards, int64_t* out_val,
                        uint64_t* out_count);
int pb200_multi_bsi_max(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth, const pb200_node* filter_prog,
                        uint32_t n_filter_nodes, const uint64_t* shards, uint32_t n_shards, int64_t* out_val,
                        uint64_t* out_count);
int pb200_multi_bsi_minmax_shards(pb200_multi* m, pb200_family bsi_family, uint32_t bit_depth,
                                  const pb200_node* filter_prog, uint32_t n_filter_nodes, int is_max,
                                  const uint64_t* shards, uint32_t n_shards, int64_t* out_vals, uint64_t* out_counts);

/* ---- one process per GPU: the cross-GPU sum of a Count batch without a collective library --------------
 * When every GPU of the box is driven by its own process (one rank per device, e.g. under torchrun), the reduce
 * function of executeCount (uint64 +, executor.go:1813) over the ranks is done by the batch's own kernel: the last CTA
 * to retire adds this GPU's partial counts into a mailbox of every rank with system-scope reductions over NVLink peer
 * memory (the peers' mailboxes are mapped through CUDA IPC; plain stores, no multicast), so no collective is launched
 * between two batches and no rank waits for another inside the launch sequence. A batch's total is picked up two
 * batches later (in stream order) or when the host fetches it.
 *   1. every rank: pb200_peer_export -> a PB200_PEER_HANDLE_BYTES handle; exchange the handles (any transport);
 *   2. every rank: pb200_peer_connect(rank, world, handles[world]); world <= 8, batches <= PB200_PEER_MAX_COUNTS queries;
 *   3. pb200_batch_launch_allreduce instead of pb200_batch_launch: after pb200_batch_fetch (or pb200_peer_flush + a
 *      stream synchronisation) the batch's device counts hold the sum over all ranks. Every rank must launch the same
 *      sequence of all-reduced batches. With PB200_OPT_PEER_ALLREDUCE = 1 pb200_count_batch_begin / _end do the same.
 * A rank that waits ~2 s for a peer's counts gives up: the fetch returns PB200_ERR_CUDA instead of hanging.
 * No reference counterpart (the reference reduces on the host, executor.go:2458-2574). */
#define PB200_PEER_HANDLE_BYTES 64
#define PB200_PEER_MAX_COUNTS 4096u
#define PB200_OPT_PEER_ALLREDUCE 4u
int pb200_peer_export(pb200_ctx* ctx, void* handle_out);
int pb200_peer_connect(pb200_ctx* ctx, int rank, int world, const void* handles);
int pb200_peer_flush(pb200_ctx* ctx);   /* queue the pick-up of every batch still in flight on the context's stream */
int pb200_batch_launch_allreduce(pb200_ctx* ctx, pb200_batch* batch);

/* Size the context's scratch (device and page-locked) once for the largest Count batch the host will issue:
 * max_queries programs of up to max_nodes_per_query nodes over max_shards shards. Without it the scratch grows on
 * demand — a cudaMalloc plus a cudaFree of the smaller buffer, which stalls the whole device — on the first calls of
 * each new size; after it no pb200_count / _count_batch / _count_batch_begin call within those bounds allocates.
 * Programs with BSI range / Shift leaves materialise temporaries sized by the call and are outside the reservation.
 * Also brings the device-side fragment tables up to date. No reference counterpart (Go allocates per call). */
int pb200_reserve(pb200_ctx* ctx, uint32_t max_queries, uint32_t max_shards, uint32_t max_nodes_per_query);

/* Timing / byte accounting of the last query call on this context. */
int pb200_last_timing(pb200_ctx* ctx, pb200_timing* out);

/* Statistics of the combining queue behind pb200_count: how many batched launches served how many single-query calls
 * since pb200_init (calls that ran alone are not counted). queries / launches = the mean width of a combined launch. */
int pb200_combine_stats(pb200_ctx* ctx, uint64_t* launches, uint64_t* queries);

#ifdef __cplusplus
}
#endif
#endif /* PILOSA_B200_H */
