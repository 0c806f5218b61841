// Device-side layout of a fragment ("device container table") and of compiled query
// programs. Shared by host code and kernels. See DESIGN.md §"Data layout in HBM".
#pragma once
#include <stdint.h>

namespace pb200 {

constexpr int kRowContainers = 16;      // 2^20 / 2^16 (fragment.go:63)
constexpr int kBitmapWords = 1024;      // roaring.go:55 bitmapN
constexpr int kBitmapBytes = 8192;
constexpr uint32_t kFullN = 65536;

enum CType : uint32_t { kNil = 0, kArray = 1, kBitmap = 2, kRun = 3 };  // roaring.go:60-65

// One roaring container resident in HBM (16 B). Replaces roaring.Container
// (container_stash.go:39-46): the payload pointer, cardinality and type tag.
struct __attribute__((aligned(16))) ContainerDesc {
  uint64_t addr;  // device address of the payload: array = n x u16 sorted; bitmap = 1024 x u64;
                  // run = len x {u16 start, u16 last}. 16 B aligned (bitmaps 128 B aligned).
  uint32_t n;     // cardinality, 0..65536 (0 = nil container)
  uint32_t meta;  // bits 0-1: CType; bits 2-14: len (array: n, run: number of runs, bitmap: 1024);
                  // bit 31: stored as a bitmap although the file held an array / run (load-time densify
                  // option), bits 15-30 then keep the file encoding's payload bytes for the byte accounting
};
static_assert(sizeof(ContainerDesc) == 16, "ContainerDesc must be 16 B");

__host__ __device__ inline uint32_t desc_type(const ContainerDesc& d) { return d.meta & 3u; }
__host__ __device__ inline uint32_t desc_len(const ContainerDesc& d) { return (d.meta >> 2) & 0x1fffu; }
__host__ __device__ inline uint32_t make_meta(uint32_t type, uint32_t len) { return type | (len << 2); }
__host__ __device__ inline bool desc_densified(const ContainerDesc& d) { return (d.meta >> 31) != 0; }
__host__ __device__ inline uint32_t desc_file_bytes(const ContainerDesc& d) { return (d.meta >> 15) & 0xffffu; }
__host__ __device__ inline uint32_t make_meta_densified(uint32_t file_payload_bytes) {
  return make_meta(2u /* kBitmap */, 1024u) | (file_payload_bytes << 15) | 0x80000000u;
}

// One fragment = all rows of one (field, view) for one shard (fragment.go:100-140).
// rows[i*16 + c] is container c of the i-th row; replaces the B+tree keyed by
// row*16 + c (containers_btree.go) and Bitmap.OffsetRange (roaring.go:535-558).
struct __attribute__((aligned(32))) FragDev {
  const uint64_t* row_ids;    // ascending row ids (nrows), unused when dense
  const ContainerDesc* rows;  // [nrows][16]
  uint32_t nrows;
  uint32_t dense;             // 1: rows is indexed directly by row id (ids 0..nrows-1)
  uint64_t pad;
};
static_assert(sizeof(FragDev) == 32, "FragDev must be 32 B");

// One family = (index, field, view): fragments indexed by shard number.
struct FamilyDev {
  const FragDev* frags;  // [cap]
  uint32_t cap;
  uint32_t pad;
};

// ---- compiled programs ------------------------------------------------------------
// The post-order pb200_node tree is compiled on the host into an accumulator machine:
// LOAD pushes the live accumulator (if any) and loads a leaf; *_LEAF combine the
// accumulator with a leaf; *_STACK pop the left operand and combine it with the
// accumulator (only bushy trees use the stack). N-ary PQL ops are left folds
// (executor.go:1668-1733,1003-1025).
enum IOp : uint8_t {
  I_LOAD = 0,
  I_AND_LEAF, I_OR_LEAF, I_XOR_LEAF, I_ANDNOT_LEAF,  // acc = acc op leaf
  I_RANDNOT_LEAF,                                    // acc = leaf & ~acc   (Not(): exists \ x)
  I_AND_STACK, I_OR_STACK, I_XOR_STACK, I_ANDNOT_STACK,  // acc = pop op acc (pop is the LEFT operand)
};

struct Instr {
  uint8_t op;
  uint8_t leaf;  // index into the query's leaf table
};

enum LeafKind : uint32_t { L_ROW = 0, L_ZERO = 1, L_TEMP = 2 };

struct __attribute__((aligned(16))) Leaf {
  uint32_t kind;    // LeafKind
  uint32_t family;  // L_ROW: family handle; L_TEMP: temp slot
  uint64_t row;     // L_ROW: row id
};

struct Prog {
  uint32_t instr_off, n_instr;  // into the batch's Instr array
  uint32_t leaf_off, n_leaves;  // into the batch's Leaf array
};

constexpr int kMaxLeaves = 32;   // per query program
constexpr int kMaxStack = 4;     // bushy-tree stack depth
constexpr int kMaxInstr = 64;

}  // namespace pb200
