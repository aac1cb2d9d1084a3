// Shared definitions of the fused trace-ingest kernel: launch parameters, the
// per-chunk shared-memory plan, the ISA hash and small warp helpers.
//
// Data layout in HBM (all arrays in file order, see include/traceingest.h):
//   text      : the shard, bytes, read exactly once, 16 B vectors
//   inst      : ti_inst_rec[n_inst]   32 B each, written once, coalesced from smem
//   addr_rows : u64[n_mem][32]        256 B rows, one coalesced 256 B store per warp
//   txn_off   : u32[n_mem+1]          CSR offsets
//   txn       : ti_txn_rec[n_txn]     32 B each
//   tb / warp : per "thread block =" / "warp =" header
//   chains    : u64[4][n_chunks+1]    decoupled look-back words (scratch)
//
// The shard is cut into fixed CHUNK-byte pieces; a chunk owns every line that
// starts inside it.  A persistent grid takes chunks in file order from an atomic
// ticket, decodes a chunk entirely in shared memory and obtains its global output
// offsets from a single-pass chained scan over the chunk descriptors, so every
// output array is dense and in file order after ONE pass over the text.
#ifndef TI_DEVICE_CUH
#define TI_DEVICE_CUH

#include <stdint.h>

#include "../../include/traceingest.h"

#if defined(__CUDACC__) || defined(TI_EMU)
#define TI_HD __host__ __device__ __forceinline__
#define TI_D __device__ __forceinline__
// rarely taken, register-hungry paths are kept out of line so that they do not set the
// register budget of the whole kernel
#ifdef TI_EMU
#define TI_D_COLD inline
#else
#define TI_D_COLD __device__ __noinline__
#endif
#else
#define TI_HD inline
#define TI_D inline
#define TI_D_COLD inline
#endif

namespace ti {

// ------------------------------------------------------------------ tunables
#ifndef TI_CHUNK
#define TI_CHUNK 18432   // = warps per CTA x 1536: a warp owns 1536 bytes of the chunk (about one line per lane)
#endif
#ifndef TI_THREADS
#define TI_THREADS 384
#endif
#ifndef TI_TXN_STAGE
#define TI_TXN_STAGE 608
#endif
#ifndef TI_TEXT_BUFS
#define TI_TEXT_BUFS 1     // text buffers; the text is dead after P6b, so one buffer takes the prefetch of the next chunk too
#endif
#ifndef TI_TMA_STORE
#define TI_TMA_STORE 1     // P8: the chunk's record block leaves shared memory as one TMA bulk store
#endif
#ifndef TI_MIN_BLOCKS
#define TI_MIN_BLOCKS 2   // resident CTAs per SM the register allocation is bounded for
#endif
constexpr int kChunk = TI_CHUNK;        // bytes of text owned by one chunk
constexpr int kThreads = TI_THREADS;    // threads per CTA
constexpr int kWarps = kThreads / 32;
constexpr int kSlack = TI_MAX_LINE;     // bytes loaded past the chunk so its last line is complete
constexpr int kTextBytes = 16 + kChunk + kSlack + 48;  // 16 B before the chunk, 48 B of '\n' sentinel after
constexpr int kRecCap = kChunk / 24;    // no tracer line is shorter than 24 bytes with its '\n'
constexpr int kSlCap = kChunk / 8 + 1;  // "#END_TB\n": no valid non-blank line is shorter than 8
constexpr int kTxnStage = TI_TXN_STAGE; // transactions staged in smem per chunk before look-back #2
constexpr int kFlushEvery = 64;         // chunks between counter flushes

// -------------------------------------------------------------- enum values
// (abstract_hardware_model.h:47-62,108-167,787-801; restated as plain constants)
enum : unsigned { OP_ALU = 1, OP_SFU, OP_TENSOR, OP_DP, OP_SP, OP_INTP, OP_ALU_SFU, OP_LOAD, OP_TLOAD,
                  OP_TSTORE, OP_STORE, OP_BRANCH, OP_BARRIER, OP_MEMBAR, OP_CALL, OP_RET, OP_EXIT,
                  OP_SPEC0 = 100 };
enum : unsigned { SP_OTHER, SP_INT, SP_IMUL24, SP_IMUL32, SP_IMUL, SP_IDIV, SP_FPMUL, SP_FPDIV, SP_FP,
                  SP_SQRT, SP_LOG, SP_SIN, SP_EXP, SP_DPMUL, SP_DPDIV, SP_DP, SP_TENSOR, SP_TEX };
enum : unsigned { SPACE_UNDEF = 0, SPACE_LOCAL = 2, SPACE_SHARED = 3, SPACE_SSTARR = 4, SPACE_CONST = 8,
                  SPACE_GLOBAL = 11 };
enum : unsigned { CACHE_UNDEF = 0, CACHE_ALL = 1, CACHE_GLOBAL = 6 };

// ------------------------------------------------------------------ ISA hash
struct IsaSlot {      // 16 bytes; id == 0 marks an empty slot
  uint64_t lo;        // mnemonic bytes 0..7, little endian, zero padded
  uint32_t hi;        // mnemonic bytes 8..11
  uint8_t id, cat;
  uint16_t pad;
};
constexpr int kIsaSlots = 512;
TI_HD uint32_t isa_hash(uint64_t lo, uint32_t hi) {
  uint64_t x = lo * 0x9E3779B97F4A7C15ull ^ (uint64_t)hi * 0xC2B2AE3D27D4EB4Full;
  x ^= x >> 29;
  return (uint32_t)(x >> 17) & (kIsaSlots - 1);
}

// -------------------------------------------------------- look-back chains
// Four chained scans run over the chunks, each a u64 array indexed by chunk, zeroed before
// launch.  A word carries its own state next to the value, so a reader needs no fence: it
// either sees a complete word or keeps polling (single-copy atomic 8-byte loads/stores).
//   bit 63: the payload is this chunk's INCLUSIVE prefix   bit 62: it is this chunk's aggregate
//   payload: two 31-bit fields, f0 | f1 << 31
//   chain 0: instructions | thread-block headers     (published after P1, resolved before the records leave)
//   chain 1: warp headers                            (same)
//   chain 2: memory instructions                     (published after P3)
//   chain 3: transactions                            (published after P7)
// Chains 2 and 3 normally resolve one iteration later, when the chunk's address rows and
// transactions leave (ti_kernel_main.cuh: finish); a chunk with warp-path instructions needs
// the row index earlier and resolves chain 2 in its own iteration as well.
constexpr int kChains = 4;
constexpr unsigned long long kValid = 1ull << 63, kAggValid = 1ull << 62;
constexpr size_t chain_bytes(size_t n_chunks) { return (size_t)kChains * 8 * (n_chunks + 1); }

struct Globals {        // zeroed before launch (first_error_off: all ones)
  uint32_t ticket;      // next chunk to hand out
  uint32_t error_flags; // TI_EF_*
  uint32_t first_error_off;
  uint32_t overflow;    // bit0 inst, bit1 mem, bit2 txn, bit3 tb, bit4 warp capacity exceeded
  uint32_t rare[4];     // how often the rare paths ran: [0] staging overflow (recompute), [1] headers not parked (early
                        // resolve), [2] chunks that resolved the row index in their own iteration (warp-path
                        // instructions), [3] descriptor list full
};

struct DecodeParams {
  const uint8_t *text;  // 16-byte aligned
  uint64_t len;
  uint32_t n_chunks;
  uint32_t trace_version;
  uint64_t shmem_base, local_base;
  // outputs + capacities (elements)
  ti_inst_rec *inst;  uint64_t cap_inst;
  uint64_t *rows;     uint64_t cap_mem;
  uint32_t *txn_off;
  ti_txn_rec *txn;    uint64_t cap_txn;
  ti_mem_desc *desc;  // compact mode (else null): one 16 B descriptor per memory instruction [cap_mem]; rows of base+stride
                      // instructions and transactions of dense ones are then NOT written (they are closed forms of it)
  ti_tb_rec *tb;      uint64_t cap_tb;
  ti_warp_rec *warp;  uint64_t cap_warp;
  // value offsets: what the preceding shards of the same kernel file produced, so that
  // indices and text offsets stored in the records of a streamed shard are file-global
  uint64_t off_inst, off_mem, off_txn, off_warp, off_text;
  // scratch
  unsigned long long *chain[kChains];   // each [n_chunks + 1], see "look-back chains"
  Globals *g;
  unsigned long long *counters;   // ti_counters viewed as u64[]
  // tables + model parameters
  const IsaSlot *isa;             // kIsaSlots entries for this kernel's architecture
  const uint8_t *spop;            // special_ops per opcode id (256 entries)
  uint16_t lat[32][2];            // <latency, initiation> per folded op_type
  int32_t coalesce_arch, mem_warp_parts, num_shmem_bank, shmem_limited_broadcast, gmem_skip_L1D;
};

// folded op_type index (1..17 -> same, 100..107 -> 18..25)
TI_HD unsigned fold_op(unsigned op) { return op >= OP_SPEC0 ? 18 + (op - OP_SPEC0) : op; }

// layout of ti_counters as u64 words
constexpr int kCntHist = 0;                         // [18][33]
constexpr int kCntOp = kCntHist + 18 * 33;          // [32]
constexpr int kCntSpace = kCntOp + 32;              // [16][2]
constexpr int kCntConst = kCntSpace + 32;
constexpr int kCntNInst = kCntConst + 1;            // n_inst, n_mem, n_txn, n_tb, n_warp, n_begin, n_end, text_bytes, smem_cycles
constexpr int kCntWords = kCntNInst + 9 + 6;
static_assert(kCntWords * 8 == sizeof(ti_counters), "ti_counters layout");
// smem mirror: the histogram part plus a few totals, u32
constexpr int kSmCnt = kCntConst + 1 + 4;           // + n_begin, n_end, smem_cycles, "insts =" lines

}  // namespace ti
#endif
