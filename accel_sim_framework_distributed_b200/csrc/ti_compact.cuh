// Compact results (include/traceingest.h: ti_view_compact): the pass behind a shard's decode that
// gathers what still has to travel to the host.
//
// In compact mode the fused kernel writes one 16 B descriptor per memory instruction, leaves out the
// 256 B rows of base+stride instructions and the transactions of dense ones, and keeps everything
// else where the full layout has it (rows of list / base+delta instructions at their memory ordinal,
// transactions at their CSR offset).  Three small kernels then make that dense again:
//   cp_count   per block of kCpBlock instructions: rows to keep, transactions to keep
//   cp_scan    exclusive scan of the block sums (one CTA)
//   cp_gather  rows -> rows_c[ordinal], transactions -> txn_c[offset], CSR of the kept transactions,
//              descriptor.base = row ordinal for the kept rows
// Traffic: the descriptors and CSR offsets once, plus whatever is kept (nothing for a base+stride
// corpus such as BASELINE config 2).
#ifndef TI_COMPACT_CUH
#define TI_COMPACT_CUH

#include "ti_device.cuh"

namespace ti {

constexpr int kCpBlock = 256;   // memory instructions per CTA, one per thread

struct CompactParams {
  ti_mem_desc *desc;            // [n_mem] in: as the decode left them; out: base = row ordinal for full rows
  const uint32_t *txn_off;      // [n_mem + 1] full CSR (values carry off_txn_full)
  const uint64_t *rows;         // full layout
  const ti_txn_rec *txn;        // full layout
  const unsigned long long *counters;   // the decode's counter block: n_mem is read on the device (the host does not know it yet)
  uint64_t cap_mem;
  uint32_t off_txn_full;        // what txn_off[0] holds (the decode's DecodeParams::off_txn)
  uint64_t *rows_c;             // [cap_rows_c][32]
  ti_txn_rec *txn_c;            // [cap_txn_c]
  uint32_t *txn_off_c;          // [n_mem + 1]
  uint64_t cap_rows_c, cap_txn_c;
  uint32_t *part;               // [2 * n_blocks + 2] block sums, then their exclusive scan; the last pair = totals
  uint32_t n_blocks;
  uint64_t off_full, off_txn_c; // file-global ordinals / offsets of what the preceding shards kept
};

TI_D uint64_t cp_n_mem(const CompactParams &C) { const uint64_t n = C.counters[kCntNInst + 1]; return n < C.cap_mem ? n : C.cap_mem; }

TI_D void cp_item(const CompactParams &C, uint64_t r, uint32_t &full, uint32_t &keep) {
  full = keep = 0;
  if (r >= cp_n_mem(C)) return;
  const uint32_t meta = C.desc[r].meta;
  full = (meta & TI_DESC_FULL_ROW) ? 1u : 0u;
  keep = (meta & TI_DESC_DENSE) ? 0u : C.txn_off[r + 1] - C.txn_off[r];
}

// block-wide exclusive scan of two small values per thread; returns the totals through tot*
TI_D void cp_block_scan(uint32_t a, uint32_t b, uint32_t &ea, uint32_t &eb, uint32_t &tota, uint32_t &totb, uint32_t *sh /* [2 * kCpBlock / 32] */) {
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t ia = a, ib = b;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t ua = __shfl_up_sync(0xffffffffu, ia, d), ub = __shfl_up_sync(0xffffffffu, ib, d);
    if ((int)lane >= d) { ia += ua; ib += ub; }
  }
  if (lane == 31) { sh[2 * w] = ia; sh[2 * w + 1] = ib; }
  __syncthreads();
  uint32_t ba = 0, bb = 0;
  tota = totb = 0;
  for (unsigned k = 0; k < (unsigned)kCpBlock / 32; ++k) {
    if (k < w) { ba += sh[2 * k]; bb += sh[2 * k + 1]; }
    tota += sh[2 * k]; totb += sh[2 * k + 1];
  }
  ea = ba + ia - a; eb = bb + ib - b;
  __syncthreads();
}

__global__ void __launch_bounds__(kCpBlock) cp_count_kernel(const CompactParams C) {
#ifdef TI_EMU
  uint32_t *sh = reinterpret_cast<uint32_t *>(emu::cur_smem());
#else
  __shared__ uint32_t sh[2 * kCpBlock / 32];
#endif
  uint32_t full, keep, ea, eb, ta, tb;
  cp_item(C, (uint64_t)blockIdx.x * kCpBlock + threadIdx.x, full, keep);
  cp_block_scan(full, keep, ea, eb, ta, tb, sh);
  if (threadIdx.x == 0) { C.part[2 * blockIdx.x] = ta; C.part[2 * blockIdx.x + 1] = tb; }
}

__global__ void __launch_bounds__(kCpBlock) cp_scan_kernel(const CompactParams C) {
#ifdef TI_EMU
  uint32_t *sh = reinterpret_cast<uint32_t *>(emu::cur_smem());
#else
  __shared__ uint32_t sh[2 * kCpBlock / 32];
#endif
  uint32_t ra = 0, rb = 0;   // running totals (uniform)
  for (uint32_t b0 = 0; b0 < C.n_blocks; b0 += kCpBlock) {
    const uint32_t b = b0 + threadIdx.x;
    const uint32_t a = b < C.n_blocks ? C.part[2 * b] : 0u, k = b < C.n_blocks ? C.part[2 * b + 1] : 0u;
    uint32_t ea, eb, ta, tb;
    cp_block_scan(a, k, ea, eb, ta, tb, sh);
    if (b < C.n_blocks) { C.part[2 * b] = ra + ea; C.part[2 * b + 1] = rb + eb; }
    ra += ta; rb += tb;
  }
  if (threadIdx.x == 0) {
    C.part[2 * C.n_blocks] = ra; C.part[2 * C.n_blocks + 1] = rb;
    C.txn_off_c[cp_n_mem(C)] = (uint32_t)(C.off_txn_c + rb);   // CSR sentinel
  }
}

__global__ void __launch_bounds__(kCpBlock) cp_gather_kernel(const CompactParams C) {
#ifdef TI_EMU
  uint32_t *sh = reinterpret_cast<uint32_t *>(emu::cur_smem());
#else
  __shared__ uint32_t sh[2 * kCpBlock / 32 + 3 * kCpBlock];
#endif
  uint32_t *s_ord = sh + 2 * kCpBlock / 32, *s_src = s_ord + kCpBlock, *s_dst = s_src + kCpBlock;   // per item: row ordinal or ~0 | txn source | txn destination
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint64_t r = (uint64_t)blockIdx.x * kCpBlock + threadIdx.x;
  uint32_t full, keep, ea, eb, ta, tb;
  cp_item(C, r, full, keep);
  cp_block_scan(full, keep, ea, eb, ta, tb, sh);
  const uint32_t ord = C.part[2 * blockIdx.x] + ea, coff = C.part[2 * blockIdx.x + 1] + eb;
  uint32_t n_keep = 0;
  const uint64_t n_mem = cp_n_mem(C);
  if (r < n_mem) {
    C.txn_off_c[r] = (uint32_t)(C.off_txn_c + coff);
    if (full) C.desc[r].base = C.off_full + ord;
    n_keep = keep;
  }
  s_ord[threadIdx.x] = full ? ord : 0xffffffffu;
  s_src[threadIdx.x] = r < n_mem ? C.txn_off[r] - C.off_txn_full : 0u;
  s_dst[threadIdx.x] = coff;
  __syncthreads();
  // rows: a warp per kept row, 8 bytes per lane
  for (unsigned i = w; i < (unsigned)kCpBlock; i += kCpBlock / 32) {
    const uint32_t o = s_ord[i];
    const uint64_t src = (uint64_t)blockIdx.x * kCpBlock + i;
    if (o != 0xffffffffu && o < C.cap_rows_c) C.rows_c[(uint64_t)o * 32 + lane] = C.rows[src * 32 + lane];
  }
  // transactions: the owning thread's count is broadcast inside its warp; two lanes per 32 B record
  for (unsigned i = 0; i < 32; ++i) {
    const uint32_t n = __shfl_sync(0xffffffffu, n_keep, i);
    if (!n) continue;
    const unsigned it = w * 32 + i;
    const uint4 *src = reinterpret_cast<const uint4 *>(C.txn + s_src[it]);
    uint4 *dst = reinterpret_cast<uint4 *>(C.txn_c + s_dst[it]);
    for (uint32_t k = lane; k < 2 * n; k += 32)
      if ((uint64_t)s_dst[it] + (k >> 1) < C.cap_txn_c) dst[k] = src[k];
  }
}

}  // namespace ti
#endif
