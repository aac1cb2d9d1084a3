// Writer side, second half: the tracer's address compressors on the device -- the inverse
// of the decode kernel's address phase.  For every memory instruction (32 addresses + active mask) it
// makes the tracer's choice of format and writes the text the tracer would print between the memory
// width and the immediate (util/tracer_nvbit/tracer_tool/tracer_tool.cu:717-750):
//     base+stride  "1 0x%llx %d "          when base_stride_compress succeeds (:599-636: ONE contiguous run of
//                                           at least two active lanes with a constant int difference, or no lane)
//     base+delta   "2 0x%llx " + "%lld "*   otherwise (:638-654)
//     list         "0 " + "0x%016lx "*      when compression is off
// A warp per instruction, lane = trace lane: format decision by ballots, every lane formats its own token,
// a shuffle scan places them.  Two passes (lengths, exclusive scan, text) so that the output is dense.
#ifndef TI_COMPRESS_CUH
#define TI_COMPRESS_CUH

#include "ti_device.cuh"

namespace ti {

struct CompressParams {
  const uint64_t *rows;     // [n][32]
  const uint32_t *masks;    // [n]
  uint64_t n;
  int enable_compress;
  uint32_t *len;            // [n] pass 1
  const uint64_t *off;      // [n + 1] pass 2
  char *out;
  uint64_t cap;
};

TI_D unsigned fmt_hex(uint64_t v, char *b, bool pad16) {   // "%llx" / "%016lx"
  unsigned nd = 1;
  for (uint64_t t = v >> 4; t; t >>= 4) ++nd;
  if (pad16) nd = 16;
  for (unsigned i = 0; i < nd; ++i) { const unsigned d = (unsigned)(v >> (4 * (nd - 1 - i))) & 15; b[i] = (char)(d < 10 ? '0' + d : 'a' + d - 10); }
  return nd;
}
TI_D unsigned fmt_dec(long long v, char *b) {              // "%lld" / "%d"
  unsigned long long m = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
  char t[20];
  unsigned nd = 0;
  do { t[nd++] = (char)('0' + m % 10); m /= 10; } while (m);
  unsigned k = 0;
  if (v < 0) b[k++] = '-';
  for (unsigned i = 0; i < nd; ++i) b[k++] = t[nd - 1 - i];
  return k;
}

// one instruction per warp (all 32 lanes call it); returns the length of its text and writes it when out != nullptr
TI_D unsigned compress_row(const CompressParams &C, uint64_t r, char *out, uint64_t room) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t m = C.masks[r];
  const uint64_t a = C.rows[r * 32 + lane];
  const unsigned first = m ? (unsigned)(__ffs(m) - 1) : 0u;
  const uint64_t prev = __shfl_up_sync(0xffffffffu, a, 1);
  const uint64_t base = m ? __shfl_sync(0xffffffffu, a, first) : 0;     // `base_addr` stays 0 when no lane is active
  const uint64_t second = __shfl_sync(0xffffffffu, a, (first + 1) & 31);
  const uint32_t below = m & ((1u << lane) - 1u);
  const uint64_t prev_active = __shfl_sync(0xffffffffu, a, below ? 31 - __clz((int)below) : 0);
  // ---- base_stride_compress (tracer_tool.cu:599-636)
  bool stride_ok = true;                                                // an empty mask leaves const_stride true
  int stride = 0;
  {
    const uint32_t inv = m ? ~(m >> first) : 0u;
    const unsigned run = !m ? 0u : inv ? (unsigned)(__ffs((int)inv) - 1) : 32u;    // first contiguous run of active lanes
    const uint32_t runmask = (run >= 32 ? 0xffffffffu : ((1u << run) - 1u)) << first;
    if (run >= 2) stride = (int)(second - base);                        // `int stride = addrs[s + 1] - addrs[s]`
    const bool in_run = m && lane > first && lane < first + run;
    const bool bad = in_run && (a - prev) != (uint64_t)(long long)stride;   // int compared with an unsigned long difference
    const bool any_bad = __any_sync(0xffffffffu, bad);
    if (m) stride_ok = run >= 2 && !(m & ~runmask) && !any_bad;
  }
  const bool act = m >> lane & 1;
  const unsigned mode = !C.enable_compress ? 0u : stride_ok ? 1u : 2u;  // address_format: list_all, base_stride, base_delta
  char tok[48];
  unsigned tl = 0;
  if (mode == 1) {
    if (lane == first) {                                                // "%u 0x%llx %d "
      tok[0] = '1'; tok[1] = ' '; tok[2] = '0'; tok[3] = 'x';
      tl = 4 + fmt_hex(base, tok + 4, false);
      tok[tl++] = ' ';
      tl += fmt_dec((long long)stride, tok + tl);
      tok[tl++] = ' ';
    }
  } else if (mode == 2) {
    if (lane == first) {                                                // "%u 0x%llx "
      tok[0] = '2'; tok[1] = ' '; tok[2] = '0'; tok[3] = 'x';
      tl = 4 + fmt_hex(base, tok + 4, false);
      tok[tl++] = ' ';
    } else if (act) {                                                   // base_delta_compress (:638-654): "%lld " to the previous active lane
      tl = fmt_dec((long long)(a - prev_active), tok);
      tok[tl++] = ' ';
    }
  } else {
    if (lane == 0) { tok[0] = '0'; tok[1] = ' '; tl = 2; }              // "%u "
    if (act) { tok[tl] = '0'; tok[tl + 1] = 'x'; tl += 2; tl += fmt_hex(a, tok + tl, true); tok[tl++] = ' '; }   // "0x%016lx "
  }
  unsigned inc = tl;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(0xffffffffu, inc, d); if ((int)lane >= d) inc += u; }
  const unsigned total = __shfl_sync(0xffffffffu, inc, 31);
  if (out) {
    const unsigned at = inc - tl;
    for (unsigned i = 0; i < tl; ++i)
      if ((uint64_t)at + i < room) out[at + i] = tok[i];
  }
  return total;
}

__global__ void __launch_bounds__(256) cmp_len_kernel(const CompressParams C) {
  const uint64_t w0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t r = w0; r < C.n; r += nw) {
    const unsigned n = compress_row(C, r, nullptr, 0);
    if ((threadIdx.x & 31) == 0) C.len[r] = n;
  }
}

__global__ void __launch_bounds__(256) cmp_write_kernel(const CompressParams C) {
  const uint64_t w0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t r = w0; r < C.n; r += nw) {
    const uint64_t at = C.off[r];
    compress_row(C, r, C.out + at, at < C.cap ? C.cap - at : 0);
  }
}

// exclusive scan of n u32 lengths into u64 offsets (off[n] = total): block sums, a one-CTA scan of them, the final pass
constexpr int kScanBlock = 1024;
__global__ void __launch_bounds__(256) scan_sum_kernel(const uint32_t *len, uint64_t n, unsigned long long *bsum) {
  unsigned long long s = 0;
  const uint64_t lo = (uint64_t)blockIdx.x * kScanBlock;
  for (uint64_t i = lo + threadIdx.x; i < lo + kScanBlock && i < n; i += blockDim.x) s += len[i];
#pragma unroll
  for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(bsum + blockIdx.x, s);
}
__global__ void scan_blocks_kernel(unsigned long long *bsum, uint64_t n_blocks) {   // one thread: the list is short
  if (threadIdx.x || blockIdx.x) return;
  unsigned long long run = 0;
  for (uint64_t b = 0; b <= n_blocks; ++b) { const unsigned long long v = b < n_blocks ? bsum[b] : 0; bsum[b] = run; run += v; }
}
__global__ void __launch_bounds__(32) scan_final_kernel(const uint32_t *len, uint64_t n, const unsigned long long *bsum, uint64_t *off, uint64_t n_blocks) {
  const unsigned lane = threadIdx.x;
  const uint64_t lo = (uint64_t)blockIdx.x * kScanBlock;
  unsigned long long run = bsum[blockIdx.x];
  for (uint64_t i0 = lo; i0 < lo + kScanBlock && i0 < n; i0 += 32) {
    const uint64_t i = i0 + lane;
    const unsigned long long v = i < n ? len[i] : 0;
    unsigned long long inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned long long u = __shfl_up_sync(0xffffffffu, inc, d); if ((int)lane >= d) inc += u; }
    if (i < n) off[i] = run + inc - v;
    run += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (blockIdx.x == 0 && lane == 0) off[n] = bsum[n_blocks];
}

}  // namespace ti
#endif
