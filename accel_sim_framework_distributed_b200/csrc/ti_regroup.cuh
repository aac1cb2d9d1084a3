// Writer side of the wire format: kernel-N.trace -> kernel-N.traceg.
//
// Replaces group_per_block of the reference's post-processing tool
// (util/tracer_nvbit/tracer_tool/traces-processing/post-traces-processing.cpp:85-194, "PTP"):
// every instruction line of a raw trace starts with four decimals "tbx tby tbz warp"; the tool
// buckets the lines by (thread block, warp), keeps file order inside a bucket, and writes one
// "#BEGIN_TB ... #END_TB" block per thread block with the four decimals stripped.  Lines that are
// empty or start with '#' or '-' are copied, in order, ahead of all blocks (PTP:110-143).
//
// The reference holds the whole kernel in a vector<vector<vector<string>>>.  Here it is a stable
// sort by key plus prefix sums, all on the device:
//   rg_count   newlines per 16 KiB chunk                      -> line ordinals (scan)
//   rg_keys    one thread per line: key = 1 + tb * W + warp (0 = copied line), payload offset
//   rs_*       stable LSD radix sort of (key, line ordinal), 8 bits per pass
//   scans      output offset of every line and of every bucket's headers
//   rg_headers / rg_lines   the text itself, written once
// Test infrastructure compiles this file with g++ against tests/simt_emu like the decoder.
#ifndef TI_REGROUP_CUH
#define TI_REGROUP_CUH

#include "ti_kernels.cuh"

namespace ti {
namespace rg {

constexpr int kRgChunk = 16384;        // bytes of text whose newlines one CTA iteration counts
constexpr int kRgThreads = 256;
constexpr int kScanTile = 2048;        // items per CTA in the scan and sort kernels (8 per thread)
constexpr int kItems = kScanTile / kRgThreads;
constexpr uint32_t kNoTb = 0xFFFFFFFFu;

struct RgGlobals { uint32_t error_flags, first_error_off, n_inst_lines, n_copy_lines; };

struct RgParams {
  const uint8_t *text;
  uint64_t len;
  uint32_t n_chunks;
  uint32_t grid[3], warps_per_tb;      // from "-grid dim" / "-block dim" (PTP:121-136), parsed on the host
  uint64_t n_tb, n_keys;               // n_keys = 1 + n_tb * warps_per_tb
  uint32_t first_inst_line_ok_after;   // byte offset where both header lines have been seen
  uint32_t *chunk_lines;               // [n_chunks + 1] newlines per chunk, then exclusive scan
  uint32_t *start;                     // [n_lines + 1] line start offsets; start[n_lines] = len + 1
  uint8_t *pdelta;                     // [n_lines] payload start - line start
  uint32_t *key;                       // [n_lines]
  uint32_t *cnt;                       // [n_keys + 1] lines per key
  uint32_t *tb_first;                  // [n_tb] line start of the first line seen of a thread block
  RgGlobals *g;
};

TI_D void rg_error(RgGlobals *g, uint32_t flag, uint32_t off) {
  atomicOr(&g->error_flags, flag);
  atomicMin(&g->first_error_off, off);
}

// ---------------------------------------------------------------- line ordinals
// A line belongs to the chunk that holds the '\n' before it; line 0 has none and is counted
// apart.  The reference's read loop (PTP:103-105, while (!eof) getline) also processes the empty
// string after a final '\n', so the number of lines is the number of newlines + 1.
__global__ void __launch_bounds__(kRgThreads) rg_count_kernel(const RgParams P) {
  TI_DYN_SMEM(smem);
  uint32_t *wsum = reinterpret_cast<uint32_t *>(smem);
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (uint32_t c = blockIdx.x; c < P.n_chunks; c += gridDim.x) {
    const uint64_t cb = (uint64_t)c * kRgChunk;
    uint32_t n = 0;
    for (uint32_t v = tid; v < (uint32_t)(kRgChunk / 16); v += kRgThreads) {
      const uint64_t p = cb + (uint64_t)v * 16;
      if (p + 16 <= P.len) {
        const uint4 x = *reinterpret_cast<const uint4 *>(P.text + p);   // the text buffer is 16 B aligned
        n += __popc(nl_mask16(x));
      } else {
        for (uint64_t q = p; q < P.len && q < p + 16; ++q) n += P.text[q] == '\n';
      }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) n += __shfl_xor_sync(kFull, n, d);
    if (lane == 0) wsum[wid] = n;
    __syncthreads();
    if (tid == 0) {
      uint32_t t = 0;
      for (int w = 0; w < kRgThreads / 32; ++w) t += wsum[w];
      P.chunk_lines[c] = t;
    }
    __syncthreads();
  }
}

TI_D bool rg_dec(const uint8_t *t, uint64_t len, uint64_t &q, uint32_t &v) {  // digits only, fits u32, then one blank
  uint64_t m = 0;
  int nd = 0;
  while (q < len) {
    const unsigned d = (unsigned)t[q] - '0';
    if (d > 9) break;
    m = m * 10 + d;
    ++q;
    if (++nd > 10) return false;
  }
  v = (uint32_t)m;
  return nd > 0 && m <= 0xffffffffull;
}

// One thread per 64 text bytes: every '\n' in them ends a line and starts the next one, which
// this thread indexes.  base ordinal of the chunk = exclusive scan of chunk_lines (+1 for line 0).
__global__ void __launch_bounds__(kRgThreads) rg_keys_kernel(const RgParams P) {
  TI_DYN_SMEM(smem);
  uint32_t *wsum = reinterpret_cast<uint32_t *>(smem);
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (uint32_t c = blockIdx.x; c < P.n_chunks; c += gridDim.x) {
    const uint64_t cb = (uint64_t)c * kRgChunk, p0 = cb + (uint64_t)tid * 64;
    uint64_t bits = 0;   // newline bitmap of this thread's 64 bytes
    for (int v = 0; v < 4; ++v) {
      const uint64_t p = p0 + (uint64_t)v * 16;
      uint32_t m = 0;
      if (p + 16 <= P.len) m = nl_mask16(*reinterpret_cast<const uint4 *>(P.text + p));
      else for (uint64_t q = p; q < P.len && q < p + 16; ++q) m |= (uint32_t)(P.text[q] == '\n') << (q - p);
      bits |= (uint64_t)m << (16 * v);
    }
    const uint32_t mine = __popcll(bits);
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (unsigned w = 0; w < wid; ++w) wbase += wsum[w];
    uint32_t ord = P.chunk_lines[c] + 1 + wbase + inc - mine;   // +1: line 0 precedes every newline
    __syncthreads();
    // line 0 of the file
    const bool first = c == 0 && tid == 0;
    for (int it = first ? -1 : 0; bits || it < 0; ++it) {
      uint64_t ls;
      uint32_t o;
      if (it < 0) { ls = 0; o = 0; }
      else { const int b = __ffsll((long long)bits) - 1; bits &= bits - 1; ls = p0 + (uint64_t)b + 1; o = ord++; }
      // classify and key the line that starts at ls (PTP:108-155)
      uint32_t key = 0;
      uint32_t delta = 0;
      const unsigned ch = ls < P.len ? P.text[ls] : (unsigned)'\n';
      if (ch != '\n' && ch != '#' && ch != '-') {
        uint64_t q = ls;
        uint32_t v4[4];
        bool ok = true;
        for (int k = 0; k < 4 && ok; ++k) {
          while (q < P.len && (P.text[q] == ' ' || P.text[q] == '\t')) ++q;
          ok = rg_dec(P.text, P.len, q, v4[k]);
        }
        // one character after the fourth number is dropped, whatever it is (PTP:152); there must be one
        ok = ok && q < P.len && P.text[q] != '\n';
        uint64_t tb = 0;
        if (ok) {
          tb = (uint64_t)v4[2] * P.grid[1] * P.grid[0] + (uint64_t)v4[1] * P.grid[0] + v4[0];
          ok = tb < P.n_tb && v4[3] < P.warps_per_tb && ls >= P.first_inst_line_ok_after && q + 1 - ls <= 255;
        }
        if (!ok) rg_error(P.g, TI_EF_BAD_STRUCTURE, (uint32_t)ls);
        else {
          key = 1u + (uint32_t)(tb * P.warps_per_tb + v4[3]);
          delta = (uint32_t)(q + 1 - ls);
          // a plain read first: after the block's first few lines the minimum no longer moves (a stale value only costs an atomic)
          if ((uint32_t)ls < *reinterpret_cast<const volatile uint32_t *>(&P.tb_first[tb])) atomicMin(&P.tb_first[tb], (uint32_t)ls);
        }
      }
      P.start[o] = (uint32_t)ls;
      P.pdelta[o] = (uint8_t)delta;
      P.key[o] = key;
      atomicAdd(&P.cnt[key], 1u);
    }
  }
}

// ---------------------------------------------------------------- exclusive scans
// out[i] = sum of f(j), j < i, for i in [0, n]; three launches: tile sums, scan of the tile sums by
// one CTA, tile-local scan + base.  f is evaluated twice.
template <class F>
TI_D uint64_t tile_sum(F f, uint64_t n, uint64_t tile, uint64_t *wsum) {
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint64_t s = 0;
  for (int k = 0; k < kItems; ++k) {
    const uint64_t i = tile * kScanTile + (uint64_t)tid * kItems + k;
    if (i < n) s += f(i);
  }
  s = warp_sum_u64(s);
  if (lane == 0) wsum[wid] = s;
  __syncthreads();
  uint64_t t = 0;
  for (int w = 0; w < kRgThreads / 32; ++w) t += wsum[w];
  __syncthreads();
  return t;
}

template <class F>
__global__ void __launch_bounds__(kRgThreads) scan_sums_kernel(F f, uint64_t n, uint64_t *tile_sums) {
  TI_DYN_SMEM(smem);
  uint64_t *wsum = reinterpret_cast<uint64_t *>(smem);
  const uint64_t n_tiles = (n + kScanTile - 1) / kScanTile;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const uint64_t t = tile_sum(f, n, tile, wsum);
    if (threadIdx.x == 0) tile_sums[tile] = t;
  }
}

__global__ void __launch_bounds__(kRgThreads) scan_tiles_kernel(uint64_t *tile_sums, uint64_t n_tiles) {  // one CTA
  TI_DYN_SMEM(smem);
  uint64_t *wsum = reinterpret_cast<uint64_t *>(smem);
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  uint64_t carry = 0;
  for (uint64_t i0 = 0; i0 < n_tiles; i0 += kRgThreads) {
    const uint64_t i = i0 + tid;
    const uint64_t v = i < n_tiles ? tile_sums[i] : 0;
    uint64_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint64_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    uint64_t wb = 0, tot = 0;
    for (int w = 0; w < kRgThreads / 32; ++w) { if ((unsigned)w < wid) wb += wsum[w]; tot += wsum[w]; }
    if (i < n_tiles) tile_sums[i] = carry + wb + inc - v;
    carry += tot;
    __syncthreads();
  }
  if (tid == 0) tile_sums[n_tiles] = carry;
}

template <class F, class O>
__global__ void __launch_bounds__(kRgThreads) scan_write_kernel(F f, uint64_t n, const uint64_t *tile_sums, O *out) {
  TI_DYN_SMEM(smem);
  uint64_t *wsum = reinterpret_cast<uint64_t *>(smem);
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint64_t n_tiles = (n + kScanTile - 1) / kScanTile;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    uint64_t v[kItems], s = 0;
    for (int k = 0; k < kItems; ++k) {
      const uint64_t i = tile * kScanTile + (uint64_t)tid * kItems + k;
      v[k] = i < n ? f(i) : 0;
      s += v[k];
    }
    uint64_t inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint64_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    uint64_t wb = 0;
    for (unsigned w = 0; w < wid; ++w) wb += wsum[w];
    uint64_t run = tile_sums[tile] + wb + inc - s;
    for (int k = 0; k < kItems; ++k) {
      const uint64_t i = tile * kScanTile + (uint64_t)tid * kItems + k;
      if (i < n) out[i] = (O)run;
      run += v[k];
    }
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = (O)tile_sums[n_tiles];
}

struct FnU32 { const uint32_t *a; TI_HD uint64_t operator()(uint64_t i) const { return a[i]; } };                  // plain array
struct FnHdr { const uint16_t *h; TI_HD uint64_t operator()(uint64_t k) const { return (uint64_t)(h[k] & 0xff) + (h[k] >> 8); } };
struct FnLine {                                                                                                    // payload + '\n' of sorted line p
  const uint32_t *perm, *start; const uint8_t *pdelta;
  TI_HD uint64_t operator()(uint64_t p) const { const uint32_t o = perm[p]; return (uint64_t)(start[o + 1] - start[o] - pdelta[o]); }
};

// ---------------------------------------------------------------- stable LSD radix sort, 8 bits per pass
// hist[d * n_tiles + tile] = items of `tile` whose digit is d; its exclusive scan (digit-major)
// is where the tile's first such item goes.
__global__ void __launch_bounds__(kRgThreads) rs_hist_kernel(const uint32_t *keys, uint64_t n, unsigned shift, uint32_t *hist, uint64_t n_tiles) {
  TI_DYN_SMEM(smem);
  uint32_t *h = reinterpret_cast<uint32_t *>(smem);   // [256]
  const unsigned tid = threadIdx.x;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    h[tid] = 0;
    __syncthreads();
    for (int k = 0; k < kItems; ++k) {
      const uint64_t i = tile * kScanTile + (uint64_t)k * kRgThreads + tid;
      if (i < n) atomicAdd(&h[(keys[i] >> shift) & 0xff], 1u);
    }
    __syncthreads();
    hist[(uint64_t)tid * n_tiles + tile] = h[tid];
    __syncthreads();
  }
}

// Warp w owns items [w * 256, (w + 1) * 256) of the tile, eight steps of 32: count per (warp,
// digit), turn the counts into start offsets warp after warp, then walk the items again and
// hand each its slot (rank among the equal digits of its step from match.any).
__global__ void __launch_bounds__(kRgThreads) rs_scatter_kernel(const uint32_t *keys, const uint32_t *vals, uint32_t *keys_out, uint32_t *vals_out,
                                                               uint64_t n, unsigned shift, const uint32_t *hist_scan, uint64_t n_tiles, int first_pass) {
  TI_DYN_SMEM(smem);
  uint32_t *wc = reinterpret_cast<uint32_t *>(smem);   // [8][256]
  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  constexpr int kW = kRgThreads / 32, kSteps = kScanTile / kRgThreads;
  for (uint64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int w = 0; w < kW; ++w) wc[w * 256 + tid] = 0;
    __syncthreads();
    uint32_t k[kSteps], v[kSteps];
    for (int s = 0; s < kSteps; ++s) {
      const uint64_t i = tile * kScanTile + (uint64_t)wid * (32 * kSteps) + (uint64_t)s * 32 + lane;
      const bool in = i < n;
      k[s] = in ? keys[i] : 0xFFFFFFFFu;
      v[s] = in ? (first_pass ? (uint32_t)i : vals[i]) : 0u;
      if (in) atomicAdd(&wc[wid * 256 + ((k[s] >> shift) & 0xff)], 1u);
    }
    __syncthreads();
    {
      uint32_t run = hist_scan[(uint64_t)tid * n_tiles + tile];
      for (int w = 0; w < kW; ++w) { const uint32_t t = wc[w * 256 + tid]; wc[w * 256 + tid] = run; run += t; }
    }
    __syncthreads();
    for (int s = 0; s < kSteps; ++s) {
      const uint64_t i = tile * kScanTile + (uint64_t)wid * (32 * kSteps) + (uint64_t)s * 32 + lane;
      const bool in = i < n;
      const unsigned d = in ? (k[s] >> shift) & 0xff : 256u + lane;   // out-of-range lanes match nobody
      const uint32_t peers = __match_any_sync(kFull, d);
      const uint32_t rank = __popc(peers & ((1u << lane) - 1));
      uint32_t base = 0;
      if (in) base = wc[wid * 256 + d];
      __syncwarp();
      if (in) {
        keys_out[base + rank] = k[s];
        vals_out[base + rank] = v[s];
        if (rank == 0) wc[wid * 256 + d] = base + __popc(peers);
      }
      __syncwarp();
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------- output text
constexpr uint32_t dec_digits(uint32_t v) {
  return v < 10 ? 1 : v < 100 ? 2 : v < 1000 ? 3 : v < 10000 ? 4 : v < 100000 ? 5 : v < 1000000 ? 6 : v < 10000000 ? 7 : v < 100000000 ? 8 : v < 1000000000 ? 9 : 10;
}
TI_D uint8_t *put_lit(uint8_t *o, const char *s) { while (*s) *o++ = (uint8_t)*s++; return o; }
TI_D uint8_t *put_dec(uint8_t *o, uint32_t v) {
  const uint32_t n = dec_digits(v);
  for (uint32_t i = 0; i < n; ++i) { o[n - 1 - i] = (uint8_t)('0' + v % 10); v /= 10; }
  return o + n;
}
TI_D void tb_coords(const RgParams &P, uint32_t line_start, uint32_t (&xyz)[3]) {   // of the first line seen (PTP:147-151)
  uint64_t q = line_start;
  for (int k = 0; k < 3; ++k) {
    while (q < P.len && (P.text[q] == ' ' || P.text[q] == '\t')) ++q;
    rg_dec(P.text, P.len, q, xyz[k]);
  }
}

// Header bytes in front of / behind the lines of key k (k >= 1: bucket (tb, warp)).  A thread
// block no line mentions is skipped entirely (PTP:159-173); one that is mentioned lists every warp,
// empty ones with "insts = 0" (PTP:174-189).
struct RgHdr { uint32_t front, back; };
TI_D RgHdr rg_hdr_bytes(const RgParams &P, uint64_t k) {
  RgHdr h = {0, 0};
  if (k == 0) return h;
  const uint64_t tb = (k - 1) / P.warps_per_tb;
  const uint32_t w = (uint32_t)((k - 1) % P.warps_per_tb);
  const uint32_t first = P.tb_first[tb];
  if (first == kNoTb) return h;
  if (w == 0) {
    uint32_t xyz[3];
    tb_coords(P, first, xyz);
    h.front += 11 + 16 + dec_digits(xyz[0]) + 1 + dec_digits(xyz[1]) + 1 + dec_digits(xyz[2]) + 1;   // "\n#BEGIN_TB\n" "\nthread block = " x,y,z "\n"
  }
  h.front += 8 + dec_digits(w) + 1 + 8 + dec_digits(P.cnt[k]) + 1;                                     // "\nwarp = " w "\n" "insts = " n "\n"
  if (w == P.warps_per_tb - 1) h.back = 9;                                                             // "\n#END_TB\n"
  return h;
}

__global__ void __launch_bounds__(kRgThreads) rg_hdrsize_kernel(const RgParams P, uint16_t *hsz) {   // front | back << 8 per key
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < P.n_keys; k += (uint64_t)gridDim.x * blockDim.x) {
    const RgHdr h = rg_hdr_bytes(P, k);
    hsz[k] = (uint16_t)(h.front | h.back << 8);
  }
}

struct RgEmit {
  RgParams P;
  const uint16_t *hsz;       // [n_keys] header bytes in front of (low byte) / behind (high byte) a key's lines
  const uint32_t *perm;      // [n_lines] line ordinals sorted by key, file order inside a key
  const uint32_t *skey;      // [n_lines] the keys in the same order
  const uint32_t *bucket;    // [n_keys + 1] exclusive scan of cnt: first sorted position of a key
  const uint64_t *line_off;  // [n_lines + 1] exclusive scan over sorted positions of (payload length + 1)
  const uint64_t *hdr_off;   // [n_keys + 1] exclusive scan over keys of (front + back)
  uint8_t *out;
  uint64_t cap;
  uint32_t *stats;           // [0] thread blocks written, [1] thread blocks without lines, [2] empty warps written
};

__global__ void __launch_bounds__(kRgThreads) rg_headers_kernel(const RgEmit E) {
  const RgParams &P = E.P;
  for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; k < P.n_keys; k += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t tb = (k - 1) / P.warps_per_tb;
    const uint32_t w = (uint32_t)((k - 1) % P.warps_per_tb);
    const uint32_t first = P.tb_first[tb];
    if (first == kNoTb) { if (w == 0) atomicAdd(&E.stats[1], 1u); continue; }
    const RgHdr h = {(uint32_t)(E.hsz[k] & 0xff), (uint32_t)(E.hsz[k] >> 8)};
    const uint64_t at = E.line_off[E.bucket[k]] + E.hdr_off[k];
    const uint64_t end = E.line_off[E.bucket[k + 1]] + E.hdr_off[k] + h.front;
    if (end + h.back > E.cap) continue;   // the host sized the buffer from the same numbers; never taken
    uint8_t *o = E.out + at;
    if (w == 0) {
      uint32_t xyz[3];
      tb_coords(P, first, xyz);
      o = put_lit(o, "\n#BEGIN_TB\n\nthread block = ");
      o = put_dec(o, xyz[0]); *o++ = ','; o = put_dec(o, xyz[1]); *o++ = ','; o = put_dec(o, xyz[2]); *o++ = '\n';
      atomicAdd(&E.stats[0], 1u);
    }
    o = put_lit(o, "\nwarp = "); o = put_dec(o, w);
    o = put_lit(o, "\ninsts = "); o = put_dec(o, P.cnt[k]); *o++ = '\n';
    if (P.cnt[k] == 0) atomicAdd(&E.stats[2], 1u);
    if (h.back) put_lit(E.out + end, "\n#END_TB\n");
  }
}

// A warp takes 32 consecutive sorted lines; each is copied by all lanes, byte per lane (adjacent
// lanes write adjacent bytes, so a 50-byte line is two or three 32 B sectors), '\n' appended.
__global__ void __launch_bounds__(kRgThreads) rg_lines_kernel(const RgEmit E, uint64_t n_lines) {
  const RgParams &P = E.P;
  const unsigned lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t p0 = warp0 * 32; p0 < n_lines; p0 += n_warps * 32) {
    const uint64_t p = p0 + lane;
    uint32_t src = 0, n = 0;
    uint64_t dst = 0;
    if (p < n_lines) {
      const uint32_t o = E.perm[p], k = E.skey[p];
      const uint32_t ls = P.start[o];
      src = ls + P.pdelta[o];
      n = P.start[o + 1] - 1 - src;                        // payload bytes, without the '\n'
      dst = E.line_off[p] + E.hdr_off[k] + (E.hsz[k] & 0xff);
    }
    const unsigned cnt = (unsigned)(n_lines - p0 < 32 ? n_lines - p0 : 32);
    for (unsigned j = 0; j < cnt; ++j) {
      const uint32_t s = __shfl_sync(kFull, src, j), m = __shfl_sync(kFull, n, j);
      const uint64_t d = __shfl_sync(kFull, dst, j);
      if (d + m + 1 > E.cap) continue;
      for (uint32_t b = lane; b <= m; b += 32) E.out[d + b] = b < m ? P.text[s + b] : (uint8_t)'\n';
    }
  }
}

}  // namespace rg
}  // namespace ti
#endif
