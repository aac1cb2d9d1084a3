// Device half of the C ABI: context, kernel objects, buffers, the launch of the
// fused decode kernel and the copies either side of it.  There is no CPU fallback:
// without a CUDA device every compute entry point returns TI_ERR_NO_DEVICE.
#ifdef TI_EMU
#include "cuda_runtime_emu.h"
#else
#include <cuda_runtime.h>
#endif

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifndef TI_EMU
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#endif

#include <algorithm>
#include <vector>

#include "ti_internal.h"
#include "ti_kernels.cuh"
#include "ti_compact.cuh"
#include "ti_compress.cuh"

// the merged ISA table (generated from the reference's ISA_Def headers)
#define TI_ISA_BEGIN static const struct IsaRow { const char *name; uint8_t id[5]; uint8_t cat[5]; } kIsaRows[] = {
#define TI_ISA_ROW(n, a, b, c, d, e, f, g, h, i, j) {n, {a, b, c, d, e}, {f, g, h, i, j}},
#define TI_ISA_END };
#define TI_SPOP_BEGIN static const uint8_t kSpopTable[] = {
#define TI_SPOP(i, v) v,
#define TI_SPOP_END };
#include "isa_tables.inc"

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      ti::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__);   \
      return TI_ERR_CUDA;                                                                          \
    }                                                                                              \
  } while (0)

namespace {

struct DevBuf {
  void *p = nullptr;
  size_t bytes = 0;
  ti_status ensure(size_t need) {
    if (need <= bytes) return TI_OK;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    CK(cudaMalloc(&p, need));
    bytes = need;
    return TI_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};
struct HostBuf {
  void *p = nullptr;
  size_t bytes = 0;
  ti_status ensure(size_t need) {
    if (need <= bytes) return TI_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; bytes = 0;
    CK(cudaMallocHost(&p, need ? need : 1));
    bytes = need;
    return TI_OK;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};

// the four look-back chains of a launch live back to back in one scratch buffer
void set_chains(ti::DecodeParams &P, void *base, uint32_t n_chunks) {
  for (int i = 0; i < ti::kChains; ++i) P.chain[i] = (unsigned long long *)base + (size_t)i * (n_chunks + 1);
}

// zero what a launch of the fused kernel starts from (ti_reset_kernel)
ti_status reset_launch_state(cudaStream_t s, void *chains, uint32_t n_chunks, void *globals, void *counters, void *txn_off) {
  const uint64_t n_words = ti::chain_bytes(n_chunks) / 8;
  const int grid = (int)std::min<uint64_t>(64, (n_words + 255) / 256 + 1);
#ifdef TI_EMU
  emu::launch(dim3(grid), dim3(256), 0, [=]() { ti::ti_reset_kernel((unsigned long long *)chains, n_words, (ti::Globals *)globals, (unsigned long long *)counters, (uint32_t *)txn_off); });
#else
  ti::ti_reset_kernel<<<grid, 256, 0, s>>>((unsigned long long *)chains, n_words, (ti::Globals *)globals, (unsigned long long *)counters, (uint32_t *)txn_off);
#endif
  CK(cudaGetLastError());
  return TI_OK;
}

int arch_column(unsigned v) {  // trace_driven.cc:100-118
  switch (v) {
    case 35: return 0;
    case 60: case 61: return 1;
    case 70: return 2;
    case 75: return 3;
    case 80: case 86: return 4;
  }
  return -1;
}

}  // namespace

struct ti_regroup_state;
static void regroup_state_free(ti_regroup_state *s);
struct ti_compress_state;
static void compress_state_free(ti_compress_state *w);

struct ti_ctx {
  ti_config cfg;
  ti_regroup_state *rg = nullptr;   // workspace of ti_regroup_* (allocated on first use)
  ti_compress_state *cmp = nullptr; // workspace of ti_compress_rows
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  ti::IsaSlot *d_isa[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  uint8_t *d_spop = nullptr;
  int grid = 0;
};

struct ti_stream_state;
static void stream_state_free(ti_stream_state *ss);

struct ti_kernel {
  ti_ctx *ctx = nullptr;
  ti_stream_state *streams = nullptr;   // ti_kernel_decode_streamed: streams, events, two device buffer sets
  bool streamed = false, compact = false;   // the last decode went through the pipelined path / left compact results
  ti_kernel_hdr hdr;
  int col = -1;
  const char *h_body = nullptr;   // host body (owned pinned file image or borrowed)
  size_t body_len = 0;
  HostBuf h_file;                 // owns the file image for ti_kernel_open_file (when it could not be mapped)
  void *map_base = nullptr;       // ... or the read-only mapping of the file (pages come and go with the page cache)
  size_t map_len = 0;
  const uint8_t *d_borrowed = nullptr;
  DevBuf d_text;
  size_t sh_begin = 0, sh_len = 0;
  bool uploaded = false, decoded = false, synced = false, retried = false;
  DevBuf inst, rows, txn_off, txn, tb, warp, chains, globals, counters, counters_sum;
  uint64_t cap_inst = 0, cap_mem = 0, cap_txn = 0, cap_tb = 0, cap_warp = 0;
  HostBuf h_inst, h_rows, h_txn_off, h_txn, h_tb, h_warp, h_small, h_desc;
  ti_counts counts;
  ti_counters host_counters;
  float decode_ms = 0;
};

static ti_status launch_decode(ti_kernel *k);

extern "C" {

int ti_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

static ti_status ctx_init(ti_ctx *c, const ti_config *cfg);

ti_status ti_ctx_create(const ti_config *cfg, ti_ctx **out) {
  if (!cfg || !out) return TI_ERR_INVALID_ARG;
  *out = nullptr;
  if (cfg->device < 0 || ti_device_count() <= cfg->device) { ti::set_error("CUDA device %d not present", cfg->device); return TI_ERR_NO_DEVICE; }
  if (cfg->coalesce_arch < 13) { ti::set_error("coalesce_arch < 13 aborts in the reference (abstract_hardware_model.cc:436)"); return TI_ERR_INVALID_ARG; }
  const int parts = cfg->mem_warp_parts;
  if (parts < 1 || parts > 32 || (32 % parts) || cfg->num_shmem_bank < 1) { ti::set_error("bad mem_warp_parts / num_shmem_bank"); return TI_ERR_INVALID_ARG; }
  const uint32_t *lats[5] = {cfg->lat_int, cfg->lat_sp, cfg->lat_dp, cfg->lat_sfu, cfg->lat_tensor};
  for (int i = 0; i < 13; ++i) {
    const uint32_t *l = i < 5 ? lats[i] : cfg->lat_spec[i - 5];
    if (l[0] > 65535 || l[1] > 255) { ti::set_error("latency > 65535 or initiation interval > 255 does not fit ti_inst_rec"); return TI_ERR_INVALID_ARG; }
  }
  ti_ctx *c = new ti_ctx;
  c->cfg = *cfg;
  const ti_status st = ctx_init(c, cfg);
  if (st != TI_OK) { ti_ctx_destroy(c); return st; }   // nothing half-built is left behind
  *out = c;
  return TI_OK;
}

static ti_status ctx_init(ti_ctx *c, const ti_config *cfg) {
  CK(cudaSetDevice(cfg->device));
  CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&c->ev0));
  CK(cudaEventCreate(&c->ev1));
  // one open-addressing table per architecture column
  for (int col = 0; col < 5; ++col) {
    std::vector<ti::IsaSlot> tab(ti::kIsaSlots);
    memset(tab.data(), 0, tab.size() * sizeof(ti::IsaSlot));
    for (const IsaRow &r : kIsaRows) {
      if (!r.id[col]) continue;
      uint64_t lo = 0;
      uint32_t hi = 0;
      const size_t n = strlen(r.name);
      for (size_t i = 0; i < n; ++i) {
        if (i < 8) lo |= (uint64_t)(uint8_t)r.name[i] << (8 * i);
        else hi |= (uint32_t)(uint8_t)r.name[i] << (8 * (i - 8));
      }
      uint32_t h = ti::isa_hash(lo, hi);
      while (tab[h].id) h = (h + 1) & (ti::kIsaSlots - 1);
      tab[h].lo = lo; tab[h].hi = hi; tab[h].id = r.id[col]; tab[h].cat = r.cat[col];
    }
    CK(cudaMalloc((void **)&c->d_isa[col], tab.size() * sizeof(ti::IsaSlot)));
    CK(cudaMemcpyAsync(c->d_isa[col], tab.data(), tab.size() * sizeof(ti::IsaSlot), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
  }
  uint8_t spop[256];
  memset(spop, 0, sizeof spop);
  memcpy(spop, kSpopTable, sizeof kSpopTable < 256 ? sizeof kSpopTable : 256);
  CK(cudaMalloc((void **)&c->d_spop, 256));
  CK(cudaMemcpyAsync(c->d_spop, spop, 256, cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  // persistent grid: every SM full of resident CTAs
  CK(cudaFuncSetAttribute(ti::ti_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ti::kSmemBytes));
  int sms = 0, per_sm = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device));
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ti::ti_decode_kernel, ti::kThreads, ti::kSmemBytes));
  if (per_sm < 1) { ti::set_error("decode kernel does not fit on an SM"); return TI_ERR_CUDA; }
  c->grid = sms * per_sm;
  return TI_OK;
}

void ti_ctx_destroy(ti_ctx *c) {
  if (!c) return;
  regroup_state_free(c->rg);
  compress_state_free(c->cmp);
  for (auto &p : c->d_isa) if (p) cudaFree(p);
  if (c->d_spop) cudaFree(c->d_spop);
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

static ti_status new_kernel(ti_ctx *ctx, const ti_kernel_hdr *hdr, ti_kernel **out) {
  if (!ctx || !hdr || !out) return TI_ERR_INVALID_ARG;
  const int col = arch_column(hdr->binary_version);
  if (col < 0) {  // trace_driven.cc:113-118 prints this and exits
    ti::set_error("unsupported binary version: %u", hdr->binary_version);
    return TI_ERR_BINARY_VERSION;
  }
  ti_kernel *k = new ti_kernel;
  k->ctx = ctx; k->hdr = *hdr; k->col = col;
  memset(&k->counts, 0, sizeof k->counts);
  memset(&k->host_counters, 0, sizeof k->host_counters);
  *out = k;
  return TI_OK;
}

ti_status ti_kernel_open_file(ti_ctx *ctx, const char *path, ti_kernel_hdr *hdr, ti_kernel **out) {
  if (!ctx || !path || !out) return TI_ERR_INVALID_ARG;
  CK(cudaSetDevice(ctx->cfg.device));
  // The file is mapped read-only rather than read into pinned memory: a caller that walks it window by window
  // (ti_kernel_decode_window_compact, the C++ drop-in) then never holds more than a window of it.
  void *map = nullptr;
  size_t n = 0;
  HostBuf img;
#ifndef TI_EMU
  {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { ti::set_error("Unable to open file: %s", path); return TI_ERR_IO; }  // trace_parser.cc:306-309
    struct stat sb;
    if (fstat(fd, &sb) == 0 && sb.st_size > 0) {
      n = (size_t)sb.st_size;
      map = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
      if (map == MAP_FAILED) map = nullptr;
    }
    close(fd);
  }
#endif
  if (!map) {
    FILE *f = fopen(path, "rb");
    if (!f) { ti::set_error("Unable to open file: %s", path); return TI_ERR_IO; }  // trace_parser.cc:306-309
    fseek(f, 0, SEEK_END);
    n = (size_t)ftell(f);
    fseek(f, 0, SEEK_SET);
    ti_status st = img.ensure(n + 16);
    if (st != TI_OK) { fclose(f); return st; }
    const size_t got = n ? fread(img.p, 1, n, f) : 0;
    fclose(f);
    if (got != n) { img.release(); ti::set_error("short read on %s", path); return TI_ERR_IO; }
  }
  const char *base = map ? (const char *)map : (const char *)img.p;
  ti_kernel_hdr h;
  ti_parse_kernel_header(base, n, &h);
  if (hdr) *hdr = h;
  ti_status st = new_kernel(ctx, &h, out);
  if (st != TI_OK) {
    img.release();
#ifndef TI_EMU
    if (map) munmap(map, n);
#endif
    return st;
  }
  (*out)->h_file = img;
  (*out)->map_base = map; (*out)->map_len = n;
  (*out)->h_body = base + h.body_offset;
  (*out)->body_len = n - h.body_offset;
  return TI_OK;
}

ti_status ti_kernel_open_buffer(ti_ctx *ctx, const char *body, size_t len, const ti_kernel_hdr *hdr, ti_kernel **out) {
  if (!body && len) return TI_ERR_INVALID_ARG;
  ti_status st = new_kernel(ctx, hdr, out);
  if (st != TI_OK) return st;
  (*out)->h_body = body;
  (*out)->body_len = len;
  return TI_OK;
}

ti_status ti_kernel_open_device(ti_ctx *ctx, const void *d_body, size_t len, const ti_kernel_hdr *hdr, ti_kernel **out) {
  if ((!d_body && len) || ((uintptr_t)d_body & 15)) { ti::set_error("device body must be 16-byte aligned"); return TI_ERR_INVALID_ARG; }
  ti_status st = new_kernel(ctx, hdr, out);
  if (st != TI_OK) return st;
  (*out)->d_borrowed = (const uint8_t *)d_body;
  (*out)->body_len = len;
  (*out)->sh_begin = 0; (*out)->sh_len = len; (*out)->uploaded = true;
  return TI_OK;
}

void ti_kernel_close(ti_kernel *k) {
  if (!k) return;
  cudaStreamSynchronize(k->ctx->stream);
  stream_state_free(k->streams);
  DevBuf *d[] = {&k->d_text, &k->inst, &k->rows, &k->txn_off, &k->txn, &k->tb, &k->warp, &k->chains, &k->globals, &k->counters, &k->counters_sum};
  for (DevBuf *b : d) b->release();
  HostBuf *h[] = {&k->h_file, &k->h_inst, &k->h_rows, &k->h_txn_off, &k->h_txn, &k->h_tb, &k->h_warp, &k->h_small, &k->h_desc};
  for (HostBuf *b : h) b->release();
#ifndef TI_EMU
  if (k->map_base) munmap(k->map_base, k->map_len);
#endif
  delete k;
}

ti_status ti_kernel_upload(ti_kernel *k, size_t begin, size_t end) {
  if (k && end == (size_t)-1) end = k->body_len;
  if (!k || begin > end || end > k->body_len) return TI_ERR_INVALID_ARG;
  CK(cudaSetDevice(k->ctx->cfg.device));   // allocations and launches below belong to the context's device
  if (end - begin > 0xfffffff0ull) { ti::set_error("shard of %zu bytes: cut it with ti_plan_shards", end - begin); return TI_ERR_TOO_LARGE; }
  k->decoded = k->synced = k->retried = k->streamed = k->compact = false;
  if (k->d_borrowed) {
    if (begin & 15) { ti::set_error("device-resident shards must start 16-byte aligned"); return TI_ERR_INVALID_ARG; }
    k->sh_begin = begin; k->sh_len = end - begin; k->uploaded = true;
    return TI_OK;
  }
  ti_status st = k->d_text.ensure(((end - begin + 15) & ~(size_t)15) + 16);
  if (st != TI_OK) return st;
  if (end > begin) CK(cudaMemcpyAsync(k->d_text.p, k->h_body + begin, end - begin, cudaMemcpyHostToDevice, k->ctx->stream));
  k->sh_begin = begin; k->sh_len = end - begin; k->uploaded = true;
  return TI_OK;
}

ti_status ti_kernel_decode(ti_kernel *k) {
  if (!k) return TI_ERR_INVALID_ARG;
  if (!k->uploaded) { ti::set_error("ti_kernel_decode before ti_kernel_upload"); return TI_ERR_STATE; }
  CK(cudaSetDevice(k->ctx->cfg.device));   // allocations and launches below belong to the context's device
  const ti_config &cfg = k->ctx->cfg;
  const double len = (double)k->sh_len;
  // default capacities: the instruction bound is exact (no tracer line is shorter than 24
  // bytes); the others are generous guesses, and a decode that overflows is re-run once
  // with the exact counts the first pass measured (ti_kernel_sync).
  if (!k->retried) {
    k->cap_inst = (uint64_t)(len * (cfg.cap_inst_per_byte > 0 ? cfg.cap_inst_per_byte : 1.0 / 24)) + ti::kRecCap;
    k->cap_mem = (uint64_t)(len * (cfg.cap_mem_per_byte > 0 ? cfg.cap_mem_per_byte : 1.0 / 32)) + ti::kRecCap;
    k->cap_txn = (uint64_t)(len * (cfg.cap_txn_per_byte > 0 ? cfg.cap_txn_per_byte : 1.0 / 8)) + 4096;
    k->cap_tb = (uint64_t)(len / 40) + 64;
    k->cap_warp = (uint64_t)(len / 19) + 64;
  }
  return launch_decode(k);
}

ti_status ti_kernel_sync(ti_kernel *k, ti_counts *counts) {
  if (!k) return TI_ERR_INVALID_ARG;
  if (!k->decoded) { ti::set_error("ti_kernel_sync before ti_kernel_decode"); return TI_ERR_STATE; }
  CK(cudaSetDevice(k->ctx->cfg.device));   // allocations and launches below belong to the context's device
  ti_ctx *c = k->ctx;
  ti_status st = k->h_small.ensure(sizeof(ti::Globals) + sizeof(ti_counters));
  if (st != TI_OK) return st;
  ti::Globals *g = (ti::Globals *)k->h_small.p;
  ti_counters *hc = (ti_counters *)((char *)k->h_small.p + sizeof(ti::Globals));
  CK(cudaMemcpyAsync(g, k->globals.p, sizeof(ti::Globals), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(hc, k->counters.p, sizeof(ti_counters), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  CK(cudaGetLastError());
  CK(cudaEventElapsedTime(&k->decode_ms, c->ev0, c->ev1));
  if (g->overflow & 32) { ti::set_error("more than 2^32-2 records in one shard: cut it with ti_plan_shards"); return TI_ERR_TOO_LARGE; }
  if (g->overflow) {
    if (k->retried) { ti::set_error("output capacity exceeded twice (flags %u)", g->overflow); return TI_ERR_CAPACITY; }
    // the first pass counted everything: size the arrays exactly and decode again
    k->retried = true;
    k->cap_inst = hc->n_inst + 1; k->cap_mem = hc->n_mem + 1; k->cap_txn = hc->n_txn + 1;
    k->cap_tb = hc->n_tb + 1; k->cap_warp = hc->n_warp + 1;
    st = launch_decode(k);
    if (st != TI_OK) return st;
    return ti_kernel_sync(k, counts);
  }
  k->host_counters = *hc;
  memset(&k->counts, 0, sizeof k->counts);
  k->counts.n_inst = hc->n_inst; k->counts.n_mem = hc->n_mem; k->counts.n_txn = hc->n_txn;
  k->counts.n_tb = hc->n_tb; k->counts.n_warp = hc->n_warp; k->counts.text_bytes = k->sh_len;
  uint32_t ef = g->error_flags;
  if (hc->n_begin_tb != hc->n_end_tb || hc->n_begin_tb != hc->n_tb || hc->reserved[0] != hc->n_warp) ef |= TI_EF_BAD_STRUCTURE;
  k->counts.error_flags = ef;
  k->counts.first_error_line_off = g->first_error_off == 0xffffffffu ? 0 : g->first_error_off;
  k->synced = ef == 0;   // a shard with grammar errors has counts and flags, but nothing to fetch
  if (counts) *counts = k->counts;
  if (ef) { ti::set_error("trace text outside the tracer grammar: flags 0x%x near shard offset %u", ef, g->first_error_off); return TI_ERR_PARSE; }
  return TI_OK;
}

ti_status ti_kernel_fetch(ti_kernel *k, ti_view *v) {
  if (!k || !v) return TI_ERR_INVALID_ARG;
  if (!k->synced) { ti::set_error("ti_kernel_fetch before a successful ti_kernel_sync"); return TI_ERR_STATE; }
  if (k->compact) { ti::set_error("ti_kernel_fetch after a compact decode: the view was returned by ti_kernel_decode_streamed_compact"); return TI_ERR_STATE; }
  CK(cudaSetDevice(k->ctx->cfg.device));   // allocations and launches below belong to the context's device
  if (k->streamed) {  // ti_kernel_decode_streamed already brought everything to the host
    v->text = k->h_body;
    v->inst = (const ti_inst_rec *)k->h_inst.p; v->addr_rows = (const uint64_t *)k->h_rows.p;
    v->txn_off = (const uint32_t *)k->h_txn_off.p; v->txn = (const ti_txn_rec *)k->h_txn.p;
    v->tb = (const ti_tb_rec *)k->h_tb.p; v->warp = (const ti_warp_rec *)k->h_warp.p;
    v->counts = k->counts;
    return TI_OK;
  }
  cudaStream_t s = k->ctx->stream;
  const ti_counts &n = k->counts;
  ti_status st;
  if ((st = k->h_inst.ensure(n.n_inst * sizeof(ti_inst_rec))) != TI_OK) return st;
  if ((st = k->h_rows.ensure(n.n_mem * 256)) != TI_OK) return st;
  if ((st = k->h_txn_off.ensure((n.n_mem + 1) * 4)) != TI_OK) return st;
  if ((st = k->h_txn.ensure(n.n_txn * sizeof(ti_txn_rec))) != TI_OK) return st;
  if ((st = k->h_tb.ensure(n.n_tb * sizeof(ti_tb_rec))) != TI_OK) return st;
  if ((st = k->h_warp.ensure(n.n_warp * sizeof(ti_warp_rec))) != TI_OK) return st;
  if (n.n_inst) CK(cudaMemcpyAsync(k->h_inst.p, k->inst.p, n.n_inst * sizeof(ti_inst_rec), cudaMemcpyDeviceToHost, s));
  if (n.n_mem) CK(cudaMemcpyAsync(k->h_rows.p, k->rows.p, n.n_mem * 256, cudaMemcpyDeviceToHost, s));
  CK(cudaMemcpyAsync(k->h_txn_off.p, k->txn_off.p, (n.n_mem + 1) * 4, cudaMemcpyDeviceToHost, s));
  if (n.n_txn) CK(cudaMemcpyAsync(k->h_txn.p, k->txn.p, n.n_txn * sizeof(ti_txn_rec), cudaMemcpyDeviceToHost, s));
  if (n.n_tb) CK(cudaMemcpyAsync(k->h_tb.p, k->tb.p, n.n_tb * sizeof(ti_tb_rec), cudaMemcpyDeviceToHost, s));
  if (n.n_warp) CK(cudaMemcpyAsync(k->h_warp.p, k->warp.p, n.n_warp * sizeof(ti_warp_rec), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  v->text = k->h_body ? k->h_body + k->sh_begin : nullptr;
  v->inst = (const ti_inst_rec *)k->h_inst.p;
  v->addr_rows = (const uint64_t *)k->h_rows.p;
  v->txn_off = (const uint32_t *)k->h_txn_off.p;
  v->txn = (const ti_txn_rec *)k->h_txn.p;
  v->tb = (const ti_tb_rec *)k->h_tb.p;
  v->warp = (const ti_warp_rec *)k->h_warp.p;
  v->counts = n;
  return TI_OK;
}

ti_status ti_kernel_dev_view(ti_kernel *k, ti_dev_view *v) {
  if (!k || !v) return TI_ERR_INVALID_ARG;
  if (!k->synced) { ti::set_error("ti_kernel_dev_view before a successful ti_kernel_sync"); return TI_ERR_STATE; }
  if (k->streamed) { ti::set_error("ti_kernel_dev_view after ti_kernel_decode_streamed: the results are on the host (ti_kernel_fetch)"); return TI_ERR_STATE; }
  v->text = k->d_borrowed ? (const void *)(k->d_borrowed + k->sh_begin) : k->d_text.p;
  v->inst = k->inst.p; v->addr_rows = k->rows.p; v->txn_off = k->txn_off.p; v->txn = k->txn.p;
  v->tb = k->tb.p; v->warp = k->warp.p; v->counters = k->counters.p;
  v->counts = k->counts;
  return TI_OK;
}

ti_status ti_kernel_counters(ti_kernel *k, ti_counters *out) {
  if (!k || !out) return TI_ERR_INVALID_ARG;
  if (!k->synced) { ti::set_error("ti_kernel_counters before a successful ti_kernel_sync"); return TI_ERR_STATE; }
  *out = k->host_counters;
  return TI_OK;
}

ti_status ti_counters_finalize(const ti_ctx *ctx, const ti_counters *c, ti_power_counters *out) {
  if (!ctx) return TI_ERR_INVALID_ARG;
  return ti_counters_finalize_cfg(&ctx->cfg, c, out);
}

ti_status ti_kernel_path_stats(ti_kernel *k, uint32_t out[4]) {
  if (!k || !out) return TI_ERR_INVALID_ARG;
  if (!k->decoded || k->streamed || !k->h_small.p) return TI_ERR_STATE;
  const ti::Globals *g = (const ti::Globals *)k->h_small.p;   // brought to the host by ti_kernel_sync
  for (int i = 0; i < 4; ++i) out[i] = g->rare[i];
  return TI_OK;
}

ti_status ti_kernel_timing(ti_kernel *k, float *decode_ms) {
  if (!k || !decode_ms) return TI_ERR_INVALID_ARG;
  if (!k->synced) return TI_ERR_STATE;
  *decode_ms = k->decode_ms;
  return TI_OK;
}

}  // extern "C"

static ti_status launch_decode(ti_kernel *k) {
  ti_ctx *c = k->ctx;
  const ti_config &cfg = c->cfg;
  cudaStream_t s = c->stream;
  const uint32_t n_chunks = (uint32_t)((k->sh_len + ti::kChunk - 1) / ti::kChunk);
  ti_status st;
  if ((st = k->inst.ensure(k->cap_inst * sizeof(ti_inst_rec))) != TI_OK) return st;
  if ((st = k->rows.ensure(k->cap_mem * 256)) != TI_OK) return st;
  if ((st = k->txn_off.ensure((k->cap_mem + 1) * 4)) != TI_OK) return st;
  if ((st = k->txn.ensure(k->cap_txn * sizeof(ti_txn_rec))) != TI_OK) return st;
  if ((st = k->tb.ensure(k->cap_tb * sizeof(ti_tb_rec))) != TI_OK) return st;
  if ((st = k->warp.ensure(k->cap_warp * sizeof(ti_warp_rec))) != TI_OK) return st;
  if ((st = k->chains.ensure(ti::chain_bytes(n_chunks))) != TI_OK) return st;
  if ((st = k->globals.ensure(sizeof(ti::Globals))) != TI_OK) return st;
  if ((st = k->counters.ensure(sizeof(ti_counters))) != TI_OK) return st;

  ti::DecodeParams P;
  memset(&P, 0, sizeof P);
  P.text = k->d_borrowed ? k->d_borrowed + k->sh_begin : (const uint8_t *)k->d_text.p;
  P.len = k->sh_len;
  P.n_chunks = n_chunks;
  P.trace_version = k->hdr.trace_version;
  P.shmem_base = k->hdr.shmem_base_addr; P.local_base = k->hdr.local_base_addr;
  P.inst = (ti_inst_rec *)k->inst.p; P.cap_inst = k->cap_inst;
  P.rows = (uint64_t *)k->rows.p; P.cap_mem = k->cap_mem;
  P.txn_off = (uint32_t *)k->txn_off.p;
  P.txn = (ti_txn_rec *)k->txn.p; P.cap_txn = k->cap_txn;
  P.tb = (ti_tb_rec *)k->tb.p; P.cap_tb = k->cap_tb;
  P.warp = (ti_warp_rec *)k->warp.p; P.cap_warp = k->cap_warp;
  set_chains(P, k->chains.p, n_chunks);
  P.g = (ti::Globals *)k->globals.p;
  P.counters = (unsigned long long *)k->counters.p;
  P.isa = c->d_isa[k->col]; P.spop = c->d_spop;
  for (unsigned slot = 0; slot < 32; ++slot) {  // trace_driven.cc:441-480
    const unsigned op = slot >= 18 ? 100 + (slot - 18) : slot;
    const uint32_t *l = nullptr;
    switch (op) {
      case ti::OP_ALU: case ti::OP_INTP: case ti::OP_BRANCH: case ti::OP_CALL: case ti::OP_RET: l = cfg.lat_int; break;
      case ti::OP_SP: l = cfg.lat_sp; break;
      case ti::OP_DP: l = cfg.lat_dp; break;
      case ti::OP_SFU: l = cfg.lat_sfu; break;
      case ti::OP_TENSOR: l = cfg.lat_tensor; break;
    }
    if (op >= 100 && op < 108) l = cfg.lat_spec[op - 100];
    P.lat[slot][0] = l ? (uint16_t)l[0] : 1; P.lat[slot][1] = l ? (uint16_t)l[1] : 1;
  }
  P.coalesce_arch = cfg.coalesce_arch; P.mem_warp_parts = cfg.mem_warp_parts; P.num_shmem_bank = cfg.num_shmem_bank;
  P.shmem_limited_broadcast = cfg.shmem_limited_broadcast; P.gmem_skip_L1D = cfg.gmem_skip_L1D;

  CK(cudaEventRecord(c->ev0, s));
  { const ti_status rs = reset_launch_state(s, k->chains.p, n_chunks, k->globals.p, k->counters.p, k->txn_off.p); if (rs != TI_OK) return rs; }
  if (n_chunks) {
    const int grid = (int)std::min<uint32_t>((uint32_t)c->grid, n_chunks);
#ifdef TI_EMU
    emu::launch(dim3(grid), dim3(ti::kThreads), ti::kSmemBytes, [=]() { ti::ti_decode_kernel(P); });
    emu::launch(dim3(2), dim3(256), 0, [=]() { ti::ti_check_kernel((const ti_warp_rec *)k->warp.p, k->cap_warp, P.counters, P.g); });
#else
    ti::ti_decode_kernel<<<grid, ti::kThreads, ti::kSmemBytes, s>>>(P);
    ti::ti_check_kernel<<<c->grid, 256, 0, s>>>((const ti_warp_rec *)k->warp.p, k->cap_warp, P.counters, P.g);
#endif
    CK(cudaGetLastError());
  }
  CK(cudaEventRecord(c->ev1, s));
  k->decoded = true;
  k->synced = false;
  return TI_OK;
}

// =====================================================================================
// ti_kernel_decode_streamed: host text -> host results with the copies overlapped.
//
// The body is cut into shards at "#BEGIN_TB" (ti_plan_shards).  Three streams run a
// software pipeline over them with two device buffer sets:
//     H2D(i+1)  ||  decode(i)  ||  D2H(i-1)
// PCIe is full duplex, so the text going up and the records coming down share the link;
// the end-to-end time approaches max(H2D, D2H) instead of their sum.  The kernel is told
// what the preceding shards produced (DecodeParams::off_*), so every index and text
// offset it stores is already file-global and the host arrays are a plain concatenation.
// =====================================================================================
namespace {

struct StreamSet {
  DevBuf text, inst, rows, txn_off, txn, tb, warp, chains, globals, counters;
  DevBuf desc, rows_c, txn_c, txn_off_c, part;   // compact results only
  HostBuf small;                           // Globals + ti_counters of the shard
  cudaEvent_t ev_h2d = nullptr, ev_dec = nullptr, ev_d2h = nullptr, t0 = nullptr, t1 = nullptr;
};

struct StreamState {
  cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
  StreamSet set[2];
  bool ready = false;
};

void fill_model_params(ti::DecodeParams &P, const ti_ctx *c, int col, const ti_kernel_hdr &hdr) {
  const ti_config &cfg = c->cfg;
  P.trace_version = hdr.trace_version;
  P.shmem_base = hdr.shmem_base_addr; P.local_base = hdr.local_base_addr;
  P.isa = c->d_isa[col]; P.spop = c->d_spop;
  for (unsigned slot = 0; slot < 32; ++slot) {  // trace_driven.cc:441-480
    const unsigned op = slot >= 18 ? 100 + (slot - 18) : slot;
    const uint32_t *l = nullptr;
    switch (op) {
      case ti::OP_ALU: case ti::OP_INTP: case ti::OP_BRANCH: case ti::OP_CALL: case ti::OP_RET: l = cfg.lat_int; break;
      case ti::OP_SP: l = cfg.lat_sp; break;
      case ti::OP_DP: l = cfg.lat_dp; break;
      case ti::OP_SFU: l = cfg.lat_sfu; break;
      case ti::OP_TENSOR: l = cfg.lat_tensor; break;
    }
    if (op >= 100 && op < 108) l = cfg.lat_spec[op - 100];
    P.lat[slot][0] = l ? (uint16_t)l[0] : 1; P.lat[slot][1] = l ? (uint16_t)l[1] : 1;
  }
  P.coalesce_arch = cfg.coalesce_arch; P.mem_warp_parts = cfg.mem_warp_parts; P.num_shmem_bank = cfg.num_shmem_bank;
  P.shmem_limited_broadcast = cfg.shmem_limited_broadcast; P.gmem_skip_L1D = cfg.gmem_skip_L1D;
}

}  // namespace

struct ti_stream_state : StreamState {};

static void stream_state_free(ti_stream_state *ss) {
  if (!ss) return;
  for (StreamSet &t : ss->set) {
    DevBuf *d[] = {&t.text, &t.inst, &t.rows, &t.txn_off, &t.txn, &t.tb, &t.warp, &t.chains, &t.globals, &t.counters, &t.desc, &t.rows_c, &t.txn_c, &t.txn_off_c, &t.part};
    for (DevBuf *b : d) b->release();
    t.small.release();
    cudaEvent_t *e[] = {&t.ev_h2d, &t.ev_dec, &t.ev_d2h, &t.t0, &t.t1};
    for (cudaEvent_t *x : e) if (*x) cudaEventDestroy(*x);
  }
  if (ss->s_h2d) cudaStreamDestroy(ss->s_h2d);
  if (ss->s_d2h) cudaStreamDestroy(ss->s_d2h);
  delete ss;
}

static ti_status decode_streamed_impl(ti_kernel *k, size_t shard_bytes, ti_view *v, ti_view_compact *vc) {
  const bool compact = vc != nullptr;
  if (!k || (!v && !vc)) return TI_ERR_INVALID_ARG;
  if (!k->h_body && k->body_len) { ti::set_error("ti_kernel_decode_streamed needs a host-resident body"); return TI_ERR_STATE; }
  if (k->body_len > 0xfffffff0ull) { ti::set_error("body of %zu bytes: opcode offsets are 32 bit", k->body_len); return TI_ERR_TOO_LARGE; }
  CK(cudaSetDevice(k->ctx->cfg.device));   // allocations and launches below belong to the context's device
  ti_ctx *c = k->ctx;
  const ti_config &cfg = c->cfg;
  if (!shard_bytes) shard_bytes = (size_t)64 << 20;
  int n_sh = (int)std::max<size_t>(1, (k->body_len + shard_bytes - 1) / shard_bytes);
  std::vector<size_t> cuts(n_sh + 1);
  ti_status st = ti_plan_shards(k->h_body, k->body_len, n_sh, cuts.data());
  if (st != TI_OK) return st;
  // ti_plan_shards returns empty tail shards when no "#BEGIN_TB" lies behind a nominal cut (a thread block longer
  // than a shard, fewer thread blocks than shards).  An empty shard launches nothing, so its buffers would hand
  // stale values (the CSR sentinel txn_off[0]) to the copies below: drop them here.
  {
    int m = 0;
    for (int i = 0; i < n_sh; ++i)
      if (cuts[i + 1] > cuts[i]) cuts[++m] = cuts[i + 1];
    if (!m) { m = 1; cuts[1] = k->body_len; }               // an empty body stays one (empty) shard
    n_sh = m;
    cuts.resize(m + 1);
  }
  size_t max_len = 0;
  for (int i = 0; i < n_sh; ++i) max_len = std::max(max_len, cuts[i + 1] - cuts[i]);

  if (!k->streams) k->streams = new ti_stream_state;
  ti_stream_state *ss = k->streams;
  if (!ss->ready) {
    CK(cudaStreamCreateWithFlags(&ss->s_h2d, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&ss->s_d2h, cudaStreamNonBlocking));
    for (StreamSet &t : ss->set) {
      CK(cudaEventCreate(&t.ev_h2d)); CK(cudaEventCreate(&t.ev_dec)); CK(cudaEventCreate(&t.ev_d2h));
      CK(cudaEventCreate(&t.t0)); CK(cudaEventCreate(&t.t1));
    }
    ss->ready = true;
  }
  // per-shard device capacities (same per-byte defaults as ti_kernel_decode)
  const double L = (double)max_len;
  const uint64_t cap_inst = (uint64_t)(L * (cfg.cap_inst_per_byte > 0 ? cfg.cap_inst_per_byte : 1.0 / 24)) + ti::kRecCap;
  const uint64_t cap_mem = (uint64_t)(L * (cfg.cap_mem_per_byte > 0 ? cfg.cap_mem_per_byte : 1.0 / 32)) + ti::kRecCap;
  const uint64_t cap_txn = (uint64_t)(L * (cfg.cap_txn_per_byte > 0 ? cfg.cap_txn_per_byte : 1.0 / 8)) + 4096;
  const uint64_t cap_tb = (uint64_t)(L / 40) + 64, cap_warp = (uint64_t)(L / 19) + 64;
  const uint32_t max_chunks = (uint32_t)((max_len + ti::kChunk - 1) / ti::kChunk);
  for (StreamSet &t : ss->set) {
    if ((st = t.text.ensure(((max_len + 15) & ~(size_t)15) + 16)) != TI_OK) return st;
    if ((st = t.inst.ensure(cap_inst * sizeof(ti_inst_rec))) != TI_OK) return st;
    if ((st = t.rows.ensure(cap_mem * 256)) != TI_OK) return st;
    if ((st = t.txn_off.ensure((cap_mem + 1) * 4)) != TI_OK) return st;
    if ((st = t.txn.ensure(cap_txn * sizeof(ti_txn_rec))) != TI_OK) return st;
    if ((st = t.tb.ensure(cap_tb * sizeof(ti_tb_rec))) != TI_OK) return st;
    if ((st = t.warp.ensure(cap_warp * sizeof(ti_warp_rec))) != TI_OK) return st;
    if ((st = t.chains.ensure(ti::chain_bytes(max_chunks))) != TI_OK) return st;
    if ((st = t.globals.ensure(sizeof(ti::Globals))) != TI_OK) return st;
    if ((st = t.counters.ensure(sizeof(ti_counters))) != TI_OK) return st;
    if ((st = t.small.ensure(sizeof(ti::Globals) + sizeof(ti_counters) + 16)) != TI_OK) return st;
    if (compact) {
      const uint32_t max_blocks = (uint32_t)((cap_mem + ti::kCpBlock - 1) / ti::kCpBlock);
      if ((st = t.desc.ensure(cap_mem * sizeof(ti_mem_desc))) != TI_OK) return st;
      if ((st = t.rows_c.ensure(cap_mem * 256)) != TI_OK) return st;
      if ((st = t.txn_c.ensure(cap_txn * sizeof(ti_txn_rec))) != TI_OK) return st;
      if ((st = t.txn_off_c.ensure((cap_mem + 1) * 4)) != TI_OK) return st;
      if ((st = t.part.ensure(((size_t)2 * max_blocks + 2) * 4)) != TI_OK) return st;
    }
  }

  cudaStream_t s_c = c->stream, s_up = ss->s_h2d, s_dn = ss->s_d2h;
#ifdef TI_STREAM_TRACE   // tuning builds: a timeline of the three streams, printed at the end (tools/e2e_diag.py)
  std::vector<cudaEvent_t> tr(6 * (size_t)n_sh + 1);
  for (cudaEvent_t &e : tr) cudaEventCreate(&e);
  cudaEventRecord(tr[6 * (size_t)n_sh], s_c);
  cudaStreamWaitEvent(s_up, tr[6 * (size_t)n_sh], 0);
#define TI_TR(i, j, s) cudaEventRecord(tr[6 * (size_t)(i) + (j)], s)
#else
#define TI_TR(i, j, s) do { } while (0)
#endif
  auto h2d = [&](int i) -> ti_status {   // text of shard i -> its buffer set
    StreamSet &t = ss->set[i & 1];
    const size_t n = cuts[i + 1] - cuts[i];
    if (i >= 2) CK(cudaStreamWaitEvent(s_up, t.ev_dec, 0));   // decode(i-2) has finished reading this buffer
    TI_TR(i, 0, s_up);
    if (n) CK(cudaMemcpyAsync(t.text.p, k->h_body + cuts[i], n, cudaMemcpyHostToDevice, s_up));
    CK(cudaEventRecord(t.ev_h2d, s_up));
    TI_TR(i, 1, s_up);
    return TI_OK;
  };

  uint64_t T_inst = 0, T_mem = 0, T_txn = 0, T_tb = 0, T_warp = 0, T_full = 0, T_keep = 0;
  ti_counters total;
  memset(&total, 0, sizeof total);
  uint32_t flags = 0, first_err = 0xffffffffu;
  float ms_sum = 0;
  bool overflow = false;
  if ((st = h2d(0)) != TI_OK) return st;
  if (n_sh > 1 && (st = h2d(1)) != TI_OK) return st;
  for (int i = 0; i < n_sh && !overflow; ++i) {
    StreamSet &t = ss->set[i & 1];
    const size_t len = cuts[i + 1] - cuts[i];
    const uint32_t n_chunks = (uint32_t)((len + ti::kChunk - 1) / ti::kChunk);
    ti::DecodeParams P;
    memset(&P, 0, sizeof P);
    fill_model_params(P, c, k->col, k->hdr);
    P.text = (const uint8_t *)t.text.p; P.len = len; P.n_chunks = n_chunks;
    P.inst = (ti_inst_rec *)t.inst.p; P.cap_inst = cap_inst;
    P.rows = (uint64_t *)t.rows.p; P.cap_mem = cap_mem;
    P.txn_off = (uint32_t *)t.txn_off.p;
    P.txn = (ti_txn_rec *)t.txn.p; P.cap_txn = cap_txn;
    P.tb = (ti_tb_rec *)t.tb.p; P.cap_tb = cap_tb;
    P.warp = (ti_warp_rec *)t.warp.p; P.cap_warp = cap_warp;
    set_chains(P, t.chains.p, n_chunks);
    P.g = (ti::Globals *)t.globals.p; P.counters = (unsigned long long *)t.counters.p;
    P.off_inst = T_inst; P.off_mem = T_mem; P.off_txn = compact ? 0 : T_txn; P.off_warp = T_warp; P.off_text = cuts[i];
    P.desc = compact ? (ti_mem_desc *)t.desc.p : nullptr;

    CK(cudaStreamWaitEvent(s_c, t.ev_h2d, 0));
    if (i >= 2) CK(cudaStreamWaitEvent(s_c, t.ev_d2h, 0));    // D2H(i-2) has drained this set's outputs
    CK(cudaEventRecord(t.t0, s_c));
    if ((st = reset_launch_state(s_c, t.chains.p, n_chunks, t.globals.p, t.counters.p, t.txn_off.p)) != TI_OK) return st;
    if (n_chunks) {
      const int grid = (int)std::min<uint32_t>((uint32_t)c->grid, n_chunks);
#ifdef TI_EMU
      emu::launch(dim3(grid), dim3(ti::kThreads), ti::kSmemBytes, [=]() { ti::ti_decode_kernel(P); });
      emu::launch(dim3(2), dim3(256), 0, [=]() { ti::ti_check_kernel(P.warp, P.cap_warp, P.counters, P.g, P.off_inst); });
#else
      ti::ti_decode_kernel<<<grid, ti::kThreads, ti::kSmemBytes, s_c>>>(P);
      ti::ti_check_kernel<<<c->grid, 256, 0, s_c>>>(P.warp, P.cap_warp, P.counters, P.g, P.off_inst);
#endif
      CK(cudaGetLastError());
    }
    uint32_t cp_blocks = 0;
    if (compact && !n_chunks) CK(cudaMemsetAsync(t.txn_off_c.p, 0, 4, s_c));   // (an empty body: one CSR sentinel)
    if (compact && n_chunks) {   // gather what still has to travel (ti_compact.cuh); n_mem is read on the device
      ti::CompactParams C;
      memset(&C, 0, sizeof C);
      cp_blocks = (uint32_t)((cap_mem + ti::kCpBlock - 1) / ti::kCpBlock);
      C.desc = (ti_mem_desc *)t.desc.p; C.txn_off = (const uint32_t *)t.txn_off.p; C.rows = (const uint64_t *)t.rows.p;
      C.txn = (const ti_txn_rec *)t.txn.p; C.counters = (const unsigned long long *)t.counters.p; C.cap_mem = cap_mem;
      C.off_txn_full = 0; C.rows_c = (uint64_t *)t.rows_c.p; C.txn_c = (ti_txn_rec *)t.txn_c.p; C.txn_off_c = (uint32_t *)t.txn_off_c.p;
      C.cap_rows_c = cap_mem; C.cap_txn_c = cap_txn; C.part = (uint32_t *)t.part.p; C.n_blocks = cp_blocks;
      C.off_full = T_full; C.off_txn_c = T_keep;
#ifdef TI_EMU
      emu::launch(dim3(cp_blocks), dim3(ti::kCpBlock), 8192, [=]() { ti::cp_count_kernel(C); });
      emu::launch(dim3(1), dim3(ti::kCpBlock), 8192, [=]() { ti::cp_scan_kernel(C); });
      emu::launch(dim3(cp_blocks), dim3(ti::kCpBlock), 8192, [=]() { ti::cp_gather_kernel(C); });
#else
      ti::cp_count_kernel<<<cp_blocks, ti::kCpBlock, 0, s_c>>>(C);
      ti::cp_scan_kernel<<<1, ti::kCpBlock, 0, s_c>>>(C);
      ti::cp_gather_kernel<<<cp_blocks, ti::kCpBlock, 0, s_c>>>(C);
#endif
      CK(cudaGetLastError());
    }
    CK(cudaEventRecord(t.t1, s_c));
    TI_TR(i, 2, s_c);
    ti::Globals *g = (ti::Globals *)t.small.p;
    ti_counters *hc = (ti_counters *)((char *)t.small.p + sizeof(ti::Globals));
    uint32_t *hkeep = (uint32_t *)((char *)t.small.p + sizeof(ti::Globals) + sizeof(ti_counters));   // rows / transactions the compact pass kept
    {
      const ti::Globals *dg = (const ti::Globals *)t.globals.p;
      const unsigned long long *dc = (const unsigned long long *)t.counters.p;
      const uint32_t *dk = cp_blocks ? (const uint32_t *)t.part.p + 2 * (size_t)cp_blocks : nullptr;
      unsigned char *hp = (unsigned char *)t.small.p;
#ifdef TI_EMU
      emu::launch(dim3(1), dim3(256), 0, [=]() { ti::ti_publish_kernel(dg, dc, dk, hp); });
#else
      ti::ti_publish_kernel<<<1, 256, 0, s_c>>>(dg, dc, dk, hp);   // pinned host memory is device-addressable under unified addressing
#endif
      CK(cudaGetLastError());
    }
    CK(cudaEventRecord(t.ev_dec, s_c));
    TI_TR(i, 3, s_c);
    if (i + 2 < n_sh && (st = h2d(i + 2)) != TI_OK) return st;  // queued behind ev_dec of this set
    CK(cudaEventSynchronize(t.ev_dec));                         // the shard's counts are on the host now
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, t.t0, t.t1));
    ms_sum += ms;
    if (g->overflow) { overflow = true; break; }
    flags |= g->error_flags;
    if (g->error_flags && first_err == 0xffffffffu) first_err = g->first_error_off + (uint32_t)cuts[i];   // shards come in file order: the first one wins
    // host arrays: grow when the running totals no longer fit (first shard: extrapolate)
    const double scale = 1.15 * (double)k->body_len / (double)std::max<size_t>(1, cuts[i + 1]);
    const uint64_t add_rows = compact ? hkeep[0] : hc->n_mem, add_txn = compact ? hkeep[1] : hc->n_txn;   // what travels
    const uint64_t at_rows = compact ? T_full : T_mem, at_txn = compact ? T_keep : T_txn;
    struct { HostBuf *b; uint64_t have, add; size_t elt; } need[] = {
        {&k->h_inst, T_inst, hc->n_inst, sizeof(ti_inst_rec)}, {&k->h_rows, at_rows, add_rows, 256},
        {&k->h_txn_off, T_mem, hc->n_mem + 1, 4},               {&k->h_txn, at_txn, add_txn, sizeof(ti_txn_rec)},
        {&k->h_tb, T_tb, hc->n_tb, sizeof(ti_tb_rec)},          {&k->h_warp, T_warp, hc->n_warp, sizeof(ti_warp_rec)},
        {&k->h_desc, compact ? T_mem : 0, compact ? hc->n_mem : 0, sizeof(ti_mem_desc)}};
    for (auto &nd : need) {
      const size_t want = (size_t)(nd.have + nd.add) * nd.elt;
      if (want <= nd.b->bytes) continue;
      CK(cudaStreamSynchronize(s_dn));                          // copies into the old buffer are done
      HostBuf bigger;
      if ((st = bigger.ensure((size_t)((double)want * scale) + 4096)) != TI_OK) return st;
      if (nd.have) memcpy(bigger.p, nd.b->p, (size_t)nd.have * nd.elt);
      nd.b->release();
      *nd.b = bigger;
    }
    // results of shard i -> host, behind its decode
    CK(cudaStreamWaitEvent(s_dn, t.ev_dec, 0));
    TI_TR(i, 4, s_dn);
    auto down = [&](HostBuf &h, uint64_t at, const DevBuf &d, uint64_t n, size_t elt) -> ti_status {
      if (n) CK(cudaMemcpyAsync((char *)h.p + at * elt, d.p, n * elt, cudaMemcpyDeviceToHost, s_dn));
      return TI_OK;
    };
    if ((st = down(k->h_inst, T_inst, t.inst, hc->n_inst, sizeof(ti_inst_rec))) != TI_OK) return st;
    if ((st = down(k->h_rows, at_rows, compact ? t.rows_c : t.rows, add_rows, 256)) != TI_OK) return st;
    if ((st = down(k->h_txn_off, T_mem, compact ? t.txn_off_c : t.txn_off, hc->n_mem + 1, 4)) != TI_OK) return st;
    if ((st = down(k->h_txn, at_txn, compact ? t.txn_c : t.txn, add_txn, sizeof(ti_txn_rec))) != TI_OK) return st;
    if (compact && (st = down(k->h_desc, T_mem, t.desc, hc->n_mem, sizeof(ti_mem_desc))) != TI_OK) return st;
    if ((st = down(k->h_tb, T_tb, t.tb, hc->n_tb, sizeof(ti_tb_rec))) != TI_OK) return st;
    if ((st = down(k->h_warp, T_warp, t.warp, hc->n_warp, sizeof(ti_warp_rec))) != TI_OK) return st;
    CK(cudaEventRecord(t.ev_d2h, s_dn));
    TI_TR(i, 5, s_dn);
    T_inst += hc->n_inst; T_mem += hc->n_mem; T_txn += hc->n_txn; T_tb += hc->n_tb; T_warp += hc->n_warp;
    T_full += hkeep[0]; T_keep += hkeep[1];
    ti_counters_add(&total, hc);
  }
  CK(cudaStreamSynchronize(s_dn));
  CK(cudaStreamSynchronize(s_up));
  CK(cudaStreamSynchronize(s_c));
#ifdef TI_STREAM_TRACE
  if (getenv("TI_STREAM_TRACE")) {
    fprintf(stderr, "shard: upload [start end]  decode kernels end / counts on host  download [start end]  (ms)\n");
    for (int i = 0; i < n_sh; ++i) {
      float v[6];
      for (int j = 0; j < 6; ++j) cudaEventElapsedTime(&v[j], tr[6 * (size_t)n_sh], tr[6 * (size_t)i + j]);
      fprintf(stderr, "%3d: up %6.2f %6.2f  dec(end) %6.2f  counts %6.2f  down %6.2f %6.2f\n", i, v[0], v[1], v[2], v[3], v[4], v[5]);
    }
  }
  for (cudaEvent_t &e : tr) cudaEventDestroy(e);
#endif
  if (overflow) {
    // a capacity guess was too small for one shard: take the one-piece path, which re-decodes with exact sizes
    if ((st = ti_kernel_upload(k, 0, k->body_len)) != TI_OK) return st;
    if ((st = ti_kernel_decode(k)) != TI_OK) return st;
    if ((st = ti_kernel_sync(k, NULL)) != TI_OK) return st;
    if (!compact) return ti_kernel_fetch(k, v);
    // compact caller: hand the full arrays out under trivial descriptors (every row kept, every transaction stored)
    ti_view fv;
    if ((st = ti_kernel_fetch(k, &fv)) != TI_OK) return st;
    if ((st = k->h_desc.ensure((size_t)(fv.counts.n_mem + 1) * sizeof(ti_mem_desc))) != TI_OK) return st;
    ti_mem_desc *hd = (ti_mem_desc *)k->h_desc.p;
    for (uint64_t r = 0; r < fv.counts.n_mem; ++r) { hd[r].base = r; hd[r].stride = 0; hd[r].meta = TI_DESC_FULL_ROW; }
    vc->text = fv.text; vc->inst = fv.inst; vc->desc = hd; vc->rows_full = fv.addr_rows; vc->txn_off = fv.txn_off; vc->txn = fv.txn;
    vc->tb = fv.tb; vc->warp = fv.warp; vc->counts = fv.counts; vc->n_full = fv.counts.n_mem; vc->n_txn_stored = fv.counts.n_txn;
    return TI_OK;
  }
  if (T_inst > 0x7ffffffeull || T_txn > 0xfffffffeull) { ti::set_error("more than 2^31 records in one kernel file"); return TI_ERR_TOO_LARGE; }
  total.n_inst = T_inst; total.n_mem = T_mem; total.n_txn = T_txn; total.n_tb = T_tb; total.n_warp = T_warp;
  total.text_bytes = k->body_len;
  k->host_counters = total;
  memset(&k->counts, 0, sizeof k->counts);
  k->counts.n_inst = T_inst; k->counts.n_mem = T_mem; k->counts.n_txn = T_txn; k->counts.n_tb = T_tb; k->counts.n_warp = T_warp;
  k->counts.text_bytes = k->body_len;
  if (total.n_begin_tb != total.n_end_tb || total.n_begin_tb != T_tb || total.reserved[0] != T_warp) flags |= TI_EF_BAD_STRUCTURE;
  k->counts.error_flags = flags;
  k->counts.first_error_line_off = first_err == 0xffffffffu ? 0 : first_err;
  k->sh_begin = 0; k->sh_len = k->body_len;
  k->decode_ms = ms_sum;
  k->synced = flags == 0; k->decoded = true; k->streamed = true; k->compact = false;   // as ti_kernel_sync: nothing to fetch after a grammar error
  if (flags) { ti::set_error("trace text outside the tracer grammar: flags 0x%x near body offset %u", flags, k->counts.first_error_line_off); return TI_ERR_PARSE; }
  if (compact) {
    k->compact = true;   // ti_kernel_fetch hands out full arrays: there are none after a compact decode
    vc->text = k->h_body;
    vc->inst = (const ti_inst_rec *)k->h_inst.p;
    vc->desc = (const ti_mem_desc *)k->h_desc.p;
    vc->rows_full = (const uint64_t *)k->h_rows.p;
    vc->txn_off = (const uint32_t *)k->h_txn_off.p;
    vc->txn = (const ti_txn_rec *)k->h_txn.p;
    vc->tb = (const ti_tb_rec *)k->h_tb.p;
    vc->warp = (const ti_warp_rec *)k->h_warp.p;
    vc->counts = k->counts;
    vc->n_full = T_full; vc->n_txn_stored = T_keep;
    return TI_OK;
  }
  v->text = k->h_body;
  v->inst = (const ti_inst_rec *)k->h_inst.p;
  v->addr_rows = (const uint64_t *)k->h_rows.p;
  v->txn_off = (const uint32_t *)k->h_txn_off.p;
  v->txn = (const ti_txn_rec *)k->h_txn.p;
  v->tb = (const ti_tb_rec *)k->h_tb.p;
  v->warp = (const ti_warp_rec *)k->h_warp.p;
  v->counts = k->counts;
  return TI_OK;
}

extern "C" ti_status ti_kernel_decode_streamed(ti_kernel *k, size_t shard_bytes, ti_view *v) {
  if (!v) return TI_ERR_INVALID_ARG;
  return decode_streamed_impl(k, shard_bytes, v, nullptr);
}

extern "C" ti_status ti_kernel_decode_streamed_compact(ti_kernel *k, size_t shard_bytes, ti_view_compact *vc) {
  if (!vc) return TI_ERR_INVALID_ARG;
  return decode_streamed_impl(k, shard_bytes, nullptr, vc);
}

extern "C" ti_status ti_kernel_host_body(ti_kernel *k, const char **body, size_t *len) {
  if (!k || !body || !len) return TI_ERR_INVALID_ARG;
  if (!k->h_body && k->body_len) { ti::set_error("the kernel object has no host-resident body"); return TI_ERR_STATE; }
  *body = k->h_body; *len = k->body_len;
  return TI_OK;
}

extern "C" ti_status ti_kernel_release_text(ti_kernel *k, size_t begin, size_t end) {
  if (!k || begin > end || end > k->body_len) return TI_ERR_INVALID_ARG;
#ifndef TI_EMU
  if (k->map_base && end > begin) {   // whole pages inside the range: the page cache keeps them, this process lets go of them
    const size_t page = (size_t)sysconf(_SC_PAGESIZE);
    const uintptr_t lo = ((uintptr_t)(k->h_body + begin) + page - 1) & ~(uintptr_t)(page - 1), hi = (uintptr_t)(k->h_body + end) & ~(uintptr_t)(page - 1);
    if (hi > lo) madvise((void *)lo, hi - lo, MADV_DONTNEED);
  }
#endif
  return TI_OK;
}

extern "C" ti_status ti_kernel_decode_window_compact(ti_kernel *k, size_t begin, size_t end, size_t shard_bytes, ti_view_compact *vc) {
  if (!k || !vc || begin > end || end > k->body_len) return TI_ERR_INVALID_ARG;
  if (!k->h_body && k->body_len) { ti::set_error("ti_kernel_decode_window_compact needs a host-resident body"); return TI_ERR_STATE; }
  // the window is decoded as a body of its own: every index and text offset in the view counts from `begin`
  const char *body = k->h_body;
  const size_t len = k->body_len;
  k->h_body = body + begin; k->body_len = end - begin;
  const ti_status st = decode_streamed_impl(k, shard_bytes, nullptr, vc);
  k->h_body = body; k->body_len = len;
  return st;
}

// =====================================================================================
// ti_decode_noalloc: the same fused decode, for callers that cannot let the library
// allocate (see include/traceingest.h).  It has no counterpart in the reference.
//
// The only memory the decode needs beyond its inputs and outputs is the set of
// per-chunk look-back tables.  For this entry point they live in the file-scope device
// array below instead of a cudaMalloc'ed buffer: (TI_NOALLOC_MAX_TEXT / chunk + 2)
// entries of 32 bytes, about 0.5 MB, resident on every GPU that loads the library
// for as long as the process lives.
//
// In host code the name `g_noalloc_scratch` is only a handle for the symbol, not an
// address the GPU can use, so the device address is looked up once per device with
// cudaGetSymbolAddress while that device is current and cached; kernels receive the
// looked-up pointer.  That first call per device is the set-up: do it before any stream
// capture starts.  All calls share the array.  Each call zeroes the entries it is about
// to use (the tables must start at "nothing published") and leaves the rest as the last
// call left them, so two calls whose GPU work can overlap would corrupt each other's
// offsets -- serialise them.
// =====================================================================================
namespace {
constexpr size_t kNoallocChunks = (size_t)(TI_NOALLOC_MAX_TEXT / ti::kChunk) + 2;
constexpr size_t kNoallocBytes = ti::chain_bytes(kNoallocChunks - 1);
static_assert(ti::kChunk != 18432 || kNoallocBytes == TI_NOALLOC_SCRATCH_BYTES,  // (tuning builds use other chunk sizes)
              "TI_NOALLOC_SCRATCH_BYTES in traceingest.h must state the real size of the scratch array");
constexpr int kMaxDevices = 64;
void *g_noalloc_addr[kMaxDevices];  // device address of the array, per device ordinal (0 = not looked up yet)
}  // namespace
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoallocBytes];

extern "C" ti_status ti_decode_noalloc(ti_ctx *c, const ti_noalloc_io *io, void *stream) {
  if (!c || !io) return TI_ERR_INVALID_ARG;
  if ((!io->d_text && io->len) || ((uintptr_t)io->d_text & 15)) { ti::set_error("d_text must be a 16-byte aligned device pointer"); return TI_ERR_INVALID_ARG; }
  if (!io->d_inst || !io->d_rows || !io->d_txn_off || !io->d_txn || !io->d_tb || !io->d_warp || !io->d_counters || !io->d_status) {
    ti::set_error("ti_decode_noalloc: every output pointer must be supplied by the caller");
    return TI_ERR_INVALID_ARG;
  }
  if (io->len > TI_NOALLOC_MAX_TEXT) {
    ti::set_error("ti_decode_noalloc: %llu bytes of text exceed the built-in scratch, which covers %llu bytes; cut the shard with ti_plan_shards",
                  (unsigned long long)io->len, (unsigned long long)TI_NOALLOC_MAX_TEXT);
    return TI_ERR_TOO_LARGE;
  }
  const int col = arch_column(io->binary_version);
  if (col < 0) { ti::set_error("unsupported binary version: %u", io->binary_version); return TI_ERR_BINARY_VERSION; }
  const ti_config &cfg = c->cfg;
  const int dev = cfg.device;
  if (dev < 0 || dev >= kMaxDevices) return TI_ERR_INVALID_ARG;
  CK(cudaSetDevice(dev));
  if (!g_noalloc_addr[dev]) {  // set-up, once per device, with that device current
    void *p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_addr[dev] = p;
  }
  cudaStream_t s = stream ? (cudaStream_t)stream : c->stream;
  const uint32_t n_chunks = (uint32_t)((io->len + ti::kChunk - 1) / ti::kChunk);
  unsigned char *scratch = (unsigned char *)g_noalloc_addr[dev];

  ti::DecodeParams P;
  memset(&P, 0, sizeof P);
  P.text = (const uint8_t *)io->d_text;
  P.len = io->len;
  P.n_chunks = n_chunks;
  P.trace_version = io->trace_version;
  P.shmem_base = io->shmem_base_addr; P.local_base = io->local_base_addr;
  P.inst = (ti_inst_rec *)io->d_inst; P.cap_inst = io->cap_inst;
  P.rows = (uint64_t *)io->d_rows; P.cap_mem = io->cap_mem;
  P.txn_off = (uint32_t *)io->d_txn_off;
  P.txn = (ti_txn_rec *)io->d_txn; P.cap_txn = io->cap_txn;
  P.tb = (ti_tb_rec *)io->d_tb; P.cap_tb = io->cap_tb;
  P.warp = (ti_warp_rec *)io->d_warp; P.cap_warp = io->cap_warp;
  set_chains(P, scratch, n_chunks);
  P.g = (ti::Globals *)io->d_status;
  P.counters = (unsigned long long *)io->d_counters;
  P.isa = c->d_isa[col]; P.spop = c->d_spop;
  for (unsigned slot = 0; slot < 32; ++slot) {  // trace_driven.cc:441-480
    const unsigned op = slot >= 18 ? 100 + (slot - 18) : slot;
    const uint32_t *l = nullptr;
    switch (op) {
      case ti::OP_ALU: case ti::OP_INTP: case ti::OP_BRANCH: case ti::OP_CALL: case ti::OP_RET: l = cfg.lat_int; break;
      case ti::OP_SP: l = cfg.lat_sp; break;
      case ti::OP_DP: l = cfg.lat_dp; break;
      case ti::OP_SFU: l = cfg.lat_sfu; break;
      case ti::OP_TENSOR: l = cfg.lat_tensor; break;
    }
    if (op >= 100 && op < 108) l = cfg.lat_spec[op - 100];
    P.lat[slot][0] = l ? (uint16_t)l[0] : 1; P.lat[slot][1] = l ? (uint16_t)l[1] : 1;
  }
  P.coalesce_arch = cfg.coalesce_arch; P.mem_warp_parts = cfg.mem_warp_parts; P.num_shmem_bank = cfg.num_shmem_bank;
  P.shmem_limited_broadcast = cfg.shmem_limited_broadcast; P.gmem_skip_L1D = cfg.gmem_skip_L1D;

  // what must start at zero: the table entries this call uses, the status block, the counters
  { const ti_status rs = reset_launch_state(s, scratch, n_chunks, io->d_status, io->d_counters, io->d_txn_off); if (rs != TI_OK) return rs; }
  if (n_chunks) {
    const int grid = (int)std::min<uint32_t>((uint32_t)c->grid, n_chunks);
#ifdef TI_EMU
    emu::launch(dim3(grid), dim3(ti::kThreads), ti::kSmemBytes, [=]() { ti::ti_decode_kernel(P); });
    emu::launch(dim3(2), dim3(256), 0, [=]() { ti::ti_check_kernel(P.warp, P.cap_warp, P.counters, P.g); });
#else
    ti::ti_decode_kernel<<<grid, ti::kThreads, ti::kSmemBytes, s>>>(P);
    ti::ti_check_kernel<<<c->grid, 256, 0, s>>>(P.warp, P.cap_warp, P.counters, P.g);
#endif
    CK(cudaGetLastError());
  }
  return TI_OK;
}

#include "ti_nccl.inc"
#include "ti_pstats.inc"
#include "ti_compress_api.inc"
#include "ti_regroup_api.inc"
#include "ti_local_api.inc"

#if defined(TI_PHASE_TIMING) && !defined(TI_EMU)
// tuning builds only: read and reset the per-phase cycle sums
extern "C" int ti_debug_phase_cycles(unsigned long long *out16) {
  if (cudaMemcpyFromSymbol(out16, ti::g_phase_cycles, sizeof(unsigned long long) * 16) != cudaSuccess) return 1;
  unsigned long long z[16] = {0};
  return cudaMemcpyToSymbol(ti::g_phase_cycles, z, sizeof z) != cudaSuccess;
}
#endif
