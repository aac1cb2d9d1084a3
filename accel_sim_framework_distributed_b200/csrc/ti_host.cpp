// Host-only half of the C ABI: everything on the path that the reference also does
// once per kernel / per command and that has no data parallelism worth a launch --
// the command list, the memcpy line, the kernel header, the shard planner and the
// floating-point finalisation of the counter block.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>

#include "../../include/traceingest.h"
#include "ti_internal.h"

namespace ti {
thread_local std::string g_last_error;
void set_error(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_last_error = buf;
}
}  // namespace ti

extern "C" {

const char *ti_strerror(ti_status s) {
  switch (s) {
    case TI_OK: return "ok";
    case TI_ERR_INVALID_ARG: return "invalid argument";
    case TI_ERR_CUDA: return "CUDA runtime error";
    case TI_ERR_IO: return "unable to open or read file";
    case TI_ERR_BINARY_VERSION: return "unsupported binary version";
    case TI_ERR_PARSE: return "trace text outside the tracer grammar";
    case TI_ERR_CAPACITY: return "output capacity exceeded";
    case TI_ERR_TOO_LARGE: return "shard too large (text > 4 GiB or > 2^32-2 records)";
    case TI_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU path)";
    case TI_ERR_STATE: return "call sequence violated";
  }
  return "unknown status";
}

const char *ti_last_error(void) { return ti::g_last_error.c_str(); }

// SM7_QV100: gpgpusim.config:68,93,126-129,155; trace.config:1-19;
// accelwattch_sass_sim.xml:48-59 (absent coefficients stay 1.0, XML_Parse.cc:4523)
void ti_config_qv100(ti_config *c) {
  memset(c, 0, sizeof *c);
  c->coalesce_arch = 70;
  c->mem_warp_parts = 1;
  c->num_shmem_bank = 32;
  c->clock_gated_lanes = 1;
  c->num_dp_units = 4;
  const uint32_t two[2] = {2, 2};
  memcpy(c->lat_int, two, 8); memcpy(c->lat_sp, two, 8); memcpy(c->lat_tensor, two, 8);
  c->lat_dp[0] = 8; c->lat_dp[1] = 4;
  c->lat_sfu[0] = 20; c->lat_sfu[1] = 8;
  for (int j = 0; j < 8; ++j) c->lat_spec[j][0] = c->lat_spec[j][1] = 4;
  c->lat_spec[1][0] = 200;
  c->lat_spec[2][0] = c->lat_spec[2][1] = 2;
  for (int j = 0; j < TI_N_COEFF; ++j) c->coeff[j] = 1.0;
  c->coeff[TI_C_INT] = 14.98768151;    c->coeff[TI_C_FP] = 0.529670751;   c->coeff[TI_C_DP] = 0.777229051;
  c->coeff[TI_C_INT_MUL] = 0.115098047; c->coeff[TI_C_FP_MUL] = 0.089517055; c->coeff[TI_C_SQRT] = 0.195089274;
  c->coeff[TI_C_LOG] = 0.125521663;    c->coeff[TI_C_SIN] = 0.13336307;   c->coeff[TI_C_EXP] = 0.36204415;
  c->coeff[TI_C_DP_MUL] = 0.1321288;   c->coeff[TI_C_TENSOR] = 0.815454621; c->coeff[TI_C_TEX] = 0.115100088;
}

// trace_parser.cc:299-377.  Header lines start with '-'; the first line starting
// with '#' ends the header and the body begins right after it.
ti_status ti_parse_kernel_header(const char *t, size_t len, ti_kernel_hdr *h) {
  if (!t || !h) return TI_ERR_INVALID_ARG;
  memset(h, 0, sizeof *h);
  snprintf(h->kernel_name, sizeof h->kernel_name, "Empty");  // kernel_trace_t ctor, trace_parser.cc:78-84
  size_t pos = 0;
  h->body_offset = len;
  while (pos < len) {
    const char *nl = (const char *)memchr(t + pos, '\n', len - pos);
    const size_t e = nl ? (size_t)(nl - t) : len;
    std::string line(t + pos, e - pos);
    pos = e + 1;
    if (line.empty()) continue;
    if (line[0] == '#') { h->body_offset = pos < len ? pos : len; break; }
    if (line[0] != '-') continue;
    // the reference keys on the first two words after '-'
    char w1[32] = "", w2[32] = "";
    sscanf(line.c_str() + 1, "%31s %31s", w1, w2);
    const size_t eq = line.find('=');
    const char *s = line.c_str();
    const std::string a = w1, b = w2;
    if (a == "kernel" && b == "name") snprintf(h->kernel_name, sizeof h->kernel_name, "%s", eq != std::string::npos && eq + 2 <= line.size() ? s + eq + 2 : "");
    else if (a == "kernel" && b == "id") sscanf(s, "-kernel id = %u", &h->kernel_id);
    else if (a == "grid" && b == "dim") sscanf(s, "-grid dim = (%u,%u,%u)", &h->grid_dim[0], &h->grid_dim[1], &h->grid_dim[2]);
    else if (a == "block" && b == "dim") sscanf(s, "-block dim = (%u,%u,%u)", &h->tb_dim[0], &h->tb_dim[1], &h->tb_dim[2]);
    else if (a == "shmem" && b == "=") sscanf(s, "-shmem = %u", &h->shmem);
    else if (a == "nregs") sscanf(s, "-nregs = %u", &h->nregs);
    else if (a == "cuda" && b == "stream") { unsigned long v = 0; sscanf(s, "-cuda stream id = %lu", &v); h->cuda_stream_id = v; }
    else if (a == "binary" && b == "version") sscanf(s, "-binary version = %u", &h->binary_version);
    else if (a == "nvbit" && b == "version") snprintf(h->nvbit_version, sizeof h->nvbit_version, "%s", eq != std::string::npos ? s + eq + 1 : "");
    else if (a == "accelsim" && b == "tracer") sscanf(s, "-accelsim tracer version = %u", &h->trace_version);
    else if (a == "shmem" && b == "base_addr" && eq != std::string::npos) h->shmem_base_addr = strtoull(s + eq + 1, NULL, 16);
    else if (a == "local" && b == "mem" && eq != std::string::npos) h->local_base_addr = strtoull(s + eq + 1, NULL, 16);
  }
  return TI_OK;
}

// trace_parser.cc:220-284: prefix match per line; unknown lines are skipped.
ti_status ti_parse_commandlist(const char *t, size_t len, ti_command *out, size_t cap, size_t *n) {
  if (!t || !n) return TI_ERR_INVALID_ARG;
  static const struct { const char *pfx; int kind; } kinds[] = {
      {"MemcpyHtoD", 2}, {"kernel", 1}, {"ncclCommInitAll", 4}, {"ncclCommDestroy", 5},
      {"ncclGroupStart", 6}, {"ncclGroupEnd", 7}, {"ncclAllReduce", 8}};
  size_t pos = 0, cnt = 0;
  while (pos < len) {
    const char *nl = (const char *)memchr(t + pos, '\n', len - pos);
    const size_t e = nl ? (size_t)(nl - t) : len;
    const size_t ll = e - pos;
    for (const auto &k : kinds) {
      const size_t pl = strlen(k.pfx);
      // the reference compares substr(0, 13) with "ncclGroupEnd" (12 chars): that never matches a longer line
      const size_t cmp = k.kind == 7 ? 13 : pl;
      const size_t have = ll < cmp ? ll : cmp;
      if (have == pl && !memcmp(t + pos, k.pfx, pl) && (k.kind != 7 || ll == pl)) {
        if (out && cnt < cap) { out[cnt].kind = k.kind; out[cnt].line_off = (uint32_t)pos; out[cnt].line_len = (uint32_t)ll; }
        ++cnt;
        break;
      }
    }
    pos = e + 1;
  }
  *n = cnt;
  return out && cnt > cap ? TI_ERR_CAPACITY : TI_OK;
}

// trace_parser.cc:286-297: "MemcpyHtoD,0x<addr>,<count>"
ti_status ti_parse_memcpy(const char *cmd, size_t len, uint64_t *addr, uint64_t *count) {
  if (!cmd || !addr || !count) return TI_ERR_INVALID_ARG;
  std::string s(cmd, len);
  const size_t c1 = s.find(','), c2 = c1 == std::string::npos ? c1 : s.find(',', c1 + 1);
  if (c2 == std::string::npos || s.find(',', c2 + 1) != std::string::npos) {
    ti::set_error("memcpy command needs exactly 3 fields (trace_parser.cc:290 asserts)");
    return TI_ERR_PARSE;
  }
  *addr = strtoull(s.c_str() + c1 + 1, NULL, 16);
  *count = strtoull(s.c_str() + c2 + 1, NULL, 10);
  return TI_OK;
}

// Cut points at "\n#BEGIN_TB": each shard owns whole thread blocks and starts on the
// blank line PTP writes before "#BEGIN_TB" (post-traces-processing.cpp:163).
ti_status ti_plan_shards(const char *body, size_t len, int n, size_t *cuts) {
  if (!body || n < 1 || !cuts) return TI_ERR_INVALID_ARG;
  cuts[0] = 0;
  cuts[n] = len;
  static const char key[] = "\n#BEGIN_TB";
  for (int k = 1; k < n; ++k) {
    size_t nominal = (size_t)((unsigned __int128)len * k / n);
    if (nominal < cuts[k - 1]) nominal = cuts[k - 1];
    const char *hit = nominal < len ? (const char *)memmem(body + nominal, len - nominal, key, sizeof key - 1) : NULL;
    cuts[k] = hit ? (size_t)(hit - body) + 1 : len;
  }
  return TI_OK;
}

// ---- host-side expansion of the compact view (ti_view_compact): the same closed forms the kernel uses.
// Row: trace_parser.cc:86-105 (base_stride_decompress: base + i * stride over the mask's first run, 0 elsewhere).
// Dense transactions: abstract_hardware_model.cc:509-586 on a contiguous, aligned run + :694-748 (size reduction).
void ti_expand_row(const ti_view_compact *v, uint64_t mem_idx, uint64_t out[32]) {
  const ti_mem_desc &d = v->desc[mem_idx];
  if (d.meta & TI_DESC_FULL_ROW) { memcpy(out, v->rows_full + d.base * 32, 256); return; }
  const unsigned first = d.meta & 31, n = d.meta >> 5 & 63;
  for (unsigned l = 0; l < 32; ++l)
    out[l] = l - first < n ? d.base + (uint64_t)(int64_t)d.stride * (uint64_t)(l - first) : 0;
}

uint32_t ti_expand_txns(const ti_view_compact *v, uint64_t mem_idx, ti_txn_rec *out) {
  const ti_mem_desc &d = v->desc[mem_idx];
  if (!(d.meta & TI_DESC_DENSE)) {
    const uint32_t b = v->txn_off[mem_idx], e = v->txn_off[mem_idx + 1];
    if (e > b) memcpy(out, v->txn + b, (size_t)(e - b) * sizeof(ti_txn_rec));
    return e - b;
  }
  const unsigned first = d.meta & 31, n = d.meta >> 5 & 63, ds = 1u << (d.meta >> 11 & 7), seg = 32u << (d.meta >> 14 & 3), wr = d.meta >> 16 & 1;
  const unsigned sh = d.meta >> 11 & 7;
  const uint64_t segmask = ~(uint64_t)(seg - 1), end = d.base + (uint64_t)n * ds;
  uint32_t k = 0;
  for (uint64_t S = d.base & segmask; S < end; S += seg, ++k) {
    const uint64_t lo = S > d.base ? S : d.base, hi = S + seg < end ? S + seg : end;
    const unsigned i0 = (unsigned)((lo - d.base) >> sh), i1 = (unsigned)((hi - d.base) >> sh);   // lanes [i0, i1) of the run
    const unsigned p0 = (unsigned)(lo & 127), p1 = p0 + (unsigned)(hi - lo);                   // bytes [p0, p1) of the 128 B line
    ti_txn_rec r;
    memset(&r, 0, sizeof r);
    r.lane_mask = (i1 - i0 == 32 ? 0xffffffffu : ((1u << (i1 - i0)) - 1)) << (first + i0);
    unsigned chunks = 0;
    for (unsigned w = 0; w < 4; ++w) {
      const unsigned a = p0 > 32 * w ? p0 : 32 * w, b = p1 < 32 * w + 32 ? p1 : 32 * w + 32;
      if (b > a) { r.byte_mask[w] = (b - a == 32 ? 0xffffffffu : ((1u << (b - a)) - 1)) << (a - 32 * w); chunks |= 1u << w; }
    }
    // memory_coalescing_arch_reduce_and_send (abstract_hardware_model.cc:694-748)
    uint64_t addr = S;
    unsigned size = seg, h = 0;
    if (seg == 128) {
      const bool l2 = chunks & 3, h2 = chunks & 12;
      if (l2 && !h2) { size = 64; h = chunks & 3; }
      else if (!l2 && h2) { addr += 64; size = 64; h = (chunks >> 2) & 3; }
    } else if (seg == 64) {
      h = (addr % 128 == 0) ? (chunks & 3) : ((chunks >> 2) & 3);
    }
    if (size == 64) {
      const bool l1 = h & 1, h1 = h & 2;
      if (l1 && !h1) size = 32;
      else if (!l1 && h1) { addr += 32; size = 32; }
    }
    r.addr = addr; r.size = (uint8_t)size; r.sector_mask = (uint8_t)chunks; r.type = wr ? 4 : 0; r.is_write = (uint8_t)wr;
    out[k] = r;
  }
  return k;
}

void ti_counters_add(ti_counters *acc, const ti_counters *x) {
  uint64_t *a = (uint64_t *)acc;
  const uint64_t *b = (const uint64_t *)x;
  for (size_t i = 0; i < sizeof(ti_counters) / 8; ++i) a[i] += b[i];
}

}  // extern "C"

// shader.h:2399-2407 -- both helpers return `unsigned`: one truncation of the double sum
static unsigned inactive_sfu(unsigned a, double c) {
  return (unsigned)((((32 - a) >> 1) * c) + (((32 - a) >> 3) * c) + (((32 - a) >> 3) * c));
}
static unsigned inactive_nonsfu(unsigned a, double c) { return (unsigned)(((32 - a) >> 1) * c); }

ti_status ti_counters_finalize_cfg(const ti_config *cfg, const ti_counters *c, ti_power_counters *o) {
  if (!cfg || !c || !o) return TI_ERR_INVALID_ARG;
  memset(o, 0, sizeof *o);
  // special_ops value -> (output slot, coefficient, uses the SFU-style inactive-lane term)
  struct Cls { int sp; double *dst; int coeff; bool sfu; };
  const Cls cls[] = {
      {1, &o->ialu, TI_C_INT, false},      {4, &o->imul, TI_C_INT_MUL, false},   {2, &o->imul24, TI_C_INT_MUL24, false},
      {3, &o->imul32, TI_C_INT_MUL32, true}, {5, &o->idiv, TI_C_INT_DIV, true},  {8, &o->fpalu, TI_C_FP, false},
      {6, &o->fpmul, TI_C_FP_MUL, false},  {7, &o->fpdiv, TI_C_FP_DIV, true},    {15, &o->dpalu, TI_C_DP, false},
      {13, &o->dpmul, TI_C_DP_MUL, false}, {14, &o->dpdiv, TI_C_DP_DIV, true},   {9, &o->sqrt_, TI_C_SQRT, true},
      {10, &o->log_, TI_C_LOG, true},      {11, &o->sin_, TI_C_SIN, true},       {12, &o->exp_, TI_C_EXP, true},
      {16, &o->tensor, TI_C_TENSOR, true}, {17, &o->tex, TI_C_TEX, true}};
  for (const Cls &k : cls) {
    const double coeff = cfg->coeff[k.coeff];
    double sum = 0;
    for (unsigned a = 0; a <= 32; ++a) {
      const uint64_t n = c->hist[k.sp][a];
      if (!n) continue;
      double per = (double)a * coeff;  // shader.h:2161-2337
      if (!cfg->clock_gated_lanes) per += (double)(k.sfu ? inactive_sfu(a, coeff) : inactive_nonsfu(a, coeff));
      sum += (double)n * per;
      o->active_exu_threads += (double)n * a;
      o->active_exu_warps += (double)n;
    }
    *k.dst = sum;
  }
  // unit routing, shader.cc:1336-1497, with each unit's inc*_stat(warp_size, latency)
  for (unsigned slot = 1; slot < 26; ++slot) {
    const uint64_t n = c->op_count[slot];
    if (!n) continue;
    const unsigned op = slot >= 18 ? 100 + (slot - 18) : slot;
    unsigned lat = 1;
    switch (op) {  // trace_driven.cc:441-480
      case 1: case 6: case 12: case 15: case 16: lat = cfg->lat_int[0]; break;
      case 5: lat = cfg->lat_sp[0]; break;
      case 4: lat = cfg->lat_dp[0]; break;
      case 2: lat = cfg->lat_sfu[0]; break;
      case 3: lat = cfg->lat_tensor[0]; break;
    }
    if (op >= 100) lat = cfg->lat_spec[op - 100][0];
    const bool to_mem = op == 8 || op == 11 || op == 14 || op == 9 || op == 10;
    if (to_mem) o->mem_accesses += (double)n * 32.0;                 // incmem_stat(32, 1)
    else if (op == 3 || op == 2 || (op == 4 && cfg->num_dp_units == 0)) o->sfu_accesses += (double)n * 32.0 * lat;
    else o->sp_accesses += (double)n * 32.0 * lat;
  }
  o->const_accesses = (double)c->const_operand_insts;
  // mem_instruction_stats, gpu-sim.cc:1548-1582
  o->n_shmem_insn = (double)(c->space_lanes[3][0] + c->space_lanes[3][1]);
  o->n_sstarr_insn = (double)(c->space_lanes[4][0] + c->space_lanes[4][1]);
  o->n_const_insn = (double)(c->space_lanes[8][0] + c->space_lanes[8][1]);
  o->n_param_insn = (double)(c->space_lanes[6][0] + c->space_lanes[6][1] + c->space_lanes[7][0] + c->space_lanes[7][1]);
  o->n_tex_insn = (double)(c->space_lanes[9][0] + c->space_lanes[9][1]);
  o->n_load_insn = (double)(c->space_lanes[11][0] + c->space_lanes[2][0]);
  o->n_store_insn = (double)(c->space_lanes[11][1] + c->space_lanes[2][1]);
  return TI_OK;
}

// post-traces-processing.cpp:44-83: kernelslist -> kernelslist.g
extern "C" ti_status ti_regroup_commandlist(const char *text, size_t len, char *out, size_t cap, size_t *out_len) {
  if ((!text && len) || !out_len) return TI_ERR_INVALID_ARG;
  size_t need = 0, p = 0;
  auto put = [&](const char *s, size_t n) { if (out && need + n <= cap) memcpy(out + need, s, n); need += n; };
  // while (!eof) getline: every '\n'-terminated piece plus the (possibly empty) rest
  while (p <= len) {
    const char *nl = p < len ? (const char *)memchr(text + p, '\n', len - p) : nullptr;
    const size_t e = nl ? (size_t)(nl - text) : len;
    const size_t n = e - p;
    if (n) {
      const char *l = text + p;
      if (n >= 6 && !memcmp(l, "Memcpy", 6)) { put(l, n); put("\n", 1); }
      else if (n >= 6 && !memcmp(l, "kernel", 6)) { put(l, n); put("g\n", 2); }
      else if (n >= 4 && !memcmp(l, "nccl", 4)) { put(l, n); put("\n", 1); }
      else { ti::set_error("Undefined command: %.*s", (int)std::min<size_t>(n, 200), l); *out_len = need; return TI_ERR_PARSE; }
    }
    if (!nl) break;
    p = e + 1;
  }
  *out_len = need;
  if (need > cap || (!out && need)) { ti::set_error("kernelslist.g needs %zu bytes", need); return TI_ERR_CAPACITY; }
  return TI_OK;
}

// the statistics that are functions of the counter block
extern "C" ti_status ti_counters_derive(const ti_counters *c, ti_derived_stats *out) {
  if (!c || !out) return TI_ERR_INVALID_ARG;
  memset(out, 0, sizeof *out);
  for (int sp = 0; sp < TI_N_SP_OPS; ++sp)
    for (int a = 0; a <= 32; ++a) out->active_lane_distro[a] += c->hist[sp][a];   // shader.cc:1045, one count per issued warp instruction
  out->smem_bank_conflict = c->smem_cycles;   // abstract_hardware_model.cc:415-418: cycles = total_accesses, reported as is
  out->uncoalesced_gmem = c->n_txn;           // abstract_hardware_model.cc:467-470: every access pushed for a global instruction
  return TI_OK;
}
