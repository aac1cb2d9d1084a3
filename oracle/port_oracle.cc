// TEST INFRASTRUCTURE -- not a product path (see oracle/README.md).
//
// Stand-alone CPU oracle built from the restatement in port_core.h / port_model.h.
//   port_oracle dump <kernel.traceg> <out.tidump> [cfg.bin]
//   port_oracle time <kernel.traceg> [cfg.bin] [parse|convert|coalesce]
// Output container: oracle/dump_format.h (same as oracle/_ref/ref_oracle).
// Exit status 5 and "error_flags" in the JSON line when a TI_EF_* flag was raised.
#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "port_model.h"
#include "qv100_cfg.h"

using namespace ti_port;

static bool read_file(const char *path, std::string &out) {
  FILE *f = fopen(path, "rb");
  if (!f) return false;
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  out.resize((size_t)n);
  size_t got = n ? fread(&out[0], 1, (size_t)n, f) : 0;
  fclose(f);
  return got == (size_t)n;
}

// TP:299-377 -- "-key = value" lines up to the first line starting with '#'
static bool parse_header(const std::string &t, ti_kernel_hdr &h) {
  memset(&h, 0, sizeof h);
  snprintf(h.kernel_name, sizeof h.kernel_name, "Empty");
  size_t pos = 0;
  while (pos < t.size()) {
    size_t nl = t.find('\n', pos);
    if (nl == std::string::npos) nl = t.size();
    std::string line = t.substr(pos, nl - pos);
    pos = nl + 1;
    if (line.empty()) continue;
    if (line[0] == '#') { h.body_offset = std::min(pos, t.size()); return true; }
    if (line[0] != '-') continue;
    const char *s = line.c_str();
    size_t eq = line.find('=');
    if (!line.compare(0, 13, "-kernel name ")) snprintf(h.kernel_name, sizeof h.kernel_name, "%s", eq + 2 <= line.size() ? s + eq + 2 : "");
    else if (!line.compare(0, 11, "-kernel id ")) sscanf(s, "-kernel id = %u", &h.kernel_id);
    else if (!line.compare(0, 10, "-grid dim ")) sscanf(s, "-grid dim = (%u,%u,%u)", &h.grid_dim[0], &h.grid_dim[1], &h.grid_dim[2]);
    else if (!line.compare(0, 11, "-block dim ")) sscanf(s, "-block dim = (%u,%u,%u)", &h.tb_dim[0], &h.tb_dim[1], &h.tb_dim[2]);
    else if (!line.compare(0, 8, "-shmem =")) sscanf(s, "-shmem = %u", &h.shmem);
    else if (!line.compare(0, 6, "-nregs")) sscanf(s, "-nregs = %u", &h.nregs);
    else if (!line.compare(0, 13, "-cuda stream ")) { unsigned long v = 0; sscanf(s, "-cuda stream id = %lu", &v); h.cuda_stream_id = v; }
    else if (!line.compare(0, 16, "-binary version ")) sscanf(s, "-binary version = %u", &h.binary_version);
    else if (!line.compare(0, 15, "-nvbit version ")) snprintf(h.nvbit_version, sizeof h.nvbit_version, "%s", s + eq + 1);
    else if (!line.compare(0, 17, "-accelsim tracer ")) sscanf(s, "-accelsim tracer version = %u", &h.trace_version);
    else if (!line.compare(0, 17, "-shmem base_addr ")) h.shmem_base_addr = strtoull(s + eq + 1, NULL, 16);
    else if (!line.compare(0, 11, "-local mem ")) h.local_base_addr = strtoull(s + eq + 1, NULL, 16);
  }
  h.body_offset = t.size();
  return true;
}

static bool exact_int(const char *&p, const char *e, int32_t &v) {
  bool neg = false;
  if (p < e && *p == '-') { neg = true; ++p; }
  if (p >= e || *p < '0' || *p > '9') return false;
  int64_t m = 0;
  int nd = 0;
  while (p < e && *p >= '0' && *p <= '9') { m = m * 10 + (*p++ - '0'); if (++nd > 10) return false; }
  if (m > 0x7fffffffLL + (neg ? 1 : 0)) return false;
  v = (int32_t)(neg ? -m : m);
  return true;
}

static bool lit(const char *&p, const char *e, const char *s) {
  size_t n = strlen(s);
  if ((size_t)(e - p) < n || memcmp(p, s, n)) return false;
  p += n;
  return true;
}

int main(int argc, char **argv) {
  if (argc < 3) { fprintf(stderr, "usage: port_oracle dump|time <traceg> ...\n"); return 2; }
  std::string mode = argv[1];
  const char *out_path = NULL, *cfg_path = NULL;
  int stages = 3;
  if (mode == "dump") { if (argc < 4) return 2; out_path = argv[3]; if (argc > 4) cfg_path = argv[4]; }
  else { if (argc > 3) cfg_path = argv[3]; if (argc > 4) stages = !strcmp(argv[4], "parse") ? 1 : !strcmp(argv[4], "convert") ? 2 : 3; }
  ti_config cfg;
  ti_oracle::qv100(&cfg);
  if (cfg_path && *cfg_path) {
    FILE *f = fopen(cfg_path, "rb");
    if (!f || fread(&cfg, sizeof cfg, 1, f) != 1) { fprintf(stderr, "bad cfg\n"); return 2; }
    fclose(f);
  }
  std::string text;
  if (!read_file(argv[2], text)) { printf("Unable to open file: %s\n", argv[2]); return 1; }  // TP:306-309
  auto t0 = std::chrono::steady_clock::now();
  ti_kernel_hdr hdr;
  parse_header(text, hdr);
  IsaView isa;
  if (!isa.init(hdr.binary_version)) { printf("unsupported binary version: %d\n", hdr.binary_version); return 4; }

  const char *body = text.data() + hdr.body_offset;
  const size_t len = text.size() - hdr.body_offset;
  const bool dump = mode == "dump";
  ti_dump d;
  d.txn_off.push_back(0);
  ti_oracle::counter_acc acc;
  uint32_t flags = 0;
  uint64_t n_inst = 0, n_mem = 0, n_txn = 0, n_tb = 0, n_warp = 0, n_begin = 0, n_end = 0;
  std::vector<ti_warp_rec> warps;  // kept in both modes for the framing check
  std::vector<ti_txn_rec> q;
  bool expect_insts = false;

  size_t pos = 0;
  while (pos < len) {  // TP:401-446
    const char *b = body + pos;
    const char *nl = (const char *)memchr(b, '\n', len - pos);
    const char *e = nl ? nl : body + len;
    pos = (size_t)(e - body) + 1;
    if (e == b) continue;                                 // TP:408
    if (e - b > TI_MAX_LINE) { flags |= TI_EF_LINE_TOO_LONG; continue; }
    const char c0 = *b;
    if (expect_insts && c0 != 'i') flags |= TI_EF_BAD_STRUCTURE;
    const char *p = b;
    if (c0 == '#') {
      if (e - b == 9 && !memcmp(b, "#BEGIN_TB", 9)) n_begin++;
      else if (e - b == 7 && !memcmp(b, "#END_TB", 7)) n_end++;
      else flags |= TI_EF_BAD_STRUCTURE;
    } else if (c0 == 't') {                               // TP:423-427
      ti_tb_rec r;
      memset(&r, 0, sizeof r);
      if (lit(p, e, "thread block = ") && exact_int(p, e, r.x) && lit(p, e, ",") && exact_int(p, e, r.y) &&
          lit(p, e, ",") && exact_int(p, e, r.z) && p == e) {
        r.text_off = (uint64_t)(b - body); r.line_len = (uint32_t)(e - b);
        r.first_warp = (uint32_t)n_warp; r.first_inst = (uint32_t)n_inst;
        if (dump) d.tb.push_back(r);
        n_tb++;
      } else flags |= TI_EF_BAD_STRUCTURE;
    } else if (c0 == 'w') {                               // TP:428-431
      int32_t w;
      if (lit(p, e, "warp = ") && exact_int(p, e, w) && p == e && w >= 0) {
        ti_warp_rec r = {(uint32_t)w, 0, (uint32_t)n_inst, 0};
        warps.push_back(r);
        n_warp++;
        expect_insts = true;
      } else flags |= TI_EF_BAD_STRUCTURE;
    } else if (c0 == 'i') {                               // TP:432-437
      int32_t n;
      if (expect_insts && lit(p, e, "insts = ") && exact_int(p, e, n) && p == e && n >= 0) warps.back().n_insts = (uint32_t)n;
      else flags |= TI_EF_BAD_STRUCTURE;
      expect_insts = false;
    } else if ((c0 >= '0' && c0 <= '9') || (c0 >= 'a' && c0 <= 'f') || (c0 >= 'A' && c0 <= 'F')) {
      Line L;
      parse_inst_line(b, e, hdr.trace_version, L);
      flags |= L.flags;
      n_inst++;
      if (L.flags || stages < 2) { if (L.has_mem) n_mem++; continue; }
      Inst I;
      convert(L, isa, cfg, hdr, I);
      flags |= I.flags;
      if (I.flags) { if (L.has_mem) n_mem++; continue; }
      unsigned cycles = 0;
      q.clear();
      if (stages >= 3) flags |= mem_accesses(cfg, I, L.mask, L.addrs, L.has_mem, q, cycles);
      const unsigned active = (unsigned)__builtin_popcount(L.mask);
      const bool is_store = I.op == kSTORE || I.op == kTSTORE || I.memory_op == 2;
      acc.issue(cfg, I.op, I.sp_op, active, I.latency, I.const_operand, I.space, is_store);
      acc.ints.smem_cycles += cycles;
      n_txn += q.size();
      if (L.has_mem) n_mem++;
      if (!dump) continue;
      ti_inst_rec r;
      memset(&r, 0, sizeof r);
      r.pc = L.pc; r.mask = L.mask;
      r.opc_off = (uint32_t)(L.opc_b - body);
      r.mem_idx = L.has_mem ? (uint32_t)(d.rows.size() / 32) : 0xFFFFFFFFu;
      r.dst = (uint8_t)L.dst;
      for (int m = 0; m < 4; ++m) r.src[m] = (uint8_t)L.src[m];
      r.opcode_id = (uint8_t)I.opcode_id; r.op = (uint8_t)I.op;
      r.pack_a = (uint8_t)(I.sp_op | ((I.oprnd_type + 1) << 5));
      r.pack_b = (uint8_t)(L.nsrc | (L.ndst << 3) | (I.memory_op << 4) | ((I.atomic ? 1 : 0) << 6) | ((I.const_operand ? 1 : 0) << 7));
      r.pack_c = (uint8_t)(I.space | (I.cache_op << 4) | ((L.has_mem ? 1 : 0) << 7));
      r.data_size = (uint8_t)I.data_size; r.cycles = (uint8_t)cycles;
      r.latency = (uint16_t)I.latency; r.initiation_interval = (uint8_t)I.init;
      r.opcode_len = (uint8_t)(L.opc_e - L.opc_b);
      d.inst.push_back(r);
      d.opcode_blob.append(L.opc_b, L.opc_e);
      d.opcode_blob.push_back('\0');
      if (L.has_mem) {
        d.rows.insert(d.rows.end(), L.addrs, L.addrs + 32);
        d.txn.insert(d.txn.end(), q.begin(), q.end());
        d.txn_off.push_back((uint32_t)d.txn.size());
      }
    } else {
      flags |= TI_EF_BAD_STRUCTURE;
    }
  }
  if (expect_insts) flags |= TI_EF_BAD_STRUCTURE;
  // count-based framing checks (the device applies the same ones)
  if (n_begin != n_end || n_begin != n_tb) flags |= TI_EF_BAD_STRUCTURE;
  for (size_t k = 0; k < warps.size(); ++k) {
    uint64_t next = k + 1 < warps.size() ? warps[k + 1].first_inst : n_inst;
    if (next - warps[k].first_inst != warps[k].n_insts) flags |= TI_EF_FRAMING;
  }
  if (warps.empty() ? n_inst != 0 : warps[0].first_inst != 0) flags |= TI_EF_FRAMING;
  double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

  acc.ints.n_inst = n_inst; acc.ints.n_mem = n_mem; acc.ints.n_txn = n_txn; acc.ints.n_tb = n_tb;
  acc.ints.n_warp = n_warp; acc.ints.n_begin_tb = n_begin; acc.ints.n_end_tb = n_end; acc.ints.text_bytes = len;
  if (dump) {
    d.warp = warps;
    d.counters = acc.ints;
    d.power = acc.dbl;
    if (!ti_dump_write(d, out_path)) { fprintf(stderr, "cannot write %s\n", out_path); return 3; }
  }
  printf("{\"impl\":\"port\",\"stages\":%d,\"seconds\":%.6f,\"n_inst\":%llu,\"n_tb\":%llu,\"n_txn\":%llu,\"error_flags\":%u}\n",
         stages, sec, (unsigned long long)n_inst, (unsigned long long)n_tb, (unsigned long long)n_txn, flags);
  return flags ? 5 : 0;
}
