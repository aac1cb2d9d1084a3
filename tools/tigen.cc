// Synthetic .traceg corpus generator (host only; part of the bench/test harness).
//
// Emits files byte-compatible with what the reference's NVBit tracer plus
// post-traces-processing write: the 12-line "-key = value" header and the format
// line (util/tracer_nvbit/tracer_tool/tracer_tool.cu:458-482), per-block grouping
// (traces-processing/post-traces-processing.cpp:161-190) and the instruction-line
// grammar incl. the tracer's choice between base+stride, base+delta and list
// address compression (tracer_tool.cu:599-654, 685-756).  The choice rule and the
// formats are re-implemented here from that description; no tracer code is used.
//
// Deterministic per (seed, thread block): the output does not depend on the number
// of generator threads.
//
//   C ABI:  tigen_generate(&params, buf, cap) -> bytes written (or needed if > cap)
//   CLI:    tigen <profile> <seed> <n_tb> <insts_per_warp> <out.traceg> [binary_version]
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

extern "C" {
typedef struct tigen_params {
  uint64_t seed;
  uint32_t profile;        // 1 backprop-like, 2 streaming mix, 3 divergent list, 4 ring step, 5 soup, 6 test_kernel
  uint32_t grid[3];
  uint32_t block[3];
  uint32_t insts_per_warp;
  uint32_t binary_version; // 35,60,61,70,75,80,86
  uint32_t kernel_id;
  uint32_t trace_version;  // 3 normally; 2 adds the four leading ids on every line
  uint32_t n_threads;      // generator threads (0 = hardware concurrency)
  uint32_t variant;        // profile-specific (profile 1: 0 layerforward, 1 adjust_weights)
  uint64_t shmem_base, local_base;
  uint32_t raw;            // 1: the tracer's raw kernel-N.trace layout (every line prefixed "tbx tby tbz warp", the warps of a
                           //    thread block interleaved, no block structure) -- the input of post-traces-processing
  uint32_t pad_;
} tigen_params;
size_t tigen_generate(const tigen_params *p, char *buf, size_t cap);
}

namespace {

struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed) {}
  uint64_t next() {  // splitmix64
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  uint32_t below(uint32_t n) { return (uint32_t)(next() % n); }
  bool chance(double p) { return (next() >> 11) * (1.0 / 9007199254740992.0) < p; }
};

#define TI_ISA_BEGIN static const struct IsaRow { const char *name; uint8_t id[5]; uint8_t cat[5]; } kIsa[] = {
#define TI_ISA_ROW(n, a, b, c, d, e, f, g, h, i, j) {n, {a, b, c, d, e}, {f, g, h, i, j}},
#define TI_ISA_END };
#define TI_SPOP_BEGIN static const uint8_t kSpop[] = {
#define TI_SPOP(i, v) v,
#define TI_SPOP_END };
#include "../accel_sim_framework_distributed_b200/csrc/isa_tables.inc"

int arch_col(uint32_t v) {
  switch (v) {
    case 35: return 0;
    case 60: case 61: return 1;
    case 70: return 2;
    case 75: return 3;
    case 80: case 86: return 4;
  }
  return -1;
}

struct Out {
  std::string s;
  void put(const char *fmt, ...) __attribute__((format(printf, 2, 3))) {
    char tmp[256];
    va_list ap;
    va_start(ap, fmt);
    int n = vsnprintf(tmp, sizeof tmp, fmt, ap);
    va_end(ap);
    s.append(tmp, n);
  }
};

// The tracer's compression choice, restated: base+stride when the active lanes are one
// contiguous run of >= 2 lanes with a constant difference (a mask of 0 also lands here
// with base 0 / stride 0); base+delta otherwise; a plain list when compression is off.
void emit_mem(Out &o, unsigned width, const uint64_t a[32], uint32_t mask, bool compress) {
  o.put("%u ", width);
  if (!compress) {
    o.put("0 ");
    for (int s = 0; s < 32; ++s)
      if (mask >> s & 1) o.put("0x%016lx ", (unsigned long)a[s]);
    return;
  }
  bool stride_ok = true;
  uint64_t base = 0;
  int stride = 0;
  if (mask) {
    int first = __builtin_ctz(mask);
    int run = 0;
    while (first + run < 32 && (mask >> (first + run) & 1)) ++run;
    uint32_t run_mask = run == 32 ? 0xffffffffu : (((1u << run) - 1) << first);
    base = a[first];
    if (run < 2 || run_mask != mask) stride_ok = false;
    else {
      stride = (int)(a[first + 1] - a[first]);
      for (int s = first + 1; s < first + run; ++s)
        if ((int)(a[s] - a[s - 1]) != stride) { stride_ok = false; break; }
    }
  }
  if (stride_ok) {
    o.put("1 0x%llx %d ", (unsigned long long)base, stride);
  } else {
    o.put("2 0x%llx ", (unsigned long long)base);
    uint64_t last = base;
    bool first = true;
    for (int s = 0; s < 32; ++s)
      if (mask >> s & 1) {
        if (!first) o.put("%lld ", (long long)(a[s] - last));
        first = false;
        last = a[s];
      }
  }
}

struct Ctx {
  const tigen_params *p;
  uint32_t warps;
  int col;
  std::vector<const char *> alu_names;  // every non-memory mnemonic valid on this arch
};

void line_prefix(Out &o, const Ctx &c, uint32_t bx, uint32_t by, uint32_t bz, uint32_t w, uint32_t pc, uint32_t mask) {
  if (c.p->trace_version < 3 || c.p->raw) o.put("%u %u %u %u ", bx, by, bz, w);
  o.put("%04x %08x ", pc, mask);
}

void emit_alu(Out &o, Rng &r, const char *opc, int ndst, int nsrc) {
  if (ndst) o.put("1 R%u ", r.below(255)); else o.put("0 ");
  o.put("%s %d ", opc, nsrc);
  for (int i = 0; i < nsrc; ++i) o.put("R%u ", r.below(255));
  o.put("0 0\n");
}

uint32_t partial_run_mask(Rng &r) {
  int a = r.below(32), b = r.below(32);
  if (a > b) std::swap(a, b);
  int n = b - a + 1;
  return n == 32 ? 0xffffffffu : (((1u << n) - 1) << a);
}

const char *width_suffix(unsigned w) {
  switch (w) { case 1: return ".U8"; case 2: return ".U16"; case 8: return ".64"; case 16: return ".128"; }
  return "";
}

// ---- profile 2: streaming mix (BASELINE config 2)
void gen_tb_stream(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  static const struct { const char *n; int d, s; } alu[10] = {
      {"IMAD.MOV.U32", 1, 2}, {"FFMA", 1, 3}, {"IADD3", 1, 3}, {"LOP3.LUT", 1, 3}, {"MUFU.RCP", 1, 1},
      {"ISETP.GE.AND", 0, 2}, {"FMUL", 1, 2}, {"SHF.R.U32.HI", 1, 3}, {"DADD", 1, 2}, {"HFMA2", 1, 3}};
  Rng r(c.p->seed * 0x100000001B3ull + tb);
  const uint32_t I = c.p->insts_per_warp;
  for (uint32_t w = 0; w < c.warps; ++w) {
    o.put("\nwarp = %u\ninsts = %u\n", w, I);
    for (uint32_t i = 0; i < I; ++i) {
      uint32_t mask = r.chance(0.05) ? partial_run_mask(r) : 0xffffffffu;
      line_prefix(o, c, bx, by, bz, w, i * 16, mask);
      uint32_t k = r.below(100);
      if (k < 60) {
        auto &a = alu[r.below(10)];
        emit_alu(o, r, a.n, a.d, a.s);
      } else {
        bool ld = k < 80;
        uint64_t idx = ((uint64_t)tb * c.warps + w) * I + i;
        uint64_t addr[32];
        for (int s = 0; s < 32; ++s) addr[s] = 0x7f0000000000ull + 128 * idx + 4 * s;
        if (ld) o.put("1 R%u LDG.E.SYS 1 R%u ", r.below(255), r.below(255));
        else o.put("0 STG.E.SYS 2 R%u R%u ", r.below(255), r.below(255));
        emit_mem(o, 4, addr, mask, true);
        o.put("0\n");
      }
    }
  }
}

// ---- profile 3: memory-divergent, list-compressed (BASELINE config 3)
void gen_tb_divergent(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  static const char *alu[4] = {"FFMA", "IADD3", "IMAD.MOV.U32", "LOP3.LUT"};
  static const unsigned widths[5] = {1, 2, 4, 8, 16};
  Rng r(c.p->seed * 0x100000001B3ull + tb);
  const uint32_t I = c.p->insts_per_warp;
  for (uint32_t w = 0; w < c.warps; ++w) {
    o.put("\nwarp = %u\ninsts = %u\n", w, I);
    for (uint32_t i = 0; i < I; ++i) {
      line_prefix(o, c, bx, by, bz, w, i * 16, 0xffffffffu);
      if (r.chance(0.5)) { emit_alu(o, r, alu[r.below(4)], 1, 3); continue; }
      unsigned wd = widths[r.below(5)];
      int perm[32];
      for (int s = 0; s < 32; ++s) perm[s] = s;
      for (int s = 31; s > 0; --s) std::swap(perm[s], perm[r.below(s + 1)]);
      uint64_t page = 0x7f0000000000ull + ((uint64_t)r.below(1u << 20) << 17);
      uint64_t addr[32];
      int share = r.chance(0.10) ? 2 + r.below(7) : 1;  // lanes per sector
      for (int s = 0; s < 32; ++s) {
        int grp = s / share;
        uint64_t sector = page + 4096ull * perm[grp] + 32ull * ((r.next() >> 8) % 128) * 0;
        sector = page + 4096ull * perm[grp];
        // one random sector per group, derived from the group so sharers agree
        Rng g(c.p->seed ^ (tb * 1315423911ull + w * 2654435761ull + i * 97 + grp));
        sector += 32ull * g.below(128);
        unsigned slots = 32 / wd;
        addr[s] = sector + (uint64_t)wd * ((s % share + g.below(slots)) % slots);
        if (r.chance(0.02) && wd > 1) addr[s] = sector + 32 - wd / 2;  // straddles into the next sector
      }
      bool ld = r.chance(0.6);
      if (ld) o.put("1 R%u LDG.E%s.SYS 1 R%u ", r.below(255), width_suffix(wd), r.below(255));
      else o.put("0 STG.E%s.SYS 2 R%u R%u ", width_suffix(wd), r.below(255), r.below(255));
      emit_mem(o, wd, addr, 0xffffffffu, false);
      o.put("0\n");
    }
  }
}

// ---- profile 1: backprop-shaped kernels (BASELINE config 1 stand-in)
void gen_tb_backprop(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  Rng r(c.p->seed * 0x100000001B3ull + tb + 77ull * c.p->variant);
  const uint32_t I = c.p->insts_per_warp;
  const uint64_t gin = 0x7f2a00000000ull, gw = 0x7f2a10000000ull, gout = 0x7f2a20000000ull;
  for (uint32_t w = 0; w < c.warps; ++w) {
    o.put("\nwarp = %u\ninsts = %u\n", w, I);
    for (uint32_t i = 0; i < I; ++i) {
      uint32_t mask = 0xffffffffu;
      if (i + 1 == I) {  // every warp ends with EXIT
        line_prefix(o, c, bx, by, bz, w, i * 16, mask);
        o.put("0 EXIT 0 0 0\n");
        continue;
      }
      uint32_t k = (i * 7 + w) % 16;
      if (k >= 12 && r.chance(0.3)) mask = 0x0000ffffu << (16 * r.below(2));
      line_prefix(o, c, bx, by, bz, w, i * 16, mask);
      uint64_t addr[32];
      uint32_t row = by * 16 + w * 2;
      switch (k) {
        case 0: case 1:
          for (int s = 0; s < 32; ++s) addr[s] = (c.p->variant ? gw : gin) + 4ull * ((row + s / 16) * 17 + s % 16 + i);
          o.put("1 R%u LDG.E.SYS 1 R%u ", r.below(64), r.below(64));
          emit_mem(o, 4, addr, mask, true); o.put("0\n"); break;
        case 2:
          for (int s = 0; s < 32; ++s) addr[s] = 4ull * ((w * 2 + s / 16) * 16 + s % 16);
          o.put("0 STS 2 R%u R%u ", r.below(64), r.below(64));
          emit_mem(o, 4, addr, mask, true); o.put("0\n"); break;
        case 3:
          o.put("0 BAR.SYNC 0 0 0\n"); break;
        case 4: case 5:
          // column walk of a 16x16 tile: 2-way .. 16-way bank conflicts depending on i
          for (int s = 0; s < 32; ++s) addr[s] = 4ull * ((s % 16) * (1u << (i % 5)) + s / 16) + 0x400;
          o.put("1 R%u LDS.U.32 1 R%u ", r.below(64), r.below(64));
          emit_mem(o, 4, addr, mask, true); o.put("0\n"); break;
        case 6: emit_alu(o, r, "FFMA", 1, 3); break;
        case 7: emit_alu(o, r, "FMUL", 1, 2); break;
        case 8: emit_alu(o, r, c.p->variant ? "MUFU.EX2" : "MUFU.RCP", 1, 1); break;
        case 9: emit_alu(o, r, "IMAD.MOV.U32", 1, 2); break;
        case 10: emit_alu(o, r, "ISETP.GE.AND", 0, 2); break;
        case 11: emit_alu(o, r, "IADD3", 1, 3); break;
        case 12:
          for (int s = 0; s < 32; ++s) addr[s] = gout + 4ull * ((uint64_t)(row + s / 16) * 4097 + s % 16) * (c.p->variant + 1);
          o.put("0 STG.E.SYS 2 R%u R%u ", r.below(64), r.below(64));
          emit_mem(o, 4, addr, mask, true); o.put("0\n"); break;
        case 13: emit_alu(o, r, "IMAD.WIDE", 1, 2); break;
        case 14: o.put("0 BRA 0 0 0\n"); break;
        default: emit_alu(o, r, "FADD", 1, 2); break;
      }
    }
  }
}

// ---- profile 4: one ring all-reduce step (recv chunk + local chunk -> send chunk)
void gen_tb_ring(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  Rng r(c.p->seed * 0x100000001B3ull + tb);
  const uint32_t I = c.p->insts_per_warp;
  const uint64_t recv = 0x7f4000000000ull + ((uint64_t)c.p->variant << 24), send = 0x7f5000000000ull + ((uint64_t)c.p->variant << 24);
  for (uint32_t w = 0; w < c.warps; ++w) {
    o.put("\nwarp = %u\ninsts = %u\n", w, I);
    for (uint32_t i = 0; i < I; ++i) {
      line_prefix(o, c, bx, by, bz, w, i * 16, 0xffffffffu);
      uint64_t elem = ((uint64_t)tb * c.warps + w) * 32 + (uint64_t)(i / 4) * c.p->grid[0] * c.warps * 32;
      uint64_t addr[32];
      switch (i % 4) {
        case 0: for (int s = 0; s < 32; ++s) addr[s] = recv + 4 * (elem + s);
          o.put("1 R%u LDG.E.SYS 1 R%u ", r.below(32), r.below(32)); emit_mem(o, 4, addr, 0xffffffffu, true); o.put("0\n"); break;
        case 1: for (int s = 0; s < 32; ++s) addr[s] = send + 4 * (elem + s);
          o.put("1 R%u LDG.E.SYS 1 R%u ", r.below(32), r.below(32)); emit_mem(o, 4, addr, 0xffffffffu, true); o.put("0\n"); break;
        case 2: emit_alu(o, r, "FADD", 1, 2); break;
        default: for (int s = 0; s < 32; ++s) addr[s] = send + 4 * (elem + s);
          o.put("0 STG.E.SYS 2 R%u R%u ", r.below(32), r.below(32)); emit_mem(o, 4, addr, 0xffffffffu, true); o.put("0\n"); break;
      }
    }
  }
}

// ---- profile 6: examples/all-reduce test_kernel (ptr[tid] = 7)
void gen_tb_testkernel(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  for (uint32_t w = 0; w < c.warps; ++w) {
    o.put("\nwarp = %u\ninsts = 5\n", w);
    uint64_t addr[32];
    for (int s = 0; s < 32; ++s) addr[s] = 0x7f6000000000ull + 4ull * ((uint64_t)(tb * c.warps + w) * 32 + s);
    line_prefix(o, c, bx, by, bz, w, 0x00, 0xffffffffu); o.put("1 R1 IMAD.MOV.U32 2 R255 R255 0 0\n");
    line_prefix(o, c, bx, by, bz, w, 0x10, 0xffffffffu); o.put("1 R2 S2R 0 0 0\n");
    line_prefix(o, c, bx, by, bz, w, 0x20, 0xffffffffu); o.put("1 R2 IMAD.WIDE 2 R2 R3 0 0\n");
    line_prefix(o, c, bx, by, bz, w, 0x30, 0xffffffffu); o.put("0 STG.E.SYS 2 R2 R5 "); emit_mem(o, 4, addr, 0xffffffffu, true); o.put("0\n");
    line_prefix(o, c, bx, by, bz, w, 0x40, 0xffffffffu); o.put("0 EXIT 0 0 0\n");
  }
}

// ---- profile 5: a soup of every line shape the reader defines (differential fuzz)
void gen_tb_soup(Out &o, const Ctx &c, uint32_t tb, uint32_t bx, uint32_t by, uint32_t bz) {
  Rng r(c.p->seed * 0x100000001B3ull + tb);
  static const unsigned widths[5] = {1, 2, 4, 8, 16};
  for (uint32_t w = 0; w < c.warps; ++w) {
    uint32_t I = r.chance(0.1) ? 0 : 1 + r.below(c.p->insts_per_warp * 2);
    o.put("\nwarp = %u\ninsts = %u\n", w, I);
    for (uint32_t i = 0; i < I; ++i) {
      uint32_t mask;
      switch (r.below(8)) {
        case 0: mask = (uint32_t)r.next(); break;
        case 1: mask = partial_run_mask(r); break;
        case 2: mask = 1u << r.below(32); break;
        case 3: mask = r.chance(0.3) ? 0 : (uint32_t)r.next() & (uint32_t)r.next(); break;
        default: mask = 0xffffffffu;
      }
      uint32_t pc = r.chance(0.02) ? 0x10000 + r.below(1u << 20) : i * 16;
      line_prefix(o, c, bx, by, bz, w, pc, mask);
      unsigned kind = r.below(100);
      if (kind < 45) {  // any non-memory mnemonic of this architecture
        const char *n = c.alu_names[r.below((uint32_t)c.alu_names.size())];
        std::string opc = n;
        if (!strcmp(n, "MUFU")) { static const char *m[8] = {".SIN", ".COS", ".EX2", ".RCP", ".RSQ", ".LG2", ".RCP64H", ".SQRT"}; opc += m[r.below(8)]; }
        else if (!strcmp(n, "IMAD")) { static const char *m[6] = {"", ".MOV", ".IADD", ".MOV.U32", ".WIDE", ".SHL.U32"}; opc += m[r.below(6)]; }
        else if (r.chance(0.3)) { static const char *m[6] = {".FTZ", ".X", ".U32", ".64", ".128", ".S16"}; opc += m[r.below(6)]; }
        emit_alu(o, r, opc.c_str(), r.below(2), r.below(5));
        continue;
      }
      unsigned wd = widths[r.below(5)];
      uint64_t addr[32];
      uint64_t base = 0x7f0000000000ull + ((r.next() & 0xfffffffull) << 2);
      bool align = true;
      switch (r.below(7)) {
        case 0: for (int s = 0; s < 32; ++s) addr[s] = base + (uint64_t)wd * s; break;                 // unit stride
        case 1: { int st = (int)r.below(600) - 300; for (int s = 0; s < 32; ++s) addr[s] = base + (int64_t)st * s; align = false; break; }
        case 2: for (int s = 0; s < 32; ++s) addr[s] = base; break;                                    // broadcast
        case 3: for (int s = 0; s < 32; ++s) addr[s] = base + (r.next() & 0xffff); align = false; break; // scattered
        case 4: for (int s = 0; s < 32; ++s) addr[s] = base + 128ull * r.below(6) + (uint64_t)wd * r.below(8); break;
        case 5: for (int s = 0; s < 32; ++s) addr[s] = base + 4096ull * (31 - s); break;               // descending
        default: for (int s = 0; s < 32; ++s) addr[s] = base + 32ull * s + (r.chance(0.2) ? 33 - wd : 0); align = false;
      }
      bool compress = !r.chance(0.25);
      char opc[96];
      if (kind < 60) {
        snprintf(opc, sizeof opc, "LDG.E%s%s", width_suffix(wd), r.chance(0.2) ? ".STRONG.GPU" : r.chance(0.5) ? ".SYS" : ".CONSTANT");
        o.put("1 R%u %s 1 R%u ", r.below(255), opc, r.below(255));
      } else if (kind < 70) {
        snprintf(opc, sizeof opc, "STG.E%s.SYS", width_suffix(wd));
        if (c.col == 0) snprintf(opc, sizeof opc, "ST.E%s", width_suffix(wd));  // Kepler: ST maps to OP_STG
        o.put("0 %s 2 R%u R%u ", opc, r.below(255), r.below(255));
      } else if (kind < 78) {  // shared
        for (int s = 0; s < 32; ++s) addr[s] &= 0xffff;
        bool st = r.chance(0.4);
        snprintf(opc, sizeof opc, "%s%s", st ? "STS" : "LDS", width_suffix(wd));
        if (st) o.put("0 %s 2 R%u R%u ", opc, r.below(255), r.below(255)); else o.put("1 R%u %s 1 R%u ", r.below(255), opc, r.below(255));
      } else if (kind < 84) {  // atomics: aligned, never straddling (AHM.cc:643)
        unsigned aw = wd < 4 ? 4 : (wd > 8 ? 8 : wd);
        wd = aw;
        for (int s = 0; s < 32; ++s) addr[s] = (addr[s] / aw) * aw;
        const char *nm = c.col >= 1 ? (r.chance(0.5) ? "ATOMG" : "RED") : (r.chance(0.5) ? "ATOM" : "RED");
        snprintf(opc, sizeof opc, "%s.E.ADD%s.STRONG.GPU", nm, aw == 8 ? ".64" : "");
        o.put("1 R%u %s 2 R%u R%u ", r.below(255), opc, r.below(255), r.below(255));
      } else if (kind < 88) {  // local
        bool st = r.chance(0.5);
        for (int s = 0; s < 32; ++s) addr[s] = c.p->local_base + 4ull * s + 512ull * r.below(8);
        snprintf(opc, sizeof opc, "%s%s", st ? "STL" : "LDL", width_suffix(wd));
        if (st) o.put("0 %s 2 R%u R%u ", opc, r.below(255), r.below(255)); else o.put("1 R%u %s 1 R%u ", r.below(255), opc, r.below(255));
      } else if (kind < 94 && c.col >= 1) {  // generic LD/ST resolved by the first active lane
        unsigned where = r.below(3);
        uint64_t b = where == 0 ? c.p->shmem_base + 64 : where == 1 ? c.p->local_base + 64 : base;
        for (int s = 0; s < 32; ++s) addr[s] = b + (addr[s] - base) % 8192;
        bool st = r.chance(0.5);
        snprintf(opc, sizeof opc, "%s.E%s", st ? "ST" : "LD", width_suffix(wd));
        if (st) o.put("0 %s 2 R%u R%u ", opc, r.below(255), r.below(255)); else o.put("1 R%u %s 1 R%u ", r.below(255), opc, r.below(255));
      } else if (kind < 97) {  // constant load: the tracer reports no addresses for it
        o.put("1 R%u LDC 1 R%u 0 0\n", r.below(255), r.below(255));
        continue;
      } else if (c.col >= 3 && r.chance(0.5)) {  // LDSM exists on Turing / Ampere only
        for (int s = 0; s < 32; ++s) addr[s] &= 0xffff;
        o.put("1 R%u LDSM.16.M88.4 1 R%u ", r.below(255), r.below(255));
        wd = 2;
      } else if (c.col >= 1) {
        for (int s = 0; s < 32; ++s) addr[s] &= 0xfffc;
        o.put("1 R%u ATOMS.ADD 2 R%u R%u ", r.below(255), r.below(255), r.below(255));
        wd = 4;
      } else {  // Kepler has no ATOMS mnemonic
        for (int s = 0; s < 32; ++s) addr[s] &= 0xfffc;
        o.put("1 R%u LDS 1 R%u ", r.below(255), r.below(255));
        wd = 4;
      }
      (void)align;
      emit_mem(o, wd, addr, mask, compress);
      if (r.chance(0.9)) o.put("%d", (int)r.below(3)); else o.s.pop_back();  // immediate may be absent (README example)
      o.put("\n");
    }
  }
}

void gen_header(Out &o, const tigen_params &p) {
  static const char *names[7] = {"", "_Z22bpnn_layerforward_CUDAPfS_S_S_ii", "_Z6streamPfS_i", "_Z6gatherPfPiS_i",
                                 "_Z9ring_stepPfS_S_i", "_Z4soupPv", "_Z11test_kernelPi"};
  const char *nm = names[p.profile < 7 ? p.profile : 0];
  if (p.profile == 1 && p.variant) nm = "_Z24bpnn_adjust_weights_cudaPfiS_iS_S_";
  o.put("-kernel name = %s\n-kernel id = %u\n-grid dim = (%u,%u,%u)\n-block dim = (%u,%u,%u)\n", nm, p.kernel_id,
        p.grid[0], p.grid[1], p.grid[2], p.block[0], p.block[1], p.block[2]);
  o.put("-shmem = %u\n-nregs = %u\n-binary version = %u\n-cuda stream id = %lu\n", p.profile == 1 ? 1088u : 0u, 32u,
        p.binary_version, 0ul);
  o.put("-shmem base_addr = 0x%016lx\n-local mem base_addr = 0x%016lx\n", (unsigned long)p.shmem_base, (unsigned long)p.local_base);
  o.put("-nvbit version = 1.5.5\n-accelsim tracer version = %u\n\n", p.trace_version);
  o.put("#traces format = [line_num] PC mask dest_num [reg_dests] opcode src_num [reg_srcs] mem_width "
        "[adrrescompress?] [mem_addresses] immediate\n\n\n");
}

}  // namespace

extern "C" size_t tigen_generate(const tigen_params *pp, char *buf, size_t cap) {
  tigen_params p = *pp;
  if (!p.trace_version) p.trace_version = 3;
  if (!p.binary_version) p.binary_version = 70;
  Ctx c;
  c.p = &p;
  c.warps = (p.block[0] * p.block[1] * p.block[2] + 31) / 32;
  c.col = arch_col(p.binary_version);
  if (c.col < 0) c.col = 2;
  for (const IsaRow &row : kIsa) {
    if (!row.id[c.col]) continue;
    unsigned cat = row.cat[c.col];
    bool ldst = cat == 8 || cat == 9 || cat == 10 || cat == 11;  // LOAD / TENSOR_LOAD / TENSOR_STORE / STORE
    unsigned id = row.id[c.col];
    if (ldst || id == TI_OP_LDC || id == TI_OP_LD || id == TI_OP_ST) continue;
    c.alu_names.push_back(row.name);
  }
  const uint64_t n_tb = (uint64_t)p.grid[0] * p.grid[1] * p.grid[2];
  unsigned nt = p.n_threads ? p.n_threads : std::max(1u, std::thread::hardware_concurrency());
  nt = (unsigned)std::min<uint64_t>(nt, std::max<uint64_t>(1, n_tb / 64));
  std::vector<Out> parts(nt);
  auto work = [&](unsigned t) {
    uint64_t lo = n_tb * t / nt, hi = n_tb * (t + 1) / nt;
    Out &o = parts[t];
    for (uint64_t tb = lo; tb < hi; ++tb) {
      uint32_t bx = (uint32_t)(tb % p.grid[0]), by = (uint32_t)(tb / p.grid[0] % p.grid[1]),
               bz = (uint32_t)(tb / ((uint64_t)p.grid[0] * p.grid[1]));
      Out tmp;
      Out &dst = o;
      Out &o = p.raw ? tmp : dst;
      o.put("\n#BEGIN_TB\n\nthread block = %u,%u,%u\n", bx, by, bz);
      switch (p.profile) {
        case 1: gen_tb_backprop(o, c, (uint32_t)tb, bx, by, bz); break;
        case 3: gen_tb_divergent(o, c, (uint32_t)tb, bx, by, bz); break;
        case 4: gen_tb_ring(o, c, (uint32_t)tb, bx, by, bz); break;
        case 5: gen_tb_soup(o, c, (uint32_t)tb, bx, by, bz); break;
        case 6: gen_tb_testkernel(o, c, (uint32_t)tb, bx, by, bz); break;
        default: gen_tb_stream(o, c, (uint32_t)tb, bx, by, bz); break;
      }
      o.put("\n#END_TB\n");
      if (p.raw) {  // keep only the instruction lines, round-robin over the warps (instruction k of every warp, then k+1 ...)
        std::vector<std::vector<std::pair<size_t, size_t>>> per_warp;
        const std::string &t = tmp.s;
        for (size_t a = 0; a < t.size();) {
          size_t e = t.find('\n', a);
          if (e == std::string::npos) e = t.size();
          if (e > a) {
            if (!t.compare(a, 7, "warp = ")) per_warp.emplace_back();
            else if (t[a] >= '0' && t[a] <= '9' && !per_warp.empty()) per_warp.back().emplace_back(a, e - a);
          }
          a = e + 1;
        }
        size_t most = 0;
        for (auto &w : per_warp) most = std::max(most, w.size());
        for (size_t k = 0; k < most; ++k)
          for (auto &w : per_warp)
            if (k < w.size()) { dst.s.append(t, w[k].first, w[k].second); dst.s.push_back('\n'); }
      }
    }
  };
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t) th.emplace_back(work, t);
  work(0);
  for (auto &t : th) t.join();
  Out hdr;
  gen_header(hdr, p);
  size_t total = hdr.s.size() + (p.raw ? 0 : 1);  // trailing blank line, as post-traces-processing leaves one
  for (auto &o : parts) total += o.s.size();
  if (total > cap || !buf) return total;
  size_t off = 0;
  memcpy(buf + off, hdr.s.data(), hdr.s.size()); off += hdr.s.size();
  for (auto &o : parts) { memcpy(buf + off, o.s.data(), o.s.size()); off += o.s.size(); }
  if (!p.raw) buf[off++] = '\n';
  return off;
}

#ifdef TIGEN_MAIN
int main(int argc, char **argv) {
  if (argc < 6) { fprintf(stderr, "usage: tigen <profile> <seed> <n_tb> <insts_per_warp> <out> [binary_version] [block_threads]\n"); return 2; }
  tigen_params p;
  memset(&p, 0, sizeof p);
  p.profile = atoi(argv[1]); p.seed = strtoull(argv[2], 0, 0);
  p.grid[0] = atoi(argv[3]); p.grid[1] = p.grid[2] = 1;
  p.insts_per_warp = atoi(argv[4]);
  p.binary_version = argc > 6 ? atoi(argv[6]) : 70;
  p.block[0] = argc > 7 ? atoi(argv[7]) : 256; p.block[1] = p.block[2] = 1;
  p.kernel_id = 1; p.trace_version = 3;
  p.shmem_base = 0x00007f8c00000000ull; p.local_base = 0x00007f8c10000000ull;
  size_t need = tigen_generate(&p, NULL, 0);
  std::vector<char> buf(need);
  size_t n = tigen_generate(&p, buf.data(), buf.size());
  FILE *f = fopen(argv[5], "wb");
  if (!f) return 3;
  fwrite(buf.data(), 1, n, f);
  fclose(f);
  fprintf(stderr, "%zu bytes\n", n);
  return 0;
}
#endif
