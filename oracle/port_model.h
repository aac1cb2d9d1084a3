// TEST INFRASTRUCTURE -- not a product path (see oracle/README.md).
// Second half of the port: record -> instruction fields (TD:128-381, 441-480) and
// the memory-access generator (AHM:284-748).  See port_core.h for conventions.
#ifndef TI_ORACLE_PORT_MODEL_H
#define TI_ORACLE_PORT_MODEL_H

#include "port_core.h"

namespace ti_port {
using namespace ti_oracle;  // enum restatements from counters_restate.h

struct IsaView {
  std::map<std::string, std::pair<uint8_t, uint8_t>> by_name;  // mnemonic -> (opcode id, op_type)
  bool init(unsigned binary_version) {
    int col = arch_column(binary_version);
    if (col < 0) return false;
    for (const IsaRow &r : kIsa)
      if (r.id[col]) by_name[r.name] = std::make_pair(r.id[col], r.cat[col]);
    return true;
  }
};

// TD:128-149
static inline int oprnd_type_of(unsigned op, unsigned sp) {
  switch (op) {
    case kSP: case kSFU: case kSPEC0 + 1: case kSPEC0 + 2: case kDP: case kLOAD: case kSTORE: return 1;
    case kINTP: case kSPEC0 + 3: return 0;
    case kALU:
      if (sp == sFP || sp == sTEX || sp == sOTHER) return 1;
      if (sp == sINT) return 0;
      return -1;
  }
  return -1;
}

// TD:441-480
static inline void latency_of(const ti_config &c, unsigned cat, unsigned &lat, unsigned &init) {
  lat = init = 1;
  switch (cat) {
    case kALU: case kINTP: case kBRANCH: case kCALL: case kRET: lat = c.lat_int[0]; init = c.lat_int[1]; break;
    case kSP: lat = c.lat_sp[0]; init = c.lat_sp[1]; break;
    case kDP: lat = c.lat_dp[0]; init = c.lat_dp[1]; break;
    case kSFU: lat = c.lat_sfu[0]; init = c.lat_sfu[1]; break;
    case kTENSOR: lat = c.lat_tensor[0]; init = c.lat_tensor[1]; break;
  }
  if (cat >= kSPEC0) { lat = c.lat_spec[cat - kSPEC0][0]; init = c.lat_spec[cat - kSPEC0][1]; }
}

struct Inst {  // the fields of trace_warp_inst_t the path fills
  unsigned opcode_id, op, sp_op, memory_op, space, cache_op, data_size, latency, init;
  int oprnd_type;
  bool atomic, const_operand;
  uint32_t flags;
};

static inline bool has_token(const std::vector<std::string> &t, const char *s) {
  for (const std::string &x : t) if (x == s) return true;
  return false;
}

// TD:151-381
static inline void convert(const Line &L, const IsaView &isa, const ti_config &cfg, const ti_kernel_hdr &k, Inst &I) {
  memset(&I, 0, sizeof I);
  I.op = kALU; I.sp_op = sOTHER; I.oprnd_type = -1;   // TD:175-185
  std::vector<std::string> toks;
  opcode_tokens(L.opc_b, L.opc_e, toks);
  std::string full(L.opc_b, L.opc_e);
  // an opcode that starts with '.' is outside the tracer grammar: flagged, not looked up
  auto it = (toks.empty() || *L.opc_b == '.') ? isa.by_name.end() : isa.by_name.find(toks[0]);
  if (it == isa.by_name.end()) { I.flags |= TI_EF_UNKNOWN_OPCODE; return; }  // TD:202-206
  I.opcode_id = it->second.first;
  const unsigned cat = it->second.second;
  I.op = cat;
  I.sp_op = kSpop[I.opcode_id];
  I.oprnd_type = oprnd_type_of(I.op, I.sp_op);  // TD:201, before the overrides below
  if (toks[0] == "MUFU") {                        // TD:208-217, whole-string compares
    if (full == "MUFU.SIN" || full == "MUFU.COS") I.sp_op = sSIN;
    if (full == "MUFU.EX2" || full == "MUFU.RCP") I.sp_op = sEXP;
    if (full == "MUFU.RSQ") I.sp_op = sSQRT;
    if (full == "MUFU.LG2") I.sp_op = sLOG;
  }
  if (toks[0] == "IMAD" && (full == "IMAD.MOV" || full == "IMAD.IADD")) I.sp_op = sINT;  // TD:219-222
  latency_of(cfg, cat, I.latency, I.init);        // TD:243
  if (L.has_mem) I.data_size = L.width;           // TD:246-250
  const unsigned id = I.opcode_id;
  bool need_width = false;
  if (id == TI_OP_LDC) {                           // TD:255-261
    I.data_size = 4; I.memory_op = 1; I.const_operand = true; I.space = spCONST; I.cache_op = 1;
  } else if (id == TI_OP_LDG || id == TI_OP_LDL) { // TD:262-276
    need_width = true; I.memory_op = 1; I.cache_op = 1;
    I.space = id == TI_OP_LDL ? spLOCAL : spGLOBAL;
    if (has_token(toks, "STRONG") && has_token(toks, "GPU")) I.cache_op = 6;
  } else if (id == TI_OP_STG || id == TI_OP_STL) { // TD:277-286
    need_width = true; I.memory_op = 2; I.cache_op = 1;
    I.space = id == TI_OP_STL ? spLOCAL : spGLOBAL;
  } else if (id == TI_OP_ATOMG || id == TI_OP_RED || id == TI_OP_ATOM) {  // TD:287-296
    need_width = true; I.memory_op = 1; I.op = kLOAD; I.space = spGLOBAL; I.atomic = true; I.cache_op = 6;
  } else if (id == TI_OP_LDS) { need_width = true; I.memory_op = 1; I.space = spSHARED;
  } else if (id == TI_OP_STS) { need_width = true; I.memory_op = 2; I.space = spSHARED;
  } else if (id == TI_OP_ATOMS) { need_width = true; I.atomic = true; I.memory_op = 1; I.space = spSHARED;
  } else if (id == TI_OP_LDSM) { need_width = true; I.space = spSHARED;
  } else if (id == TI_OP_LD || id == TI_OP_ST) {   // TD:317-354
    need_width = true;
    I.memory_op = id == TI_OP_LD ? 1 : 2;
    if (k.shmem_base_addr == 0 || k.local_base_addr == 0) I.space = spSHARED;
    else if (L.mask && L.has_mem) {
      uint64_t a = L.addrs[__builtin_ctz(L.mask)];
      if (a >= k.shmem_base_addr && a < k.local_base_addr) I.space = spSHARED;
      else if (a >= k.local_base_addr && a < k.local_base_addr + (1u << 14)) { I.space = spLOCAL; I.cache_op = 1; }
      else { I.space = spGLOBAL; I.cache_op = 1; }
    }
  } else if (id == TI_OP_HADD2 || id == TI_OP_HADD2_32I || id == TI_OP_HFMA2 || id == TI_OP_HFMA2_32I ||
             id == TI_OP_HMUL2_32I || id == TI_OP_HSET2 || id == TI_OP_HSETP2) {
    I.init /= 2;                                   // TD:366-375
  }
  if (need_width && I.data_size == 0) I.flags |= TI_EF_ZERO_WIDTH;  // the asserts at TD:264-319
}

struct Txn { uint64_t block; uint32_t lanes, bytes[4]; unsigned chunks; };

static inline void set_bytes(Txn &t, unsigned idx, unsigned n) {
  for (unsigned i = 0; i < n; ++i)
    if (idx + i < 128) t.bytes[(idx + i) >> 5] |= 1u << ((idx + i) & 31);
}

// AHM:694-748
static inline void reduce_and_send(const Txn &t, unsigned seg, unsigned type, bool wr, std::vector<ti_txn_rec> &out) {
  uint64_t addr = t.block;
  unsigned size = seg, q = t.chunks, h = 0;
  if (seg == 128) {
    bool lo = q & 3, hi = q & 12;
    if (lo && !hi) { size = 64; h = q & 3; }
    else if (!lo && hi) { addr += 64; size = 64; h = (q >> 2) & 3; }
  } else if (seg == 64) {
    h = (addr % 128 == 0) ? (q & 3) : ((q >> 2) & 3);
  }
  if (size == 64) {
    bool lo = h & 1, hi = h & 2;
    if (lo && !hi) size = 32;
    else if (!lo && hi) { addr += 32; size = 32; }
  }
  ti_txn_rec r;
  memset(&r, 0, sizeof r);
  r.addr = addr; r.lane_mask = t.lanes; memcpy(r.byte_mask, t.bytes, 16);
  r.size = (uint8_t)size; r.sector_mask = (uint8_t)t.chunks; r.type = (uint8_t)type; r.is_write = wr;
  out.push_back(r);
}

static inline unsigned segment_size(const ti_config &c, unsigned data_size, unsigned cache_op) {
  bool sector = false;                              // AHM:482-493
  if (c.coalesce_arch >= 20 && c.coalesce_arch < 39) sector = c.gmem_skip_L1D || cache_op == 6;
  else if (c.coalesce_arch >= 40) sector = true;
  switch (data_size) {                              // AHM:495-507
    case 1: return 32;
    case 2: return sector ? 32 : 64;
    case 4: case 8: case 16: return sector ? 32 : 128;
  }
  return 0;
}

// AHM:284-473 for the spaces SASS traces can reach; returns flags.  `cycles` is the
// shared-memory bank-conflict count (AHM:329-420).
static inline uint32_t mem_accesses(const ti_config &c, const Inst &I, uint32_t mask, const uint64_t a[32],
                                    bool has_mem, std::vector<ti_txn_rec> &out, unsigned &cycles) {
  cycles = 0;
  bool is_load = I.op == kLOAD || I.op == kTLOAD || I.memory_op == 1;
  bool is_store = I.op == kSTORE || I.op == kTSTORE || I.memory_op == 2;
  if (!(is_load || is_store)) return 0;                               // TD:620
  if (!(I.op == kLOAD || I.op == kTLOAD || I.op == kSTORE || I.op == kTSTORE)) return 0;  // AHM:286-288
  if (mask == 0) return 0;                                            // AHM:289
  if (!has_mem) return TI_EF_UNDEF_SPACE;                             // AHM:296 assert
  const unsigned parts = c.mem_warp_parts, sub = 32 / parts;
  if (I.space == spSHARED || I.space == spSSTARR) {
    unsigned total = 0;
    for (unsigned p = 0; p < parts; ++p) {
      std::map<unsigned, std::map<uint64_t, unsigned>> banks;
      for (unsigned t = p * sub; t < (p + 1) * sub; ++t)
        if (mask >> t & 1) banks[(unsigned)((a[t] / 4) % c.num_shmem_bank)][a[t] & ~3ull]++;
      unsigned worst = 0;
      if (c.shmem_limited_broadcast) {
        bool found = false; unsigned bbank = 0; uint64_t bword = 0;
        for (auto &b : banks) {
          for (auto &w : b.second) if (w.second > 1) { found = true; bbank = b.first; bword = w.first; break; }
          if (found) break;
        }
        for (auto &b : banks) {
          unsigned n = 0;
          for (auto &w : b.second) n += w.second;
          if (found && b.first == bbank) n -= b.second[bword] - 1;
          worst = std::max(worst, n);
        }
      } else {
        for (auto &b : banks) worst = std::max(worst, (unsigned)b.second.size());
      }
      total += worst;
    }
    cycles = total;
    return 0;
  }
  if (I.space == spLOCAL) return 0;  // rewritten per SM at issue time (TD:593-601): outside the trace-only path
  if (I.space != spGLOBAL) return TI_EF_UNDEF_SPACE;                  // AHM:318-321
  const bool wr = is_store;
  const unsigned type = wr ? 4 : 0;                                   // GLOBAL_ACC_W / GLOBAL_ACC_R
  const unsigned seg = segment_size(c, I.data_size, I.cache_op);
  if (seg == 0) return TI_EF_BAD_WIDTH;
  const unsigned ds = I.data_size;
  for (unsigned p = 0; p < parts; ++p) {
    if (I.atomic) {                                                   // AHM:588-692
      std::map<uint64_t, std::vector<Txn>> m;
      for (unsigned t = p * sub; t < (p + 1) * sub; ++t) {
        if (!(mask >> t & 1)) continue;
        uint64_t addr = a[t], block = addr & ~(uint64_t)(seg - 1);
        if (block != ((addr + ds - 1) & ~(uint64_t)(seg - 1))) return TI_EF_ATOMIC_STRADDLE;
        unsigned idx = addr & 127;
        std::vector<Txn> &lst = m[block];
        Txn *tx = NULL;
        for (Txn &x : lst) {
          bool hit = false;
          for (unsigned i = idx; i <= idx + ds - 1; ++i) hit |= (x.bytes[i >> 5] >> (i & 31)) & 1;
          if (!hit) { tx = &x; break; }
        }
        if (!tx) { Txn n; memset(&n, 0, sizeof n); n.block = block; lst.push_back(n); tx = &lst.back(); }
        tx->chunks |= 1u << (idx / 32);
        tx->lanes |= 1u << t;
        set_bytes(*tx, idx, ds);
      }
      for (auto &kv : m) for (Txn &x : kv.second) reduce_and_send(x, seg, type, wr, out);
    } else {                                                          // AHM:475-586
      std::map<uint64_t, Txn> m;
      for (unsigned t = p * sub; t < (p + 1) * sub; ++t) {
        if (!(mask >> t & 1) || a[t] == 0) continue;                  // AHM:534-537
        uint64_t addr = a[t];
        for (int pass = 0; pass < 2; ++pass) {
          uint64_t block = addr & ~(uint64_t)(seg - 1);
          Txn &x = m[block];
          x.block = block;
          x.chunks |= 1u << ((addr & 127) / 32);
          x.lanes |= 1u << t;
          set_bytes(x, addr & 127, ds);
          if (pass == 1 || block == ((addr + ds - 1) & ~(uint64_t)(seg - 1))) break;
          addr = addr + ds - 1;                                       // AHM:559-571
        }
      }
      for (auto &kv : m) reduce_and_send(kv.second, seg, type, wr, out);
    }
  }
  return 0;
}

}  // namespace ti_port
#endif
