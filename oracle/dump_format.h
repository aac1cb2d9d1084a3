// TEST INFRASTRUCTURE -- not a product path.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build or run anything
// under oracle/.
//
// Container written by both CPU oracles (oracle/_ref/ref_oracle, built from the
// reference's own translation units, and oracle/port_oracle, the restatement) so
// the tests can compare them with each other and with the CUDA library using one
// numpy loader (tests/tiformat.py).  Array element layouts are the C-ABI ones in
// include/traceingest.h.
//
//   char     magic[8]  = "TIDUMP2\0"
//   uint64_t n_inst, n_mem, n_txn, n_tb, n_warp, opcode_blob_bytes
//   ti_inst_rec  inst[n_inst]
//   uint64_t     rows[n_mem][32]
//   uint32_t     txn_off[n_mem + 1]   (padded to a multiple of 8 bytes)
//   ti_txn_rec   txn[n_txn]
//   ti_tb_rec    tb[n_tb]
//   ti_warp_rec  warp[n_warp]
//   ti_counters  counters
//   ti_power_counters power   (doubles accumulated in file order, reference style)
//   char         opcode_blob[opcode_blob_bytes]  (NUL-terminated opcode strings, in order)
#ifndef TI_ORACLE_DUMP_FORMAT_H
#define TI_ORACLE_DUMP_FORMAT_H

#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../include/traceingest.h"

struct ti_dump {
  std::vector<ti_inst_rec> inst;
  std::vector<uint64_t> rows;      // 32 per memory instruction
  std::vector<uint32_t> txn_off;   // CSR, n_mem + 1
  std::vector<ti_txn_rec> txn;
  std::vector<ti_tb_rec> tb;
  std::vector<ti_warp_rec> warp;
  ti_counters counters;
  ti_power_counters power;
  std::string opcode_blob;
};

static inline bool ti_dump_write(const ti_dump &d, const char *path) {
  FILE *f = fopen(path, "wb");
  if (!f) return false;
  const char magic[8] = {'T', 'I', 'D', 'U', 'M', 'P', '2', 0};
  uint64_t hdr[6] = {d.inst.size(),      d.rows.size() / 32, d.txn.size(),
                     d.tb.size(),        d.warp.size(),      d.opcode_blob.size()};
  fwrite(magic, 1, 8, f);
  fwrite(hdr, 8, 6, f);
  if (!d.inst.empty()) fwrite(d.inst.data(), sizeof(ti_inst_rec), d.inst.size(), f);
  if (!d.rows.empty()) fwrite(d.rows.data(), 8, d.rows.size(), f);
  fwrite(d.txn_off.data(), 4, d.txn_off.size(), f);
  if (d.txn_off.size() & 1) {
    uint32_t pad = 0;
    fwrite(&pad, 4, 1, f);
  }
  if (!d.txn.empty()) fwrite(d.txn.data(), sizeof(ti_txn_rec), d.txn.size(), f);
  if (!d.tb.empty()) fwrite(d.tb.data(), sizeof(ti_tb_rec), d.tb.size(), f);
  if (!d.warp.empty()) fwrite(d.warp.data(), sizeof(ti_warp_rec), d.warp.size(), f);
  fwrite(&d.counters, sizeof(ti_counters), 1, f);
  fwrite(&d.power, sizeof(ti_power_counters), 1, f);
  fwrite(d.opcode_blob.data(), 1, d.opcode_blob.size(), f);
  return fclose(f) == 0;
}

#endif
