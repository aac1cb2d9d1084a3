// TEST INFRASTRUCTURE -- not a product path (see oracle/README.md).
//
// CPU restatement of the AccelWattch access-counter increments that every
// issued warp instruction triggers in the reference.  shader.cc/shader.h drag
// in the whole timing model, so they cannot be linked into the oracle; this
// file follows them line for line instead and is fed with fields produced by
// the reference's own parse_from_trace_struct (oracle/_ref) or by the port.
//
// Followed code (relative to gpu-simulator/gpgpu-sim/src/gpgpu-sim/):
//   scheduler_unit::cycle unit routing        shader.cc:1336-1497
//   sp/dp/int/specialized/sfu/tensor ::issue  shader.cc:2350-2368, 2464-2498
//   ldst_unit::issue                          shader.cc:2636-2661
//   pipelined_simd_unit::issue -> incexecstat shader.cc:2539-2548, 3226-3290
//   inc*_stat, inactive_lanes_accesses_*      shader.h:2159-2360, 2399-2407
//   mem_instruction_stats                     gpu-sim.cc:1548-1582
#ifndef TI_ORACLE_COUNTERS_RESTATE_H
#define TI_ORACLE_COUNTERS_RESTATE_H

#include <string.h>

#include "../include/traceingest.h"

namespace ti_oracle {

// enum values restated from abstract_hardware_model.h:108-167, 47-62
enum { kALU = 1, kSFU, kTENSOR, kDP, kSP, kINTP, kALU_SFU, kLOAD, kTLOAD, kTSTORE,
       kSTORE, kBRANCH, kBARRIER, kMEMBAR, kCALL, kRET, kEXIT, kSPEC0 = 100 };
enum { sOTHER, sINT, sIMUL24, sIMUL32, sIMUL, sIDIV, sFPMUL, sFPDIV, sFP, sSQRT,
       sLOG, sSIN, sEXP, sDPMUL, sDPDIV, sDP, sTENSOR, sTEX };
enum { spUNDEF = 0, spREG, spLOCAL, spSHARED, spSSTARR, spPARAM_U, spPARAM_K,
       spPARAM_L, spCONST, spTEX, spSURF, spGLOBAL, spGENERIC, spINSTR };

static inline unsigned fold_op(unsigned op) { return op >= kSPEC0 ? 18 + (op - kSPEC0) : op; }

// shader.h:2399-2407 -- both helpers return `unsigned`, so the double sum is
// truncated once per call.
static inline unsigned inactive_sfu(unsigned a, double c) {
  return (unsigned)((((32 - a) >> 1) * c) + (((32 - a) >> 3) * c) + (((32 - a) >> 3) * c));
}
static inline unsigned inactive_nonsfu(unsigned a, double c) {
  return (unsigned)(((32 - a) >> 1) * c);
}

struct counter_acc {
  ti_counters ints;
  ti_power_counters dbl;
  counter_acc() {
    memset(&ints, 0, sizeof ints);
    memset(&dbl, 0, sizeof dbl);
  }

  // One issued warp instruction.  `op` is the final op_type (after the atomics
  // flip), `is_store` is inst_t::is_store().
  void issue(const ti_config &cfg, unsigned op, unsigned sp_op, unsigned active,
             unsigned latency, bool const_operand, unsigned space, bool is_store) {
    ints.hist[sp_op][active]++;
    ints.op_count[fold_op(op)]++;
    // --- which unit does the scheduler hand it to (shader.cc:1336-1497)
    bool to_mem = op == kLOAD || op == kSTORE || op == kMEMBAR || op == kTLOAD || op == kTSTORE;
    if (to_mem) {
      // ldst_unit::issue -> mem_instruction_stats (gpu-sim.cc:1548-1582)
      ints.space_lanes[space][is_store ? 1 : 0] += active;
      switch (space) {
        case spSHARED: dbl.n_shmem_insn += active; break;
        case spSSTARR: dbl.n_sstarr_insn += active; break;
        case spCONST: dbl.n_const_insn += active; break;
        case spPARAM_K: case spPARAM_L: dbl.n_param_insn += active; break;
        case spTEX: dbl.n_tex_insn += active; break;
        case spGLOBAL: case spLOCAL:
          if (is_store) dbl.n_store_insn += active; else dbl.n_load_insn += active;
          break;
        default: break;  // undefined_space / reg_space
      }
      // incmem_stat(warp_size, 1): shader.h:2350-2360; the inactive term is 0 for 32 lanes
      dbl.mem_accesses = dbl.mem_accesses + (double)32 * 1.0;
    } else if (op == kTENSOR || op == kSFU || (op == kDP && cfg.num_dp_units == 0)) {
      dbl.sfu_accesses = dbl.sfu_accesses + (double)32 * (double)latency;   // incsfu_stat
    } else {
      dbl.sp_accesses = dbl.sp_accesses + (double)32 * (double)latency;     // incsp_stat
    }
    // --- incexecstat (shader.cc:3226-3290); runs for every unit
    double *slot = 0; double c = 0; bool sfu_like = false;
    switch (sp_op) {
      case sINT:    slot = &dbl.ialu;   c = cfg.coeff[TI_C_INT]; break;
      case sIMUL:   slot = &dbl.imul;   c = cfg.coeff[TI_C_INT_MUL]; break;
      case sIMUL24: slot = &dbl.imul24; c = cfg.coeff[TI_C_INT_MUL24]; break;
      case sIMUL32: slot = &dbl.imul32; c = cfg.coeff[TI_C_INT_MUL32]; sfu_like = true; break;
      case sIDIV:   slot = &dbl.idiv;   c = cfg.coeff[TI_C_INT_DIV]; sfu_like = true; break;
      case sFP:     slot = &dbl.fpalu;  c = cfg.coeff[TI_C_FP]; break;
      case sFPMUL:  slot = &dbl.fpmul;  c = cfg.coeff[TI_C_FP_MUL]; break;
      case sFPDIV:  slot = &dbl.fpdiv;  c = cfg.coeff[TI_C_FP_DIV]; sfu_like = true; break;
      case sDP:     slot = &dbl.dpalu;  c = cfg.coeff[TI_C_DP]; break;
      case sDPMUL:  slot = &dbl.dpmul;  c = cfg.coeff[TI_C_DP_MUL]; break;
      case sDPDIV:  slot = &dbl.dpdiv;  c = cfg.coeff[TI_C_DP_DIV]; sfu_like = true; break;
      case sSQRT:   slot = &dbl.sqrt_;  c = cfg.coeff[TI_C_SQRT]; sfu_like = true; break;
      case sLOG:    slot = &dbl.log_;   c = cfg.coeff[TI_C_LOG]; sfu_like = true; break;
      case sSIN:    slot = &dbl.sin_;   c = cfg.coeff[TI_C_SIN]; sfu_like = true; break;
      case sEXP:    slot = &dbl.exp_;   c = cfg.coeff[TI_C_EXP]; sfu_like = true; break;
      case sTENSOR: slot = &dbl.tensor; c = cfg.coeff[TI_C_TENSOR]; sfu_like = true; break;
      case sTEX:    slot = &dbl.tex;    c = cfg.coeff[TI_C_TEX]; sfu_like = true; break;
      default: break;  // OTHER_OP: no class counter
    }
    if (slot) {
      if (!cfg.clock_gated_lanes)
        *slot = *slot + (double)active * c +
                (sfu_like ? inactive_sfu(active, c) : inactive_nonsfu(active, c));
      else
        *slot = *slot + (double)active * c;
      dbl.active_exu_threads += active;
      dbl.active_exu_warps += 1;
    }
    if (const_operand) {
      dbl.const_accesses = dbl.const_accesses + 1;
      ints.const_operand_insts++;
    }
  }
};

}  // namespace ti_oracle
#endif
