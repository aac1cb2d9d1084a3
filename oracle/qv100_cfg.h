// TEST INFRASTRUCTURE -- not a product path (see oracle/README.md).
// SM7_QV100 parameters of the path, as the reference's tested config sets them:
//   gpu-simulator/gpgpu-sim/configs/tested-cfgs/SM7_QV100/gpgpusim.config:68,93,126-129,155
//   gpu-simulator/configs/tested-cfgs/SM7_QV100/trace.config:1-19 (spec_op_4..8 keep "4,4", TD.cc:417-425)
//   gpu-simulator/gpgpu-sim/configs/tested-cfgs/SM7_QV100/accelwattch_sass_sim.xml:48-59
//   absent coefficients keep 1.0 (gpgpu-sim/src/accelwattch/XML_Parse.cc:4523)
#ifndef TI_ORACLE_QV100_CFG_H
#define TI_ORACLE_QV100_CFG_H
#include <string.h>
#include "../include/traceingest.h"

namespace ti_oracle {
static inline void qv100(ti_config *c) {
  memset(c, 0, sizeof *c);
  c->coalesce_arch = 70;
  c->mem_warp_parts = 1;
  c->num_shmem_bank = 32;
  c->shmem_limited_broadcast = 0;
  c->gmem_skip_L1D = 0;
  c->clock_gated_lanes = 1;
  c->num_dp_units = 4;
  c->lat_int[0] = 2;    c->lat_int[1] = 2;
  c->lat_sp[0] = 2;     c->lat_sp[1] = 2;
  c->lat_dp[0] = 8;     c->lat_dp[1] = 4;
  c->lat_sfu[0] = 20;   c->lat_sfu[1] = 8;
  c->lat_tensor[0] = 2; c->lat_tensor[1] = 2;
  for (int j = 0; j < 8; ++j) { c->lat_spec[j][0] = 4; c->lat_spec[j][1] = 4; }
  c->lat_spec[1][0] = 200; c->lat_spec[1][1] = 4;
  c->lat_spec[2][0] = 2;   c->lat_spec[2][1] = 2;
  for (int j = 0; j < TI_N_COEFF; ++j) c->coeff[j] = 1.0;
  c->coeff[TI_C_INT] = 14.98768151;
  c->coeff[TI_C_FP] = 0.529670751;
  c->coeff[TI_C_DP] = 0.777229051;
  c->coeff[TI_C_INT_MUL] = 0.115098047;
  c->coeff[TI_C_FP_MUL] = 0.089517055;
  c->coeff[TI_C_SQRT] = 0.195089274;
  c->coeff[TI_C_LOG] = 0.125521663;
  c->coeff[TI_C_SIN] = 0.13336307;
  c->coeff[TI_C_EXP] = 0.36204415;
  c->coeff[TI_C_DP_MUL] = 0.1321288;
  c->coeff[TI_C_TENSOR] = 0.815454621;
  c->coeff[TI_C_TEX] = 0.115100088;
}
}  // namespace ti_oracle
#endif
