"""numpy views of the C-ABI record layouts declared in include/traceingest.h."""
import numpy as np

INST = np.dtype([
    ("pc", "<u4"), ("mask", "<u4"), ("opc_off", "<u4"), ("mem_idx", "<u4"),
    ("dst", "u1"), ("src", "u1", (4,)), ("opcode_id", "u1"), ("op", "u1"),
    ("pack_a", "u1"), ("pack_b", "u1"), ("pack_c", "u1"), ("data_size", "u1"),
    ("cycles", "u1"), ("latency", "<u2"), ("initiation_interval", "u1"), ("opcode_len", "u1"),
])
TXN = np.dtype([
    ("addr", "<u8"), ("lane_mask", "<u4"), ("byte_mask", "<u4", (4,)),
    ("size", "u1"), ("sector_mask", "u1"), ("type", "u1"), ("is_write", "u1"),
])
TB = np.dtype([
    ("text_off", "<u8"), ("x", "<i4"), ("y", "<i4"), ("z", "<i4"),
    ("first_warp", "<u4"), ("first_inst", "<u4"), ("line_len", "<u4"),
])
MEM_DESC = np.dtype([("base", "<u8"), ("stride", "<i4"), ("meta", "<u4")])   # ti_mem_desc
WARP = np.dtype([("warp_id", "<u4"), ("n_insts", "<u4"), ("first_inst", "<u4"), ("reserved", "<u4")])
COUNTERS = np.dtype([
    ("hist", "<u8", (18, 33)), ("op_count", "<u8", (32,)), ("space_lanes", "<u8", (16, 2)),
    ("const_operand_insts", "<u8"),
    ("n_inst", "<u8"), ("n_mem", "<u8"), ("n_txn", "<u8"), ("n_tb", "<u8"), ("n_warp", "<u8"),
    ("n_begin_tb", "<u8"), ("n_end_tb", "<u8"), ("text_bytes", "<u8"), ("smem_cycles", "<u8"),
    ("reserved", "<u8", (6,)),
])
POWER_FIELDS = [
    "ialu", "imul", "imul24", "imul32", "idiv", "fpalu", "fpmul", "fpdiv", "dpalu", "dpmul", "dpdiv",
    "sqrt_", "log_", "sin_", "exp_", "tensor", "tex", "sp_accesses", "sfu_accesses", "mem_accesses",
    "const_accesses", "active_exu_threads", "active_exu_warps", "n_load_insn", "n_store_insn",
    "n_shmem_insn", "n_sstarr_insn", "n_const_insn", "n_param_insn", "n_tex_insn",
]
POWER = np.dtype([(n, "<f8") for n in POWER_FIELDS])

assert INST.itemsize == 32 and TXN.itemsize == 32 and TB.itemsize == 32 and WARP.itemsize == 16


# ---------------------------------------------------------------- ti_config
CONFIG = np.dtype([
    ("device", "<i4"), ("coalesce_arch", "<i4"), ("mem_warp_parts", "<i4"), ("num_shmem_bank", "<i4"),
    ("shmem_limited_broadcast", "<i4"), ("gmem_skip_L1D", "<i4"), ("clock_gated_lanes", "<i4"),
    ("num_dp_units", "<i4"),
    ("lat_int", "<u4", (2,)), ("lat_sp", "<u4", (2,)), ("lat_dp", "<u4", (2,)), ("lat_sfu", "<u4", (2,)),
    ("lat_tensor", "<u4", (2,)), ("lat_spec", "<u4", (8, 2)),
    ("coeff", "<f8", (17,)),
    ("cap_inst_per_byte", "<f8"), ("cap_mem_per_byte", "<f8"), ("cap_txn_per_byte", "<f8"),
])
assert CONFIG.itemsize == 296

COEFF_ORDER = ["INT", "INT_MUL", "INT_MUL24", "INT_MUL32", "INT_DIV", "FP", "DP", "FP_MUL", "FP_DIV",
               "DP_MUL", "DP_DIV", "SQRT", "LOG", "SIN", "EXP", "TENSOR", "TEX"]


def qv100_config(**over):
    """SM7_QV100 parameters (values cited in oracle/qv100_cfg.h), with overrides."""
    c = np.zeros(1, dtype=CONFIG)[0]
    c["coalesce_arch"] = 70
    c["mem_warp_parts"] = 1
    c["num_shmem_bank"] = 32
    c["clock_gated_lanes"] = 1
    c["num_dp_units"] = 4
    c["lat_int"] = (2, 2)
    c["lat_sp"] = (2, 2)
    c["lat_dp"] = (8, 4)
    c["lat_sfu"] = (20, 8)
    c["lat_tensor"] = (2, 2)
    spec = np.full((8, 2), 4, dtype=np.uint32)
    spec[1] = (200, 4)
    spec[2] = (2, 2)
    c["lat_spec"] = spec
    co = dict.fromkeys(COEFF_ORDER, 1.0)
    co.update(INT=14.98768151, FP=0.529670751, DP=0.777229051, INT_MUL=0.115098047, FP_MUL=0.089517055,
              SQRT=0.195089274, LOG=0.125521663, SIN=0.13336307, EXP=0.36204415, DP_MUL=0.1321288,
              TENSOR=0.815454621, TEX=0.115100088)
    c["coeff"] = [co[k] for k in COEFF_ORDER]
    for k, v in over.items():
        c[k] = v
    return c

# ti_local_issue
LOCAL_ISSUE = np.dtype([("inst", "<u4"), ("sid", "<u4"), ("tid0", "<u4")])
