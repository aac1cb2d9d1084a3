"""Hand-written .traceg images for the corner cases SURVEY 8(c) lists.  Every VALID
case is something the reference parses without tripping an assert; the ERROR cases
are lines outside the tracer grammar, one TI_EF_* flag each."""

HEADER = """-kernel name = _Z4edgePv
-kernel id = 7
-grid dim = (2,1,1)
-block dim = (64,1,1)
-shmem = 0
-nregs = 16
-binary version = {bv}
-cuda stream id = 0
-shmem base_addr = 0x{shmem:016x}
-local mem base_addr = 0x{local:016x}
-nvbit version = 1.5.5
-accelsim tracer version = {tv}

#traces format = [line_num] PC mask dest_num [reg_dests] opcode src_num [reg_srcs] mem_width [adrrescompress?] [mem_addresses] immediate

"""

SHMEM = 0x00007F8C00000000
LOCAL = 0x00007F8C10000000


def image(tbs, bv=70, tv=3, shmem=SHMEM, local=LOCAL, tail="\n"):
    """tbs: list of (x, y, z, [ [lines of warp 0], [lines of warp 1], ... ])"""
    out = [HEADER.format(bv=bv, tv=tv, shmem=shmem, local=local)]
    for x, y, z, warps in tbs:
        warps = list(warps) + [[]] * (2 - len(warps))  # block dim is 64: the tracer lists both warps
        out.append("\n#BEGIN_TB\n\nthread block = %d,%d,%d\n" % (x, y, z))
        for w, lines in enumerate(warps):
            out.append("\nwarp = %d\ninsts = %d\n" % (w, len(lines)))
            for ln in lines:
                pre = "%d %d %d %d " % (x, y, z, w) if tv < 3 else ""
                out.append(pre + ln + "\n")
        out.append("\n#END_TB\n")
    return ("".join(out) + tail).encode()


def lst(addrs):
    return " ".join("0x%016x" % a for a in addrs)


W0 = [
    # the README example line: no immediate field at the end
    "0000 ffffffff 1 R1 IMAD.MOV.U32 2 R255 R255 0",
    # base+stride, full mask, then partial contiguous run
    "0010 ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x7f0000001000 4 0",
    "0020 00ffff00 1 R2 LDG.E.SYS 1 R4 4 1 0x7f0000001000 4 0",
    # stride mode with a NON-contiguous mask: lanes after the first gap read address 0 (trace_parser.cc:86-105)
    "0030 f0f0f0f0 0 STG.E.SYS 2 R4 R6 4 1 0x7f0000002000 4 0",
    # negative stride, 64-bit wrap-around of base + stride
    "0040 ffffffff 1 R2 LDG.E.64.SYS 1 R4 8 1 0x7f0000003000 -8 0",
    "0050 0000000f 1 R2 LDG.E.SYS 1 R4 4 1 0xfffffffffffffffc 4 0",
    # stride mode, mask 0 (tracer writes base 0 stride 0)
    "0060 00000000 1 R2 LDG.E.SYS 1 R4 4 1 0x0 0 0",
    # base+delta: mixed signs, single active lane (no deltas), missing deltas read as 0
    "0070 0000001f 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004000 4 -4 128 4096 0",
    "0080 00010000 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004100 0",
    "0090 000000ff 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004200 4 4",
    # list mode: scattered, lane with address 0 generates no access (abstract_hardware_model.cc:534-537)
    "00a0 0000000f 1 R2 LDG.E.SYS 1 R4 4 0 " + lst([0x7F0000005000, 0, 0x7F0000005004, 0x7F0000009000]) + " 0",
    # list mode with fewer tokens than active lanes: the rest read as 0
    "00b0 000000ff 1 R2 LDG.E.SYS 1 R4 4 0 " + lst([0x7F0000005100, 0x7F0000005104]),
    # straddling accesses: 8 bytes starting 4 before a sector end; 16 bytes crossing a 128 B line
    "00c0 00000003 1 R2 LDG.E.64.SYS 1 R4 8 0 " + lst([0x7F000000601C, 0x7F000000603C]) + " 0",
    "00d0 00000001 1 R2 LDG.E.128.SYS 1 R4 16 0 " + lst([0x7F0000006078]) + " 0",
    # widths: U8, U16, S16 (falls back to 4), 64, 128; STRONG.GPU bypasses L1 (cache op)
    "00e0 ffffffff 1 R2 LDG.E.U8.SYS 1 R4 1 1 0x7f0000007000 1 0",
    "00f0 ffffffff 1 R2 LDG.E.U16.SYS 1 R4 2 1 0x7f0000007100 2 0",
    "0100 ffffffff 1 R2 LDG.E.S16.SYS 1 R4 2 1 0x7f0000007200 2 0",
    "0110 ffffffff 1 R2 LDG.E.128.STRONG.GPU 1 R4 16 1 0x7f0000007300 16 0",
    "0120 ffffffff 1 R2 LDG.E.STRONG.SYS 1 R4 4 1 0x7f0000007400 4 0",
    # atomics: colliding lanes (same word) -> one transaction each; distinct words merge
    "0130 0000000f 1 R2 ATOMG.E.ADD.STRONG.GPU 2 R4 R6 4 0 " + lst([0x7F0000008000] * 3 + [0x7F0000008004]) + " 0",
    "0140 ffffffff 0 RED.E.ADD.STRONG.GPU 2 R4 R6 4 1 0x7f0000008100 4 0",
    # shared memory: conflict-free, 2-way, 32-way, broadcast
    "0150 ffffffff 1 R2 LDS.U.32 1 R4 4 1 0x100 4 0",
    "0160 ffffffff 1 R2 LDS.U.32 1 R4 4 1 0x100 8 0",
    "0170 ffffffff 1 R2 LDS.U.32 1 R4 4 1 0x100 128 0",
    "0180 ffffffff 1 R2 LDS.U.32 1 R4 4 1 0x100 0 0",
    "0190 ffffffff 0 STS.64 2 R4 R6 8 1 0x200 8 0",
    "01a0 0000ffff 1 R2 ATOMS.ADD 2 R4 R6 4 1 0x300 4 0",
    # generic LD/ST resolved by the first active lane: shared window, local window, global
    "01b0 ffffffff 1 R2 LD.E.SYS 1 R4 4 1 0x%x 4 0" % (SHMEM + 0x40),
    "01c0 ffffffff 1 R2 LD.E.SYS 1 R4 4 1 0x%x 4 0" % (LOCAL + 0x40),
    "01d0 fffffffe 0 ST.E.SYS 2 R4 R6 4 1 0x7f0000009000 4 0",
    # local space is decoded, not coalesced (rewritten per SM at issue time)
    "01e0 ffffffff 1 R2 LDL 1 R4 4 1 0x%x 4 0" % (LOCAL + 0x100),
    # LDC with and without addresses; BAR; MEMBAR; EXIT
    "01f0 ffffffff 1 R2 LDC 1 R4 0 0",
    "0200 ffffffff 1 R2 LDC.64 1 R4 8 1 0x160 0 0",
    "0210 ffffffff 0 BAR.SYNC 0 0 0",
    "0220 ffffffff 0 MEMBAR.SC.GPU 0 0 0",
    # MUFU / IMAD power-class overrides compare the WHOLE opcode (trace_driven.cc:208-222)
    "0230 ffffffff 1 R2 MUFU.SIN 1 R4 0 0",
    "0240 ffffffff 1 R2 MUFU.EX2 1 R4 0 0",
    "0250 ffffffff 1 R2 MUFU.RSQ 1 R4 0 0",
    "0260 ffffffff 1 R2 MUFU.LG2 1 R4 0 0",
    "0270 ffffffff 1 R2 MUFU.RCP64H 1 R4 0 0",
    "0280 ffffffff 1 R2 IMAD.MOV 2 R4 R5 0 0",
    "0290 ffffffff 1 R2 IMAD.IADD 2 R4 R5 0 0",
    "02a0 ffffffff 1 R2 IMAD.MOV.U32 2 R4 R5 0 0",
    # half precision halves the initiation interval; HMUL2 does not (trace_driven.cc:366-375)
    "02b0 ffffffff 1 R2 HFMA2.MMA 3 R4 R5 R6 0 0",
    "02c0 ffffffff 1 R2 HMUL2 2 R4 R5 0 0",
    # 0x-prefixed and upper-case hex, tabs and doubled spaces, trailing blanks, 5-digit pc
    "0x02d0 0xFFFFFFFF 1 R2 FFMA 3 R4 R5 R6 0 0",
    "02e0\tffffffff  1 R2   FADD 2 R4 R5 0 0  ",
    "12f00 ffffffff 1 R2 FADD 2 R4 R5 0 0",
    # register token with a suffix sscanf("R%d") ignores; four sources
    "0300 ffffffff 1 R2.reuse FFMA 4 R4 R5 R6 R7 0 0",
    # partial mask on an ALU op; mask 0 on a store
    "0310 0000ffff 1 R2 DADD 2 R4 R6 0 0",
    "0320 00000000 0 STG.E.SYS 2 R4 R6 4 1 0x0 0 0",
    # unknown address mode keeps zero addresses (value-initialised inst_memadd_info_t)
    "0330 ffffffff 1 R2 LDG.E.SYS 1 R4 4 3 0x7f000000a000 4 0",
    # leading zeros beyond the field width, zero-padded counts and register numbers
    "0000000350 00000000ffffffff 01 R007 FADD 02 R004 R5 00 0",
    "0000000000000360 ffffffff 1 R2 LDG.E.64.SYS 1 R4 8 1 0x00000000007f000000b000 8 0",
    # a width token after several modifiers; U-prefixed width; digits-only long token
    "0370 ffffffff 1 R2 LDG.E.CONSTANT.EF.128.SYS 1 R4 16 1 0x7f000000c000 16 0",
    "0380 ffffffff 1 R2 LDG.E.LTC128B.U16.CONSTANT 1 R4 2 1 0x7f000000d000 2 0",
    "0340 ffffffff 0 EXIT 0 0 0",
]

# One active lane: the tracer's base+stride compressor gives up (no second lane to take a stride from,
# tracer_tool.cu:609-614), so the line comes in base+delta form without deltas: base, then the immediate.
W_SINGLE = [
    "0000 00000001 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004100 0",
    "0010 80000000 0 STG.E.SYS 2 R4 R6 4 2 0x7f0000004104 0",
    # straddles a 32 B sector / a 128 B line
    "0020 00010000 1 R2 LDG.E.64.SYS 1 R4 8 2 0x7f000000411c 0",
    "0030 00000100 1 R2 LDG.E.128.SYS 1 R4 16 2 0x7f0000004178 0",
    # no immediate behind the base; a negative and a long one; upper-case 0X; address 0 (no access)
    "0040 00000002 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004200",
    "0050 00000004 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004204 -12",
    "0060 00000008 1 R2 LDG.E.SYS 1 R4 4 2 0X7F0000004208 1234567890",
    "0070 00000010 1 R2 LDG.E.SYS 1 R4 4 2 0x0 0",
    # the same through a byte and a half-word width; an unaligned word
    "0080 00000020 1 R2 LDG.E.U8.SYS 1 R4 1 2 0x7f0000004301 0",
    "0090 00000040 0 STG.E.U16.SYS 2 R4 R6 2 2 0x7f0000004302 0",
    "00a0 00000080 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004306 0",
    # not the global space / atomic / generic: these stay with the warp path
    "00b0 00000001 1 R2 LDS.U.32 1 R4 4 2 0x100 0",
    "00c0 00000001 1 R2 ATOMG.E.ADD.STRONG.GPU 2 R4 R6 4 2 0x7f0000004400 0",
    "00d0 00000001 1 R2 LD.E.SYS 1 R4 4 2 0x7f0000004500 0",
    # two active lanes in base+delta form next to them
    "00e0 00000101 1 R2 LDG.E.SYS 1 R4 4 2 0x7f0000004600 256 0",
    "00f0 ffffffff 0 EXIT 0 0 0",
]

VALID = {
    # name: (image bytes, config overrides)
    "edge_single_lane": (image([(0, 0, 0, [W_SINGLE, W_SINGLE[:4]]), (1, 0, 0, [W_SINGLE[4:], []])]), {}),
    "edge_single_lane_fermi": (image([(0, 0, 0, [W_SINGLE])]), dict(coalesce_arch=20)),
    "edge_v3": (image([(0, 0, 0, [W0, []]), (1, 0, 0, [W0[:5], W0[5:9]])]), {}),
    "edge_v2_prefix": (image([(0, 0, 0, [W0, []])], tv=2), {}),
    "edge_no_bases": (image([(0, 0, 0, [W0])], shmem=0, local=0), {}),
    "edge_no_trailing_newline": (image([(0, 0, 0, [W0[:3]])], tail="")[:-1], {}),
    "edge_fermi_segments": (image([(0, 0, 0, [W0])]), dict(coalesce_arch=20)),
    "edge_subwarps_broadcast": (image([(0, 0, 0, [W0])]), dict(mem_warp_parts=4, shmem_limited_broadcast=1, num_shmem_bank=16, clock_gated_lanes=0)),
    "edge_empty_kernel": (image([]), {}),
    # header-dense: ~150 thread blocks per 16 KiB chunk, more than the kernel's header parking area
    # holds, across several chunks; a base+stride instruction every 7th block keeps rows in play
    "edge_many_headers": (image([(i % 50, i // 50, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0"] if i % 3 else [],
                                                       ["0010 ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x7f00000%05x 4 0" % (i * 128)] if i % 7 == 0 else []])
                                 for i in range(700)]), {}),
    "edge_ampere": (image([(0, 0, 0, [[l for l in W0 if "MEMBAR" not in l] + ["0350 ffffffff 1 R2 LDSM.16.M88.4 1 R4 2 1 0x400 16 0"]])], bv=86), {}),
}


def _one(line):
    return image([(0, 0, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0", line]])])


LONG_LINE = "0000 ffffffff 1 R1 LDG.E.SYS 1 R2 4 0 " + " ".join(["0x00007f0000000000"] * 32) + " " + "0" * 600

ERRORS = {
    # name: (image bytes, expected TI_EF_* bit name)
    "bad_number_pc": (_one("00zz ffffffff 1 R1 FADD 2 R2 R3 0 0"), "BAD_NUMBER"),
    "bad_number_mask_overflow": (_one("0010 1ffffffff 1 R1 FADD 2 R2 R3 0 0"), "BAD_NUMBER"),
    "bad_number_address": (_one("0010 00000001 1 R1 LDG.E.SYS 1 R2 4 0 0x7f00zz 0"), "BAD_NUMBER"),
    "bad_number_stride": (_one("0010 ffffffff 1 R1 LDG.E.SYS 1 R2 4 1 0x7f0000001000 4x 0"), "BAD_NUMBER"),
    "bad_reg": (_one("0010 ffffffff 1 P1 FADD 2 R2 R3 0 0"), "BAD_REG"),
    "bad_reg_range": (_one("0010 ffffffff 1 R300 FADD 2 R2 R3 0 0"), "BAD_REG"),
    "too_many_dst": (_one("0010 ffffffff 2 R1 R2 FADD 2 R2 R3 0 0"), "TOO_MANY_REGS"),
    "too_many_src": (_one("0010 ffffffff 1 R1 FADD 5 R2 R3 R4 R5 R6 0 0"), "TOO_MANY_REGS"),
    "unknown_opcode": (_one("0010 ffffffff 1 R1 FROBNICATE.X 2 R2 R3 0 0"), "UNKNOWN_OPCODE"),
    "unknown_opcode_long": (_one("0010 ffffffff 1 R1 ABCDEFGHIJKLMNOP 2 R2 R3 0 0"), "UNKNOWN_OPCODE"),
    "zero_width": (_one("0010 ffffffff 1 R1 LDG.E.4.SYS 1 R2 4 1 0x7f0000001000 4 0"), "ZERO_WIDTH"),
    "zero_width_no_mem": (_one("0010 ffffffff 1 R1 LDG.E.SYS 1 R2 0 0"), "ZERO_WIDTH"),
    "truncated": (_one("0010 ffffffff 1 R1 FADD 2 R2 R3"), "TRUNCATED"),
    "bad_width": (_one("0010 ffffffff 1 R1 LDG.E.24.SYS 1 R2 3 1 0x7f0000001000 4 0"), "BAD_WIDTH"),
    "atomic_straddle": (_one("0010 00000001 1 R1 ATOMG.E.ADD.64.STRONG.GPU 2 R2 R3 8 0 0x00007f000000101c 0"), "ATOMIC_STRADDLE"),
    "long_opcode": (_one("0010 ffffffff 1 R1 FADD." + "X" * 300 + " 2 R2 R3 0 0"), "LONG_OPCODE"),
    "line_too_long": (_one(LONG_LINE), "LINE_TOO_LONG"),
    "undef_space": (image([(0, 0, 0, [["0010 ffffffff 1 R1 LDGSTS.E 1 R2 0 0"]])], bv=86), "UNDEF_SPACE"),
    "bad_structure_junk": (image([(0, 0, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0"]])]).replace(b"#END_TB", b"#END_OF_TB"), "BAD_STRUCTURE"),
    "bad_structure_no_insts": (image([(0, 0, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0"]])]).replace(b"insts = 1\n", b""), "BAD_STRUCTURE"),
    "framing_short": (image([(0, 0, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0"]])]).replace(b"insts = 1", b"insts = 2"), "FRAMING"),
    "framing_long": (image([(0, 0, 0, [["0000 ffffffff 1 R1 FADD 2 R2 R3 0 0"] * 3])]).replace(b"insts = 3", b"insts = 2"), "FRAMING"),
}
