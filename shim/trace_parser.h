// Interface header of the B200 trace parser shim.
//
// trace-driven/trace_driven.{h,cc} and main.cc of the reference include "trace_parser.h" and use
// the names below (reference: gpu-simulator/trace-parser/trace_parser.h:12-123).  This file
// declares the same names with the same meaning -- nothing else is shared with the reference
// header -- so that those callers compile against the shim unchanged; trace_parser_b200.cc
// implements them over the C ABI of libtraceingest.so (include/traceingest.h).
//
// As with the reference's header, <cstdint>, <bitset> and <fstream> must already be visible
// (the reference's .cc files pull in <bits/stdc++.h> first).
#ifndef TRACE_PARSER_H
#define TRACE_PARSER_H

#include <assert.h>
#include <stdio.h>
#include <stdlib.h>

#include <string>
#include <vector>

// -- limits of one trace line (tracer_tool.cu writes at most one destination and four sources)
#define MAX_SRC 4
#define MAX_DST 1
#define WARP_SIZE 32

// -- address compression of a memory instruction: the mode digit after the width field
enum address_format { list_all = 0, base_stride = 1, base_delta = 2 };
// -- declared by the reference and unused on this path; kept so that includers still compile
enum address_scope { L1_CACHE = 1, L2_CACHE, SYS_MEM };
enum address_space { GLOBAL_MEM = 1, SHARED_MEM, LOCAL_MEM, TEX_MEM };

// -- one line of kernelslist.g; the five NCCL kinds are this fork's addition
enum command_type { kernel_launch = 1, cpu_gpu_mem_copy, gpu_cpu_mem_copy, ncclCommInitAll, ncclCommDestroy, ncclGroupStart, ncclGroupEnd, ncclAllReduce };
struct trace_command { std::string command_string; command_type m_type; };

// -- per-lane addresses of one memory instruction (= one 256-byte row of ti_view.addr_rows)
struct inst_memadd_info_t {
  uint64_t addrs[WARP_SIZE];
  int32_t width;
  // host-side decompressors, only reached through inst_trace_t::parse_from_string
  void base_delta_decompress(unsigned long long base_address, const std::vector<long long> &deltas, const std::bitset<WARP_SIZE> &mask);
  void base_stride_decompress(unsigned long long base_address, int stride, const std::bitset<WARP_SIZE> &mask);
};

// -- one warp instruction as trace_driven.cc reads it; filled from a 32-byte ti_inst_rec
struct inst_trace_t {
  unsigned m_pc, mask;
  unsigned reg_dsts_num, reg_dest[MAX_DST];
  std::string opcode;
  unsigned reg_srcs_num, reg_src[MAX_SRC];
  inst_memadd_info_t *memadd_info;   // owned; NULL for instructions without addresses

  inst_trace_t();
  inst_trace_t(const inst_trace_t &b);   // copies nothing, like the reference's (records are resized in place)
  ~inst_trace_t();

  std::vector<std::string> get_opcode_tokens() const;
  unsigned get_datawidth_from_opcode(const std::vector<std::string> &opcode) const;
  bool check_opcode_contain(const std::vector<std::string> &opcode, std::string param) const;
  // single-line host decode for callers that ask for it explicitly; the bulk path is on the GPU
  bool parse_from_string(std::string trace, unsigned tracer_version);
};

// -- header of one kernel-N.traceg (spellings "verion" as in the reference: callers use them)
struct kernel_trace_t {
  std::string kernel_name, nvbit_verion;
  unsigned kernel_id;
  unsigned grid_dim_x, grid_dim_y, grid_dim_z, tb_dim_x, tb_dim_y, tb_dim_z;
  unsigned shmem, nregs;
  unsigned long cuda_stream_id;
  unsigned binary_verion, trace_verion;
  unsigned long long shmem_base_addr, local_base_addr;
  std::ifstream *ifs;   // callers only pass it back to get_next_threadblock_traces: it names the device-side kernel object

  kernel_trace_t();
};

class trace_parser {
  struct impl;
  impl *m;
  std::string kernellist_filename;

 public:
  explicit trace_parser(const char *kernellist_filepath);
  ~trace_parser();

  // kernelslist.g -> commands; kernel paths get the list's directory in front
  std::vector<trace_command> parse_commandlist_file();
  // "MemcpyHtoD,0x<addr>,<bytes>"
  void parse_memcpy_info(const std::string &memcpy_command, size_t &add, size_t &count);
  // header lines of a kernel trace, echoed to stdout; the object lives until kernel_finalizer
  kernel_trace_t *parse_kernel_info(const std::string &kerneltraces_filepath);
  void kernel_finalizer(kernel_trace_t *trace_info);
  // next "#BEGIN_TB ... #END_TB" block in file order: one vector per warp of the thread block
  void get_next_threadblock_traces(std::vector<std::vector<inst_trace_t> *> threadblock_traces, unsigned trace_version, std::ifstream *ifs);
};

#endif
