// TEST INFRASTRUCTURE -- one driver, two builds: against the REFERENCE's
// trace-parser/trace_parser.{h,cc} (oracle/_ref/tp_ref) and against the drop-in shim
// (shim/trace_parser.h + trace_parser_b200.cc + libtraceingest).  It uses nothing but
// the public trace_parser interface the simulator uses (main.cc:64-175,
// trace_driven.cc:568-586) and prints every field it can see; tests diff the two
// outputs byte for byte, stdout echo lines included.
#include <bits/stdc++.h>

#include "trace_parser.h"

// `tp_driver time <kernelslist.g>`: the same walk without printing -- wall seconds of parse_kernel_info +
// get_next_threadblock_traces over every thread block of every kernel, instructions seen, peak resident set.
// (stdout of the parser itself, its header and "thread block" echo lines, is left to the caller to discard.)
static int time_mode(const char *list) {
  const auto t0 = std::chrono::steady_clock::now();
  trace_parser tracer(list);
  unsigned long long n_inst = 0, n_mem = 0, n_tb_seen = 0, sum = 0;
  for (const trace_command &c : tracer.parse_commandlist_file()) {
    if (c.m_type != command_type::kernel_launch) continue;
    kernel_trace_t *k = tracer.parse_kernel_info(c.command_string);
    const unsigned n_warps = (k->tb_dim_x * k->tb_dim_y * k->tb_dim_z + 31) / 32;
    std::vector<std::vector<inst_trace_t>> warps(n_warps);
    std::vector<std::vector<inst_trace_t> *> ptrs;
    for (auto &w : warps) ptrs.push_back(&w);
    const unsigned long long n_tb = (unsigned long long)k->grid_dim_x * k->grid_dim_y * k->grid_dim_z;
    for (unsigned long long tb = 0; tb < n_tb; ++tb) {
      tracer.get_next_threadblock_traces(ptrs, k->trace_verion, k->ifs);
      ++n_tb_seen;
      for (auto &w : warps)
        for (const inst_trace_t &t : w) {
          ++n_inst;
          sum += t.m_pc + t.mask + t.opcode.size();
          if (t.memadd_info) { ++n_mem; sum += t.memadd_info->addrs[0] + t.memadd_info->addrs[31]; }
        }
    }
    tracer.kernel_finalizer(k);
  }
  const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  // VmHWM of this address space (ru_maxrss would carry the parent's high-water mark across fork + exec)
  double hwm_mb = 0, anon_mb = 0;
  if (FILE *f = fopen("/proc/self/status", "r")) {
    char line[256];
    while (fgets(line, sizeof line, f)) {
      long kb;
      if (sscanf(line, "VmHWM: %ld kB", &kb) == 1) hwm_mb = kb / 1024.0;
      if (sscanf(line, "RssAnon: %ld kB", &kb) == 1) anon_mb = kb / 1024.0;
    }
    fclose(f);
  }
  fprintf(stderr, "{\"seconds\": %.3f, \"n_inst\": %llu, \"n_mem\": %llu, \"n_tb\": %llu, \"checksum\": %llu, \"peak_rss_mb\": %.1f, \"rss_anon_at_end_mb\": %.1f}\n", sec,
          n_inst, n_mem, n_tb_seen, sum, hwm_mb, anon_mb);
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  if (argc > 2 && !strcmp(argv[1], "time")) return time_mode(argv[2]);
  trace_parser tracer(argv[1]);
  std::vector<trace_command> cmds = tracer.parse_commandlist_file();
  for (const trace_command &c : cmds) {
    // paths differ between runs only by the list's directory, which both builds share
    printf("CMD %d %s\n", (int)c.m_type, c.command_string.c_str());
    if (c.m_type == command_type::cpu_gpu_mem_copy) {
      size_t a = 0, n = 0;
      tracer.parse_memcpy_info(c.command_string, a, n);
      printf("MEMCPY %zx %zu\n", a, n);
    } else if (c.m_type == command_type::kernel_launch) {
      kernel_trace_t *k = tracer.parse_kernel_info(c.command_string);
      printf("KERNEL name=%s id=%u grid=%u,%u,%u block=%u,%u,%u shmem=%u nregs=%u stream=%lu bin=%u ver=%u nvbit=[%s] sb=%llx lb=%llx\n",
             k->kernel_name.c_str(), k->kernel_id, k->grid_dim_x, k->grid_dim_y, k->grid_dim_z, k->tb_dim_x, k->tb_dim_y,
             k->tb_dim_z, k->shmem, k->nregs, k->cuda_stream_id, k->binary_verion, k->trace_verion, k->nvbit_verion.c_str(),
             k->shmem_base_addr, k->local_base_addr);
      const unsigned n_warps = (k->tb_dim_x * k->tb_dim_y * k->tb_dim_z + 31) / 32;
      std::vector<std::vector<inst_trace_t>> warps(n_warps);
      std::vector<std::vector<inst_trace_t> *> ptrs;
      for (auto &w : warps) ptrs.push_back(&w);
      const unsigned n_tb = k->grid_dim_x * k->grid_dim_y * k->grid_dim_z;
      for (unsigned tb = 0; tb < n_tb + 1; ++tb) {  // one call past the end, like a drained kernel
        tracer.get_next_threadblock_traces(ptrs, k->trace_verion, k->ifs);
        fflush(stdout);
        for (unsigned w = 0; w < n_warps; ++w) {
          printf("W %u %zu\n", w, warps[w].size());
          for (const inst_trace_t &t : warps[w]) {
            printf("I %x %x %u", t.m_pc, t.mask, t.reg_dsts_num);
            for (unsigned i = 0; i < t.reg_dsts_num; ++i) printf(" R%u", t.reg_dest[i]);
            printf(" %s %u", t.opcode.c_str(), t.reg_srcs_num);
            for (unsigned i = 0; i < t.reg_srcs_num; ++i) printf(" R%u", t.reg_src[i]);
            if (t.memadd_info) {
              printf(" M %d", t.memadd_info->width);
              for (int s = 0; s < 32; ++s) printf(" %llx", (unsigned long long)t.memadd_info->addrs[s]);
            }
            printf("\n");
          }
        }
      }
      tracer.kernel_finalizer(k);
    }
  }
  return 0;
}
