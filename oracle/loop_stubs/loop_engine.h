// TEST INFRASTRUCTURE -- a deterministic stand-in for the timing model, so that the reference's command loop
// (gpu-simulator/main.cc:75-197, compiled where it lies) and its restatement (shim/command_loop.h) can be run
// over the same inputs and their event logs compared.  Everything main.cc touches is declared here with the
// reference's names; the real classes need the whole simulator (bison / flex generated sources).
//
// The "GPU": a launched kernel pulls all of its thread blocks through the trace_parser it was created with,
// then stays busy for 3 + (instructions % 5) cycles; at most TI_LOOP_MAX_RUNNING (2) kernels run at once.
#ifndef TI_LOOP_ENGINE_H
#define TI_LOOP_ENGINE_H
#include <assert.h>
#include <locale.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>

#include <deque>
#include <iostream>
#include <string>
#include <vector>

struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
typedef void *option_parser_t;
inline option_parser_t option_parser_create() { return NULL; }
inline void option_parser_cmdline(option_parser_t, int, const char **) {}
inline void option_parser_print(option_parser_t, FILE *) {}
inline void icnt_reg_options(option_parser_t) {}
inline void print_splash() {}
static const char *g_accelsim_version = "loop-stub";

inline unsigned env_u(const char *name, unsigned dflt) { const char *e = getenv(name); return e ? (unsigned)atoi(e) : dflt; }

class gpgpu_context;
struct gpgpu_ptx_sim_info { unsigned smem, regs; };
class trace_config {
 public:
  void reg_options(option_parser_t) {}
  void parse_config() {}
  const char *get_traces_filename() { return getenv("TI_LOOP_KERNELSLIST"); }
};
class trace_function_info {
 public:
  trace_function_info(const gpgpu_ptx_sim_info &, gpgpu_context *) {}
  void set_name(const char *n) { name = n; }
  std::string name;
};

class trace_kernel_info_t {
 public:
  trace_kernel_info_t(dim3 grid, dim3 block, trace_function_info *f, trace_parser *parser, trace_config *, kernel_trace_t *info)
      : m_grid(grid), m_block(block), m_f(f), m_parser(parser), m_info(info), m_launched(false) { m_uid = ++next_uid(); }
  static unsigned &next_uid() { static unsigned u = 0; return u; }   // kernel_info_t::m_next_uid starts at 1
  unsigned long get_cuda_stream_id() { return m_info->cuda_stream_id; }
  bool was_launched() { return m_launched; }
  void set_launched() { m_launched = true; }
  std::string get_name() const { return m_f->name; }
  unsigned get_uid() const { return m_uid; }
  kernel_trace_t *get_trace_info() { return m_info; }
  trace_function_info *entry() { return m_f; }
  dim3 m_grid, m_block;
  trace_function_info *m_f;
  trace_parser *m_parser;
  kernel_trace_t *m_info;
  bool m_launched;
  unsigned m_uid;
};

struct shader_core_config { bool gpgpu_concurrent_kernel_sm; };
class gpgpu_sim_config {
 public:
  explicit gpgpu_sim_config(gpgpu_context *) {}
  void reg_options(option_parser_t) {}
  void init() {}
  unsigned get_max_concurrent_kernel() const { return env_u("TI_LOOP_WINDOW", 4); }
  unsigned get_nccl_allreduce_latency() const { return env_u("TI_LOOP_NCCL_LATENCY", 100); }
};

class gpgpu_sim {
 public:
  gpgpu_sim(const gpgpu_sim_config &c, gpgpu_context *) : gpu_tot_sim_cycle(0), m_cfg(c) { m_shader.gpgpu_concurrent_kernel_sm = env_u("TI_LOOP_CONCURRENT", 1) != 0; }
  virtual ~gpgpu_sim() {}
  void init() {}
  const shader_core_config *getShaderCoreConfig() { return &m_shader; }
  const gpgpu_sim_config &get_config() const { return m_cfg; }
  void perf_memcpy_to_gpu(size_t addr, size_t count) { printf("ENGINE memcpy %zx %zu\n", addr, count); }
  bool can_start_kernel() { return m_running.size() < env_u("TI_LOOP_MAX_RUNNING", 2); }
  void launch(trace_kernel_info_t *k) {
    const unsigned n_warps = (k->m_block.x * k->m_block.y * k->m_block.z + 31) / 32;
    std::vector<std::vector<inst_trace_t> > warps(n_warps);
    std::vector<std::vector<inst_trace_t> *> ptrs;
    for (size_t w = 0; w < warps.size(); ++w) ptrs.push_back(&warps[w]);
    unsigned long long n_inst = 0, sum = 0;
    const unsigned long long n_tb = (unsigned long long)k->m_grid.x * k->m_grid.y * k->m_grid.z;
    for (unsigned long long tb = 0; tb < n_tb; ++tb) {
      k->m_parser->get_next_threadblock_traces(ptrs, k->m_info->trace_verion, k->m_info->ifs);
      for (size_t w = 0; w < warps.size(); ++w)
        for (size_t i = 0; i < warps[w].size(); ++i) { ++n_inst; sum += warps[w][i].m_pc + warps[w][i].mask + (warps[w][i].memadd_info ? warps[w][i].memadd_info->addrs[0] : 0); }
    }
    printf("ENGINE launch uid=%u insts=%llu checksum=%llu\n", k->get_uid(), n_inst, sum);
    run r = {k->get_uid(), 3 + (unsigned)(n_inst % 5)};
    m_running.push_back(r);
  }
  bool active() { return !m_running.empty() || !m_done.empty(); }
  void cycle() {
    ++gpu_tot_sim_cycle;
    for (size_t i = 0; i < m_running.size();) {
      if (--m_running[i].left == 0) { m_done.push_back(m_running[i].uid); m_running.erase(m_running.begin() + i); }
      else ++i;
    }
  }
  void deadlock_check() {}
  bool cycle_insn_cta_max_hit() { return gpu_tot_sim_cycle >= env_u("TI_LOOP_MAX_CYCLES", 1000000); }
  unsigned finished_kernel() {
    if (m_done.empty()) return 0;
    const unsigned u = m_done.front();
    m_done.pop_front();
    return u;
  }
  void print_stats() { printf("ENGINE stats tot_cycle=%llu running=%zu\n", gpu_tot_sim_cycle, m_running.size()); }
  void update_stats() {}
  unsigned long long gpu_tot_sim_cycle;

 private:
  struct run { unsigned uid, left; };
  std::vector<run> m_running;
  std::deque<unsigned> m_done;
  shader_core_config m_shader;
  gpgpu_sim_config m_cfg;
};
class trace_gpgpu_sim : public gpgpu_sim {
 public:
  trace_gpgpu_sim(const gpgpu_sim_config &c, gpgpu_context *ctx) : gpgpu_sim(c, ctx) {}
};

class stream_manager {
 public:
  stream_manager(gpgpu_sim *, bool) {}
  void stop_all_running_kernels() { printf("ENGINE stop_all_running_kernels\n"); }
};
struct cuda_sim_stub { bool g_cuda_launch_blocking; void ptx_opcocde_latency_options(option_parser_t) {} };
struct GPGPUsim_ctx_stub {
  gpgpu_sim *g_the_gpu;
  gpgpu_sim_config *g_the_gpu_config;
  stream_manager *g_stream_manager;
  time_t g_simulation_starttime;
};
class gpgpu_context {
 public:
  gpgpu_context() { the_gpgpusim = new GPGPUsim_ctx_stub(); func_sim = new cuda_sim_stub(); func_sim->g_cuda_launch_blocking = false; }
  void ptx_reg_options(option_parser_t) {}
  void print_simulation_time() {}
  GPGPUsim_ctx_stub *the_gpgpusim;
  cuda_sim_stub *func_sim;
};
#endif
