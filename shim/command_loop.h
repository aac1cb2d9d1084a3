// The caller of the trace parser: Accel-Sim's command loop, gpu-simulator/main.cc:75-197, restated over
// the drop-in so that a host which does not build main.cc (another front end, a test bench) drives the path the way the
// simulator does.  What it keeps from the reference, line for line in behaviour:
//   * the window: with -gpgpu_concurrent_kernel_sm the headers of up to get_max_concurrent_kernel() kernels are parsed
//     ahead (parse_kernel_info + create_kernel_info), else one (main.cc:75-77, 101-107);
//   * launch rule: a kernel of the window starts when its CUDA stream is not busy, the GPU can take one and it was not
//     launched before; its stream is busy until it finishes (main.cc:109-122, 171-190);
//   * the other commands: memcpy -> perf_memcpy_to_gpu, ncclAllReduce -> gpu_tot_sim_cycle += nccl_allreduce_latency,
//     the four other NCCL commands are only announced (main.cc:91-97, 123-141);
//   * the inner loop: cycle() while active and no kernel finished; a finished kernel frees its stream, is handed to
//     kernel_finalizer and leaves the window; statistics are printed after it (main.cc:146-195);
//   * every stdout line.
// It is a template over the simulator's types, so the same code runs over gpgpu_sim / trace_kernel_info_t in a real
// tree and over the stand-ins of oracle/loop_stubs in the tests (which compare it with the reference's own main.cc).
#ifndef TI_COMMAND_LOOP_H
#define TI_COMMAND_LOOP_H

#include <assert.h>
#include <stdio.h>

#include <iostream>
#include <vector>

#include "trace_parser.h"

// MakeKernel: trace_kernel_info_t *(kernel_trace_t *)  -- main.cc's create_kernel_info with its context bound
template <class Sim, class Context, class KernelInfo, class MakeKernel>
void ti_command_loop(trace_parser &tracer, Sim *sim, Context *context, MakeKernel make_kernel) {
  const bool concurrent_kernel_sm = sim->getShaderCoreConfig()->gpgpu_concurrent_kernel_sm;
  const unsigned window_size = concurrent_kernel_sm ? sim->get_config().get_max_concurrent_kernel() : 1;
  assert(window_size > 0);
  std::vector<trace_command> commandlist = tracer.parse_commandlist_file();
  std::vector<unsigned long> busy_streams;
  std::vector<KernelInfo *> window;
  window.reserve(window_size);

  unsigned i = 0;
  while (i < commandlist.size() || !window.empty()) {
    // (the reference indexes commandlist[i] even when only the window is left to drain: a default command there)
    const command_type type = i < commandlist.size() ? commandlist[i].m_type : command_type::kernel_launch;
    if (type == command_type::cpu_gpu_mem_copy) {
      size_t addr, count;
      tracer.parse_memcpy_info(commandlist[i].command_string, addr, count);
      std::cout << "launching memcpy command : " << commandlist[i].command_string << std::endl;
      sim->perf_memcpy_to_gpu(addr, count);
      i++;
      continue;
    } else if (type == command_type::kernel_launch) {
      while (window.size() < window_size && i < commandlist.size()) {   // headers of the next kernels of the window
        KernelInfo *k = make_kernel(tracer.parse_kernel_info(commandlist[i].command_string));
        window.push_back(k);
        std::cout << "Header info loaded for kernel command : " << commandlist[i].command_string << std::endl;
        i++;
      }
      for (KernelInfo *k : window) {                                    // launch what may run
        bool stream_busy = false;
        for (unsigned long s : busy_streams) stream_busy = stream_busy || s == k->get_cuda_stream_id();
        if (!stream_busy && sim->can_start_kernel() && !k->was_launched()) {
          std::cout << "launching kernel name: " << k->get_name() << " uid: " << k->get_uid() << std::endl;
          sim->launch(k);
          k->set_launched();
          busy_streams.push_back(k->get_cuda_stream_id());
        }
      }
    } else if (type == command_type::ncclAllReduce) {
      const unsigned latency = sim->get_config().get_nccl_allreduce_latency();
      std::cout << "ncclAllReduce was run! Latency: " << latency << " cycles." << std::endl;
      sim->gpu_tot_sim_cycle += latency;
      i++;
    } else if (type == command_type::ncclCommInitAll) {
      std::cout << "ncclCommInitAll was run!" << std::endl;
      i++;
    } else if (type == command_type::ncclCommDestroy) {
      std::cout << "ncclCommDestroy was run!" << std::endl;
      i++;
    } else if (type == command_type::ncclGroupStart) {
      std::cout << "ncclGroupStart was run!" << std::endl;
      i++;
    } else if (type == command_type::ncclGroupEnd) {
      std::cout << "ncclGroupEnd was run!" << std::endl;
      i++;
    } else if (window.empty())
      assert(0 && "Undefined Command");

    bool active = false, sim_cycles = false;
    unsigned finished_kernel_uid = 0;
    do {
      if (!sim->active()) break;
      if (sim->active()) {
        sim->cycle();
        sim_cycles = true;
        sim->deadlock_check();
      } else if (sim->cycle_insn_cta_max_hit()) {
        context->the_gpgpusim->g_stream_manager->stop_all_running_kernels();
        break;
      }
      active = sim->active();
      finished_kernel_uid = sim->finished_kernel();
    } while (active && !finished_kernel_uid);

    if (finished_kernel_uid) {                                           // the finished kernel leaves the window
      KernelInfo *k = NULL;
      for (unsigned j = 0; j < window.size(); j++) {
        k = window.at(j);
        if (k->get_uid() == finished_kernel_uid) {
          for (size_t l = 0; l < busy_streams.size(); l++)
            if (busy_streams.at(l) == k->get_cuda_stream_id()) { busy_streams.erase(busy_streams.begin() + l); break; }
          tracer.kernel_finalizer(k->get_trace_info());
          delete k->entry();
          delete k;
          window.erase(window.begin() + j);
          break;
        }
      }
      assert(k);
      sim->print_stats();
    }
    if (sim_cycles) {
      sim->update_stats();
      context->print_simulation_time();
    }
    if (sim->cycle_insn_cta_max_hit()) {
      printf("GPGPU-Sim: ** break due to reaching the maximum cycles (or instructions) **\n");
      fflush(stdout);
      break;
    }
  }
  printf("GPGPU-Sim: *** simulation thread exiting ***\n");
  printf("GPGPU-Sim: *** exit detected ***\n");
  fflush(stdout);
}

#endif
