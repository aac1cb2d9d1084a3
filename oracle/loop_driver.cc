// TEST INFRASTRUCTURE -- the restated command loop (shim/command_loop.h) over the stand-in timing model of
// oracle/loop_stubs, set up the way main.cc:55-74 sets the real one up.  tests/test_command_loop.py compares its
// stdout with the reference's own main.cc built over the same stand-ins.
#include "trace_driven.h"   // loop_stubs: the trace parser header + the stand-ins

#include "command_loop.h"

int main() {
  std::cout << "Accel-Sim [build " << g_accelsim_version << "]";
  gpgpu_context *ctx = new gpgpu_context();
  trace_config tconfig;
  srand(1);
  fprintf(stdout, "GPGPU-Sim: Configuration options:\n\n");
  ctx->the_gpgpusim->g_the_gpu_config = new gpgpu_sim_config(ctx);
  ctx->the_gpgpusim->g_the_gpu = new trace_gpgpu_sim(*ctx->the_gpgpusim->g_the_gpu_config, ctx);
  ctx->the_gpgpusim->g_stream_manager = new stream_manager(ctx->the_gpgpusim->g_the_gpu, false);
  gpgpu_sim *sim = ctx->the_gpgpusim->g_the_gpu;
  sim->init();
  trace_parser tracer(tconfig.get_traces_filename());
  tconfig.parse_config();
  ti_command_loop<gpgpu_sim, gpgpu_context, trace_kernel_info_t>(tracer, sim, ctx, [&](kernel_trace_t *info) {
    gpgpu_ptx_sim_info pinfo;
    pinfo.smem = info->shmem;
    pinfo.regs = info->nregs;
    trace_function_info *f = new trace_function_info(pinfo, ctx);
    f->set_name(info->kernel_name.c_str());
    return new trace_kernel_info_t(dim3(info->grid_dim_x, info->grid_dim_y, info->grid_dim_z), dim3(info->tb_dim_x, info->tb_dim_y, info->tb_dim_z), f,
                                   &tracer, &tconfig, info);
  });
  return 0;
}
