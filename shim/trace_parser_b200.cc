// C++ host side above the C ABI: the reference's trace_parser interface
// (gpu-simulator/trace-parser/trace_parser.cc:216-447) implemented on top of
// libtraceingest.so.  Behaviour kept: stdout echo of header and "thread block" lines
// in order (trace_parser.cc:311,370,427), exit(1) on unopenable files (:224-227,
// :306-309), abort() where the reference asserts, ownership of kernel_trace_t / ifs
// (:301-302, :379-385), thread blocks handed out in file order (:387-447).
//
// What changes: get_next_threadblock_traces decodes the kernel file on the GPU one
// WINDOW at a time (256 MiB of text by default, TI_SHIM_WINDOW_MB; a window starts at
// "#BEGIN_TB") and every call then only copies one thread block of finished records
// into the caller's vectors.  The next window is decoded when the current one is used up
// and reuses its buffers, so the drop-in holds one window of results whatever the size
// of the file -- the reference holds one thread block per call (trace_parser.cc:387-447).
// The results travel in compact form (ti_view_compact): the 256-byte address row of a
// base+stride instruction is expanded here, on hand-out, from its 16-byte descriptor.
#include <bits/stdc++.h>

#include "trace_parser.h"

#include "../include/traceingest.h"

namespace {

ti_config g_cfg;
bool g_cfg_set = false;

[[noreturn]] void die(const char *what, ti_status st) {
  // the reference has no error codes: it asserts (TP.cc:151,160,417-439; TD.cc:205,264-319)
  fprintf(stderr, "trace_parser(b200): %s: %s -- %s\n", what, ti_strerror(st), ti_last_error());
  fflush(stdout);
  abort();
}

bool all_digits(const std::string &s, size_t from) {
  if (s.size() <= from) return false;
  for (size_t i = from; i < s.size(); ++i)
    if (s[i] < '0' || s[i] > '9') return false;
  return true;
}

struct KernelState {
  ti_kernel *k = nullptr;
  size_t body_len = 0;
  const char *body = nullptr;          // host body (the library's mapping of the file)
  ti_view_compact view;                // the current window
  std::vector<size_t> cuts;            // window boundaries in the body, all at "#BEGIN_TB"
  size_t next_window = 0;
  bool planned = false;
  uint64_t next_tb = 0;                // next thread block of the current window
};

// get_datawidth_from_opcode(get_opcode_tokens()) without building the token vector: the first '.'-separated token that
// is all digits, or 'U' followed by digits, gives the width in bits (trace_parser.cc:16-20,51-76); 4 bytes otherwise
unsigned width_from_opcode(const char *p, size_t n) {
  size_t b = 0;
  while (b < n) {
    size_t e = b;
    while (e < n && p[e] != '.') ++e;
    const size_t d = b + (p[b] == 'U' ? 1 : 0);
    if (d < e) {
      unsigned long v = 0;
      size_t i = d;
      for (; i < e && p[i] >= '0' && p[i] <= '9'; ++i) v = v * 10 + (unsigned long)(p[i] - '0');
      if (i == e) return (unsigned)v / 8;
    }
    b = e + 1;
  }
  return 4;
}

size_t window_bytes() {
  if (const char *b = getenv("TI_SHIM_WINDOW_BYTES")) { const long long v = atoll(b); if (v > 0) return (size_t)v; }   // (tests: tiny windows)
  const char *e = getenv("TI_SHIM_WINDOW_MB");
  const long mb = e ? atol(e) : 256;
  return (size_t)(mb > 0 ? mb : 256) << 20;
}

}  // namespace

// Optional: coalescer / latency / power parameters for the device-side stages that go
// beyond inst_trace_t (transactions, counters).  Defaults to the SM7_QV100 values.
extern "C" void trace_parser_b200_set_config(const ti_config *cfg) { g_cfg = *cfg; g_cfg_set = true; }

struct trace_parser::impl {
  ti_ctx *ctx = nullptr;
  std::unordered_map<std::ifstream *, KernelState> kernels;
  double s_ctx = 0, s_decode = 0;      // TI_SHIM_TIMING=1: seconds in ti_ctx_create / in the window decodes, printed by the destructor
  ti_ctx *context() {
    if (!ctx) {
      if (!g_cfg_set) ti_config_qv100(&g_cfg);
      const auto t0 = std::chrono::steady_clock::now();
      ti_status st = ti_ctx_create(&g_cfg, &ctx);
      s_ctx += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      if (st != TI_OK) die("ti_ctx_create", st);
    }
    return ctx;
  }
};

// ---- inst_trace_t / kernel_trace_t value semantics (trace_parser.cc:31-42, 78-84)
inst_trace_t::inst_trace_t() { memadd_info = NULL; }
inst_trace_t::~inst_trace_t() { delete memadd_info; }
inst_trace_t::inst_trace_t(const inst_trace_t &) { memadd_info = NULL; }  // the reference copies nothing either

kernel_trace_t::kernel_trace_t() {
  kernel_name = "Empty";
  shmem_base_addr = 0;
  local_base_addr = 0;
  binary_verion = 0;
  trace_verion = 0;
}

// trace_parser.cc:52-76, used by trace_driven.cc:188,272
std::vector<std::string> inst_trace_t::get_opcode_tokens() const {
  std::vector<std::string> out;
  size_t b = 0;
  while (b <= opcode.size()) {
    size_t e = opcode.find('.', b);
    if (e == std::string::npos) e = opcode.size();
    if (e > b) out.push_back(opcode.substr(b, e - b));
    b = e + 1;
  }
  return out;
}

bool inst_trace_t::check_opcode_contain(const std::vector<std::string> &toks, std::string param) const {
  return std::find(toks.begin(), toks.end(), param) != toks.end();
}

unsigned inst_trace_t::get_datawidth_from_opcode(const std::vector<std::string> &toks) const {
  for (const std::string &t : toks) {
    if (all_digits(t, 0)) return (unsigned)std::stoi(t) / 8;
    if (t[0] == 'U' && all_digits(t, 1)) return (unsigned)strtoul(t.c_str() + 1, NULL, 10) / 8;
  }
  return 4;
}

// trace_parser.cc:86-125: kept for API completeness (the device does this in bulk)
void inst_memadd_info_t::base_stride_decompress(unsigned long long base, int stride, const std::bitset<WARP_SIZE> &mask) {
  int s = 0;
  for (; s < WARP_SIZE && !mask.test(s); ++s) addrs[s] = 0;
  unsigned long long a = base;
  for (bool first = true; s < WARP_SIZE && mask.test(s); ++s, first = false) {
    if (!first) a += stride;
    addrs[s] = a;
  }
  // the lane that ends the run keeps whatever it held; every later lane is zeroed
  for (++s; s < WARP_SIZE; ++s) addrs[s] = 0;
}

void inst_memadd_info_t::base_delta_decompress(unsigned long long base, const std::vector<long long> &deltas,
                                               const std::bitset<WARP_SIZE> &mask) {
  long long last = 0;
  size_t k = 0;
  bool first = true;
  for (int s = 0; s < WARP_SIZE; ++s) {
    if (!mask.test(s)) { addrs[s] = 0; continue; }
    if (first) { last = (long long)base; first = false; }
    else { assert(k < deltas.size()); last += deltas[k++]; }
    addrs[s] = (uint64_t)last;
  }
}

// One line on the host, for callers that use inst_trace_t stand-alone.
bool inst_trace_t::parse_from_string(std::string trace, unsigned trace_version) {
  std::istringstream ss(trace);
  std::string tok;
  if (trace_version < 3) { unsigned a; for (int i = 0; i < 4; ++i) ss >> std::dec >> a; }
  ss >> std::hex >> m_pc >> mask >> std::dec >> reg_dsts_num;
  assert(reg_dsts_num <= MAX_DST);
  for (unsigned i = 0; i < reg_dsts_num; ++i) { ss >> tok; sscanf(tok.c_str(), "R%u", &reg_dest[i]); }
  ss >> opcode >> reg_srcs_num;
  assert(reg_srcs_num <= MAX_SRC);
  for (unsigned i = 0; i < reg_srcs_num; ++i) { ss >> tok; sscanf(tok.c_str(), "R%u", &reg_src[i]); }
  unsigned width = 0, mode = 0;
  ss >> width;
  if (!width) return true;
  std::bitset<WARP_SIZE> bits(mask);
  memadd_info = new inst_memadd_info_t();
  memadd_info->width = (int32_t)get_datawidth_from_opcode(get_opcode_tokens());
  ss >> mode;
  if (mode == list_all) {
    for (int s = 0; s < WARP_SIZE; ++s) if (bits.test(s)) ss >> std::hex >> memadd_info->addrs[s];
  } else if (mode == base_stride) {
    unsigned long long base = 0; int stride = 0;
    ss >> std::hex >> base >> std::dec >> stride;
    memadd_info->base_stride_decompress(base, stride, bits);
  } else if (mode == base_delta) {
    unsigned long long base = 0; std::vector<long long> deltas;
    ss >> std::hex >> base >> std::dec;
    for (int s = 0; s < WARP_SIZE; ++s) if (bits.test(s)) { long long d = 0; ss >> d; deltas.push_back(d); }
    memadd_info->base_delta_decompress(base, deltas, bits);
  }
  return true;
}

trace_parser::trace_parser(const char *kernellist_filepath) : m(new impl) { kernellist_filename = kernellist_filepath; }

trace_parser::~trace_parser() {
  if (getenv("TI_SHIM_TIMING")) fprintf(stderr, "trace_parser(b200): ti_ctx_create %.3f s, window decodes %.3f s\n", m->s_ctx, m->s_decode);
  for (auto &kv : m->kernels) ti_kernel_close(kv.second.k);
  if (m->ctx) ti_ctx_destroy(m->ctx);
  delete m;
}

std::vector<trace_command> trace_parser::parse_commandlist_file() {
  std::ifstream fs(kernellist_filename);
  if (!fs.is_open()) {
    std::cout << "Unable to open file: " << kernellist_filename << std::endl;
    exit(1);
  }
  std::string text((std::istreambuf_iterator<char>(fs)), std::istreambuf_iterator<char>());
  std::string directory(kernellist_filename);
  const size_t slash = directory.rfind('/');
  if (slash != std::string::npos) directory = directory.substr(0, slash);
  size_t n = 0;
  ti_parse_commandlist(text.data(), text.size(), NULL, 0, &n);
  std::vector<ti_command> cmds(n ? n : 1);
  ti_parse_commandlist(text.data(), text.size(), cmds.data(), cmds.size(), &n);
  std::vector<trace_command> out;
  for (size_t i = 0; i < n; ++i) {
    trace_command c;
    c.m_type = (command_type)cmds[i].kind;
    c.command_string = text.substr(cmds[i].line_off, cmds[i].line_len);
    if (c.m_type == kernel_launch) c.command_string = directory + "/" + c.command_string;
    out.push_back(c);
  }
  return out;
}

void trace_parser::parse_memcpy_info(const std::string &cmd, size_t &address, size_t &count) {
  uint64_t a = 0, n = 0;
  ti_status st = ti_parse_memcpy(cmd.data(), cmd.size(), &a, &n);
  if (st != TI_OK) die("parse_memcpy_info", st);  // assert(params.size() == 3)
  address = a;
  count = n;
}

kernel_trace_t *trace_parser::parse_kernel_info(const std::string &path) {
  ti_kernel_hdr h;
  ti_kernel *k = nullptr;
  ti_status st = ti_kernel_open_file(m->context(), path.c_str(), &h, &k);
  if (st == TI_ERR_IO) {
    std::cout << "Unable to open file: " << path << std::endl;
    exit(1);
  }
  kernel_trace_t *info = new kernel_trace_t;
  info->ifs = new std::ifstream;
  info->ifs->open(path.c_str());
  std::cout << "Processing kernel " << path << std::endl;
  // echo the header lines the way the reference does while it parses them
  std::string line;
  while (std::getline(*info->ifs, line)) {
    if (line.empty()) continue;
    if (line[0] == '#') break;
    if (line[0] == '-') std::cout << line << std::endl;
  }
  info->kernel_name = h.kernel_name;
  info->kernel_id = h.kernel_id;
  info->grid_dim_x = h.grid_dim[0]; info->grid_dim_y = h.grid_dim[1]; info->grid_dim_z = h.grid_dim[2];
  info->tb_dim_x = h.tb_dim[0]; info->tb_dim_y = h.tb_dim[1]; info->tb_dim_z = h.tb_dim[2];
  info->shmem = h.shmem; info->nregs = h.nregs; info->cuda_stream_id = h.cuda_stream_id;
  info->binary_verion = h.binary_version; info->trace_verion = h.trace_version;
  info->nvbit_verion = h.nvbit_version;
  info->shmem_base_addr = h.shmem_base_addr; info->local_base_addr = h.local_base_addr;
  if (st == TI_ERR_BINARY_VERSION) {
    // trace_kernel_info_t's constructor reports this and exits (trace_driven.cc:113-118);
    // keep the header usable until then
    return info;
  }
  if (st != TI_OK) die("parse_kernel_info", st);
  m->kernels[info->ifs].k = k;
  return info;
}

void trace_parser::kernel_finalizer(kernel_trace_t *trace_info) {
  assert(trace_info);
  assert(trace_info->ifs);
  auto it = m->kernels.find(trace_info->ifs);
  if (it != m->kernels.end()) { ti_kernel_close(it->second.k); m->kernels.erase(it); }
  if (trace_info->ifs->is_open()) trace_info->ifs->close();
  delete trace_info->ifs;
  delete trace_info;
}

void trace_parser::get_next_threadblock_traces(std::vector<std::vector<inst_trace_t> *> threadblock_traces,
                                               unsigned /*trace_version: the header's value is used*/, std::ifstream *ifs) {
  for (auto *v : threadblock_traces) v->clear();
  auto it = m->kernels.find(ifs);
  if (it == m->kernels.end()) return;
  KernelState &s = it->second;
  if (!s.planned) {
    ti_status st = ti_kernel_host_body(s.k, &s.body, &s.body_len);
    if (st != TI_OK) die("get_next_threadblock_traces", st);
    const int n_win = (int)std::max<size_t>(1, (s.body_len + window_bytes() - 1) / window_bytes());
    s.cuts.assign(n_win + 1, 0);
    st = ti_plan_shards(s.body, s.body_len, n_win, s.cuts.data());
    if (st != TI_OK) die("ti_plan_shards", st);
    s.planned = true;
    memset(&s.view, 0, sizeof s.view);
  }
  while (s.next_tb >= s.view.counts.n_tb) {   // the window is used up (or none decoded yet): next non-empty one
    if (s.next_window > 0) ti_kernel_release_text(s.k, s.cuts[s.next_window - 1], s.cuts[s.next_window]);   // the window just used up
    if (s.next_window + 1 >= s.cuts.size()) return;  // the reference's loop ends at EOF with nothing filled
    const size_t b = s.cuts[s.next_window], e = s.cuts[s.next_window + 1];
    ++s.next_window;
    if (e == b) continue;
    const auto t0 = std::chrono::steady_clock::now();
    ti_status st = ti_kernel_decode_window_compact(s.k, b, e, 0, &s.view);  // copies and kernel overlapped inside
    m->s_decode += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (st != TI_OK) die("get_next_threadblock_traces", st);
    s.next_tb = 0;
  }
  const ti_view_compact &v = s.view;
  const uint64_t k = s.next_tb++;
  const ti_tb_rec &tb = v.tb[k];
  std::cout << std::string(v.text + tb.text_off, tb.line_len) << std::endl;  // trace_parser.cc:427
  const uint64_t w_end = k + 1 < v.counts.n_tb ? v.tb[k + 1].first_warp : v.counts.n_warp;
  for (uint64_t w = tb.first_warp; w < w_end; ++w) {
    const ti_warp_rec &wr = v.warp[w];
    if (wr.warp_id >= threadblock_traces.size()) die("warp id beyond the block's warps", TI_ERR_PARSE);
    std::vector<inst_trace_t> &dst = *threadblock_traces[wr.warp_id];
    dst.resize(wr.n_insts);  // trace_parser.cc:435
    for (uint32_t i = 0; i < wr.n_insts; ++i) {
      const ti_inst_rec &r = v.inst[wr.first_inst + i];
      inst_trace_t &t = dst[i];
      t.m_pc = r.pc;
      t.mask = r.mask;
      t.reg_dsts_num = TI_NDST(&r);
      if (t.reg_dsts_num) t.reg_dest[0] = r.dst;
      t.opcode.assign(v.text + r.opc_off, r.opcode_len);
      t.reg_srcs_num = TI_NSRC(&r);
      for (unsigned j = 0; j < t.reg_srcs_num; ++j) t.reg_src[j] = r.src[j];
      if (TI_HAS_MEM(&r)) {
        t.memadd_info = new inst_memadd_info_t();
        ti_expand_row(&v, r.mem_idx, (uint64_t *)t.memadd_info->addrs);   // 16 B descriptor -> inst_memadd_info_t::addrs
        // inst_trace_t keeps the opcode-derived width (the LDC override belongs to the next stage)
        t.memadd_info->width = (int32_t)width_from_opcode(v.text + r.opc_off, r.opcode_len);
      }
    }
  }
}
