// Drop-in for the reference's post-processing binary
// (util/tracer_nvbit/tracer_tool/traces-processing/post-traces-processing.cpp): same command
// line (path of a kernelslist), same files written (kernel-N.traceg beside every
// kernel-N.trace, kernelslist.g), same stdout.  The regrouping itself runs on the GPU through
// the C ABI (ti_regroup_buffer); this file only does what the reference's main() and its
// printing do.  Build: g++ -std=c++17 -I include shim/post_traces_processing_b200.cc
//                      -L<csrc> -l:libtraceingest.so
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "traceingest.h"

using namespace std;

static ti_ctx *g_ctx = nullptr;

// group_per_block (PTP:85-194)
static void group_per_block(const char *filepath) {
  FILE *f = fopen(filepath, "rb");
  if (!f) {  // PTP:92-95
    cout << "Unable to open file: " << filepath << endl;
    return;
  }
  cout << "Processing file " << filepath << endl;  // PTP:97
  vector<char> raw;
  fseek(f, 0, SEEK_END);
  const long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  raw.resize(n > 0 ? (size_t)n : 0);
  if (n > 0 && fread(raw.data(), 1, (size_t)n, f) != (size_t)n) { fclose(f); cout << "Unable to open file: " << filepath << endl; return; }
  fclose(f);

  const void *out = nullptr;
  size_t out_len = 0;
  ti_regroup_stats st;
  if (ti_regroup_buffer(g_ctx, raw.data(), raw.size(), &out, &out_len, &st) != TI_OK) {
    // where the reference indexes past a vector (undefined behaviour, usually a crash)
    fprintf(stderr, "post-traces-processing: %s: %s\n", filepath, ti_last_error());
    abort();
  }
  FILE *g = fopen((string(filepath) + "g").c_str(), "wb");
  if (g) { fwrite(out, 1, out_len, g); fclose(g); }

  // the warnings of PTP:159-181, in the reference's order
  const uint32_t *cnt, *first;
  uint64_t n_tb;
  uint32_t warps;
  if (ti_regroup_buckets(g_ctx, &cnt, &first, &n_tb, &warps) != TI_OK) return;
  for (uint64_t i = 0; i < n_tb; ++i) {
    if (first[i] == 0xFFFFFFFFu || warps == 0) {
      cout << "Warning: Thread block 0,0,0 is empty" << endl;  // the coordinates of a block no line mentions were never set
      continue;
    }
    unsigned x = 0, y = 0, z = 0;
    sscanf(raw.data() + first[i], "%u %u %u", &x, &y, &z);
    for (uint32_t j = 0; j < warps; ++j)
      if (cnt[1 + i * warps + j] == 0) cout << "Warning: Warp " << j << " in thread block" << x << "," << y << "," << z << " is empty" << endl;
  }
}

int main(int argc, char **argv) {
  string kernellist_filepath;
  if (argc == 1) {
    cout << "File path is missing\n";
    return 0;
  } else if (argc == 2 || argc == 3) {
    kernellist_filepath = argv[1];
  } else {
    cout << "Too Many Arguemnts!\n";
    return 0;
  }
  ifstream ifs(kernellist_filepath.c_str());
  ofstream ofs((kernellist_filepath + ".g").c_str());
  if (!ifs.is_open()) {
    cout << "Unable to open file: " << kernellist_filepath << endl;
    return 0;
  }
  ti_config cfg;
  ti_config_qv100(&cfg);
  if (ti_ctx_create(&cfg, &g_ctx) != TI_OK) { fprintf(stderr, "post-traces-processing: %s\n", ti_last_error()); return 1; }

  string directory(kernellist_filepath);
  const size_t last_slash_idx = directory.rfind('/');
  if (string::npos != last_slash_idx) directory = directory.substr(0, last_slash_idx);

  // PTP:62-80, line by line so that kernelslist.g is cut where the reference cuts it on an unknown command
  string line;
  while (!ifs.eof()) {
    getline(ifs, line);
    if (line.empty()) continue;
    char one[8];
    size_t need = 0;
    const string with_nl = line + "\n";
    const ti_status e = ti_regroup_commandlist(with_nl.data(), with_nl.size(), one, 0, &need);
    if (e == TI_ERR_PARSE) {
      cout << "Undefined command: " << line << endl;
      return 0;
    }
    if (line.substr(0, 6) == "kernel") group_per_block((directory + "/" + line).c_str());
    vector<char> buf(need + 1);
    ti_regroup_commandlist(with_nl.data(), with_nl.size(), buf.data(), buf.size(), &need);
    ofs.write(buf.data(), (streamsize)need);
    ofs.flush();
  }
  ti_ctx_destroy(g_ctx);
  return 0;
}
