// TEST INFRASTRUCTURE -- not a product path.  Oracle of ti_compress_rows: the REFERENCE's own
// base_stride_compress / base_delta_compress, cut out of util/tracer_nvbit/tracer_tool/tracer_tool.cu where
// it lies (oracle/Makefile writes the two functions to oracle/_ref/compress_fns.inc; the file itself needs
// NVBit and cannot be compiled), driven by the format choice and the fprintf calls of tracer_tool.cu:717-750.
//
//   compress_ref <in.bin> <enable_compress>      in.bin: u64 n, then n x { u32 mask, u32 pad, u64 addrs[32] }
//   stdout: one line per instruction = the text between the memory width and the immediate
#include <bitset>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

bool base_stride_compress(const uint64_t *addrs, const std::bitset<32> &mask, uint64_t &base_addr, int &stride);
void base_delta_compress(const uint64_t *addrs, const std::bitset<32> &mask, uint64_t &base_addr, std::vector<long long> &deltas);
#include "_ref/compress_fns.inc"

enum address_format { list_all = 0, base_stride = 1, base_delta = 2 };   // tracer_tool.cu, same values as trace_parser.h

int main(int argc, char **argv) {
  if (argc < 3) return 2;
  FILE *f = fopen(argv[1], "rb");
  if (!f) return 3;
  const bool enable_compress = atoi(argv[2]) != 0;
  uint64_t n = 0;
  if (fread(&n, 8, 1, f) != 1) return 3;
  for (uint64_t i = 0; i < n; ++i) {
    struct { uint32_t mask, pad; uint64_t addrs[32]; } rec;
    if (fread(&rec, sizeof rec, 1, f) != 1) return 3;
    std::bitset<32> mask(rec.mask);
    // tracer_tool.cu:717-750
    bool base_stride_success = false;
    uint64_t base_addr = 0;
    int stride = 0;
    std::vector<long long> deltas;
    if (enable_compress) {
      base_stride_success = base_stride_compress(rec.addrs, mask, base_addr, stride);
      if (!base_stride_success) base_delta_compress(rec.addrs, mask, base_addr, deltas);
    }
    if (base_stride_success && enable_compress) {
      printf("%u 0x%llx %d ", base_stride, (unsigned long long)base_addr, stride);
    } else if (!base_stride_success && enable_compress) {
      printf("%u 0x%llx ", base_delta, (unsigned long long)base_addr);
      for (size_t s = 0; s < deltas.size(); s++) printf("%lld ", deltas[s]);
    } else {
      printf("%u ", list_all);
      for (int s = 0; s < 32; s++)
        if (mask.test(s)) printf("0x%016lx ", rec.addrs[s]);
    }
    printf("\n");
  }
  fclose(f);
  return 0;
}
