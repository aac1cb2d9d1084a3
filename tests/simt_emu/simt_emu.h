// TEST INFRASTRUCTURE -- a lock-step SIMT emulator for debugging device code on a
// machine without a GPU.  It compiles the *same* kernel source (csrc/*.cuh) with
// g++ by supplying the CUDA vocabulary the kernels use: threadIdx/blockIdx, dynamic
// shared memory, __syncthreads, the *_sync warp collectives, atomics, and a launcher
// that runs every thread of every CTA as a coroutine (ucontext) on one OS thread.
//
// It is NOT a product path and NOT a fallback: nothing in the package loads it, only
// tests/ build it (tests/simt_emu/build.py) to check kernel logic against the oracle
// before GPU time is spent.  Memory is sequentially consistent here, so it cannot
// prove the absence of races or missing fences -- the -m gpu tests and
// compute-sanitizer do that.
#ifndef TI_SIMT_EMU_H
#define TI_SIMT_EMU_H

#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ucontext.h>

#include <functional>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))

struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
struct uint3 { unsigned x, y, z; };
struct uint2 { unsigned x, y; };
struct __attribute__((aligned(16))) uint4 { unsigned x, y, z, w; };
struct __attribute__((aligned(16))) ulonglong2 { unsigned long long x, y; };
static inline uint4 make_uint4(unsigned a, unsigned b, unsigned c, unsigned d) { uint4 v = {a, b, c, d}; return v; }
static inline uint2 make_uint2(unsigned a, unsigned b) { uint2 v = {a, b}; return v; }
static inline ulonglong2 make_ulonglong2(unsigned long long a, unsigned long long b) { ulonglong2 v = {a, b}; return v; }

namespace emu {

enum Op { BALLOT, SHFL_IDX, SHFL_UP, SHFL_DOWN, SHFL_XOR, MATCH_ANY, RED_OR, RED_ADD, RED_MIN, RED_MAX, RED_AND, SYNCWARP };

struct Slot {           // one in-flight warp collective, keyed by its member mask
  uint32_t mask = 0, arrived = 0, remaining = 0;
  bool done = false;
  int op = -1;
  uint64_t in[32], arg[32], out[32];
};

struct Warp { Slot slots[40]; };

struct Cta {
  std::vector<unsigned char> smem;
  std::vector<Warp> warps;
  unsigned bar_count = 0, bar_gen = 0, live = 0;
  int and_acc = 1;
  // named barriers (bar.sync id, n / bar.red.and id, n): id 1..15, n = participating threads
  unsigned nbar_count[16] = {0}, nbar_gen[16] = {0};
  int nbar_and[16] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};
};

struct Thread {
  ucontext_t ctx;
  char *stack = nullptr;
  bool done = false;
  unsigned tid = 0, cta = 0;
};

struct Machine {
  std::vector<Cta> ctas;
  std::vector<Thread> threads;
  ucontext_t sched;
  Thread *cur = nullptr;
  std::function<void()> body;
  uint64_t progress = 0;
  dim3 grid, block;
};

Machine &M();
void yield_();
uint64_t collective(uint32_t mask, int op, uint64_t in, uint64_t arg);
void launch(dim3 grid, dim3 block, size_t smem_bytes, std::function<void()> body);
unsigned char *cur_smem();

}  // namespace emu

extern uint3 threadIdx, blockIdx;
extern dim3 blockDim, gridDim;

// ------------------------------------------------------------ synchronisation
void __syncthreads();
int __syncthreads_and(int pred);
namespace emu {
void named_bar_sync(unsigned id, unsigned n);
int named_bar_and(unsigned id, unsigned n, int pred);
}
static inline void __syncwarp(unsigned mask = 0xffffffffu) { emu::collective(mask, emu::SYNCWARP, 0, 0); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __nanosleep(unsigned) { emu::M().cur ? emu::yield_() : (void)0; }

// ------------------------------------------------------------ warp collectives
static inline unsigned __ballot_sync(unsigned mask, int pred) {
  return (unsigned)emu::collective(mask, emu::BALLOT, pred != 0, 0);
}
static inline int __any_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) != 0; }
static inline int __all_sync(unsigned mask, int pred) { return __ballot_sync(mask, pred) == mask; }
template <class T> static inline T emu_shfl(unsigned mask, int op, T v, unsigned arg) {
  static_assert(sizeof(T) <= 8, "shuffle payload");
  uint64_t raw = 0;
  memcpy(&raw, &v, sizeof(T));
  raw = emu::collective(mask, op, raw, arg);
  T r;
  memcpy(&r, &raw, sizeof(T));
  return r;
}
template <class T> static inline T __shfl_sync(unsigned m, T v, int src, int = 32) { return emu_shfl(m, emu::SHFL_IDX, v, (unsigned)src); }
template <class T> static inline T __shfl_up_sync(unsigned m, T v, unsigned d, int = 32) { return emu_shfl(m, emu::SHFL_UP, v, d); }
template <class T> static inline T __shfl_down_sync(unsigned m, T v, unsigned d, int = 32) { return emu_shfl(m, emu::SHFL_DOWN, v, d); }
template <class T> static inline T __shfl_xor_sync(unsigned m, T v, int x, int = 32) { return emu_shfl(m, emu::SHFL_XOR, v, (unsigned)x); }
template <class T> static inline unsigned __match_any_sync(unsigned m, T v) {
  uint64_t raw = 0;
  memcpy(&raw, &v, sizeof(T));
  return (unsigned)emu::collective(m, emu::MATCH_ANY, raw, 0);
}
static inline unsigned __reduce_or_sync(unsigned m, unsigned v) { return (unsigned)emu::collective(m, emu::RED_OR, v, 0); }
static inline unsigned __reduce_and_sync(unsigned m, unsigned v) { return (unsigned)emu::collective(m, emu::RED_AND, v, 0); }
static inline unsigned __reduce_add_sync(unsigned m, unsigned v) { return (unsigned)emu::collective(m, emu::RED_ADD, v, 0); }
static inline unsigned __reduce_min_sync(unsigned m, unsigned v) { return (unsigned)emu::collective(m, emu::RED_MIN, v, 0); }
static inline unsigned __reduce_max_sync(unsigned m, unsigned v) { return (unsigned)emu::collective(m, emu::RED_MAX, v, 0); }

// ------------------------------------------------------------------- integer
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline unsigned __brev(unsigned v) {
  unsigned r = 0;
  for (int i = 0; i < 32; ++i) r |= ((v >> i) & 1u) << (31 - i);
  return r;
}
template <class T> static inline T __ldg(const T *p) { return *p; }
template <class T> static inline T __ldcg(const T *p) { return *p; }

// ------------------------------------------------------------------- atomics
template <class T, class U> static inline T atomicAdd(T *p, U v) { T o = *p; *p = (T)(o + (T)v); emu::M().progress++; return o; }
template <class T, class U> static inline T atomicOr(T *p, U v) { T o = *p; *p = (T)(o | (T)v); emu::M().progress++; return o; }
template <class T, class U> static inline T atomicAnd(T *p, U v) { T o = *p; *p = (T)(o & (T)v); emu::M().progress++; return o; }
template <class T, class U> static inline T atomicMax(T *p, U v) { T o = *p; if ((T)v > o) *p = (T)v; emu::M().progress++; return o; }
template <class T, class U> static inline T atomicMin(T *p, U v) { T o = *p; if ((T)v < o) *p = (T)v; emu::M().progress++; return o; }
template <class T, class U> static inline T atomicExch(T *p, U v) { T o = *p; *p = (T)v; emu::M().progress++; return o; }
template <class T, class U> static inline T atomicCAS(T *p, U c, U v) { T o = *p; if (o == (T)c) *p = (T)v; emu::M().progress++; return o; }

#endif
