// TEST INFRASTRUCTURE -- see simt_emu.h.
#include "simt_emu.h"

#include <sys/mman.h>

uint3 threadIdx, blockIdx;
dim3 blockDim, gridDim;

namespace emu {

static Machine g_m;
Machine &M() { return g_m; }

static const size_t kStack = 512 * 1024;

static void enter(Thread *t) {
  Machine &m = g_m;
  m.cur = t;
  threadIdx.x = t->tid; threadIdx.y = threadIdx.z = 0;
  blockIdx.x = t->cta; blockIdx.y = blockIdx.z = 0;
}

void yield_() {
  Machine &m = g_m;
  Thread *t = m.cur;
  swapcontext(&t->ctx, &m.sched);
  enter(t);
}

unsigned char *cur_smem() { return g_m.ctas[g_m.cur->cta].smem.data(); }

static void finish(Slot &s) {
  const uint32_t mask = s.mask;
  uint64_t red = 0;
  bool first = true;
  for (int l = 0; l < 32; ++l) {
    if (!(mask >> l & 1)) continue;
    uint32_t v = (uint32_t)s.in[l];
    switch (s.op) {
      case BALLOT: red |= (uint64_t)(s.in[l] & 1) << l; break;
      case RED_OR: red |= v; break;
      case RED_AND: red = first ? v : (red & v); break;
      case RED_ADD: red = (uint32_t)(red + v); break;
      case RED_MIN: red = first ? v : (v < red ? v : red); break;
      case RED_MAX: red = first ? v : (v > red ? v : red); break;
      default: break;
    }
    first = false;
  }
  for (int l = 0; l < 32; ++l) {
    if (!(mask >> l & 1)) continue;
    int src = l;
    switch (s.op) {
      case SHFL_IDX: src = (int)(s.arg[l] & 31); break;
      case SHFL_UP: src = l - (int)s.arg[l]; break;
      case SHFL_DOWN: src = l + (int)s.arg[l]; break;
      case SHFL_XOR: src = l ^ (int)(s.arg[l] & 31); break;
      default: break;
    }
    switch (s.op) {
      case SHFL_IDX: case SHFL_UP: case SHFL_DOWN: case SHFL_XOR:
        if (src < 0 || src > 31) src = l;
        if (!(mask >> src & 1)) {
          // reading a lane outside the member mask is undefined in CUDA: poison it
          s.out[l] = 0xDEADBEEFDEADBEEFull;
        } else s.out[l] = s.in[src];
        break;
      case MATCH_ANY: {
        uint32_t g = 0;
        for (int k = 0; k < 32; ++k)
          if ((mask >> k & 1) && s.in[k] == s.in[l]) g |= 1u << k;
        s.out[l] = g;
        break;
      }
      case SYNCWARP: s.out[l] = 0; break;
      default: s.out[l] = red; break;
    }
  }
}

uint64_t collective(uint32_t mask, int op, uint64_t in, uint64_t arg) {
  Machine &m = g_m;
  Thread *t = m.cur;
  const unsigned lane = t->tid & 31, wid = t->tid >> 5;
  if (!(mask >> lane & 1)) {
    fprintf(stderr, "emu: lane %u calls a collective with mask %08x that excludes it\n", lane, mask);
    abort();
  }
  Warp &w = m.ctas[t->cta].warps[wid];
  for (;;) {
    Slot *s = nullptr, *freeslot = nullptr;
    for (Slot &x : w.slots) {
      if (x.mask == mask && x.op != -1) { s = &x; break; }
      if (x.op == -1 && !freeslot) freeslot = &x;
    }
    if (s && (s->done || (s->arrived >> lane & 1))) { yield_(); continue; }  // previous round still draining
    if (s && s->op != op) {
      fprintf(stderr, "emu: divergent collectives in warp %u (mask %08x): op %d vs %d\n", wid, mask, s->op, op);
      abort();
    }
    if (!s) {
      if (!freeslot) { fprintf(stderr, "emu: out of collective slots\n"); abort(); }
      s = freeslot;
      s->mask = mask; s->op = op; s->arrived = 0; s->done = false; s->remaining = 0;
    }
    s->in[lane] = in; s->arg[lane] = arg; s->arrived |= 1u << lane;
    while (!s->done && s->arrived != mask) yield_();
    if (!s->done) { finish(*s); s->done = true; s->remaining = mask; m.progress++; }
    uint64_t r = s->out[lane];
    s->remaining &= ~(1u << lane);
    if (!s->remaining) { s->op = -1; s->done = false; s->arrived = 0; s->mask = 0; }
    return r;
  }
}

static void trampoline() {
  Machine &m = g_m;
  Thread *t = m.cur;
  m.body();
  t = m.cur;
  t->done = true;
  m.ctas[t->cta].live--;
  m.progress++;
  swapcontext(&t->ctx, &m.sched);
}

void launch(dim3 grid, dim3 block, size_t smem_bytes, std::function<void()> body) {
  Machine &m = g_m;
  const unsigned nt = block.x, nc = grid.x;
  if (nt % 32) { fprintf(stderr, "emu: block size must be a multiple of 32\n"); abort(); }
  m.grid = grid; m.block = block; m.body = body;
  gridDim = grid; blockDim = block;
  m.ctas.assign(nc, Cta());
  for (Cta &c : m.ctas) { c.smem.assign(smem_bytes + 64, 0xCD); c.warps.assign(nt / 32, Warp()); c.live = nt; }
  m.threads.assign((size_t)nc * nt, Thread());
  for (unsigned c = 0; c < nc; ++c)
    for (unsigned i = 0; i < nt; ++i) {
      Thread &t = m.threads[(size_t)c * nt + i];
      t.tid = i; t.cta = c;
      t.stack = (char *)mmap(NULL, kStack, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
      getcontext(&t.ctx);
      t.ctx.uc_stack.ss_sp = t.stack;
      t.ctx.uc_stack.ss_size = kStack;
      t.ctx.uc_link = &m.sched;
      makecontext(&t.ctx, trampoline, 0);
    }
  size_t live = m.threads.size();
  uint64_t last = m.progress;
  int idle = 0;
  while (live) {
    live = 0;
    for (Thread &t : m.threads) {
      if (t.done) continue;
      enter(&t);
      swapcontext(&m.sched, &t.ctx);
      if (!t.done) ++live;
    }
    if (m.progress == last) {
      if (++idle > 20000) { fprintf(stderr, "emu: deadlock -- no thread made progress\n"); abort(); }
    } else { idle = 0; last = m.progress; }
  }
  for (Thread &t : m.threads) munmap(t.stack, kStack);
  m.threads.clear();
  m.cur = nullptr;
}

}  // namespace emu

void __syncthreads() {
  emu::Machine &m = emu::M();
  emu::Cta &c = m.ctas[m.cur->cta];
  const unsigned gen = c.bar_gen;
  if (++c.bar_count == m.block.x) {
    if (c.live != m.block.x) { fprintf(stderr, "emu: __syncthreads after some threads exited\n"); abort(); }
    c.bar_count = 0;
    c.bar_gen++;
    m.progress++;
    return;
  }
  while (c.bar_gen == gen) emu::yield_();
}

// barrier + AND-reduction of a predicate over the CTA (three plain barriers: vote, read, reset)
int __syncthreads_and(int pred) {
  emu::Machine &m = emu::M();
  emu::Cta &c = m.ctas[m.cur->cta];
  if (!pred) c.and_acc = 0;
  __syncthreads();
  const int r = c.and_acc;
  __syncthreads();
  if (m.cur->tid == 0) c.and_acc = 1;
  __syncthreads();
  return r;
}

// Named barriers: the first `n` arrivals at barrier `id` release each other.  The AND variant
// is three rounds (vote, read, reset) like __syncthreads_and; the last thread to leave the
// read round resets the accumulator.
namespace emu {
void named_bar_sync(unsigned id, unsigned n) {
  if (id == 0) { __syncthreads(); return; }
  Machine &m = M();
  Cta &c = m.ctas[m.cur->cta];
  const unsigned gen = c.nbar_gen[id];
  if (++c.nbar_count[id] == n) {
    c.nbar_count[id] = 0;
    c.nbar_gen[id]++;
    m.progress++;
    return;
  }
  while (c.nbar_gen[id] == gen) yield_();
}
int named_bar_and(unsigned id, unsigned n, int pred) {
  if (id == 0) return __syncthreads_and(pred);
  Machine &m = M();
  Cta &c = m.ctas[m.cur->cta];
  if (!pred) c.nbar_and[id] = 0;
  named_bar_sync(id, n);
  const int r = c.nbar_and[id];
  named_bar_sync(id, n);
  c.nbar_and[id] = 1;       // every participant writes the same value; all have read
  named_bar_sync(id, n);
  return r;
}
}  // namespace emu
