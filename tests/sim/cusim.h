// cusim.h -- TEST INFRASTRUCTURE ONLY.  A tiny single-OS-thread, fiber-based simulator of the CUDA execution
// model, just large enough to run this repo's device code (compressed_image_b200/csrc/*.cuh) on the CPU of the
// dev container, which has no GPU.  It exists so the warp-synchronous codec kernels can be functionally tested
// and debugged here before GPU minutes are spent; it is never part of the product and proves nothing about
// performance.  Every CUDA thread is a fiber with its own stack; warp collectives (__shfl_sync, __ballot_sync,
// __syncwarp ...) and __syncthreads are rendezvous points.  Lanes run maximally skewed between rendezvous
// points, which makes a missing __syncwarp() show up as a wrong answer instead of hiding behind lock-step luck.
#pragma once
#include <algorithm>
#include <cassert>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#define __device__
#define __global__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __noinline__ __attribute__((noinline))
#define __launch_bounds__(...)
#define __constant__ static
#define B2G_SIM 1

struct dim3 {
  unsigned x, y, z;
  dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
struct uint4 { uint32_t x, y, z, w; };
struct uint2 { uint32_t x, y; };
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{x, y, z, w}; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{x, y}; }

extern "C" void cusim_switch(void** from_sp, void* to_sp);
asm(R"(
.text
.globl cusim_switch
.type cusim_switch,@function
cusim_switch:
  pushq %rbp
  pushq %rbx
  pushq %r12
  pushq %r13
  pushq %r14
  pushq %r15
  movq %rsp, (%rdi)
  movq %rsi, %rsp
  popq %r15
  popq %r14
  popq %r13
  popq %r12
  popq %rbx
  popq %rbp
  ret
.size cusim_switch,.-cusim_switch
)");

namespace cusim {

enum Op : uint8_t { OP_NONE, OP_SYNCWARP, OP_SHFL, OP_BALLOT, OP_ANY, OP_ALL, OP_REDUCE, OP_MATCH };

struct Warp {
  uint64_t slot[2][32];
  uint8_t op[2][32];
  uint32_t gen = 0;
  int arrived = 0, nalive = 0;
};

struct Cta {
  dim3 bid;
  uint8_t* smem = nullptr;
  uint32_t gen = 0;
  int arrived = 0, nalive = 0;
  std::vector<Warp> warps;
};

struct Fiber {
  void* sp = nullptr;
  uint8_t* stack = nullptr;
  Cta* cta = nullptr;
  Warp* warp = nullptr;
  dim3 tid;
  int lane = 0;
  bool done = false, spin = false;
  volatile uint32_t* wait_gen = nullptr;
  uint32_t wait_val = 0;
};

struct State {
  void* main_sp = nullptr;
  Fiber* cur = nullptr;
  dim3 grid, block;
  const std::function<void()>* body = nullptr;
  uint64_t collectives = 0;
};
inline State& S() { static State s; return s; }

inline void yield_to_main() { cusim_switch(&S().cur->sp, S().main_sp); }

inline void block_on(volatile uint32_t* gen, uint32_t val) {
  Fiber* f = S().cur;
  f->wait_gen = gen;
  f->wait_val = val;
  while (*gen == val) yield_to_main();
  f->wait_gen = nullptr;
}

inline void fiber_exit_accounting(Fiber* f) {
  Warp* w = f->warp;
  Cta* c = f->cta;
  w->nalive--;
  if (w->nalive > 0 && w->arrived == w->nalive) { w->arrived = 0; w->gen++; }
  c->nalive--;
  if (c->nalive > 0 && c->arrived == c->nalive) { c->arrived = 0; c->gen++; }
}

static void fiber_main() {
  Fiber* f = S().cur;
  (*S().body)();
  f->done = true;
  fiber_exit_accounting(f);
  yield_to_main();
  abort();
}

// rendezvous of all live lanes of the calling warp; returns the parity whose slots hold this round's values
inline int warp_rendezvous(uint64_t v, Op op) {
  Fiber* f = S().cur;
  Warp* w = f->warp;
  S().collectives++;
  int p = (int)(w->gen & 1);
  uint32_t g = w->gen;
  w->slot[p][f->lane] = v;
  w->op[p][f->lane] = op;
  if (++w->arrived == w->nalive) {
    w->arrived = 0;
    w->gen = g + 1;
  } else {
    block_on(&w->gen, g);
  }
  if (w->op[p][0] != op && w->nalive == 32) {
    fprintf(stderr, "cusim: divergent warp collective (lane %d op %d vs lane0 op %d)\n", f->lane, op, w->op[p][0]);
    abort();
  }
  return p;
}

inline void cta_barrier() {
  Fiber* f = S().cur;
  Cta* c = f->cta;
  uint32_t g = c->gen;
  if (++c->arrived == c->nalive) {
    c->arrived = 0;
    c->gen = g + 1;
  } else {
    block_on(&c->gen, g);
  }
}

inline void spin_yield() {
  S().cur->spin = true;
  yield_to_main();
  S().cur->spin = false;
}

template <class F>
void launch(dim3 grid, dim3 block, size_t smem_bytes, F&& fn) {
  State& s = S();
  std::function<void()> body = fn;
  s.body = &body;
  s.grid = grid;
  s.block = block;
  const int nthreads = (int)(block.x * block.y * block.z);
  const int nwarps = (nthreads + 31) / 32;
  const int nctas = (int)(grid.x * grid.y * grid.z);
  const size_t kStack = 256 * 1024;
  std::vector<Cta> ctas(nctas);
  std::vector<Fiber> fibers((size_t)nctas * nthreads);
  std::vector<uint8_t*> smems;
  for (int c = 0; c < nctas; c++) {
    Cta& C = ctas[c];
    C.bid = dim3(c % grid.x, (c / grid.x) % grid.y, c / (grid.x * grid.y));
    C.smem = (uint8_t*)aligned_alloc(1024, ((smem_bytes + 1023) / 1024 + 1) * 1024);
    memset(C.smem, 0xCD, smem_bytes);
    C.warps.resize(nwarps);
    C.nalive = nthreads;
    for (int t = 0; t < nthreads; t++) {
      Fiber& f = fibers[(size_t)c * nthreads + t];
      f.cta = &C;
      f.warp = &C.warps[t / 32];
      f.warp->nalive++;
      f.lane = t % 32;
      f.tid = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
      f.stack = (uint8_t*)aligned_alloc(64, kStack);
      uintptr_t top = ((uintptr_t)f.stack + kStack) & ~(uintptr_t)15;
      void** sp = (void**)top;
      *--sp = nullptr;               // fake return address of fiber_main
      *--sp = (void*)&fiber_main;    // `ret` in cusim_switch jumps here
      for (int i = 0; i < 6; i++) *--sp = nullptr;
      f.sp = (void*)sp;
    }
  }
  size_t alive = fibers.size();
  while (alive > 0) {
    bool progress = false;
    for (int c = 0; c < nctas; c++) {
      for (int w = 0; w < nwarps; w++) {
        // keep running this warp's lanes while any of them can make progress without the rest of the CTA
        for (;;) {
          bool ran_nonspin = false, spun = false;
          int lo = w * 32, hi = std::min(nthreads, lo + 32);
          for (int t = lo; t < hi; t++) {
            Fiber& f = fibers[(size_t)c * nthreads + t];
            if (f.done) continue;
            if (f.wait_gen && *f.wait_gen == f.wait_val) continue;
            s.cur = &f;
            cusim_switch(&s.main_sp, f.sp);
            progress = true;
            if (f.done) alive--;
            else if (!f.spin) ran_nonspin = true;
            else spun = true;
          }
          // a lane that polls is waiting for another warp: give the rest of the grid a turn (a polling loop may hold
          // warp collectives, so "only spinning lanes ran" alone would never become true)
          if (!ran_nonspin || spun) break;
        }
      }
    }
    if (!progress) {
      fprintf(stderr, "cusim: deadlock (%zu fibers alive, none runnable)\n", alive);
      abort();
    }
  }
  for (auto& f : fibers) free(f.stack);
  for (auto& C : ctas) free(C.smem);
  s.body = nullptr;
  s.cur = nullptr;
}

struct TidProxy { operator dim3() const { return S().cur->tid; } };
}  // namespace cusim

#define threadIdx (cusim::S().cur->tid)
#define blockIdx (cusim::S().cur->cta->bid)
#define blockDim (cusim::S().block)
#define gridDim (cusim::S().grid)
#define warpSize 32

// ---- synchronisation + warp collectives ----------------------------------------------------------------------
static inline void __syncthreads() { cusim::cta_barrier(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { cusim::warp_rendezvous(0, cusim::OP_SYNCWARP); }
static inline unsigned __activemask() { return 0xffffffffu; }

template <class T>
static inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
  static_assert(sizeof(T) <= 8, "shfl type");
  uint64_t bits = 0;
  memcpy(&bits, &v, sizeof(T));
  int p = cusim::warp_rendezvous(bits, cusim::OP_SHFL);
  int lane = cusim::S().cur->lane;
  int base = lane & ~(width - 1);
  int s = base + (src & (width - 1));
  T out;
  memcpy(&out, &cusim::S().cur->warp->slot[p][s], sizeof(T));
  return out;
}
template <class T>
static inline T __shfl_up_sync(unsigned, T v, unsigned delta, int width = 32) {
  uint64_t bits = 0;
  memcpy(&bits, &v, sizeof(T));
  int p = cusim::warp_rendezvous(bits, cusim::OP_SHFL);
  int lane = cusim::S().cur->lane;
  int s = lane - (int)delta;
  if (s < (lane & ~(width - 1))) s = lane;
  T out;
  memcpy(&out, &cusim::S().cur->warp->slot[p][s], sizeof(T));
  return out;
}
template <class T>
static inline T __shfl_down_sync(unsigned, T v, unsigned delta, int width = 32) {
  uint64_t bits = 0;
  memcpy(&bits, &v, sizeof(T));
  int p = cusim::warp_rendezvous(bits, cusim::OP_SHFL);
  int lane = cusim::S().cur->lane;
  int s = lane + (int)delta;
  if (s > (lane | (width - 1))) s = lane;
  T out;
  memcpy(&out, &cusim::S().cur->warp->slot[p][s], sizeof(T));
  return out;
}
template <class T>
static inline T __shfl_xor_sync(unsigned, T v, int m, int width = 32) {
  uint64_t bits = 0;
  memcpy(&bits, &v, sizeof(T));
  int p = cusim::warp_rendezvous(bits, cusim::OP_SHFL);
  int lane = cusim::S().cur->lane;
  int s = lane ^ m;
  if (s > (lane | (width - 1))) s = lane;
  T out;
  memcpy(&out, &cusim::S().cur->warp->slot[p][s], sizeof(T));
  return out;
}
static inline unsigned __ballot_sync(unsigned, int pred) {
  int p = cusim::warp_rendezvous(pred ? 1 : 0, cusim::OP_BALLOT);
  unsigned r = 0;
  cusim::Warp* w = cusim::S().cur->warp;
  for (int i = 0; i < 32; i++)
    if (w->slot[p][i] & 1) r |= 1u << i;
  return r;
}
// __match_any_sync(full mask, v): the lanes holding the same value
static inline unsigned cusim_match_any(unsigned v) {
  int p = cusim::warp_rendezvous(v, cusim::OP_MATCH);
  unsigned r = 0;
  cusim::Warp* w = cusim::S().cur->warp;
  for (int i = 0; i < 32; i++)
    if ((unsigned)w->slot[p][i] == v) r |= 1u << i;
  return r;
}
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, pred) == 0xffffffffu; }
static inline unsigned __reduce_add_sync(unsigned, unsigned v) {
  int p = cusim::warp_rendezvous(v, cusim::OP_REDUCE);
  unsigned r = 0;
  for (int i = 0; i < 32; i++) r += (unsigned)cusim::S().cur->warp->slot[p][i];
  return r;
}
static inline unsigned __reduce_max_sync(unsigned, unsigned v) {
  int p = cusim::warp_rendezvous(v, cusim::OP_REDUCE);
  unsigned r = 0;
  for (int i = 0; i < 32; i++) r = std::max(r, (unsigned)cusim::S().cur->warp->slot[p][i]);
  return r;
}
static inline unsigned __reduce_min_sync(unsigned, unsigned v) {
  int p = cusim::warp_rendezvous(v, cusim::OP_REDUCE);
  unsigned r = 0xffffffffu;
  for (int i = 0; i < 32; i++) r = std::min(r, (unsigned)cusim::S().cur->warp->slot[p][i]);
  return r;
}
static inline unsigned __reduce_or_sync(unsigned, unsigned v) {
  int p = cusim::warp_rendezvous(v, cusim::OP_REDUCE);
  unsigned r = 0;
  for (int i = 0; i < 32; i++) r |= (unsigned)cusim::S().cur->warp->slot[p][i];
  return r;
}

// ---- scalar intrinsics -------------------------------------------------------------------------------------
static inline int __popc(unsigned v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline int __clzll(long long v) { return v ? __builtin_clzll((unsigned long long)v) : 64; }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline int __ffsll(long long v) { return __builtin_ffsll(v); }
static inline unsigned __brev(unsigned v) {
  unsigned r = 0;
  for (int i = 0; i < 32; i++) r |= ((v >> i) & 1u) << (31 - i);
  return r;
}
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned s) {
  uint64_t v = ((uint64_t)b << 32) | a;
  unsigned r = 0;
  for (int i = 0; i < 4; i++) {
    unsigned sel = (s >> (4 * i)) & 0xF;
    unsigned byte = (unsigned)((v >> (8 * (sel & 7))) & 0xFF);
    if (sel & 8) byte = (byte & 0x80) ? 0xFF : 0x00;
    r |= byte << (8 * i);
  }
  return r;
}
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) {
  uint64_t v = ((uint64_t)hi << 32) | lo;
  return (unsigned)(v >> (sh & 31));
}
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned sh) {
  uint64_t v = ((uint64_t)hi << 32) | lo;
  return (unsigned)((v << (sh & 31)) >> 32);
}
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((uint64_t)a * b) >> 32); }
template <class T> static inline T __ldg(const T* p) { return *p; }
template <class T> static inline T __ldcs(const T* p) { return *p; }
template <class T> static inline void __stcs(T* p, T v) { *p = v; }
static inline void __nanosleep(unsigned) { cusim::spin_yield(); }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __trap() { fprintf(stderr, "cusim: __trap()\n"); abort(); }
using std::max;
using std::min;

// atomics: one OS thread, fibers switch only at rendezvous points -> plain read-modify-write is atomic here
template <class T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
template <class T> static inline T atomicSub(T* p, T v) { T o = *p; *p = o - v; return o; }
template <class T> static inline T atomicExch(T* p, T v) { T o = *p; *p = v; return o; }
template <class T> static inline T atomicMin(T* p, T v) { T o = *p; *p = std::min(o, v); return o; }
template <class T> static inline T atomicMax(T* p, T v) { T o = *p; *p = std::max(o, v); return o; }
template <class T> static inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }
template <class T> static inline T atomicAnd(T* p, T v) { T o = *p; *p = o & v; return o; }
template <class T> static inline T atomicCAS(T* p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }
