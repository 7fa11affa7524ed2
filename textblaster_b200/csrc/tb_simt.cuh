// tb_simt.cuh -- the handful of warp-level primitives the kernel bodies use.
//
// Device build: thin wrappers over the CUDA intrinsics (full-warp masks only).
// TB_HOST_EMUL build (tests/host_emul, test infrastructure only, never linked into the shipped
// library): the lanes of one warp (or the threads of one CTA) are fibers of one host thread, run round-robin;
// every collective is a barrier (= a switch to the next lane) plus an exchange through a small mailbox,
// atomics are GCC builtins.  This lets the exact kernel bodies run against the oracle on a machine without a GPU.
#pragma once
#include <stdint.h>

#include "tb_common.cuh"

#if !defined(TB_HOST_EMUL)

#define SIMT_FULL 0xFFFFFFFFu
TB_D uint32_t simt_shfl(uint32_t v, int src) { return __shfl_sync(SIMT_FULL, v, src); }
TB_D uint64_t simt_shfl(uint64_t v, int src) { return __shfl_sync(SIMT_FULL, v, src); }
TB_D uint32_t simt_shfl_up(uint32_t v, int o) { return __shfl_up_sync(SIMT_FULL, v, o); }
TB_D uint64_t simt_shfl_xor(uint64_t v, int o) { return __shfl_xor_sync(SIMT_FULL, v, o); }
TB_D uint32_t simt_sum(uint32_t v) { return __reduce_add_sync(SIMT_FULL, v); }
TB_D uint32_t simt_or(uint32_t v) { return __reduce_or_sync(SIMT_FULL, v); }
TB_D bool simt_any(bool p) { return __any_sync(SIMT_FULL, p) != 0; }
TB_D uint32_t simt_ballot(bool p) { return __ballot_sync(SIMT_FULL, p); }
TB_D uint32_t simt_match_any64(uint64_t v) { return __match_any_sync(SIMT_FULL, (unsigned long long)v); }  // lanes holding the same value
TB_D uint32_t simt_max(uint32_t v) { return __reduce_max_sync(SIMT_FULL, v); }
TB_D void simt_sync() { __syncwarp(); }
TB_D uint32_t simt_cas(uint32_t* a, uint32_t cmp, uint32_t val) { return atomicCAS(a, cmp, val); }
TB_D uint32_t simt_atomic_add(uint32_t* a, uint32_t v) { return atomicAdd(a, v); }
TB_D uint32_t simt_atomic_min(uint32_t* a, uint32_t v) { return atomicMin(a, v); }
TB_D uint32_t simt_atomic_max(uint32_t* a, uint32_t v) { return atomicMax(a, v); }
TB_D uint32_t simt_atomic_or(uint32_t* a, uint32_t v) { return atomicOr(a, v); }
// a word other lanes of this warp have updated with simt_atomic_or since the last simt_sync
TB_D uint32_t simt_load_fresh(const uint32_t* a) { return __ldcg(a); }
TB_D int simt_popc(uint32_t v) { return __popc(v); }
TB_D int simt_ffs(uint32_t v) { return __ffs(v); }

#else  // ---------------------------------------------------------------- host emulation

#include <cstring>

// Defined by the harness: suspends the calling fiber and runs the next one of the group.
void simt_fiber_yield();
// A counting barrier over n fibers of one host thread: the last to arrive opens it, the others hand over until then.
// (Groups of different size -- the warps of a CTA and the CTA itself -- may synchronise independently.)
struct SimtBar {
  int n, count = 0, gen = 0;
  explicit SimtBar(int n_) : n(n_) {}
  void arrive_and_wait() {
    const int g = gen;
    if (++count == n) { count = 0; gen++; }
    else while (gen == g) simt_fiber_yield();
  }
};
struct SimtWarp {
  SimtBar bar{32};
  uint64_t box[32];
};
extern thread_local SimtWarp* simt_warp;
extern thread_local int simt_lane;
struct SimtBlock {  // one emulated CTA of n threads
  int n;
  SimtBar bar;
  uint64_t box[1024];
  explicit SimtBlock(int n_) : n(n_), bar(n_) {}
};
extern thread_local SimtBlock* simt_block;

inline uint64_t simt_exchange_(uint64_t v, int src_lane, bool valid) {
  SimtWarp* w = simt_warp;
  w->box[simt_lane] = v;
  w->bar.arrive_and_wait();
  uint64_t r = valid ? w->box[src_lane] : v;
  w->bar.arrive_and_wait();
  return r;
}
inline uint32_t simt_shfl(uint32_t v, int src) { return (uint32_t)simt_exchange_(v, src & 31, true); }
inline uint64_t simt_shfl(uint64_t v, int src) { return simt_exchange_(v, src & 31, true); }
inline uint32_t simt_shfl_up(uint32_t v, int o) { return (uint32_t)simt_exchange_(v, simt_lane - o, simt_lane - o >= 0); }
inline uint64_t simt_shfl_xor(uint64_t v, int o) { return simt_exchange_(v, simt_lane ^ o, true); }
inline uint32_t simt_sum(uint32_t v) {
  SimtWarp* w = simt_warp;
  w->box[simt_lane] = v;
  w->bar.arrive_and_wait();
  uint32_t s = 0;
  for (int i = 0; i < 32; i++) s += (uint32_t)w->box[i];
  w->bar.arrive_and_wait();
  return s;
}
inline uint32_t simt_or(uint32_t v) {
  SimtWarp* w = simt_warp;
  w->box[simt_lane] = v;
  w->bar.arrive_and_wait();
  uint32_t s = 0;
  for (int i = 0; i < 32; i++) s |= (uint32_t)w->box[i];
  w->bar.arrive_and_wait();
  return s;
}
inline bool simt_any(bool p) { return simt_or(p ? 1u : 0u) != 0; }
inline uint32_t simt_match_any64(uint64_t v) {
  SimtWarp* w = simt_warp;
  w->box[simt_lane] = v;
  w->bar.arrive_and_wait();
  uint32_t m = 0;
  for (int i = 0; i < 32; i++) if (w->box[i] == v) m |= 1u << i;
  w->bar.arrive_and_wait();
  return m;
}
inline uint32_t simt_max(uint32_t v) {
  SimtWarp* w = simt_warp;
  w->box[simt_lane] = v;
  w->bar.arrive_and_wait();
  uint32_t s = 0;
  for (int i = 0; i < 32; i++) s = (uint32_t)w->box[i] > s ? (uint32_t)w->box[i] : s;
  w->bar.arrive_and_wait();
  return s;
}
inline uint32_t simt_ballot(bool p) { return simt_or(p ? (1u << simt_lane) : 0u); }
inline void simt_sync() { simt_warp->bar.arrive_and_wait(); }
inline uint32_t simt_cas(uint32_t* a, uint32_t cmp, uint32_t val) {
  __atomic_compare_exchange_n(a, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
  return cmp;
}
inline uint32_t simt_atomic_add(uint32_t* a, uint32_t v) { return __atomic_fetch_add(a, v, __ATOMIC_SEQ_CST); }
inline uint32_t simt_atomic_or(uint32_t* a, uint32_t v) { return __atomic_fetch_or(a, v, __ATOMIC_SEQ_CST); }
inline uint32_t simt_atomic_min(uint32_t* a, uint32_t v) {
  uint32_t old = __atomic_load_n(a, __ATOMIC_SEQ_CST);
  while (v < old && !__atomic_compare_exchange_n(a, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}
inline uint32_t simt_atomic_max(uint32_t* a, uint32_t v) {
  uint32_t old = __atomic_load_n(a, __ATOMIC_SEQ_CST);
  while (v > old && !__atomic_compare_exchange_n(a, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return old;
}
inline uint32_t simt_load_fresh(const uint32_t* a) { return __atomic_load_n(a, __ATOMIC_SEQ_CST); }
inline int simt_popc(uint32_t v) { return __builtin_popcount(v); }
inline int simt_ffs(uint32_t v) { return __builtin_ffs((int)v); }

#endif

// ---- composite helpers (same code on both builds) --------------------------------------------
TB_D uint64_t simt_sum64(uint64_t v) {
  for (int o = 16; o; o >>= 1) v += simt_shfl_xor(v, o);
  return v;
}
TB_D uint64_t simt_max64(uint64_t v) {  // two 32-bit reductions: the largest high word, then the largest low word among its holders
  const uint32_t hi = simt_max((uint32_t)(v >> 32));
  const uint32_t lo = simt_max((uint32_t)(v >> 32) == hi ? (uint32_t)v : 0u);
  return ((uint64_t)hi << 32) | lo;
}
TB_D int64_t simt_shfl_up64(int64_t v, int o) {
  uint32_t lo = simt_shfl_up((uint32_t)(uint64_t)v, o), hi = simt_shfl_up((uint32_t)((uint64_t)v >> 32), o);
  return (int64_t)(((uint64_t)hi << 32) | lo);
}
// exclusive prefix sum across the warp; *total gets the warp sum
TB_D uint32_t simt_exscan(uint32_t v, int lane, uint32_t* total) {
  uint32_t x = v;
  for (int o = 1; o < 32; o <<= 1) { uint32_t y = simt_shfl_up(x, o); if (lane >= o) x += y; }
  *total = simt_shfl(x, 31);
  return x - v;
}
