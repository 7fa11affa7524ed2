// tb_kernels.cu -- __global__ wrappers and launchers for the warp bodies in tb_bodies.cuh.
// One warp per document, 4 warps per CTA; documents are independent, so the hardware block
// scheduler does the load balancing (grid = ceil(n_docs / 4) CTAs, many waves over the 148 SMs).
#include <cuda_runtime.h>
#include <stdint.h>

#include "tb_bodies.cuh"

namespace {

// ASCII row of the cinfo table + the fast path's per-byte class flags (tb_fastwords.cuh) into shared memory
__device__ __forceinline__ void load_ascii(uint16_t* s_ascii, const UcTables& uc) {
  uint8_t* af = (uint8_t*)(s_ascii + 128);
  for (uint32_t i = threadIdx.x; i < 128; i += blockDim.x) {
    uint16_t ci = uc.stage2[(uint32_t)uc.stage1[0] * 256u + i];
    s_ascii[i] = ci;
    af[i] = (uint8_t)tb_ascii_flags(ci);
  }
  __syncthreads();
}
// per-byte class table of the mask-based word passes (tb_tile.cuh), built once per CTA
__device__ __forceinline__ void load_lut(LutEntry* s_lut, const UcTables& uc, int kind, uint64_t aux) {
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
    uint32_t ci = i < 128 ? uc.stage2[(uint32_t)uc.stage1[0] * 256u + i] : 0u;
    s_lut[i] = tb_lut_entry(kind, i, ci, (kind == TB_LUT_C4A || kind == TB_LUT_FW) ? (uint64_t)((uc.wslead[i >> 5] >> (i & 31)) & 1u) : aux);
  }
  __syncthreads();
}
// CTAs (of one warp) per SM each persistent text pass is built and launched for: the register budget follows from it
#ifndef TB_RES_GQ
#define TB_RES_GQ 12
#endif
#ifndef TB_RES_GR_SCAN
#define TB_RES_GR_SCAN 16
#endif
#ifndef TB_RES_C4_A
#define TB_RES_C4_A 24
#endif
#ifndef TB_RES_C4_B
#define TB_RES_C4_B 20
#endif
#ifndef TB_RES_C4_C
#define TB_RES_C4_C 32
#endif
#ifndef TB_RES_C4_D
#define TB_RES_C4_D 24
#endif
// persistent one-warp CTAs pull work items from a cursor (zeroed by the host before the launch): first the
// listed long documents, then all documents in index order (listed ones are skipped there).
// Returns -1 when the work is exhausted, -2 for an item to skip.
__device__ __forceinline__ int64_t next_doc(uint32_t* cursor, const KBatch& a) {
  uint32_t i = 0;
  if ((threadIdx.x & 31) == 0) i = atomicAdd(cursor, 1u);
  i = __shfl_sync(0xFFFFFFFFu, i, 0);
  const uint32_t listed = a.order_counts[0];
  if (i < listed) return (int64_t)a.order[i];
  const int64_t d = (int64_t)i - listed;
  if (d >= a.n) return -1;
  return a.sizecls[d] ? -2 : d;
}
// Size class of a document: 0 below 8 KiB (not listed), then one class per doubling up to 7 (512 KiB and more).
__device__ __forceinline__ uint32_t size_class(int64_t len) {
  if (len <= TB_LONG_DOC_BYTES) return 0u;
  const uint32_t c = (uint32_t)(63 - __clzll((unsigned long long)(len - 1))) - 12u;  // (8K, 16K] -> 1, (16K, 32K] -> 2, ...
  return c > 7u ? 7u : c;
}
// Work order, longest documents first (one class per doubling of the length: the last document to start is then at
// most twice as long as any that could have been started instead).  Pass 0 counts the classes (counts[8 + c]),
// pass 1 places every listed document behind the longer classes; counts[0] = number of listed documents.
__global__ void k_order(const int64_t* rs, const int64_t* re, int64_t n, uint32_t* order, uint32_t* counts, uint8_t* sizecls, int pass) {
  int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n) return;
  const uint32_t c = size_class(re[d] - rs[d]);
  if (pass == 0) {
    sizecls[d] = (uint8_t)c;
    if (c) atomicAdd(&counts[8 + c], 1u);
    return;
  }
  if (!c) { if (d == 0) { uint32_t t = 0; for (uint32_t k = 1; k < 8; k++) t += counts[8 + k]; counts[0] = t; } return; }
  uint32_t base = 0;
  for (uint32_t k = c + 1; k < 8; k++) base += counts[8 + k];
  order[base + atomicAdd(&counts[c], 1u)] = (uint32_t)d;
  if (d == 0) { uint32_t t = 0; for (uint32_t k = 1; k < 8; k++) t += counts[8 + k]; counts[0] = t; }
}
__device__ __forceinline__ int64_t warp_doc() { return ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; }
__device__ __forceinline__ int64_t thread_doc() { return (int64_t)blockIdx.x * blockDim.x + threadIdx.x; }

__global__ void k_init_ranges(const void* off, int off64, int64_t doc0, int64_t n, int64_t base, int64_t* rs,
                              int64_t* re, uint8_t* alive, int8_t* failed) {
  int64_t d = thread_doc();
  if (d < n) body_init_ranges(off, off64, doc0, base, rs, re, alive, failed, d);
}
__global__ void __launch_bounds__(TB_BLOCK) k_validate_utf8(const uint8_t* t, const int64_t* rs, const int64_t* re,
                                                             int64_t n, int32_t* bad) {
  int64_t d = warp_doc();
  if (d < n) body_validate_utf8(t, rs, re, bad, d, threadIdx.x & 31);
}
// FEED: the tiles of long documents arrive by bulk copy (tb_feed.cuh); the plain build carries none of that code
template <bool FEED>
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_GQ) k_gq(KBatch a, SwTable sw, GqCfg cfg, int step, tb_gq_stats* out, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  __shared__ LutEntry s_lut[256];
  load_ascii(s_ascii, a.uc);
  load_lut(s_lut, a.uc, TB_LUT_GQ, ((uint64_t)sw.secondmask << 32) | sw.firstmask);
  __shared__ uint32_t s_cand[2 * TB_GQ_STAGE];
  __shared__ SwEntry s_sw[64];  // small stop-word tables are probed in shared memory
  if (sw.mask < 64u) {
    for (uint32_t i = threadIdx.x; i <= sw.mask; i += blockDim.x) s_sw[i] = sw.e[i];
    __syncthreads();
    sw.e = s_sw;
  }
  __shared__ __align__(FEED ? TB_TILE : 16) uint8_t s_feed[FEED ? TB_FEED_SMEM : 16];  // the ring is aligned on the tile size (tile_classify relies on it)
  typename FeedOf<FEED>::type feed;
  feed_init(feed, s_feed, threadIdx.x & 31);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_gq(a, s_ascii, s_lut, s_cand, sw, cfg, step, out, d, threadIdx.x & 31, &feed);
  }
}
__global__ void k_trim(KBatch a, int64_t* ts, int64_t* te) {
  __shared__ uint16_t s_ascii[192];
  load_ascii(s_ascii, a.uc);
  int64_t d = thread_doc();
  if (d < a.n) body_trim(a, s_ascii, ts, te, d);
}
template <bool FEED>
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_GR_SCAN) k_gr_scan(KBatch a, const int64_t* ts, const int64_t* te, GrScratch s) {
  __shared__ uint16_t s_ascii[192];
  __shared__ LutEntry s_lut[256];
  load_ascii(s_ascii, a.uc);
  load_lut(s_lut, a.uc, TB_LUT_GR, 0u);
  __shared__ __align__(8) uint32_t s_stage[TB_GR_STAGE_U32];
  __shared__ __align__(FEED ? TB_TILE : 16) uint8_t s_feed[FEED ? TB_FEED_SMEM : 16];  // the ring is aligned on the tile size (tile_classify relies on it)
  typename FeedOf<FEED>::type feed;
  feed_init(feed, s_feed, threadIdx.x & 31);
  for (;;) {
    const int64_t d = next_doc(&s.counters[TB_GR_CNT_SCAN], a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_gr_scan(a, s_ascii, s_lut, s_stage, ts, te, s, d, threadIdx.x & 31, &feed);
  }
}
// Repetition pass 2, persistent CTAs pulling document ids of their tier from the lists the scan pass built.
// Warp tiers: RES CTAs of one warp per SM (the register budget follows from it).
template <int TIER, int RES>
__global__ void __launch_bounds__(32, RES) k_gr_rep_w(KBatch a, const int64_t* ts, const int64_t* te, GrScratch s, GrCfg cfg,
                                                      int step, tb_gr_stats* out) {
  constexpr uint32_t CAP = TB_REP_CAP0 << TIER;
  __shared__ __align__(16) uint8_t smem[TB_REP_WARP_SMEM(CAP)];
  const int lane = threadIdx.x;
  const uint32_t count = s.counters[TB_GR_CNT_LEN(TIER)];
  const uint32_t* list = s.lists + (uint64_t)TIER * (uint64_t)s.n;
  for (;;) {
    uint32_t i = 0;
    if (lane == 0) i = atomicAdd(&s.counters[TB_GR_CNT_CUR(TIER)], 1u);
    i = __shfl_sync(0xFFFFFFFFu, i, 0);
    if (i >= count) break;
    body_gr_rep_warp(a, smem, CAP, ts, te, s, cfg, step, out, (int64_t)list[i], lane);
    __syncwarp();
  }
}
// CTA tiers: capacity / 8 threads, the working arrays in dynamic shared memory
template <int TIER>
__global__ void __launch_bounds__((TB_REP_CAP0 << TIER) / 8) k_gr_rep_c(KBatch a, const int64_t* ts, const int64_t* te, GrScratch s,
                                                                        GrCfg cfg, int step, tb_gr_stats* out) {
  constexpr uint32_t CAP = TB_REP_CAP0 << TIER;
  extern __shared__ __align__(16) uint8_t dsmem[];
  __shared__ uint32_t s_i;
  const uint32_t count = s.counters[TB_GR_CNT_LEN(TIER)];
  const uint32_t* list = s.lists + (uint64_t)TIER * (uint64_t)s.n;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_i = atomicAdd(&s.counters[TB_GR_CNT_CUR(TIER)], 1u);
    __syncthreads();
    uint32_t i = s_i;
    if (i >= count) break;
    body_gr_rep_cta(a, dsmem, CAP, ts, te, s, cfg, step, out, (int64_t)list[i], (int)threadIdx.x);
  }
}
__global__ void __launch_bounds__(TB_REP_X_THREADS) k_gr_rep_x(KBatch a, const int64_t* ts, const int64_t* te, GrScratch s,
                                                               GrCfg cfg, int step, tb_gr_stats* out) {
  __shared__ uint64_t s_red[64 + 40];  // the CTA collectives' scratch and the per-level results of the warps' tasks (tb_rep.cuh)
  __shared__ uint32_t s_i;
  const uint32_t count = s.counters[TB_GR_CNT_LEN(TB_REP_TIER_X)];
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) s_i = atomicAdd(&s.counters[TB_GR_CNT_CUR(TB_REP_TIER_X)], 1u);
    __syncthreads();
    uint32_t i = s_i;
    if (i >= count) break;
    body_gr_rep_x(a, s_red, ts, te, s, cfg, step, out, (int64_t)s.lists[(uint64_t)TB_REP_TIER_X * (uint64_t)s.n + i], (int)threadIdx.x);
  }
}
template <bool FEED>
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_C4_A) k_c4_a(KBatch a, C4Cfg cfg, C4Scratch s, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  __shared__ LutEntry s_lut[256];
  __shared__ __align__(16) uint8_t s_stage[TB_STAGE_BYTES];
  load_ascii(s_ascii, a.uc);
  load_lut(s_lut, a.uc, TB_LUT_C4A, 0u);
  __shared__ __align__(FEED ? TB_TILE : 16) uint8_t s_feed[FEED ? TB_FEED_SMEM : 16];  // the ring is aligned on the tile size (tile_classify relies on it)
  typename FeedOf<FEED>::type feed;
  feed_init(feed, s_feed, threadIdx.x & 31);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_c4_a(a, s_ascii, s_lut, s_stage, cfg, s, d, threadIdx.x & 31, &feed);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, 16) k_entities(KBatch a, const EntEntry* tab, uint32_t ntab, uint32_t* amp, uint32_t* drop,
                                                           uint8_t* D, int64_t* drs, int64_t* dre, uint32_t* cursor) {
  __shared__ __align__(16) uint8_t s_stage[TB_SENT_STAGE_BYTES];
  for (;;) {  // runs before the work order of the chain exists: documents in index order
    uint32_t i = 0;
    if ((threadIdx.x & 31) == 0) i = atomicAdd(cursor, 1u);
    i = __shfl_sync(0xFFFFFFFFu, i, 0);
    if ((int64_t)i >= a.n) break;
    body_entities(a, s_stage, tab, ntab, amp, drop, D, drs, dre, (int64_t)i, threadIdx.x & 31);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, 16) k_c4_sent(KBatch a, const int64_t* ts, const int64_t* te, uint32_t* bits, uint8_t* S,
                                                          int64_t* srs, int64_t* sre, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  __shared__ __align__(16) uint8_t s_stage[TB_SENT_STAGE_BYTES];
  load_ascii(s_ascii, a.uc);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_c4_sent(a, s_ascii, s_stage, ts, te, bits, S, srs, sre, d, threadIdx.x & 31);
    else if ((threadIdx.x & 31) == 0) { srs[d] = tb_sent_off(a.rs[d], d); sre[d] = srs[d]; }
  }
}
template <bool FEED>
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_C4_B) k_c4_b(KBatch a, C4Cfg cfg, C4Scratch s, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  __shared__ LutEntry s_lut[256];
  load_ascii(s_ascii, a.uc);
  load_lut(s_lut, a.uc, TB_LUT_C4B, 0u);
  __shared__ __align__(FEED ? TB_TILE : 16) uint8_t s_feed[FEED ? TB_FEED_SMEM : 16];  // the ring is aligned on the tile size (tile_classify relies on it)
  typename FeedOf<FEED>::type feed;
  feed_init(feed, s_feed, threadIdx.x & 31);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d] && !s.docflags[d] && s.nlines[d] != 0) body_c4_b(a, s_ascii, s_lut, cfg, s, d, threadIdx.x & 31, &feed);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_C4_C) k_c4_c(KBatch a, C4Cfg cfg, C4Scratch s, uint32_t* cursor) {
  __shared__ __align__(16) uint8_t s_stage[TB_STAGE_BYTES];
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    body_c4_c(a, s_stage, cfg, s, d, threadIdx.x & 31);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, TB_RES_C4_D) k_c4_d(KBatch a, C4Cfg cfg, C4Scratch s, int step, tb_c4_stats* out,
                                                       uint8_t* rewritten, int64_t* rs_out, int64_t* re_out, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  load_ascii(s_ascii, a.uc);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_c4_d(a, s_ascii, cfg, s, step, out, rewritten, rs_out, re_out, d, threadIdx.x & 31);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, 24) k_fw_scan(KBatch a, FwCfg cfg, FwScratch s, uint32_t* cursor) {
  __shared__ uint16_t s_ascii[192];
  __shared__ LutEntry s_lut[256];
  load_ascii(s_ascii, a.uc);
  load_lut(s_lut, a.uc, TB_LUT_FW, 0u);
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_fw_scan(a, s_ascii, s_lut, cfg, s, d, threadIdx.x & 31);
  }
}
__global__ void __launch_bounds__(TB_BLOCK, 16) k_fw_fin(KBatch a, FwCfg cfg, FwScratch s, int step, tb_fw_stats* out, uint32_t* cursor) {
  __shared__ uint32_t s_tab[2 * TB_FW_SMEM_LINES];
  for (;;) {
    const int64_t d = next_doc(cursor, a);
    if (d == -1) break;
    if (d < 0) continue;
    if (a.alive[d]) body_fw_fin(a, s_tab, cfg, s, step, out, d, threadIdx.x & 31);
    __syncwarp();
  }
}
__global__ void __launch_bounds__(TB_BLOCK) k_c4_runs(int64_t n, const int64_t* rs0, const int64_t* re0, const uint8_t* rewritten, const uint32_t* kf,
                                                      uint32_t mis, const int64_t* frs, const int64_t* fre, int64_t* nruns, const int64_t* run_off,
                                                      tb_run_dev* runs, int64_t cap, int emit) {
  int64_t d = warp_doc();
  if (d < n) body_c4_runs(rs0, re0, rewritten, kf, mis, frs, fre, nruns, run_off, runs, cap, emit != 0, d, threadIdx.x & 31);
}
__global__ void k_final_len(int64_t n, const int64_t* rs0, const int64_t* re0, const int64_t* rs, const int64_t* re,
                            const uint8_t* rewritten, int64_t* len) {
  int64_t d = thread_doc();
  if (d < n) body_final_len(rs0, re0, rs, re, rewritten, len, d);
}
// single-block exclusive scan over the documents of a chunk (not a hot kernel)
__global__ void k_scan_i64(const int64_t* in, int64_t* out, int64_t n) {
  __shared__ int64_t s_w[32];
  __shared__ int64_t s_carry;
  int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int64_t base = 0; base < n; base += blockDim.x) {
    int64_t i = base + threadIdx.x;
    int64_t v = i < n ? in[i] : 0, x = v;
    for (int o = 1; o < 32; o <<= 1) { int64_t y = __shfl_up_sync(0xFFFFFFFFu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_w[wid] = x;
    __syncthreads();
    if (wid == 0) {
      int64_t w = lane < (int)(blockDim.x >> 5) ? s_w[lane] : 0, y = w;
      for (int o = 1; o < 32; o <<= 1) { int64_t z = __shfl_up_sync(0xFFFFFFFFu, y, o); if (lane >= o) y += z; }
      s_w[lane] = y - w;
    }
    __syncthreads();
    int64_t carry = s_carry;
    if (i < n) out[i] = carry + s_w[wid] + x - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) s_carry = carry + s_w[wid] + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) out[n] = s_carry;
}
__global__ void __launch_bounds__(TB_BLOCK) k_pack(int64_t n, const uint8_t* orig, const int64_t* rs0, const uint8_t* cur, const uint8_t* cur2,
                                                   const int64_t* rs, const uint8_t* rewritten, const int64_t* out_off,
                                                   uint8_t* out) {
  int64_t d = warp_doc();
  if (d < n) body_pack(orig, rs0, cur, cur2, rs, rewritten, out_off, out, d, threadIdx.x & 31);
}

inline int persistent_grid(int64_t n, int sm_count, int per_sm) {
  int64_t g = (int64_t)sm_count * per_sm;
  return (int)(n < g ? n : g);
}
inline int blocks_for_warps(int64_t n) { return (int)((n * 32 + TB_BLOCK - 1) / TB_BLOCK); }
inline int blocks_for_threads(int64_t n) { return (int)((n + 255) / 256); }

}  // namespace

void tb_launch_order(cudaStream_t st, const int64_t* rs, const int64_t* re, int64_t n, uint32_t* order, uint32_t* counts, uint8_t* sizecls) {
  if (!n) return;
  k_order<<<(int)((n + 255) / 256), 256, 0, st>>>(rs, re, n, order, counts, sizecls, 0);
  k_order<<<(int)((n + 255) / 256), 256, 0, st>>>(rs, re, n, order, counts, sizecls, 1);
}
void tb_launch_init_ranges(cudaStream_t st, const void* off, int off64, int64_t doc0, int64_t n, int64_t base,
                           int64_t* rs, int64_t* re, uint8_t* alive, int8_t* failed) {
  if (n) k_init_ranges<<<blocks_for_threads(n), 256, 0, st>>>(off, off64, doc0, n, base, rs, re, alive, failed);
}
void tb_launch_validate_utf8(cudaStream_t st, const uint8_t* text, const int64_t* rs, const int64_t* re, int64_t n,
                             int32_t* bad) {
  if (n) k_validate_utf8<<<blocks_for_warps(n), TB_BLOCK, 0, st>>>(text, rs, re, n, bad);
}
void tb_launch_gq(cudaStream_t st, const KBatch& b, const SwTable& sw, const GqCfg& cfg, int step, tb_gq_stats* out,
                  int sm_count, uint32_t* cursor) {
  if (b.n) {
    if (b.feed_min_tiles) k_gq<true><<<persistent_grid(b.n, sm_count, TB_RES_GQ), TB_BLOCK, 0, st>>>(b, sw, cfg, step, out, cursor);
    else k_gq<false><<<persistent_grid(b.n, sm_count, TB_RES_GQ), TB_BLOCK, 0, st>>>(b, sw, cfg, step, out, cursor);
  }
}
void tb_launch_trim(cudaStream_t st, const KBatch& b, int64_t* ts, int64_t* te) {
  if (b.n) k_trim<<<blocks_for_threads(b.n), 256, 0, st>>>(b, ts, te);
}
void tb_launch_gr_scan(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, const GrScratch& s, int sm_count) {
  if (b.n) {
    if (b.feed_min_tiles) k_gr_scan<true><<<persistent_grid(b.n, sm_count, TB_RES_GR_SCAN), TB_BLOCK, 0, st>>>(b, ts, te, s);
    else k_gr_scan<false><<<persistent_grid(b.n, sm_count, TB_RES_GR_SCAN), TB_BLOCK, 0, st>>>(b, ts, te, s);
  }
}
void tb_launch_gr_rep(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, const GrScratch& s,
                      const GrCfg& cfg, int step, tb_gr_stats* out, int sm_count, int tier) {
  if (!b.n) return;
  // grids: what fits an SM by shared memory (227 KB) and registers, a little over for the tiers whose CTAs are small
  if (tier == 0) k_gr_rep_w<0, 32><<<sm_count * 32, 32, 0, st>>>(b, ts, te, s, cfg, step, out);
  else if (tier == 1) k_gr_rep_w<1, 17><<<sm_count * 17, 32, 0, st>>>(b, ts, te, s, cfg, step, out);
  else if (tier == 2) k_gr_rep_c<2><<<sm_count * 8, tb_rep_threads(2), TB_REP_CTA_SMEM(tb_rep_cap(2)), st>>>(b, ts, te, s, cfg, step, out);
  else if (tier == 3) k_gr_rep_c<3><<<sm_count * 4, tb_rep_threads(3), TB_REP_CTA_SMEM(tb_rep_cap(3)), st>>>(b, ts, te, s, cfg, step, out);
  else if (tier == 4) k_gr_rep_c<4><<<sm_count * 2, tb_rep_threads(4), TB_REP_CTA_SMEM(tb_rep_cap(4)), st>>>(b, ts, te, s, cfg, step, out);
  else if (tier == 5) k_gr_rep_c<5><<<sm_count, tb_rep_threads(5), TB_REP_CTA_SMEM(tb_rep_cap(5)), st>>>(b, ts, te, s, cfg, step, out);
  else k_gr_rep_x<<<sm_count * TB_REP_X_PER_SM, TB_REP_X_THREADS, 0, st>>>(b, ts, te, s, cfg, step, out);  // global-memory tables: several CTAs per SM hide the latency
}
void tb_launch_fw_scan(cudaStream_t st, const KBatch& b, const FwCfg& cfg, const FwScratch& s, int sm_count, uint32_t* cursor) {
  if (b.n) k_fw_scan<<<persistent_grid(b.n, sm_count, 24), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
}
void tb_launch_fw_fin(cudaStream_t st, const KBatch& b, const FwCfg& cfg, const FwScratch& s, int step, tb_fw_stats* out, int sm_count,
                      uint32_t* cursor) {
  if (b.n) k_fw_fin<<<persistent_grid(b.n, sm_count, 16), TB_BLOCK, 0, st>>>(b, cfg, s, step, out, cursor);
}
#if defined(TB_REP_PROFILE)
// measurement build: the clocks of the repetition pass per tier and phase (tb_rep.cuh), read out and cleared
extern "C" int tb_debug_rep_profile(unsigned long long* out) {
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out, tb_rep_prof, sizeof(unsigned long long) * 8 * 32) != cudaSuccess) return 1;
  unsigned long long zero[8 * 32] = {};
  return cudaMemcpyToSymbol(tb_rep_prof, zero, sizeof zero) == cudaSuccess ? 0 : 1;
}
#endif
int tb_kernels_init() {
  cudaError_t e = cudaFuncSetAttribute(k_gr_rep_c<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TB_REP_CTA_SMEM(tb_rep_cap(3)));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gr_rep_c<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TB_REP_CTA_SMEM(tb_rep_cap(4)));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_gr_rep_c<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TB_REP_CTA_SMEM(tb_rep_cap(5)));
  return (int)e;
}
void tb_launch_c4_a(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor) {
  if (b.n) {
    if (b.feed_min_tiles) k_c4_a<true><<<persistent_grid(b.n, sm_count, TB_RES_C4_A), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
    else k_c4_a<false><<<persistent_grid(b.n, sm_count, TB_RES_C4_A), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
  }
}
void tb_launch_entities(cudaStream_t st, const KBatch& b, const EntEntry* tab, uint32_t ntab, uint32_t* amp, uint32_t* drop,
                        uint8_t* D, int64_t* drs, int64_t* dre, int sm_count, uint32_t* cursor) {
  if (b.n) k_entities<<<persistent_grid(b.n, sm_count, 16), TB_BLOCK, 0, st>>>(b, tab, ntab, amp, drop, D, drs, dre, cursor);
}
void tb_launch_c4_sent(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, uint32_t* bits, uint8_t* S,
                       int64_t* srs, int64_t* sre, int sm_count, uint32_t* cursor) {
  if (b.n) k_c4_sent<<<persistent_grid(b.n, sm_count, 16), TB_BLOCK, 0, st>>>(b, ts, te, bits, S, srs, sre, cursor);
}
void tb_launch_c4_b(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor) {
  if (b.n) {
    if (b.feed_min_tiles) k_c4_b<true><<<persistent_grid(b.n, sm_count, TB_RES_C4_B), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
    else k_c4_b<false><<<persistent_grid(b.n, sm_count, TB_RES_C4_B), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
  }
}
void tb_launch_c4_c(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor) {
  if (b.n) k_c4_c<<<persistent_grid(b.n, sm_count, TB_RES_C4_C), TB_BLOCK, 0, st>>>(b, cfg, s, cursor);
}
void tb_launch_c4_d(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int step, tb_c4_stats* out,
                    uint8_t* rewritten, int64_t* rs_out, int64_t* re_out, int sm_count, uint32_t* cursor) {
  // the new ranges go to rs_out / re_out (the chain's rs / re), only for documents that were rewritten
  if (b.n) k_c4_d<<<persistent_grid(b.n, sm_count, TB_RES_C4_D), TB_BLOCK, 0, st>>>(b, cfg, s, step, out, rewritten, rs_out, re_out, cursor);
}
void tb_launch_c4_runs(cudaStream_t st, int64_t n, const int64_t* rs0, const int64_t* re0, const uint8_t* rewritten, const uint32_t* kf,
                       uint32_t mis, const int64_t* frs, const int64_t* fre, int64_t* nruns, const int64_t* run_off, void* runs, int64_t cap, int emit) {
  if (n) k_c4_runs<<<blocks_for_warps(n), TB_BLOCK, 0, st>>>(n, rs0, re0, rewritten, kf, mis, frs, fre, nruns, run_off, (tb_run_dev*)runs, cap, emit);
}
void tb_launch_final_len(cudaStream_t st, int64_t n, const int64_t* rs0, const int64_t* re0, const int64_t* rs,
                         const int64_t* re, const uint8_t* rewritten, int64_t* len) {
  if (n) k_final_len<<<blocks_for_threads(n), 256, 0, st>>>(n, rs0, re0, rs, re, rewritten, len);
}
void tb_launch_scan_i64(cudaStream_t st, const int64_t* in, int64_t* out, int64_t n) {
  k_scan_i64<<<1, 1024, 0, st>>>(in, out, n);
}
void tb_launch_pack(cudaStream_t st, int64_t n, const uint8_t* orig, const int64_t* rs0, const uint8_t* cur, const uint8_t* cur2,
                    const int64_t* rs, const uint8_t* rewritten, const int64_t* out_off, uint8_t* out) {
  if (n) k_pack<<<blocks_for_warps(n), TB_BLOCK, 0, st>>>(n, orig, rs0, cur, cur2, rs, rewritten, out_off, out);
}
