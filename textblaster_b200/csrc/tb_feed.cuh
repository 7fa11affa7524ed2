// tb_feed.cuh -- bulk-copy staging of the text tiles of long documents.
//
// The text passes walk a document tile by tile (1 KiB, one 32-byte stretch per lane).  For a short document the two
// 16-byte loads per lane are all there is to it.  For a long one the warp asks the copy engine for whole tiles ahead of
// time instead: lane 0 issues one `cp.async.bulk` (global -> shared, 1 KiB, completion counted on an mbarrier) per tile
// into a small ring in shared memory, and the lanes read their stretches from there when the tile's barrier has flipped.
// The loads of tile k + 1 are then in flight while tile k is worked on, and no lane spends issue slots on address
// arithmetic and guarded vector loads for them.  Tiles that are not wholly inside the text buffer (its first and last
// bytes) are not requested: the caller falls back to the per-lane loads for those.
//
// One warp per CTA (TB_BLOCK = 32), so "the warp" and "the CTA" are the same thing here.  The barriers live for the whole
// kernel: every requested tile is waited for before the document is left, so the phase bits carry over from one
// document to the next.
#pragma once
#include <stdint.h>

#include "tb_common.cuh"

#ifndef TB_TILE
#define TB_STRETCH 32
#define TB_TILE (32 * TB_STRETCH)
#endif

#define TB_FEED_STAGES 2
#define TB_FEED_SMEM (TB_FEED_STAGES * TB_TILE + TB_FEED_STAGES * 8)
#define TB_FEED_NONE 0xFFFFFFFFu

// What the kernels without bulk staging are built with: nothing is staged, and every call folds away.
struct NoFeed {
  TB_HD void begin(const uint8_t*, int64_t, int64_t, int64_t, int64_t, int, uint32_t) {}
  TB_HD uint32_t tile(int64_t) const { return 0xFFFFFFFFu; }
  TB_HD void done(int64_t, int) {}
  TB_HD void drain(int64_t) {}
};

#if defined(__CUDA_ARCH__)
struct TileFeed {
  uint32_t buf;      // shared-memory address of the ring (TB_FEED_STAGES tiles, 128-byte aligned)
  uint32_t bar;      // shared-memory address of the mbarriers
  const uint8_t* t;
  int64_t ext_lo, ext_hi, end;
  int64_t origin;    // base of tile 0 of the current document
  uint32_t phase;    // bit s: the parity stage s flips to next
  bool on;

  // once per kernel; `shared` = TB_FEED_SMEM bytes, 128-byte aligned
  __device__ __forceinline__ void init(uint8_t* shared, int lane) {
    buf = (uint32_t)__cvta_generic_to_shared(shared);
    bar = buf + TB_FEED_STAGES * TB_TILE;
    phase = 0; on = false;
    if (lane == 0) {
      for (int s = 0; s < TB_FEED_STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar + 8u * s) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
  }
  __device__ __forceinline__ bool wanted(int64_t base) const { return base >= ext_lo && base + TB_TILE <= ext_hi && base < end; }
  __device__ __forceinline__ void request(int64_t base, int s) {  // lane 0 only
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar + 8u * s), "r"((uint32_t)TB_TILE) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(buf + (uint32_t)(s * TB_TILE)),
                 "l"(t + base), "r"((uint32_t)TB_TILE), "r"(bar + 8u * s)
                 : "memory");
  }
  // start of a document: first tile at `first` (tb_tile_origin), text ends at `re`
  // min_tiles: documents of at least this many tiles are staged (0: none)
  __device__ __forceinline__ void begin(const uint8_t* text, int64_t lo, int64_t hi, int64_t first, int64_t re, int lane, uint32_t min_tiles) {
    t = text; ext_lo = lo; ext_hi = hi; end = re; origin = first;
    on = min_tiles != 0u && re - first >= (int64_t)min_tiles * TB_TILE;
    if (on && lane == 0)
      for (int s = 0; s < TB_FEED_STAGES; s++) if (wanted(first + (int64_t)s * TB_TILE)) request(first + (int64_t)s * TB_TILE, s);
  }
  // shared-memory address of the staged copy of the tile at `base`, or TB_FEED_NONE (not staged: read it from global
  // memory).  Every lane calls this.
  __device__ __forceinline__ uint32_t tile(int64_t base) {
    if (!on || !wanted(base)) return TB_FEED_NONE;
    const int s = (int)(((base - origin) / TB_TILE) % TB_FEED_STAGES);
    const uint32_t par = (phase >> s) & 1u;
    uint32_t ok;
    do {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar + 8u * s), "r"(par) : "memory");
    } while (!ok);
    phase ^= 1u << s;
    return buf + (uint32_t)(s * TB_TILE);
  }
  // the document is left early after done(base): wait for the tiles that have been requested beyond `base`
  __device__ __forceinline__ void drain(int64_t base) {
    if (!on) return;
    for (int k = 1; k <= TB_FEED_STAGES; k++) (void)tile(base + (int64_t)k * TB_TILE);
    on = false;
  }
  // the tile at `base` has been read by every lane: its stage takes the tile TB_FEED_STAGES further on.  Every lane calls this.
  __device__ __forceinline__ void done(int64_t base, int lane) {
    if (!on) return;
    __syncwarp();
    const int64_t nxt = base + (int64_t)TB_FEED_STAGES * TB_TILE;
    if (lane == 0 && wanted(nxt)) request(nxt, (int)(((base - origin) / TB_TILE) % TB_FEED_STAGES));
  }
};
#else
struct TileFeed {  // host emulation: nothing is staged
  bool on = false;
  void init(uint8_t*, int) {}
  void begin(const uint8_t*, int64_t, int64_t, int64_t, int64_t, int, uint32_t) {}
  uint32_t tile(int64_t) { return TB_FEED_NONE; }
  void done(int64_t, int) {}
  void drain(int64_t) {}
};
#endif

template <bool FEED> struct FeedOf { using type = NoFeed; };
template <> struct FeedOf<true> { using type = TileFeed; };
TB_HD void feed_init(NoFeed&, uint8_t*, int) {}
TB_HD void feed_init(TileFeed& f, uint8_t* shared, int lane) { f.init(shared, lane); }
