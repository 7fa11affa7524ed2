// tb_tile.cuh -- bit-parallel view of one lane's 32-byte stretch.
//
// The text passes of GopherQuality, GopherRepetition (scan) and C4 (pass B) all need the UAX #29 word
// segments of the text (/root/reference src/utils/text.rs:103-181).  Instead of walking segment by segment,
// a lane turns its 32 bytes into 32-bit class masks (one shared-memory table lookup per byte, a nibble
// transpose), derives the mask of word boundaries with a dozen shifts and ANDs, and works on whole masks:
// counts are popcounts, per-segment properties are segmented prefix-ORs, hashes run in one converged
// 32-iteration loop.  Segments that cross a stretch boundary are stitched by a warp scan over "open
// fragment" summaries, and across tiles by a carry.
//
// Boundary decisions that involve a non-ASCII byte or a '\r' are "dirty": for those positions the general
// rule machine (tb_wordmachine.cuh) supplies the decision and the class bits of the character, so the
// masks are exact for any text.  The only thing the mask form cannot express is a segment that mixes
// White_Space and other characters (it would need trimming, text.rs:125-137) -- U+202F next to a letter,
// white space followed by a combining mark.  Such a document is flagged TILE_EXOTIC and redone by the
// segment-walking bodies, which stay in the library as the general path.
#pragma once
#include "tb_fastwords.cuh"
#include "tb_wordmachine.cuh"
#include "tb_feed.cuh"

#ifndef TB_STRETCH
#define TB_STRETCH 32
#define TB_TILE (32 * TB_STRETCH)
#endif

// class slots of the lookup table, low word (the same for every kernel)
#define TL_A 0    // ALetter
#define TL_N 1    // Numeric
#define TL_MLQ 2  // MidLetter | MidNumLet | Single_Quote
#define TL_MNQ 3  // MidNum | MidNumLet | Single_Quote
#define TL_E 4    // ExtendNumLet
#define TL_S 5    // WSegSpace
#define TL_NA 6   // non-ASCII byte or '\r': decisions around it come from the rule machine
#define TL_WC 7   // neither PUNCTUATION nor White_Space (text.rs:141); for non-ASCII set by the rule machine
// high word (per kernel)
#define TH_CONT 0  // UTF-8 continuation byte
#define TH_NL 1    // '\n'
#define TH_HASH 2  // GopherQuality: '#'
#define TH_DOT 3   //                '.'
#define TH_DASH 4  //                '-'
#define TH_SWF 5   //                byte can start a stop word (SwTable::firstmask)
#define TH_E2 6    //                0xE2, lead byte of U+2026 and U+2022
#define TH_SW2 7   //                byte can be the second byte of a stop word (SwTable::secondmask)
#define TH_P1 2    // C4 pass B: first / second / third letter of a javascript / policy pattern
#define TH_P2 3
#define TH_P3 4
// C4 pass A uses its own slot layout (no word rules there)
#define CA_WS 0     // ASCII White_Space (includes '\n')
#define CA_NL 1     // '\n'
#define CA_NA 2     // non-ASCII byte
#define CA_CONT 3   // UTF-8 continuation byte
#define CA_LBR 4    // '['
#define CA_CURLY 5  // '{' or '}'
#define CA_LL 6     // 'l' or 'L'
#define CA_LO 7     // 'o' or 'O'
#define CH_WSLEAD 0 // lead byte of a non-ASCII character that may be White_Space (UcTables::wslead)
#define FH_WS 2      // FineWeb: ASCII White_Space
#define FH_WSLEAD 3  //          lead byte of a non-ASCII character that may be White_Space
#define FH_CR 4      //          '\r'
enum { TB_LUT_GQ = 0, TB_LUT_GR = 1, TB_LUT_C4B = 2, TB_LUT_C4A = 3, TB_LUT_FW = 4 };

struct __attribute__((aligned(8))) LutEntry { uint32_t lo, hi; };

// table entry of byte value b; ci = cinfo of the code point b (only used for b < 128); aux: SwTable::firstmask in the low
// and SwTable::secondmask in the high word (TB_LUT_GQ), or "b is a lead byte that may begin a White_Space character"
// (TB_LUT_C4A, TB_LUT_FW)
TB_HD LutEntry tb_lut_entry(int kind, uint32_t b, uint32_t ci, uint64_t aux) {
  const uint32_t firstmask = (uint32_t)aux, secondmask = (uint32_t)(aux >> 32);
  LutEntry e; e.lo = 0; e.hi = 0;
  if (kind == TB_LUT_C4A) {
    if (b >= 0x80u) {
      e.lo |= 1u << (4 * CA_NA);
      if ((b & 0xC0u) == 0x80u) e.lo |= 1u << (4 * CA_CONT);
      else if (aux) e.hi |= 1u << (4 * CH_WSLEAD);
    } else {
      if (ci & CI_WS) e.lo |= 1u << (4 * CA_WS);
      if (b == '\n') e.lo |= 1u << (4 * CA_NL);
      if (b == '[') e.lo |= 1u << (4 * CA_LBR);
      if (b == '{' || b == '}') e.lo |= 1u << (4 * CA_CURLY);
      if ((b | 0x20u) == 'l') e.lo |= 1u << (4 * CA_LL);
      if ((b | 0x20u) == 'o') e.lo |= 1u << (4 * CA_LO);
    }
    return e;
  }
  if (b >= 0x80u) {
    e.lo |= 1u << (4 * TL_NA);
    if ((b & 0xC0u) == 0x80u) e.hi |= 1u << (4 * TH_CONT);
  } else if (b == '\r') {
    e.lo |= 1u << (4 * TL_NA);
  } else {
    uint32_t f = tb_ascii_flags(ci);
    if (f & AF_A) e.lo |= 1u << (4 * TL_A);
    if (f & AF_N) e.lo |= 1u << (4 * TL_N);
    if (f & AF_MLQ) e.lo |= 1u << (4 * TL_MLQ);
    if (f & AF_MNQ) e.lo |= 1u << (4 * TL_MNQ);
    if (f & AF_E) e.lo |= 1u << (4 * TL_E);
    if (f & AF_S) e.lo |= 1u << (4 * TL_S);
    if (f & AF_WC) e.lo |= 1u << (4 * TL_WC);
  }
  if (b == '\n') e.hi |= 1u << (4 * TH_NL);
  if (kind == TB_LUT_GQ) {
    if (b == '#') e.hi |= 1u << (4 * TH_HASH);
    if (b == '.') e.hi |= 1u << (4 * TH_DOT);
    if (b == '-') e.hi |= 1u << (4 * TH_DASH);
    if (b == 0xE2u) e.hi |= 1u << (4 * TH_E2);
    // a non-ASCII first byte may change under to_lowercase(): always a candidate
    uint32_t fl = (b | 0x20u) - 'a';
    if (b >= 0x80u || (firstmask & (fl < 26u ? (1u << fl) : (1u << 26)))) e.hi |= 1u << (4 * TH_SWF);
    if (b >= 0x80u || (secondmask & (fl < 26u ? (1u << fl) : (1u << 26)))) e.hi |= 1u << (4 * TH_SW2);
  } else if (kind == TB_LUT_FW) {
    if (b < 0x80u) {
      if (ci & CI_WS) e.hi |= 1u << (4 * FH_WS);
      if (b == '\r') e.hi |= 1u << (4 * FH_CR);
    } else if ((b & 0xC0u) != 0x80u && aux) {
      e.hi |= 1u << (4 * FH_WSLEAD);
    }
  } else if (kind == TB_LUT_C4B) {
    if (b < 0x80u) {
      uint32_t l = b | 0x20u;
      if (l == 'c' || l == 'j' || l == 'p' || l == 't' || l == 'u') e.hi |= 1u << (4 * TH_P1);
      if (l == 'a' || l == 'e' || l == 'r' || l == 'o' || l == 's') e.hi |= 1u << (4 * TH_P2);
      if (l == 'v' || l == 'r' || l == 'i' || l == 'o' || l == 'e') e.hi |= 1u << (4 * TH_P3);
    }
  }
  return e;
}

// ---- small bit helpers -----------------------------------------------------------------------
TB_HD uint32_t tb_funnel_l(uint32_t lo, uint32_t hi, uint32_t n) {  // high word of (hi:lo) << n, 0 < n < 32
#if defined(__CUDA_ARCH__)
  return __funnelshift_l(lo, hi, n);
#else
  return (hi << n) | (lo >> (32u - n));
#endif
}
TB_HD uint32_t tb_funnel_r1(uint32_t lo, uint32_t hi) {  // (lo >> 1) with bit 0 of hi entering at the top
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, 1);
#else
  return (lo >> 1) | (hi << 31);
#endif
}
TB_HD int tb_ctz32(uint32_t v) {  // 32 for v == 0
#if defined(__CUDA_ARCH__)
  return v ? __ffs((int)v) - 1 : 32;
#else
  return v ? __builtin_ctz(v) : 32;
#endif
}
TB_HD int tb_clz32(uint32_t v) {
#if defined(__CUDA_ARCH__)
  return __clz((int)v);
#else
  return v ? __builtin_clz(v) : 32;
#endif
}
TB_HD int tb_popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}
TB_HD uint32_t tb_below(int n) { return n >= 32 ? 0xFFFFFFFFu : ((1u << n) - 1u); }       // bits [0, n)
TB_HD uint32_t tb_upto(int n) { return n >= 31 ? 0xFFFFFFFFu : ((2u << n) - 1u); }        // bits [0, n]
TB_HD uint32_t tb_prmt(uint32_t a, uint32_t b, uint32_t sel) {
#if defined(__CUDA_ARCH__)
  return __byte_perm(a, b, sel);
#else
  uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
  return r;
#endif
}
// 8 x 8 transpose of a nibble matrix held in eight words: out[c] nibble k = in[k] nibble c
TB_HD void tb_transpose_nibbles(uint32_t* t) {
#pragma unroll
  for (int k = 0; k < 8; k += 2) {
    uint32_t a = t[k], b = t[k + 1];
    t[k] = (a & 0x0F0F0F0Fu) | ((b & 0x0F0F0F0Fu) << 4);
    t[k + 1] = ((a >> 4) & 0x0F0F0F0Fu) | (b & 0xF0F0F0F0u);
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    if (k & 2) continue;
    uint32_t a = t[k], b = t[k + 2];
    t[k] = tb_prmt(a, b, 0x6240);      // a0 b0 a2 b2
    t[k + 2] = tb_prmt(a, b, 0x7351);  // a1 b1 a3 b3
  }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    uint32_t a = t[k], b = t[k + 4];
    t[k] = tb_prmt(a, b, 0x5410);      // a.lo16 b.lo16
    t[k + 4] = tb_prmt(a, b, 0x7632);  // a.hi16 b.hi16
  }
}
// Segmented prefix-OR inside one word.  B: bit i set = a segment starts at i.
TB_HD uint32_t tb_spread_up(uint32_t x, uint32_t B) {  // towards higher bits
  uint32_t m = ~B;
  x |= (x << 1) & m; m &= m << 1;
  x |= (x << 2) & m; m &= m << 2;
  x |= (x << 4) & m; m &= m << 4;
  x |= (x << 8) & m; m &= m << 8;
  x |= (x << 16) & m;
  return x;
}
TB_HD uint32_t tb_spread_down(uint32_t x, uint32_t B) {  // towards lower bits
  uint32_t m = ~B >> 1;
  x |= (x >> 1) & m; m &= m >> 1;
  x |= (x >> 2) & m; m &= m >> 2;
  x |= (x >> 4) & m; m &= m >> 4;
  x |= (x >> 8) & m; m &= m >> 8;
  x |= (x >> 16) & m;
  return x;
}
TB_HD uint32_t tb_spread(uint32_t x, uint32_t B) { return tb_spread_up(x, B) | tb_spread_down(x, B); }

// ---- lazily reduced Mersenne-31 hash step (same residues as Hash2::push; bases below 2^29) ----
TB_HD uint32_t tb_hpush(uint32_t a, uint32_t base, uint32_t v) {  // a < 2^32, v <= 256 -> result < 2^32
  uint64_t x = (uint64_t)a * base + v;
  return (uint32_t)(x >> 31) + ((uint32_t)x & TB_P31);
}
TB_HD uint32_t tb_hfold(uint32_t a) {  // -> <= p
  a = (a & TB_P31) + (a >> 31);
  return (a & TB_P31) + (a >> 31);
}

// ---- the tile ------------------------------------------------------------------------------------
#define TILE_EXOTIC 1u  // a segment mixes White_Space and other characters: the walking bodies take the document

struct Tile {
  int64_t p0;        // position of bit 0
  uint32_t w[8];     // the 32 bytes (0 outside the range)
  uint32_t valid;    // bytes inside [rs, re)
  uint32_t m[8];     // TL_* masks
  uint32_t h[8];     // TH_* masks
  uint32_t lctx;     // low-word classes of bytes p0-2 (bit 0 of each nibble) and p0-1 (bit 1)
  uint32_t rctx;     // low-word classes of byte p0+32 (bit 0 of each nibble)
  uint32_t l1hi;     // high-word table entry of byte p0-1
  uint32_t brk;      // bit i: a segment starts at p0+i (all invalid positions are set)
  uint32_t dirty;    // decisions the masks cannot make
  uint32_t al;       // Alphabetic characters (lead bytes)
};

TB_HD void tb_load32(const uint8_t* p, uint32_t* w) {
#if defined(__CUDA_ARCH__)
  const uint4 x = __ldg((const uint4*)p), y = __ldg((const uint4*)(p + 16));
  w[0] = x.x; w[1] = x.y; w[2] = x.z; w[3] = x.w; w[4] = y.x; w[5] = y.y; w[6] = y.z; w[7] = y.w;
#else
  for (int k = 0; k < 8; k++) w[k] = (uint32_t)p[4 * k] | ((uint32_t)p[4 * k + 1] << 8) | ((uint32_t)p[4 * k + 2] << 16) | ((uint32_t)p[4 * k + 3] << 24);
#endif
}
TB_HD uint32_t tb_byte_shared(uint32_t addr) {  // one byte at a shared-memory address (device only)
#if defined(__CUDA_ARCH__)
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
#else
  (void)addr;
  return 0;
#endif
}
// first tile position of a range: the stretch of lane 0 is the 32-byte aligned block that holds rs
TB_HD int64_t tb_tile_origin(const TextRange& R) { return R.rs - (int64_t)((uintptr_t)(R.t + R.rs) & 31u); }

// Load the stretch at p0 and look its bytes up: T.w, T.valid, T.m[] (and T.h[] if HI).  Returns false for a
// stretch without any byte of the range.  [ext_lo, ext_hi): extent of the buffer that vector loads may touch.
// the 32 bytes at p0 (0 outside [rs, re)); returns the mask of bytes inside the range
// `staged`: shared-memory address of the stretch in a tile the copy engine brought in (tb_feed.cuh), or TB_FEED_NONE
TB_HD uint32_t tile_load(const TextRange& R, int64_t ext_lo, int64_t ext_hi, int64_t p0, uint32_t* w, uint32_t staged = TB_FEED_NONE) {
  const uint8_t* t = R.t;
  const int64_t a = R.rs - p0, b = R.re - p0;
  const uint32_t vlo = a <= 0 ? 0xFFFFFFFFu : (a >= 32 ? 0u : ~((1u << (uint32_t)a) - 1u));
  const uint32_t vhi = b >= 32 ? 0xFFFFFFFFu : (b <= 0 ? 0u : ((1u << (uint32_t)b) - 1u));
  const uint32_t V = vlo & vhi;
  if (V == 0u) {
#pragma unroll
    for (int k = 0; k < 8; k++) w[k] = 0;
  } else if (staged != TB_FEED_NONE || (p0 >= ext_lo && p0 + 32 <= ext_hi)) {
#if defined(__CUDA_ARCH__)
    if (staged != TB_FEED_NONE) {
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(staged));
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]) : "r"(staged + 16u));
    } else
#endif
      tb_load32(t + p0, w);
    if (V != 0xFFFFFFFFu) {
#pragma unroll
      for (int k = 0; k < 8; k++) {
        uint32_t nib = (V >> (4 * k)) & 15u;
        w[k] &= ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
      }
    }
  } else {
#pragma unroll
    for (int k = 0; k < 8; k++) w[k] = 0;
    for (int i = 0; i < 32; i++) if ((V >> i) & 1u) w[i >> 2] |= tb_byte(t, p0 + i) << (8 * (i & 3));
  }
  return V;
}
// mask of the bytes equal to c (SWAR compare, no table)
TB_HD uint32_t tile_eq_mask(const uint32_t* w, uint32_t c) {
  const uint32_t cc = c * 0x01010101u;
  uint32_t m = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const uint32_t z = w[k] ^ cc;
    const uint32_t t = ~(((z & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | z | 0x7F7F7F7Fu);  // 0x80 where the byte of z is zero
    m |= ((((t >> 7) * 0x00204081u) >> 21) & 15u) << (4 * k);
  }
  return m;
}

// Load the stretch at p0 and look its bytes up: T.w, T.valid, T.m[] (and T.h[] if HI).  Returns false for a
// stretch without any byte of the range.  [ext_lo, ext_hi): extent of the buffer that vector loads may touch.
template <bool HI>
TB_HD bool tile_gather(const LutEntry* lut, const TextRange& R, int64_t ext_lo, int64_t ext_hi, int64_t p0, Tile& T, uint32_t staged = TB_FEED_NONE) {
  T.p0 = p0;
  const uint32_t V = tile_load(R, ext_lo, ext_hi, p0, T.w, staged);
  T.valid = V;
  if (V == 0u) {
#pragma unroll
    for (int k = 0; k < 8; k++) { T.m[k] = 0; T.h[k] = 0; }
    T.lctx = T.rctx = T.l1hi = 0; T.brk = 0xFFFFFFFFu; T.dirty = 0; T.al = 0;
    return false;
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const uint32_t x = T.w[k];
    const LutEntry e0 = lut[x & 0xFFu], e1 = lut[(x >> 8) & 0xFFu], e2 = lut[(x >> 16) & 0xFFu], e3 = lut[x >> 24];
    T.m[k] = e0.lo | (e1.lo << 1) | (e2.lo << 2) | (e3.lo << 3);
    if (HI) T.h[k] = e0.hi | (e1.hi << 1) | (e2.hi << 2) | (e3.hi << 3);
  }
  tb_transpose_nibbles(T.m);
  if (HI) tb_transpose_nibbles(T.h);
  if (V != 0xFFFFFFFFu) {
#pragma unroll
    for (int k = 0; k < 8; k++) { T.m[k] &= V; if (HI) T.h[k] &= V; }
  }
  return true;
}

// Load and classify for the word passes: tile_gather plus the boundary mask.
template <bool HI>
TB_HD void tile_classify(const LutEntry* lut, const TextRange& R, int64_t ext_lo, int64_t ext_hi, int64_t p0, Tile& T, uint32_t staged = TB_FEED_NONE) {
  const uint8_t* t = R.t;
  if (!tile_gather<HI>(lut, R, ext_lo, ext_hi, p0, T, staged)) return;
  const uint32_t V = T.valid;
  // context bytes
  {
    const int64_t q2 = p0 - 2, q1 = p0 - 1, qr = p0 + 32;
    uint32_t l2 = 0, l1 = 0, l1h = 0, r1 = 0;
    // (in a staged tile -- the ring is aligned on the tile size -- the neighbours' bytes are in shared memory too,
    // except in front of the first stretch and behind the last)
    const bool sl = staged != TB_FEED_NONE && (staged & (TB_TILE - 1u)) != 0u, sr = staged != TB_FEED_NONE && (staged & (TB_TILE - 1u)) != TB_TILE - 32u;
    if (q2 >= R.rs && q2 < R.re) l2 = lut[sl ? tb_byte_shared(staged - 2u) : tb_byte(t, q2)].lo;
    if (q1 >= R.rs && q1 < R.re) { LutEntry e = lut[sl ? tb_byte_shared(staged - 1u) : tb_byte(t, q1)]; l1 = e.lo; l1h = e.hi; }
    if (qr >= R.rs && qr < R.re) r1 = lut[sr ? tb_byte_shared(staged + 32u) : tb_byte(t, qr)].lo;
    T.lctx = l2 | (l1 << 1); T.rctx = r1; T.l1hi = l1h;
  }
#define TILE_L(c) (T.lctx << (30 - 4 * (c)))
#define TILE_SHL1(c) tb_funnel_l(TILE_L(c), T.m[c], 1)
#define TILE_SHL2(c) tb_funnel_l(TILE_L(c), T.m[c], 2)
#define TILE_SHR1(c) tb_funnel_r1(T.m[c], T.rctx >> (4 * (c)))
  const uint32_t A = T.m[TL_A], N = T.m[TL_N], Q = T.m[TL_MLQ], K = T.m[TL_MNQ], E = T.m[TL_E], S = T.m[TL_S], NA = T.m[TL_NA];
  const uint32_t A1 = TILE_SHL1(TL_A), A2 = TILE_SHL2(TL_A), Ar = TILE_SHR1(TL_A);
  const uint32_t N1 = TILE_SHL1(TL_N), N2 = TILE_SHL2(TL_N), Nr = TILE_SHR1(TL_N);
  const uint32_t Q1 = TILE_SHL1(TL_MLQ), K1 = TILE_SHL1(TL_MNQ), E1 = TILE_SHL1(TL_E), S1 = TILE_SHL1(TL_S);
  const uint32_t AN = A | N, AN1 = A1 | N1;
  const uint32_t nb = (AN1 & AN)                            // WB5, WB8, WB9, WB10
                      | (A1 & Q & Ar) | (A2 & Q1 & A)       // WB6, WB7
                      | (N1 & K & Nr) | (N2 & K1 & N)       // WB12, WB11
                      | ((AN1 | E1) & E) | (E1 & AN)        // WB13a, WB13b
                      | (S1 & S);                           // WB3d
  T.brk = ~nb | ~V;
  T.dirty = (NA | TILE_SHL1(TL_NA) | TILE_SHL2(TL_NA) | TILE_SHR1(TL_NA)) & V;
  T.al = A;
#undef TILE_L
#undef TILE_SHL1
#undef TILE_SHL2
#undef TILE_SHR1
}

// Decisions and character classes of the dirty positions, from the rule machine (out of line: the rule
// machine is large and most stretches never need it).  fl: TILE_EXOTIC and/or CI_COMPLEX (a character of a
// dictionary-segmented script was seen).
struct DirtyFix { uint32_t span, brk, wc, al, fl; };  // span: positions decided; brk / wc / al: their bits
TB_HD_NOINLINE void tile_dirty_bits(const Classifier& C, const TextRange& R, int64_t p0, uint32_t dirty, DirtyFix& o) {
  const uint8_t* t = R.t;
  uint32_t dm = dirty;
  o.span = dirty; o.brk = 0; o.wc = 0; o.al = 0; o.fl = 0;
  while (dm) {
    const int s = tb_ctz32(dm);
    int hi = s + tb_ctz32(~(dm >> s));
    if (hi > 32) hi = 32;
    dm &= ~(tb_below(hi) & ~tb_below(s));
    int64_t q = p0 + s;
    const int64_t q1 = p0 + hi;
    while (q < q1 && q > R.rs && tb_is_cont(tb_byte(t, q))) q++;
    if (q >= q1) continue;
    WbState st;
    wb_sync(C, R, q, st);
    uint32_t prev_ws = 0;
    if (q > R.rs) { int64_t k = C.prev_lead(t, q, R.rs); int l0; uint32_t c0; prev_ws = C.at(t, k, &l0, &c0) & CI_WS; }
    int64_t i = q;
    while (i < q1) {
      int len; uint32_t cp;
      const uint32_t ci = C.at(t, i, &len, &cp), cls = CI_WB(ci);
      const uint32_t bit = 1u << (uint32_t)(i - p0);
      if (wb_break_before(C, R, st, cls, ci, i, len)) o.brk |= bit;
      else if ((ci & CI_WS) != prev_ws) o.fl |= TILE_EXOTIC;
      if (cp >= 0x80u) {
        if (!(ci & (CI_PUNCT | CI_WS))) o.wc |= bit;
        if (ci & CI_ALPHA) o.al |= bit;
        o.fl |= ci & CI_COMPLEX;
      }
      prev_ws = ci & CI_WS;
      wb_advance(st, cls);
      i += len;
    }
  }
}
TB_HD uint32_t tile_fix_dirty(const Classifier& C, const TextRange& R, Tile& T) {
  DirtyFix f;
  tile_dirty_bits(C, R, T.p0, T.dirty, f);
  T.brk = (T.brk & ~f.span) | f.brk;
  T.m[TL_WC] |= f.wc;
  T.al |= f.al;
  return f.fl;
}

// ---- open-fragment summaries (segments that cross stretch boundaries) ------------------------------
// nbytes, and in `meta`: character count (bits 0..26), "first byte can start a stop word" (27),
// "holds a non-ASCII byte" (28), "holds a word character" (29), "holds an alphabetic character" (30).
#define FR_SWF (1u << 27)
#define FR_NA (1u << 28)
#define FR_WC (1u << 29)
#define FR_AL (1u << 30)
#define FR_CNT 0x07FFFFFFu
struct Frag { uint32_t nbytes, meta; };
TB_HD Frag frag_cat(Frag x, Frag y) {  // x followed by y
  Frag r;
  r.nbytes = x.nbytes + y.nbytes;
  uint32_t cnt = (x.meta & FR_CNT) + (y.meta & FR_CNT);
  if (cnt > FR_CNT) cnt = FR_CNT;
  r.meta = cnt | ((x.meta | y.meta) & (FR_NA | FR_WC | FR_AL)) | ((x.nbytes ? x.meta : y.meta) & FR_SWF);
  return r;
}
TB_HD Frag frag_of(const Tile& T, uint32_t range, uint32_t first_hi_swf) {  // summary of the bytes in `range`
  Frag f;
  f.nbytes = (uint32_t)tb_popc32(T.valid & range);
  f.meta = (uint32_t)tb_popc32(T.valid & ~T.h[TH_CONT] & range);
  if (T.m[TL_NA] & range) f.meta |= FR_NA;
  if (T.m[TL_WC] & range) f.meta |= FR_WC;
  if (T.al & range) f.meta |= FR_AL;
  if (first_hi_swf) f.meta |= FR_SWF;
  return f;
}
