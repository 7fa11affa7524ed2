// tb_fastwords.cuh -- bit-parallel fast path of the word pass for pure-ASCII stretches.
//
// For text without non-ASCII bytes (and without '\r') the UAX #29 word rules involve no ignorable
// characters, so every boundary decision depends on at most the two bytes before and the byte after
// it.  A lane classifies a 48-byte window (2 bytes of left context, its 32-byte stretch, 14 bytes of
// right context) into eight 48-bit class masks with one shared-memory table lookup per byte, derives
// the break mask with a dozen shifts and ANDs (WB3d, WB5-WB13b; WB3a/3b fall out because newline
// characters carry no class bit), and then only iterates over the segments that start in its
// stretch.  Anything else -- a non-ASCII byte or '\r' in the window, or a segment that does not
// close inside the window -- goes through the general state machine (tb_wordmachine.cuh), segment
// by segment, so the result is identical by construction.  Same ownership rule as the general
// path: a lane emits the words whose segment starts in [p0, p1).
#pragma once
#include "tb_wordmachine.cuh"

// per-ASCII-byte class flags, derived from the cinfo table
#define AF_A 0x01u    // ALetter
#define AF_N 0x02u    // Numeric
#define AF_MLQ 0x04u  // MidLetter | MidNumLet | Single_Quote
#define AF_MNQ 0x08u  // MidNum | MidNumLet | Single_Quote
#define AF_E 0x10u    // ExtendNumLet
#define AF_S 0x20u    // WSegSpace
#define AF_NL 0x40u   // LF | Newline (informational; '\r' never reaches the fast path)
#define AF_WC 0x80u   // neither PUNCTUATION nor White_Space (text.rs:141)

TB_HD uint32_t tb_ascii_flags(uint32_t ci) {
  uint32_t c = CI_WB(ci), f = 0;
  if (c == WB_ALETTER) f |= AF_A;
  if (c == WB_NUMERIC) f |= AF_N;
  if (c == WB_MIDLETTER || c == WB_MIDNUMLET || c == WB_SQ) f |= AF_MLQ;
  if (c == WB_MIDNUM || c == WB_MIDNUMLET || c == WB_SQ) f |= AF_MNQ;
  if (c == WB_EXTENDNUMLET) f |= AF_E;
  if (c == WB_WSEG) f |= AF_S;
  if (c == WB_LF || c == WB_NEWLINE) f |= AF_NL;
  if (!(ci & (CI_PUNCT | CI_WS))) f |= AF_WC;
  return f;
}
// classes that the fast path does not model must not occur among ASCII code points
TB_HD bool tb_ascii_class_ok(uint32_t ci) {
  uint32_t c = CI_WB(ci);
  return c == WB_OTHER || c == WB_CR || c == WB_LF || c == WB_NEWLINE || c == WB_ALETTER || c == WB_SQ || c == WB_DQ ||
         c == WB_MIDNUMLET || c == WB_MIDLETTER || c == WB_MIDNUM || c == WB_NUMERIC || c == WB_EXTENDNUMLET || c == WB_WSEG;
}

#define TB_FW_WIN 64
#define TB_FW_LCTX 2

TB_HD uint32_t tb_funnel_r(uint32_t lo, uint32_t hi, uint32_t bytes) {
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, bytes * 8u);
#else
  return (uint32_t)((((uint64_t)hi << 32) | lo) >> (bytes * 8u));
#endif
}
TB_HD int tb_ctz64(uint64_t v) {
#if defined(__CUDA_ARCH__)
  return __ffsll((long long)v) - 1;
#else
  return __builtin_ctzll(v);
#endif
}
TB_HD int tb_popc64(uint64_t v) {
#if defined(__CUDA_ARCH__)
  return __popcll(v);
#else
  return __builtin_popcountll(v);
#endif
}

struct AsciiWin {
  uint64_t A, WC, brk;
  uint64_t dirty;  // decisions that involve a non-ASCII byte, a '\r', or bytes beyond the window: not final
  int64_t base;    // byte position of bit 0
};

// Build the window for the stretch starting at p0 (64 bytes: 2 of left context, the 32-byte stretch, 30 of
// right context).  Bytes outside the range read as 0.  Decisions whose four-byte neighbourhood holds a
// non-ASCII byte or a '\r' are flagged dirty instead of failing the whole window.
TB_HD void tb_ascii_window(const Classifier& C, const TextRange& R, int64_t p0, AsciiWin& w) {
  const uint8_t* af = C.aflags();
  const int64_t base = p0 - TB_FW_LCTX;
  const uintptr_t addr = (uintptr_t)(R.t + base);
  const uint32_t mis = (uint32_t)(addr & 3u);
  const uint8_t* a0 = (const uint8_t*)(addr - mis);
  const int64_t q0 = base - (int64_t)mis;  // byte position of a0
  uint32_t raw[TB_FW_WIN / 4 + 1];
#pragma unroll
  for (int k = 0; k <= TB_FW_WIN / 4; k++) {
    int64_t q = q0 + 4 * k;
    uint32_t v = 0;
    if (q >= R.rs && q + 4 <= R.re) v = TB_LDG((const uint32_t*)(a0 + 4 * k));
    else if (q + 4 > R.rs && q < R.re) {
      for (int r = 0; r < 4; r++) { int64_t pos = q + r; if (pos >= R.rs && pos < R.re) v |= tb_byte(R.t, pos) << (8 * r); }
    }
    raw[k] = v;
  }
  uint64_t mA = 0, mN = 0, mMLQ = 0, mMNQ = 0, mE = 0, mS = 0, mWC = 0, mNA = 0;
#pragma unroll
  for (int k = 0; k < TB_FW_WIN / 4; k++) {
    uint32_t x = mis ? tb_funnel_r(raw[k], raw[k + 1], mis) : raw[k];
    uint32_t hi = x & 0x80808080u;                              // non-ASCII bytes
    uint32_t z = (x & 0x7F7F7F7Fu) ^ 0x0D0D0D0Du;               // '\r' (WB3 CR x LF is left to the general machine)
    uint32_t cr = (((z + 0x7F7F7F7Fu) | z) & 0x80808080u) ^ 0x80808080u;
    uint32_t na = (hi | cr) >> 7;
    uint32_t fw = (uint32_t)af[x & 0x7Fu] | ((uint32_t)af[(x >> 8) & 0x7Fu] << 8) | ((uint32_t)af[(x >> 16) & 0x7Fu] << 16) |
                  ((uint32_t)af[(x >> 24) & 0x7Fu] << 24);
#define TB_FW_GATHER(mask, v) mask |= (uint64_t)((((v) & 0x01010101u) * 0x01020408u) >> 24) << (4 * k)
    TB_FW_GATHER(mA, fw); TB_FW_GATHER(mN, fw >> 1); TB_FW_GATHER(mMLQ, fw >> 2); TB_FW_GATHER(mMNQ, fw >> 3);
    TB_FW_GATHER(mE, fw >> 4); TB_FW_GATHER(mS, fw >> 5); TB_FW_GATHER(mWC, fw >> 7); TB_FW_GATHER(mNA, na);
#undef TB_FW_GATHER
  }
  const uint64_t AN = mA | mN, ANE = AN | mE;
  const uint64_t nb = ((AN << 1) & AN)                                                      // WB5, WB8, WB9, WB10
                      | ((mA << 1) & mMLQ & (mA >> 1)) | ((mA << 2) & (mMLQ << 1) & mA)     // WB6, WB7
                      | ((mN << 1) & mMNQ & (mN >> 1)) | ((mN << 2) & (mMNQ << 1) & mN)     // WB12, WB11
                      | ((ANE << 1) & mE) | ((mE << 1) & AN)                                // WB13a, WB13b
                      | ((mS << 1) & mS);                                                   // WB3d
  w.A = mA; w.WC = mWC; w.brk = ~nb; w.base = base;
  // a decision at bit j looks at bytes j-2 .. j+1: it is final only if none of them is non-ASCII / '\r' and
  // byte j+1 is inside the window (or beyond the end of the text, which ends inside the window)
  const int64_t left = R.re - base;
  const uint32_t lim = left < TB_FW_WIN ? (uint32_t)left : (uint32_t)(TB_FW_WIN - 2);
  w.dirty = mNA | (mNA << 1) | (mNA << 2) | (mNA >> 1) | ~((2ull << lim) - 1ull);
}

// Walk the stretch [p0, p1): segments whose boundary decisions are all clean come straight from the masks;
// dirty runs and segments that touch a dirty decision are handed to the general machine (tb_walk_words),
// always in text order.
template <int MODE, class Sink>
TB_HD uint32_t tb_walk_words_fast(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1, Sink& sink) {
  AsciiWin w;
  tb_ascii_window(C, R, p0, w);
  const uint32_t s0 = TB_FW_LCTX, s1 = (uint32_t)(p1 - w.base);  // own segment starts: bits [s0, s1)
  const uint64_t ownrange = ((1ull << s1) - 1ull) & ~((1ull << s0) - 1ull);
  const uint64_t dm = w.dirty & ownrange;
  // work items in text order: clean segment starts, and the first bit of every dirty run
  uint64_t todo = (w.brk & ~w.dirty & ownrange) | (dm & ~(dm << 1));
  const uint64_t stop = w.brk | w.dirty;
  uint32_t flags = 0;
  while (todo) {
    const uint32_t s = (uint32_t)tb_ctz64(todo);
    todo &= todo - 1;
    if ((w.dirty >> s) & 1ull) {  // dirty run [s, hi): the general machine emits the segments that start in it
      const uint32_t hi = s + (uint32_t)tb_ctz64(~(dm >> s));
      flags |= tb_walk_words<MODE>(C, R, w.base + s, w.base + hi, sink);
      continue;
    }
    const uint64_t rest = (stop >> (s + 1)) << (s + 1);  // never empty: the bits beyond the window are dirty
    const uint32_t e = (uint32_t)tb_ctz64(rest);
    if ((w.dirty >> e) & 1ull) {  // the segment runs into a dirty decision
      flags |= tb_walk_words<MODE>(C, R, w.base + s, w.base + s + 1, sink);
      continue;
    }
    const uint64_t seg = ((1ull << e) - 1ull) & ~((1ull << s) - 1ull);
    if (!(w.WC & seg)) continue;  // no word character: not a word (text.rs:139-157)
    WordInfo wi;
    wi.start = w.base + s; wi.end = w.base + e; wi.nchars = e - s;
    wi.alpha = (w.A & seg) ? 1u : 0u;
    if (MODE & WM_NEED_HASH) { const uint64_t pw1 = TB_LDG(C.u.powtab + 1); wi.hash.init(); for (int64_t i = wi.start; i < wi.end; i++) wi.hash.push(tb_byte(R.t, i), pw1); }
    if (MODE & WM_NEED_LOWKEY) {
      wi.key.init();
      const uint32_t len = e - s;
      uint32_t fb = (tb_byte(R.t, wi.start) | 0x20u) - 'a';
      fb = fb < 26u ? (1u << fb) : (1u << 26);
      if (len <= C.key_maxlen && ((C.key_lenmask >> (len < 31u ? len : 31u)) & 1u) && (C.key_firstmask & fb)) {
        for (int64_t i = wi.start; i < wi.end; i++) { uint32_t b = tb_byte(R.t, i); wi.key.push((b - 'A') < 26u ? b + 32u : b); }
      }  // else: cannot equal any stop word; an empty key never matches (sw_lookup)
    }
    sink.on_word(wi);
  }
  return flags;
}
