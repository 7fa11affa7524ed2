// tb_wordmachine.cuh -- per-lane UAX #29 word segmentation + TextBlaster word filtering.
//
// Replaces (per document range): icu WordSegmenter::new_auto().segment_str (rule-based part) and
// the segment filter of split_into_words, /root/reference src/utils/text.rs:103-181.
//
// Parallel decomposition: a document range [rs, re) is cut into fixed stretches; the lane that
// owns stretch [p0, p1) emits exactly the segments whose first byte lies in [p0, p1).  To do that
// it (1) re-synchronises the rule state by scanning back to the nearest code point after which the
// WB rules need no further left context, (2) walks forward, and (3) keeps walking past p1 until
// the segment it owns is closed.  The rule state is (prev literal class, last two non-ignored
// classes, parity of the trailing Regional_Indicator run) -- see the rule list in SURVEY.md App. B.
#pragma once
#include "tb_common.cuh"

#define WM_NEED_ALPHA 1
#define WM_NEED_HASH 2
#define WM_NEED_LOWKEY 4

#define WB_NLMASK (BIT(WB_CR) | BIT(WB_LF) | BIT(WB_NEWLINE))
#define WB_EFZMASK (BIT(WB_EXTEND) | BIT(WB_FORMAT) | BIT(WB_ZWJ))
#define WB_AHLMASK (BIT(WB_ALETTER) | BIT(WB_HEBREW))
#define WB_MLQMASK (BIT(WB_MIDLETTER) | BIT(WB_MIDNUMLET) | BIT(WB_SQ))   // MidLetter | MidNumLetQ
#define WB_MNQMASK (BIT(WB_MIDNUM) | BIT(WB_MIDNUMLET) | BIT(WB_SQ))      // MidNum | MidNumLetQ
// classes after which a forward walk needs no older context (not ignorable, not Mid*/quote, not RI)
#define WB_NOSYNCMASK (WB_EFZMASK | WB_MLQMASK | BIT(WB_MIDNUM) | BIT(WB_DQ) | BIT(WB_RI))

struct LowKey {  // lower-cased word: first 16 bytes packed, byte length, FNV-style fold of bytes past 16
  uint64_t lo, hi, ext;
  uint32_t len;
  uint32_t sigma;  // word contains U+03A3: final-sigma rule needs the slow path
  TB_HD void init() { lo = 0; hi = 0; ext = 0; len = 0; sigma = 0; }
  TB_HD void push(uint32_t b) {
    if (len < 8) lo |= (uint64_t)b << (8 * len);
    else if (len < 16) hi |= (uint64_t)b << (8 * (len - 8));
    else ext = (ext ^ (uint64_t)(b + 1u)) * 0x100000001B3ull;
    len++;
  }
  TB_HD void push_cp(uint32_t cp) {
    if (cp < 0x80u) push(cp);
    else if (cp < 0x800u) { push(0xC0u | (cp >> 6)); push(0x80u | (cp & 0x3Fu)); }
    else if (cp < 0x10000u) { push(0xE0u | (cp >> 12)); push(0x80u | ((cp >> 6) & 0x3Fu)); push(0x80u | (cp & 0x3Fu)); }
    else { push(0xF0u | (cp >> 18)); push(0x80u | ((cp >> 12) & 0x3Fu)); push(0x80u | ((cp >> 6) & 0x3Fu)); push(0x80u | (cp & 0x3Fu)); }
  }
};

struct WordInfo {
  int64_t start, end;  // trimmed byte range
  uint32_t nchars;
  uint32_t alpha;
  Hash2 hash;          // over the trimmed bytes (WM_NEED_HASH)
  LowKey key;          // to_lowercase() of the word (WM_NEED_LOWKEY)
};

struct WbState {
  uint32_t prev, l1, l2, ri_odd;
  TB_HD void reset() { prev = WB_SOT; l1 = WB_SOT; l2 = WB_SOT; ri_odd = 0; }
};

TB_HD uint32_t wb_join_simple(uint32_t a) {  // classes b with  a x b  unconditionally (WB5,7a,8,9,10,13,13a,13b)
  switch (a) {
    case WB_ALETTER: return WB_AHLMASK | BIT(WB_NUMERIC) | BIT(WB_EXTENDNUMLET);
    case WB_HEBREW: return WB_AHLMASK | BIT(WB_NUMERIC) | BIT(WB_EXTENDNUMLET) | BIT(WB_SQ);
    case WB_NUMERIC: return BIT(WB_NUMERIC) | WB_AHLMASK | BIT(WB_EXTENDNUMLET);
    case WB_KATAKANA: return BIT(WB_KATAKANA) | BIT(WB_EXTENDNUMLET);
    case WB_EXTENDNUMLET: return BIT(WB_EXTENDNUMLET) | WB_AHLMASK | BIT(WB_NUMERIC) | BIT(WB_KATAKANA);
    default: return 0;
  }
}

// class of the next non-ignored code point at or after `pos` (WB4 skipping), WB_SOT if none
TB_HD uint32_t wb_lookahead(const Classifier& C, const TextRange& R, int64_t pos) {
  while (pos < R.re) {
    int len; uint32_t cp;
    uint32_t c = CI_WB(C.at(R.t, pos, &len, &cp));
    if (!(BIT(c) & WB_EFZMASK)) return c;
    pos += len;
  }
  return WB_SOT;
}

// Is there a word boundary before the code point (class cls, info ci) at `i`, whose encoded length is len?
TB_HD bool wb_break_before(const Classifier& C, const TextRange& R, const WbState& s, uint32_t cls, uint32_t ci,
                           int64_t i, int len) {
  uint32_t prev = s.prev;
  if (prev == WB_SOT) return true;                                        // WB1
  if (prev == WB_CR && cls == WB_LF) return false;                        // WB3
  if ((BIT(prev) | BIT(cls)) & WB_NLMASK) return true;                    // WB3a, WB3b
  if (prev == WB_ZWJ && (ci & CI_EXTPICT)) return false;                  // WB3c
  if (prev == WB_WSEG && cls == WB_WSEG) return false;                    // WB3d
  if (BIT(cls) & WB_EFZMASK) return false;                                // WB4
  uint32_t a = s.l1, z = s.l2;
  if ((wb_join_simple(a) >> cls) & 1u) return false;
  uint32_t ma = a < 31 ? BIT(a) : 0u, mz = z < 31 ? BIT(z) : 0u, mb = BIT(cls);
  if ((ma & WB_AHLMASK) && (mb & WB_MLQMASK)) {                           // WB6
    uint32_t r = wb_lookahead(C, R, i + len);
    return !(r < 31 && (BIT(r) & WB_AHLMASK));
  }
  if (a == WB_HEBREW && cls == WB_DQ) return wb_lookahead(C, R, i + len) != WB_HEBREW;   // WB7b
  if (a == WB_NUMERIC && (mb & WB_MNQMASK)) return wb_lookahead(C, R, i + len) != WB_NUMERIC;   // WB12
  if ((mz & WB_AHLMASK) && (ma & WB_MLQMASK) && (mb & WB_AHLMASK)) return false;   // WB7
  if (z == WB_HEBREW && a == WB_DQ && cls == WB_HEBREW) return false;     // WB7c
  if (z == WB_NUMERIC && (ma & WB_MNQMASK) && cls == WB_NUMERIC) return false;   // WB11
  if (a == WB_RI && cls == WB_RI && s.ri_odd) return false;               // WB15, WB16
  return true;                                                            // WB999
}

TB_HD void wb_advance(WbState& s, uint32_t cls) {
  bool ignored = (BIT(cls) & WB_EFZMASK) && s.prev != WB_SOT && !(BIT(s.prev) & WB_NLMASK);
  if (!ignored) {
    s.ri_odd = (cls == WB_RI) ? ((s.l1 == WB_RI) ? (s.ri_odd ^ 1u) : 1u) : 0u;
    s.l2 = s.l1;
    s.l1 = cls;
  }
  s.prev = cls;
}

// Establish the rule state in front of the code point at q (rs <= q < re) by scanning back to a
// synchronisation point and replaying.
TB_HD void wb_sync(const Classifier& C, const TextRange& R, int64_t q, WbState& s) {
  s.reset();
  if (q <= R.rs) return;
  int64_t j = q;
  for (;;) {
    j = C.prev_lead(R.t, j, R.rs);
    int len; uint32_t cp;
    uint32_t cls = CI_WB(C.at(R.t, j, &len, &cp));
    if (j <= R.rs) break;  // replay from the true start of text
    if (!(BIT(cls) & WB_NOSYNCMASK)) break;
  }
  while (j < q) {
    int len; uint32_t cp;
    uint32_t cls = CI_WB(C.at(R.t, j, &len, &cp));
    wb_advance(s, cls);
    j += len;
  }
}

// ---- lower-casing one code point into a LowKey (str::to_lowercase semantics, final sigma deferred)
TB_HD void lowkey_push(const Classifier& C, LowKey& k, uint32_t cp, uint32_t ci) {
  if (cp < 0x80u) { k.push((cp - 'A') < 26u ? cp + 32u : cp); return; }
  if (!(ci & CI_LOWERS)) { k.push_cp(cp); return; }
  if (cp == 0x130u) { k.push('i'); k.push(0xCCu); k.push(0x87u); return; }  // U+0130 -> "i" + U+0307
  if (cp == 0x3A3u) { k.sigma = 1; k.push_cp(0x3C3u); return; }
  k.push_cp(tb_simple_lower(C.u, cp));
}

// Exact to_lowercase() of the word [start, end) including the Final_Sigma condition
// (alloc::str::to_lowercase / map_uppercase_sigma).  Slow path: only words containing U+03A3.
TB_HD_NOINLINE void lowkey_slow(const Classifier& C, const uint8_t* t, int64_t start, int64_t end, LowKey& k) {
  k.init();
  int64_t i = start;
  while (i < end) {
    int len; uint32_t cp;
    uint32_t ci = C.at(t, i, &len, &cp);
    if (cp == 0x3A3u) {
      bool before = false, after = false;
      int64_t j = i;
      while (j > start) {
        j = C.prev_lead(t, j, start);
        int l2; uint32_t c2; C.at(t, j, &l2, &c2);
        if (tb_in_ranges(C.u.caseign, C.u.n_caseign, c2)) continue;
        before = tb_in_ranges(C.u.cased, C.u.n_cased, c2);
        break;
      }
      j = i + len;
      while (j < end) {
        int l2; uint32_t c2; C.at(t, j, &l2, &c2);
        j += l2;
        if (tb_in_ranges(C.u.caseign, C.u.n_caseign, c2)) continue;
        after = tb_in_ranges(C.u.cased, C.u.n_cased, c2);
        break;
      }
      k.push_cp((before && !after) ? 0x3C2u : 0x3C3u);
    } else {
      lowkey_push(C, k, cp, ci);
    }
    i += len;
  }
  k.sigma = 0;
}

// Segment accumulator: trims White_Space at both ends and tracks the word features.
template <int MODE>
struct SegAcc {
  bool open, seen, in_ws, wordchar;
  uint32_t alpha;
  int64_t start, end;
  uint32_t nchars_run, nchars;
  Hash2 h_run, h_snap;
  LowKey k_run, k_snap;
  TB_HD void begin() { open = true; seen = false; in_ws = false; wordchar = false; alpha = 0; nchars_run = 0; nchars = 0; start = end = 0; }
  TB_HD void add(const Classifier& C, const uint8_t* t, int64_t i, int len, uint32_t cp, uint32_t ci) {
    bool ws = (ci & CI_WS) != 0;
    if (ws) {
      if (!seen) return;  // leading whitespace of the segment
      if (!in_ws) {       // possible trailing whitespace: remember the state before it
        in_ws = true;
        if (MODE & WM_NEED_HASH) h_snap = h_run;
        if (MODE & WM_NEED_LOWKEY) k_snap = k_run;
        nchars = nchars_run;
      }
    } else {
      if (!seen) {
        seen = true; start = i; nchars_run = 0;
        if (MODE & WM_NEED_HASH) h_run.init();
        if (MODE & WM_NEED_LOWKEY) k_run.init();
      }
      in_ws = false;
      end = i + len;
      if (!(ci & CI_PUNCT)) wordchar = true;
      if (MODE & WM_NEED_ALPHA) alpha |= (ci & CI_ALPHA);
    }
    nchars_run++;
    if (MODE & WM_NEED_HASH) { const uint64_t pw1 = TB_LDG(C.u.powtab + 1); for (int k = 0; k < len; k++) h_run.push(tb_byte(t, i + k), pw1); }
    if (MODE & WM_NEED_LOWKEY) lowkey_push(C, k_run, cp, ci);
  }
  // returns true and fills w if the closed segment is a word (text.rs:139-157)
  TB_HD bool finish(WordInfo& w) {
    open = false;
    if (!seen || !wordchar) return false;
    w.start = start; w.end = end;
    w.nchars = in_ws ? nchars : nchars_run;
    w.alpha = alpha ? 1u : 0u;
    if (MODE & WM_NEED_HASH) w.hash = in_ws ? h_snap : h_run;
    if (MODE & WM_NEED_LOWKEY) w.key = in_ws ? k_snap : k_run;
    return true;
  }
};

// Walk the stretch [p0, p1) of range R; sink.on_word(w) for each word whose segment starts in it.
// Returns the OR of the cinfo words of every code point starting in [p0, p1) (document flags).
template <int MODE, class Sink>
TB_HD uint32_t tb_walk_words(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1, Sink& sink) {
  const uint8_t* t = R.t;
  int64_t q = p0;
  if (q > R.rs) while (q < R.re && tb_is_cont(tb_byte(t, q))) q++;
  if (q >= p1 || q >= R.re) return 0;
  WbState s;
  wb_sync(C, R, q, s);
  SegAcc<MODE> acc;
  acc.open = false;
  uint32_t seen_flags = 0;
  int64_t i = q;
  while (i < R.re) {
    int len; uint32_t cp;
    uint32_t ci = C.at(t, i, &len, &cp);
    uint32_t cls = CI_WB(ci);
    if (wb_break_before(C, R, s, cls, ci, i, len)) {
      if (acc.open) { WordInfo w; if (acc.finish(w)) sink.on_word(w); }
      if (i >= p1) break;
      acc.begin();
    } else if (!acc.open && i >= p1) {
      break;
    }
    if (i < p1) seen_flags |= ci;
    if (acc.open) acc.add(C, t, i, len, cp, ci);
    wb_advance(s, cls);
    i += len;
  }
  if (acc.open) { WordInfo w; if (acc.finish(w)) sink.on_word(w); }
  return seen_flags;
}
