// tb_sentmachine.cuh -- per-lane UAX #29 sentence segmentation, counting the non-empty trimmed
// pieces that split_into_sentences returns (/root/reference src/utils/text.rs:59-101; consumed by
// C4QualityFilter only through .len(), src/pipeline/filters/c4_filters.rs:261-262).
//
// Same ownership scheme as tb_wordmachine.cuh: the lane owning [p0, p1) counts the sentence
// segments that START in its stretch and contain at least one non-White_Space character.
#pragma once
#include "tb_common.cuh"

#define SB_PARASEPMASK (BIT(SB_SEP) | BIT(SB_CR) | BIT(SB_LF))
#define SB_SATERMMASK (BIT(SB_STERM) | BIT(SB_ATERM))
#define SB_EFMASK (BIT(SB_EXTEND) | BIT(SB_FORMAT))
// after one of these the rules need no older context
#define SB_NOSYNCMASK (SB_SATERMMASK | BIT(SB_CLOSE) | BIT(SB_SP) | SB_EFMASK)

struct SbState {
  uint32_t prev, l1;
  uint32_t st;  // 0 none; 1 ATerm; 2 ATerm Close+; 3 ATerm Close* Sp+; 4 STerm; 5 STerm Close+; 6 STerm Close* Sp+
  uint32_t before_letter;  // (Upper|Lower) immediately before the ATerm (SB7)
  TB_HD void reset() { prev = SB_SOT; l1 = SB_SOT; st = 0; before_letter = 0; }
};

// SB8 look-ahead: skip  not(OLetter|Upper|Lower|ParaSep|SATerm)  then require Lower
TB_HD bool sb_lower_ahead(const Classifier& C, const TextRange& R, int64_t pos) {
  const uint32_t stop = BIT(SB_OLETTER) | BIT(SB_UPPER) | BIT(SB_LOWER) | SB_PARASEPMASK | SB_SATERMMASK;
  while (pos < R.re) {
    int len; uint32_t cp;
    uint32_t c = CI_SB(C.at(R.t, pos, &len, &cp));
    if (BIT(c) & stop) return c == SB_LOWER;
    pos += len;
  }
  return false;
}

TB_HD bool sb_break_before(const Classifier& C, const TextRange& R, const SbState& s, uint32_t cls, int64_t i) {
  if (s.prev == SB_SOT) return true;                                   // SB1
  if (s.prev == SB_CR && cls == SB_LF) return false;                   // SB3
  if (BIT(s.prev) & SB_PARASEPMASK) return true;                       // SB4
  uint32_t mb = BIT(cls);
  if (mb & SB_EFMASK) return false;                                    // SB5
  uint32_t st = s.st;
  if (st == 0) return false;                                           // SB998
  if (st == 1 && cls == SB_NUMERIC) return false;                      // SB6
  if (st == 1 && s.before_letter && cls == SB_UPPER) return false;     // SB7
  if (st <= 3 && sb_lower_ahead(C, R, i)) return false;                // SB8
  if (mb & (BIT(SB_SCONTINUE) | SB_SATERMMASK)) return false;          // SB8a
  if ((st == 1 || st == 2 || st == 4 || st == 5) && (mb & (BIT(SB_CLOSE) | BIT(SB_SP) | SB_PARASEPMASK))) return false;  // SB9
  if (mb & (BIT(SB_SP) | SB_PARASEPMASK)) return false;                // SB10
  return true;                                                         // SB11
}

TB_HD void sb_advance(SbState& s, uint32_t cls) {
  bool ignored = (BIT(cls) & SB_EFMASK) && s.prev != SB_SOT && !(BIT(s.prev) & SB_PARASEPMASK);
  if (!ignored) {
    if (cls == SB_ATERM) { s.before_letter = (s.l1 == SB_UPPER || s.l1 == SB_LOWER); s.st = 1; }
    else if (cls == SB_STERM) s.st = 4;
    else if (cls == SB_CLOSE && (s.st == 1 || s.st == 2 || s.st == 4 || s.st == 5)) s.st = s.st < 4 ? 2 : 5;
    else if (cls == SB_SP && s.st != 0) s.st = s.st < 4 ? 3 : 6;
    else s.st = 0;
    s.l1 = cls;
  }
  s.prev = cls;
}

TB_HD void sb_sync(const Classifier& C, const TextRange& R, int64_t q, SbState& s) {
  s.reset();
  if (q <= R.rs) return;
  int64_t j = q;
  for (;;) {
    j = C.prev_lead(R.t, j, R.rs);
    int len; uint32_t cp;
    uint32_t cls = CI_SB(C.at(R.t, j, &len, &cp));
    if (j <= R.rs) break;
    if (!(BIT(cls) & SB_NOSYNCMASK)) break;
  }
  while (j < q) {
    int len; uint32_t cp;
    uint32_t cls = CI_SB(C.at(R.t, j, &len, &cp));
    sb_advance(s, cls);
    j += len;
  }
}

// Number of sentence pieces (non-empty after trim) whose segment starts in [p0, p1)  -- reference walk,
// one rule evaluation per code point.  Kept as the definition the event-driven version is tested against.
TB_HD uint32_t tb_count_sentences_walk(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1) {
  const uint8_t* t = R.t;
  int64_t q = p0;
  if (q > R.rs) while (q < R.re && tb_is_cont(tb_byte(t, q))) q++;
  if (q >= p1 || q >= R.re) return 0;
  SbState s;
  sb_sync(C, R, q, s);
  uint32_t count = 0;
  bool open = false;  // we own a segment in which no non-whitespace was seen yet
  int64_t i = q;
  while (i < R.re) {
    int len; uint32_t cp;
    uint32_t ci = C.at(t, i, &len, &cp);
    uint32_t cls = CI_SB(ci);
    if (sb_break_before(C, R, s, cls, i)) {
      if (i >= p1) break;
      open = true;
    } else if (!open && i >= p1) {
      break;
    }
    if (open && !(ci & CI_WS)) { count++; open = false; }
    sb_advance(s, cls);
    i += len;
  }
  return count;
}

// Event-driven count.  A sentence boundary can only arise (a) at the start of text, (b) right after a
// paragraph separator (SB4) or (c) at the end of a  SATerm Close* Sp*  run (SB11, unless SB6-SB10 veto).
// So instead of stepping the rule machine over every code point, a lane looks for the terminators and
// separators that start in its stretch and evaluates the rules only around them, with the same
// sb_sync / sb_break_before / sb_advance functions.  Every white-space character has class Sp or is a
// separator, so a segment opened by SB11 always begins with a non-white-space character; one opened
// by SB4 counts iff a non-white-space character follows before the next separator.
// phase 2: evaluate the rules at the event positions of the stretch that starts at p0 (bit k = byte p0 + k)
TB_HD uint32_t tb_count_sentences_events(const Classifier& C, const TextRange& R, int64_t p0, uint32_t ev) {
  const uint8_t* t = R.t;
  uint32_t count = 0;
  while (ev) {
#if defined(__CUDA_ARCH__)
    const uint32_t bit = (uint32_t)__ffs((int)ev) - 1u;
#else
    const uint32_t bit = (uint32_t)__builtin_ctz(ev);
#endif
    ev &= ev - 1;
    const int64_t i = p0 + bit;
    int len; uint32_t cp;
    const uint32_t ci = C.at(t, i, &len, &cp);
    const uint32_t cls = CI_SB(ci);
    if (BIT(cls) & SB_PARASEPMASK) {                                   // (b)
      int64_t j = i + len;
      if (!(cls == SB_CR && j < R.re && tb_byte(t, j) == '\n')) {     // SB3: CR x LF
        while (j < R.re) { int l2; uint32_t c2; uint32_t i2 = C.at(t, j, &l2, &c2); if (BIT(CI_SB(i2)) & SB_PARASEPMASK) break; if (!(i2 & CI_WS)) { count++; break; } j += l2; }
      }
    } else if (BIT(cls) & SB_SATERMMASK) {                             // (c)
      SbState s;
      sb_sync(C, R, i, s);
      sb_advance(s, cls);
      int64_t j = i + len;
      while (j < R.re && s.st != 0) {
        int l2; uint32_t c2;
        uint32_t k2 = CI_SB(C.at(t, j, &l2, &c2));
        if (sb_break_before(C, R, s, k2, j)) { count++; break; }
        if (BIT(k2) & SB_SATERMMASK) break;  // the next terminator evaluates what follows it
        sb_advance(s, k2);
        j += l2;
      }
    }
  }
  return count;
}
// Same events, but reporting WHERE each non-empty piece after the first begins: the position of its first
// non-white-space character (for SB11 breaks that is the break position itself).  Used by the
// sentence-wise line mode of C4 (split_paragraph = false, c4_filters.rs:152-155).  `mark(q)` is called once
// per piece; R is the trimmed text (text.rs:60).
template <class Mark>
TB_HD void tb_sentence_starts_events(const Classifier& C, const TextRange& R, int64_t p0, uint32_t ev, Mark& mark) {
  const uint8_t* t = R.t;
  while (ev) {
#if defined(__CUDA_ARCH__)
    const uint32_t bit = (uint32_t)__ffs((int)ev) - 1u;
#else
    const uint32_t bit = (uint32_t)__builtin_ctz(ev);
#endif
    ev &= ev - 1;
    const int64_t i = p0 + bit;
    int len; uint32_t cp;
    const uint32_t ci = C.at(t, i, &len, &cp);
    const uint32_t cls = CI_SB(ci);
    if (BIT(cls) & SB_PARASEPMASK) {                                   // SB4
      int64_t j = i + len;
      if (!(cls == SB_CR && j < R.re && tb_byte(t, j) == '\n')) {     // SB3: CR x LF
        while (j < R.re) { int l2; uint32_t c2; uint32_t i2 = C.at(t, j, &l2, &c2); if (BIT(CI_SB(i2)) & SB_PARASEPMASK) break; if (!(i2 & CI_WS)) { mark(j); break; } j += l2; }
      }
    } else if (BIT(cls) & SB_SATERMMASK) {                             // SB11
      SbState s;
      sb_sync(C, R, i, s);
      sb_advance(s, cls);
      int64_t j = i + len;
      while (j < R.re && s.st != 0) {
        int l2; uint32_t c2;
        uint32_t k2 = CI_SB(C.at(t, j, &l2, &c2));
        if (sb_break_before(C, R, s, k2, j)) { mark(j); break; }
        if (BIT(k2) & SB_SATERMMASK) break;
        sb_advance(s, k2);
        j += l2;
      }
    }
  }
}
// (a): the segment that starts the text counts iff a non-white-space character comes before the first separator
TB_HD uint32_t tb_count_sentences_start(const Classifier& C, const TextRange& R) {
  const uint8_t* t = R.t;
  int64_t j = R.rs;
  while (j < R.re) { int len; uint32_t cp; uint32_t ci = C.at(t, j, &len, &cp); if (BIT(CI_SB(ci)) & SB_PARASEPMASK) break; if (!(ci & CI_WS)) return 1u; j += len; }
  return 0u;
}
TB_HD uint32_t tb_count_sentences(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1) {
  const uint8_t* t = R.t;
  uint32_t count = 0;
  if (p0 <= R.rs && R.rs < p1) count += tb_count_sentences_start(C, R);
  // phase 1 (converged): which bytes of the stretch can matter -- ASCII terminators / separators and
  // the lead bytes of non-ASCII characters; phase 2: evaluate the rules at those positions only
  uint32_t ev = 0;
  for (int64_t j = p0 < R.rs ? R.rs : p0; j < p1; j++) {
    uint32_t b = tb_byte(t, j);
    bool hit = (b == '.') | (b == '!') | (b == '?') | (b == '\n') | (b == '\r') | (b >= 0xC0u);
    ev |= (hit ? 1u : 0u) << (uint32_t)(j - p0);
  }
  return count + tb_count_sentences_events(C, R, p0, ev);
}
