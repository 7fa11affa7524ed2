// tb_lanes.cuh -- the per-lane (per 32-byte stretch) work of every filter pass.  No warp
// collectives in here: kernels (tb_kernels.cu) combine lane results with shuffles / ballots; the
// host emulation harness (tests/host_emul, test-only) combines them serially.
#pragma once
#include "tb_common.cuh"
#include "tb_sentmachine.cuh"
#include "tb_fastwords.cuh"
#include "tb_wordmachine.cuh"

#define TB_STRETCH 32
#define TB_TILE (32 * TB_STRETCH)

// ---------------------------------------------------------------------------------------------
// Stop words (GopherQualityFilter, gopher_quality.rs:184-187): exact lookup of the lower-cased
// word.  Key = first 16 bytes + byte length (+ FNV-1a of the bytes beyond 16).
// ---------------------------------------------------------------------------------------------
struct SwEntry { uint64_t lo, hi, ext; uint32_t len, used; };
struct SwTable {
  const SwEntry* e; uint32_t mask, maxlen;
  uint32_t lenmask;    // bit min(len, 31) set for every entry length: cheap reject before building a key
  uint32_t firstmask;  // bit (first byte | 0x20) - 'a' for letters, bit 26 for anything else
  uint32_t secondmask; // the same for the entries' second bytes; all ones if some entry is a single byte (no filter then)
};
TB_HD uint32_t sw_first_bit(uint32_t b) { uint32_t f = (b | 0x20u) - 'a'; return f < 26u ? (1u << f) : (1u << 26); }

TB_HD uint64_t sw_slot_hash(uint64_t lo, uint64_t hi, uint64_t ext, uint32_t len) {
  uint64_t x = lo ^ (hi * 0x9E3779B97F4A7C15ull) ^ (ext * 0xC2B2AE3D27D4EB4Full) ^ len;
  uint32_t h = ((uint32_t)x ^ (uint32_t)(x >> 32)) * 0x9E3779B1u;
  return h ^ (h >> 15);
}
TB_HD bool sw_lookup(const SwTable& sw, const LowKey& k) {
  if (k.len == 0 || k.len > sw.maxlen) return false;
  uint32_t s = (uint32_t)sw_slot_hash(k.lo, k.hi, k.ext, k.len) & sw.mask;
  for (;;) {
    SwEntry e = sw.e[s];
    if (!e.used) return false;
    if (e.len == k.len && e.lo == k.lo && e.hi == k.hi && e.ext == k.ext) return true;
    s = (s + 1) & sw.mask;
  }
}

// ---------------------------------------------------------------------------------------------
// GopherQualityFilter pass (gopher_quality.rs:73-187)
// ---------------------------------------------------------------------------------------------
struct GqAcc {
  uint32_t n_words, n_alpha, n_stop, n_hash, n_ell, n_lines, n_bullet, n_ell_lines, flags;
  uint64_t sum_chars;
  TB_HD void zero() { n_words = n_alpha = n_stop = n_hash = n_ell = n_lines = n_bullet = n_ell_lines = flags = 0; sum_chars = 0; }
};
struct GqSink {
  const Classifier* C; const uint8_t* t; const SwTable* sw; GqAcc* a;
  TB_HD void on_word(WordInfo& w) {
    a->n_words++; a->sum_chars += w.nchars; a->n_alpha += w.alpha;
    if (w.key.sigma) lowkey_slow(*C, t, w.start, w.end, w.key);
    if (sw_lookup(*sw, w.key)) a->n_stop++;
  }
};

// line.trim_end().ends_with("...") || ends_with("…") for the line that ends at e (position of its
// '\n', or re).  gopher_quality.rs:152-158
TB_HD_NOINLINE bool gq_line_ends_with_ellipsis(const Classifier& C, const TextRange& R, int64_t e) {
  const uint8_t* t = R.t;
  int64_t j = e;
  while (j > R.rs) {
    int64_t k = C.prev_lead(t, j, R.rs);
    uint32_t b = tb_byte(t, k);
    if (b == '\n') break;
    int len; uint32_t cp;
    if (!(C.at(t, k, &len, &cp) & CI_WS)) break;
    j = k;
  }
  if (j - 3 < R.rs) return false;
  uint32_t a = tb_byte(t, j - 3), b = tb_byte(t, j - 2), c = tb_byte(t, j - 1);
  return (a == '.' && b == '.' && c == '.') || (a == 0xE2 && b == 0x80 && c == 0xA6);
}
// line.trim_start().starts_with('•') || starts_with('-') for the line starting at s.  :134-141
TB_HD_NOINLINE bool gq_line_is_bullet(const Classifier& C, const TextRange& R, int64_t s) {
  const uint8_t* t = R.t;
  int64_t j = s;
  while (j < R.re) {
    uint32_t b = tb_byte(t, j);
    if (b == '\n') return false;
    int len; uint32_t cp;
    if (!(C.at(t, j, &len, &cp) & CI_WS)) return b == '-' || cp == 0x2022u;
    j += len;
  }
  return false;
}

TB_HD void gq_lane(const Classifier& C, const TextRange& R, const SwTable& sw, int64_t p0, int64_t p1, GqAcc& acc) {
  const uint8_t* t = R.t;
  GqSink sink{&C, t, &sw, &acc};
  acc.flags |= tb_walk_words_fast<WM_NEED_ALPHA | WM_NEED_LOWKEY>(C, R, p0, p1, sink) & CI_COMPLEX;
  uint32_t prev = p0 > R.rs ? tb_byte(t, p0 - 1) : 0x100u;
  for (int64_t i = p0; i < p1; i++) {
    uint32_t b = tb_byte(t, i);
    if (b == '#') acc.n_hash++;                                        // :104
    else if (b == '.') {                                               // :115 matches("...")
      if (prev != '.') {
        int64_t j = i + 1;
        while (j < R.re && tb_byte(t, j) == '.') j++;
        acc.n_ell += (uint32_t)((j - i) / 3);
      }
    } else if (b == 0xE2) {                                            // :115 matches("…")
      if (i + 2 < R.re && tb_byte(t, i + 1) == 0x80 && tb_byte(t, i + 2) == 0xA6) acc.n_ell++;
    }
    if (prev == '\n' || prev == 0x100u) {                              // a line starts here (:126)
      acc.n_lines++;
      if (gq_line_is_bullet(C, R, i)) acc.n_bullet++;
    }
    if (b == '\n') { if (gq_line_ends_with_ellipsis(C, R, i)) acc.n_ell_lines++; }
    else if (i == R.re - 1) { if (gq_line_ends_with_ellipsis(C, R, R.re)) acc.n_ell_lines++; }
    prev = b;
  }
}

// ---------------------------------------------------------------------------------------------
// GopherRepetitionFilter scan pass (gopher_rep.rs:57-58,80-84,120-124,159)
// ---------------------------------------------------------------------------------------------
#define TB_LANE_MAXW 23  // a 32-byte stretch starts at most 22 words (two per three bytes; +1 slack)
struct GrLane {
  uint32_t nw;
  uint64_t wh[TB_LANE_MAXW];
  uint32_t wl[TB_LANE_MAXW];
  uint64_t ws[TB_LANE_MAXW];
  // '\n'-delimited pieces of the stretch: head (before the first '\n'), complete inner pieces,
  // tail (after the last '\n').  Without any '\n' the whole stretch is `tail`.
  uint32_t has_nl, first_nl;     // first_nl: offset of the first '\n'
  Hash2 head; uint32_t head_len;
  Hash2 tail; uint32_t tail_len;
  uint32_t np;
  uint64_t ph[16]; uint32_t pl[16]; uint32_t pend[16];  // inner pieces: hash, length, end offset
  uint32_t nchars, flags;
};
struct GrSink {
  GrLane* o;
  const uint8_t* t;
  TB_HD void on_word(WordInfo& w) {
    if (o->nw < TB_LANE_MAXW) {
      o->wh[o->nw] = w.hash.pack(); o->wl[o->nw] = (uint32_t)(w.end - w.start);
      uint64_t sig = 0;
      for (int64_t i = w.end - w.start > 16 ? w.end - 16 : w.start; i < w.end; i++) sig = tb_sig_push(sig, tb_byte(t, i));
      o->ws[o->nw] = sig;
    }
    o->nw++;
  }
};
TB_HD void gr_lane(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1, GrLane& o) {
  const uint8_t* t = R.t;
  o.nw = 0;
  GrSink sink{&o, t};
  o.flags = tb_walk_words_fast<WM_NEED_HASH>(C, R, p0, p1, sink) & CI_COMPLEX;
  o.has_nl = 0; o.first_nl = 0; o.np = 0; o.nchars = 0;
  o.head.init(); o.head_len = 0;
  Hash2 cur; cur.init(); uint32_t cur_len = 0;
  for (int64_t i = p0; i < p1; i++) {
    uint32_t b = tb_byte(t, i);
    if (!tb_is_cont(b)) o.nchars++;
    if (b == '\n') {
      if (!o.has_nl) { o.has_nl = 1; o.first_nl = (uint32_t)(i - p0); o.head = cur; o.head_len = cur_len; }
      else if (cur_len) { o.ph[o.np] = cur.pack(); o.pl[o.np] = cur_len; o.pend[o.np] = (uint32_t)(i - p0); o.np++; }
      cur.init(); cur_len = 0;
    } else { cur.push(b, TB_LDG(C.u.powtab + 1)); cur_len++; }
  }
  o.tail = cur; o.tail_len = cur_len;
}
// was the piece starting at s (s > ts) preceded by a run of >= 2 newlines? (paragraph_exp \n{2,})
TB_HD uint32_t gr_gap2(const TextRange& R, int64_t s) {
  return (s - 2 >= R.rs && tb_byte(R.t, s - 1) == '\n' && tb_byte(R.t, s - 2) == '\n') ? 1u : 0u;
}

// ---------------------------------------------------------------------------------------------
// C4QualityFilter (c4_filters.rs:147-295)
// ---------------------------------------------------------------------------------------------
// case-folded compare against an ASCII lower-case pattern, as `s.to_lowercase().contains(pat)`
// sees it: A-Z fold to a-z and U+212A KELVIN SIGN lower-cases to 'k'; no other character
// lower-cases to an ASCII letter (U+0130 gives "i" + U+0307, which can never complete a pattern).
TB_HD_NOINLINE bool c4_fold_match(const uint8_t* t, int64_t i, int64_t re, const char* pat, int n) {
  int64_t j = i;
  for (int k = 0; k < n; k++) {
    if (j >= re) return false;
    uint32_t b = tb_byte(t, j), c = (uint32_t)(unsigned char)pat[k];
    uint32_t fb = (b - 'A') < 26u ? b + 32u : b;
    if (fb == c) { j++; continue; }
    if (c == 'k' && b == 0xE2 && j + 2 < re && tb_byte(t, j + 1) == 0x84 && tb_byte(t, j + 2) == 0xAA) { j += 3; continue; }
    return false;
  }
  return true;
}

// CITATION_REGEX \[\d+(?:,\s*\d+)*\] anchored at the '[' at i; returns the end or -1.  The
// reference applies it per line, so '\n' never matches \s here.  c4_filters.rs:33,195-199
TB_HD_NOINLINE int64_t c4_citation_end(const Classifier& C, const TextRange& R, int64_t i) {
  const uint8_t* t = R.t;
  int64_t j = i + 1;
  int len; uint32_t cp;
  int nd = 0;
  while (j < R.re && (C.at(t, j, &len, &cp) & CI_ND)) { j += len; nd++; }
  if (!nd) return -1;
  for (;;) {
    if (j < R.re && tb_byte(t, j) == ',') {
      int64_t k = j + 1;
      while (k < R.re && tb_byte(t, k) != '\n' && (C.at(t, k, &len, &cp) & CI_WS)) k += len;
      int m = 0;
      while (k < R.re && (C.at(t, k, &len, &cp) & CI_ND)) { k += len; m++; }
      if (m) { j = k; continue; }
    }
    break;
  }
  if (j < R.re && tb_byte(t, j) == ']') return j + 1;
  return -1;
}

#define C4F_LOREM 1u
#define C4F_CURLY 2u
struct C4aLane {
  uint32_t nkept;
  uint8_t bytes[TB_STRETCH + 4];
  uint32_t n_nl;    // '\n' bytes in [p0, p1)
  uint32_t flags;   // C4F_*
  uint64_t kmask;   // kept bytes: bit i = byte p0 + i (a character that begins in the stretch can end up to 3 bytes behind it)
};
// Pass A: original text -> "processed lines" (each line trimmed, citations removed), still joined
// by the original '\n' (the one that ends the document is dropped).  c4_filters.rs:150-151,193-199
TB_HD void c4a_lane(const Classifier& C, const TextRange& R, int64_t p0, int64_t p1, bool remove_citations,
                    bool want_lorem, C4aLane& o) {
  const uint8_t* t = R.t;
  o.nkept = 0; o.n_nl = 0; o.flags = 0; o.kmask = 0;
  for (int64_t i = p0; i < p1; i++) {
    uint32_t b = tb_byte(t, i);
    if (b == '\n') o.n_nl++;
    else if (b == '{' || b == '}') o.flags |= C4F_CURLY;
    else if (want_lorem && (b | 0x20u) == 'l' && c4_fold_match(t, i, R.re, "lorem ipsum", 11)) o.flags |= C4F_LOREM;
  }
  int64_t q = p0;
  if (q > R.rs) while (q < R.re && tb_is_cont(tb_byte(t, q))) q++;
  if (q >= p1 || q >= R.re) return;
  int len; uint32_t cp;
  uint32_t ci = C.at(t, q, &len, &cp);
  // state in front of q: inside a citation?  has the line already shown a non-whitespace char?
  int64_t cit_end = -1;
  if (remove_citations && q > R.rs && cp != '\n' && ((ci & (CI_ND | CI_WS)) || cp == ',' || cp == ']')) {
    int64_t j = q;
    while (j > R.rs) {
      int64_t k = C.prev_lead(t, j, R.rs);
      int l2; uint32_t c2;
      uint32_t i2 = C.at(t, k, &l2, &c2);
      if (c2 == '[') { int64_t e = c4_citation_end(C, R, k); if (e > q) cit_end = e; break; }
      if (c2 == '\n' || !((i2 & (CI_ND | CI_WS)) || c2 == ',')) break;
      j = k;
    }
  }
  bool seen_nonws = false;
  if ((ci & CI_WS) && cp != '\n') {
    int64_t j = q;
    while (j > R.rs) {
      int64_t k = C.prev_lead(t, j, R.rs);
      int l2; uint32_t c2;
      uint32_t i2 = C.at(t, k, &l2, &c2);
      if (c2 == '\n') break;
      if (!(i2 & CI_WS)) { seen_nonws = true; break; }
      j = k;
    }
  }
  bool run_decided = false, run_keep = false;
  int64_t i = q;
  while (i < p1) {
    ci = C.at(t, i, &len, &cp);
    bool keep;
    if (cp == '\n') {
      keep = (i != R.re - 1);
      seen_nonws = false; run_decided = false;
    } else if (ci & CI_WS) {
      if (!run_decided) {
        run_decided = true; run_keep = false;
        if (seen_nonws) {
          int64_t j = i + len;
          while (j < R.re) {
            int l2; uint32_t c2;
            uint32_t i2 = C.at(t, j, &l2, &c2);
            if (c2 == '\n') break;
            if (!(i2 & CI_WS)) { run_keep = true; break; }
            j += l2;
          }
        }
      }
      keep = run_keep && i >= cit_end;
    } else {
      seen_nonws = true; run_decided = false;
      if (i < cit_end) keep = false;
      else if (remove_citations && cp == '[') {
        int64_t e = c4_citation_end(C, R, i);
        if (e > 0) { cit_end = e; keep = false; } else keep = true;
      } else keep = true;
    }
    if (keep) for (int k = 0; k < len; k++) { o.bytes[o.nkept++] = (uint8_t)tb_byte(t, i + k); o.kmask |= 1ull << (uint32_t)(i + k - p0); }
    i += len;
  }
}

// Pass B flags per processed line
#define C4L_TERM 1u     // last char in END_PUNCTUATION (c4_filters.rs:20,220-223)
#define C4L_ELLIPSIS 2u // ends_with("...") (:224)
#define C4L_JS 4u       // contains "javascript" (:243)
#define C4L_POLICY 8u   // contains a POLICY_SUBSTRINGS entry (:24-31,248)

// flags for the line that ends at e (exclusive), looking only at its last bytes
TB_HD_NOINLINE uint32_t c4_line_end_flags(const Classifier& C, const TextRange& R, int64_t e) {
  const uint8_t* t = R.t;
  if (e <= R.rs || tb_byte(t, e - 1) == '\n') return 0;  // empty line
  int64_t k = C.prev_lead(t, e, R.rs);
  int len; uint32_t cp;
  C.at(t, k, &len, &cp);
  uint32_t f = 0;
  if (cp == '.' || cp == '!' || cp == '?' || cp == '"' || cp == '\'' || cp == 0x201Du) f |= C4L_TERM;
  if (e - 3 >= R.rs && tb_byte(t, e - 1) == '.' && tb_byte(t, e - 2) == '.' && tb_byte(t, e - 3) == '.') f |= C4L_ELLIPSIS;
  return f;
}
// javascript / policy substrings starting at byte i (lower-cased compare)
TB_HD_NOINLINE uint32_t c4_pattern_flags(const uint8_t* t, int64_t i, int64_t re, uint32_t b, bool js, bool policy) {
  uint32_t fb = b | 0x20u;
  // cheap reject: only j / t / p / c / u can start a pattern, and the next two bytes must continue it
  // ("jav", "ter", "pri", "coo", "use"); the exact (Kelvin-sign aware) compare runs on the few survivors
  const uint32_t d = fb - 'c';  // c, j, p, t, u as bits of a small mask
  if (d > (uint32_t)('u' - 'c') || !((0x62081u >> d) & 1u) || i + 2 >= re) return 0;
  uint32_t b1 = tb_byte(t, i + 1) | 0x20u, b2 = tb_byte(t, i + 2) | 0x20u;
  uint32_t tri = (fb << 16) | (b1 << 8) | b2;
  if (tri != 0x6A6176u && tri != 0x746572u && tri != 0x707269u && tri != 0x636F6Fu && tri != 0x757365u) return 0;
  if (js && fb == 'j') return c4_fold_match(t, i, re, "javascript", 10) ? C4L_JS : 0;
  if (!policy) return 0;
  switch (fb) {
    case 't': return c4_fold_match(t, i, re, "terms of use", 12) ? C4L_POLICY : 0;
    case 'p': return c4_fold_match(t, i, re, "privacy policy", 14) ? C4L_POLICY : 0;
    case 'c': return c4_fold_match(t, i, re, "cookie policy", 13) ? C4L_POLICY : 0;
    case 'u': return (c4_fold_match(t, i, re, "uses cookies", 12) || c4_fold_match(t, i, re, "use of cookies", 14) ||
                      c4_fold_match(t, i, re, "use cookies", 11)) ? C4L_POLICY : 0;
    default: return 0;
  }
}

// per-line decision, in the reference's order (c4_filters.rs:207-250). returns 0 keep, 1 too_long_word,
// 2 no_terminal_punc, 3 too_few_words, 4 javascript/policy (dropped, no counter)
struct C4Cfg {
  uint32_t remove_citations, filter_no_terminal_punct, filter_lorem_ipsum, filter_javascript, filter_curly_bracket, filter_policy;
  uint64_t min_num_sentences, min_words_per_line, max_word_length;
};
TB_HD int c4_line_decision(const C4Cfg& p, uint32_t wc, uint32_t maxc, uint32_t fl) {
  if (p.max_word_length > 0 && (uint64_t)maxc > p.max_word_length) return 1;
  if (p.filter_no_terminal_punct && (!(fl & C4L_TERM) || (fl & C4L_ELLIPSIS))) return 2;
  if (p.min_words_per_line > 0 && (uint64_t)wc < p.min_words_per_line) return 3;
  if (p.filter_javascript && (fl & C4L_JS)) return 4;
  if (p.filter_policy && (fl & C4L_POLICY)) return 4;
  return 0;
}
