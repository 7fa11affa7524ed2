// tb_bodies.cuh -- warp-level bodies of the sm_100a kernels (one warp per document).
//
// Execution shape shared by every text pass: the document range is walked in 1 KiB tiles, each lane
// owning one 32-byte stretch of the tile (tb_lanes.cuh holds the per-lane work).  Lane results are
// combined with shuffles, reductions and (segmented) warp scans through the tb_simt.cuh primitives,
// so the same bodies run under the host emulation harness (tests/host_emul, test-only).
// No tensor cores (byte/integer work); every pass reads its text once from HBM.
#pragma once
#include "tb_kernels.cuh"
#include "tb_simt.cuh"

#define TB_BLOCK 128

// ---------------------------------------------------------------------------------------------
TB_D void body_init_ranges(const void* off, int off64, int64_t doc0, int64_t base, int64_t* rs, int64_t* re,
                           uint8_t* alive, int8_t* failed, int64_t d) {
  int64_t a, b;
  if (off64) { a = ((const int64_t*)off)[doc0 + d]; b = ((const int64_t*)off)[doc0 + d + 1]; }
  else { a = ((const int32_t*)off)[doc0 + d]; b = ((const int32_t*)off)[doc0 + d + 1]; }
  rs[d] = a - base; re[d] = b - base; alive[d] = 1; failed[d] = -1;
}

// UTF-8 well-formedness (RFC 3629 / Unicode Table 3-7), one warp per document
TB_D void body_validate_utf8(const uint8_t* t, const int64_t* rs, const int64_t* re, int32_t* bad, int64_t d, int lane) {
  int64_t s = rs[d], e = re[d];
  bool err = false;
  for (int64_t i = s + lane; i < e; i += 32) {
    uint32_t b = t[i];
    if (b < 0x80) continue;
    if ((b & 0xC0) == 0x80) {  // continuation: must be covered by a lead within the 3 previous bytes
      bool ok = false;
      for (int k = 1; k <= 3 && i - k >= s; k++) {
        uint32_t p = t[i - k];
        if ((p & 0xC0) == 0x80) continue;
        ok = (p >= 0xC2) && (k < tb_u8len(p));
        break;
      }
      if (!ok) err = true;
      continue;
    }
    int len = tb_u8len(b);
    if (b < 0xC2 || b > 0xF4 || i + len > e) { err = true; continue; }
    bool cont_ok = true;
    for (int k = 1; k < len; k++) if ((t[i + k] & 0xC0) != 0x80) cont_ok = false;
    if (!cont_ok) { err = true; continue; }
    uint32_t b1 = t[i + 1];
    if (b == 0xE0 && b1 < 0xA0) err = true;
    if (b == 0xED && b1 > 0x9F) err = true;
    if (b == 0xF0 && b1 < 0x90) err = true;
    if (b == 0xF4 && b1 > 0x8F) err = true;
  }
  if (simt_any(err) && lane == 0) simt_atomic_add((uint32_t*)bad, 1u);
}

// ---------------------------------------------------------------------------------------------
// GopherQualityFilter: one pass, then the threshold checks of gopher_quality.rs:198-295
// ---------------------------------------------------------------------------------------------
TB_D void body_gq(const KBatch& a, const uint16_t* s_ascii, const SwTable& sw, const GqCfg& cfg, int step,
                  tb_gq_stats* out, int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, a.rs[d], a.re[d]};
  GqAcc acc; acc.zero();
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    if (p0 < R.re) { int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re; gq_lane(C, R, sw, p0, p1, acc); }
  }
  uint32_t n_words = simt_sum(acc.n_words), n_alpha = simt_sum(acc.n_alpha), n_stop = simt_sum(acc.n_stop),
           n_hash = simt_sum(acc.n_hash), n_ell = simt_sum(acc.n_ell), n_lines = simt_sum(acc.n_lines),
           n_bullet = simt_sum(acc.n_bullet), n_ell_lines = simt_sum(acc.n_ell_lines), flags = simt_or(acc.flags);
  uint64_t sum_chars = simt_sum64(acc.sum_chars);
  if (lane) return;
  tb_gq_stats st;
  st.n_words = n_words; st.sum_word_chars = sum_chars; st.n_hash = n_hash; st.n_ellipsis_units = n_ell;
  st.n_lines = n_lines; st.n_bullet_lines = n_bullet; st.n_ellipsis_lines = n_ell_lines; st.n_alpha_words = n_alpha;
  st.n_stop_words = n_stop; st.flags = flags ? TB_DOCFLAG_COMPLEX_SCRIPT : 0;
  double avg = n_words ? (double)sum_chars / (double)n_words : 0.0;
  double nw = (double)(n_words ? n_words : 1u), nl = (double)(n_lines ? n_lines : 1u);
  uint32_t r = 0;
  if ((cfg.has & TB_GQ_HAS_MIN_DOC_WORDS) && (uint64_t)n_words < cfg.min_doc_words) r |= TB_GQ_REASON_SHORT_DOC;
  if ((cfg.has & TB_GQ_HAS_MAX_DOC_WORDS) && (uint64_t)n_words > cfg.max_doc_words) r |= TB_GQ_REASON_LONG_DOC;
  if ((cfg.has & TB_GQ_HAS_MIN_AVG_WORD_LENGTH) && avg < cfg.min_avg_word_length) r |= TB_GQ_REASON_BELOW_AVG;
  if ((cfg.has & TB_GQ_HAS_MAX_AVG_WORD_LENGTH) && n_words > 0 && avg > cfg.max_avg_word_length) r |= TB_GQ_REASON_ABOVE_AVG;
  if (cfg.has & TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO) {
    if ((double)n_hash / nw > cfg.max_symbol_word_ratio) r |= TB_GQ_REASON_TOO_MANY_HASHES;
    if ((double)n_ell / nw > cfg.max_symbol_word_ratio) r |= TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS;
  }
  if ((cfg.has & TB_GQ_HAS_MAX_BULLET_LINES_RATIO) && (double)n_bullet / nl > cfg.max_bullet_lines_ratio) r |= TB_GQ_REASON_TOO_MANY_BULLETS;
  if ((cfg.has & TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO) && (double)n_ell_lines / nl > cfg.max_ellipsis_lines_ratio) r |= TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES;
  if ((cfg.has & TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO) && (double)n_alpha / nw < cfg.max_non_alpha_words_ratio) r |= TB_GQ_REASON_BELOW_ALPHA;
  if ((cfg.has & TB_GQ_HAS_MIN_STOP_WORDS) && cfg.min_stop_words > 0 && (uint64_t)n_stop < cfg.min_stop_words) r |= TB_GQ_REASON_TOO_FEW_STOP_WORDS;
  st.reason_bits = r;
  out[d] = st;
  if (r) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
}

// str::trim() of one live range (thread per document: leading / trailing White_Space is short)
TB_D void body_trim(const KBatch& a, const uint16_t* s_ascii, int64_t* ts, int64_t* te, int64_t d) {
  Classifier C{s_ascii, a.uc};
  int64_t s = a.rs[d], e = a.re[d];
  if (a.alive[d]) {
    while (s < e) { int len; uint32_t cp; if (!(C.at(a.text, s, &len, &cp) & CI_WS)) break; s += len; }
    while (e > s) { int64_t k = C.prev_lead(a.text, e, s); int len; uint32_t cp; if (!(C.at(a.text, k, &len, &cp) & CI_WS)) break; e = k; }
  }
  ts[d] = s; te[d] = e;
}

// ---------------------------------------------------------------------------------------------
// GopherRepetitionFilter, pass 1: words + line pieces of the trimmed text
// ---------------------------------------------------------------------------------------------
struct HL { uint32_t a, b, len; };  // partial piece: hash residues + byte length
TB_D HL hl_cat(HL x, HL y, const uint64_t* powtab) {  // x followed by y
  uint64_t pw = tb_pow_pack(powtab, y.len);
  HL r;
  r.a = tb_am(tb_mm(x.a, (uint32_t)(pw >> 32)), y.a);
  r.b = tb_am(tb_mm(x.b, (uint32_t)pw), y.b);
  r.len = x.len + y.len;
  return r;
}
TB_D HL hl_shfl_up(HL v, int o) { HL r; r.a = simt_shfl_up(v.a, o); r.b = simt_shfl_up(v.b, o); r.len = simt_shfl_up(v.len, o); return r; }
TB_D HL hl_shfl(HL v, int src) { HL r; r.a = simt_shfl(v.a, src); r.b = simt_shfl(v.b, src); r.len = simt_shfl(v.len, src); return r; }
TB_D uint64_t hl_pack(HL v) { return ((uint64_t)tb_canon31(v.a) << 32) | tb_canon31(v.b); }

TB_D void body_gr_scan(const KBatch& a, const uint16_t* s_ascii, const int64_t* ts, const int64_t* te, const GrScratch& s,
                       int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, ts[d], te[d]};
  const uint64_t ub = (uint64_t)(R.rs >> 1) + (uint64_t)d;
  uint64_t* wh = s.wh + ub; uint32_t* wl = s.wl + ub;
  uint64_t* lh = s.lh + ub; uint32_t* ll = s.ll + ub; uint8_t* lg = s.lg + ub;
  uint32_t W = 0, NL = 0, flags = 0;
  uint64_t nchars = 0;
  HL carry{0, 0, 0};  // open line piece entering the tile
  GrLane o;
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    bool active = p0 < R.re;
    int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re;
    if (active) gr_lane(C, R, p0, p1, o);
    else { o.nw = 0; o.has_nl = 0; o.np = 0; o.nchars = 0; o.flags = 0; o.tail.init(); o.tail_len = 0; o.head.init(); o.head_len = 0; o.first_nl = 0; }
    flags |= o.flags; nchars += o.nchars;
    // ---- words: dense, ordered
    uint32_t tot, pre = simt_exscan(o.nw, lane, &tot);
    for (uint32_t k = 0; k < o.nw; k++) { wh[W + pre + k] = o.wh[k]; wl[W + pre + k] = o.wl[k]; }
    W += tot;
    // ---- line pieces: segmented scan of the open-piece value; a lane with a '\n' restarts it
    HL v{o.tail.a, o.tail.b, o.tail_len};
    uint32_t f = o.has_nl;
    for (int off = 1; off < 32; off <<= 1) {
      HL vu = hl_shfl_up(v, off);
      uint32_t fu = simt_shfl_up(f, off);
      if (lane >= off && !f) { v = hl_cat(vu, v, a.uc.powtab); f = fu; }
    }
    if (!f) v = hl_cat(carry, v, a.uc.powtab);  // no '\n' at or before this lane in the tile
    HL cin = hl_shfl_up(v, 1);
    if (lane == 0) cin = carry;
    // pieces completed in this lane: (cin + head) at the first '\n', then the inner ones
    uint32_t ncomp = 0;
    HL first{0, 0, 0};
    if (o.has_nl) {
      HL hd{o.head.a, o.head.b, o.head_len};
      first = hl_cat(cin, hd, a.uc.powtab);
      ncomp = (first.len ? 1u : 0u) + o.np;
    }
    uint32_t ltot, lpre = simt_exscan(ncomp, lane, &ltot);
    if (ncomp) {
      uint32_t k = NL + lpre;
      if (first.len) {
        int64_t st = p0 + o.first_nl - first.len;
        lh[k] = hl_pack(first); ll[k] = first.len; lg[k] = (uint8_t)gr_gap2(R, st); k++;
      }
      for (uint32_t j = 0; j < o.np; j++) {
        int64_t st = p0 + o.pend[j] - o.pl[j];
        lh[k] = o.ph[j]; ll[k] = o.pl[j]; lg[k] = (uint8_t)gr_gap2(R, st); k++;
      }
    }
    NL += ltot;
    carry = hl_shfl(v, 31);
  }
  if (carry.len) {  // the piece that runs to the end of the text
    if (lane == 0) { lh[NL] = hl_pack(carry); ll[NL] = carry.len; lg[NL] = (uint8_t)gr_gap2(R, R.re - carry.len); }
    NL++;
  }
  flags = simt_or(flags);
  nchars = simt_sum64(nchars);
  if (lane == 0) {
    GrDoc g; g.char_len = nchars; g.W = W; g.NL = NL; g.flags = flags ? TB_DOCFLAG_COMPLEX_SCRIPT : 0; g.empty = (R.re <= R.rs);
    s.doc[d] = g;
  }
}

// ---------------------------------------------------------------------------------------------
// GopherRepetitionFilter, pass 2: duplicate lines / paragraphs, top and duplicate n-grams.
// Open-addressing tables of element indices (keys live in the arrays; equality = 62-bit hash +
// byte length); in shared memory when they fit, else in the document's global work area.
// ---------------------------------------------------------------------------------------------
#define T_EMPTY 0xFFFFFFFFu
TB_D uint32_t tab_size(uint32_t count) {
  uint32_t t = TB_REP_MIN_TABLE;
  while (t < 2u * count) t <<= 1;
  return t;
}
TB_D void tab_clear(uint32_t* tab, uint32_t T, int lane) {
  for (uint32_t i = lane; i < T; i += 32) tab[i] = T_EMPTY;
  simt_sync();
}

// find_duplicates (text.rs:197-208) over n pieces: dup_elems = n - distinct, dup_bytes = total - distinct bytes
TB_D void dedup_pieces(uint32_t* tab, const uint64_t* h, const uint32_t* len, uint32_t n, int lane, uint32_t* dup_elems,
                       uint64_t* dup_bytes) {
  uint32_t T = tab_size(n);
  tab_clear(tab, T, lane);
  uint32_t distinct = 0; uint64_t dbytes = 0, total = 0;
  for (uint32_t i = lane; i < n; i += 32) {
    uint64_t hi = h[i]; uint32_t li = len[i];
    uint32_t sl = (uint32_t)tb_mix64(hi ^ ((uint64_t)li << 40)) & (T - 1);
    total += li;
    for (;;) {
      uint32_t old = simt_cas(&tab[sl], T_EMPTY, i);
      if (old == T_EMPTY) { distinct++; dbytes += li; break; }
      if (h[old] == hi && len[old] == li) break;
      sl = (sl + 1) & (T - 1);
    }
  }
  simt_sync();
  distinct = simt_sum(distinct); dbytes = simt_sum64(dbytes); total = simt_sum64(total);
  *dup_elems = n - distinct; *dup_bytes = total - dbytes;
}

TB_D void body_gr_rep(const KBatch& a, uint32_t* s_tab, const int64_t* ts, const int64_t* te, const GrScratch& s,
                      const GrCfg& cfg, int step, tb_gr_stats* out, int64_t d, int lane) {
  const GrDoc g = s.doc[d];
  tb_gr_stats st;
  st.reason_bits = 0; st.flags = g.flags; st.char_len = g.char_len; st.n_paras = 0; st.para_dup_elems = 0; st.para_dup_bytes = 0;
  st.n_lines = 0; st.line_dup_elems = 0; st.line_dup_bytes = 0; st.n_words = 0;
  for (int k = 0; k < TB_MAX_NGRAMS; k++) { st.top_bytes[k] = 0; st.dup_bytes[k] = 0; }
  if (g.empty) {  // gopher_rep.rs:60-74
    if (lane == 0) { st.reason_bits = TB_GR_REASON_EMPTY; out[d] = st; a.alive[d] = 0; a.failed[d] = (int8_t)step; }
    return;
  }
  const int64_t rs = ts[d], re = te[d];
  const uint64_t ub = (uint64_t)(rs >> 1) + (uint64_t)d;
  const uint32_t U = (uint32_t)((re - rs) >> 1) + 1u;
  const uint64_t* wh = s.wh + ub; const uint32_t* wl = s.wl + ub;
  const uint64_t* lh = s.lh + ub; const uint32_t* ll = s.ll + ub; const uint8_t* lg = s.lg + ub;
  // work area layout (32-bit words): C[2U] J[2U] gtab[4U+16] cnt[U] fo[n_dup*U] woff[U+1]
  uint32_t* work = s.work + ub * s.K + (uint64_t)d * TB_REP_DOC_EXTRA;
  uint64_t* Cc = (uint64_t*)work;
  uint64_t* Jj = Cc + U;
  uint32_t* gtab = (uint32_t*)(Jj + U);
  uint32_t* cnt = gtab + 4u * U + TB_REP_MIN_TABLE;
  uint32_t* fo = cnt + U;
  uint32_t* woff = fo + (uint32_t)cfg.n_dup * U;
  const uint32_t W = g.W, NL = g.NL;
  const uint32_t maxcount = W > NL ? W : NL;
  uint32_t* tab = tab_size(maxcount) <= TB_REP_SMEM_TABLE ? s_tab : gtab;
  const double text_char_len = (double)(g.char_len ? g.char_len : 1);
  uint32_t reasons = 0;

  // ---- lines (gopher_rep.rs:120-156)
  uint32_t de; uint64_t db;
  dedup_pieces(tab, lh, ll, NL, lane, &de, &db);
  st.n_lines = NL; st.line_dup_elems = de; st.line_dup_bytes = db;
  {
    double r1 = (double)de / (double)(NL ? NL : 1u), r2 = (double)db / text_char_len;
    if ((cfg.has & TB_GR_HAS_DUP_LINE_FRAC) && r1 > cfg.dup_line_frac) reasons |= TB_GR_REASON_DUP_LINE_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_LINE_CHAR_FRAC) && r2 > cfg.dup_line_char_frac) reasons |= TB_GR_REASON_DUP_LINE_CHAR_FRAC;
  }
  // ---- paragraphs (gopher_rep.rs:80-116): fold the ordered line pieces; a single '\n' joins
  // lines inside a paragraph, a run of >= 2 separates paragraphs.
  uint32_t NP = 0;
  if (lane == 0) {
    Hash2 cur; cur.init(); uint32_t curlen = 0;
    for (uint32_t k = 0; k < NL; k++) {
      Hash2 hk = tb_unpack(lh[k]);
      if (k == 0) { cur = hk; curlen = ll[k]; continue; }
      if (lg[k]) { Cc[NP] = cur.pack(); cnt[NP] = curlen; NP++; cur = hk; curlen = ll[k]; }
      else { cur.push('\n'); cur = tb_hcat(cur, hk, tb_pow_pack(a.uc.powtab, ll[k])); curlen += 1 + ll[k]; }
    }
    if (NL) { Cc[NP] = cur.pack(); cnt[NP] = curlen; NP++; }
  }
  NP = simt_shfl(NP, 0);
  simt_sync();
  dedup_pieces(tab, Cc, cnt, NP, lane, &de, &db);
  st.n_paras = NP; st.para_dup_elems = de; st.para_dup_bytes = db;
  {
    double r1 = (double)de / (double)(NP ? NP : 1u), r2 = (double)db / text_char_len;
    if ((cfg.has & TB_GR_HAS_DUP_PARA_FRAC) && r1 > cfg.dup_para_frac) reasons |= TB_GR_REASON_DUP_PARA_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_PARA_CHAR_FRAC) && r2 > cfg.dup_para_char_frac) reasons |= TB_GR_REASON_DUP_PARA_CHAR_FRAC;
  }
  simt_sync();

  // ---- n-grams (gopher_rep.rs:159-196)
  uint32_t nmax = 0;
  for (int k = 0; k < cfg.n_top; k++) nmax = cfg.top_n[k] > nmax ? cfg.top_n[k] : nmax;
  for (int k = 0; k < cfg.n_dup; k++) nmax = cfg.dup_n[k] > nmax ? cfg.dup_n[k] : nmax;
  if (nmax > W) nmax = W;
  uint64_t top_val[TB_MAX_NGRAMS];
  uint32_t dup_any = 0;  // bit k: dup_n_grams[k] has at least one repeated key
  for (int k = 0; k < TB_MAX_NGRAMS; k++) top_val[k] = 0;
  if (nmax) {
    // byte offsets of the words in the concatenation; C = concat hash, J = " "-joined hash
    uint32_t run = 0;
    for (uint32_t b0 = 0; b0 < W; b0 += 32) {
      uint32_t i = b0 + lane, v = i < W ? wl[i] : 0, tot, pre = simt_exscan(v, lane, &tot);
      if (i < W) { woff[i] = run + pre; Cc[i] = wh[i]; Jj[i] = wh[i]; }
      run += tot;
    }
    if (lane == 0) woff[W] = run;
    simt_sync();
    for (uint32_t n = 1; n <= nmax; n++) {
      const uint32_t cn = W - n + 1;  // number of n-grams
      if (n > 1) {
        for (uint32_t i = lane; i < cn; i += 32) {
          uint32_t j = i + n - 1;
          Hash2 hj = tb_unpack(wh[j]);
          uint64_t pw = tb_pow_pack(a.uc.powtab, wl[j]);
          Cc[i] = tb_hcat(tb_unpack(Cc[i]), hj, pw).pack();
          Hash2 jj = tb_unpack(Jj[i]); jj.push(' ');
          Jj[i] = tb_hcat(jj, hj, pw).pack();
        }
        simt_sync();
      }
      const uint32_t T = tab_size(cn);
      for (int k = 0; k < cfg.n_top; k++) {
        if (cfg.top_n[k] != n) continue;
        // find_top_duplicate (text.rs:211-238): max count, then the longest key among those
        tab_clear(tab, T, lane);
        for (uint32_t i = lane; i < cn; i += 32) cnt[i] = 0;
        simt_sync();
        for (uint32_t i = lane; i < cn; i += 32) {
          // key length = concatenated bytes + (n - 1) spaces; equal joined strings <=> equal hash + length
          uint32_t len = woff[i + n] - woff[i] + n - 1;
          uint64_t key = Jj[i];
          uint32_t sl = (uint32_t)tb_mix64(key ^ ((uint64_t)len << 40)) & (T - 1);
          for (;;) {
            uint32_t old = simt_cas(&tab[sl], T_EMPTY, i);
            if (old == T_EMPTY) { simt_atomic_add(&cnt[i], 1u); break; }
            if (Jj[old] == key && woff[old + n] - woff[old] + n - 1 == len) { simt_atomic_add(&cnt[old], 1u); break; }
            sl = (sl + 1) & (T - 1);
          }
        }
        simt_sync();
        uint64_t best = 0;
        for (uint32_t i = lane; i < cn; i += 32) {
          uint32_t c = cnt[i];
          if (c) { uint64_t v = ((uint64_t)c << 32) | (woff[i + n] - woff[i] + n - 1); best = v > best ? v : best; }
        }
        top_val[k] = simt_max64(best);
        simt_sync();
      }
      for (int k = 0; k < cfg.n_dup; k++) {
        if (cfg.dup_n[k] != n) continue;
        // first occurrence (smallest position) of every concatenated key -> fo_k[i]
        uint32_t* fok = fo + (uint32_t)k * U;
        tab_clear(tab, T, lane);
        for (uint32_t i = lane; i < cn; i += 32) {
          uint32_t len = woff[i + n] - woff[i];
          uint64_t key = Cc[i];
          uint32_t sl = (uint32_t)tb_mix64(key ^ ((uint64_t)len << 40)) & (T - 1);
          for (;;) {
            uint32_t old = simt_cas(&tab[sl], T_EMPTY, i);
            if (old == T_EMPTY) break;
            if (Cc[old] == key && woff[old + n] - woff[old] == len) { simt_atomic_min(&tab[sl], i); break; }
            sl = (sl + 1) & (T - 1);
          }
          fok[i] = sl;
        }
        simt_sync();
        bool any = false;
        for (uint32_t i = lane; i < cn; i += 32) { uint32_t f = tab[fok[i]]; fok[i] = f; any |= (f != i); }
        if (simt_any(any)) dup_any |= 1u << k;
        simt_sync();
      }
    }
  }
  // find_all_duplicate (text.rs:241-259): the greedy scan is sequential; one lane per configured n.
  uint64_t dupv = 0;
  if (lane < cfg.n_dup && ((dup_any >> lane) & 1u)) {
    const uint32_t n = cfg.dup_n[lane];
    uint32_t* fok = fo + (uint32_t)lane * U;
    const uint32_t SEEN = 0x80000000u;
    uint32_t idx = 0;
    while (idx + n <= W) {
      uint32_t c = fok[idx] & ~SEEN;
      uint32_t v = fok[c];
      if (v & SEEN) { dupv += woff[idx + n] - woff[idx]; idx += n; }
      else { fok[c] = v | SEEN; idx += 1; }
    }
  }
  simt_sync();
  for (int k = 0; k < cfg.n_dup; k++) {
    uint64_t v = simt_shfl(dupv, k);
    st.dup_bytes[k] = v;
    if (cfg.dup_n[k] > 0 && (double)v / text_char_len > cfg.dup_frac[k]) reasons |= 1u << (TB_GR_REASON_DUP_SHIFT + k);
  }
  for (int k = 0; k < cfg.n_top; k++) {
    uint32_t c = (uint32_t)(top_val[k] >> 32), len = (uint32_t)top_val[k];
    uint64_t v = c > 1 ? (uint64_t)len * c : 0;
    st.top_bytes[k] = v;
    if (cfg.top_n[k] > 0 && (double)v / text_char_len > cfg.top_frac[k]) reasons |= 1u << (TB_GR_REASON_TOP_SHIFT + k);
  }
  if (lane == 0) {
    st.n_words = W; st.reason_bits = reasons;
    out[d] = st;
    if (reasons) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
  }
}

// ---------------------------------------------------------------------------------------------
// C4QualityFilter
// ---------------------------------------------------------------------------------------------
// Pass A: doc-level checks, processed lines P, line count; zeroes the per-line slots.
TB_D void body_c4_a(const KBatch& a, const uint16_t* s_ascii, const C4Cfg& cfg, const C4Scratch& s, int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, a.rs[d], a.re[d]};
  uint8_t* out = s.P + R.rs;
  const uint64_t lb = (uint64_t)R.rs + (uint64_t)d;
  uint32_t flags = 0, nl = 0;
  int64_t outpos = 0;
  C4aLane o;
  if (lane == 0) { s.lwc[lb] = 0; s.lmax[lb] = 0; s.lflags[lb] = 0; }
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re;
    if (p0 < R.re) c4a_lane(C, R, p0, p1, cfg.remove_citations != 0, cfg.filter_lorem_ipsum != 0, o);
    else { o.nkept = 0; o.n_nl = 0; o.flags = 0; }
    flags |= o.flags;
    uint32_t tot, pre = simt_exscan(o.nkept, lane, &tot);
    for (uint32_t k = 0; k < o.nkept; k++) out[outpos + pre + k] = o.bytes[k];
    outpos += tot;
    // zero the slot of every line that starts after a '\n' of this stretch
    uint32_t ntot, npre = simt_exscan(o.n_nl, lane, &ntot);
    for (uint32_t k = 0; k < o.n_nl; k++) { uint64_t li = lb + nl + npre + k + 1; s.lwc[li] = 0; s.lmax[li] = 0; s.lflags[li] = 0; }
    nl += ntot;
  }
  flags = simt_or(flags);
  if (lane == 0) {
    uint32_t lines = nl + ((R.re > R.rs && a.text[R.re - 1] != '\n') ? 1u : 0u);  // str::lines().count()
    s.plen[d] = outpos; s.nlines[d] = lines;
    uint32_t df = 0;
    if (cfg.filter_lorem_ipsum && (flags & C4F_LOREM)) df |= C4F_LOREM;
    if (cfg.filter_curly_bracket && (flags & C4F_CURLY)) df |= C4F_CURLY;
    s.docflags[d] = df;
    s.prs[d] = R.rs; s.pre[d] = R.rs + outpos;
  }
}

// Pass B: words per processed line (count, longest), terminal punctuation, javascript / policy.
struct C4bSink {
  uint32_t* lwc; uint32_t* lmax; uint64_t lbase; int64_t p0, p1; uint32_t nlmask;
  TB_D void on_word(WordInfo& w) {
    int64_t x = w.start < p1 ? w.start : p1;
    uint32_t below = (x - p0) >= 32 ? nlmask : (nlmask & ((1u << (uint32_t)(x - p0)) - 1u));
    uint64_t li = lbase + simt_popc(below);
    simt_atomic_add(&lwc[li], 1u);
    simt_atomic_max(&lmax[li], w.nchars);
  }
};
TB_D void body_c4_b(const KBatch& a, const uint16_t* s_ascii, const C4Cfg& cfg, const C4Scratch& s, int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{s.P, s.prs[d], s.pre[d]};
  const uint64_t lb = (uint64_t)a.rs[d] + (uint64_t)d;
  const uint8_t* t = R.t;
  uint32_t nl = 0;
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re;
    uint32_t nlmask = 0;
    if (p0 < R.re) for (int64_t i = p0; i < p1; i++) if (t[i] == '\n') nlmask |= 1u << (uint32_t)(i - p0);
    uint32_t tot, pre = simt_exscan((uint32_t)simt_popc(nlmask), lane, &tot);
    if (p0 < R.re) {
      uint64_t lbase = lb + nl + pre;
      C4bSink sink{s.lwc, s.lmax, lbase, p0, p1, nlmask};
      tb_walk_words<0>(C, R, p0, p1, sink);
      uint32_t line = 0, fl = 0;
      for (int64_t i = p0; i < p1; i++) {
        uint32_t b = t[i];
        if (b == '\n') {
          fl |= c4_line_end_flags(C, R, i);
          if (fl) simt_atomic_or(&s.lflags[lbase + line], fl);
          fl = 0; line++;
        } else if (b < 0x80) {
          fl |= c4_pattern_flags(t, i, R.re, b, cfg.filter_javascript != 0, cfg.filter_policy != 0);
        }
        if (i == R.re - 1 && b != '\n') fl |= c4_line_end_flags(C, R, R.re);
      }
      if (fl) simt_atomic_or(&s.lflags[lbase + line], fl);
    }
    nl += tot;
  }
}

// Pass C: drop lines, join the kept ones with '\n' (c4_filters.rs:207-258) -> F; line counters.
TB_D void body_c4_c(const KBatch& a, const C4Cfg& cfg, const C4Scratch& s, int64_t d, int lane) {
  if (!a.alive[d] || s.docflags[d]) { if (lane == 0) { s.frs[d] = a.rs[d]; s.fre[d] = a.rs[d]; } return; }
  const int64_t rs = s.prs[d], re = s.pre[d];
  const uint8_t* t = s.P;
  uint8_t* out = s.F + a.rs[d];
  const uint64_t lb = (uint64_t)a.rs[d] + (uint64_t)d;
  const uint32_t m = s.nlines[d];
  uint32_t c1 = 0, c2 = 0, c3 = 0, kept = 0;  // this lane's share of the counters
  uint32_t nl = 0, kept_run = 0;              // lines terminated / kept before this tile
  int64_t outpos = 0;
  if (m) {
    for (int64_t base = rs; base < re; base += TB_TILE) {
      int64_t p0 = base + lane * TB_STRETCH;
      int64_t p1 = p0 + TB_STRETCH < re ? p0 + TB_STRETCH : re;
      uint32_t nlmask = 0;
      if (p0 < re) for (int64_t i = p0; i < p1; i++) if (t[i] == '\n') nlmask |= 1u << (uint32_t)(i - p0);
      uint32_t tot, pre = simt_exscan((uint32_t)simt_popc(nlmask), lane, &tot);
      // decisions of the lines terminated in this stretch
      uint32_t keepmask = 0, nkept_here = 0;
      const uint64_t l0 = lb + nl + pre;  // line that contains the first byte of the stretch
      {
        uint32_t mm = nlmask, j = 0;
        while (mm) {
          uint32_t bit = (uint32_t)simt_ffs(mm) - 1; mm &= mm - 1;
          int dec = c4_line_decision(cfg, s.lwc[l0 + j], s.lmax[l0 + j], s.lflags[l0 + j]);
          if (dec == 0) { keepmask |= 1u << bit; nkept_here++; kept++; }
          else if (dec == 1) c1++; else if (dec == 2) c2++; else if (dec == 3) c3++;
          j++;
        }
      }
      uint32_t ktot, kpre = simt_exscan(nkept_here, lane, &ktot);
      // emit: bytes of kept lines; a '\n' in front of a kept line that follows another kept line
      uint8_t buf[TB_STRETCH];
      uint32_t nb = 0;
      if (p0 < re) {
        uint32_t j = 0, kept_before = kept_run + kpre;
        bool keep_cur = c4_line_decision(cfg, s.lwc[l0], s.lmax[l0], s.lflags[l0]) == 0;
        for (int64_t i = p0; i < p1; i++) {
          uint32_t b = t[i];
          if (b == '\n') {
            if ((keepmask >> (uint32_t)(i - p0)) & 1u) kept_before++;
            j++;
            keep_cur = c4_line_decision(cfg, s.lwc[l0 + j], s.lmax[l0 + j], s.lflags[l0 + j]) == 0;
            if (keep_cur && kept_before) buf[nb++] = '\n';
          } else if (keep_cur) buf[nb++] = (uint8_t)b;
        }
      }
      uint32_t btot, bpre = simt_exscan(nb, lane, &btot);
      for (uint32_t k = 0; k < nb; k++) out[outpos + bpre + k] = buf[k];
      outpos += btot; nl += tot; kept_run += ktot;
    }
    if (lane == 0) {  // the last line has no terminator
      int dec = c4_line_decision(cfg, s.lwc[lb + m - 1], s.lmax[lb + m - 1], s.lflags[lb + m - 1]);
      if (dec == 0) kept++; else if (dec == 1) c1++; else if (dec == 2) c2++; else if (dec == 3) c3++;
    }
  }
  c1 = simt_sum(c1); c2 = simt_sum(c2); c3 = simt_sum(c3); kept = simt_sum(kept);
  if (lane == 0) {
    s.cnt[4 * d + 0] = c1; s.cnt[4 * d + 1] = c2; s.cnt[4 * d + 2] = c3; s.cnt[4 * d + 3] = kept;
    s.frs[d] = a.rs[d]; s.fre[d] = a.rs[d] + outpos;
  }
}

// Pass D: sentence count of the (trimmed) rewritten content and the document decision.
TB_D void body_c4_d(const KBatch& a, const uint16_t* s_ascii, const C4Cfg& cfg, const C4Scratch& s, int step,
                    tb_c4_stats* out, uint8_t* rewritten, int64_t* rs_out, int64_t* re_out, int64_t d, int lane) {
  tb_c4_stats st;
  st.n_lines = 0; st.too_long_word = 0; st.no_terminal_punc = 0; st.too_few_words = 0; st.n_kept_lines = 0;
  st.n_sentences = 0; st.rewritten = 0; st.reason_bits = 0;
  uint32_t df = s.docflags[d];
  if (df) {  // early exit, content untouched (c4_filters.rs:175-187)
    if (lane == 0) {
      if (df & C4F_LOREM) st.reason_bits |= TB_C4_REASON_LOREM_IPSUM;
      if (df & C4F_CURLY) st.reason_bits |= TB_C4_REASON_CURLY_BRACKET;
      out[d] = st; a.alive[d] = 0; a.failed[d] = (int8_t)step;
    }
    return;
  }
  Classifier C{s_ascii, a.uc};
  TextRange R{s.F, s.frs[d], s.fre[d]};  // already trimmed by body_trim
  uint32_t ns = 0;
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    if (p0 < R.re) { int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re; ns += tb_count_sentences(C, R, p0, p1); }
  }
  ns = simt_sum(ns);
  if (lane) return;
  st.n_lines = s.nlines[d]; st.too_long_word = s.cnt[4 * d]; st.no_terminal_punc = s.cnt[4 * d + 1];
  st.too_few_words = s.cnt[4 * d + 2]; st.n_kept_lines = s.cnt[4 * d + 3]; st.n_sentences = ns; st.rewritten = 1;
  if (cfg.min_num_sentences > 0 && (uint64_t)ns < cfg.min_num_sentences) st.reason_bits |= TB_C4_REASON_TOO_FEW_SENTENCES;
  out[d] = st;
  rewritten[d] = 1;
  rs_out[d] = R.rs; re_out[d] = R.re;
  if (st.reason_bits) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
}

// ---------------------------------------------------------------------------------------------
// final content
// ---------------------------------------------------------------------------------------------
TB_D void body_final_len(const int64_t* rs0, const int64_t* re0, const int64_t* rs, const int64_t* re,
                         const uint8_t* rewritten, int64_t* len, int64_t d) {
  len[d] = rewritten[d] ? re[d] - rs[d] : re0[d] - rs0[d];
}
TB_D void body_pack(const uint8_t* orig, const int64_t* rs0, const uint8_t* cur, const int64_t* rs, const uint8_t* rewritten,
                    const int64_t* out_off, uint8_t* out, int64_t d, int lane) {
  const uint8_t* src = rewritten[d] ? cur + rs[d] : orig + rs0[d];
  int64_t o = out_off[d], len = out_off[d + 1] - o;
  for (int64_t i = lane; i < len; i += 32) out[o + i] = src[i];
}
