// tb_bodies_fw.cuh -- FineWebQualityFilter (/root/reference src/pipeline/filters/fineweb_quality.rs:71-226),
// the first "next" row of SURVEY 8(f) that is a filter.  Same shape as the other text passes: one warp per
// document, class masks per 32-byte stretch (tb_tile.cuh).  Pass 1 (body_fw_scan) counts words, newlines
// and characters and emits the non-blank lines of `content.lines()` as (hash, byte length); pass 2
// (body_fw_fin) runs find_duplicates over them (text.rs:197-208) and applies the checks in the
// reference's order.  Included at the end of tb_bodies.cuh.
#pragma once

TB_HD bool fw_is_stop_char(const FwCfg& cfg, uint32_t cp) {
  if (cfg.n_stop_chars < 0)  // END_PUNCTUATION, fineweb_quality.rs:26
    return cp == '.' || cp == '!' || cp == '?' || cp == '"' || cp == '\'' || cp == 0x201Du;
  for (int i = 0; i < cfg.n_stop_chars; i++) if (cfg.stop_chars[i] == cp) return true;
  return false;
}
// line.trim_end().chars().last() of the line that ends at e (a '\n' or the end of the text) is a stop
// character (:93-104).  Only called for non-blank lines.
TB_HD_NOINLINE bool fw_line_stop(const Classifier& C, const TextRange& R, const FwCfg& cfg, int64_t e) {
  const uint8_t* t = R.t;
  int64_t j = e;
  while (j > R.rs) {
    const int64_t k = C.prev_lead(t, j, R.rs);
    int len; uint32_t cp;
    const uint32_t ci = C.at(t, k, &len, &cp);
    if (cp == '\n') return false;
    if (!(ci & CI_WS)) return fw_is_stop_char(cfg, cp);
    j = k;
  }
  return false;
}

// open line: hash residues, byte length, character count
struct FL { uint32_t a, b, len, nc; };
TB_D FL fl_cat(FL x, FL y, const uint64_t* powtab) {
  HL h = hl_cat(HL{x.a, x.b, x.len}, HL{y.a, y.b, y.len}, powtab);
  return FL{h.a, h.b, h.len, x.nc + y.nc};
}
TB_D FL fl_shfl_up(FL v, int o) {
  FL r; r.a = simt_shfl_up(v.a, o); r.b = simt_shfl_up(v.b, o); r.len = simt_shfl_up(v.len, o); r.nc = simt_shfl_up(v.nc, o);
  return r;
}
TB_D FL fl_shfl(FL v, int src) {
  FL r; r.a = simt_shfl(v.a, src); r.b = simt_shfl(v.b, src); r.len = simt_shfl(v.len, src); r.nc = simt_shfl(v.nc, src);
  return r;
}

struct FwCountSink { uint32_t n; TB_D void on_word(WordInfo&) { n++; } };
// general path of the word count (documents the mask form cannot express)
TB_D_NOINLINE uint32_t fw_count_words_walk(const Classifier& C, const TextRange& R, int lane) {
  FwCountSink sink{0u};
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    const int64_t p0 = base + lane * TB_STRETCH;
    if (p0 < R.re) { const int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re; tb_walk_words_fast<0>(C, R, p0, p1, sink); }
  }
  return sink.n;
}

// ---------------------------------------------------------------------------------------------
// Pass 1: words, newlines, characters; non-blank lines -> (hash, byte length); stop / short line counts
// ---------------------------------------------------------------------------------------------
TB_D void body_fw_scan(const KBatch& a, const uint16_t* s_ascii, const LutEntry* lut, const FwCfg& cfg, const FwScratch& s,
                       int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, a.rs[d], a.re[d]};
  const uint8_t* t = R.t;
  const uint64_t ub = tb_lunit(R.rs, d);
  uint64_t* lh = s.lh + ub; uint32_t* ll = s.ll + ub;
  const uint64_t* powtab = a.uc.powtab;
  const uint32_t base1 = (uint32_t)(TB_LDG(powtab + 1) >> 32), base2 = (uint32_t)TB_LDG(powtab + 1);  // the context's hash bases
  const uint32_t below_lane = tb_below(lane);
  uint32_t NLc = 0;                                                   // lines emitted so far (uniform)
  uint32_t n_stop = 0, n_short = 0, n_words = 0, n_nl = 0, flags = 0;  // this lane's share
  uint64_t n_chars = 0;
  Frag wcarry{0u, 0u};
  FL lcarry{0u, 0u, 0u, 0u};
  uint32_t lcarry_nb = 0;  // the open line already showed a non-white-space character
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    Tile T;
    tile_classify<true>(lut, R, a.ext_lo, a.ext_hi, base + lane * TB_STRETCH, T);
    flags |= T.dirty ? tile_fix_dirty(C, R, T) : 0u;
    const int64_t p0 = T.p0;
    const uint32_t V = T.valid, B = T.brk, WC = T.m[TL_WC], CONT = T.h[TH_CONT], NLm = T.h[TH_NL], LEAD = V & ~CONT;
    uint32_t WSm = T.h[FH_WS];
    if (T.m[TL_NA]) {  // non-ASCII White_Space: only a few lead bytes can begin one
      const int fv = tb_ctz32(V);
      if (((CONT >> fv) & 1u) && p0 + fv > R.rs) {  // the stretch begins inside a character
        const int64_t k = C.prev_lead(t, p0 + fv, R.rs);
        int len; uint32_t cp;
        if (C.at(t, k, &len, &cp) & CI_WS) { const int64_t e = k + len - p0; WSm |= tb_below(e > 32 ? 32 : (int)e) & ~tb_below(fv); }
      }
      uint32_t cand = T.h[FH_WSLEAD];
      while (cand) {
        const int i = tb_ctz32(cand);
        cand &= cand - 1u;
        int len; uint32_t cp;
        if (C.at(t, p0 + i, &len, &cp) & CI_WS) WSm |= tb_below(i + len > 32 ? 32 : i + len) & ~tb_below(i);
      }
    }
    const uint32_t NW = V & ~WSm;
    n_chars += (uint32_t)tb_popc32(LEAD & ~NLm);  // :158-159
    n_nl += (uint32_t)tb_popc32(NLm);             // :189
    // ---- words (:188), counted where they end
    {
      const uint32_t H = tb_spread(WC, B);
      const int fb = tb_ctz32(B);
      const uint32_t headm = tb_below(fb), tailm = B ? ~tb_below(31 - tb_clz32(B)) : 0xFFFFFFFFu;
      n_words += (uint32_t)tb_popc32(B & H & ~headm & ~tailm);
      Frag cin = frag_scan(frag_of(T, tailm, 0u), B != 0u, lane, wcarry);
      if (B && (frag_cat(cin, frag_of(T, headm, 0u)).meta & FR_WC)) n_words++;
    }
    // ---- lines (:72-77): a '\r' directly in front of a '\n' is not part of the line (str::lines)
    const uint32_t NLx = NLm | ~V;  // every byte outside the text ends a line
    uint32_t SKIP = T.h[FH_CR] & (NLm >> 1);
    if ((T.h[FH_CR] >> 31) && p0 + 32 < R.re && tb_byte(t, p0 + 32) == '\n') SKIP |= 0x80000000u;
    const int fn = tb_ctz32(NLx);
    const uint32_t tailLm = NLx ? ~tb_upto(31 - tb_clz32(NLx)) : 0xFFFFFFFFu;
    const uint32_t nl_lanes = simt_ballot(NLx != 0u), tnb_lanes = simt_ballot((NW & tailLm) != 0u);
    uint32_t in_nb;
    {
      const uint32_t lower = nl_lanes & below_lane;
      if (lower) in_nb = (tnb_lanes & below_lane & ~tb_below(31 - tb_clz32(lower))) != 0u;
      else in_nb = lcarry_nb | ((tnb_lanes & below_lane) != 0u);
    }
    const uint32_t first_line = (NLx && (in_nb || (NW & tb_below(fn)))) ? 1u : 0u;
    const uint32_t SP = tb_spread_up(NW, NLx | (NLx << 1));    // a non-white-space byte at or before, in the same line
    const uint32_t LEq = NLx & (SP << 1) & ~tb_upto(fn);       // later line ends that close a non-blank line
    uint32_t ltot, lpre = simt_exscan((uint32_t)tb_popc32(LEq) + first_line, lane, &ltot);
    uint32_t la = 0, lb_ = 0, ln = 0, lc = 0, hla = 0, hlb = 0, hln = 0, hlc = 0;
    uint32_t lidx = NLc + lpre + first_line;
    {
      uint32_t wq[8];
#pragma unroll
      for (int k = 0; k < 8; k++) wq[k] = T.w[k];
      uint32_t NLq = NLx, LEqq = LEq, SKq = SKIP, LDq = LEAD;
      int fnq = fn;
#pragma unroll 1
      for (int k = 0; k < 8; k++) {
        const uint32_t x = wq[0];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          if ((NLq >> j) & 1u) {
            if (j == fnq) { hla = la; hlb = lb_; hln = ln; hlc = lc; }
            else if ((LEqq >> j) & 1u) {
              lh[lidx] = hl_pack(hl_fold(la, lb_, ln)); ll[lidx] = ln; lidx++;
              if ((uint64_t)lc <= cfg.short_line_length) n_short++;
              if (fw_line_stop(C, R, cfg, p0 + 4 * k + j)) n_stop++;
            }
            la = 0; lb_ = 0; ln = 0; lc = 0;
          } else if (!((SKq >> j) & 1u)) {
            const uint32_t byte = (x >> (8 * j)) & 0xFFu;
            la = tb_hpush(la, base1, byte + 1u); lb_ = tb_hpush(lb_, base2, byte + 1u); ln++;
            lc += (LDq >> j) & 1u;
          }
        }
#pragma unroll
        for (int q = 0; q < 7; q++) wq[q] = wq[q + 1];
        NLq >>= 4; LEqq >>= 4; SKq >>= 4; LDq >>= 4; fnq -= 4;
      }
    }
    // ---- the line that was open at the start of the stretch
    {
      FL v{tb_hfold(la), tb_hfold(lb_), ln, lc};
      uint32_t f = NLx != 0u;
      for (int off = 1; off < 32; off <<= 1) {
        FL vu = fl_shfl_up(v, off);
        uint32_t fu = simt_shfl_up(f, off);
        if (lane >= off && !f) { v = fl_cat(vu, v, powtab); f = fu; }
      }
      if (!f) v = fl_cat(lcarry, v, powtab);
      FL cin = fl_shfl_up(v, 1);
      if (lane == 0) cin = lcarry;
      lcarry = fl_shfl(v, 31);
      if (first_line) {
        FL first = fl_cat(cin, FL{tb_hfold(hla), tb_hfold(hlb), hln, hlc}, powtab);
        const uint32_t k = NLc + lpre;
        lh[k] = hl_pack(HL{first.a, first.b, first.len}); ll[k] = first.len;
        if ((uint64_t)first.nc <= cfg.short_line_length) n_short++;
        if (fw_line_stop(C, R, cfg, p0 + fn)) n_stop++;
      }
      if (nl_lanes) lcarry_nb = (tnb_lanes & ~tb_below(31 - tb_clz32(nl_lanes))) != 0u;
      else lcarry_nb |= tnb_lanes != 0u;
    }
    NLc += ltot;
  }
  if (lcarry_nb) {  // the last line has no terminator
    if (lane == 0) {
      lh[NLc] = hl_pack(HL{lcarry.a, lcarry.b, lcarry.len}); ll[NLc] = lcarry.len;
      if ((uint64_t)lcarry.nc <= cfg.short_line_length) n_short++;
      if (fw_line_stop(C, R, cfg, R.re)) n_stop++;
    }
    NLc++;
  }
  if (lane == 0 && (wcarry.meta & FR_WC)) n_words++;
  flags = simt_or(flags);
  if ((flags & TILE_EXOTIC) || (a.dbg & TB_DBG_FORCE_WALK)) n_words = fw_count_words_walk(C, R, lane);
  n_stop = simt_sum(n_stop); n_short = simt_sum(n_short); n_words = simt_sum(n_words); n_nl = simt_sum(n_nl);
  n_chars = simt_sum64(n_chars);
  if (lane == 0) {
    FwDoc fd;
    fd.n_chars = n_chars; fd.n_lines = NLc; fd.n_stop = n_stop; fd.n_short = n_short; fd.n_words = n_words; fd.n_newlines = n_nl;
    fd.flags = (flags & CI_COMPLEX) ? TB_DOCFLAG_COMPLEX_SCRIPT : 0u;
    s.doc[d] = fd;
  }
}

// ---------------------------------------------------------------------------------------------
// Pass 2: find_duplicates over the non-blank lines (:150-162), then the checks in the reference's order;
// like the reference, counters behind the first failing check stay 0.  `smem_tab`: 1024 words per warp.
// ---------------------------------------------------------------------------------------------
#define TB_FW_SMEM_LINES 512u
TB_D void body_fw_fin(const KBatch& a, uint32_t* smem_tab, const FwCfg& cfg, const FwScratch& s, int step, tb_fw_stats* out,
                      int64_t d, int lane) {
  const FwDoc fd = s.doc[d];
  const uint64_t ub = tb_lunit(a.rs[d], d);
  const uint32_t NL = fd.n_lines;
  uint32_t de = 0; uint64_t db = 0;
  if (NL) {
    WarpGroup g{lane};
    uint32_t* tab = NL <= TB_FW_SMEM_LINES ? smem_tab : s.tab + 2 * ub + 16 * (uint64_t)d;
    rep_dedup(g, tab, s.lh + ub, s.ll + ub, NL, &de, &db);
  }
  if (lane) return;
  tb_fw_stats st;
  st.reason_bits = 0; st.flags = fd.flags; st.n_lines = NL; st.n_stop_lines = 0; st.n_short_lines = 0; st.n_words = 0;
  st.n_newlines = 0; st._pad = 0; st.dup_line_bytes = 0; st.n_chars_no_newline = 0;
  uint32_t r = 0;
  if (NL == 0) {
    r = TB_FW_REASON_EMPTY;                                                                      // :79-90
  } else {
    st.n_stop_lines = fd.n_stop;
    const double punct = (double)fd.n_stop / (double)NL;
    if (punct < cfg.line_punct_thr && !(punct == 0.0 && cfg.line_punct_exclude_zero)) {
      r = TB_FW_REASON_LINE_PUNCT;                                                               // :106-123
    } else {
      st.n_short_lines = fd.n_short;
      if ((double)fd.n_short / (double)NL > cfg.short_line_thr) {
        r = TB_FW_REASON_SHORT_LINE;                                                             // :131-146
      } else {
        st.dup_line_bytes = db; st.n_chars_no_newline = fd.n_chars;
        const double dup = fd.n_chars > 0 ? (double)db / (double)fd.n_chars : 0.0;               // :165-169
        if (dup > cfg.char_duplicates_ratio) {
          r = TB_FW_REASON_CHAR_DUP;                                                             // :171-185
        } else {
          st.n_words = fd.n_words; st.n_newlines = fd.n_newlines;
          if (fd.n_words == 0) { if (fd.n_newlines > 0) r = TB_FW_REASON_LIST_NO_WORDS; }         // :191-204
          else if ((double)fd.n_newlines / (double)fd.n_words > cfg.new_line_ratio) r = TB_FW_REASON_LIST_RATIO;   // :206-222
        }
      }
    }
  }
  st.reason_bits = r;
  out[d] = st;
  if (r) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
}
