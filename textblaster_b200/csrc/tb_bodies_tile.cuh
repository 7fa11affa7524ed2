// tb_bodies_tile.cuh -- mask-based document loops of the three word passes (tb_tile.cuh).  Each returns false
// when the document holds a segment the mask form cannot express; the caller then runs the walking loop.
// Included at the end of tb_bodies.cuh.
#pragma once

// segmented inclusive scan of open-fragment summaries; `f`: the lane holds a segment start
TB_D Frag frag_shfl_up(Frag v, int o) { Frag r; r.nbytes = simt_shfl_up(v.nbytes, o); r.meta = simt_shfl_up(v.meta, o); return r; }
TB_D Frag frag_shfl(Frag v, int src) { Frag r; r.nbytes = simt_shfl(v.nbytes, src); r.meta = simt_shfl(v.meta, src); return r; }
// returns the fragment open at the start of this lane's stretch; updates carry to the one open after lane 31
TB_D Frag frag_scan(Frag tail, uint32_t f, int lane, Frag& carry) {
  Frag v = tail;
  for (int off = 1; off < 32; off <<= 1) {
    Frag vu = frag_shfl_up(v, off);
    uint32_t fu = simt_shfl_up(f, off);
    if (lane >= off && !f) { v = frag_cat(vu, v); f = fu; }
  }
  if (!f) v = frag_cat(carry, v);
  Frag cin = frag_shfl_up(v, 1);
  if (lane == 0) cin = carry;
  carry = frag_shfl(v, 31);
  return cin;
}

// ---------------------------------------------------------------------------------------------
// GopherQualityFilter (gopher_quality.rs:73-187)
// ---------------------------------------------------------------------------------------------
// is the word [s, e) a stop word?  (:184-187: stop_words.contains(&word.to_lowercase()))
TB_D_NOINLINE bool gq_is_stop_word(const Classifier& C, const uint8_t* t, const SwTable& sw, int64_t s, int64_t e, bool pure_ascii) {
  const uint64_t len = (uint64_t)(e - s);
  if (pure_ascii) {
    if (len > sw.maxlen || !((sw.lenmask >> (len < 31u ? (uint32_t)len : 31u)) & 1u)) return false;
  } else if (len > 3ull * sw.maxlen) {  // to_lowercase() shrinks a character by at most a factor of three
    return false;
  }
  LowKey k; k.init();
  if (pure_ascii && len <= 8) {  // the common case: the key is the lower-cased word itself
    // eight bytes from the word's start (three aligned loads, realigned), cut to its length, A-Z folded in place
    const uintptr_t addr = (uintptr_t)(t + s);
    const uint32_t* q = (const uint32_t*)(addr & ~(uintptr_t)3);
    const uint32_t mis = (uint32_t)(addr & 3u), l = (uint32_t)len;
    uint32_t w0 = TB_LDG(q), w1 = l + mis > 4u ? TB_LDG(q + 1) : 0u, w2 = l + mis > 8u ? TB_LDG(q + 2) : 0u;
    uint32_t lo = mis ? tb_funnel_r(w0, w1, mis) : w0, hi = mis ? tb_funnel_r(w1, w2, mis) : w1;
    if (l < 4u) { lo &= (1u << (8u * l)) - 1u; hi = 0u; }
    else if (l == 4u) hi = 0u;
    else if (l < 8u) hi &= (1u << (8u * (l - 4u))) - 1u;
    lo |= (((lo + 0x3F3F3F3Fu) & ~(lo + 0x25252525u)) & 0x80808080u) >> 2;  // bytes are ASCII: no carries between them
    hi |= (((hi + 0x3F3F3F3Fu) & ~(hi + 0x25252525u)) & 0x80808080u) >> 2;
    k.lo = ((uint64_t)hi << 32) | lo; k.len = l;
    return sw_lookup(sw, k);
  }
  int64_t i = s;
  while (i < e) { int l; uint32_t cp; uint32_t ci = C.at(t, i, &l, &cp); lowkey_push(C, k, cp, ci); i += l; }
  if (k.sigma) lowkey_slow(C, t, s, e, k);
  return sw_lookup(sw, k);
}

// cheap reject by byte length (exact for pure-ASCII words; to_lowercase() shrinks a character at most threefold)
TB_D bool gq_len_may_match(const SwTable& sw, uint32_t len, bool pure_ascii) {
  if (pure_ascii) return len <= sw.maxlen && ((sw.lenmask >> (len < 31u ? len : 31u)) & 1u);
  return (uint64_t)len <= 3ull * sw.maxlen;
}
// `stage`: 2 * TB_GQ_STAGE words of per-warp scratch (shared memory)
template <class Feed>
TB_D bool gq_doc_tiles(const KBatch& a, const LutEntry* lut, uint32_t* stage, const Classifier& C, const SwTable& sw, const TextRange& R,
                       GqAcc& acc, int lane, Feed& feed) {
  const uint8_t* t = R.t;
  uint32_t* st_pos = stage; uint32_t* st_len = stage + TB_GQ_STAGE;
  Frag carry{0u, 0u};
  feed.begin(t, a.ext_lo, a.ext_hi, tb_tile_origin(R), R.re, lane, a.feed_min_tiles);
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    Tile T;
    const uint32_t sm = feed.tile(base);
    tile_classify<true>(lut, R, a.ext_lo, a.ext_hi, base + lane * TB_STRETCH, T, sm == TB_FEED_NONE ? sm : sm + (uint32_t)(lane * TB_STRETCH));
    feed.done(base, lane);  // (the stretch is in registers now)
    const uint32_t fl = T.dirty ? tile_fix_dirty(C, R, T) : 0u;
    if (simt_any((fl & TILE_EXOTIC) != 0u)) { feed.drain(base); return false; }
    acc.flags |= fl & CI_COMPLEX;
    const int64_t p0 = T.p0;
    const uint32_t V = T.valid, B = T.brk, WC = T.m[TL_WC], LEAD = V & ~T.h[TH_CONT], NAm = T.m[TL_NA], SWF = T.h[TH_SWF];
    const uint32_t H = tb_spread(WC, B), HA = tb_spread(T.al, B);
    const int fb = tb_ctz32(B);                // first segment start of the stretch (32: none)
    const int lb = B ? 31 - tb_clz32(B) : 0;   // last one
    const uint32_t headm = tb_below(fb), tailm = B ? ~tb_below(lb) : 0xFFFFFFFFu, inner = ~headm & ~tailm;
    // ---- segments that start and end inside the stretch
    acc.n_words += (uint32_t)tb_popc32(B & H & inner);
    acc.sum_chars += (uint32_t)tb_popc32(LEAD & H & inner);
    acc.n_alpha += (uint32_t)tb_popc32(B & HA & inner);
    // ---- the segment that was open when the stretch began ends at its first segment start
    Frag tail = frag_of(T, tailm, B ? ((SWF >> lb) & 1u) : 0u);
    Frag cin = frag_scan(tail, B != 0u, lane, carry);
    Frag first{0u, 0u};
    bool first_cand = false;
    if (B) {
      first = frag_cat(cin, frag_of(T, headm, 0u));
      if (first.meta & FR_WC) {
        acc.n_words++; acc.sum_chars += first.meta & FR_CNT;
        if (first.meta & FR_AL) acc.n_alpha++;
        first_cand = (first.meta & FR_SWF) && gq_len_may_match(sw, first.nbytes, !(first.meta & FR_NA));
      }
    }
    // ---- stop-word candidates (first letter and length fit): collected per tile, then tested with all lanes busy
    {
      // first byte fits, and the byte after it can be the second byte of an entry (a one-byte word passes or fails on its
      // length below either way: secondmask is all ones as soon as the list has a one-byte entry)
      const uint32_t cand = B & H & inner & SWF & ((T.h[TH_SW2] | ~V) >> 1);  // (a byte beyond the text filters nothing)
      uint32_t keepc = 0;
      for (uint32_t c = cand; c; c &= c - 1u) {
        const int sb = tb_ctz32(c);
        const int e = tb_ctz32(B & ~tb_upto(sb));
        if (gq_len_may_match(sw, (uint32_t)(e - sb), !(NAm & tb_below(e) & ~tb_below(sb)))) keepc |= 1u << sb;
      }
      uint32_t ctot, k = simt_exscan((uint32_t)tb_popc32(keepc) + (first_cand ? 1u : 0u), lane, &ctot);
      if (first_cand) {
        st_pos[k] = (uint32_t)(p0 + fb - (int64_t)first.nbytes - R.rs);
        st_len[k] = first.nbytes | ((first.meta & FR_NA) ? 0u : 0x80000000u);
        k++;
      }
      for (uint32_t c = keepc; c; c &= c - 1u) {
        const int sb = tb_ctz32(c);
        const int e = tb_ctz32(B & ~tb_upto(sb));
        st_pos[k] = (uint32_t)(p0 + sb - R.rs);
        st_len[k] = (uint32_t)(e - sb) | ((NAm & tb_below(e) & ~tb_below(sb)) ? 0u : 0x80000000u);
        k++;
      }
      simt_sync();
      for (uint32_t j = (uint32_t)lane; j < ctot; j += 32) {
        const int64_t sp = R.rs + (int64_t)st_pos[j];
        const uint32_t l = st_len[j];
        if (gq_is_stop_word(C, t, sw, sp, sp + (int64_t)(l & 0x7FFFFFFFu), (l >> 31) != 0u)) acc.n_stop++;
      }
      simt_sync();
    }
    // ---- symbols (:104,:115)
    acc.n_hash += (uint32_t)tb_popc32(T.h[TH_HASH]);
    {
      const uint32_t DOT = T.h[TH_DOT];
      const uint32_t before = (DOT << 1) | ((T.l1hi >> (4 * TH_DOT)) & 1u);
      uint32_t runs = DOT & ~before & ((DOT >> 1) | 0x80000000u);  // starts of dot runs that may be >= 3 long
      while (runs) {
        const int64_t i = p0 + tb_ctz32(runs);
        runs &= runs - 1u;
        int64_t j = i + 1;
        while (j < R.re && tb_byte(t, j) == '.') j++;
        acc.n_ell += (uint32_t)((j - i) / 3);
      }
      uint32_t e2 = T.h[TH_E2];
      while (e2) {
        const int64_t i = p0 + tb_ctz32(e2);
        e2 &= e2 - 1u;
        if (i + 2 < R.re && tb_byte(t, i + 1) == 0x80 && tb_byte(t, i + 2) == 0xA6) acc.n_ell++;
      }
    }
    // ---- lines (:126-158)
    {
      const uint32_t NLm = T.h[TH_NL];
      uint32_t LS = ((NLm << 1) | ((T.l1hi >> (4 * TH_NL)) & 1u)) & V;
      if (R.rs >= p0 && R.rs < p0 + 32 && R.rs < R.re) LS |= 1u << (uint32_t)(R.rs - p0);
      acc.n_lines += (uint32_t)tb_popc32(LS);
      acc.n_bullet += (uint32_t)tb_popc32(LS & T.h[TH_DASH]);
      uint32_t bl = LS & ~(WC & ~NAm) & ~T.h[TH_DASH];  // a line that starts with an ASCII word character is no bullet line
      while (bl) {
        const int i = tb_ctz32(bl);
        bl &= bl - 1u;
        if (gq_line_is_bullet(C, R, p0 + i)) acc.n_bullet++;
      }
      const uint32_t WC1 = (WC << 1) | ((T.lctx >> (4 * TL_WC + 1)) & 1u);
      uint32_t ends = NLm & ~WC1;  // a line whose last byte is an ASCII word character cannot end with an ellipsis
      while (ends) {
        const int i = tb_ctz32(ends);
        ends &= ends - 1u;
        if (gq_line_ends_with_ellipsis(C, R, p0 + i)) acc.n_ell_lines++;
      }
      const int64_t last = R.re - 1;
      if (last >= p0 && last < p0 + 32 && last >= R.rs) {
        const uint32_t bit = 1u << (uint32_t)(last - p0);
        if (!(NLm & bit) && !(WC & bit & ~NAm) && gq_line_ends_with_ellipsis(C, R, R.re)) acc.n_ell_lines++;
      }
    }
  }
  if (lane == 0 && (carry.meta & FR_WC)) {  // the segment that runs to the end of the text
    acc.n_words++; acc.sum_chars += carry.meta & FR_CNT;
    if (carry.meta & FR_AL) acc.n_alpha++;
    if ((carry.meta & FR_SWF) && gq_is_stop_word(C, t, sw, R.re - (int64_t)carry.nbytes, R.re, !(carry.meta & FR_NA))) acc.n_stop++;
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// C4QualityFilter pass B (c4_filters.rs:207-250): words per processed line, longest word, line flags
// ---------------------------------------------------------------------------------------------
template <class Feed>
TB_D bool c4b_doc_tiles(const KBatch& a, const LutEntry* lut, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s,
                        const TextRange& R, uint64_t lb, int lane, Feed& feed) {
  const uint8_t* t = R.t;
  const bool check_len = cfg.max_word_length > 0;
  const bool small = check_len && cfg.max_word_length < 32;  // words inside one stretch can be too long
  const bool pats = cfg.filter_javascript != 0 || cfg.filter_policy != 0;
  Frag carry{0u, 0u};
  uint32_t nl = 0;
  feed.begin(t, a.ext_lo, a.ext_hi, tb_tile_origin(R), R.re, lane, a.feed_min_tiles);
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    Tile T;
    const uint32_t sm = feed.tile(base);
    tile_classify<true>(lut, R, a.ext_lo, a.ext_hi, base + lane * TB_STRETCH, T, sm == TB_FEED_NONE ? sm : sm + (uint32_t)(lane * TB_STRETCH));
    feed.done(base, lane);
    const uint32_t fl = T.dirty ? tile_fix_dirty(C, R, T) : 0u;
    if (simt_any((fl & TILE_EXOTIC) != 0u)) { feed.drain(base); return false; }
    const int64_t p0 = T.p0;
    const uint32_t V = T.valid, B = T.brk, WC = T.m[TL_WC], LEAD = V & ~T.h[TH_CONT], NLm = T.h[TH_NL];
    uint32_t tot, pre = simt_exscan((uint32_t)tb_popc32(NLm), lane, &tot);
    const uint64_t lbase = lb + nl + pre;  // line of the first byte of the stretch
    const uint32_t H = tb_spread(WC, B);
    const int fb = tb_ctz32(B);
    const int lbk = B ? 31 - tb_clz32(B) : 0;
    const uint32_t headm = tb_below(fb), tailm = B ? ~tb_below(lbk) : 0xFFFFFFFFu, inner = ~headm & ~tailm;
    Frag tail = frag_of(T, tailm, 0u);
    Frag cin = frag_scan(tail, B != 0u, lane, carry);
    // word ends: bit j = a word ends just before p0 + j (its line: the '\n' count in front of j)
    uint32_t WE = B & (H << 1) & ~tb_upto(fb);
    if (B) {
      Frag first = frag_cat(cin, frag_of(T, headm, 0u));
      if (first.meta & FR_WC) {
        WE |= 1u << fb;
        const uint32_t cnt = first.meta & FR_CNT;
        if (check_len && (small || cnt > 31u)) simt_atomic_max(&s.lmax[lbase + (uint32_t)tb_popc32(NLm & headm)], cnt);
      }
    }
    {
      uint32_t rem = WE, nlm = NLm;
      uint64_t li = lbase;
      while (rem) {
        const uint32_t part = rem & tb_upto(tb_ctz32(nlm));
        if (part) simt_atomic_add(&s.lwc[li], (uint32_t)tb_popc32(part));
        rem &= ~part;
        nlm &= nlm - 1u;
        li++;
      }
    }
    if (small) {
      uint32_t starts = B & H & inner;
      while (starts) {
        const int sb = tb_ctz32(starts);
        starts &= starts - 1u;
        const int e = tb_ctz32(B & ~tb_upto(sb));
        const uint32_t cnt = (uint32_t)tb_popc32(LEAD & tb_below(e) & ~tb_below(sb));
        simt_atomic_max(&s.lmax[lbase + (uint32_t)tb_popc32(NLm & tb_below(sb))], cnt);
      }
    }
    // line flags: terminal punctuation of the lines that end here, javascript / policy patterns
    {
      uint32_t m = NLm, k = 0;
      while (m) {
        const int i = tb_ctz32(m);
        m &= m - 1u;
        const uint32_t f = c4_line_end_flags(C, R, p0 + i);
        if (f) simt_atomic_or(&s.lflags[lbase + k], f);
        k++;
      }
      const int64_t last = R.re - 1;
      if (last >= p0 && last < p0 + 32 && last >= R.rs && !((NLm >> (uint32_t)(last - p0)) & 1u)) {
        const uint32_t f = c4_line_end_flags(C, R, R.re);
        if (f) simt_atomic_or(&s.lflags[lbase + k], f);
      }
      if (pats) {
        uint32_t cand = T.h[TH_P1] & ((T.h[TH_P2] >> 1) | 0x80000000u) & ((T.h[TH_P3] >> 2) | 0xC0000000u);
        while (cand) {
          const int i = tb_ctz32(cand);
          cand &= cand - 1u;
          const uint32_t f = c4_pattern_flags(t, p0 + i, R.re, tb_byte(t, p0 + i), cfg.filter_javascript != 0, cfg.filter_policy != 0);
          if (f) simt_atomic_or(&s.lflags[lbase + (uint32_t)tb_popc32(NLm & tb_below(i))], f);
        }
      }
    }
    nl += tot;
  }
  if (lane == 0 && (carry.meta & FR_WC)) {  // the word that runs to the end of the text
    simt_atomic_add(&s.lwc[lb + nl], 1u);
    if (check_len) simt_atomic_max(&s.lmax[lb + nl], carry.meta & FR_CNT);
  }
  return true;
}

// ---------------------------------------------------------------------------------------------
// GopherRepetitionFilter, pass 1 (gopher_rep.rs:57-58,80-84,120-124,159): word and line-piece hashes
// ---------------------------------------------------------------------------------------------
TB_D HL hl_fold(uint32_t a, uint32_t b, uint32_t len) { HL r; r.a = tb_hfold(a); r.b = tb_hfold(b); r.len = len; return r; }
// hash of a segment from the prefix hashes at its two ends: pe - ps * B^len mod p (pe, ps lazily reduced, pw <= p) -> <= p
TB_D uint32_t tb_hseg(uint32_t pe, uint32_t ps, uint32_t pw) {
  const uint32_t r = tb_hfold(pe) + (TB_P31 - tb_mm(tb_hfold(ps), pw));
  return (r & TB_P31) + (r >> 31);
}
// the low nibbles of the four bytes of x, the first byte (bits 0-7) in the top nibble of the 16-bit result
TB_D uint32_t tb_nib16(uint32_t x) {
  x &= 0x0F0F0F0Fu;
  x = ((x & 0x000000FFu) << 24) | ((x & 0x0000FF00u) << 8) | ((x >> 8) & 0x0000FF00u) | (x >> 24);  // bytes reversed
  x = (x | (x >> 4)) & 0x00FF00FFu;
  return (x | (x >> 8)) & 0xFFFFu;
}
// suffix signature (tb_sig_push) of the bytes [s, e) of a stretch, from its nibble string (byte j in nibble 31 - j of the
// 128-bit number n[3] : n[2] : n[1] : n[0]): the nibbles of the last min(16, e - s) bytes, the last byte lowest
TB_D uint64_t tb_nib_sig(const uint32_t* n, int s, int e) {
  if (e <= s) return 0ull;
  const int sh = 4 * (32 - e);  // the string shifted right by this has byte e - 1 in its low nibble
  const int w = sh >> 5, r = sh & 31;
  uint32_t x0 = n[0], x1 = n[1], x2 = n[2], x3 = n[3];
  if (w & 2) { x0 = x2; x1 = x3; x2 = 0u; x3 = 0u; }
  if (w & 1) { x0 = x1; x1 = x2; x2 = x3; }
  const uint32_t lo = r ? (x0 >> r) | (x1 << (32 - r)) : x0, hi = r ? (x1 >> r) | (x2 << (32 - r)) : x1;
  return (((uint64_t)hi << 32) | lo) & tb_sig_mask((uint32_t)(e - s));
}
// word fragments also carry their suffix signature (tb_sig_push)
struct HLS { HL h; uint64_t sig; };
TB_D HLS hls_cat(HLS x, HLS y, const uint64_t* powtab) { HLS r; r.sig = tb_sig_cat(x.sig, y.sig, y.h.len); r.h = hl_cat(x.h, y.h, powtab); return r; }
TB_D HLS hls_shfl_up(HLS v, int o) { HLS r; r.h = hl_shfl_up(v.h, o); r.sig = (uint64_t)simt_shfl_up64((int64_t)v.sig, o); return r; }
TB_D HLS hls_shfl(HLS v, int src) { HLS r; r.h = hl_shfl(v.h, src); r.sig = simt_shfl(v.sig, src); return r; }

template <class Feed>
TB_D bool gr_doc_tiles(const KBatch& a, const LutEntry* lut, uint32_t* stage, const Classifier& C, const TextRange& R,
                       const GrScratch& s, int64_t d, int lane, GrTotals& out, Feed& feed) {
  uint64_t* st_s = (uint64_t*)stage;
  uint32_t* st_a = stage + 2 * TB_GR_STAGE_WORDS; uint32_t* st_b = stage + 3 * TB_GR_STAGE_WORDS; uint32_t* st_n = stage + 4 * TB_GR_STAGE_WORDS;
  uint64_t* pfx = (uint64_t*)(stage + 5 * TB_GR_STAGE_WORDS);  // [32][32]: the prefix hashes of the lanes' stretches, a column per lane
  const uint64_t uw = tb_wunit(R.rs, d), ul = tb_lunit(R.rs, d);
  uint64_t* wh = s.wh + uw; uint32_t* wl = s.wl + uw; uint64_t* wsg = s.ws + uw;
  uint64_t* lh = s.lh + ul; uint32_t* ll = s.ll + ul; uint8_t* lg = s.lg + ul;
  const uint64_t* powtab = a.uc.powtab;
  const uint32_t base1 = (uint32_t)(TB_LDG(powtab + 1) >> 32), base2 = (uint32_t)TB_LDG(powtab + 1);  // the context's hash bases
  const uint32_t below_lane = tb_below(lane);
  uint32_t W = 0, NLc = 0, flags = 0;
  uint64_t nchars = 0;
  HLS wcarry{{0u, 0u, 0u}, 0ull};
  HL lcarry{0u, 0u, 0u};
  uint64_t sig_carry = 0;  // signature of the text's words up to the last one stored (its last 16 bytes)
  uint32_t wcarry_wc = 0;
  feed.begin(R.t, a.ext_lo, a.ext_hi, tb_tile_origin(R), R.re, lane, a.feed_min_tiles);
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    Tile T;
    const uint32_t sm = feed.tile(base);
    tile_classify<true>(lut, R, a.ext_lo, a.ext_hi, base + lane * TB_STRETCH, T, sm == TB_FEED_NONE ? sm : sm + (uint32_t)(lane * TB_STRETCH));
    feed.done(base, lane);
    const uint32_t fl = T.dirty ? tile_fix_dirty(C, R, T) : 0u;
    if (simt_any((fl & TILE_EXOTIC) != 0u)) { feed.drain(base); return false; }
    flags |= fl & CI_COMPLEX;
    const int64_t p0 = T.p0;
    const uint32_t V = T.valid, B = T.brk, WC = T.m[TL_WC];
    const uint32_t NLx = T.h[TH_NL] | ~V;  // every byte outside the text ends a piece
    nchars += (uint32_t)tb_popc32(V & ~T.h[TH_CONT]);
    const uint32_t H = tb_spread(WC, B);
    const int fb = tb_ctz32(B);
    const int lbk = B ? 31 - tb_clz32(B) : 0;
    const uint32_t headm = tb_below(fb), tailm = B ? ~tb_below(lbk) : 0xFFFFFFFFu;
    // ---- which lanes close a word with their first segment start: "word character seen" of the open fragment
    const uint32_t brk_lanes = simt_ballot(B != 0u), twc_lanes = simt_ballot((WC & tailm) != 0u);
    uint32_t in_wc;
    {
      const uint32_t lower = brk_lanes & below_lane;
      if (lower) in_wc = (twc_lanes & below_lane & ~tb_below(31 - tb_clz32(lower))) != 0u;
      else in_wc = wcarry_wc | ((twc_lanes & below_lane) != 0u);
    }
    const uint32_t first_word = (B && (in_wc || (WC & headm))) ? 1u : 0u;
    const uint32_t WEin = B & (H << 1) & ~tb_upto(fb);  // word ends after the first segment start
    uint32_t wtot, wpre = simt_exscan((uint32_t)tb_popc32(WEin) + first_word, lane, &wtot);
    if (wtot > TB_GR_STAGE_WORDS) { feed.drain(base); return false; }  // more words than the staging arrays hold: the walking path takes the document
    // ---- line pieces: is the piece that is open at the start of the stretch non-empty?
    const int fn = tb_ctz32(NLx);
    const int ln_ = NLx ? 31 - tb_clz32(NLx) : -1;
    const uint32_t nl_lanes = simt_ballot(NLx != 0u), tnz_lanes = simt_ballot(ln_ < 31);
    uint32_t in_nz;
    {
      const uint32_t lower = nl_lanes & below_lane;
      if (lower) in_nz = (tnz_lanes & below_lane & ~tb_below(31 - tb_clz32(lower))) != 0u;
      else in_nz = (lcarry.len != 0u) | ((tnz_lanes & below_lane) != 0u);
    }
    const uint32_t first_piece = (NLx && (in_nz || fn > 0)) ? 1u : 0u;
    const uint32_t PEin = NLx & ~(NLx << 1) & ~tb_upto(fn);  // a later '\n' that ends a non-empty piece
    uint32_t ltot, lpre = simt_exscan((uint32_t)tb_popc32(PEin) + first_piece, lane, &ltot);
    // ---- the stretch's 32 prefix hashes P_1 .. P_32 (P_j covers the bytes [0, j); no restarts, no branches) go to the
    // lane's column of `pfx`; the hash of any segment [s, e) of the stretch is then P_e - P_s * B^(e-s), taken only
    // where a word or a line piece ends.  Its length is e - s, its signature a piece of the stretch's nibble string.
    uint64_t* const pcol = pfx + lane;  // P_j at pcol[(j - 1) * 32]
    uint32_t P32a, P32b;
    {
      uint32_t wq[8];
#pragma unroll
      for (int k = 0; k < 8; k++) wq[k] = T.w[k];
      uint32_t pa = 0, pb = 0;
      uint64_t* pc = pcol;
#pragma unroll 1
      for (int k = 0; k < 8; k++) {  // (eight rounds of four bytes; the words rotate so that the round body is one piece of code)
        const uint32_t x = wq[0];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const uint32_t byte = (x >> (8 * j)) & 0xFFu;
          pa = tb_hpush(pa, base1, byte + 1u); pb = tb_hpush(pb, base2, byte + 1u);
          pc[j * 32] = ((uint64_t)pa << 32) | pb;
        }
        pc += 4 * 32;
#pragma unroll
        for (int q = 0; q < 7; q++) wq[q] = wq[q + 1];
      }
      P32a = pa; P32b = pb;
    }
    uint32_t nib[4];  // the low nibbles of the 32 bytes, byte 0 in the top nibble of nib[3], byte 31 in the low nibble of nib[0]
#pragma unroll
    for (int k = 0; k < 4; k++) nib[3 - k] = (tb_nib16(T.w[2 * k]) << 16) | tb_nib16(T.w[2 * k + 1]);
    // (a lane reads back its own column only: no barrier)
    uint32_t wa, wb, wn, la, lb_, ln;
    uint32_t hwa = 0, hwb = 0, hwn = 0, hla = 0, hlb = 0, hln = 0;
    uint64_t wsig, hwsig = 0;  // suffix signature of the fragment after the last segment start / of the one before the first
    {
      // ---- words: the fragment before the first segment start, the words that end inside the stretch, the open tail
      if (B) {
        if (fb) { const uint64_t pe = pcol[(fb - 1) * 32]; hwa = (uint32_t)(pe >> 32); hwb = (uint32_t)pe; }
        hwn = (uint32_t)fb; hwsig = tb_nib_sig(nib, 0, fb);
        uint32_t widx = wpre + first_word;  // slot in the tile's staging arrays
        uint32_t we = WEin;
#pragma unroll 1
        while (we) {
          const int e = tb_ctz32(we);
          we &= we - 1u;
          const int sp = 31 - tb_clz32(B & tb_below(e));  // the segment start before e (e > fb: there is one)
          const uint64_t pe = pcol[(e - 1) * 32], ps = sp ? pcol[(sp - 1) * 32] : 0ull, pw = TB_LDG(powtab + (e - sp));
          st_a[widx] = tb_hseg((uint32_t)(pe >> 32), (uint32_t)(ps >> 32), (uint32_t)(pw >> 32));
          st_b[widx] = tb_hseg((uint32_t)pe, (uint32_t)ps, (uint32_t)pw);
          st_n[widx] = (uint32_t)(e - sp); st_s[widx] = tb_nib_sig(nib, sp, e);
          widx++;
        }
        const uint64_t ps = lbk ? pcol[(lbk - 1) * 32] : 0ull, pw = TB_LDG(powtab + (32 - lbk));
        wa = tb_hseg(P32a, (uint32_t)(ps >> 32), (uint32_t)(pw >> 32)); wb = tb_hseg(P32b, (uint32_t)ps, (uint32_t)pw);
        wn = (uint32_t)(32 - lbk); wsig = tb_nib_sig(nib, lbk, 32);
      } else {
        wa = P32a; wb = P32b; wn = 32u; wsig = tb_nib_sig(nib, 0, 32);
      }
      // ---- line pieces ('\n' bytes and the bytes outside the text belong to no piece)
      if (NLx) {
        if (fn) { const uint64_t pe = pcol[(fn - 1) * 32]; hla = (uint32_t)(pe >> 32); hlb = (uint32_t)pe; }
        hln = (uint32_t)fn;
        uint32_t lidx = NLc + lpre + first_piece;
        uint32_t pe_ = PEin;
#pragma unroll 1
        while (pe_) {
          const int e = tb_ctz32(pe_);
          pe_ &= pe_ - 1u;
          const int sp = 32 - tb_clz32(NLx & tb_below(e));  // the byte after the '\n' before e (e > fn: there is one)
          const uint64_t pe = pcol[(e - 1) * 32], ps = pcol[(sp - 1) * 32], pw = TB_LDG(powtab + (e - sp));
          HL piece; piece.a = tb_hseg((uint32_t)(pe >> 32), (uint32_t)(ps >> 32), (uint32_t)(pw >> 32));
          piece.b = tb_hseg((uint32_t)pe, (uint32_t)ps, (uint32_t)pw); piece.len = (uint32_t)(e - sp);
          lh[lidx] = hl_pack(piece); ll[lidx] = piece.len; lg[lidx] = (uint8_t)gr_gap2(R, p0 + sp);
          lidx++;
        }
        if (ln_ < 31) {
          const uint64_t ps = pcol[ln_ * 32], pw = TB_LDG(powtab + (31 - ln_));
          la = tb_hseg(P32a, (uint32_t)(ps >> 32), (uint32_t)(pw >> 32)); lb_ = tb_hseg(P32b, (uint32_t)ps, (uint32_t)pw);
        } else { la = 0; lb_ = 0; }
        ln = (uint32_t)(31 - ln_);
      } else {
        la = P32a; lb_ = P32b; ln = 32u;
      }
    }
    // ---- words: hash of the fragment open at the start of the stretch
    {
      HLS v{hl_fold(wa, wb, wn), wsig};
      HLS cin;
      if (brk_lanes == 0xFFFFFFFFu) {  // every lane restarts: the open fragment is the previous lane's tail
        cin = hls_shfl_up(v, 1);
        if (lane == 0) cin = wcarry;
        wcarry = hls_shfl(v, 31);
      } else {
        uint32_t f = B != 0u;
        for (int off = 1; off < 32; off <<= 1) {
          HLS vu = hls_shfl_up(v, off);
          uint32_t fu = simt_shfl_up(f, off);
          if (lane >= off && !f) { v = hls_cat(vu, v, powtab); f = fu; }
        }
        if (!f) v = hls_cat(wcarry, v, powtab);
        cin = hls_shfl_up(v, 1);
        if (lane == 0) cin = wcarry;
        wcarry = hls_shfl(v, 31);
      }
      if (first_word) {
        HLS first = hls_cat(cin, HLS{hl_fold(hwa, hwb, hwn), hwsig}, powtab);
        st_a[wpre] = first.h.a; st_b[wpre] = first.h.b; st_n[wpre] = first.h.len; st_s[wpre] = first.sig;
      }
      // the tile's words, in order: reduce and store them with all lanes busy
      simt_sync();
      // (the signature stored with a word covers the last 16 bytes of all the words up to it: short words take
      // the missing bytes from the words before them -- in the staging arrays, or before this tile: sig_carry)
      uint64_t last_sig = 0;
      for (uint32_t j = (uint32_t)lane; j < wtot; j += 32) {
        wh[W + j] = hl_pack(hl_fold(st_a[j], st_b[j], st_n[j])); wl[W + j] = st_n[j];
        uint64_t sig = st_s[j];
        uint32_t filled = st_n[j], k = j;
        while (filled < 16u && k > 0u) { k--; sig |= st_s[k] << (4u * filled); filled += st_n[k]; }
        if (filled < 16u) sig |= sig_carry << (4u * filled);
        wsg[W + j] = sig; last_sig = sig;
      }
      if (wtot) sig_carry = simt_shfl(last_sig, (int)((wtot - 1u) & 31u));
      simt_sync();
      // "word character seen" of the fragment open after lane 31
      if (brk_lanes) wcarry_wc = (twc_lanes & ~tb_below(31 - tb_clz32(brk_lanes))) != 0u;
      else wcarry_wc |= twc_lanes != 0u;
    }
    // ---- line pieces
    {
      HL v = hl_fold(la, lb_, ln);
      uint32_t f = NLx != 0u;
      for (int off = 1; off < 32; off <<= 1) {
        HL vu = hl_shfl_up(v, off);
        uint32_t fu = simt_shfl_up(f, off);
        if (lane >= off && !f) { v = hl_cat(vu, v, powtab); f = fu; }
      }
      if (!f) v = hl_cat(lcarry, v, powtab);
      HL cin = hl_shfl_up(v, 1);
      if (lane == 0) cin = lcarry;
      lcarry = hl_shfl(v, 31);
      if (first_piece) {
        HL first = hl_cat(cin, hl_fold(hla, hlb, hln), powtab);
        const uint32_t k = NLc + lpre;
        lh[k] = hl_pack(first); ll[k] = first.len; lg[k] = (uint8_t)gr_gap2(R, p0 + fn - (int64_t)first.len);
      }
    }
    W += wtot; NLc += ltot;
  }
  if (wcarry_wc && wcarry.h.len) {  // the word / the piece that runs to the end of the text
    if (lane == 0) { wh[W] = hl_pack(wcarry.h); wl[W] = wcarry.h.len; wsg[W] = tb_sig_cat(sig_carry, wcarry.sig, wcarry.h.len); }
    W++;
  }
  if (lcarry.len) {
    if (lane == 0) { lh[NLc] = hl_pack(lcarry); ll[NLc] = lcarry.len; lg[NLc] = (uint8_t)gr_gap2(R, R.re - lcarry.len); }
    NLc++;
  }
  out.W = W; out.NL = NLc; out.flags = flags; out.nchars = nchars;
  return true;
}

// ---------------------------------------------------------------------------------------------
// Compacting store: the bytes of the tile selected by K, in order, to dst.  They are gathered in a
// per-warp staging buffer first (whole words when the stretch is kept entirely) and leave it as
// aligned 16-byte stores.  Returns the number of bytes written.
// ---------------------------------------------------------------------------------------------
TB_D void tb_copy16(uint8_t* d, const uint8_t* s) {
#if defined(__CUDA_ARCH__)
  *(uint4*)d = *(const uint4*)s;
#else
  memcpy(d, s, 16);
#endif
}
TB_D uint32_t tile_store_kept(const uint32_t* w, uint32_t K, uint8_t* stage, uint8_t* dst, int lane) {
  uint32_t tot;
  const uint32_t pre = simt_exscan((uint32_t)tb_popc32(K), lane, &tot);
  const uint32_t sh = (uint32_t)((uintptr_t)dst & 15u);  // stage + sh + k and dst + k are aligned alike
  uint8_t* sp = stage + sh + pre;
  if (K == 0xFFFFFFFFu) {
    const uint32_t m = (uint32_t)((uintptr_t)sp & 3u);
    if (m == 0u) {
      uint32_t* q = (uint32_t*)sp;
#pragma unroll
      for (int k = 0; k < 8; k++) q[k] = w[k];
    } else {
      const uint32_t hb = 4u - m;
      for (uint32_t j = 0; j < hb; j++) sp[j] = (uint8_t)(w[0] >> (8u * j));
      uint32_t* q = (uint32_t*)(sp + hb);
#pragma unroll
      for (int k = 0; k < 7; k++) q[k] = tb_funnel_r(w[k], w[k + 1], hb);
      for (uint32_t j = 0; j < m; j++) sp[32u - m + j] = (uint8_t)(w[7] >> (8u * (hb + j)));
    }
  } else if (K) {
#pragma unroll
    for (int i = 0; i < 32; i++)
      if ((K >> i) & 1u) *sp++ = (uint8_t)(w[i >> 2] >> (8 * (i & 3)));
  }
  simt_sync();
  const uint8_t* src = stage + sh;
  uint32_t head = (16u - sh) & 15u;
  if (head > tot) head = tot;
  if ((uint32_t)lane < head) dst[lane] = src[lane];
  const uint32_t body = (tot - head) & ~15u;
  for (uint32_t c = head + 16u * (uint32_t)lane; c < head + body; c += 512u) tb_copy16(dst + c, src + c);
  for (uint32_t i = head + body + (uint32_t)lane; i < tot; i += 32u) dst[i] = src[i];
  simt_sync();
  return tot;
}

// ---------------------------------------------------------------------------------------------
// C4QualityFilter pass A (c4_filters.rs:150-151,175-199): document flags, line count, and the processed
// lines P: every line trimmed, citations removed, still joined by the original '\n'.
// ---------------------------------------------------------------------------------------------
template <class Feed>
TB_D void c4a_doc_tiles(const KBatch& a, const LutEntry* lut, uint8_t* stage, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s,
                        const TextRange& R, int64_t d, int lane, Feed& feed) {
  const uint8_t* t = R.t;
  uint8_t* out = s.P + R.rs;
  const uint64_t lb = (uint64_t)R.rs + (uint64_t)d;
  const uint32_t below_lane = tb_below(lane), upto_lane = tb_upto(lane);
  uint32_t flags = 0, nl = 0;
  int64_t outpos = 0;
  uint32_t carry_seen = 0;  // the current line showed a non-white-space character before this tile
  int64_t carry_cit = -1;   // end of a citation that runs into this tile
  if (lane == 0) { s.lwc[lb] = 0; s.lmax[lb] = 0; s.lflags[lb] = 0; }
  feed.begin(t, a.ext_lo, a.ext_hi, tb_tile_origin(R), R.re, lane, a.feed_min_tiles);
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    Tile T;
    const int64_t p0 = base + lane * TB_STRETCH;
    const uint32_t sm = feed.tile(base);
    tile_gather<true>(lut, R, a.ext_lo, a.ext_hi, p0, T, sm == TB_FEED_NONE ? sm : sm + (uint32_t)(lane * TB_STRETCH));
    feed.done(base, lane);
    const uint32_t V = T.valid, NLm = T.m[CA_NL], NAm = T.m[CA_NA];
    uint32_t WSm = T.m[CA_WS];
    if (NAm) {  // non-ASCII White_Space: only a few lead bytes can begin one
      const int fv = tb_ctz32(V);
      if (((T.m[CA_CONT] >> fv) & 1u) && p0 + fv > R.rs) {  // the stretch begins inside a character
        const int64_t k = C.prev_lead(t, p0 + fv, R.rs);
        int len; uint32_t cp;
        if (C.at(t, k, &len, &cp) & CI_WS) { const int64_t e = k + len - p0; WSm |= tb_below(e > 32 ? 32 : (int)e) & ~tb_below(fv); }
      }
      uint32_t cand = T.h[CH_WSLEAD];
      while (cand) {
        const int i = tb_ctz32(cand);
        cand &= cand - 1u;
        int len; uint32_t cp;
        if (C.at(t, p0 + i, &len, &cp) & CI_WS) WSm |= tb_below(i + len > 32 ? 32 : i + len) & ~tb_below(i);
      }
    }
    const uint32_t NW = V & ~WSm;
    if (T.m[CA_CURLY]) flags |= C4F_CURLY;
    if (cfg.filter_lorem_ipsum) {
      uint32_t c = T.m[CA_LL] & ((T.m[CA_LO] >> 1) | 0x80000000u);
      while (c) {
        const int i = tb_ctz32(c);
        c &= c - 1u;
        if (c4_fold_match(t, p0 + i, R.re, "lorem ipsum", 11)) flags |= C4F_LOREM;
      }
    }
    // ---- citations: every '[' is tried on its own (a match holds no '[' after its first byte)
    uint32_t cit = 0;
    if (cfg.remove_citations) {
      int64_t my_end = -1;
      uint32_t c = T.m[CA_LBR];
      while (c) {
        const int i = tb_ctz32(c);
        c &= c - 1u;
        const int64_t e = c4_citation_end(C, R, p0 + i);
        if (e > 0) { cit |= tb_below(e - p0 > 32 ? 32 : (int)(e - p0)) & ~tb_below(i); if (e > my_end) my_end = e; }
      }
      if (simt_any(my_end > p0 + 32) || carry_cit > base) {  // a citation runs past its stretch
        int64_t v = my_end;
        for (int off = 1; off < 32; off <<= 1) { int64_t u = simt_shfl_up64(v, off); if (lane >= off && u > v) v = u; }
        int64_t inc = simt_shfl_up64(v, 1);
        if (lane == 0) inc = -1;
        if (carry_cit > inc) inc = carry_cit;
        if (inc > p0) cit |= tb_below(inc - p0 > 32 ? 32 : (int)(inc - p0));
        const int64_t last = (int64_t)simt_shfl((uint64_t)v, 31);
        if (last > carry_cit) carry_cit = last;
      }
    }
    // ---- which white space is inside its (trimmed) line: a non-white-space character on both sides
    const uint32_t LSb = NLm << 1;  // line starts inside the stretch
    const int fnl = tb_ctz32(NLm);
    const uint32_t headm = tb_upto(fnl), tailm = NLm ? ~tb_upto(31 - tb_clz32(NLm)) : 0xFFFFFFFFu;
    uint32_t seenL = tb_spread_up(NW, LSb), seenR = tb_spread_down(NW, LSb);
    const uint32_t nl_lanes = simt_ballot(NLm != 0u), tnw_lanes = simt_ballot((NW & tailm) != 0u);
    {
      const uint32_t lower = nl_lanes & below_lane;
      uint32_t in_seen;
      if (lower) in_seen = (tnw_lanes & below_lane & ~tb_below(31 - tb_clz32(lower))) != 0u;
      else in_seen = carry_seen | ((tnw_lanes & below_lane) != 0u);
      if (in_seen) seenL |= headm;
    }
    {
      const uint32_t ev = NW | NLm;
      const uint32_t ev_lanes = simt_ballot(ev != 0u), nwfirst_lanes = simt_ballot((ev & (0u - ev) & NW) != 0u);
      const uint32_t higher = ev_lanes & ~upto_lane;
      const uint32_t unres = V & WSm & ~NLm & tailm & ~seenR;
      uint32_t beyond = 0;
      if (simt_any(unres != 0u && higher == 0u)) {  // the answer lies beyond the tile: look ahead (white-space runs are short)
        int64_t j = base + TB_TILE;
        while (j < R.re && tb_is_cont(tb_byte(t, j))) j++;
        while (j < R.re) {
          int len; uint32_t cp;
          const uint32_t ci = C.at(t, j, &len, &cp);
          if (cp == '\n') break;
          if (!(ci & CI_WS)) { beyond = 1; break; }
          j += len;
        }
      }
      const uint32_t after_nw = higher ? ((nwfirst_lanes >> tb_ctz32(higher)) & 1u) : beyond;
      if (after_nw) seenR |= tailm;
    }
    if (nl_lanes) carry_seen = (tnw_lanes & ~tb_below(31 - tb_clz32(nl_lanes))) != 0u;
    else carry_seen |= tnw_lanes != 0u;
    // ---- kept bytes
    uint32_t K = (NW | (WSm & ~NLm & seenL & seenR)) & ~cit & V;
    uint32_t nlk = NLm;
    {
      const int64_t last = R.re - 1;
      if (last >= p0 && last < p0 + 32) nlk &= ~(1u << (uint32_t)(last - p0));  // the '\n' that ends the document
    }
    K |= nlk;
    tb_bits_store(s.ka, p0, (uint32_t)((uintptr_t)t & 31u), V, K);
    outpos += tile_store_kept(T.w, K, stage, out + outpos, lane);
    // zero the slot of every line that starts after a '\n' of this stretch
    const uint32_t cnt = (uint32_t)tb_popc32(NLm);
    uint32_t ntot, npre = simt_exscan(cnt, lane, &ntot);
    for (uint32_t k = 0; k < cnt; k++) { const uint64_t li = lb + nl + npre + k + 1; s.lwc[li] = 0; s.lmax[li] = 0; s.lflags[li] = 0; }
    nl += ntot;
  }
  flags = simt_or(flags);
  if (lane == 0) {
    const uint32_t lines = nl + ((R.re > R.rs && t[R.re - 1] != '\n') ? 1u : 0u);  // str::lines().count()
    s.plen[d] = outpos; s.nlines[d] = lines;
    uint32_t df = 0;
    if (cfg.filter_lorem_ipsum && (flags & C4F_LOREM)) df |= C4F_LOREM;
    if (cfg.filter_curly_bracket && (flags & C4F_CURLY)) df |= C4F_CURLY;
    s.docflags[d] = df;
    s.prs[d] = R.rs; s.pre[d] = R.rs + outpos;
  }
}
