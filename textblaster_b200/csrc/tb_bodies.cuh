// tb_bodies.cuh -- warp-level bodies of the sm_100a kernels (one warp per document).
//
// Execution shape shared by every text pass: the document range is walked in 1 KiB tiles, each lane
// owning one 32-byte stretch of the tile (tb_lanes.cuh holds the per-lane work).  Lane results are
// combined with shuffles, reductions and (segmented) warp scans through the tb_simt.cuh primitives,
// so the same bodies run under the host emulation harness (tests/host_emul, test-only).
// No tensor cores (byte/integer work); every pass reads its text once from HBM.
#pragma once
#include "tb_kernels.cuh"
#include "tb_rep.cuh"
#include "tb_simt.cuh"
#include "tb_tile.cuh"
#include "tb_entities.cuh"

#if defined(TB_HOST_EMUL)
extern uint64_t tbe_stat_tiles, tbe_stat_walk;  // documents taken by the mask path / by the walking path (lane 0 counts)
#define TBE_COUNT(var, lane) do { if ((lane) == 0) __atomic_fetch_add(&(var), 1, __ATOMIC_RELAXED); } while (0)
#else
#define TBE_COUNT(var, lane) do { } while (0)
#endif

#define TB_BLOCK 32  // one warp per CTA: a document never waits for a longer neighbour in its CTA

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
TB_D_NOINLINE void gq_doc_walk(const Classifier& C, const TextRange& R, const SwTable& sw, GqAcc& acc, int lane) {
  for (int64_t base = R.rs; base < R.re; base += TB_TILE) {
    int64_t p0 = base + lane * TB_STRETCH;
    if (p0 < R.re) { int64_t p1 = p0 + TB_STRETCH < R.re ? p0 + TB_STRETCH : R.re; gq_lane(C, R, sw, p0, p1, acc); }
  }
}
#define TB_GQ_STAGE (23 * 32)  // stop-word candidates one tile can hold: a stretch ends at most 22 words (two per three bytes) + 1
template <class Feed>
TB_D bool gq_doc_tiles(const KBatch& a, const LutEntry* lut, uint32_t* stage, const Classifier& C, const SwTable& sw,
                       const TextRange& R, GqAcc& acc, int lane, Feed& feed);
// `stage`: 2 * TB_GQ_STAGE words of per-warp scratch (shared memory)
// `feed`: the warp's bulk-copy ring for the tiles of long documents (TileFeed, tb_feed.cuh), or NoFeed: plain loads only
template <class Feed = NoFeed>
TB_D void body_gq(const KBatch& a, const uint16_t* s_ascii, const LutEntry* lut, uint32_t* stage, const SwTable& sw, const GqCfg& cfg,
                  int step, tb_gq_stats* out, int64_t d, int lane, Feed* feed = nullptr) {
  Feed nofeed{};
  Feed& F = feed ? *feed : nofeed;
  Classifier C{s_ascii, a.uc};
  C.key_maxlen = sw.maxlen; C.key_lenmask = sw.lenmask; C.key_firstmask = sw.firstmask;
  TextRange R{a.text, a.rs[d], a.re[d]};
  GqAcc acc; acc.zero();
  if ((a.dbg & TB_DBG_FORCE_WALK) || !gq_doc_tiles(a, lut, stage, C, sw, R, acc, lane, F)) { acc.zero(); gq_doc_walk(C, R, sw, acc, lane); TBE_COUNT(tbe_stat_walk, lane); }
  else TBE_COUNT(tbe_stat_tiles, lane);
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

struct GrTotals { uint32_t W, NL, flags; uint64_t nchars; };

// general path: segment-walking lanes (tb_lanes.cuh)
TB_D_NOINLINE void gr_doc_walk(const KBatch& a, const Classifier& C, const TextRange& R, const GrScratch& s, int64_t d, int lane,
                      GrTotals& out) {
  const uint64_t uw = tb_wunit(R.rs, d), ul = tb_lunit(R.rs, d);
  uint64_t* wh = s.wh + uw; uint32_t* wl = s.wl + uw; uint64_t* wsg = s.ws + uw;
  uint64_t* lh = s.lh + ul; uint32_t* ll = s.ll + ul; uint8_t* lg = s.lg + ul;
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
    for (uint32_t k = 0; k < o.nw; k++) { wh[W + pre + k] = o.wh[k]; wl[W + pre + k] = o.wl[k]; wsg[W + pre + k] = o.ws[k]; }
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
  // the signatures become those of the text up to each word (as the mask path stores them): one sequential pass
  simt_sync();
  if (lane == 0) for (uint32_t e = 1; e < W; e++) wsg[e] = tb_sig_cat(wsg[e - 1], wsg[e], wl[e]);
  simt_sync();
  out.W = W; out.NL = NL; out.flags = flags; out.nchars = nchars;
}
// Words one tile can stage (five 32-bit words each).  A tile can end up to 23 * 32 words (two words per three
// bytes, "a\u00a9a\u00a9..."); a tile with more than this many -- twelve per stretch, under three bytes per word and
// separator -- hands the document to the walking path instead of tripling every warp's shared memory.
#ifndef TB_GR_STAGE_WORDS
#define TB_GR_STAGE_WORDS (12 * 32)
#endif
#define TB_GR_STAGE_U32 (5 * TB_GR_STAGE_WORDS + 2 * 32 * 32)
template <class Feed>
TB_D bool gr_doc_tiles(const KBatch& a, const LutEntry* lut, uint32_t* stage, const Classifier& C, const TextRange& R,
                       const GrScratch& s, int64_t d, int lane, GrTotals& out, Feed& feed);

// `stage`: TB_GR_STAGE_U32 words of per-warp scratch (shared memory, 8-byte aligned): the staging arrays of a tile's
// words and the prefix-hash columns of the lanes' stretches
template <class Feed = NoFeed>
TB_D void body_gr_scan(const KBatch& a, const uint16_t* s_ascii, const LutEntry* lut, uint32_t* stage, const int64_t* ts,
                       const int64_t* te, const GrScratch& s, int64_t d, int lane, Feed* feed = nullptr) {
  Feed nofeed{};
  Feed& F = feed ? *feed : nofeed;
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, ts[d], te[d]};
  GrTotals o;
  if ((a.dbg & TB_DBG_FORCE_WALK) || !gr_doc_tiles(a, lut, stage, C, R, s, d, lane, o, F)) { gr_doc_walk(a, C, R, s, d, lane, o); TBE_COUNT(tbe_stat_walk, lane); }
  else TBE_COUNT(tbe_stat_tiles, lane);
  const uint32_t flags = simt_or(o.flags);
  const uint64_t nchars = simt_sum64(o.nchars);
  if (lane == 0) {
    const uint32_t W = o.W, NL = o.NL;
    GrDoc g; g.char_len = nchars; g.W = W; g.NL = NL; g.flags = flags ? TB_DOCFLAG_COMPLEX_SCRIPT : 0; g.empty = (R.re <= R.rs);
    s.doc[d] = g;
    uint32_t big = W > NL ? W : NL;
    const uint32_t tier = tb_rep_tier(big);
    uint32_t k = simt_atomic_add(&s.counters[TB_GR_CNT_LEN(tier)], 1u);
    s.lists[(uint64_t)tier * (uint64_t)s.n + k] = (uint32_t)d;
  }
}

// ---------------------------------------------------------------------------------------------
// GopherRepetitionFilter, pass 2 (tb_rep.cuh): tiers by document size
// ---------------------------------------------------------------------------------------------
// shared-memory layout of one document's working arrays for capacity CAP (8-byte aligned pieces first)
#define TB_REP_SMEM(CAP) ((CAP) * 8 + ((CAP) + 2) * 4 + 2 * (CAP) * 4 + ((CAP) / 32 + 2) * 4 + (CAP) * 2 + (CAP) * 2)
#define TB_REP_WARP_SMEM(CAP) TB_REP_SMEM(CAP)              // a warp tier
#define TB_REP_CTA_SMEM(CAP) (TB_REP_SMEM(CAP) + (64 + 40) * 8)    // a CTA tier: + the scratch of the CTA's collectives (64 words) and the per-level results of the warps' tasks
struct RepCarve { RepArrays A; uint32_t* tab; uint16_t* cnt; uint16_t* list; uint8_t* end; };
TB_D RepCarve rep_carve(uint8_t* smem, uint32_t cap) {
  RepCarve c;
  c.A.H = (uint64_t*)smem;
  c.A.woff = (uint32_t*)(c.A.H + cap);
  c.tab = c.A.woff + cap + 2;
  c.A.flag = c.tab + 2 * cap;
  c.cnt = (uint16_t*)(c.A.flag + cap / 32 + 2);
  c.list = c.cnt + cap;
  c.end = (uint8_t*)(c.list + cap);
  return c;
}

// a warp tier: one warp, working arrays of capacity `cap` carved from `smem` (TB_REP_WARP_SMEM(cap) bytes, 8-byte aligned)
TB_D void body_gr_rep_warp(const KBatch& a, uint8_t* smem, uint32_t cap, const int64_t* ts, const int64_t* te, const GrScratch& s,
                           const GrCfg& cfg, int step, tb_gr_stats* out, int64_t d, int lane) {
  RepCarve c = rep_carve(smem, cap);
  WarpGroup g{lane};
  rep_body<WarpGroup, uint16_t>(g, a, ts, te, s, cfg, step, out, d, c.A, c.tab, c.cnt, c.list);
}
// a CTA tier: one CTA, working arrays of capacity `cap` carved from `smem` (TB_REP_CTA_SMEM(cap) bytes)
TB_D void body_gr_rep_cta(const KBatch& a, uint8_t* smem, uint32_t cap, const int64_t* ts, const int64_t* te, const GrScratch& s,
                          const GrCfg& cfg, int step, tb_gr_stats* out, int64_t d, int tid) {
  RepCarve c = rep_carve(smem, cap);
  BlockGroup g{tid, (uint64_t*)(((uintptr_t)c.end + 7) & ~(uintptr_t)7)};
  rep_body<BlockGroup, uint16_t>(g, a, ts, te, s, cfg, step, out, d, c.A, c.tab, c.cnt, c.list);
}
// tier X: one CTA, working arrays in the document's region of the global work area; `red` = 64 + 40 u64 of shared scratch
TB_D void body_gr_rep_x(const KBatch& a, uint64_t* red, const int64_t* ts, const int64_t* te, const GrScratch& s,
                        const GrCfg& cfg, int step, tb_gr_stats* out, int64_t d, int tid) {
  const int64_t rs = ts[d], re = te[d];
  const uint64_t ub = tb_wunit(rs, d);
  const uint32_t U = tb_wunits_of(re - rs);  // >= number of words, >= number of line pieces
  uint32_t* work = s.work + ub * TB_REP_K + (uint64_t)d * TB_REP_DOC_EXTRA;
  RepArrays A;
  A.H = (uint64_t*)work;                   // U entries
  A.woff = (uint32_t*)(A.H + U);           // U + 2
  uint32_t* tab = A.woff + U + 2;          // 2 U + TB_REP_MIN_TABLE
  uint32_t* cnt = tab + 2u * U + TB_REP_MIN_TABLE;
  uint32_t* list = cnt + U;
  A.flag = list + U;                       // U / 32 + 1
  BlockGroup g{tid, red};
  rep_body<BlockGroup, uint32_t>(g, a, ts, te, s, cfg, step, out, d, A, tab, cnt, list);
}

// ---------------------------------------------------------------------------------------------
// C4QualityFilter
// ---------------------------------------------------------------------------------------------
// Pass A: doc-level checks, processed lines P, line count; zeroes the per-line slots.
#define TB_STAGE_BYTES (TB_TILE + 64)  // per-warp staging buffer of the compacting stores (tile_store_kept)
template <class Feed>
TB_D void c4a_doc_tiles(const KBatch& a, const LutEntry* lut, uint8_t* stage, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s,
                        const TextRange& R, int64_t d, int lane, Feed& feed);
TB_D_NOINLINE void c4a_doc_walk(const KBatch& a, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s, const TextRange& R, int64_t d, int lane);
template <class Feed = NoFeed>
TB_D void body_c4_a(const KBatch& a, const uint16_t* s_ascii, const LutEntry* lut, uint8_t* stage, const C4Cfg& cfg, const C4Scratch& s,
                    int64_t d, int lane, Feed* feed = nullptr) {
  Feed nofeed{};
  Feed& F = feed ? *feed : nofeed;
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, a.rs[d], a.re[d]};
  if (a.dbg & TB_DBG_FORCE_WALK) { c4a_doc_walk(a, C, cfg, s, R, d, lane); TBE_COUNT(tbe_stat_walk, lane); }
  else { c4a_doc_tiles(a, lut, stage, C, cfg, s, R, d, lane, F); TBE_COUNT(tbe_stat_tiles, lane); }
}
TB_D_NOINLINE void c4a_doc_walk(const KBatch& a, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s, const TextRange& R, int64_t d, int lane) {
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
    else { o.nkept = 0; o.n_nl = 0; o.flags = 0; o.kmask = 0; }
    flags |= o.flags;
    if (o.kmask) {  // the kept bytes into the bitmap of pass A (the stretches of this path are not aligned: always OR)
      const uint64_t k0 = (uint64_t)p0 + (uint32_t)((uintptr_t)a.text & 31u);
      const uint32_t sh = (uint32_t)(k0 & 31u);
      const uint64_t lo = o.kmask << sh;                       // bits of the words k0 >> 5 and + 1 ...
      const uint64_t hi = sh ? (o.kmask >> (64u - sh)) : 0ull;  // ... and + 2
      if ((uint32_t)lo) simt_atomic_or(&s.ka[k0 >> 5], (uint32_t)lo);
      if ((uint32_t)(lo >> 32)) simt_atomic_or(&s.ka[(k0 >> 5) + 1], (uint32_t)(lo >> 32));
      if ((uint32_t)hi) simt_atomic_or(&s.ka[(k0 >> 5) + 2], (uint32_t)hi);
    }
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
struct C4bSink {  // words arrive in text order: aggregate per line in registers, one pair of atomics per line and lane
  uint32_t* lwc; uint32_t* lmax; uint64_t lbase; int64_t p0, p1; uint32_t nlmask;
  uint64_t cur; uint32_t cnt, mx;
  TB_D void flush() {
    if (cnt) { simt_atomic_add(&lwc[cur], cnt); simt_atomic_max(&lmax[cur], mx); }
    cnt = 0; mx = 0;
  }
  TB_D void on_word(WordInfo& w) {
    int64_t x = w.start < p1 ? w.start : p1;
    uint32_t below = (x - p0) >= 32 ? nlmask : (nlmask & ((1u << (uint32_t)(x - p0)) - 1u));
    uint64_t li = lbase + simt_popc(below);
    if (li != cur) { flush(); cur = li; }
    cnt++; mx = w.nchars > mx ? w.nchars : mx;
  }
};
TB_D_NOINLINE void c4b_doc_walk(const Classifier& C, const C4Cfg& cfg, const C4Scratch& s, const TextRange& R, uint64_t lb, int lane) {
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
      C4bSink sink{s.lwc, s.lmax, lbase, p0, p1, nlmask, lbase, 0u, 0u};
      tb_walk_words_fast<0>(C, R, p0, p1, sink);
      sink.flush();
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
template <class Feed>
TB_D bool c4b_doc_tiles(const KBatch& a, const LutEntry* lut, const Classifier& C, const C4Cfg& cfg, const C4Scratch& s,
                        const TextRange& R, uint64_t lb, int lane, Feed& feed);
template <class Feed = NoFeed>
TB_D void body_c4_b(const KBatch& a, const uint16_t* s_ascii, const LutEntry* lut, const C4Cfg& cfg, const C4Scratch& s, int64_t d,
                    int lane, Feed* feed = nullptr) {
  Feed nofeed{};
  Feed& F = feed ? *feed : nofeed;
  Classifier C{s_ascii, a.uc};
  TextRange R{s.P, s.prs[d], s.pre[d]};
  const uint64_t lb = (uint64_t)a.rs[d] + (uint64_t)d;
  if (!(a.dbg & TB_DBG_FORCE_WALK) && c4b_doc_tiles(a, lut, C, cfg, s, R, lb, lane, F)) { TBE_COUNT(tbe_stat_tiles, lane); return; }
  TBE_COUNT(tbe_stat_walk, lane);
  // general path: clear what the mask path may have added, then walk
  simt_sync();
  for (uint32_t k = (uint32_t)lane; k <= s.nlines[d]; k += 32) { s.lwc[lb + k] = 0; s.lmax[lb + k] = 0; s.lflags[lb + k] = 0; }
  simt_sync();
  c4b_doc_walk(C, cfg, s, R, lb, lane);
}

// Pass C: drop lines, join the kept ones with '\n' (c4_filters.rs:207-258) -> F; line counters.
// It walks the text pass A read, not P: the processed lines are the bytes pass A marked in `ka`, and the '\n' bytes
// (all of them marked but the document's last) sit where they sat.  So every byte of F is a byte of that text, and
// the bitmap `kf` of the bytes that went into F is all a caller needs to rebuild the content from its own copy
// of the text (TB_OPT_RETURN_CONTENT = 2, k_c4_runs).
TB_D uint32_t tile_store_kept(const uint32_t* w, uint32_t K, uint8_t* stage, uint8_t* dst, int lane);
TB_D void body_c4_c(const KBatch& a, uint8_t* stage, const C4Cfg& cfg, const C4Scratch& s, int64_t d, int lane) {
  if (!a.alive[d] || s.docflags[d]) { if (lane == 0) { s.frs[d] = a.rs[d]; s.fre[d] = a.rs[d]; } return; }
  TextRange R{a.text, a.rs[d], a.re[d]};
  const uint32_t mis = (uint32_t)((uintptr_t)a.text & 31u);
  uint8_t* out = s.F + a.rs[d];
  const uint64_t lb = (uint64_t)a.rs[d] + (uint64_t)d;
  const uint32_t m = s.nlines[d];
  uint32_t c1 = 0, c2 = 0, c3 = 0, kept = 0;  // this lane's share of the counters
  uint32_t nl = 0, kept_run = 0;              // lines terminated / kept before this tile
  int64_t outpos = 0;
  if (m) {
    for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
      uint32_t w[8];
      const int64_t p0 = base + lane * TB_STRETCH;
      const uint32_t V = tile_load(R, a.ext_lo, a.ext_hi, p0, w);
      uint32_t NLm = tile_eq_mask(w, '\n') & V;
      {
        const int64_t last = R.re - 1;  // the '\n' that ends the document ends no line of P (pass A dropped it)
        if (last >= p0 && last < p0 + 32) NLm &= ~(1u << (uint32_t)(last - p0));
      }
      const uint32_t KA = V ? (s.ka[((uint64_t)p0 + mis) >> 5] & V) : 0u;
      const uint32_t cnt = (uint32_t)tb_popc32(NLm);
      uint32_t tot, pre = simt_exscan(cnt, lane, &tot);
      const uint64_t l0 = lb + nl + pre;  // line that holds the first byte of the stretch
      // decisions of the lines this stretch touches: bit j = line l0 + j is kept
      uint64_t dec = 0;  // up to 33 lines
      if (V) {
        for (uint32_t j = 0; j <= cnt; j++) {
          const int r = c4_line_decision(cfg, s.lwc[l0 + j], s.lmax[l0 + j], s.lflags[l0 + j]);
          if (r == 0) dec |= 1ull << j;
          if (j < cnt) { if (r == 0) kept++; else if (r == 1) c1++; else if (r == 2) c2++; else if (r == 3) c3++; }  // lines that end here
        }
      }
      uint32_t ktot, kpre = simt_exscan((uint32_t)tb_popc64(dec & ((1ull << cnt) - 1ull)), lane, &ktot);
      // bytes of kept lines; a '\n' goes in front of a kept line that follows another kept line
      uint32_t K = 0, mm = NLm;
      int start = 0;
      for (uint32_t j = 0; j < cnt; j++) {
        const int bit = tb_ctz32(mm);
        mm &= mm - 1u;
        if ((dec >> j) & 1u) K |= tb_below(bit) & ~tb_below(start);
        if (((dec >> (j + 1)) & 1u) && kept_run + kpre + (uint32_t)tb_popc64(dec & ((2ull << j) - 1ull)) > 0u) K |= 1u << bit;
        start = bit + 1;
      }
      if ((dec >> cnt) & 1u) K |= ~tb_below(start);
      K &= KA;  // (the '\n' bytes between lines are marked in ka)
      if (s.kf) tb_bits_store(s.kf, p0, mis, V, K);
      outpos += tile_store_kept(w, K, stage, out + outpos, lane);
      nl += tot; kept_run += ktot;
    }
    if (lane == 0) {  // the last line has no terminator
      int r = c4_line_decision(cfg, s.lwc[lb + m - 1], s.lmax[lb + m - 1], s.lflags[lb + m - 1]);
      if (r == 0) kept++; else if (r == 1) c1++; else if (r == 2) c2++; else if (r == 3) c3++;
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
  if (lane == 0 && R.rs < R.re) ns = tb_count_sentences_start(C, R);
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    // events of the stretch: ASCII terminators / separators and lead bytes of non-ASCII characters (SWAR masks)
    uint32_t w[8];
    const int64_t p0 = base + lane * TB_STRETCH;
    const uint32_t V = tile_load(R, a.ext_lo, a.ext_hi, p0, w);
    if (!V) continue;
    uint32_t hi = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) hi |= (((((w[k] & (w[k] << 1)) & 0x80808080u) >> 7) * 0x00204081u >> 21) & 15u) << (4 * k);
    const uint32_t ev = (tile_eq_mask(w, '.') | tile_eq_mask(w, '!') | tile_eq_mask(w, '?') | tile_eq_mask(w, '\n') |
                         tile_eq_mask(w, '\r') | hi) & V;
    ns += tb_count_sentences_events(C, R, p0, ev);
  }
  ns = simt_sum(ns);
  if (lane) return;
  st.n_lines = s.nlines[d]; st.too_long_word = s.cnt[4 * d]; st.no_terminal_punc = s.cnt[4 * d + 1];
  st.too_few_words = s.cnt[4 * d + 2]; st.n_kept_lines = s.cnt[4 * d + 3]; st.n_sentences = ns; st.rewritten = 1;
  if (cfg.min_num_sentences > 0 && (uint64_t)ns < cfg.min_num_sentences) st.reason_bits |= TB_C4_REASON_TOO_FEW_SENTENCES;
  out[d] = st;
  rewritten[d] = s.ftag;
  rs_out[d] = R.rs; re_out[d] = R.re;
  if (st.reason_bits) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
}

// Sentence-wise lines (split_paragraph = false: `lines = split_into_sentences(content)`, c4_filters.rs:152-155).
// The document is rewritten into S so that str::lines() of the result yields exactly the reference's pieces
// (up to outer white space, which pass A trims per line like the reference does, :193), and the line passes
// above then run on S unchanged:
//   * only the trimmed text is copied (text.rs:60);
//   * every '\n' is dropped -- a piece always ends right after one (SB4), so none is interior to a piece;
//   * a '\n' goes in front of the first non-white-space character of every piece but the first; pieces that
//     are white space only (dropped by text.rs:80-86) thereby melt into the end of the line before them.
// S can be longer than the text: a break that follows no white space ("a?b") adds a byte.  Such a break
// closes a piece of at least two bytes, so doc d fits into [soff(d), soff(d + 1)), soff(d) = rs + rs / 2 + 2 d.
#define TB_SENT_STAGE_BYTES (TB_TILE + TB_TILE / 2 + 32)
TB_HD int64_t tb_sent_off(int64_t rs, int64_t d) { return rs + (rs >> 1) + 2 * d; }
struct SentMark {
  uint32_t* bits; uint32_t mis;
  TB_D void operator()(int64_t q) { const uint64_t k = (uint64_t)q + mis; simt_atomic_or(&bits[k >> 5], 1u << (uint32_t)(k & 31u)); }
};
TB_D void body_c4_sent(const KBatch& a, const uint16_t* s_ascii, uint8_t* stage, const int64_t* ts, const int64_t* te,
                       uint32_t* bits, uint8_t* S, int64_t* srs, int64_t* sre, int64_t d, int lane) {
  Classifier C{s_ascii, a.uc};
  TextRange R{a.text, ts[d], te[d]};
  const int64_t o0 = tb_sent_off(a.rs[d], d);
  if (R.rs >= R.re) { if (lane == 0) { srs[d] = o0; sre[d] = o0; } return; }
  SentMark mark{bits, (uint32_t)((uintptr_t)a.text & 31u)};
  // pass 1: where the pieces begin (bit per byte; a mark can land in a later stretch or tile)
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    uint32_t w[8];
    const int64_t p0 = base + lane * TB_STRETCH;
    const uint32_t V = tile_load(R, a.ext_lo, a.ext_hi, p0, w);
    if (!V) continue;
    uint32_t hi = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) hi |= (((((w[k] & (w[k] << 1)) & 0x80808080u) >> 7) * 0x00204081u >> 21) & 15u) << (4 * k);
    const uint32_t ev = (tile_eq_mask(w, '.') | tile_eq_mask(w, '!') | tile_eq_mask(w, '?') | tile_eq_mask(w, '\n') |
                         tile_eq_mask(w, '\r') | hi) & V;
    tb_sentence_starts_events(C, R, p0, ev, mark);
  }
  simt_sync();
  // pass 2: copy without the '\n' bytes, a '\n' in front of every marked byte
  uint8_t* out = S + o0;
  int64_t outpos = 0;
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    uint32_t w[8];
    const int64_t p0 = base + lane * TB_STRETCH;
    const uint32_t V = tile_load(R, a.ext_lo, a.ext_hi, p0, w);
    uint32_t INS = 0;
    if (V) INS = simt_load_fresh(&bits[((uint64_t)p0 + mark.mis) >> 5]) & V;
    const uint32_t KEEP = V & ~tile_eq_mask(w, '\n');
    uint32_t tot, pre = simt_exscan((uint32_t)(tb_popc32(KEEP) + tb_popc32(INS)), lane, &tot);
    uint32_t o = pre;
#pragma unroll
    for (int k = 0; k < 32; k++) {
      if ((INS >> k) & 1u) stage[o++] = '\n';
      if ((KEEP >> k) & 1u) stage[o++] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
    }
    simt_sync();
    for (uint32_t k = (uint32_t)lane; k < tot; k += 32) out[outpos + k] = stage[k];
    simt_sync();
    outpos += tot;
  }
  if (lane == 0) { srs[d] = o0; sre[d] = o0 + outpos; }
}

// ---------------------------------------------------------------------------------------------
// HTML entities (tb_entities.cuh): the decoded copy of a document, the text every step then sees
// ---------------------------------------------------------------------------------------------
// Pass 1: every '&' is a candidate, parsed by the lane that owns it; for an entity the '&' is marked in `amp`
// and the rest of its span in `drop` (bit per byte of the text buffer; a span can reach into later stretches).
// Pass 2: copy, with the replacement in place of a marked '&' and without the dropped bytes.  Only "&nLt;" and
// "&nGt;" have a replacement longer than the entity (by one byte); tb_sent_off leaves room for far more.
TB_D_NOINLINE uint32_t ent_parse_at(const TextRange& R, int64_t a, const EntEntry* tab, uint32_t ntab, uint32_t* span, uint64_t* repl) {
  return ent_parse(R.t, a, R.re, tab, ntab, span, repl);
}
TB_D void body_entities(const KBatch& a, uint8_t* stage, const EntEntry* tab, uint32_t ntab, uint32_t* amp, uint32_t* drop,
                        uint8_t* D, int64_t* drs, int64_t* dre, int64_t d, int lane) {
  TextRange R{a.text, a.rs[d], a.re[d]};
  int64_t o0 = tb_sent_off(a.rs[d], d);
  // The copy starts at the same address modulo 32 as the text (a document of 128 bytes or more has the room):
  // as long as no entity has been met, a stretch is then copied with two aligned 16-byte stores.
  const bool same_phase = R.re - R.rs >= 128;
  if (same_phase) o0 += (int64_t)(((uintptr_t)(R.t + R.rs) - (uintptr_t)(D + o0)) & 31u);
  const uint32_t mis = (uint32_t)((uintptr_t)a.text & 31u);
  uint8_t* out = D + o0;
  int64_t outpos = 0;
  for (int64_t base = tb_tile_origin(R); base < R.re; base += TB_TILE) {
    uint32_t w[8];
    const int64_t p0 = base + lane * TB_STRETCH;
    const uint32_t V = tile_load(R, a.ext_lo, a.ext_hi, p0, w);
    // marks: a span only reaches forward (into this tile or the next), so the marks of a tile are complete once
    // the candidates of the tiles up to it have been parsed
    uint32_t cand = tile_eq_mask(w, '&') & V;
    if (simt_any(cand != 0u)) {
      while (cand) {
        const int i = tb_ctz32(cand);
        cand &= cand - 1u;
        uint32_t span; uint64_t repl;
        if (!ent_parse_at(R, p0 + i, tab, ntab, &span, &repl)) continue;
        const uint64_t k0 = (uint64_t)(p0 + i) + mis;
        simt_atomic_or(&amp[k0 >> 5], 1u << (uint32_t)(k0 & 31u));
        for (uint64_t k = k0 + 1; k < k0 + span; k++) simt_atomic_or(&drop[k >> 5], 1u << (uint32_t)(k & 31u));
      }
      simt_sync();
    }
    uint32_t AMP = 0, DROP = 0;
    if (V) { const uint64_t wi = ((uint64_t)p0 + mis) >> 5; AMP = simt_load_fresh(&amp[wi]) & V; DROP = simt_load_fresh(&drop[wi]) & V; }
    if (same_phase && outpos == (base < R.rs ? 0 : base - R.rs) && !simt_any((AMP | DROP) != 0u)) {
      // nothing decoded up to the end of this tile: the copy is in phase with the text
      uint8_t* q = out + (p0 - R.rs);
      if (V == 0xFFFFFFFFu) {
#if defined(__CUDA_ARCH__)
        ((uint4*)q)[0] = make_uint4(w[0], w[1], w[2], w[3]); ((uint4*)q)[1] = make_uint4(w[4], w[5], w[6], w[7]);
#else
        memcpy(q, w, 32);
#endif
      } else {
        for (uint32_t m = V; m; m &= m - 1u) { const int k = tb_ctz32(m); q[k] = (uint8_t)tb_byte(R.t, p0 + k); }
      }
      outpos += (int64_t)simt_sum((uint32_t)tb_popc32(V));
      continue;
    }
    const uint32_t KEEP = V & ~DROP & ~AMP;
    uint32_t n = (uint32_t)tb_popc32(KEEP);
    for (uint32_t m = AMP; m; m &= m - 1u) { uint32_t span; uint64_t repl; n += ent_parse_at(R, p0 + tb_ctz32(m), tab, ntab, &span, &repl); }
    uint32_t tot, pre = simt_exscan(n, lane, &tot);
    uint32_t o = pre;
    if (AMP == 0u) {
#pragma unroll
      for (int k = 0; k < 32; k++) if ((KEEP >> k) & 1u) stage[o++] = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
    } else {  // rare: a stretch with an entity in it
      for (int k = 0; k < 32; k++) {
        if ((AMP >> k) & 1u) {
          uint32_t span; uint64_t repl;
          const uint32_t rl = ent_parse_at(R, p0 + k, tab, ntab, &span, &repl);
          for (uint32_t q = 0; q < rl; q++) stage[o++] = (uint8_t)(repl >> (8 * q));
        } else if ((KEEP >> k) & 1u) stage[o++] = (uint8_t)tb_byte(R.t, p0 + k);
      }
    }
    simt_sync();
    for (uint32_t k = (uint32_t)lane; k < tot; k += 32) out[outpos + k] = stage[k];
    simt_sync();
    outpos += tot;
  }
  if (lane == 0) { drs[d] = o0; dre[d] = o0 + outpos; }
}

// ---------------------------------------------------------------------------------------------
// final content as runs of the caller's own text (TB_OPT_RETURN_CONTENT = 2)
// ---------------------------------------------------------------------------------------------
// C4's output is a subsequence of its input (c4_filters.rs:192-258: lines trimmed, citations cut, kept lines joined by
// the '\n' that stood in front of them).  Instead of the rewritten bytes the host gets, per document, the maximal runs
// [begin, end) of input bytes (relative to the document's start) whose concatenation is the content.  A document C4
// did not rewrite is one run, its whole text.  For a rewritten one the runs are read off the bitmap `kf` of pass C,
// cut to the bytes that survive the final trim (c4_filters.rs:258): F holds `total` bytes, the trimmed content is
// its bytes [lo, hi).  One warp per document, a bitmap word per lane; emit = false only counts (nruns[d]).
struct tb_run_dev { uint32_t begin, end; };
TB_D void body_c4_runs(const int64_t* rs0, const int64_t* re0, const uint8_t* rewritten, const uint32_t* kf, uint32_t mis,
                       const int64_t* frs, const int64_t* fre, int64_t* nruns, const int64_t* run_off, tb_run_dev* runs,
                       int64_t cap, bool emit, int64_t d, int lane) {
  const int64_t rs = rs0[d], re = re0[d];
  if (!rewritten[d]) {
    if (lane == 0) {
      if (!emit) nruns[d] = re > rs ? 1 : 0;
      else if (re > rs && run_off[d] < cap) runs[run_off[d]] = tb_run_dev{0u, (uint32_t)(re - rs)};
    }
    return;
  }
  const int64_t lo = frs[d] - rs, hi = lo + (fre[d] - frs[d]);  // ranks (in F) of the bytes that stay
  const uint64_t w0 = ((uint64_t)rs + mis) >> 5, w1 = ((uint64_t)re + mis) >> 5;  // one word past the text closes a run that reaches its end
  int64_t rank = 0, nb = 0, ne = 0;  // kept bytes / run begins / run ends before this round of 32 words
  uint32_t carry = 0;                // the bit in front of this round's first word
  const int64_t base = run_off ? run_off[d] : 0;
  for (uint64_t wb = w0; wb <= w1; wb += 32) {
    const uint64_t wi = wb + (uint64_t)lane;
    uint32_t w = 0;
    if (wi <= w1 && rs < re) {
      w = kf[wi];
      const int64_t p0 = (int64_t)(wi << 5) - mis;  // position of bit 0
      if (p0 < rs) w &= ~tb_below((int)(rs - p0 > 32 ? 32 : rs - p0));
      if (p0 + 32 > re) w &= tb_below((int)(re - p0 < 0 ? 0 : re - p0));
    }
    // keep the bits whose rank lies in [lo, hi)
    const uint32_t c = (uint32_t)tb_popc32(w);
    uint32_t tot, pre = simt_exscan(c, lane, &tot);
    const int64_t r0 = rank + pre;
    if (c) {
      if (r0 + c <= lo || r0 >= hi) w = 0;
      else {
        for (int64_t k = r0; k < lo; k++) w &= w - 1u;                     // drop the lowest bits below lo
        int64_t over = r0 + c - hi;                                         // ... and the highest beyond hi
        while (over-- > 0) w &= ~(0x80000000u >> tb_clz32(w));
      }
    }
    rank += tot;
    uint32_t prev_top = simt_shfl_up(w >> 31, 1);
    if (lane == 0) prev_top = carry;
    carry = simt_shfl(w >> 31, 31);
    const uint32_t before = (w << 1) | prev_top;
    uint32_t begins = w & ~before, ends = ~w & before;  // a run begins at a set bit after a clear one, and ends in front of a clear bit after a set one
    uint32_t tb_, te_, pb = simt_exscan((uint32_t)tb_popc32(begins), lane, &tb_), pe = simt_exscan((uint32_t)tb_popc32(ends), lane, &te_);
    if (emit) {
      const int64_t p0 = (int64_t)(wi << 5) - mis - rs;
      int64_t k = base + nb + pb;
      for (; begins; begins &= begins - 1u, k++) if (k < cap) runs[k].begin = (uint32_t)(p0 + tb_ctz32(begins));
      k = base + ne + pe;
      for (; ends; ends &= ends - 1u, k++) if (k < cap) runs[k].end = (uint32_t)(p0 + tb_ctz32(ends));
    }
    nb += tb_; ne += te_;
  }
  if (!emit && lane == 0) nruns[d] = nb;
}

// ---------------------------------------------------------------------------------------------
// final content
// ---------------------------------------------------------------------------------------------
TB_D void body_final_len(const int64_t* rs0, const int64_t* re0, const int64_t* rs, const int64_t* re,
                         const uint8_t* rewritten, int64_t* len, int64_t d) {
  len[d] = rewritten[d] ? re[d] - rs[d] : re0[d] - rs0[d];
}
TB_D void body_pack(const uint8_t* orig, const int64_t* rs0, const uint8_t* cur, const uint8_t* cur2, const int64_t* rs, const uint8_t* rewritten,
                    const int64_t* out_off, uint8_t* out, int64_t d, int lane) {
  const uint8_t* src = rewritten[d] == 2 ? cur2 + rs[d] : rewritten[d] ? cur + rs[d] : orig + rs0[d];  // (the buffer the document was last rewritten into)
  const int64_t o = out_off[d], len = out_off[d + 1] - o;
  uint8_t* dst = out + o;
#if defined(__CUDA_ARCH__)
  // 16-byte stores on the destination's alignment; the source words are realigned with funnel shifts
  int64_t head = (int64_t)((16u - (uint32_t)((uintptr_t)dst & 15u)) & 15u);
  if (head > len) head = len;
  if (lane < head) dst[lane] = src[lane];
  const int64_t body = (len - head) & ~(int64_t)15;
  const uint8_t* s2 = src + head;
  uint8_t* d2 = dst + head;
  const uint32_t mis = (uint32_t)((uintptr_t)s2 & 3u);
  const uint32_t* sw = (const uint32_t*)(s2 - mis);
  for (int64_t c = (int64_t)lane * 16; c < body; c += 512) {
    const uint32_t* q = sw + (c >> 2);
    uint32_t w0 = q[0], w1 = q[1], w2 = q[2], w3 = q[3];
    uint4 v;
    if (mis) {
      const uint32_t w4 = q[4], sh = 8u * mis;  // the fifth word still overlaps byte c + 15 of the source
      v.x = __funnelshift_r(w0, w1, sh); v.y = __funnelshift_r(w1, w2, sh); v.z = __funnelshift_r(w2, w3, sh); v.w = __funnelshift_r(w3, w4, sh);
    } else { v.x = w0; v.y = w1; v.z = w2; v.w = w3; }
    *(uint4*)(d2 + c) = v;
  }
  for (int64_t i = head + body + lane; i < len; i += 32) dst[i] = src[i];
#else
  for (int64_t i = lane; i < len; i += 32) dst[i] = src[i];
#endif
}

#include "tb_bodies_tile.cuh"
#include "tb_bodies_fw.cuh"
