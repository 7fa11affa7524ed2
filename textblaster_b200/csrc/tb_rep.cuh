// tb_rep.cuh -- GopherRepetitionFilter pass 2: duplicate lines / paragraphs, top n-grams and
// duplicate n-grams of one document, written once against a "group" of cooperating threads:
//   tier S  one warp  per document, working arrays in shared memory        (W, NL <= 512)
//   tier M  one CTA   per document, working arrays in shared memory        (W, NL <= 2048)
//   tier L  one CTA   per document (one per SM), arrays in shared memory   (W, NL <= 8192)
//   tier X  one CTA   per document, working arrays in the global work area (anything larger; L2 resident)
// Replaces text.rs:197-259 as called from gopher_rep.rs:86-196.
//
// Open-addressing tables hold element indices; the keys live in the arrays (equality = 62-bit
// polynomial hash + byte length).  find_all_duplicate's greedy scan is order dependent; it is
// reduced to a short sequential pass over the *candidate* positions only (those whose key already
// occurred earlier), see rep_dup_walk.
#pragma once
#include "tb_kernels.cuh"
#include "tb_simt.cuh"

#define TB_REP_S_CAP 512u
#define TB_REP_M_CAP 2048u
#define TB_REP_L_CAP 8192u
#define TB_REP_M_THREADS 256
#define TB_REP_L_THREADS 1024
#define TB_REP_X_THREADS 512

// ---- thread groups ---------------------------------------------------------------------------
struct WarpGroup {
  int tid;
  static constexpr int kSize = 32;
  TB_D int size() const { return 32; }
  TB_D void sync() const { simt_sync(); }
  TB_D uint32_t sum(uint32_t v) const { return simt_sum(v); }
  TB_D uint64_t sum64(uint64_t v) const { return simt_sum64(v); }
  TB_D uint64_t max64(uint64_t v) const { return simt_max64(v); }
  TB_D bool any(bool p) const { simt_sync(); return simt_any(p); }  // also a barrier with memory ordering, like the CTA's
  TB_D uint32_t exscan(uint32_t v, uint32_t* total) const { return simt_exscan(v, tid, total); }
  TB_D uint32_t bcast(uint32_t v) const { return simt_shfl(v, 0); }
};

#if !defined(TB_HOST_EMUL)
// One CTA; `red` is 64 u64 of shared scratch, used as two alternating halves so that a collective needs a
// single barrier: a half is rewritten only after every thread has passed the barrier of the collective
// that used the other half.
struct BlockGroup {
  int tid;
  uint64_t* red;
  mutable int phase = 0;
  TB_D int size() const { return (int)blockDim.x; }
  TB_D void sync() const { __syncthreads(); }
  TB_D uint64_t* publish(uint64_t v) const {  // lane 0 of every warp posts v; returns the half holding the posts
    uint64_t* r = red + 32 * phase;
    phase ^= 1;
    if ((tid & 31) == 0) r[tid >> 5] = v;
    __syncthreads();
    return r;
  }
  TB_D uint64_t sum64(uint64_t v) const {
    const uint64_t* r = publish(simt_sum64(v));
    const int lane = tid & 31, nw = (int)blockDim.x >> 5;
    return simt_sum64(lane < nw ? r[lane] : 0ull);
  }
  TB_D uint32_t sum(uint32_t v) const { return (uint32_t)sum64(v); }
  TB_D uint64_t max64(uint64_t v) const {
    const uint64_t* r = publish(simt_max64(v));
    const int lane = tid & 31, nw = (int)blockDim.x >> 5;
    return simt_max64(lane < nw ? r[lane] : 0ull);
  }
  TB_D bool any(bool p) const { return __syncthreads_or(p ? 1 : 0) != 0; }
  TB_D uint32_t exscan(uint32_t v, uint32_t* total) const {
    uint32_t wt, pre = simt_exscan(v, tid & 31, &wt);
    const uint64_t* r = publish(wt);
    const int lane = tid & 31, w = tid >> 5, nw = (int)blockDim.x >> 5;
    uint32_t tot, wpre = simt_exscan(lane < nw ? (uint32_t)r[lane] : 0u, lane, &tot);
    *total = tot;
    return simt_shfl(wpre, w) + pre;
  }
  TB_D uint32_t bcast(uint32_t v) const {
    uint64_t* r = red + 32 * phase;
    phase ^= 1;
    if (tid == 0) r[0] = v;
    __syncthreads();
    return (uint32_t)r[0];
  }
};
#else
// host emulation: N threads share one barrier + mailbox (tb_simt.cuh SimtBlock)
struct BlockGroup {
  int tid;
  uint64_t* red;  // unused
  mutable int phase = 0;
  TB_D int size() const { return simt_block->n; }
  TB_D void sync() const { simt_block->bar.arrive_and_wait(); }
  TB_D uint64_t sum64(uint64_t v) const {
    SimtBlock* b = simt_block; b->box[tid] = v; b->bar.arrive_and_wait();
    uint64_t s = 0; for (int i = 0; i < b->n; i++) s += b->box[i];
    b->bar.arrive_and_wait(); return s;
  }
  TB_D uint32_t sum(uint32_t v) const { return (uint32_t)sum64(v); }
  TB_D uint64_t max64(uint64_t v) const {
    SimtBlock* b = simt_block; b->box[tid] = v; b->bar.arrive_and_wait();
    uint64_t s = 0; for (int i = 0; i < b->n; i++) s = b->box[i] > s ? b->box[i] : s;
    b->bar.arrive_and_wait(); return s;
  }
  TB_D bool any(bool p) const { return sum64(p ? 1 : 0) != 0; }
  TB_D uint32_t exscan(uint32_t v, uint32_t* total) const {
    SimtBlock* b = simt_block; b->box[tid] = v; b->bar.arrive_and_wait();
    uint32_t base = 0, tot = 0;
    for (int i = 0; i < b->n; i++) { if (i < tid) base += (uint32_t)b->box[i]; tot += (uint32_t)b->box[i]; }
    b->bar.arrive_and_wait(); *total = tot; return base;
  }
  TB_D uint32_t bcast(uint32_t v) const {
    SimtBlock* b = simt_block; if (tid == 0) b->box[0] = v; b->bar.arrive_and_wait();
    uint32_t r = (uint32_t)b->box[0]; b->bar.arrive_and_wait(); return r;
  }
};
#endif

// ---- tables ----------------------------------------------------------------------------------
template <class IdxT> struct IdxTraits;
template <> struct IdxTraits<uint16_t> { static constexpr uint32_t kEmpty = 0xFFFFu; };
template <> struct IdxTraits<uint32_t> { static constexpr uint32_t kEmpty = 0xFFFFFFFFu; };

// table slots are 32-bit (native atomics); the per-position counters may be 16-bit in shared memory
TB_D uint32_t idx_cas(uint32_t* a, uint32_t cmp, uint32_t val) { return simt_cas(a, cmp, val); }
TB_D void idx_min(uint32_t* a, uint32_t val) { simt_atomic_min(a, val); }
TB_D void idx_inc(uint32_t* a) { simt_atomic_add(a, 1u); }
TB_D void idx_inc(uint16_t* a) {
  uint32_t* w = (uint32_t*)((uintptr_t)a & ~(uintptr_t)3);
  simt_atomic_add(w, ((uintptr_t)a & 2) ? 0x10000u : 1u);  // counts stay < 65536, no carry across halves
}

// table size: load factor 0.5; any size works because slots are picked by multiply-high
TB_D uint32_t rep_tab_size(uint32_t count) { return 2u * count < TB_REP_MIN_TABLE ? TB_REP_MIN_TABLE : 2u * count; }
TB_D uint32_t rep_slot(uint64_t key, uint32_t T) {  // the two residues are already well mixed
  uint32_t h = ((uint32_t)(key >> 32) * 0x9E3779B1u) ^ (uint32_t)key;
  h ^= h >> 15;
  return (uint32_t)(((uint64_t)h * T) >> 32);
}
TB_D uint32_t rep_next(uint32_t sl, uint32_t T) { return sl + 1 == T ? 0u : sl + 1; }
template <class G>
TB_D void rep_fill(const G& g, uint32_t* a, uint32_t n, uint32_t v) {
  for (uint32_t i = g.tid; i < n; i += g.size()) a[i] = v;
  g.sync();
}

// find_duplicates (text.rs:197-208) over n pieces: dup_elems = n - distinct, dup_bytes = total - distinct bytes
template <class G>
TB_D void rep_dedup(const G& g, uint32_t* tab, const uint64_t* h, const uint32_t* len, uint32_t n, uint32_t* dup_elems,
                    uint64_t* dup_bytes) {
  const uint32_t EMPTY = 0xFFFFFFFFu;
  uint32_t T = rep_tab_size(n);
  rep_fill(g, tab, T, EMPTY);
  uint32_t distinct = 0; uint64_t dbytes = 0, total = 0;
  for (uint32_t i = g.tid; i < n; i += g.size()) {
    uint64_t hi = h[i]; uint32_t li = len[i];
    uint32_t sl = rep_slot(hi, T);
    total += li;
    for (;;) {
      uint32_t old = idx_cas(&tab[sl], EMPTY, i);
      if (old == EMPTY) { distinct++; dbytes += li; break; }
      if (h[old] == hi) break;  // both residues agree: equal strings (the byte offset +1 makes lengths agree too)
      sl = rep_next(sl, T);
    }
  }
  g.sync();
  distinct = g.sum(distinct); dbytes = g.sum64(dbytes); total = g.sum64(total);
  *dup_elems = n - distinct; *dup_bytes = total - dbytes;
}

// The greedy scan of find_all_duplicate (text.rs:241-259) given fo[i] = first position with the same
// key.  Positions whose key is new (fo[i] == i) only ever insert; they are pre-marked "seen".  A hit at
// q skips q+1..q+n-1: first occurrences inside that window were never visited, so their mark is
// removed; a later occurrence of such a key then inserts instead of hitting.  Only candidates
// (fo[q] < q), in increasing order, need the sequential pass.
template <class G, class IdxT>
TB_D uint64_t rep_dup_walk(const G& g, IdxT* fo, uint32_t* cand, const uint32_t* woff, uint32_t cn, uint32_t n) {
  const uint32_t EMPTY = IdxTraits<IdxT>::kEmpty;
  // ordered compaction of the candidates
  uint32_t ncand = 0;
  for (uint32_t b0 = 0; b0 < cn; b0 += g.size()) {
    uint32_t i = b0 + g.tid;
    uint32_t is = (i < cn && fo[i] != i) ? 1u : 0u, tot, pre = g.exscan(is, &tot);
    if (is) cand[ncand + pre] = i;
    ncand += tot;
  }
  g.sync();
  uint64_t rep = 0;
  if (g.tid == 0) {
    uint32_t skip_until = 0;
    for (uint32_t k = 0; k < ncand; k++) {
      uint32_t q = cand[k];
      if (q < skip_until) continue;
      uint32_t c = fo[q];
      if (fo[c] == c) {  // key seen: hit
        rep += woff[q + n] - woff[q];
        uint32_t e = q + n < cn ? q + n : cn;
        for (uint32_t p = q + 1; p < e; p++) if (fo[p] == p) fo[p] = (IdxT)EMPTY;
        skip_until = q + n;
      } else {
        fo[c] = (IdxT)c;  // its first occurrence was skipped: this one inserts
      }
    }
  }
  g.sync();
  return rep;  // valid on tid 0
}

struct RepArrays {
  // per-document working arrays (shared memory for tiers S/M, global work area for tier L)
  uint64_t* H;     // [cap] current n-gram key hash / paragraph hashes
  uint32_t* woff;  // [cap + 1] byte offsets of the words in their concatenation / paragraph lengths
  uint32_t* flag;  // [cap / 32 + 1] top n-grams: positions whose n-gram repeats (the next level's candidates)
};

template <class G, class IdxT>
TB_D void rep_body(const G& g, const KBatch& a, const int64_t* ts, const int64_t* te, const GrScratch& s, const GrCfg& cfg,
                   int step, tb_gr_stats* out, int64_t d, RepArrays A, uint32_t* tab, IdxT* cnt, IdxT* list) {
  (void)te;
  const uint32_t EMPTY = 0xFFFFFFFFu;
  const GrDoc gd = s.doc[d];
  tb_gr_stats st;
  st.reason_bits = 0; st.flags = gd.flags; st.char_len = gd.char_len; st.n_paras = 0; st.para_dup_elems = 0; st.para_dup_bytes = 0;
  st.n_lines = 0; st.line_dup_elems = 0; st.line_dup_bytes = 0; st.n_words = 0;
  for (int k = 0; k < TB_MAX_NGRAMS; k++) { st.top_bytes[k] = 0; st.dup_bytes[k] = 0; }
  if (gd.empty) {  // gopher_rep.rs:60-74
    if (g.tid == 0) { st.reason_bits = TB_GR_REASON_EMPTY; out[d] = st; a.alive[d] = 0; a.failed[d] = (int8_t)step; }
    return;
  }
  const int64_t rs = ts[d];
  const uint64_t uw = tb_wunit(rs, d), ul = tb_lunit(rs, d);
  const uint64_t* wh_g = s.wh + uw; const uint32_t* wl_g = s.wl + uw;
  const uint64_t* lh = s.lh + ul; const uint32_t* ll = s.ll + ul; const uint8_t* lg = s.lg + ul;
  const uint32_t W = gd.W, NL = gd.NL;
  const double text_char_len = (double)(gd.char_len ? gd.char_len : 1);
  uint32_t reasons = 0;

  // ---- lines (gopher_rep.rs:120-156)
  uint32_t de; uint64_t db;
  rep_dedup(g, tab, lh, ll, NL, &de, &db);
  st.n_lines = NL; st.line_dup_elems = de; st.line_dup_bytes = db;
  {
    double r1 = (double)de / (double)(NL ? NL : 1u), r2 = (double)db / text_char_len;
    if ((cfg.has & TB_GR_HAS_DUP_LINE_FRAC) && r1 > cfg.dup_line_frac) reasons |= TB_GR_REASON_DUP_LINE_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_LINE_CHAR_FRAC) && r2 > cfg.dup_line_char_frac) reasons |= TB_GR_REASON_DUP_LINE_CHAR_FRAC;
  }
  // ---- paragraphs (gopher_rep.rs:80-116): fold the ordered line pieces; a single '\n' joins lines
  // inside a paragraph, a run of >= 2 separates paragraphs.
  uint32_t NP = 0;
  if (g.tid == 0) {
    Hash2 cur; cur.init(); uint32_t curlen = 0;
    for (uint32_t k = 0; k < NL; k++) {
      Hash2 hk = tb_unpack(lh[k]);
      if (k == 0) { cur = hk; curlen = ll[k]; continue; }
      if (lg[k]) { A.H[NP] = cur.pack(); A.woff[NP] = curlen; NP++; cur = hk; curlen = ll[k]; }
      else { cur.push('\n'); cur = tb_hcat(cur, hk, tb_pow_pack(a.uc.powtab, ll[k])); curlen += 1 + ll[k]; }
    }
    if (NL) { A.H[NP] = cur.pack(); A.woff[NP] = curlen; NP++; }
  }
  NP = g.bcast(NP);
  g.sync();
  rep_dedup(g, tab, A.H, A.woff, NP, &de, &db);
  st.n_paras = NP; st.para_dup_elems = de; st.para_dup_bytes = db;
  {
    double r1 = (double)de / (double)(NP ? NP : 1u), r2 = (double)db / text_char_len;
    if ((cfg.has & TB_GR_HAS_DUP_PARA_FRAC) && r1 > cfg.dup_para_frac) reasons |= TB_GR_REASON_DUP_PARA_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_PARA_CHAR_FRAC) && r2 > cfg.dup_para_char_frac) reasons |= TB_GR_REASON_DUP_PARA_CHAR_FRAC;
  }
  g.sync();

  // ---- n-grams (gopher_rep.rs:159-196)
  uint32_t ntop_max = 0, ndup_max = 0;
  uint64_t top_levels = 0, dup_levels = 0;  // bit n: some entry of the list asks for n-grams of n words (n <= TB_MAX_NGRAM_N)
  for (int k = 0; k < cfg.n_top; k++) { ntop_max = cfg.top_n[k] > ntop_max ? cfg.top_n[k] : ntop_max; top_levels |= 1ull << cfg.top_n[k]; }
  for (int k = 0; k < cfg.n_dup; k++) { ndup_max = cfg.dup_n[k] > ndup_max ? cfg.dup_n[k] : ndup_max; dup_levels |= 1ull << cfg.dup_n[k]; }
  if (ntop_max > W) ntop_max = W;
  if (ndup_max > W) ndup_max = W;
  if (ntop_max || ndup_max) {
    // byte offsets of the words in the concatenation; word hashes into the working array
    uint32_t run = 0;
    for (uint32_t b0 = 0; b0 < W; b0 += g.size()) {
      uint32_t i = b0 + g.tid, v = i < W ? wl_g[i] : 0, tot, pre = g.exscan(v, &tot);
      if (i < W) A.woff[i] = run + pre;
      run += tot;
    }
    if (g.tid == 0) A.woff[W] = run;
    g.sync();
  }
  const uint64_t* wh = wh_g;  // read in place (coalesced, once per pass)
  const uint32_t* woff = A.woff;
  // -- top n-grams: keys are the " "-joined strings (find_top_duplicate, text.rs:211-238).  A joined key
  // determines its word sequence, so an n-gram that occurs twice has an (n-1)-gram prefix that occurs twice:
  // after the first counted level only the positions whose n-gram repeated are extended and counted again.
  {
    const uint32_t fw = W / 32 + 1;
    uint32_t count = W;     // positions still in play: all, or list[0 .. count)
    bool pruned = false;
    uint32_t cur = 0;       // level of the keys in A.H (0: none yet)
    for (uint32_t n = 1; n <= ntop_max && count; n++) {
      if (!((top_levels >> n) & 1ull)) continue;  // only configured levels run a pass; the keys jump levels in registers
      const uint32_t cn = W - n + 1;
      const uint32_t T = rep_tab_size(count);
      for (uint32_t i = g.tid; i < T; i += g.size()) tab[i] = EMPTY;
      for (uint32_t i = g.tid; i < fw; i += g.size()) A.flag[i] = 0;
      for (uint32_t k = g.tid; k < count; k += g.size()) {
        const uint32_t i = pruned ? (uint32_t)list[k] : k;
        cnt[i] = 0;
        if (i >= cn) continue;
        Hash2 h = tb_unpack(cur ? A.H[i] : TB_LDG(wh + i));
        for (uint32_t j = i + (cur ? cur : 1u); j < i + n; j++) {
          h.push(' ');
          h = tb_hcat(h, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(a.uc.powtab, woff[j + 1] - woff[j]));
        }
        A.H[i] = h.pack();
      }
      cur = n;
      g.sync();
      bool hit = false;
      for (uint32_t k = g.tid; k < count; k += g.size()) {
        const uint32_t i = pruned ? (uint32_t)list[k] : k;
        if (i >= cn) continue;
        const uint64_t key = A.H[i];
        uint32_t sl = rep_slot(key, T);
        for (;;) {
          uint32_t old = idx_cas(&tab[sl], EMPTY, i);
          if (old == EMPTY) { idx_inc(&cnt[i]); break; }
          if (A.H[old] == key) {
            idx_inc(&cnt[old]); hit = true;
            simt_atomic_or(&A.flag[i >> 5], 1u << (i & 31)); simt_atomic_or(&A.flag[old >> 5], 1u << (old & 31));
            break;
          }
          sl = rep_next(sl, T);
        }
      }
      uint64_t best = 0;
      const bool any_hit = g.any(hit);  // (a barrier: the counts are complete)
      if (any_hit) {
        for (uint32_t k = g.tid; k < count; k += g.size()) {
          const uint32_t i = pruned ? (uint32_t)list[k] : k;
          if (i >= cn) continue;
          const uint32_t c = cnt[i];
          // key length = concatenated bytes + (n - 1) spaces
          if (c > 1) { uint64_t v = ((uint64_t)c << 32) | (woff[i + n] - woff[i] + n - 1); best = v > best ? v : best; }
        }
        best = g.max64(best);
      }
      {
        const uint32_t c = (uint32_t)(best >> 32), len = (uint32_t)best;
        const uint64_t v = c > 1 ? (uint64_t)len * c : 0;
        for (int k = 0; k < cfg.n_top; k++) {
          if (cfg.top_n[k] != n) continue;
          st.top_bytes[k] = v;
          if ((double)v / text_char_len > cfg.top_frac[k]) reasons |= 1u << (TB_GR_REASON_TOP_SHIFT + k);
        }
      }
      // survivors: positions whose n-gram occurred at least twice, in order
      uint32_t ncount = 0;
      if (any_hit && n < ntop_max) {
        for (uint32_t b0 = 0; b0 < cn; b0 += g.size()) {
          const uint32_t i = b0 + g.tid;
          const uint32_t is = (i < cn && ((A.flag[i >> 5] >> (i & 31)) & 1u)) ? 1u : 0u;
          uint32_t tot, pre = g.exscan(is, &tot);
          if (is) list[ncount + pre] = (IdxT)i;
          ncount += tot;
        }
      }
      count = ncount; pruned = true;
      g.sync();
    }
  }
  // -- duplicate n-grams: keys are the concatenated strings (find_all_duplicate, text.rs:241-259)
  uint32_t dcur = 0;  // level of the keys in A.H (0: none yet)
  for (uint32_t n = 1; n <= ndup_max; n++) {
    if (!((dup_levels >> n) & 1ull)) continue;  // only configured levels run a pass; the keys jump levels in registers
    const uint32_t cn = W - n + 1;
    {
      const uint32_t T = rep_tab_size(cn);
      // one sweep: clear the table, bring the keys from level dcur to level n
#pragma unroll 2
      for (uint32_t i = g.tid; i < T; i += g.size()) {
        tab[i] = EMPTY;
        if (i >= cn) continue;
        Hash2 h = tb_unpack(dcur ? A.H[i] : TB_LDG(wh + i));
        for (uint32_t j = i + (dcur ? dcur : 1u); j < i + n; j++)
          h = tb_hcat(h, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(a.uc.powtab, woff[j + 1] - woff[j]));
        A.H[i] = h.pack();
      }
      dcur = n;
      g.sync();
      bool dup = false;
      for (uint32_t i = g.tid; i < cn; i += g.size()) {
        const uint64_t key = A.H[i];
        uint32_t sl = rep_slot(key, T);
        for (;;) {
          uint32_t old = idx_cas(&tab[sl], EMPTY, i);
          if (old == EMPTY) break;
          if (A.H[old] == key) { idx_min(&tab[sl], i); dup = true; break; }
          sl = rep_next(sl, T);
        }
        cnt[i] = (IdxT)sl;  // the slot of this key; resolved to its first position below
      }
      uint64_t v = 0;
      if (g.any(dup)) {  // (a barrier: the table is complete)
        // fo[i] = smallest position carrying the same key
        for (uint32_t i = g.tid; i < cn; i += g.size()) cnt[i] = (IdxT)tab[cnt[i]];
        g.sync();
        v = rep_dup_walk(g, cnt, tab, woff, cn, n);  // the table is dead now: reuse it for the candidate list
      }
      if (g.tid == 0) {
        for (int k = 0; k < cfg.n_dup; k++) {
          if (cfg.dup_n[k] != n) continue;
          st.dup_bytes[k] = v;
          if ((double)v / text_char_len > cfg.dup_frac[k]) reasons |= 1u << (TB_GR_REASON_DUP_SHIFT + k);
        }
      }
      // no barrier here: the next level's sweep only follows g.any() above or the one that ends rep_dup_walk
    }
  }
  if (g.tid == 0) {
    st.n_words = W; st.reason_bits = reasons;
    out[d] = st;
    if (reasons) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
  }
}
