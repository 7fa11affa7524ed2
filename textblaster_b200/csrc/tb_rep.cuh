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
//
// Duplicate n-grams, pruned.  Their keys are bare concatenations, so levels cannot be pruned by one
// another ("no where" == "now here" without any shared shorter n-gram).  But two equal keys are equal
// byte strings, so they end in the same k bytes for every k up to their length.  With m = the shortest
// n-gram of the smallest pruned level (in bytes) and k = min(16, m), every n-gram of every pruned level is
// at least k bytes long: a position whose last k key bytes (4 bits of each: the suffix signatures the scan
// pass stores per word, tb_sig_push) occur nowhere else in the document can neither repeat nor be repeated
// at any of those levels.  One cheap table pass over the signatures finds the word ends that can; the
// levels then hash, insert and walk those only -- a few percent of the positions of ordinary prose.
#pragma once
#include "tb_kernels.cuh"
#include "tb_simt.cuh"

// -DTB_REP_PROFILE (a measurement build, tools/build_variant.sh): thread 0 of every document's group adds the clocks it
// spent in each phase of rep_body to tb_rep_prof[tier][phase]; tb_debug_rep_profile (tb_kernels.cu) reads them out.
#if defined(TB_REP_PROFILE) && !defined(TB_HOST_EMUL)
static __device__ unsigned long long tb_rep_prof[8][32];
#define REP_PROF_BEGIN(tier_) const int prof_tier_ = (tier_); long long prof_t0_ = clock64();
#define REP_PROF(ph) { if (g.tid == 0) { const long long t1_ = clock64(); atomicAdd(&tb_rep_prof[prof_tier_][ph], (unsigned long long)(t1_ - prof_t0_)); prof_t0_ = t1_; } }
#define REP_PROF_COUNT(slot) { if (g.tid == 0) atomicAdd(&tb_rep_prof[prof_tier_][slot], 1ull); }
#else
#define REP_PROF_BEGIN(tier_)
#define REP_PROF(ph) {}
#define REP_PROF_COUNT(slot) {}
#endif

#if defined(TB_HOST_EMUL)
extern uint64_t tbe_stat_rep_words, tbe_stat_rep_items, tbe_stat_rep_pruned;  // word ends / candidates among them / documents pruned
extern uint64_t tbe_stat_rep_task_small, tbe_stat_rep_task_table;             // levels a warp of a CTA tier worked off alone: in registers / with a table
#endif

#ifndef TB_REP_X_THREADS
#define TB_REP_X_THREADS 512
#endif
#ifndef TB_REP_X_PER_SM
#define TB_REP_X_PER_SM 3
#endif
#define TB_REP_TASK_ITEMS 256u  // CTA tiers: up to this many pruned word ends, a warp works a whole level off by itself (capacity / 8 threads = capacity / 256 warps share the arrays)
#define TB_REP_PRUNE_FROM 4u  // duplicate n-gram levels from this order on run over the signature candidates only
#define TB_REP_SIG_MIN 6u     // ... if the shortest n-gram of the first such level has at least this many bytes

// ---- thread groups ---------------------------------------------------------------------------
struct WarpGroup {
  int tid;
  static constexpr int kSize = 32;
  TB_D int size() const { return 32; }
  TB_D void sync() const { simt_sync(); }
  TB_D uint32_t sum(uint32_t v) const { return simt_sum(v); }
  TB_D uint64_t sum64(uint64_t v) const { return simt_sum64(v); }
  TB_D uint64_t max64(uint64_t v) const { return simt_max64(v); }
  TB_D uint32_t max32(uint32_t v) const { return simt_max(v); }
  TB_D bool any(bool p) const { simt_sync(); return simt_any(p); }  // also a barrier with memory ordering, like the CTA's
  TB_D uint32_t exscan(uint32_t v, uint32_t* total) const { return simt_exscan(v, tid, total); }
  TB_D uint32_t bcast(uint32_t v) const { return simt_shfl(v, 0); }
  TB_D uint32_t warp_ballot(bool p) const { return simt_ballot(p); }
};

#if !defined(TB_HOST_EMUL)
// One CTA; `red` is 64 u64 of shared scratch, used as two alternating halves so that a collective needs a
// single barrier: a half is rewritten only after every thread has passed the barrier of the collective
// that used the other half.
struct BlockGroup {
  int tid;
  uint64_t* red;
  mutable int phase = 0;
  static constexpr int kSize = 0;  // (not one warp)
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
  TB_D uint32_t max32(uint32_t v) const { return (uint32_t)max64(v); }
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
  TB_D uint32_t warp_ballot(bool p) const { return __ballot_sync(0xFFFFFFFFu, p); }  // over the 32 threads tid & ~31 .. | 31
};
#else
// host emulation: N threads share one barrier + mailbox (tb_simt.cuh SimtBlock)
struct BlockGroup {
  int tid;
  uint64_t* red;  // the collectives go through the emulated CTA's mailbox; the words behind the first 64 are used as on the device
  mutable int phase = 0;
  static constexpr int kSize = 0;
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
  TB_D uint32_t max32(uint32_t v) const { return (uint32_t)max64(v); }
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
  TB_D uint32_t warp_ballot(bool p) const {
    SimtBlock* b = simt_block; b->box[tid] = p ? 1 : 0; b->bar.arrive_and_wait();
    uint32_t r = 0; const int w0 = tid & ~31;
    for (int i = 0; i < 32 && w0 + i < b->n; i++) r |= (uint32_t)b->box[w0 + i] << i;
    b->bar.arrive_and_wait(); return r;
  }
};
#endif

// num / den > thr in double precision (the reference's f64 comparison); out of line: a double division is some forty
// instructions and the repetition pass has a dozen of them, all rarely on the critical path
TB_HD_NOINLINE bool rep_frac_gt_slow(uint64_t num, double den, double thr) { return (double)num / den > thr; }
// (most numerators are zero -- nothing repeated -- and 0 / den is +0.0 for the positive denominators used here: no division then)
TB_HD bool rep_frac_gt(uint64_t num, double den, double thr) { return num == 0 ? 0.0 > thr : rep_frac_gt_slow(num, den, thr); }

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
// slot for a key that is not a hash (the suffix signatures: four bits of each byte): the top half of one 64-bit product
// depends on every bit of the key
TB_D uint32_t rep_slot_raw(uint64_t key, uint32_t T) {
  const uint32_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32);
  return (uint32_t)(((uint64_t)h * T) >> 32);
}
TB_D uint32_t rep_next(uint32_t sl, uint32_t T) { return sl + 1 == T ? 0u : sl + 1; }
template <class G>
TB_D void rep_fill(const G& g, uint32_t* a, uint32_t n, uint32_t v) {
#pragma unroll 1
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
#pragma unroll 1
  for (uint32_t i = g.tid; i < n; i += g.size()) {
    uint64_t hi = h[i]; uint32_t li = len[i];
    uint32_t sl = rep_slot(hi, T);
    total += li;
#pragma unroll 1
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

// Ordered list of the set bits of a bitmap: out[0 .. count) = indices i < nbits with bit i set, increasing.
// One prefix sum over the bitmap's words (a thread per word), then every thread expands its own word.
// span > 1: only the indices i whose bits i, i + 1, ..., i + span - 1 are all set (span <= 32; bits at and beyond nbits
// count as clear).
template <class G, class IdxT>
TB_D uint32_t rep_compact_bits(const G& g, const uint32_t* bits, uint32_t nbits, IdxT* out, uint32_t span = 1) {
  const uint32_t nwords = (nbits + 31u) >> 5;
  const uint32_t lastmask = (nbits & 31u) ? (1u << (nbits & 31u)) - 1u : 0xFFFFFFFFu;
  uint32_t run = 0;
#pragma unroll 1
  for (uint32_t b0 = 0; b0 < nwords; b0 += (uint32_t)g.size()) {
    const uint32_t wi = b0 + (uint32_t)g.tid;
    uint32_t w = wi < nwords ? bits[wi] : 0u;
    if (wi == nwords - 1u) w &= lastmask;
    if (span > 1u && w) {
      uint32_t nx = wi + 1u < nwords ? bits[wi + 1u] : 0u;
      if (wi + 2u == nwords) nx &= lastmask;
      const uint64_t both = ((uint64_t)nx << 32) | w;
#pragma unroll 1
      for (uint32_t s = 1; s < span && w; s++) w &= (uint32_t)(both >> s);
    }
    uint32_t tot, o = run + g.exscan((uint32_t)simt_popc(w), &tot);
#pragma unroll 1
    while (w) { out[o++] = (IdxT)(wi * 32u + (uint32_t)(simt_ffs(w) - 1)); w &= w - 1u; }
    run += tot;
  }
  g.sync();
  return run;
}

// The greedy scan of find_all_duplicate (text.rs:241-259) over `items` n-grams given in increasing position:
// item k is the n-gram that ENDS with word E(k) (E(k) = elist[k], or k itself without a list), i.e. starts at
// q = E(k) - (n - 1); fo[k] = first item with the same key.  Items whose key is new (fo[k] == k) only ever insert;
// they are pre-marked "seen".  A hit at q skips q+1..q+n-1: first occurrences inside that window were never
// visited, so their mark is removed; a later occurrence of such a key then inserts instead of hitting.  Only
// the candidates (fo[k] != k; bitmap `cbits`), in increasing order, need the sequential pass.  n-grams that are
// not among the items are unique in the document: nothing ever refers to them.
template <class G, class IdxT>
TB_D uint64_t rep_dup_walk(const G& g, IdxT* fo, IdxT* cand, const uint32_t* cbits, const uint32_t* woff, const IdxT* elist,
                           uint32_t items, uint32_t n) {
  const uint32_t EMPTY = IdxTraits<IdxT>::kEmpty;
  const uint32_t ncand = rep_compact_bits(g, cbits, items, cand);
  uint64_t rep = 0;
  if (g.tid == 0) {
    uint32_t skip_until = 0;
#pragma unroll 1
    for (uint32_t kk = 0; kk < ncand; kk++) {
      const uint32_t k = cand[kk];
      const uint32_t e = elist ? (uint32_t)elist[k] : k, q = e + 1u - n;
      if (q < skip_until) continue;
      const uint32_t c = fo[k];
      if (fo[c] == c) {  // key seen: hit
        rep += woff[e + 1] - woff[q];
#pragma unroll 1
        for (uint32_t p = k + 1; p < items && (elist ? (uint32_t)elist[p] : p) < e + n; p++) if (fo[p] == p) fo[p] = (IdxT)EMPTY;
        skip_until = q + n;
      } else {
        fo[c] = (IdxT)c;  // its first occurrence was skipped: this one inserts
      }
    }
  }
  g.sync();
  return rep;  // valid on tid 0
}

// One level of find_all_duplicate over the items list[0 .. items) (word ends, increasing; NULL: item i ends with word i):
// the keys of level n in H (brought up from level dcur, 0: from scratch -- new words go in front), a table of first
// items per key, the candidates, the walk.  Returns the repeated bytes (valid on tid 0).  tab: rep_tab_size(items)
// words; H, cnt: one entry per item; flag: a bit per item.  Ends with a barrier of the group.
template <class G, class IdxT>
TB_D uint64_t rep_dup_level(const G& g, const uint64_t* wh, const uint32_t* woff, const uint64_t* powtab, const IdxT* list, uint32_t items,
                            uint32_t n, uint32_t dcur, uint64_t* H, uint32_t* tab, IdxT* cnt, uint32_t* flag) {
  const uint32_t EMPTY = 0xFFFFFFFFu;
  const uint32_t T = rep_tab_size(items);
  // one sweep: clear the table, build the keys
#pragma unroll 2
  for (uint32_t i = g.tid; i < T; i += g.size()) {
    tab[i] = EMPTY;
    if (i >= items) continue;
    const uint32_t e = list ? (uint32_t)list[i] : i;
    if (e + 1u < n) continue;
    const uint32_t q = e + 1u - n;  // first word of the n-gram
    Hash2 h = tb_unpack(TB_LDG(wh + q));
    const uint32_t stop = dcur ? e + 1u - dcur : e + 1u;  // the words [stop, e] are in the stored key already
#pragma unroll 1
    for (uint32_t j = q + 1; j < stop; j++) h = tb_hcat(h, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(powtab, woff[j + 1] - woff[j]));
    if (dcur) h = tb_hcat(h, tb_unpack(H[i]), tb_pow_pack(powtab, woff[e + 1] - woff[stop]));
    H[i] = h.pack();
  }
  g.sync();
  bool dup = false;
#pragma unroll 1
  for (uint32_t i = g.tid; i < items; i += g.size()) {
    const uint32_t e = list ? (uint32_t)list[i] : i;
    if (e + 1u < n) { cnt[i] = (IdxT)EMPTY; continue; }
    const uint64_t key = H[i];
    uint32_t sl = rep_slot(key, T);
#pragma unroll 1
    for (;;) {
      uint32_t old = idx_cas(&tab[sl], EMPTY, i);
      if (old == EMPTY) break;
      if (H[old] == key) { idx_min(&tab[sl], i); dup = true; break; }
      sl = rep_next(sl, T);
    }
    cnt[i] = (IdxT)sl;  // the slot of this key; resolved to its first item below
  }
  uint64_t v = 0;
  if (g.any(dup)) {  // (a barrier: the table is complete)
    // fo[i] = first item carrying the same key; the items that are not the first of their key form the bitmap
    // of candidates (assembled a word per warp from ballots: no atomics)
    const uint32_t rounds = (items + (uint32_t)g.size() - 1u) / (uint32_t)g.size();
#pragma unroll 1
    for (uint32_t r = 0; r < rounds; r++) {
      const uint32_t i = r * (uint32_t)g.size() + (uint32_t)g.tid;
      uint32_t f = i;
      if (i < items && cnt[i] != (IdxT)EMPTY) f = tab[cnt[i]];
      const uint32_t bits = g.warp_ballot(f != i);
      if (i < items) cnt[i] = (IdxT)f;
      if ((g.tid & 31) == 0 && i < items) flag[i >> 5] = bits;
    }
    g.sync();
    v = rep_dup_walk(g, cnt, (IdxT*)tab, flag, woff, list, items, n);  // the table is dead now: it holds the candidate list
  }
  return v;
}

// ---- sets of at most 32 elements, one per lane of a warp, compared in registers (no table) --------------------
// lowest lane holding the same key; lanes that hold no element (holder = false) match nobody.  Keys are packed
// pairs of 31-bit residues: bits 63 and 31 are clear, so the stand-in of a non-holder equals no key.
TB_D uint32_t rep_small_first(uint64_t key, bool holder, int lane) {
  const uint32_t m = simt_match_any64(holder ? key : (0x8000000000000000ull | (uint64_t)(uint32_t)lane));
  return (uint32_t)(simt_ffs(m) - 1);
}
// find_duplicates (text.rs:197-208) over n <= 32 pieces, piece `lane` on lane `lane`
TB_D void rep_dedup_small(uint64_t h, uint32_t len, uint32_t n, int lane, uint32_t* dup_elems, uint64_t* dup_bytes) {
  const bool holder = (uint32_t)lane < n;
  const uint32_t first = rep_small_first(h, holder, lane);  // (every lane takes part in the match)
  const bool dup = holder && first != (uint32_t)lane;
  *dup_elems = (uint32_t)simt_popc(simt_ballot(dup));
  const uint32_t l = dup ? len : 0u;  // 32 lengths below 2^32: two 32-bit sums of 16-bit halves cannot overflow
  *dup_bytes = (uint64_t)simt_sum(l & 0xFFFFu) + ((uint64_t)simt_sum(l >> 16) << 16);
}
// Lines and paragraphs of a document with NL <= 32 line pieces (gopher_rep.rs:80-156), by one warp.  The fold of the
// pieces into paragraphs runs on every lane alike (uniform loads); lane p keeps paragraph p.
TB_D void rep_lines_small(const uint64_t* lh, const uint32_t* ll, const uint8_t* lg, uint32_t NL, const uint64_t* powtab, int lane,
                          uint32_t* line_de, uint64_t* line_db, uint32_t* NPo, uint32_t* para_de, uint64_t* para_db) {
  const bool holder = (uint32_t)lane < NL;
  const uint64_t myline = holder ? lh[lane] : 0ull;
  const uint32_t mylen = holder ? ll[lane] : 0u;
  rep_dedup_small(myline, mylen, NL, lane, line_de, line_db);
  // paragraphs: a piece that follows a run of >= 2 newlines (lg) opens one, the others join the paragraph before them
  // with a '\n' in between.  Joining is associative: an inclusive segmented scan over the lanes leaves every
  // paragraph's hash and length on the lane of its last piece.
  const bool head = holder && (lane == 0 || lg[lane] != 0);
  Hash2 v = tb_unpack(myline);
  uint32_t vl = mylen, f = head ? 1u : 0u;
#pragma unroll 1
  for (int off = 1; off < (int)NL; off <<= 1) {  // (no piece has more than NL - 1 pieces before it)
    const uint32_t ua = simt_shfl_up(v.a, off), ub = simt_shfl_up(v.b, off), ul = simt_shfl_up(vl, off), uf = simt_shfl_up(f, off);
    if (lane >= off && !f && holder) {
      Hash2 x; x.a = ua; x.b = ub;
      x.push('\n', TB_LDG(powtab + 1));
      v = tb_hcat(x, v, tb_pow_pack(powtab, vl));
      vl += ul + 1u; f = uf;
    }
  }
  const uint32_t heads = simt_ballot(head);
  const bool last = holder && ((uint32_t)lane + 1u == NL || ((heads >> ((lane + 1) & 31)) & 1u));  // the next piece opens a paragraph
  const uint32_t NP = (uint32_t)simt_popc(heads);
  *NPo = NP;
  // find_duplicates over the paragraphs, each on the lane of its last piece (the order of the lanes is their order)
  const uint32_t first = rep_small_first(v.pack(), last, lane);
  const bool dup = last && first != (uint32_t)lane;
  *para_de = (uint32_t)simt_popc(simt_ballot(dup));
  const uint32_t l = dup ? vl : 0u;
  *para_db = (uint64_t)simt_sum(l & 0xFFFFu) + ((uint64_t)simt_sum(l >> 16) << 16);
}
// Paragraphs of a document with any number of line pieces, by one warp: 32 pieces at a time, the segmented scan of
// rep_lines_small within the chunk, the paragraph that is open at the chunk's start carried in.  Writes hash and
// length of paragraph p to H[p] / len[p] in order; returns their number.
TB_D uint32_t rep_paragraphs_scan(const uint64_t* lh, const uint32_t* ll, const uint8_t* lg, uint32_t NL, const uint64_t* powtab, int lane,
                                  uint64_t* H, uint32_t* len) {
  uint32_t NP = 0;
  Hash2 cv; cv.init(); uint32_t cl = 0;  // the open paragraph before this chunk (none for the first)
#pragma unroll 1
  for (uint32_t c0 = 0; c0 < NL; c0 += 32) {
    const uint32_t k = c0 + (uint32_t)lane;
    const bool holder = k < NL;
    const bool head = holder && (k == 0 || lg[k] != 0);
    Hash2 v = tb_unpack(holder ? lh[k] : 0ull);
    uint32_t vl = holder ? ll[k] : 0u, f = head ? 1u : 0u;
#pragma unroll 1
    for (int off = 1; off < 32; off <<= 1) {
      const uint32_t ua = simt_shfl_up(v.a, off), ub = simt_shfl_up(v.b, off), ul = simt_shfl_up(vl, off), uf = simt_shfl_up(f, off);
      if (lane >= off && !f && holder) {
        Hash2 x; x.a = ua; x.b = ub;
        x.push('\n', TB_LDG(powtab + 1));
        v = tb_hcat(x, v, tb_pow_pack(powtab, vl));
        vl += ul + 1u; f = uf;
      }
    }
    if (holder && !f && c0) {  // no paragraph opens at or before this piece in the chunk: it continues the carried one
      Hash2 x = cv;
      x.push('\n', TB_LDG(powtab + 1));
      v = tb_hcat(x, v, tb_pow_pack(powtab, vl));
      vl += cl + 1u;
    }
    const uint32_t heads = simt_ballot(head);
    const bool next_head = lane < 31 ? ((heads >> (lane + 1)) & 1u) != 0u : (k + 1u < NL && lg[k + 1u] != 0);
    const bool last = holder && (k + 1u == NL || next_head);
    const uint32_t lasts = simt_ballot(last);
    if (last) { const uint32_t p = NP + (uint32_t)simt_popc(lasts & ((1u << lane) - 1u)); H[p] = v.pack(); len[p] = vl; }
    NP += (uint32_t)simt_popc(lasts);
    cv.a = simt_shfl(v.a, 31); cv.b = simt_shfl(v.b, 31); cl = simt_shfl(vl, 31);
  }
  return NP;
}
// Duplicate n-grams of every configured level from n_from on (text.rs:241-259) when at most 32 word ends are in play:
// item `lane` is the n-gram that ends with word e (valid: lane < items).  Keys grow to the left in registers; equal
// keys are found with one match per level; the order-dependent walk over the candidates runs on all lanes alike,
// the "first item with my key" values staying in the lanes' registers (see rep_dup_walk for the rules).
// Writes st.dup_bytes / reason bits through `sink(level, bytes)`; returns nothing (valid on every lane).
template <class Sink>
TB_D void rep_dup_small(const uint64_t* wh, const uint32_t* woff, const uint64_t* powtab, uint32_t e, uint32_t items, uint32_t n_from,
                        uint32_t ndup_max, uint64_t dup_levels, int lane, Sink sink) {
  const uint32_t NONE = 0xFFFFFFFFu;
  const bool holder = (uint32_t)lane < items;
  // every level reaches one word further back from e: the ten or so word hashes an item needs over all levels sit in one
  // or two cache lines, asked for now so that the levels' loads find them in L1
  if (holder) { tb_prefetch(wh + (e + 1u >= ndup_max ? e + 1u - ndup_max : 0u)); tb_prefetch(wh + e); }
  uint32_t dcur = 0;
  Hash2 h; h.init();
#pragma unroll 1
  for (uint32_t n = n_from; n <= ndup_max; n++) {
    if (!((dup_levels >> n) & 1ull)) continue;
    const bool valid = holder && e + 1u >= n;
    const uint32_t q = valid ? e + 1u - n : 0u;
    if (valid) {
      const uint32_t stop = dcur ? e + 1u - dcur : e + 1u;  // the words [stop, e] are in h already
      Hash2 p = tb_unpack(TB_LDG(wh + q));
#pragma unroll 1
      for (uint32_t j = q + 1; j < stop; j++) p = tb_hcat(p, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(powtab, woff[j + 1] - woff[j]));
      h = dcur ? tb_hcat(p, h, tb_pow_pack(powtab, woff[e + 1] - woff[stop])) : p;
    }
    dcur = n;
    uint32_t fo = rep_small_first(h.pack(), valid, lane);  // first item with my key
    uint32_t cm = simt_ballot(valid && fo != (uint32_t)lane);  // candidates, in position order
    uint64_t rep = 0;
    if (cm) {
      const uint32_t len = valid ? woff[e + 1] - woff[q] : 0u;
      uint32_t skip_until = 0;
#pragma unroll 1
      while (cm) {
        const int k = simt_ffs(cm) - 1;
        cm &= cm - 1u;
        const uint32_t ek = simt_shfl(e, k), qk = ek + 1u - n;
        const uint32_t c = simt_shfl(fo, k);
        const uint32_t fc = simt_shfl(fo, (int)(c & 31u));
        const uint32_t lk = simt_shfl(len, k);
        if (qk < skip_until) continue;
        if (fc == c) {  // key seen: hit; first occurrences inside the skipped window lose their mark
          rep += lk;
          if (valid && lane > k && e < ek + n && fo == (uint32_t)lane) fo = NONE;
          skip_until = qk + n;
        } else if ((uint32_t)lane == c) {
          fo = c;  // its first occurrence was skipped: this one inserts
        }
      }
    }
    sink(n, rep);
  }
}

// Top n-grams of the levels above `cur` (text.rs:211-238) when at most 32 positions are still in play: position i
// (its " "-joined key of level cur in h) on lane `lane`, valid for lane < count.  A level's count per key is the size
// of its match group; the positions whose n-gram repeated stay in play.
template <class Sink>
TB_D void rep_top_small(const uint64_t* wh, const uint32_t* woff, const uint64_t* powtab, uint32_t i, Hash2 h, uint32_t count, uint32_t cur,
                        uint32_t ntop_max, uint64_t top_levels, uint32_t W, int lane, Sink sink) {
  bool valid = (uint32_t)lane < count;
  if (valid) { tb_prefetch(wh + i + cur); tb_prefetch(wh + (i + ntop_max < W ? i + ntop_max : W - 1u)); }  // the words the levels will append
#pragma unroll 1
  for (uint32_t n = cur + 1; n <= ntop_max; n++) {
    if (!((top_levels >> n) & 1ull)) continue;
    const bool v = valid && i + n <= W;
    if (v) {
#pragma unroll 1
      for (uint32_t j = i + cur; j < i + n; j++) {
        h.push(' ', TB_LDG(powtab + 1));
        h = tb_hcat(h, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(powtab, woff[j + 1] - woff[j]));
      }
    }
    cur = n;
    const uint32_t m = simt_match_any64(v ? h.pack() : (0x8000000000000000ull | (uint64_t)(uint32_t)lane));
    const uint32_t c = v ? (uint32_t)simt_popc(m) : 0u;
    const uint32_t cmax = simt_max(c);
    uint64_t val = 0;
    if (cmax > 1) val = (uint64_t)simt_max((v && c == cmax) ? woff[i + n] - woff[i] + n - 1 : 0u) * cmax;  // key length = bytes + (n - 1) spaces
    sink(n, val);
    valid = v && c > 1;
    if (cmax <= 1) break;  // nothing repeats: the higher levels stay 0
  }
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
  REP_PROF_BEGIN(sizeof(IdxT) == 4 ? TB_REP_TIER_X : (int)tb_rep_tier(gd.W > gd.NL ? gd.W : gd.NL))
  // The document's record was zeroed before the pass (the caller clears the statistics of a chunk): thread 0 writes the
  // fields that are not zero straight into it -- no 248-byte copy of the record in every thread's local memory.
  tb_gr_stats* const o = out + d;
  const bool writer = g.tid == 0;
  if (gd.empty) {  // gopher_rep.rs:60-74
    if (writer) { o->reason_bits = TB_GR_REASON_EMPTY; o->flags = gd.flags; o->char_len = gd.char_len; a.alive[d] = 0; a.failed[d] = (int8_t)step; }
    return;
  }
  const int64_t rs = ts[d];
  const uint64_t uw = tb_wunit(rs, d), ul = tb_lunit(rs, d);
  const uint64_t* wh_g = s.wh + uw; const uint32_t* wl_g = s.wl + uw; const uint64_t* wsg = s.ws + uw;
  const uint64_t* lh = s.lh + ul; const uint32_t* ll = s.ll + ul; const uint8_t* lg = s.lg + ul;
  const uint32_t W = gd.W, NL = gd.NL;
  const double text_char_len = (double)(gd.char_len ? gd.char_len : 1);
  uint32_t reasons = 0;

  // ---- lines (gopher_rep.rs:120-156) and paragraphs (:80-116)
  uint32_t de = 0; uint64_t db = 0;
  uint32_t NP = 0;
  if (NL <= 32u) {
    // few pieces: the first warp compares them in registers (no table, no barrier); only thread 0 needs the results
    if (g.tid < 32) {
      uint32_t pde; uint64_t pdb;
      rep_lines_small(lh, ll, lg, NL, a.uc.powtab, g.tid, &de, &db, &NP, &pde, &pdb);
      if (writer) {
        o->n_lines = NL; o->line_dup_elems = de; o->line_dup_bytes = db;
        o->n_paras = NP; o->para_dup_elems = pde; o->para_dup_bytes = pdb;
      }
      if ((cfg.has & TB_GR_HAS_DUP_PARA_FRAC) && rep_frac_gt(pde, (double)(NP ? NP : 1u), cfg.dup_para_frac)) reasons |= TB_GR_REASON_DUP_PARA_FRAC;
      if ((cfg.has & TB_GR_HAS_DUP_PARA_CHAR_FRAC) && rep_frac_gt(pdb, text_char_len, cfg.dup_para_char_frac)) reasons |= TB_GR_REASON_DUP_PARA_CHAR_FRAC;
      if ((cfg.has & TB_GR_HAS_DUP_LINE_FRAC) && rep_frac_gt(de, (double)(NL ? NL : 1u), cfg.dup_line_frac)) reasons |= TB_GR_REASON_DUP_LINE_FRAC;
      if ((cfg.has & TB_GR_HAS_DUP_LINE_CHAR_FRAC) && rep_frac_gt(db, text_char_len, cfg.dup_line_char_frac)) reasons |= TB_GR_REASON_DUP_LINE_CHAR_FRAC;
    }
  } else {
  rep_dedup(g, tab, lh, ll, NL, &de, &db);
  if (writer) { o->n_lines = NL; o->line_dup_elems = de; o->line_dup_bytes = db; }
  {
    if ((cfg.has & TB_GR_HAS_DUP_LINE_FRAC) && rep_frac_gt(de, (double)(NL ? NL : 1u), cfg.dup_line_frac)) reasons |= TB_GR_REASON_DUP_LINE_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_LINE_CHAR_FRAC) && rep_frac_gt(db, text_char_len, cfg.dup_line_char_frac)) reasons |= TB_GR_REASON_DUP_LINE_CHAR_FRAC;
  }
  // ---- paragraphs (gopher_rep.rs:80-116): fold the ordered line pieces; a single '\n' joins lines
  // inside a paragraph, a run of >= 2 separates paragraphs.
  if (g.tid < 32) NP = rep_paragraphs_scan(lh, ll, lg, NL, a.uc.powtab, g.tid, A.H, A.woff);  // (the first warp; the others wait at the broadcast)
  NP = g.bcast(NP);
  g.sync();
  rep_dedup(g, tab, A.H, A.woff, NP, &de, &db);
  if (writer) { o->n_paras = NP; o->para_dup_elems = de; o->para_dup_bytes = db; }
  {
    if ((cfg.has & TB_GR_HAS_DUP_PARA_FRAC) && rep_frac_gt(de, (double)(NP ? NP : 1u), cfg.dup_para_frac)) reasons |= TB_GR_REASON_DUP_PARA_FRAC;
    if ((cfg.has & TB_GR_HAS_DUP_PARA_CHAR_FRAC) && rep_frac_gt(db, text_char_len, cfg.dup_para_char_frac)) reasons |= TB_GR_REASON_DUP_PARA_CHAR_FRAC;
  }
  g.sync();
  }

  REP_PROF(1)  // lines and paragraphs
  // ---- n-grams (gopher_rep.rs:159-196)
  uint32_t ntop_max = cfg.ntop_max, ndup_max = cfg.ndup_max;
  const uint64_t top_levels = cfg.top_levels, dup_levels = cfg.dup_levels;  // bit n: some entry of the list asks for n-grams of n words
  if (ntop_max > W) ntop_max = W;
  if (ndup_max > W) ndup_max = W;
  if (ntop_max || ndup_max) {
    // byte offsets of the words in the concatenation: the lengths come into place first (coalesced, the loads in flight
    // together), then every thread sums a run of consecutive words, one prefix sum over the threads, and the run is
    // rewritten as offsets
#pragma unroll 8
    for (uint32_t i = g.tid; i < W; i += g.size()) A.woff[i + 1] = TB_LDG(wl_g + i);
    g.sync();
    const uint32_t per = (W + (uint32_t)g.size() - 1u) / (uint32_t)g.size();
    const uint32_t i0 = (uint32_t)g.tid * per < W ? (uint32_t)g.tid * per : W, i1 = i0 + per < W ? i0 + per : W;
    uint32_t sum = 0;
#pragma unroll 1
    for (uint32_t i = i0; i < i1; i++) sum += A.woff[i + 1];
    uint32_t tot, run = g.exscan(sum, &tot);
#pragma unroll 1
    for (uint32_t i = i0; i < i1; i++) { run += A.woff[i + 1]; A.woff[i + 1] = run; }
    if (g.tid == 0) A.woff[0] = 0;
    g.sync();
  }
  REP_PROF(2)  // word offsets
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
#pragma unroll 1
    for (uint32_t n = 1; n <= ntop_max && count; n++) {
      if (!((top_levels >> n) & 1ull)) continue;  // only configured levels run a pass; the keys jump levels in registers
      if (pruned && count <= 32u) {
        // few positions left: the first warp takes them into registers and finishes the list on its own
        uint32_t i = 0; Hash2 h; h.init();
        if ((uint32_t)g.tid < count) { i = (uint32_t)list[g.tid]; h = tb_unpack(A.H[i]); }
        g.sync();  // (the working arrays are free for what follows)
        if (g.tid < 32)
          rep_top_small(wh, woff, a.uc.powtab, i, h, count, cur, ntop_max, top_levels, W, g.tid, [&](uint32_t lvl, uint64_t v) {
#pragma unroll 1
            for (int k = 0; k < cfg.n_top; k++) {
              if (cfg.top_n[k] != lvl) continue;
              if (writer && v) o->top_bytes[k] = v;
              if (rep_frac_gt(v, text_char_len, cfg.top_frac[k])) reasons |= 1u << (TB_GR_REASON_TOP_SHIFT + k);
            }
          });
        break;
      }
      const uint32_t cn = W - n + 1;
      const uint32_t T = rep_tab_size(count);
      if (sizeof(IdxT) == 2 && !pruned) {
        // The first counted level of a shared-memory tier runs over every position: the word hashes come into shared
        // memory first (the table's space, with all the loads in flight together), so that the key loop waits for no
        // global load; the table is cleared once the keys are made.
        uint64_t* const whs = (uint64_t*)tab;
#pragma unroll 4
        for (uint32_t i = g.tid; i < W; i += g.size()) whs[i] = TB_LDG(wh + i);
#pragma unroll 1
        for (uint32_t i = g.tid; i < fw; i += g.size()) A.flag[i] = 0;
        g.sync();
#pragma unroll 1
        for (uint32_t i = g.tid; i < W; i += g.size()) {
          cnt[i] = 0;
          if (i >= cn) continue;
          Hash2 h = tb_unpack(whs[i]);
#pragma unroll 1
          for (uint32_t j = i + 1u; j < i + n; j++) {
            h.push(' ', TB_LDG(a.uc.powtab + 1));
            h = tb_hcat(h, tb_unpack(whs[j]), tb_pow_pack(a.uc.powtab, woff[j + 1] - woff[j]));
          }
          A.H[i] = h.pack();
        }
        g.sync();
#pragma unroll 1
        for (uint32_t i = g.tid; i < T; i += g.size()) tab[i] = EMPTY;
      } else {
#pragma unroll 1
        for (uint32_t i = g.tid; i < T; i += g.size()) tab[i] = EMPTY;
#pragma unroll 1
        for (uint32_t i = g.tid; i < fw; i += g.size()) A.flag[i] = 0;
#pragma unroll 1
        for (uint32_t k = g.tid; k < count; k += g.size()) {
          const uint32_t i = pruned ? (uint32_t)list[k] : k;
          cnt[i] = 0;
          if (i >= cn) continue;
          Hash2 h = tb_unpack(cur ? A.H[i] : TB_LDG(wh + i));
#pragma unroll 1
          for (uint32_t j = i + (cur ? cur : 1u); j < i + n; j++) {
            h.push(' ', TB_LDG(a.uc.powtab + 1));
            h = tb_hcat(h, tb_unpack(TB_LDG(wh + j)), tb_pow_pack(a.uc.powtab, woff[j + 1] - woff[j]));
          }
          A.H[i] = h.pack();
        }
      }
      cur = n;
      g.sync();
      if (!pruned) { REP_PROF(12) }
      bool hit = false;
#pragma unroll 1
      for (uint32_t k = g.tid; k < count; k += g.size()) {
        const uint32_t i = pruned ? (uint32_t)list[k] : k;
        if (i >= cn) continue;
        const uint64_t key = A.H[i];
        uint32_t sl = rep_slot(key, T);
#pragma unroll 1
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
      if (!pruned) { REP_PROF(13) }
      if (any_hit) {
#pragma unroll 1
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
#pragma unroll 1
        for (int k = 0; k < cfg.n_top; k++) {
          if (cfg.top_n[k] != n) continue;
          if (writer && v) o->top_bytes[k] = v;
          if (rep_frac_gt(v, text_char_len, cfg.top_frac[k])) reasons |= 1u << (TB_GR_REASON_TOP_SHIFT + k);
        }
      }
      if (!pruned) { REP_PROF(14) }
      // survivors, in order: an n'-gram (n' > n) that occurs twice holds n' - n + 1 n-grams that occur twice, so only
      // the positions i whose n-grams at i, i + 1, ..., i + (n' - n) all repeated can repeat at the next counted level n'
      uint32_t ncount = 0;
      if (any_hit && n < ntop_max) {
        uint32_t nn = n + 1;
#pragma unroll 1
        while (nn <= ntop_max && !((top_levels >> nn) & 1ull)) nn++;
        ncount = rep_compact_bits(g, A.flag, cn, list, nn - n + 1u > 32u ? 32u : nn - n + 1u);
      }
      if (!pruned) { REP_PROF(7) }  // (the first counted level, over every position)
      count = ncount; pruned = true;
      g.sync();
    }
  }
  REP_PROF(3)  // top n-grams
  // -- duplicate n-grams: keys are the concatenated strings (find_all_duplicate, text.rs:241-259).
  // An item is the n-gram that ends with word e; its key grows to the left from level to level.  Levels below
  // TB_REP_PRUNE_FROM run over every word end; from there on only over the ends whose suffix signature repeats.
  {
    const uint32_t fw = W / 32 + 1;
    uint32_t n_p = 0;  // first pruned level
#pragma unroll 1
    for (uint32_t n = TB_REP_PRUNE_FROM; n <= ndup_max && !n_p; n++) if ((dup_levels >> n) & 1ull) n_p = n;
    bool pruned = false;   // the items are list[0 .. items)
    uint32_t items = W;
    uint32_t dcur = 0;     // level of the keys in A.H (0: none yet)
#pragma unroll 1
    for (uint32_t n = 1; n <= ndup_max; n++) {
      if (!((dup_levels >> n) & 1ull)) continue;  // only configured levels run a pass; the keys jump levels in registers
      if (n == n_p) {
        // shortest n-gram of this level, in bytes
        uint32_t m = 0xFFFFFFFFu;
#pragma unroll 1
        for (uint32_t i = g.tid; i + n <= W; i += g.size()) { const uint32_t v = woff[i + n] - woff[i]; m = v < m ? v : m; }
        m = 0xFFFFFFFFu - g.max32(0xFFFFFFFFu - m);
        const uint32_t k = m < 16u ? m : 16u;
        REP_PROF(15)
        if (k >= TB_REP_SIG_MIN) {
          // word ends whose last k key bytes occur at another word end too (first end per value in the table)
          const uint64_t mask = tb_sig_mask(k);
          const uint32_t T = rep_tab_size(W - n + 1);
#pragma unroll 1
          for (uint32_t i = g.tid; i < T; i += g.size()) tab[i] = EMPTY;
#pragma unroll 1
          for (uint32_t i = g.tid; i < fw; i += g.size()) A.flag[i] = 0;
          // the last k bytes of the words up to e (the scan pass stores 16 of them), into the key array: the probes
          // below compare keys, and must not wait for global memory each time
#pragma unroll 4
          for (uint32_t e = n - 1 + g.tid; e < W; e += g.size()) A.H[e] = TB_LDG(wsg + e) & mask;
          g.sync();
          REP_PROF(16)
#pragma unroll 1
          for (uint32_t e = n - 1 + g.tid; e < W; e += g.size()) {
            const uint64_t key = A.H[e];
            uint32_t sl = rep_slot_raw(key, T);
#pragma unroll 1
            for (;;) {
              uint32_t old = idx_cas(&tab[sl], EMPTY, e);
              if (old == EMPTY) break;
              if (A.H[old] == key) { simt_atomic_or(&A.flag[e >> 5], 1u << (e & 31)); simt_atomic_or(&A.flag[old >> 5], 1u << (old & 31)); break; }
              sl = rep_next(sl, T);
            }
          }
          g.sync();
          REP_PROF(17)
          items = rep_compact_bits(g, A.flag, W, list);
          pruned = true; dcur = 0;
          REP_PROF(4)  // signature pass
          REP_PROF_COUNT(items <= 32u ? 9 : items <= TB_REP_TASK_ITEMS ? 10 : 11)
#if defined(TB_HOST_EMUL)
          if (g.tid == 0) { tbe_stat_rep_words += W; tbe_stat_rep_items += items; tbe_stat_rep_pruned++; }
#endif
        }
      }
      if (items == 0) continue;  // nothing can repeat at this level: dup_bytes stays 0
      if constexpr (G::kSize != 32) if (pruned && items <= TB_REP_TASK_ITEMS) {
        // A CTA, and few word ends in play: given the list, the remaining levels are independent of one another.  A warp
        // takes a level at a time and works it off on its own -- in registers for up to 32 items, else with a table in
        // its share of the working arrays -- so that the order-dependent walks of the levels run side by side instead
        // of one after the other with the rest of the CTA waiting.
        const int w = g.tid >> 5, lane = g.tid & 31, nw = g.size() >> 5;
        const WarpGroup wg{lane};
        uint64_t* const res = g.red + 64;  // [TB_MAX_NGRAM_N + 1] repeated bytes per level
        uint32_t t = 0;
#pragma unroll 1
        for (uint32_t n2 = n; n2 <= ndup_max; n2++) {
          if (!((dup_levels >> n2) & 1ull)) continue;
          if ((int)(t++ % (uint32_t)nw) != w) continue;
          uint64_t v = 0;
#if defined(TB_HOST_EMUL)
          if (lane == 0) { if (items <= 32u) tbe_stat_rep_task_small++; else tbe_stat_rep_task_table++; }
#endif
          if (items <= 32u)
            rep_dup_small(wh, woff, a.uc.powtab, (uint32_t)lane < items ? (uint32_t)list[lane] : 0u, items, n2, n2, 1ull << n2, lane,
                          [&](uint32_t, uint64_t val) { v = val; });
          else
            v = rep_dup_level<WarpGroup, IdxT>(wg, wh, woff, a.uc.powtab, list, items, n2, 0u, A.H + (uint32_t)w * TB_REP_TASK_ITEMS,
                                               tab + (uint32_t)w * 2u * TB_REP_TASK_ITEMS, cnt + (uint32_t)w * TB_REP_TASK_ITEMS,
                                               A.flag + (uint32_t)w * (TB_REP_TASK_ITEMS / 32u));
          if (lane == 0) res[n2] = v;
        }
        g.sync();
        if (g.tid == 0) {
#pragma unroll 1
          for (int k = 0; k < cfg.n_dup; k++) {
            const uint32_t lvl = cfg.dup_n[k];
            if (lvl < n || lvl > ndup_max) continue;
            const uint64_t v = res[lvl];
            if (writer && v) o->dup_bytes[k] = v;
            if (rep_frac_gt(v, text_char_len, cfg.dup_frac[k])) reasons |= 1u << (TB_GR_REASON_DUP_SHIFT + k);
          }
        }
        break;
      }
      if (pruned && items <= 32u) {
        // few word ends in play: the first warp takes them into registers and finishes the list on its own
        if (g.tid < 32)
          rep_dup_small(wh, woff, a.uc.powtab, (uint32_t)g.tid < items ? (uint32_t)list[g.tid] : 0u, items, n, ndup_max, dup_levels, g.tid,
                        [&](uint32_t lvl, uint64_t v) {
#pragma unroll 1
                          for (int k = 0; k < cfg.n_dup; k++) {
                            if (cfg.dup_n[k] != lvl) continue;
                            if (writer && v) o->dup_bytes[k] = v;
                            if (rep_frac_gt(v, text_char_len, cfg.dup_frac[k])) reasons |= 1u << (TB_GR_REASON_DUP_SHIFT + k);
                          }
                        });
        break;
      }
      const uint64_t v = rep_dup_level<G, IdxT>(g, wh, woff, a.uc.powtab, pruned ? list : (const IdxT*)nullptr, items, n, dcur, A.H, tab, cnt, A.flag);
      dcur = n;
      if (g.tid == 0) {
#pragma unroll 1
        for (int k = 0; k < cfg.n_dup; k++) {
          if (cfg.dup_n[k] != n) continue;
          if (writer && v) o->dup_bytes[k] = v;
          if (rep_frac_gt(v, text_char_len, cfg.dup_frac[k])) reasons |= 1u << (TB_GR_REASON_DUP_SHIFT + k);
        }
      }
      // no barrier here: the next level's sweep only follows g.any() above or the one that ends rep_dup_walk
    }
  }
  REP_PROF(5)  // duplicate n-gram levels
  if (g.tid == 0) {
    o->n_words = W; o->reason_bits = reasons; o->flags = gd.flags; o->char_len = gd.char_len;
    if (reasons) { a.alive[d] = 0; a.failed[d] = (int8_t)step; }
  }
  REP_PROF(6)  // result
  REP_PROF_COUNT(8)
}
