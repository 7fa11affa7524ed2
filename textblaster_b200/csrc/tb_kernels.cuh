// tb_kernels.cuh -- device-side parameter blocks and host launch prototypes (internal).
#pragma once
#include <stdint.h>
#if defined(TB_HOST_EMUL)
typedef void* cudaStream_t;  // host emulation harness (tests/host_emul): no CUDA runtime
#else
#include <cuda_runtime.h>
#endif

#include "../../include/textblaster_b200.h"
#include "tb_lanes.cuh"
#include "tb_tile.cuh"
#include "tb_simt.cuh"

// Current view of the batch chunk on the device.
struct KBatch {
  const uint8_t* text;   // buffer the ranges index into
  const int64_t* rs;     // [n] range start (sot)
  const int64_t* re;     // [n] range end (eot)
  uint8_t* alive;        // [n] 1 while the document is still in the pipeline
  int8_t* failed;        // [n] index of the step that filtered it, -1
  int64_t n;
  UcTables uc;
  int64_t ext_lo, ext_hi;  // positions of `text` (and of the C4 work buffers) that 32-byte vector loads may touch
  uint32_t dbg;            // TB_DBG_*
  // Work order of the persistent kernels (built by tb_launch_order): long documents go first, longest class first, so
  // that none of them is started when everything else is already done.  order[0 .. counts[0]): the listed documents
  // (8 KiB and more); sizecls[d] != 0 marks a listed document.  counts: 16 words, zero before tb_launch_order.
  const uint32_t* order; const uint32_t* order_counts; const uint8_t* sizecls;
  uint32_t feed_min_tiles;  // documents of at least this many 1 KiB tiles get their tiles by bulk copy (tb_feed.cuh); 0: none
};
#define TB_LONG_DOC_BYTES (8 << 10)
#define TB_XLONG_DOC_BYTES (32 << 10)
#define TB_DBG_FORCE_WALK 1u  // run the segment-walking bodies for every document (tests)

struct GqCfg {
  uint32_t has;
  uint64_t min_doc_words, max_doc_words, min_stop_words;
  double min_avg_word_length, max_avg_word_length, max_symbol_word_ratio, max_bullet_lines_ratio,
      max_ellipsis_lines_ratio, max_non_alpha_words_ratio;
};

struct GrCfg {
  uint32_t has;
  int32_t n_top, n_dup;
  double dup_line_frac, dup_para_frac, dup_line_char_frac, dup_para_char_frac;
  uint32_t top_n[TB_MAX_NGRAMS];
  double top_frac[TB_MAX_NGRAMS];
  uint32_t dup_n[TB_MAX_NGRAMS];
  double dup_frac[TB_MAX_NGRAMS];
  // derived once per pipeline (tb_grcfg_finish): largest configured order and the set of configured orders per list
  uint32_t ntop_max, ndup_max;
  uint64_t top_levels, dup_levels;  // bit n: some entry asks for n-grams of n words (n <= TB_MAX_NGRAM_N)
};
inline void tb_grcfg_finish(GrCfg& c) {
  c.ntop_max = c.ndup_max = 0; c.top_levels = c.dup_levels = 0;
  for (int k = 0; k < c.n_top; k++) { if (c.top_n[k] > c.ntop_max) c.ntop_max = c.top_n[k]; c.top_levels |= 1ull << c.top_n[k]; }
  for (int k = 0; k < c.n_dup; k++) { if (c.dup_n[k] > c.ndup_max) c.ndup_max = c.dup_n[k]; c.dup_levels |= 1ull << c.dup_n[k]; }
}

// per-document result of the repetition scan pass
struct GrDoc {
  uint64_t char_len;
  uint32_t W, NL, flags, empty;
};

// Per-document scratch is addressed from the document's byte offset, without a prefix sum over the batch.
//   words: a text of L bytes holds at most (2L + 1) / 3 of them -- only the letters and digits are one-byte
//   word characters and two of those never break apart (WB5, WB8-10), so two adjacent words take at least
//   three bytes ("a\u00a9a\u00a9a": 5 words in 7 bytes).  Document d owns the units [tb_wunit(rs, d), tb_wunit(rs', d + 1)),
//   at least floor(2L / 3) + 1 of them.
//   line pieces: non-empty and separated by '\n', at most (L + 1) / 2: units from tb_lunit.
TB_HD uint64_t tb_wunit(int64_t rs, int64_t d) { return (uint64_t)((2 * rs) / 3) + (uint64_t)d; }
TB_HD uint64_t tb_lunit(int64_t rs, int64_t d) { return (uint64_t)(rs >> 1) + (uint64_t)d; }
TB_HD uint32_t tb_wunits_of(int64_t len) { return (uint32_t)((2 * len) / 3) + 1u; }  // word slots a text of len bytes needs at most

// Tiers by document size (max of words and line pieces): six in shared memory, capacity doubling from tier to
// tier, and tier X (global work area) above.  Tiers 0 and 1 are one warp per document; from tier 2 on a CTA of
// capacity / 8 threads.  A document in flight costs its tier's shared memory and nothing more, so the many
// short documents run many to an SM: the pass is bound by the latency of its dependent phases, not by issue.
#define TB_REP_NT 7          // number of tiers, X included
#define TB_REP_TIER_X 6
#define TB_REP_CAP0 256u     // capacity of tier t: TB_REP_CAP0 << t
#define TB_REP_WARP_TIERS 2  // tiers 0, 1: one warp
TB_HD uint32_t tb_rep_cap(int tier) { return TB_REP_CAP0 << tier; }
TB_HD int tb_rep_threads(int tier) { return tier < TB_REP_WARP_TIERS ? 32 : (int)(tb_rep_cap(tier) >> 3); }
TB_HD uint32_t tb_rep_tier(uint32_t big) {
  uint32_t t = 0;
  while (t < TB_REP_TIER_X && big > tb_rep_cap((int)t)) t++;
  return t;
}
// GrScratch.counters: list lengths, work cursors, the scan pass's cursor
#define TB_GR_CNT_LEN(t) (t)
#define TB_GR_CNT_CUR(t) (8 + (t))
#define TB_GR_CNT_SCAN 16
#define TB_GR_COUNTERS 32

// scratch of the repetition filter; word arrays are indexed from tb_wunit(rs[d], d), line arrays from tb_lunit(rs[d], d)
struct GrScratch {
  uint64_t* wh; uint32_t* wl;               // words: hash, byte length
  uint64_t* ws;                             //        suffix signature of the words up to and including this one (tb_sig_push over
                                            //        their bytes: the last 16 bytes of the concatenation that ends with the word)
  uint64_t* lh; uint32_t* ll; uint8_t* lg;  // line pieces: hash, byte length, "preceded by >= 2 newlines"
  GrDoc* doc;                               // [n]
  uint32_t* work;                           // tier-X work area, TB_REP_K words per word unit + TB_REP_DOC_EXTRA per doc
  uint32_t* lists;                          // [TB_REP_NT][n] document ids per tier, filled by the scan pass
  uint32_t* counters;                       // [TB_GR_COUNTERS]: list lengths, work cursors, scan cursor (TB_GR_CNT_*, tb_rep.cuh)
  int64_t n;
};
#define TB_REP_K 8u           // tier-L words per unit: H[2U] woff[U+2] tab[2U+16] cnt[U] list[U] flag[U/32+1]
#define TB_REP_DOC_EXTRA 24u  // words: the constant parts of the above, padded to keep 8-byte alignment
#define TB_REP_MIN_TABLE 16u

// FineWebQualityFilter
struct FwCfg {
  double line_punct_thr, short_line_thr, char_duplicates_ratio, new_line_ratio;
  uint64_t short_line_length;
  uint32_t line_punct_exclude_zero;
  int32_t n_stop_chars;  // < 0: default set
  uint32_t stop_chars[TB_FW_MAX_STOP_CHARS];
};
// per-document result of the scan pass
struct FwDoc {
  uint64_t n_chars;  // chars other than '\n'
  uint32_t n_lines, n_stop, n_short, n_words, n_newlines, flags;
};
// non-blank lines of document d: hash / byte length at unit tb_lunit(rs[d], d); tab: 2 words per unit + 16 per document
struct FwScratch {
  uint64_t* lh; uint32_t* ll;
  FwDoc* doc;
  uint32_t* tab;
};

// scratch of the C4 filter
struct C4Scratch {
  uint8_t* P;          // processed lines, doc d at [rs0[d], rs0[d] + plen[d])
  uint8_t* F;          // final content, doc d inside [rs0[d], ...)
  int64_t* plen;       // [n]
  uint32_t* nlines;    // [n] lines of the original content
  uint32_t* docflags;  // [n] C4F_*
  uint32_t* lwc;       // per line: word count          (indexed rs0[d] + d + line)
  uint32_t* lmax;      // per line: longest word, chars
  uint32_t* lflags;    // per line: C4L_*
  int64_t* prs; int64_t* pre;   // [n] ranges of P
  int64_t* frs; int64_t* fre;   // [n] ranges of F (untrimmed), then trimmed in place
  uint32_t* cnt;       // [n][4] too_long_word, no_terminal_punc, too_few_words, kept
  // One bit per byte of the text C4 reads (bit (pos + mis) of the bitmap, mis = address of the text buffer & 31, so
  // that a 32-byte stretch is one word): ka = the byte is part of its processed line (pass A: line trimmed,
  // citations cut, '\n' kept except the document's last); kf = the byte is part of the joined kept lines (pass C,
  // before the final trim; only written when kf != NULL).  Words shared by two documents are OR-ed: zero them first.
  uint32_t* ka; uint32_t* kf;
  uint8_t ftag = 1;    // what pass D writes into rewritten[d]: which of the two C4 output buffers F is (1, 2): the C4 steps of a chain write them in turn
};
// store the word of a stretch into such a bitmap (V: the bytes of the stretch that belong to this document)
TB_D void tb_bits_store(uint32_t* bits, int64_t p0, uint32_t mis, uint32_t V, uint32_t K) {
  if (!V) return;
  uint32_t* w = bits + (((uint64_t)p0 + mis) >> 5);
  if (V == 0xFFFFFFFFu) *w = K; else if (K) simt_atomic_or(w, K);
}

void tb_launch_init_ranges(cudaStream_t st, const void* off, int off64, int64_t doc0, int64_t n, int64_t base,
                           int64_t* rs, int64_t* re, uint8_t* alive, int8_t* failed);
// gq, gr_scan and c4_b run as persistent one-warp CTAs; `cursor` (gr_scan: s.counters[8]) must be zero at launch
// order / counts (zeroed by the caller) / sizecls for KBatch, from the original ranges
void tb_launch_order(cudaStream_t st, const int64_t* rs, const int64_t* re, int64_t n, uint32_t* order, uint32_t* counts, uint8_t* sizecls);
void tb_launch_gq(cudaStream_t st, const KBatch& b, const SwTable& sw, const GqCfg& cfg, int step, tb_gq_stats* out,
                  int sm_count, uint32_t* cursor);
void tb_launch_trim(cudaStream_t st, const KBatch& b, int64_t* ts, int64_t* te);
void tb_launch_gr_scan(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, const GrScratch& s, int sm_count);
void tb_launch_gr_rep(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, const GrScratch& s,
                      const GrCfg& cfg, int step, tb_gr_stats* out, int sm_count, int tier);  // tier 0 S, 1 M, 2 L, 3 X
void tb_launch_fw_scan(cudaStream_t st, const KBatch& b, const FwCfg& cfg, const FwScratch& s, int sm_count, uint32_t* cursor);
void tb_launch_fw_fin(cudaStream_t st, const KBatch& b, const FwCfg& cfg, const FwScratch& s, int step, tb_fw_stats* out, int sm_count,
                      uint32_t* cursor);
int tb_kernels_init();  // per-process function attributes (dynamic shared memory opt-in)
void tb_launch_c4_a(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor);
void tb_launch_c4_b(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor);
void tb_launch_c4_c(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int sm_count, uint32_t* cursor);
void tb_launch_c4_d(cudaStream_t st, const KBatch& b, const C4Cfg& cfg, const C4Scratch& s, int step, tb_c4_stats* out,
                    uint8_t* rewritten, int64_t* rs_out, int64_t* re_out, int sm_count, uint32_t* cursor);
// HTML entities decoded into D (ranges drs / dre): the pre-pass of TB_OPT_DECODE_HTML_ENTITIES
struct EntEntry;
void tb_launch_entities(cudaStream_t st, const KBatch& b, const EntEntry* tab, uint32_t ntab, uint32_t* amp, uint32_t* drop,
                        uint8_t* D, int64_t* drs, int64_t* dre, int sm_count, uint32_t* cursor);
// split_paragraph = false: the document rewritten with one sentence per line (S, ranges srs / sre)
void tb_launch_c4_sent(cudaStream_t st, const KBatch& b, const int64_t* ts, const int64_t* te, uint32_t* bits, uint8_t* S,
                       int64_t* srs, int64_t* sre, int sm_count, uint32_t* cursor);
void tb_launch_validate_utf8(cudaStream_t st, const uint8_t* text, const int64_t* rs, const int64_t* re, int64_t n,
                             int32_t* bad);
// final content as runs of the input text (TB_OPT_RETURN_CONTENT = 2): emit = 0 counts the runs of every document into
// nruns, emit = 1 writes them (8 bytes each: begin, end) at run_off[d]; at most cap runs are written
void tb_launch_c4_runs(cudaStream_t st, int64_t n, const int64_t* rs0, const int64_t* re0, const uint8_t* rewritten, const uint32_t* kf,
                       uint32_t mis, const int64_t* frs, const int64_t* fre, int64_t* nruns, const int64_t* run_off, void* runs, int64_t cap, int emit);
// final content: lengths -> exclusive scan -> packed copy
void tb_launch_final_len(cudaStream_t st, int64_t n, const int64_t* rs0, const int64_t* re0, const int64_t* rs,
                         const int64_t* re, const uint8_t* rewritten, int64_t* len);
void tb_launch_scan_i64(cudaStream_t st, const int64_t* in, int64_t* out, int64_t n);  // out[n] = total
void tb_launch_pack(cudaStream_t st, int64_t n, const uint8_t* orig, const int64_t* rs0, const uint8_t* cur, const uint8_t* cur2,
                    const int64_t* rs, const uint8_t* rewritten, const int64_t* out_off, uint8_t* out);
