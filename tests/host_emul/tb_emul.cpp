// tb_emul.cpp -- TEST INFRASTRUCTURE ONLY (never linked into or loaded by the product).
//
// Runs the exact warp bodies of the CUDA kernels (textblaster_b200/csrc/tb_bodies.cuh) on the host:
// 32 fibers play the lanes of a warp, warp collectives become barriers (tb_simt.cuh, TB_HOST_EMUL).
// It exists so that the kernel logic can be checked against the oracle on a machine with no GPU
// (`pytest -m "not gpu"`); the GPU tests check the real kernels through the C ABI.  Scratch memory
// is filled with garbage first, like cudaMalloc'd memory, to expose reads of unwritten slots.
#define TB_HOST_EMUL 1
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "../../textblaster_b200/csrc/tb_bodies.cuh"
#include "../../textblaster_b200/csrc/tb_hostutil.h"
#include "../../textblaster_b200/csrc/unicode_tables.inc"
#include "../../textblaster_b200/csrc/entity_table.inc"

thread_local SimtWarp* simt_warp = nullptr;
thread_local int simt_lane = 0;
thread_local SimtBlock* simt_block = nullptr;
uint64_t tbe_stat_tiles = 0, tbe_stat_walk = 0;
uint64_t tbe_stat_rep_words = 0, tbe_stat_rep_items = 0, tbe_stat_rep_pruned = 0;
uint64_t tbe_stat_rep_task_small = 0, tbe_stat_rep_task_table = 0;
extern "C" void tbe_rep_stats(uint64_t* words, uint64_t* items, uint64_t* pruned, int reset) {
  *words = tbe_stat_rep_words; *items = tbe_stat_rep_items; *pruned = tbe_stat_rep_pruned;
  if (reset) tbe_stat_rep_words = tbe_stat_rep_items = tbe_stat_rep_pruned = 0;
}
extern "C" void tbe_rep_task_stats(uint64_t* small, uint64_t* table, int reset) {
  *small = tbe_stat_rep_task_small; *table = tbe_stat_rep_task_table;
  if (reset) tbe_stat_rep_task_small = tbe_stat_rep_task_table = 0;
}
extern "C" void tbe_stats(uint64_t* tiles, uint64_t* walk) { *tiles = tbe_stat_tiles; *walk = tbe_stat_walk; }
// the chunk plan of the host executor (tb_hostutil.h): sizes into out[0 .. cap), returns their number
extern "C" int64_t tbe_chunk_targets(int64_t total_bytes, int64_t chunk_bytes, int device_input, int64_t* out, int64_t cap) {
  std::vector<int64_t> t = tb_chunk_targets(total_bytes, chunk_bytes, device_input != 0);
  for (size_t i = 0; i < t.size() && (int64_t)i < cap; i++) out[i] = t[i];
  return (int64_t)t.size();
}

namespace {

// ---- fibers: the lanes of a group share one host thread and run round-robin between barriers ------------
#if defined(__x86_64__)
extern "C" void tbe_switch(void** save_sp, void* load_sp);
asm(".text\n.globl tbe_switch\n.type tbe_switch,@function\ntbe_switch:\n"
    "  pushq %rbp\n  pushq %rbx\n  pushq %r12\n  pushq %r13\n  pushq %r14\n  pushq %r15\n"
    "  movq %rsp, (%rdi)\n  movq %rsi, %rsp\n"
    "  popq %r15\n  popq %r14\n  popq %r13\n  popq %r12\n  popq %rbx\n  popq %rbp\n  ret\n"
    ".size tbe_switch,.-tbe_switch\n");
#else
#error "the host emulation harness switches fibers with x86-64 code"
#endif
struct FiberGroup {
  int n = 0, cur = 0, done = 0;
  std::vector<void*> sp;
  std::vector<char*> stacks;
  std::vector<char> finished;
  void* main_sp = nullptr;
  const std::function<void(int)>* body = nullptr;
  SimtWarp* warps = nullptr;  // block mode: the warp-level state of fibers 32 w .. 32 w + 31
};
FiberGroup* g_fibers = nullptr;  // one group at a time (the harness is single-threaded)
constexpr size_t kFiberStack = 256 << 10;
// make fiber `nx` the current one (its warp-level context with it) -- called right before switching to it
void fiber_select(FiberGroup* g, int nx) {
  g->cur = nx; simt_lane = nx & 31;
  if (g->warps) simt_warp = &g->warps[nx >> 5];
}
int fiber_next(FiberGroup* g, int me) {
  int nx = me;
  do { nx = (nx + 1) % g->n; } while (g->finished[nx] && nx != me);
  return nx;
}
void fiber_entry() {
  FiberGroup* g = g_fibers;
  const int me = g->cur;
  (*g->body)(me);
  g->done++; g->finished[me] = 1;
  void* dummy;
  if (g->done == g->n) tbe_switch(&dummy, g->main_sp);
  const int nx = fiber_next(g, me);
  fiber_select(g, nx);
  tbe_switch(&dummy, g->sp[nx]);
  abort();
}
// run body(0) .. body(n - 1) as fibers; a barrier inside hands over to the other fibers until all have arrived
void run_fibers(int n, const std::function<void(int)>& body, SimtWarp* warps = nullptr) {
  FiberGroup g;
  g.n = n; g.body = &body; g.sp.resize(n); g.stacks.resize(n); g.finished.assign(n, 0); g.warps = warps;
  for (int i = 0; i < n; i++) {
    char* st = (char*)malloc(kFiberStack);
    g.stacks[i] = st;
    uintptr_t top = ((uintptr_t)st + kFiberStack) & ~(uintptr_t)15;
    void** q = (void**)top;
    *--q = nullptr;                 // return address of fiber_entry (never used)
    *--q = (void*)&fiber_entry;     // `ret` of the first switch lands here
    for (int k = 0; k < 6; k++) *--q = nullptr;
    g.sp[i] = q;
  }
  g_fibers = &g;
  fiber_select(&g, 0);
  tbe_switch(&g.main_sp, g.sp[0]);
  g_fibers = nullptr;
  for (char* st : g.stacks) free(st);
}
}  // namespace
void simt_fiber_yield() {
  FiberGroup* g = g_fibers;
  const int me = g->cur, nx = fiber_next(g, me);
  if (nx == me) return;
  fiber_select(g, nx);
  tbe_switch(&g->sp[me], g->sp[nx]);
}
namespace {

// Launch `fn(d, lane)` for every document as one emulated warp (32 lanes in lock step at the
// collectives).  `pred(d)` is the warp-uniform early-exit test of the __global__ wrapper.
void run_warps(int64_t n, const std::function<bool(int64_t)>& pred, const std::function<void(int64_t, int)>& fn) {
  SimtWarp warp;
  simt_warp = &warp;
  run_fibers(32, [&](int lane) {
    for (int64_t d = 0; d < n; d++) {
      bool go = pred(d);
      warp.bar.arrive_and_wait();  // every lane has sampled the predicate before anyone runs
      if (!go) continue;
      fn(d, lane);
      warp.bar.arrive_and_wait();  // kernels never share state between warps mid-flight
    }
  });
}

// One emulated CTA of `nthreads` threads per listed document (block-level collectives only).
void run_blocks(const uint32_t* list, uint32_t count, int nthreads, const std::function<void(int64_t, int)>& fn) {
  SimtBlock blk(nthreads);
  simt_block = &blk;
  std::vector<SimtWarp> warps((size_t)(nthreads + 31) / 32);  // warp-level collectives inside the CTA (full warps only)
  run_fibers(nthreads, [&](int t) {
    for (uint32_t i = 0; i < count; i++) { fn((int64_t)list[i], t); blk.bar.arrive_and_wait(); }
  }, warps.data());
}
void run_warps_list(const uint32_t* list, uint32_t count, const std::function<void(int64_t, int)>& fn) {
  SimtWarp warp;
  simt_warp = &warp;
  run_fibers(32, [&](int lane) {
    for (uint32_t i = 0; i < count; i++) { fn((int64_t)list[i], lane); warp.bar.arrive_and_wait(); }
  });
}

template <class T> T* garbage(size_t n) {
  T* p = (T*)malloc(n * sizeof(T) + 64);
  memset(p, 0xCD, n * sizeof(T) + 64);
  return p;
}

size_t stat_size(int type) {
  return type == TB_STEP_C4_QUALITY ? sizeof(tb_c4_stats) : type == TB_STEP_GOPHER_QUALITY ? sizeof(tb_gq_stats)
         : type == TB_STEP_FINEWEB_QUALITY ? sizeof(tb_fw_stats) : sizeof(tb_gr_stats);
}

}  // namespace

extern "C" int tbe_run_chain(const tb_step_cfg* steps, size_t n_steps, const uint8_t* utf8, const int64_t* offsets,
                             int64_t n, uint8_t* keep, int8_t* failed_out, void** stats_out, uint8_t* content_out,
                             int64_t* content_off_out, int decode_entities) {
  std::vector<void*> frees;
  auto G8 = [&](size_t c) { auto* p = garbage<uint8_t>(c); frees.push_back(p); return p; };
  auto G32 = [&](size_t c) { auto* p = garbage<uint32_t>(c); frees.push_back(p); return p; };
  auto G64 = [&](size_t c) { auto* p = garbage<int64_t>(c); frees.push_back(p); return p; };
  const int64_t base = n ? offsets[0] : 0;
  const int64_t nb_in = n ? offsets[n] - base : 0;
  const int64_t nb = decode_entities ? (int64_t)tb_grown_bytes(nb_in, n) : nb_in;  // extent of the text the chain sees
  bool any_sent = false;
  for (size_t s = 0; s < n_steps; s++) any_sent |= steps[s].type == TB_STEP_C4_QUALITY && !steps[s].u.c4.split_paragraph;
  const size_t n1 = (size_t)n + 1, cbx = any_sent ? tb_grown_bytes(nb, n) : (size_t)nb;
  const size_t wunits = tb_word_units(cbx, (size_t)n), units = tb_line_units(cbx, (size_t)n);
  // tables
  std::vector<uint64_t> pw(TB_POWTAB_ALL);
  tb_fill_powtab(pw.data(), TB_B1_DEFAULT, TB_B2_DEFAULT);
  UcTables uc{tb_uc_stage1, tb_uc_stage2, &tb_uc_lower[0][0], &tb_uc_cased[0][0], &tb_uc_caseign[0][0], pw.data(),
              TB_UC_NLOWER, TB_UC_NCASED, TB_UC_NCASEIGN, {0}};
  tb_compute_wslead(tb_uc_stage1, tb_uc_stage2, uc.wslead);
  uint16_t ascii[192];
  for (int i = 0; i < 128; i++) {
    ascii[i] = tb_uc_stage2[(uint32_t)tb_uc_stage1[0] * 256u + i];
    ((uint8_t*)(ascii + 128))[i] = (uint8_t)tb_ascii_flags(ascii[i]);
    if (!tb_ascii_class_ok(ascii[i])) return 7;  // the ASCII fast path would not be exact
  }

  uint8_t* text = G8((size_t)nb_in + 64);
  if (nb_in) memcpy(text, utf8 + base, (size_t)nb_in);
  int64_t *rs0 = G64(n1), *re0 = G64(n1), *rs = G64(n1), *re = G64(n1), *ts = G64(n1), *te = G64(n1), *len = G64(n1), *out_off = G64(n1 + 1);
  uint8_t *alive = G8(n1), *rewritten = G8(n1);
  int8_t* failed = (int8_t*)G8(n1);
  for (int64_t d = 0; d < n; d++) body_init_ranges(offsets, 1, 0, base, rs0, re0, alive, failed, d);
  if (decode_entities) {  // the pre-pass of TB_OPT_DECODE_HTML_ENTITIES (run_impl): the decoded copy becomes the text
    const size_t bw = (size_t)nb_in / 32 + 4;
    uint32_t *amp = G32(bw), *drop = G32(bw);
    memset(amp, 0, bw * 4); memset(drop, 0, bw * 4);
    uint8_t* D = G8((size_t)nb + 64);
    int64_t *drs = G64(n1), *dre = G64(n1);
    uint8_t* estage = (uint8_t*)G64(TB_SENT_STAGE_BYTES / 8 + 2);
    KBatch eb{text, rs0, re0, alive, failed, n, uc, 0, nb_in, 0u};
    run_warps(n, [&](int64_t) { return true; }, [&](int64_t d, int lane) { body_entities(eb, estage, tb_ent_table, TB_ENT_COUNT, amp, drop, D, drs, dre, d, lane); });
    memcpy(rs0, drs, (size_t)n * 8); memcpy(re0, dre, (size_t)n * 8);
    text = D;
  }
  memcpy(rs, rs0, (size_t)n * 8); memcpy(re, re0, (size_t)n * 8);
  memset(rewritten, 0, (size_t)n);
  const char* fw = getenv("TBE_FORCE_WALK");
  KBatch b{text, rs, re, alive, failed, n, uc, 0, nb, (fw && fw[0] == '1') ? TB_DBG_FORCE_WALK : 0u};
  std::vector<LutEntry> lut(256);
  auto build_lut = [&](int kind, uint64_t firstmask) {
    for (uint32_t i = 0; i < 256; i++) lut[i] = tb_lut_entry(kind, i, i < 128 ? ascii[i] : 0u, (kind == TB_LUT_C4A || kind == TB_LUT_FW) ? (uint64_t)((uc.wslead[i >> 5] >> (i & 31)) & 1u) : firstmask);
  };
  auto alive_pred = [&](int64_t d) { return alive[d] != 0; };
  bool has_c4 = false;
  int n_c4 = 0, c4_seen = 0;
  for (size_t s = 0; s < n_steps; s++) n_c4 += steps[s].type == TB_STEP_C4_QUALITY;
  uint8_t* Fbuf[2] = {nullptr, nullptr};
  const uint32_t* runs_kf = nullptr; const int64_t* runs_frs = nullptr; const int64_t* runs_fre = nullptr;  // for the runs check below
  for (size_t s = 0; s < n_steps; s++) {
    const tb_step_cfg& sc = steps[s];
    uint8_t* stats = G8(n1 * stat_size(sc.type));
    if (sc.type == TB_STEP_GOPHER_QUALITY) {
      const tb_gq_params& q = sc.u.gq;
      GqCfg cfg{q.has, q.min_doc_words, q.max_doc_words, q.min_stop_words, q.min_avg_word_length, q.max_avg_word_length,
                q.max_symbol_word_ratio, q.max_bullet_lines_ratio, q.max_ellipsis_lines_ratio, q.max_non_alpha_words_ratio};
      uint32_t mask, maxlen, lenmask, firstmask, secondmask;
      std::vector<SwEntry> tab = tb_build_stop_table(tb_stop_words_of(q), &mask, &maxlen, &lenmask, &firstmask, &secondmask);
      SwTable sw{tab.data(), mask, maxlen, lenmask, firstmask, secondmask};
      build_lut(TB_LUT_GQ, ((uint64_t)secondmask << 32) | firstmask);
      uint32_t* gstage = G32(2 * TB_GQ_STAGE);
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_gq(b, ascii, lut.data(), gstage, sw, cfg, (int)s, (tb_gq_stats*)stats, d, lane); });
    } else if (sc.type == TB_STEP_GOPHER_REPETITION) {
      const tb_gr_params& q = sc.u.gr;
      GrCfg g{};
      g.has = q.has; g.n_top = q.n_top; g.n_dup = q.n_dup;
      g.dup_line_frac = q.dup_line_frac; g.dup_para_frac = q.dup_para_frac;
      g.dup_line_char_frac = q.dup_line_char_frac; g.dup_para_char_frac = q.dup_para_char_frac;
      for (int k = 0; k < TB_MAX_NGRAMS; k++) { g.top_n[k] = q.top_n[k]; g.top_frac[k] = q.top_frac[k]; g.dup_n[k] = q.dup_n[k]; g.dup_frac[k] = q.dup_frac[k]; }
      tb_grcfg_finish(g);
      GrScratch gs{(uint64_t*)G64(wunits), G32(wunits), (uint64_t*)G64(wunits), (uint64_t*)G64(units), G32(units), G8(units),
                   (GrDoc*)G8(n1 * sizeof(GrDoc)), G32(wunits * TB_REP_K + n1 * TB_REP_DOC_EXTRA), G32(TB_REP_NT * n1), G32(TB_GR_COUNTERS), n};
      memset(gs.counters, 0, TB_GR_COUNTERS * 4);
      memset(stats, 0, (size_t)n * sizeof(tb_gr_stats));  // (pass 2 writes the non-zero fields only)
      for (int64_t d = 0; d < n; d++) body_trim(b, ascii, ts, te, d);
      build_lut(TB_LUT_GR, 0u);
      uint32_t* stage = (uint32_t*)G64(TB_GR_STAGE_U32 / 2 + 1);
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_gr_scan(b, ascii, lut.data(), stage, ts, te, gs, d, lane); });
      for (int tier = 0; tier < TB_REP_TIER_X; tier++) {
        const uint32_t cap = tb_rep_cap(tier);
        uint8_t* smem = (uint8_t*)G64(TB_REP_CTA_SMEM(cap) / 8 + 2);
        if (tier < TB_REP_WARP_TIERS)
          run_warps_list(gs.lists + (int64_t)tier * n, gs.counters[TB_GR_CNT_LEN(tier)],
                         [&](int64_t d, int lane) { body_gr_rep_warp(b, smem, cap, ts, te, gs, g, (int)s, (tb_gr_stats*)stats, d, lane); });
        else
          run_blocks(gs.lists + (int64_t)tier * n, gs.counters[TB_GR_CNT_LEN(tier)], tb_rep_threads(tier),
                     [&](int64_t d, int tid) { body_gr_rep_cta(b, smem, cap, ts, te, gs, g, (int)s, (tb_gr_stats*)stats, d, tid); });
      }
      uint64_t* red_x = (uint64_t*)G64(64 + 40);
      run_blocks(gs.lists + (int64_t)TB_REP_TIER_X * n, gs.counters[TB_GR_CNT_LEN(TB_REP_TIER_X)], TB_REP_X_THREADS, [&](int64_t d, int tid) { body_gr_rep_x(b, red_x, ts, te, gs, g, (int)s, (tb_gr_stats*)stats, d, tid); });
    } else if (sc.type == TB_STEP_FINEWEB_QUALITY) {
      const tb_fw_params& q = sc.u.fw;
      FwCfg f{};
      f.line_punct_thr = q.line_punct_thr; f.short_line_thr = q.short_line_thr; f.char_duplicates_ratio = q.char_duplicates_ratio;
      f.new_line_ratio = q.new_line_ratio; f.short_line_length = q.short_line_length;
      f.line_punct_exclude_zero = q.line_punct_exclude_zero != 0; f.n_stop_chars = q.n_stop_chars;
      for (int k = 0; k < TB_FW_MAX_STOP_CHARS; k++) f.stop_chars[k] = q.stop_chars[k];
      FwScratch fs{(uint64_t*)G64(units), G32(units), (FwDoc*)G8(n1 * sizeof(FwDoc)), G32(units * 2 + n1 * 16)};
      build_lut(TB_LUT_FW, 0u);
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_fw_scan(b, ascii, lut.data(), f, fs, d, lane); });
      uint32_t* ftab = G32(2 * TB_FW_SMEM_LINES);
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_fw_fin(b, ftab, f, fs, (int)s, (tb_fw_stats*)stats, d, lane); });
    } else if (sc.type == TB_STEP_C4_QUALITY) {
      C4Cfg cfg = tb_c4cfg_of(sc.u.c4);
      const bool sent = !sc.u.c4.split_paragraph;
      if (n_c4 > 1 && sent) return 5;  // (several C4 steps: only with split_paragraph = true, as in tb_pipeline_create)
      const size_t cb = sent ? tb_grown_bytes(nb, n) : (size_t)nb, clines = cb + n1 + 1;
      const size_t bw = cb / 32 + 4;
      // the C4 steps of a chain write the two output buffers in turn (a step reads the one before it), like enqueue_chain
      if (!Fbuf[c4_seen & 1]) Fbuf[c4_seen & 1] = G8(cb + 64);
      C4Scratch cs{G8(cb + 64), Fbuf[c4_seen & 1], G64(n1), G32(n1), G32(n1), G32(clines), G32(clines), G32(clines),
                   G64(n1), G64(n1), G64(n1), G64(n1), G32(4 * n1), G32(bw), G32(bw), (uint8_t)(1 + (c4_seen & 1))};
      c4_seen++;
      memset(cs.ka, 0, bw * 4); memset(cs.kf, 0, bw * 4);
      const KBatch b_in = b;
      if (sent) {  // one sentence per line into S (the launch sequence of enqueue_chain)
        uint32_t* bits = G32((size_t)nb / 32 + 4);
        memset(bits, 0, ((size_t)nb / 32 + 4) * 4);
        uint8_t* S = G8(cb + 64);
        int64_t *srs = G64(n1), *sre = G64(n1);
        for (int64_t d = 0; d < n; d++) body_trim(b, ascii, ts, te, d);
        uint8_t* sstage = (uint8_t*)G64(TB_SENT_STAGE_BYTES / 8 + 2);
        for (int64_t d = 0; d < n; d++) if (!alive[d]) { srs[d] = tb_sent_off(rs[d], d); sre[d] = srs[d]; }
        run_warps(n, alive_pred, [&](int64_t d, int lane) { body_c4_sent(b, ascii, sstage, ts, te, bits, S, srs, sre, d, lane); });
        b.text = S; b.rs = srs; b.re = sre; b.ext_hi = (int64_t)cb;
      }
      build_lut(TB_LUT_C4A, 0u);
      uint8_t* cstage = (uint8_t*)G64(TB_STAGE_BYTES / 8 + 2);
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_c4_a(b, ascii, lut.data(), cstage, cfg, cs, d, lane); });
      build_lut(TB_LUT_C4B, 0u);
      run_warps(n, [&](int64_t d) { return alive[d] && !cs.docflags[d] && cs.nlines[d] != 0; },
                [&](int64_t d, int lane) { body_c4_b(b, ascii, lut.data(), cfg, cs, d, lane); });
      run_warps(n, [&](int64_t) { return true; }, [&](int64_t d, int lane) { body_c4_c(b, cstage, cfg, cs, d, lane); });
      { KBatch fb = b; fb.text = cs.F; fb.rs = cs.frs; fb.re = cs.fre; for (int64_t d = 0; d < n; d++) body_trim(fb, ascii, cs.frs, cs.fre, d); }
      run_warps(n, alive_pred, [&](int64_t d, int lane) { body_c4_d(b, ascii, cfg, cs, (int)s, (tb_c4_stats*)stats, rewritten, rs, re, d, lane); });
      b.rs = b_in.rs; b.re = b_in.re;
      b.text = cs.F;
      has_c4 = true;
      if (!sent && !decode_entities && n_c4 == 1) { runs_kf = cs.kf; runs_frs = cs.frs; runs_fre = cs.fre; }
    } else {
      return 1;
    }
    memcpy(stats_out[s], stats, (size_t)n * stat_size(sc.type));
  }
  memcpy(keep, alive, (size_t)n);
  memcpy(failed_out, failed, (size_t)n);
  if ((has_c4 || decode_entities) && content_out) {
    for (int64_t d = 0; d < n; d++) body_final_len(rs0, re0, rs, re, rewritten, len, d);
    int64_t acc = 0;
    for (int64_t d = 0; d < n; d++) { out_off[d] = acc; acc += len[d]; }
    out_off[n] = acc;
    run_warps(n, [&](int64_t) { return true; },
              [&](int64_t d, int lane) { body_pack(text, rs0, has_c4 ? Fbuf[0] : b.text, Fbuf[1], rs, rewritten, out_off, content_out, d, lane); });
    memcpy(content_off_out, out_off, n1 * 8);
    if (runs_kf) {
      // TB_OPT_RETURN_CONTENT = 2: the runs of the input text (count, offsets, write, as enqueue_chain does) must
      // reproduce the packed content byte for byte
      const uint32_t mis = (uint32_t)((uintptr_t)text & 31u);
      int64_t *nruns = G64(n1), *roff = G64(n1 + 1);
      run_warps(n, [&](int64_t) { return true; }, [&](int64_t d, int lane) {
        body_c4_runs(rs0, re0, rewritten, runs_kf, mis, runs_frs, runs_fre, nruns, nullptr, nullptr, 0, false, d, lane); });
      int64_t tot = 0;
      for (int64_t d = 0; d < n; d++) { roff[d] = tot; tot += nruns[d]; }
      roff[n] = tot;
      tb_run_dev* runs = (tb_run_dev*)G64((size_t)tot + 1);
      run_warps(n, [&](int64_t) { return true; }, [&](int64_t d, int lane) {
        body_c4_runs(rs0, re0, rewritten, runs_kf, mis, runs_frs, runs_fre, nullptr, roff, runs, tot, true, d, lane); });
      for (int64_t d = 0; d < n; d++) {
        int64_t o = out_off[d];
        for (int64_t k = roff[d]; k < roff[d + 1]; k++) {
          if (runs[k].end <= runs[k].begin || (int64_t)runs[k].end > re0[d] - rs0[d]) return 9;
          if (k > roff[d] && runs[k].begin <= runs[k - 1].end) return 9;  // runs are maximal and in order
          const int64_t len = (int64_t)runs[k].end - runs[k].begin;
          if (o + len > out_off[d + 1] || memcmp(content_out + o, text + rs0[d] + runs[k].begin, (size_t)len) != 0) return 9;
          o += len;
        }
        if (o != out_off[d + 1]) return 9;
      }
    }
  }
  for (void* p : frees) free(p);
  return (has_c4 || decode_entities) ? 100 : 0;
}
