// tb_api.cu -- the C ABI (include/textblaster_b200.h): context, pipeline construction with the
// reference's validation rules, and the batch executor (chunked, double-buffered H2D on a copy
// stream, kernels on a compute stream, results to pinned host memory).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <emmintrin.h>
#endif
#include <nvtx3/nvToolsExt.h>  // header-only; ranges show up under Nsight Systems / Compute, and cost nothing without a tool attached

#include "tb_hostutil.h"
#include "tb_kernels.cuh"
#include "unicode_tables.inc"
#include "tb_entities.cuh"
#include "entity_table.inc"

static_assert(TB_WB_OTHER == WB_OTHER && TB_WB_CR == WB_CR && TB_WB_LF == WB_LF && TB_WB_NEWLINE == WB_NEWLINE &&
              TB_WB_EXTEND == WB_EXTEND && TB_WB_ZWJ == WB_ZWJ && TB_WB_REGIONAL_INDICATOR == WB_RI &&
              TB_WB_FORMAT == WB_FORMAT && TB_WB_KATAKANA == WB_KATAKANA && TB_WB_HEBREW_LETTER == WB_HEBREW &&
              TB_WB_ALETTER == WB_ALETTER && TB_WB_SINGLE_QUOTE == WB_SQ && TB_WB_DOUBLE_QUOTE == WB_DQ &&
              TB_WB_MIDNUMLET == WB_MIDNUMLET && TB_WB_MIDLETTER == WB_MIDLETTER && TB_WB_MIDNUM == WB_MIDNUM &&
              TB_WB_NUMERIC == WB_NUMERIC && TB_WB_EXTENDNUMLET == WB_EXTENDNUMLET && TB_WB_WSEGSPACE == WB_WSEG,
              "Word_Break numbering");
static_assert(TB_SB_OTHER == SB_OTHER && TB_SB_CR == SB_CR && TB_SB_LF == SB_LF && TB_SB_SEP == SB_SEP &&
              TB_SB_EXTEND == SB_EXTEND && TB_SB_FORMAT == SB_FORMAT && TB_SB_SP == SB_SP && TB_SB_LOWER == SB_LOWER &&
              TB_SB_UPPER == SB_UPPER && TB_SB_OLETTER == SB_OLETTER && TB_SB_NUMERIC == SB_NUMERIC &&
              TB_SB_ATERM == SB_ATERM && TB_SB_SCONTINUE == SB_SCONTINUE && TB_SB_STERM == SB_STERM &&
              TB_SB_CLOSE == SB_CLOSE, "Sentence_Break numbering");

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      return fail(e_ == cudaErrorMemoryAllocation ? TB_ERR_OUT_OF_MEMORY : TB_ERR_CUDA,            \
                  std::string(#call) + ": " + cudaGetErrorString(e_));                             \
    }                                                                                              \
  } while (0)

// format!("{}", f64): shortest round-trip digits, no exponent, no trailing ".0"
std::string rust_f64(double v) {
  if (std::isnan(v)) return "NaN";
  if (std::isinf(v)) return v > 0 ? "inf" : "-inf";
  char buf[64];
  for (int p = 1; p <= 17; p++) {
    snprintf(buf, sizeof buf, "%.*g", p, v);
    if (strtod(buf, nullptr) == v) break;
  }
  std::string s = buf;
  if (s.find('e') != std::string::npos || s.find('E') != std::string::npos) {
    // expand the exponent form
    char big[512];
    snprintf(big, sizeof big, "%.340f", v);
    std::string t = big;
    // trim to the shortest prefix that round-trips
    size_t dot = t.find('.');
    size_t keep = dot;
    for (size_t n = dot; n <= t.size(); n++) { if (strtod(t.substr(0, n).c_str(), nullptr) == v) { keep = n; break; } }
    s = t.substr(0, keep);
    if (!s.empty() && s.back() == '.') s.pop_back();
  }
  return s;
}

struct DevBuf {  // grow-only device allocation
  void* p = nullptr; size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return TB_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, bytes); want = bytes; }
    if (e != cudaSuccess) { p = nullptr; return fail(TB_ERR_OUT_OF_MEMORY, "cudaMalloc of " + std::to_string(bytes) + " bytes failed"); }
    cap = want;
    return TB_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};
struct HostBuf {  // grow-only pinned host allocation
  void* p = nullptr; size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return TB_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    if (cudaMallocHost(&p, want) != cudaSuccess) { cudaGetLastError(); p = nullptr; return fail(TB_ERR_OUT_OF_MEMORY, "cudaMallocHost failed"); }
    cap = want;
    return TB_OK;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};

size_t stat_size(int type) {
  return type == TB_STEP_C4_QUALITY ? sizeof(tb_c4_stats) : type == TB_STEP_GOPHER_QUALITY ? sizeof(tb_gq_stats)
         : type == TB_STEP_FINEWEB_QUALITY ? sizeof(tb_fw_stats) : sizeof(tb_gr_stats);
}

struct StepHost {
  int type;
  C4Cfg c4;
  bool c4_sentences = false;  // split_paragraph = false: lines are sentences (c4_filters.rs:152-155)
  GqCfg gq;
  GrCfg gr;
  FwCfg fw;
  SwTable sw{nullptr, 0, 0, 0, 0, 0};
  DevBuf sw_dev;
};

struct PlainBuf {  // grow-only pageable host allocation (not initialised)
  void* p = nullptr; size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return TB_OK;
    free(p);
    p = malloc(bytes + bytes / 8 + 256); cap = p ? bytes + bytes / 8 + 256 : 0;
    return p ? TB_OK : fail(TB_ERR_OUT_OF_MEMORY, "malloc of " + std::to_string(bytes) + " bytes failed");
  }
  void release() { free(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return (T*)p; }
};

struct ResultSet {  // host-side results of one run
  HostBuf keep, failed, content, content_off;
  PlainBuf kept, kept_off;  // TB_OPT_GATHER_KEPT
  HostBuf bad;              // TB_OPT_VALIDATE_UTF8: the device's count of malformed documents lands here (pinned: the copy does not block)
  HostBuf stats[TB_MAX_STEPS];
  bool in_use = false;
};

// Pageable input (an Arrow buffer as pyarrow or the arrow crate allocate it): cudaMemcpyAsync from it goes through the
// driver's own small staging buffer at a fraction of the link rate.  The library stages it itself: one thread per call
// copies the text piece by piece into a ring of pinned slots (several workers per piece) and sends every slot on with one
// cudaMemcpyAsync on the copy stream, so the link sees pinned memory and the copies overlap the kernels as before.
struct Stager {
  static constexpr int kSlots = 3;
  static constexpr size_t kSlotBytes = 64ull << 20;
  struct Job { int64_t ci; const uint8_t* src; size_t bytes; uint8_t* dst; const void* off_src; size_t off_bytes; void* off_dst; cudaEvent_t wait_before, record_after; };
  std::mutex mu; std::condition_variable cv; std::deque<Job> q; bool closed = false;
  int64_t recorded = -1;  // last chunk whose copies are all enqueued and whose event is recorded
  std::thread th;
  HostBuf* slots = nullptr; cudaEvent_t* slot_ev = nullptr; bool* slot_used = nullptr;
  cudaStream_t copy = nullptr; int T = 4, device = 0; cudaError_t err = cudaSuccess;
  static void par_copy(uint8_t* d, const uint8_t* s, size_t n, int T) {
    if (T <= 1 || n < (8u << 20)) { memcpy(d, s, n); return; }
    std::vector<std::thread> th;
    const size_t per = (n / T + 4095) & ~(size_t)4095;
    for (int t = 1; t < T; t++) { const size_t a = std::min(n, per * t), b = std::min(n, per * (t + 1)); if (b > a) th.emplace_back([=] { memcpy(d + a, s + a, b - a); }); }
    memcpy(d, s, std::min(n, per));
    for (auto& x : th) x.join();
  }
  void start() {
    th = std::thread([this] {
      cudaSetDevice(device);
      int next = 0;
      for (;;) {
        Job j;
        { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return closed || !q.empty(); }); if (q.empty()) return; j = q.front(); q.pop_front(); }
        cudaError_t e = cudaStreamWaitEvent(copy, j.wait_before, 0);  // the device buffer is free again
        for (size_t o = 0; o < j.bytes && e == cudaSuccess; o += kSlotBytes) {
          const size_t n = std::min(kSlotBytes, j.bytes - o);
          if (slot_used[next]) e = cudaEventSynchronize(slot_ev[next]);  // its last copy has left the slot
          if (e != cudaSuccess) break;
          par_copy(slots[next].as<uint8_t>(), j.src + o, n, T);
          e = cudaMemcpyAsync(j.dst + o, slots[next].p, n, cudaMemcpyHostToDevice, copy);
          if (e == cudaSuccess) e = cudaEventRecord(slot_ev[next], copy);
          slot_used[next] = true;
          next = (next + 1) % kSlots;
        }
        if (e == cudaSuccess) e = cudaMemcpyAsync(j.off_dst, j.off_src, j.off_bytes, cudaMemcpyHostToDevice, copy);
        if (e == cudaSuccess) e = cudaEventRecord(j.record_after, copy);
        { std::lock_guard<std::mutex> lk(mu); if (e != cudaSuccess) err = e; recorded = j.ci; }
        cv.notify_all();
      }
    });
  }
  void push(const Job& j) { { std::lock_guard<std::mutex> lk(mu); q.push_back(j); } cv.notify_all(); }
  cudaError_t wait_recorded(int64_t ci) { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return recorded >= ci || err != cudaSuccess; }); return err; }
  void finish() { if (!th.joinable()) return; { std::lock_guard<std::mutex> lk(mu); closed = true; } cv.notify_all(); th.join(); }
  ~Stager() { finish(); }
};

struct NvtxRange {  // one named range on the calling thread (a call, a chunk, a step of the chain)
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

struct KTime { const char* name; double ms; int64_t launches; };

}  // namespace

struct tb_ctx {
  int device = 0, sm_count = 148;
  UcTables uc{};
  DevBuf t_stage1, t_stage2, t_lower, t_cased, t_caseign, t_pow, t_ent;
};

struct tb_pipeline {
  tb_ctx* ctx = nullptr;
  // every pipeline has streams of its own: pipelines of one context run beside each other on the device, a call waits
  // for its own work only, and the per-kernel event times of one pipeline hold none of another's kernels
  cudaStream_t compute = nullptr, copy = nullptr, d2h = nullptr, side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  std::vector<StepHost> steps;
  int64_t opt_return_content = 1, opt_validate = 0, opt_chunk = 1024ll << 20, opt_debug = 0, opt_decode = 0, opt_gather = 0, opt_ramp = 1, opt_stage = 4, opt_bulk = 0;
  HostBuf stage_slots[Stager::kSlots]; cudaEvent_t stage_ev[Stager::kSlots] = {nullptr, nullptr, nullptr}; bool stage_used[Stager::kSlots] = {false, false, false};
  std::vector<cudaEvent_t> ev_gather;  // one per chunk of a call (TB_OPT_GATHER_KEPT)
  std::mutex mu;
  // device buffers (grow-only)
  DevBuf text[2], offs[2], packed[2];
  DevBuf rs0, re0, rs, re, ts, te, rewritten, len, bad, cursors, order, sizecls;
  // per-chunk results, double-buffered: the copies to the host run beside the next chunk's kernels
  DevBuf alive[2], failed[2], out_off[2], stats[TB_MAX_STEPS][2];
  // GopherRepetition scratch
  DevBuf g_wh, g_wl, g_ws, g_lh, g_ll, g_lg, g_doc, g_work, g_lists, g_counters;
  // FineWeb scratch
  DevBuf f_lh, f_ll, f_doc, f_tab;
  // C4 scratch
  DevBuf e_D, e_amp, e_drop, e_drs, e_dre;  // HTML entity pre-pass
  DevBuf c_S, c_bits, c_srs, c_sre;  // sentence-wise lines
  DevBuf c_ka, c_kf;                 // kept-byte bitmaps of passes A and C (tb_kernels.cuh C4Scratch)
  DevBuf c_F2;                       // second C4 output buffer: the C4 steps of a chain write F and F2 in turn (a step reads the one before it)
  DevBuf c_P, c_F, c_plen, c_nlines, c_docflags, c_lwc, c_lmax, c_lflags, c_prs, c_pre, c_frs, c_fre, c_cnt;
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr}, ev_d2h[2] = {nullptr, nullptr}, ev_small[2] = {nullptr, nullptr};
  std::vector<ResultSet*> results;
  // kernel timing
  std::vector<cudaEvent_t> ev_pool;
  struct Pending { int k; cudaEvent_t a, b; };
  std::vector<Pending> pending;
  std::vector<KTime> ktimes;
  size_t ev_next = 0;
  int launches_this_run = 0;
  int n_c4 = 0;  // C4QualityFilter steps in the chain
  double ms_this_run = 0;
};

namespace {

int upload(DevBuf& b, const void* src, size_t bytes) {
  int rc = b.ensure(bytes ? bytes : 1);
  if (rc) return rc;
  if (bytes) CK(cudaMemcpy(b.p, src, bytes, cudaMemcpyHostToDevice));
  return TB_OK;
}

int ktime_index(tb_pipeline* p, const char* name) {
  for (size_t i = 0; i < p->ktimes.size(); i++) if (strcmp(p->ktimes[i].name, name) == 0) return (int)i;
  p->ktimes.push_back(KTime{name, 0.0, 0});
  return (int)p->ktimes.size() - 1;
}
cudaEvent_t next_event(tb_pipeline* p) {
  if (p->ev_next == p->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); p->ev_pool.push_back(e); }
  return p->ev_pool[p->ev_next++];
}
struct Timed {  // brackets one kernel launch with events on the compute stream
  tb_pipeline* p; int k; cudaEvent_t a, b;
  Timed(tb_pipeline* p_, const char* name) : p(p_) {
    k = ktime_index(p, name); a = next_event(p); b = next_event(p);
    cudaEventRecord(a, p->compute);
  }
  ~Timed() { cudaEventRecord(b, p->compute); p->pending.push_back({k, a, b}); p->launches_this_run++; }
};
void resolve_times(tb_pipeline* p) {
  for (auto& q : p->pending) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, q.a, q.b) == cudaSuccess) { p->ktimes[q.k].ms += ms; p->ktimes[q.k].launches++; p->ms_this_run += ms; }
  }
  p->pending.clear();
  p->ev_next = 0;
}

// ---- validation: src/config/pipeline.rs:82-99, :118-158, :179-257 -----------------------------
int validate_step(const tb_step_cfg& s) {
  if (s.type == TB_STEP_C4_QUALITY) {
    const tb_c4_params& p = s.u.c4;
    if (p.min_num_sentences == 0) return fail(TB_ERR_CONFIG_VALIDATION, "C4QualityParams: min_num_sentences must be greater than 0");
    if (p.min_words_per_line == 0) return fail(TB_ERR_CONFIG_VALIDATION, "C4QualityParams: min_words_per_line must be greater than 0");
    if (p.max_word_length == 0) return fail(TB_ERR_CONFIG_VALIDATION, "C4QualityParams: max_word_length must be greater than 0");
    return TB_OK;
  }
  if (s.type == TB_STEP_GOPHER_REPETITION) {
    const tb_gr_params& p = s.u.gr;
    struct { const char* name; uint32_t bit; double v; } fr[4] = {{"dup_line_frac", TB_GR_HAS_DUP_LINE_FRAC, p.dup_line_frac},
                                                                  {"dup_para_frac", TB_GR_HAS_DUP_PARA_FRAC, p.dup_para_frac},
                                                                  {"dup_line_char_frac", TB_GR_HAS_DUP_LINE_CHAR_FRAC, p.dup_line_char_frac},
                                                                  {"dup_para_char_frac", TB_GR_HAS_DUP_PARA_CHAR_FRAC, p.dup_para_char_frac}};
    for (auto& f : fr)
      if ((p.has & f.bit) && !(f.v >= 0.0 && f.v <= 1.0))
        return fail(TB_ERR_CONFIG_VALIDATION, std::string("GopherRepetitionParams: ") + f.name + " must be between 0.0 and 1.0, got " + rust_f64(f.v));
    if (p.n_top < 0 || p.n_top > TB_MAX_NGRAMS || p.n_dup < 0 || p.n_dup > TB_MAX_NGRAMS)
      return fail(TB_ERR_INVALID_ARGUMENT, "GopherRepetitionFilter: at most " + std::to_string(TB_MAX_NGRAMS) + " entries per n-gram list");
    for (int which = 0; which < 2; which++) {
      const char* name = which ? "dup_n_grams" : "top_n_grams";
      int n = which ? p.n_dup : p.n_top;
      for (int i = 0; i < n; i++) {
        uint32_t size = which ? p.dup_n[i] : p.top_n[i];
        double fraction = which ? p.dup_frac[i] : p.top_frac[i];
        if (size == 0)
          return fail(TB_ERR_CONFIG_VALIDATION, std::string("GopherRepetitionParams: n-gram size in ") + name + " at index " + std::to_string(i) + " must be greater than 0");
        if (!(fraction >= 0.0 && fraction <= 1.0))
          return fail(TB_ERR_CONFIG_VALIDATION, std::string("GopherRepetitionParams: n-gram fraction in ") + name + " at index " + std::to_string(i) + " must be between 0.0 and 1.0, got " + rust_f64(fraction));
        if (size > TB_MAX_NGRAM_N) return fail(TB_ERR_UNSUPPORTED, "GopherRepetitionFilter: n-gram order above " + std::to_string(TB_MAX_NGRAM_N));
      }
    }
    return TB_OK;
  }
  if (s.type == TB_STEP_GOPHER_QUALITY) {
    const tb_gq_params& p = s.u.gq;
    bool hmin = p.has & TB_GQ_HAS_MIN_DOC_WORDS, hmax = p.has & TB_GQ_HAS_MAX_DOC_WORDS;
    if (hmin && p.min_doc_words == 0) return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: min_doc_words must be greater than 0");
    if (hmax && p.max_doc_words == 0) return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: max_doc_words must be greater than 0");
    if (hmin && hmax && p.min_doc_words > p.max_doc_words)
      return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: min_doc_words (" + std::to_string(p.min_doc_words) + ") cannot be greater than max_doc_words (" + std::to_string(p.max_doc_words) + ")");
    bool amin = p.has & TB_GQ_HAS_MIN_AVG_WORD_LENGTH, amax = p.has & TB_GQ_HAS_MAX_AVG_WORD_LENGTH;
    if (amin && !(p.min_avg_word_length > 0.0)) return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: min_avg_word_length must be greater than 0.0");
    if (amax && !(p.max_avg_word_length > 0.0)) return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: max_avg_word_length must be greater than 0.0");
    if (amin && amax && p.min_avg_word_length > p.max_avg_word_length)
      return fail(TB_ERR_CONFIG_VALIDATION, "GopherQualityParams: min_avg_word_length (" + rust_f64(p.min_avg_word_length) + ") cannot be greater than max_avg_word_length (" + rust_f64(p.max_avg_word_length) + ")");
    struct { const char* name; uint32_t bit; double v; } rr[4] = {{"max_symbol_word_ratio", TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO, p.max_symbol_word_ratio},
                                                                  {"max_bullet_lines_ratio", TB_GQ_HAS_MAX_BULLET_LINES_RATIO, p.max_bullet_lines_ratio},
                                                                  {"max_ellipsis_lines_ratio", TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO, p.max_ellipsis_lines_ratio},
                                                                  {"max_non_alpha_words_ratio", TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO, p.max_non_alpha_words_ratio}};
    for (auto& r : rr)
      if ((p.has & r.bit) && r.v < 0.0)
        return fail(TB_ERR_CONFIG_VALIDATION, std::string("GopherQualityParams: ") + r.name + " must be non-negative, got " + rust_f64(r.v));
    if ((p.has & TB_GQ_HAS_STOP_WORDS) && (p.n_stop_words < 0 || (p.n_stop_words > 0 && (!p.stop_words_utf8 || !p.stop_words_offsets))))
      return fail(TB_ERR_INVALID_ARGUMENT, "GopherQualityFilter: stop word list pointers missing");
    return TB_OK;
  }
  if (s.type == TB_STEP_FINEWEB_QUALITY) {  // src/config/pipeline.rs:324-348
    const tb_fw_params& p = s.u.fw;
    struct { const char* name; double v; } rr[4] = {{"line_punct_thr", p.line_punct_thr}, {"short_line_thr", p.short_line_thr},
                                                    {"char_duplicates_ratio", p.char_duplicates_ratio}, {"new_line_ratio", p.new_line_ratio}};
    for (auto& r : rr)
      if (!(r.v >= 0.0 && r.v <= 1.0))
        return fail(TB_ERR_CONFIG_VALIDATION, std::string("FineWebQualityFilterParams: ") + r.name + " must be between 0.0 and 1.0, got " + rust_f64(r.v));
    if (p.short_line_length == 0) return fail(TB_ERR_CONFIG_VALIDATION, "FineWebQualityFilterParams: short_line_length must be greater than 0");
    if (p.n_stop_chars > TB_FW_MAX_STOP_CHARS) return fail(TB_ERR_INVALID_ARGUMENT, "FineWebQualityFilter: at most " + std::to_string(TB_FW_MAX_STOP_CHARS) + " stop_chars");
    return TB_OK;
  }
  return fail(TB_ERR_INVALID_ARGUMENT, "unknown step type " + std::to_string(s.type));
}

int build_stop_table(StepHost& sh, const tb_gq_params& p) {
  uint32_t mask = 0, maxlen = 0, lenmask = 0, firstmask = 0, secondmask = 0;
  std::vector<SwEntry> tab = tb_build_stop_table(tb_stop_words_of(p), &mask, &maxlen, &lenmask, &firstmask, &secondmask);
  int rc = upload(sh.sw_dev, tab.data(), tab.size() * sizeof(SwEntry));
  if (rc) return rc;
  sh.sw = SwTable{sh.sw_dev.as<SwEntry>(), mask, maxlen, lenmask, firstmask, secondmask};
  return TB_OK;
}

struct Chunk { int64_t d0, d1, b0, b1; };

inline int64_t off_at(const void* offsets, int is64, int64_t i) {
  return is64 ? ((const int64_t*)offsets)[i] : (int64_t)((const int32_t*)offsets)[i];
}

// Sequential writer for the gathered content.  The runs are short (hundreds of bytes) and the destination is written once
// and not read again soon: plain stores would first fetch every destination line (read for ownership), a third of the
// memory traffic of the gather.  So the bytes are collected in a small cache-resident block and leave it as whole
// 64-byte lines with streaming stores; only the partial lines at the two ends of a thread's range are copied plainly.
struct StreamWriter {
  uint8_t* dst;          // destination base
  int64_t pos;           // next offset to write
  int fill = 0, cap = 0; // bytes in buf / bytes buf may take before the next flush
  alignas(64) uint8_t buf[4096];
  StreamWriter(uint8_t* d, int64_t o) : dst(d), pos(o) {
    const int head = (int)((64 - ((uintptr_t)(d + o) & 63)) & 63);  // up to the next line boundary: plain
    cap = head ? head : (int)sizeof buf;
  }
  void flush() {
    uint8_t* q = dst + pos;
#if defined(__x86_64__)
    if (((uintptr_t)q & 63) == 0) {
      const int lines = fill & ~63;
      for (int k = 0; k < lines; k += 16) _mm_stream_si128((__m128i*)(q + k), _mm_load_si128((const __m128i*)(buf + k)));
      if (fill > lines) memcpy(q + lines, buf + lines, (size_t)(fill - lines));
    } else
#endif
      memcpy(q, buf, (size_t)fill);
    pos += fill; fill = 0; cap = (int)sizeof buf;
  }
  void write(const uint8_t* src, size_t len) {
    while (len) {
      const size_t k = std::min(len, (size_t)(cap - fill));
      memcpy(buf + fill, src, k);
      fill += (int)k; src += k; len -= k;
      if (fill == cap) flush();
    }
  }
  void finish() {
    if (fill) flush();
#if defined(__x86_64__)
    _mm_sfence();
#endif
  }
};

// The content of the documents d in [d0, d1) with select[d] != 0 (NULL: all), rebuilt from their runs and the input
// column: bytes to dst + byte_base (dst may be NULL: sizes only), offsets to dst_off[k_base + 1 ...].  The documents
// are dealt to T threads in contiguous shares; pass 1 sizes every share, pass 2 writes.  `last_end`: run_offsets[d1]
// (passed by value: the entry is shared with the next chunk's results, which may be arriving).  Returns through nk / nb
// the number of documents and bytes written.
struct GatherSrc { const tb_run* runs; const int64_t* ro; const uint8_t* utf8; const void* offsets; int is64; const uint8_t* select; };
void gather_range(const GatherSrc& g, int64_t d0, int64_t d1, int64_t last_end, uint8_t* dst, int64_t* dst_off, int64_t k_base,
                  int64_t byte_base, int T, int64_t* nk, int64_t* nb) {
  const int64_t n = d1 - d0;
  T = (int)std::max<int64_t>(1, std::min<int64_t>(T, n / 1024 + 1));
  std::vector<int64_t> share_docs(T + 1, 0), share_bytes(T + 1, 0);
  auto ro_end = [&](int64_t d) { return d + 1 == d1 ? last_end : g.ro[d + 1]; };
  auto each = [&](auto&& fn) {
    if (T == 1) { fn(0); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(fn, t);
    fn(0);
    for (auto& x : th) x.join();
  };
  each([&](int t) {
    int64_t cd = 0, cb = 0;
    for (int64_t d = d0 + n * t / T, e = d0 + n * (t + 1) / T; d < e; d++) {
      if (g.select && !g.select[d]) continue;
      cd++;
      for (int64_t k = g.ro[d], ke = ro_end(d); k < ke; k++) cb += (int64_t)g.runs[k].end - (int64_t)g.runs[k].begin;
    }
    share_docs[t + 1] = cd; share_bytes[t + 1] = cb;
  });
  for (int t = 0; t < T; t++) { share_docs[t + 1] += share_docs[t]; share_bytes[t + 1] += share_bytes[t]; }
  *nk = share_docs[T]; *nb = share_bytes[T];
  each([&](int t) {
    int64_t k = k_base + share_docs[t], o = byte_base + share_bytes[t];
    StreamWriter w(dst, o);
    // Every run starts somewhere else in the input, so a copy begins with a miss.  A second cursor runs a few runs
    // ahead of the copy (across documents) and asks for the first lines of each run it passes.
    const int64_t d_end = d0 + n * (t + 1) / T;
    int64_t pd = d0 + n * t / T, pr = -1, ahead = 0;
    auto prefetch_next = [&]() {
      for (;;) {
        if (pd >= d_end) return;
        if (pr < 0) { if (g.select && !g.select[pd]) { pd++; continue; } pr = g.ro[pd]; }
        if (pr >= ro_end(pd)) { pd++; pr = -1; continue; }
        const uint8_t* nx = g.utf8 + off_at(g.offsets, g.is64, pd) + g.runs[pr].begin;
        __builtin_prefetch(nx); __builtin_prefetch(nx + 64); __builtin_prefetch(nx + 128); __builtin_prefetch(nx + 192);
        pr++; ahead++;
        return;
      }
    };
    static const bool plain = [] { const char* e = getenv("TB_GATHER_COPY"); return e && !strcmp(e, "plain"); }();  // (measurement knob: memcpy straight into the destination)
    if (dst) for (int i = 0; i < 6; i++) prefetch_next();
    for (int64_t d = d0 + n * t / T; d < d_end; d++) {
      if (g.select && !g.select[d]) continue;
      const uint8_t* src = g.utf8 + off_at(g.offsets, g.is64, d);
      for (int64_t r = g.ro[d], re = ro_end(d); r < re; r++) {
        const size_t len = (size_t)g.runs[r].end - g.runs[r].begin;
        if (dst) {
          prefetch_next();
          if (plain) memcpy(dst + o, src + g.runs[r].begin, len);
          else w.write(src + g.runs[r].begin, len);
        }
        o += (int64_t)len;
      }
      dst_off[++k] = o;
    }
    if (dst) w.finish();
  });
}

// TB_OPT_GATHER_KEPT: one thread per call that takes the chunks in order as their runs reach the host (an event per
// chunk on the d2h stream) and gathers the kept documents' content with T workers, behind the device.
struct Gatherer {
  struct Item { int64_t d0, d1, last_end; cudaEvent_t ev; };
  std::mutex mu; std::condition_variable cv; std::deque<Item> q; bool closed = false;
  std::thread th;
  GatherSrc src{}; uint8_t* dst = nullptr; int64_t* dst_off = nullptr; int T = 1, device = 0;
  int64_t nk = 0, nb = 0; cudaError_t err = cudaSuccess; double busy_ms = 0;
  void start() {
    th = std::thread([this] {
      cudaSetDevice(device);
      for (;;) {
        Item it;
        { std::unique_lock<std::mutex> lk(mu); cv.wait(lk, [&] { return closed || !q.empty(); }); if (q.empty()) return; it = q.front(); q.pop_front(); }
        cudaError_t e = cudaEventSynchronize(it.ev);
        if (e != cudaSuccess) { err = e; continue; }
        int64_t k = 0, b = 0;
        const auto t0 = std::chrono::steady_clock::now();
        gather_range(src, it.d0, it.d1, it.last_end, dst, dst_off, nk, nb, T, &k, &b);
        busy_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        nk += k; nb += b;
      }
    });
  }
  void push(const Item& it) { { std::lock_guard<std::mutex> lk(mu); q.push_back(it); } cv.notify_one(); }
  void finish() { if (!th.joinable()) return; { std::lock_guard<std::mutex> lk(mu); closed = true; } cv.notify_one(); th.join(); }
  ~Gatherer() { finish(); }
};

ResultSet* acquire_results(tb_pipeline* p) {
  for (auto* r : p->results) if (!r->in_use) { r->in_use = true; return r; }
  auto* r = new ResultSet();
  r->in_use = true;
  p->results.push_back(r);
  return r;
}

// Device buffers for a chunk of n docs / nb bytes.
int ensure_chunk_buffers(tb_pipeline* p, int64_t n, int64_t nb, bool host_input) {
  int rc;
  size_t n1 = (size_t)n + 1;
#define ENS(buf, bytes) if ((rc = (buf).ensure(bytes))) return rc
  if (host_input) { for (int i = 0; i < 2; i++) { ENS(p->text[i], (size_t)nb + 64); ENS(p->offs[i], n1 * 8); } }
  // with sentence-wise lines the rewritten text can outgrow the input (tb_sent_off)
  bool sent = false;
  for (auto& sh : p->steps) sent |= sh.type == TB_STEP_C4_QUALITY && sh.c4_sentences;
  size_t cb = (size_t)nb;  // extent of the buffers the chain works in
  if (p->opt_decode) {
    cb = tb_grown_bytes((int64_t)cb, n);
    ENS(p->e_D, cb + 64); ENS(p->e_amp, ((size_t)nb / 32 + 4) * 4); ENS(p->e_drop, ((size_t)nb / 32 + 4) * 4); ENS(p->e_drs, n1 * 8); ENS(p->e_dre, n1 * 8);
  }
  const size_t cb_in = cb;  // what the sentence-wise pre-pass of C4 reads
  if (sent) cb = tb_grown_bytes((int64_t)cb, n);
  for (int i = 0; i < 2; i++) ENS(p->packed[i], cb + 8 * n1 + 64);  // the content, or its runs: at most one per 8 bytes of text + one per document
  ENS(p->rs0, n1 * 8); ENS(p->re0, n1 * 8); ENS(p->rs, n1 * 8); ENS(p->re, n1 * 8); ENS(p->ts, n1 * 8); ENS(p->te, n1 * 8);
  for (int i = 0; i < 2; i++) { ENS(p->alive[i], n1); ENS(p->failed[i], n1); ENS(p->out_off[i], (n1 + 1) * 8); }
  ENS(p->rewritten, n1); ENS(p->len, n1 * 8); ENS(p->bad, 16); ENS(p->cursors, 256); ENS(p->order, n1 * 4); ENS(p->sizecls, n1);
  // word / line slots are indexed by the position of the document in its buffer (tb_wunit / tb_lunit)
  const size_t wunits = tb_word_units(cb, (size_t)n), units = tb_line_units(cb, (size_t)n);
  for (size_t s = 0; s < p->steps.size(); s++) {
    const StepHost& sh = p->steps[s];
    for (int i = 0; i < 2; i++) ENS(p->stats[s][i], n1 * stat_size(sh.type));
    if (sh.type == TB_STEP_GOPHER_REPETITION) {
      ENS(p->g_wh, wunits * 8); ENS(p->g_wl, wunits * 4); ENS(p->g_ws, wunits * 8); ENS(p->g_lh, units * 8); ENS(p->g_ll, units * 4); ENS(p->g_lg, units);
      ENS(p->g_doc, n1 * sizeof(GrDoc));
      ENS(p->g_work, (wunits * TB_REP_K + n1 * TB_REP_DOC_EXTRA) * 4);
      ENS(p->g_lists, TB_REP_NT * n1 * 4); ENS(p->g_counters, TB_GR_COUNTERS * 4);
    } else if (sh.type == TB_STEP_FINEWEB_QUALITY) {
      ENS(p->f_lh, units * 8); ENS(p->f_ll, units * 4); ENS(p->f_doc, n1 * sizeof(FwDoc)); ENS(p->f_tab, (units * 2 + n1 * 16) * 4);
    } else if (sh.type == TB_STEP_C4_QUALITY) {
      size_t lines = cb + n1 + 1;
      ENS(p->c_P, cb + 64); ENS(p->c_F, cb + 64);
      if (p->n_c4 > 1) ENS(p->c_F2, cb + 64);
      ENS(p->c_ka, (cb / 32 + 4) * 4); ENS(p->c_kf, (cb / 32 + 4) * 4);
      if (sh.c4_sentences) { ENS(p->c_S, cb + 64); ENS(p->c_bits, (cb_in / 32 + 4) * 4); ENS(p->c_srs, n1 * 8); ENS(p->c_sre, n1 * 8); }
      ENS(p->c_plen, n1 * 8); ENS(p->c_nlines, n1 * 4); ENS(p->c_docflags, n1 * 4);
      ENS(p->c_lwc, lines * 4); ENS(p->c_lmax, lines * 4); ENS(p->c_lflags, lines * 4);
      ENS(p->c_prs, n1 * 8); ENS(p->c_pre, n1 * 8); ENS(p->c_frs, n1 * 8); ENS(p->c_fre, n1 * 8); ENS(p->c_cnt, n1 * 16);
    }
  }
#undef ENS
  return TB_OK;
}

// Enqueue every kernel of the chain for one chunk whose text is at `text` (device) with ranges in
// rs0/re0 (already initialised).  Leaves packed content in p->packed[pb] and offsets in out_off.
// content_mode: 0 nothing is returned, 1 the packed content, 2 runs of the input text (one C4 step, lines = str::lines)
int enqueue_chain(tb_pipeline* p, const uint8_t* text, int64_t n, int64_t nbytes, int pb, int content_mode, bool* has_c4) {
  const bool want_content = content_mode != 0;
  tb_ctx* c = p->ctx;
  cudaStream_t st = p->compute;
  CK(cudaMemcpyAsync(p->rs.p, p->rs0.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(p->re.p, p->re0.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemsetAsync(p->rewritten.p, 0, (size_t)n, st));
  CK(cudaMemsetAsync(p->cursors.p, 0, 256, st));
  uint32_t* cursors = p->cursors.as<uint32_t>();  // [0, 4 * TB_MAX_STEPS): kernel cursors; [40, 56): counts of the work order; [56..): pre-pass cursors
  KBatch b{text, p->rs.as<int64_t>(), p->re.as<int64_t>(), p->alive[pb].as<uint8_t>(), p->failed[pb].as<int8_t>(), n, c->uc, 0, nbytes,
           (uint32_t)p->opt_debug, p->order.as<uint32_t>(), cursors + 40, p->sizecls.as<uint8_t>(), (uint32_t)p->opt_bulk};
  { Timed t(p, "k_order"); tb_launch_order(st, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), n, p->order.as<uint32_t>(), cursors + 40, p->sizecls.as<uint8_t>()); }
  *has_c4 = false;
  int c4_seen = 0;
  if (p->opt_validate) {  // p->bad counts over the whole run (cleared by run_impl)
    { Timed t(p, "k_validate_utf8"); tb_launch_validate_utf8(st, text, b.rs, b.re, n, p->bad.as<int32_t>()); }
  }
  for (size_t s = 0; s < p->steps.size(); s++) {
    StepHost& sh = p->steps[s];
    NvtxRange nvtx_step(tb_step_name(sh.type));
    if (sh.type == TB_STEP_GOPHER_QUALITY) {
      Timed t(p, "k_gq");
      tb_launch_gq(st, b, sh.sw, sh.gq, (int)s, p->stats[s][pb].as<tb_gq_stats>(), c->sm_count, cursors + s);
    } else if (sh.type == TB_STEP_GOPHER_REPETITION) {
      GrScratch g{p->g_wh.as<uint64_t>(), p->g_wl.as<uint32_t>(), p->g_ws.as<uint64_t>(), p->g_lh.as<uint64_t>(), p->g_ll.as<uint32_t>(),
                  p->g_lg.as<uint8_t>(), p->g_doc.as<GrDoc>(), p->g_work.as<uint32_t>(), p->g_lists.as<uint32_t>(),
                  p->g_counters.as<uint32_t>(), n};
      CK(cudaMemsetAsync(p->g_counters.p, 0, TB_GR_COUNTERS * 4, st));
      CK(cudaMemsetAsync(p->stats[s][pb].p, 0, (size_t)n * sizeof(tb_gr_stats), st));  // pass 2 writes the non-zero fields only
      { Timed t(p, "k_trim"); tb_launch_trim(st, b, p->ts.as<int64_t>(), p->te.as<int64_t>()); }
      { Timed t(p, "k_gr_scan"); tb_launch_gr_scan(st, b, p->ts.as<int64_t>(), p->te.as<int64_t>(), g, c->sm_count); }
      // the few documents of tier X (each a long chain of passes on one CTA) run beside the other tiers
      CK(cudaEventRecord(p->ev_fork, st));
      CK(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
      tb_launch_gr_rep(p->side, b, p->ts.as<int64_t>(), p->te.as<int64_t>(), g, sh.gr, (int)s, p->stats[s][pb].as<tb_gr_stats>(), c->sm_count, TB_REP_TIER_X);
      CK(cudaEventRecord(p->ev_join, p->side));
      p->launches_this_run++;
      static const char* const kTierName[TB_REP_TIER_X] = {"k_gr_rep_w256", "k_gr_rep_w512", "k_gr_rep_c1k", "k_gr_rep_c2k", "k_gr_rep_c4k", "k_gr_rep_c8k"};
      for (int tier = TB_REP_TIER_X - 1; tier >= 0; tier--) {  // the few large documents first
        Timed t(p, kTierName[tier]);
        tb_launch_gr_rep(st, b, p->ts.as<int64_t>(), p->te.as<int64_t>(), g, sh.gr, (int)s, p->stats[s][pb].as<tb_gr_stats>(), c->sm_count, tier);
      }
      { Timed t(p, "k_gr_rep_x(join)"); CK(cudaStreamWaitEvent(st, p->ev_join, 0)); }
    } else if (sh.type == TB_STEP_FINEWEB_QUALITY) {
      FwScratch fs{p->f_lh.as<uint64_t>(), p->f_ll.as<uint32_t>(), p->f_doc.as<FwDoc>(), p->f_tab.as<uint32_t>()};
      { Timed t(p, "k_fw_scan"); tb_launch_fw_scan(st, b, sh.fw, fs, c->sm_count, cursors + s); }
      { Timed t(p, "k_fw_fin"); tb_launch_fw_fin(st, b, sh.fw, fs, (int)s, p->stats[s][pb].as<tb_fw_stats>(), c->sm_count, cursors + TB_MAX_STEPS + s); }
    } else {
      uint8_t* const Fout = (c4_seen++ & 1) ? p->c_F2.as<uint8_t>() : p->c_F.as<uint8_t>();  // (never the buffer this step reads)
      C4Scratch cs{p->c_P.as<uint8_t>(), Fout, p->c_plen.as<int64_t>(), p->c_nlines.as<uint32_t>(),
                   p->c_docflags.as<uint32_t>(), p->c_lwc.as<uint32_t>(), p->c_lmax.as<uint32_t>(), p->c_lflags.as<uint32_t>(),
                   p->c_prs.as<int64_t>(), p->c_pre.as<int64_t>(), p->c_frs.as<int64_t>(), p->c_fre.as<int64_t>(), p->c_cnt.as<uint32_t>(),
                   p->c_ka.as<uint32_t>(), content_mode == 2 ? p->c_kf.as<uint32_t>() : nullptr,
                   (uint8_t)(Fout == p->c_F.as<uint8_t>() ? 1 : 2)};
      {  // the bitmaps are OR-ed into where two documents share a word
        const size_t bw = ((size_t)(sh.c4_sentences ? tb_grown_bytes(nbytes, n) : (size_t)nbytes) / 32 + 4) * 4;
        CK(cudaMemsetAsync(p->c_ka.p, 0, bw, st));
        if (cs.kf) CK(cudaMemsetAsync(p->c_kf.p, 0, bw, st));
      }
      const KBatch b_in = b;  // rs / re of the chain: pass D writes the ranges of the rewritten documents there
      if (sh.c4_sentences) {  // one sentence per line into S; the line passes then run on S
        CK(cudaMemsetAsync(p->c_bits.p, 0, ((size_t)nbytes / 32 + 4) * 4, st));
        { Timed t(p, "k_trim"); tb_launch_trim(st, b, p->ts.as<int64_t>(), p->te.as<int64_t>()); }
        { Timed t(p, "k_c4_sent"); tb_launch_c4_sent(st, b, p->ts.as<int64_t>(), p->te.as<int64_t>(), p->c_bits.as<uint32_t>(), p->c_S.as<uint8_t>(),
                                                     p->c_srs.as<int64_t>(), p->c_sre.as<int64_t>(), c->sm_count, cursors + 58); }
        b.text = p->c_S.as<uint8_t>(); b.rs = p->c_srs.as<int64_t>(); b.re = p->c_sre.as<int64_t>();
        b.ext_hi = (int64_t)tb_grown_bytes(nbytes, n);
      }
      { Timed t(p, "k_c4_a"); tb_launch_c4_a(st, b, sh.c4, cs, c->sm_count, cursors + TB_MAX_STEPS + s); }
      { Timed t(p, "k_c4_b"); tb_launch_c4_b(st, b, sh.c4, cs, c->sm_count, cursors + s); }
      { Timed t(p, "k_c4_c"); tb_launch_c4_c(st, b, sh.c4, cs, c->sm_count, cursors + 2 * TB_MAX_STEPS + s); }
      {  // content = kept.join("\n").trim()   (c4_filters.rs:258)
        KBatch fb = b; fb.text = cs.F; fb.rs = cs.frs; fb.re = cs.fre;
        Timed t(p, "k_trim"); tb_launch_trim(st, fb, cs.frs, cs.fre);
      }
      { Timed t(p, "k_c4_d"); tb_launch_c4_d(st, b, sh.c4, cs, (int)s, p->stats[s][pb].as<tb_c4_stats>(), p->rewritten.as<uint8_t>(), (int64_t*)b_in.rs, (int64_t*)b_in.re, c->sm_count, cursors + 3 * TB_MAX_STEPS + s); }
      b.rs = b_in.rs; b.re = b_in.re;
      b.text = cs.F;  // every document still alive now lives in F
      *has_c4 = true;
    }
  }
  if (content_mode == 2) {  // runs of the caller's text: count, offsets, write (the text of the chain's first step is the caller's)
    const uint32_t mis = (uint32_t)((uintptr_t)text & 31u);
    const int64_t cap = nbytes / 8 + n;  // runs packed[pb] holds
    { Timed t(p, "k_c4_runs"); tb_launch_c4_runs(st, n, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), p->rewritten.as<uint8_t>(), p->c_kf.as<uint32_t>(), mis,
                                                 p->c_frs.as<int64_t>(), p->c_fre.as<int64_t>(), p->len.as<int64_t>(), nullptr, nullptr, cap, 0); }
    { Timed t(p, "k_scan_i64"); tb_launch_scan_i64(st, p->len.as<int64_t>(), p->out_off[pb].as<int64_t>(), n); }
    { Timed t(p, "k_c4_runs"); tb_launch_c4_runs(st, n, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), p->rewritten.as<uint8_t>(), p->c_kf.as<uint32_t>(), mis,
                                                 p->c_frs.as<int64_t>(), p->c_fre.as<int64_t>(), nullptr, p->out_off[pb].as<int64_t>(), p->packed[pb].p, cap, 1); }
  } else if (want_content) {  // a C4 step rewrote documents, or the entity pre-pass made `text` a copy of its own
    { Timed t(p, "k_final_len"); tb_launch_final_len(st, n, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), b.rs, b.re, p->rewritten.as<uint8_t>(), p->len.as<int64_t>()); }
    { Timed t(p, "k_scan_i64"); tb_launch_scan_i64(st, p->len.as<int64_t>(), p->out_off[pb].as<int64_t>(), n); }
    { Timed t(p, "k_pack"); tb_launch_pack(st, n, text, p->rs0.as<int64_t>(), *has_c4 ? p->c_F.as<uint8_t>() : b.text, p->c_F2.as<uint8_t>(), b.rs, p->rewritten.as<uint8_t>(), p->out_off[pb].as<int64_t>(), p->packed[pb].as<uint8_t>()); }
  }
  CK(cudaGetLastError());
  return TB_OK;
}

int run_impl(tb_pipeline* p, const uint8_t* utf8, const void* offsets, int is64, int64_t n_docs, bool device_input,
             const int64_t* offsets_dev, tb_batch_out* out) {
  tb_ctx* c = p->ctx;
  CK(cudaSetDevice(c->device));
  NvtxRange nvtx_call(device_input ? "tb_pipeline_run_batch_device" : "tb_pipeline_run_batch");
  static const bool trace = getenv("TB_TRACE") != nullptr;  // host-side timeline of one call on stderr
  const auto t_entry = std::chrono::steady_clock::now();
  auto since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_entry).count(); };
  memset(out, 0, sizeof *out);
  out->n_docs = n_docs;
  out->n_steps = (int32_t)p->steps.size();
  p->launches_this_run = 0; p->ms_this_run = 0;
  bool any_c4 = false;
  for (auto& s : p->steps) any_c4 |= s.type == TB_STEP_C4_QUALITY;
  bool sent_mode = false;
  for (auto& s : p->steps) sent_mode |= s.type == TB_STEP_C4_QUALITY && s.c4_sentences;
  // what comes back as the content: nothing, the bytes, or runs of the caller's own text.  Runs need the text C4 reads to
  // be the caller's (no entity pre-pass, lines = str::lines); otherwise the bytes are returned.
  int content_mode = (any_c4 || p->opt_decode) ? (int)p->opt_return_content : 0;
  if (content_mode == 2 && (p->opt_decode || sent_mode || p->n_c4 != 1)) content_mode = 1;
  const bool want_content = content_mode != 0;
  const size_t esz = content_mode == 2 ? sizeof(tb_run) : 1;  // bytes per element of the content buffer

  size_t content_cap = (size_t)0;
  ResultSet* rsx = acquire_results(p);
  out->_owner = rsx;
  int rc;
  size_t n1 = (size_t)n_docs + 1;
  if ((rc = rsx->keep.ensure(n1)) || (rc = rsx->failed.ensure(n1))) return rc;
  for (size_t s = 0; s < p->steps.size(); s++) if ((rc = rsx->stats[s].ensure(n1 * stat_size(p->steps[s].type)))) return rc;
  const int64_t base0 = n_docs ? off_at(offsets, is64, 0) : 0;
  const int64_t total_bytes = n_docs ? off_at(offsets, is64, n_docs) - base0 : 0;
  content_cap = (size_t)total_bytes;
  if (p->opt_decode) content_cap = tb_grown_bytes((int64_t)content_cap, n_docs);
  if (sent_mode) content_cap = tb_grown_bytes((int64_t)content_cap, n_docs);
  if (want_content) { if ((rc = rsx->content.ensure(content_cap + 8 * n1 + 64)) || (rc = rsx->content_off.ensure(n1 * 8))) return rc; }
  {  // (one branch-free sweep per offset width: a million offsets cost a fraction of a millisecond)
    int bad = 0;
    if (is64) { const int64_t* o = (const int64_t*)offsets; for (int64_t i = 0; i < n_docs; i++) bad |= o[i + 1] < o[i]; }
    else { const int32_t* o = (const int32_t*)offsets; for (int64_t i = 0; i < n_docs; i++) bad |= o[i + 1] < o[i]; }
    if (bad) return fail(TB_ERR_INVALID_ARGUMENT, "offsets must be non-decreasing");
  }
  const int gather_T = (content_mode == 2 && !device_input) ? (int)std::min<int64_t>(p->opt_gather, 64) : 0;
  Gatherer gth;  // (joins its thread when this frame is left, whichever way)
  if (gather_T > 0) { if ((rc = rsx->kept.ensure((size_t)total_bytes + 64)) || (rc = rsx->kept_off.ensure((n1 + 1) * 8))) return rc; }

  // chunk plan.  Large chunks cost the least device time (every launch of the chain ends with a tail: about
  // 90 us per launch, measured), but the copy of the first chunk to the device and the copy of the last
  // chunk's results back are not hidden behind kernels.  So the sizes ramp: they start small
  // and double (the copy engine moves text ~2.5x faster than the kernels consume it, so the next chunk is
  // always there in time), they stay at the configured size, and they shrink by thirds at the end (the
  // results of a chunk leave while the next, smaller one is computed).
  std::vector<Chunk> chunks;
  int64_t max_n = 0, max_b = 0;
  {
    const int64_t C = std::max<int64_t>(p->opt_chunk, 1);
    // (a caller that streams batch after batch through several pipelines has the copies of one batch hidden behind the
    // kernels of another: no ramps then, TB_OPT_CHUNK_RAMP = 0)
    // (a chunk should hold some thousand documents whatever their length: a persistent kernel over a few hundred long
    // documents leaves most of the GPU idle while the longest of them is worked off)
    const int64_t floor_bytes = n_docs ? (int64_t)((double)total_bytes / (double)n_docs * 4096.0) : 0;
    const std::vector<int64_t> targets = p->opt_ramp ? tb_chunk_targets(total_bytes, C, device_input, floor_bytes) : std::vector<int64_t>{C};
    size_t ti = 0;
    for (int64_t d = 0; d < n_docs;) {
      const int64_t b0 = off_at(offsets, is64, d);
      const int64_t target = targets[std::min(ti, targets.size() - 1)];
      ti++;
      // the last document that still ends within the target (the offsets are sorted: bisection), at least one document
      int64_t lo = d + 1, hi = n_docs;
      while (lo < hi) { const int64_t m = lo + (hi - lo + 1) / 2; if (off_at(offsets, is64, m) - b0 <= target) lo = m; else hi = m - 1; }
      int64_t e = lo;
      if (ti >= targets.size()) {  // rounding left a few documents over: they join the last chunk if that keeps it <= C
        if (off_at(offsets, is64, n_docs) - b0 <= C) e = n_docs;
      }
      chunks.push_back({d, e, b0, off_at(offsets, is64, e)});
      max_n = std::max(max_n, e - d); max_b = std::max(max_b, off_at(offsets, is64, e) - b0);
      d = e;
    }
  }
  if ((rc = ensure_chunk_buffers(p, max_n, max_b, !device_input))) return rc;
  if (trace) fprintf(stderr, "[tb] %.2f ms: planned %zu chunks, buffers ready\n", since(), chunks.size());

  uint8_t* h_keep = rsx->keep.as<uint8_t>();
  int8_t* h_failed = rsx->failed.as<int8_t>();
  int64_t* h_coff = rsx->content_off.as<int64_t>();
  uint8_t* h_content = rsx->content.as<uint8_t>();
  int64_t content_base = 0;
  if (p->opt_validate) { if ((rc = rsx->bad.ensure(16))) return rc; *rsx->bad.as<int32_t>() = 0; }
  volatile int32_t* h_bad = p->opt_validate ? rsx->bad.as<int32_t>() : nullptr;

  // pageable text goes through the library's pinned ring (Stager); pinned or registered text is copied from where it is
  bool staged = false;
  Stager stager;
  if (!device_input && p->opt_stage > 0 && total_bytes >= (4ll << 20)) {
    cudaPointerAttributes attr;
    const cudaError_t e = cudaPointerGetAttributes(&attr, utf8 + base0);
    if (e != cudaSuccess) cudaGetLastError();
    staged = e != cudaSuccess || attr.type == cudaMemoryTypeUnregistered;
  }
  if (staged) {
    for (int i = 0; i < Stager::kSlots; i++) {
      if ((rc = p->stage_slots[i].ensure(Stager::kSlotBytes))) return rc;
      if (!p->stage_ev[i]) CK(cudaEventCreateWithFlags(&p->stage_ev[i], cudaEventDisableTiming));
      p->stage_used[i] = false;
    }
    stager.slots = p->stage_slots; stager.slot_ev = p->stage_ev; stager.slot_used = p->stage_used;
    stager.copy = p->copy; stager.T = (int)std::min<int64_t>(p->opt_stage, 16); stager.device = c->device;
    stager.start();
    if (trace) fprintf(stderr, "[tb] %.2f ms: pageable input: staged through %d pinned slots of %zu MB with %d copy threads\n", since(), Stager::kSlots, Stager::kSlotBytes >> 20, stager.T);
  }
  auto enqueue_h2d = [&](size_t ci) -> int {
    const Chunk& ch = chunks[ci];
    int buf = (int)(ci & 1);
    if (device_input) return TB_OK;
    if (staged) {
      const size_t osz = is64 ? 8 : 4;
      stager.push({(int64_t)ci, utf8 + ch.b0, (size_t)(ch.b1 - ch.b0), p->text[buf].as<uint8_t>(), (const uint8_t*)offsets + (size_t)ch.d0 * osz,
                   (size_t)(ch.d1 - ch.d0 + 1) * osz, p->offs[buf].p, p->ev_done[buf], p->ev_h2d[buf]});
      return TB_OK;
    }
    CK(cudaStreamWaitEvent(p->copy, p->ev_done[buf], 0));
    int64_t nb = ch.b1 - ch.b0, n = ch.d1 - ch.d0;
    if (nb) CK(cudaMemcpyAsync(p->text[buf].p, utf8 + ch.b0, (size_t)nb, cudaMemcpyHostToDevice, p->copy));
    size_t osz = is64 ? 8 : 4;
    CK(cudaMemcpyAsync(p->offs[buf].p, (const uint8_t*)offsets + (size_t)ch.d0 * osz, (size_t)(n + 1) * osz, cudaMemcpyHostToDevice, p->copy));
    CK(cudaEventRecord(p->ev_h2d[buf], p->copy));
    return TB_OK;
  };

  // Results of a chunk leave on the d2h stream while the compute stream already runs the next chunk; the
  // host only waits (for the content size) one chunk behind.
  auto finish = [&](size_t ci) -> int {
    const Chunk& ch = chunks[ci];
    const int buf = (int)(ci & 1);
    const int64_t n = ch.d1 - ch.d0;
    CK(cudaEventSynchronize(p->ev_small[buf]));
    if (trace) fprintf(stderr, "[tb] %.2f ms: chunk %zu (%lld docs, %.0f MB) results on the host\n", since(), ci, (long long)n, (ch.b1 - ch.b0) / 1048576.0);
    if (p->opt_validate && *h_bad) return fail(TB_ERR_INVALID_UTF8, "input contains " + std::to_string((int)*h_bad) + " document(s) with invalid UTF-8 (first chunk doc " + std::to_string(ch.d0) + ")");
    if (want_content) {
      int64_t tot = h_coff[ch.d0 + n];
      if (content_mode == 2 && tot > (ch.b1 - ch.b0) / 8 + n)
        return fail(TB_ERR_UNSUPPORTED, "content too fragmented for TB_OPT_RETURN_CONTENT = 2 (more than one run per 8 bytes of text in a chunk): use 1");
      for (int64_t i = 0; i <= n; i++) h_coff[ch.d0 + i] += content_base;  // chunk-relative -> absolute
      if (tot) CK(cudaMemcpyAsync(h_content + (size_t)content_base * esz, p->packed[buf].p, (size_t)tot * esz, cudaMemcpyDeviceToHost, p->d2h));
      content_base += tot;
      if (gather_T > 0) {  // the chunk's runs are on their way: its kept documents can be gathered once they have landed
        while (p->ev_gather.size() <= ci) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); p->ev_gather.push_back(e); }
        CK(cudaEventRecord(p->ev_gather[ci], p->d2h));
        gth.push({ch.d0, ch.d1, content_base, p->ev_gather[ci]});
      }
    }
    CK(cudaEventRecord(p->ev_d2h[buf], p->d2h));
    return TB_OK;
  };

  if (gather_T > 0) {
    gth.src = GatherSrc{(const tb_run*)h_content, h_coff, utf8 + 0, offsets, is64, h_keep};
    gth.dst = rsx->kept.as<uint8_t>(); gth.dst_off = rsx->kept_off.as<int64_t>(); gth.dst_off[0] = 0;
    gth.T = gather_T; gth.device = c->device;
    gth.start();
  }
  if (p->opt_validate) CK(cudaMemsetAsync(p->bad.p, 0, 4, p->compute));
  if (!chunks.empty()) if ((rc = enqueue_h2d(0))) return rc;
  for (size_t ci = 0; ci < chunks.size(); ci++) {
    const Chunk& ch = chunks[ci];
    int buf = (int)(ci & 1);
    int64_t n = ch.d1 - ch.d0;
    if (ci + 1 < chunks.size()) if ((rc = enqueue_h2d(ci + 1))) return rc;
    // result buffers and packed[buf] may still be read by the copies of chunk ci-2
    CK(cudaStreamWaitEvent(p->compute, p->ev_d2h[buf], 0));
    const uint8_t* text;
    if (device_input) {
      text = utf8 + ch.b0;
      tb_launch_init_ranges(p->compute, offsets_dev, 1, ch.d0, n, ch.b0, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), p->alive[buf].as<uint8_t>(), p->failed[buf].as<int8_t>());
    } else {
      if (staged) {  // the event is recorded by the staging thread: it must have been, before the compute stream can wait on it
        const cudaError_t e = stager.wait_recorded((int64_t)ci);
        if (e != cudaSuccess) return fail(TB_ERR_CUDA, std::string("staging thread: ") + cudaGetErrorString(e));
      }
      CK(cudaStreamWaitEvent(p->compute, p->ev_h2d[buf], 0));
      text = p->text[buf].as<uint8_t>();
      tb_launch_init_ranges(p->compute, p->offs[buf].p, is64, 0, n, ch.b0, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), p->alive[buf].as<uint8_t>(), p->failed[buf].as<int8_t>());
    }
    p->launches_this_run++;
    int64_t text_extent = ch.b1 - ch.b0;
    if (p->opt_decode) {  // the decoded copy becomes the text, and its ranges the documents' original ranges
      const size_t bw = ((size_t)text_extent / 32 + 4) * 4;
      CK(cudaMemsetAsync(p->e_amp.p, 0, bw, p->compute));
      CK(cudaMemsetAsync(p->e_drop.p, 0, bw, p->compute));
      CK(cudaMemsetAsync(p->cursors.p, 0, 256, p->compute));
      KBatch eb{text, p->rs0.as<int64_t>(), p->re0.as<int64_t>(), p->alive[buf].as<uint8_t>(), p->failed[buf].as<int8_t>(), n, c->uc, 0, text_extent,
                0u, nullptr, nullptr, nullptr, 0u};
      { Timed t(p, "k_entities");
        tb_launch_entities(p->compute, eb, c->t_ent.as<EntEntry>(), TB_ENT_COUNT, p->e_amp.as<uint32_t>(), p->e_drop.as<uint32_t>(), p->e_D.as<uint8_t>(),
                           p->e_drs.as<int64_t>(), p->e_dre.as<int64_t>(), c->sm_count, p->cursors.as<uint32_t>() + 59); }
      CK(cudaMemcpyAsync(p->rs0.p, p->e_drs.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, p->compute));
      CK(cudaMemcpyAsync(p->re0.p, p->e_dre.p, (size_t)n * 8, cudaMemcpyDeviceToDevice, p->compute));
      text = p->e_D.as<uint8_t>();
      text_extent = (int64_t)tb_grown_bytes(text_extent, n);
    }
    bool has_c4 = false;
    {
      NvtxRange nvtx_chunk("chunk: enqueue chain");
      if ((rc = enqueue_chain(p, text, n, text_extent, buf, content_mode, &has_c4))) return rc;
    }
    CK(cudaEventRecord(p->ev_done[buf], p->compute));
    // the previous chunk's content goes out first (its copy must not queue behind this chunk's kernels)
    if (ci > 0) if ((rc = finish(ci - 1))) return rc;
    // small results
    CK(cudaStreamWaitEvent(p->d2h, p->ev_done[buf], 0));
    CK(cudaMemcpyAsync(h_keep + ch.d0, p->alive[buf].p, (size_t)n, cudaMemcpyDeviceToHost, p->d2h));
    CK(cudaMemcpyAsync(h_failed + ch.d0, p->failed[buf].p, (size_t)n, cudaMemcpyDeviceToHost, p->d2h));
    for (size_t s = 0; s < p->steps.size(); s++) {
      size_t sz = stat_size(p->steps[s].type);
      CK(cudaMemcpyAsync(rsx->stats[s].as<uint8_t>() + (size_t)ch.d0 * sz, p->stats[s][buf].p, (size_t)n * sz, cudaMemcpyDeviceToHost, p->d2h));
    }
    if (want_content) CK(cudaMemcpyAsync(h_coff + ch.d0, p->out_off[buf].p, (size_t)(n + 1) * 8, cudaMemcpyDeviceToHost, p->d2h));
    if (p->opt_validate) CK(cudaMemcpyAsync((void*)h_bad, p->bad.p, 4, cudaMemcpyDeviceToHost, p->d2h));
    CK(cudaEventRecord(p->ev_small[buf], p->d2h));
  }
  if (trace) fprintf(stderr, "[tb] %.2f ms: all chunks enqueued\n", since());
  if (!chunks.empty()) if ((rc = finish(chunks.size() - 1))) return rc;
  if (trace) fprintf(stderr, "[tb] %.2f ms: last chunk's kernels done\n", since());
  CK(cudaStreamSynchronize(p->compute));
  resolve_times(p);
  CK(cudaStreamSynchronize(p->d2h));
  CK(cudaStreamSynchronize(p->copy));
  out->keep = h_keep;
  out->failed_step = h_failed;
  for (size_t s = 0; s < p->steps.size(); s++) out->step_stats[s] = rsx->stats[s].p;
  if (want_content) {
    if (n_docs == 0) h_coff[0] = 0;
    out->content_kind = content_mode;
    if (content_mode == 2) { out->runs = (const tb_run*)h_content; out->run_offsets = h_coff; out->n_runs = content_base; }
    else { out->content_rewritten = 1; out->content = h_content; out->content_offsets = h_coff; out->content_bytes = content_base; }
    if (gather_T > 0) {
      gth.finish();
      if (trace) fprintf(stderr, "[tb] %.2f ms: kept documents gathered (%.1f ms of gathering in all)\n", since(), gth.busy_ms);
      if (gth.err != cudaSuccess) return fail(TB_ERR_CUDA, std::string("gather thread: ") + cudaGetErrorString(gth.err));
      out->kept_content = rsx->kept.as<uint8_t>(); out->kept_offsets = rsx->kept_off.as<int64_t>(); out->n_kept = gth.nk;
    }
  }
  out->gpu_ms = p->ms_this_run;
  out->gpu_launches = p->launches_this_run;
  if (trace) fprintf(stderr, "[tb] %.2f ms: done (kernels %.2f ms)\n", since(), p->ms_this_run);
  return TB_OK;
}

}  // namespace

extern "C" {

int tb_abi_version(void) { return TB_ABI_VERSION; }
const char* tb_last_error(const tb_ctx*) { return g_err.c_str(); }
const char* tb_step_name(int32_t t) {
  switch (t) {
    case TB_STEP_C4_QUALITY: return "C4QualityFilter";
    case TB_STEP_GOPHER_QUALITY: return "GopherQualityFilter";
    case TB_STEP_GOPHER_REPETITION: return "GopherRepetitionFilter";
    case TB_STEP_FINEWEB_QUALITY: return "FineWebQualityFilter";
    default: return "";
  }
}

int tb_ctx_create(int device, tb_ctx** out) {
  if (!out) return fail(TB_ERR_INVALID_ARGUMENT, "out is NULL");
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); return fail(TB_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  if (device < 0 || device >= count) return fail(TB_ERR_INVALID_ARGUMENT, "device index out of range");
  CK(cudaSetDevice(device));
  auto* c = new tb_ctx();
  c->device = device;
  { cudaDeviceProp prop; if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount; }
  if (tb_kernels_init() != 0) { cudaGetLastError(); delete c; return fail(TB_ERR_CUDA, "cudaFuncSetAttribute(dynamic shared memory) failed: not an sm_100-class device?"); }
  int rc;
  if ((rc = upload(c->t_stage1, tb_uc_stage1, sizeof tb_uc_stage1)) || (rc = upload(c->t_stage2, tb_uc_stage2, sizeof tb_uc_stage2)) ||
      (rc = upload(c->t_lower, tb_uc_lower, sizeof tb_uc_lower)) || (rc = upload(c->t_cased, tb_uc_cased, sizeof tb_uc_cased)) ||
      (rc = upload(c->t_caseign, tb_uc_caseign, sizeof tb_uc_caseign))) { delete c; return rc; }
  // the two bases of the string hashes, drawn per context (TB_HASH_SEED fixes them for a reproducible run): an input
  // cannot be built to collide under bases it does not know, and the collision bound of tb_common.cuh is over this draw
  std::vector<uint64_t> pw(TB_POWTAB_ALL);
  {
    uint64_t seed;
    if (const char* e = getenv("TB_HASH_SEED")) seed = strtoull(e, nullptr, 0);
    else { std::random_device rd; seed = ((uint64_t)rd() << 32) ^ rd() ^ (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count(); }
    const uint64_t r1 = tb_mix64(seed + 0x9E3779B97F4A7C15ull), r2 = tb_mix64(seed + 2 * 0x9E3779B97F4A7C15ull);
    const uint32_t half = 1u << (TB_BASE_BITS - 1);
    uint32_t b1 = half + (uint32_t)(r1 % half), b2 = half + (uint32_t)(r2 % half);  // in [2^28, 2^29)
    if (b2 == b1) b2 = half + (b2 + 1u - half) % half;
    tb_fill_powtab(pw.data(), b1, b2);
  }
  if ((rc = upload(c->t_pow, pw.data(), pw.size() * 8))) { delete c; return rc; }
  if ((rc = upload(c->t_ent, tb_ent_table, sizeof tb_ent_table))) { delete c; return rc; }
  c->uc = UcTables{c->t_stage1.as<uint16_t>(), c->t_stage2.as<uint16_t>(), c->t_lower.as<uint32_t>(), c->t_cased.as<uint32_t>(),
                   c->t_caseign.as<uint32_t>(), c->t_pow.as<uint64_t>(), TB_UC_NLOWER, TB_UC_NCASED, TB_UC_NCASEIGN, {0}};
  tb_compute_wslead(tb_uc_stage1, tb_uc_stage2, c->uc.wslead);
  *out = c;
  return TB_OK;
}

void tb_ctx_destroy(tb_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  for (DevBuf* b : {&c->t_stage1, &c->t_stage2, &c->t_lower, &c->t_cased, &c->t_caseign, &c->t_pow, &c->t_ent}) b->release();
  delete c;
}

int tb_pipeline_create(tb_ctx* ctx, const tb_step_cfg* steps, size_t n_steps, tb_pipeline** out) {
  if (!ctx || !out || (n_steps && !steps)) return fail(TB_ERR_INVALID_ARGUMENT, "NULL argument");
  *out = nullptr;
  if (n_steps > TB_MAX_STEPS) return fail(TB_ERR_INVALID_ARGUMENT, "at most " + std::to_string(TB_MAX_STEPS) + " steps");
  int n_c4 = 0;
  for (size_t i = 0; i < n_steps; i++) {
    int rc = validate_step(steps[i]);
    if (rc) return rc;
    n_c4 += steps[i].type == TB_STEP_C4_QUALITY;
  }
  if (n_c4 > 1) {  // several C4 steps: each reads the output of the one before it; sentence-wise lines grow the text, once
    for (size_t i = 0; i < n_steps; i++)
      if (steps[i].type == TB_STEP_C4_QUALITY && !steps[i].u.c4.split_paragraph)
        return fail(TB_ERR_UNSUPPORTED, "several C4QualityFilter steps in one pipeline: only with split_paragraph = true");
  }
  CK(cudaSetDevice(ctx->device));
  auto* p = new tb_pipeline();
  p->ctx = ctx;
  p->n_c4 = n_c4;
  p->steps.resize(n_steps);
  for (size_t i = 0; i < n_steps; i++) {
    StepHost& sh = p->steps[i];
    sh.type = steps[i].type;
    if (sh.type == TB_STEP_C4_QUALITY) {
      const tb_c4_params& q = steps[i].u.c4;
      sh.c4 = tb_c4cfg_of(q);
      sh.c4_sentences = !q.split_paragraph;
    } else if (sh.type == TB_STEP_GOPHER_QUALITY) {
      const tb_gq_params& q = steps[i].u.gq;
      sh.gq = GqCfg{q.has, q.min_doc_words, q.max_doc_words, q.min_stop_words, q.min_avg_word_length, q.max_avg_word_length,
                    q.max_symbol_word_ratio, q.max_bullet_lines_ratio, q.max_ellipsis_lines_ratio, q.max_non_alpha_words_ratio};
      int rc = build_stop_table(sh, q);
      if (rc) { tb_pipeline_destroy(p); return rc; }
    } else if (sh.type == TB_STEP_FINEWEB_QUALITY) {
      const tb_fw_params& q = steps[i].u.fw;
      FwCfg f{};
      f.line_punct_thr = q.line_punct_thr; f.short_line_thr = q.short_line_thr; f.char_duplicates_ratio = q.char_duplicates_ratio;
      f.new_line_ratio = q.new_line_ratio; f.short_line_length = q.short_line_length;
      f.line_punct_exclude_zero = q.line_punct_exclude_zero != 0; f.n_stop_chars = q.n_stop_chars;
      for (int k = 0; k < TB_FW_MAX_STOP_CHARS; k++) f.stop_chars[k] = q.stop_chars[k];
      sh.fw = f;
    } else {
      const tb_gr_params& q = steps[i].u.gr;
      GrCfg g{};
      g.has = q.has; g.n_top = q.n_top; g.n_dup = q.n_dup;
      g.dup_line_frac = q.dup_line_frac; g.dup_para_frac = q.dup_para_frac;
      g.dup_line_char_frac = q.dup_line_char_frac; g.dup_para_char_frac = q.dup_para_char_frac;
      for (int k = 0; k < TB_MAX_NGRAMS; k++) { g.top_n[k] = q.top_n[k]; g.top_frac[k] = q.top_frac[k]; g.dup_n[k] = q.dup_n[k]; g.dup_frac[k] = q.dup_frac[k]; }
      tb_grcfg_finish(g);
      sh.gr = g;
    }
  }
  CK(cudaStreamCreateWithFlags(&p->compute, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&p->copy, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&p->d2h, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
  for (int i = 0; i < 2; i++) {
    CK(cudaEventCreateWithFlags(&p->ev_h2d[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&p->ev_done[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&p->ev_d2h[i], cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&p->ev_small[i], cudaEventDisableTiming));
  }
  *out = p;
  return TB_OK;
}

void tb_pipeline_destroy(tb_pipeline* p) {
  if (!p) return;
  cudaSetDevice(p->ctx->device);
  cudaDeviceSynchronize();
  for (DevBuf* b : {&p->text[0], &p->text[1], &p->offs[0], &p->offs[1], &p->packed[0], &p->packed[1], &p->rs0, &p->re0, &p->rs,
                    &p->re, &p->ts, &p->te, &p->alive[0], &p->alive[1], &p->failed[0], &p->failed[1], &p->rewritten, &p->len, &p->out_off[0], &p->out_off[1], &p->bad, &p->cursors, &p->order, &p->sizecls, &p->g_wh,
                    &p->g_wl, &p->g_ws, &p->g_lh, &p->g_ll, &p->g_lg, &p->g_doc, &p->g_work, &p->g_lists, &p->g_counters, &p->f_lh, &p->f_ll, &p->f_doc, &p->f_tab, &p->e_D, &p->e_amp, &p->e_drop, &p->e_drs, &p->e_dre, &p->c_S, &p->c_bits, &p->c_srs, &p->c_sre, &p->c_ka, &p->c_kf, &p->c_P, &p->c_F, &p->c_F2, &p->c_plen, &p->c_nlines,
                    &p->c_docflags, &p->c_lwc, &p->c_lmax, &p->c_lflags, &p->c_prs, &p->c_pre, &p->c_frs, &p->c_fre, &p->c_cnt})
    b->release();
  for (auto& b : p->stats) { b[0].release(); b[1].release(); }
  for (auto& s : p->steps) s.sw_dev.release();
  for (auto* r : p->results) {
    r->keep.release(); r->failed.release(); r->content.release(); r->content_off.release(); r->kept.release(); r->kept_off.release(); r->bad.release();
    for (auto& s : r->stats) s.release();
    delete r;
  }
  for (int i = 0; i < 2; i++) {
    if (p->ev_h2d[i]) cudaEventDestroy(p->ev_h2d[i]);
    if (p->ev_done[i]) cudaEventDestroy(p->ev_done[i]);
    if (p->ev_d2h[i]) cudaEventDestroy(p->ev_d2h[i]);
    if (p->ev_small[i]) cudaEventDestroy(p->ev_small[i]);
  }
  if (p->compute) cudaStreamDestroy(p->compute);
  if (p->copy) cudaStreamDestroy(p->copy);
  if (p->d2h) cudaStreamDestroy(p->d2h);
  if (p->side) cudaStreamDestroy(p->side);
  if (p->ev_fork) cudaEventDestroy(p->ev_fork);
  if (p->ev_join) cudaEventDestroy(p->ev_join);
  for (auto e : p->ev_pool) cudaEventDestroy(e);
  for (auto e : p->ev_gather) cudaEventDestroy(e);
  for (int i = 0; i < Stager::kSlots; i++) { p->stage_slots[i].release(); if (p->stage_ev[i]) cudaEventDestroy(p->stage_ev[i]); }
  delete p;
}

int tb_pipeline_set_option(tb_pipeline* p, int32_t option, int64_t value) {
  if (!p) return fail(TB_ERR_INVALID_ARGUMENT, "NULL pipeline");
  std::lock_guard<std::mutex> lk(p->mu);
  switch (option) {
    case TB_OPT_RETURN_CONTENT:
      if (value < 0 || value > 2) return fail(TB_ERR_INVALID_ARGUMENT, "TB_OPT_RETURN_CONTENT: 0 (nothing), 1 (content) or 2 (runs of the input text)");
      p->opt_return_content = value; return TB_OK;
    case TB_OPT_VALIDATE_UTF8: p->opt_validate = value != 0; return TB_OK;
    case TB_OPT_GENERAL_WORD_PATH: p->opt_debug = value != 0 ? TB_DBG_FORCE_WALK : 0; return TB_OK;
    case TB_OPT_DECODE_HTML_ENTITIES: p->opt_decode = value != 0; return TB_OK;
    case TB_OPT_CHUNK_RAMP: p->opt_ramp = value != 0; return TB_OK;
    case TB_OPT_BULK_STAGE_TILES:
      if (value < 0 || value > (1 << 30)) return fail(TB_ERR_INVALID_ARGUMENT, "TB_OPT_BULK_STAGE_TILES: smallest document, in 1 KiB tiles, whose text is staged by bulk copies; 0 = off");
      p->opt_bulk = value; return TB_OK;
    case TB_OPT_STAGE_THREADS:
      if (value < 0) return fail(TB_ERR_INVALID_ARGUMENT, "TB_OPT_STAGE_THREADS: number of host threads, 0 = off");
      p->opt_stage = value; return TB_OK;
    case TB_OPT_GATHER_KEPT:
      if (value < 0) return fail(TB_ERR_INVALID_ARGUMENT, "TB_OPT_GATHER_KEPT: number of host threads, 0 = off");
      p->opt_gather = value; return TB_OK;
    case TB_OPT_CHUNK_BYTES:
      if (value < (1 << 16)) return fail(TB_ERR_INVALID_ARGUMENT, "chunk size must be at least 64 KiB");
      p->opt_chunk = value; return TB_OK;
    default: return fail(TB_ERR_INVALID_ARGUMENT, "unknown option");
  }
}

int tb_pipeline_run_batch(tb_pipeline* p, const uint8_t* utf8, const void* offsets, int offsets_are_i64, int64_t n_docs,
                          tb_batch_out* out) {
  if (!p || !out || n_docs < 0 || (n_docs > 0 && !offsets)) return fail(TB_ERR_INVALID_ARGUMENT, "NULL or negative argument");
  std::lock_guard<std::mutex> lk(p->mu);
  int rc = run_impl(p, utf8, offsets, offsets_are_i64, n_docs, false, nullptr, out);
  if (rc) { cudaDeviceSynchronize(); cudaGetLastError(); if (out->_owner) { ((ResultSet*)out->_owner)->in_use = false; out->_owner = nullptr; } }
  return rc;
}

int tb_pipeline_run_batch_device(tb_pipeline* p, const uint8_t* utf8_dev, const int64_t* offsets_dev, int64_t n_docs,
                                 int64_t n_bytes, tb_batch_out* out) {
  if (!p || !out || n_docs < 0 || (n_docs > 0 && !offsets_dev)) return fail(TB_ERR_INVALID_ARGUMENT, "NULL or negative argument");
  (void)n_bytes;
  std::lock_guard<std::mutex> lk(p->mu);
  CK(cudaSetDevice(p->ctx->device));
  std::vector<int64_t> h_off((size_t)n_docs + 1, 0);
  if (n_docs) CK(cudaMemcpy(h_off.data(), offsets_dev, ((size_t)n_docs + 1) * 8, cudaMemcpyDeviceToHost));
  int rc = run_impl(p, utf8_dev, h_off.data(), 1, n_docs, true, offsets_dev, out);
  if (rc) { cudaDeviceSynchronize(); cudaGetLastError(); if (out->_owner) { ((ResultSet*)out->_owner)->in_use = false; out->_owner = nullptr; } }
  return rc;
}

void tb_batch_out_release(tb_pipeline* p, tb_batch_out* out) {
  if (!p || !out || !out->_owner) return;
  std::lock_guard<std::mutex> lk(p->mu);
  ((ResultSet*)out->_owner)->in_use = false;
  memset(out, 0, sizeof *out);
}

int tb_gather_content(const tb_batch_out* out, const uint8_t* utf8, const void* offsets, int offsets_are_i64, const uint8_t* select,
                      uint8_t* dst, int64_t dst_capacity, int64_t* dst_offsets, int64_t* dst_bytes, int nthreads) {
  if (!out || !dst_offsets || (out->n_docs > 0 && !offsets)) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gather_content: NULL argument");
  if (out->content_kind != 2 || !out->run_offsets) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gather_content: the batch holds no runs (TB_OPT_RETURN_CONTENT = 2)");
  const int64_t n = out->n_docs;
  const tb_run* runs = out->runs;
  const int64_t* ro = out->run_offsets;
  const int T = nthreads > 0 ? std::min(nthreads, 64) : (int)std::min<int64_t>(8, n / 32768 + 1);
  GatherSrc src{runs, ro, utf8, offsets, offsets_are_i64, select};
  int64_t nk = 0, acc = 0;
  dst_offsets[0] = 0;
  if (dst) {  // sizes first: the destination must hold them
    std::vector<int64_t> tmp((size_t)n + 2);
    gather_range(src, 0, n, n ? ro[n] : 0, nullptr, tmp.data(), 0, 0, T, &nk, &acc);
    if (acc > dst_capacity) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gather_content: dst holds " + std::to_string(dst_capacity) + " bytes, " + std::to_string(acc) + " needed");
    if (acc > 0 && !utf8) return fail(TB_ERR_INVALID_ARGUMENT, "tb_gather_content: utf8 is NULL");
  }
  gather_range(src, 0, n, n ? ro[n] : 0, dst, dst_offsets, 0, 0, T, &nk, &acc);
  if (dst_bytes) *dst_bytes = acc;
  return TB_OK;
}

int tb_pipeline_kernel_times(tb_pipeline* p, int32_t max_entries, const char** names, double* ms, int64_t* launches) {
  if (!p) return 0;
  std::lock_guard<std::mutex> lk(p->mu);
  int n = (int)std::min<size_t>(p->ktimes.size(), (size_t)std::max(0, max_entries));
  for (int i = 0; i < n; i++) { names[i] = p->ktimes[i].name; ms[i] = p->ktimes[i].ms; launches[i] = p->ktimes[i].launches; }
  return (int)p->ktimes.size();
}
void tb_pipeline_reset_kernel_times(tb_pipeline* p) {
  if (!p) return;
  std::lock_guard<std::mutex> lk(p->mu);
  for (auto& k : p->ktimes) { k.ms = 0; k.launches = 0; }
}

}  // extern "C"
