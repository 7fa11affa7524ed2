// textblaster_b200.hpp -- C++ host-side mirror of the reference's step interface for the accelerated path,
// header-only, on top of the C ABI (textblaster_b200.h).  The reference is Rust; a Rust toolchain is not
// available in this build image, so this is the compiled-language host side: same type names, argument
// order and error behaviour as
//   TextDocument                 src/data_model.rs:5-17
//   PipelineError variants       src/error.rs:9-66   (ConfigValidationError, DocumentFiltered, StepError, Unexpected)
//   trait ProcessingStep         src/executor.rs:8-15
//   PipelineExecutor             src/executor.rs:17-70
//   C4QualityFilter::new         src/pipeline/filters/c4_filters.rs:112-138
//   GopherQualityFilter::new     src/pipeline/filters/gopher_quality.rs:34-60
//   GopherRepetitionFilter::new  src/pipeline/filters/gopher_rep.rs:25-43
//   FineWebQualityFilter::new    src/pipeline/filters/fineweb_quality.rs:41-62
// Reason strings and metadata keys are rebuilt from the integer counters with the reference's format
// strings (c4_filters.rs:167-293, gopher_quality.rs:200-315, gopher_rep.rs:61-218, fineweb_quality.rs:79-222).
// There is no CPU path: constructing a step without a CUDA device throws Unexpected.
#pragma once
#include <cstdint>
#include <cstdio>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <variant>
#include <vector>

#include "textblaster_b200.h"

namespace textblaster {

struct TextDocument {  // src/data_model.rs:5-17
  std::string id, source, content;
  std::map<std::string, std::string> metadata;
  std::optional<int32_t> added;                        // days, carried through untouched
  std::optional<std::pair<int64_t, int64_t>> created;  // (start, end) in microseconds, carried through untouched
};

// ---- errors ------------------------------------------------------------------------------------
class PipelineError : public std::runtime_error {  // enum PipelineError, src/error.rs:9-10
 public:
  explicit PipelineError(const std::string& what) : std::runtime_error(what) {}
};
class ConfigValidationError : public PipelineError {  // src/error.rs:62-63
 public:
  explicit ConfigValidationError(const std::string& m) : PipelineError("Configuration validation error: " + m), message(m) {}
  std::string message;
};
class Unexpected : public PipelineError {  // src/error.rs:65-66; also CUDA / library failures
 public:
  explicit Unexpected(const std::string& m) : PipelineError("Unexpected error: " + m), message(m) {}
  std::string message;
};
class DocumentFiltered : public PipelineError {  // src/error.rs:33-37: the document travels inside the error
 public:
  DocumentFiltered(TextDocument doc, const std::string& r)
      : PipelineError("Document '" + doc.id + "' filtered out: " + r), document(std::move(doc)), reason(r) {}
  TextDocument document;
  std::string reason;
};
class StepError : public PipelineError {  // src/error.rs:39-43, how the executor wraps a step's error (executor.rs:37-46)
 public:
  StepError(const std::string& step, std::shared_ptr<DocumentFiltered> src)
      : PipelineError("Error in processing step '" + step + "': " + src->what()), step_name(step), source(std::move(src)) {}
  std::string step_name;
  std::shared_ptr<DocumentFiltered> source;
};

namespace detail {
inline std::string f2(double v) { char b[64]; snprintf(b, sizeof b, "%.2f", v); return b; }
inline std::string f4(double v) { char b[64]; snprintf(b, sizeof b, "%.4f", v); return b; }
inline void check(int rc) {
  if (rc == TB_OK) return;
  const char* m = tb_last_error(nullptr);
  std::string msg = m ? m : "";
  if (rc == TB_ERR_CONFIG_VALIDATION) throw ConfigValidationError(msg);
  throw Unexpected(msg);
}
struct Ctx {  // one context per device per process
  tb_ctx* h = nullptr;
  explicit Ctx(int device) { check(tb_ctx_create(device, &h)); }
  ~Ctx() { tb_ctx_destroy(h); }
};
inline std::shared_ptr<Ctx> ctx_for(int device) {
  static std::map<int, std::weak_ptr<Ctx>> cache;
  auto it = cache.find(device);
  if (it != cache.end()) if (auto sp = it->second.lock()) return sp;
  auto sp = std::make_shared<Ctx>(device);
  cache[device] = sp;
  return sp;
}
}  // namespace detail

// ---- step configuration (src/config/pipeline.rs:66-115, 161-178, 311-322) --------------------
struct StepSpec {           // a tb_step_cfg plus the storage its pointers refer to
  tb_step_cfg cfg{};
  std::string stop_bytes;   // GopherQuality stop words, concatenated
  std::vector<int32_t> stop_offsets;
  std::vector<std::pair<uint32_t, double>> top_n_grams, dup_n_grams;
  bool custom_stop_words = false;
  void fix_pointers() {
    if (cfg.type == TB_STEP_GOPHER_QUALITY && custom_stop_words) {
      cfg.u.gq.stop_words_utf8 = (const uint8_t*)stop_bytes.data();
      cfg.u.gq.stop_words_offsets = stop_offsets.data();
    }
  }
};

// ---- reason strings / metadata from the counters ------------------------------------------------
namespace detail {
using Meta = std::map<std::string, std::string>;
inline std::string join(const std::vector<std::string>& v) {
  std::string s;
  for (size_t i = 0; i < v.size(); i++) { if (i) s += "; "; s += v[i]; }
  return s;
}
// each returns the reason ("" = passed) and adds the keys the step writes
inline std::string c4_outcome(const tb_c4_params& p, const tb_c4_stats& st, Meta& md) {  // c4_filters.rs:165-293
  std::vector<std::string> r;
  if (st.reason_bits & TB_C4_REASON_LOREM_IPSUM) r.push_back("lorem_ipsum");
  if (st.reason_bits & TB_C4_REASON_CURLY_BRACKET) r.push_back("curly_bracket");
  if (st.reason_bits & TB_C4_REASON_TOO_FEW_SENTENCES)
    r.push_back("too_few_sentences (found " + std::to_string(st.n_sentences) + ", required " + std::to_string(p.min_num_sentences) + ")");
  if (r.empty()) { md["c4_filter_status"] = "passed"; return ""; }
  const std::string rs = join(r);
  md["c4_filter_status"] = "filtered"; md["c4_filter_reasons"] = rs;
  if (st.rewritten) {
    if (st.too_long_word) md["line-filter-too_long_word"] = std::to_string(st.too_long_word);
    if (st.no_terminal_punc) md["line-filter-no_terminal_punc"] = std::to_string(st.no_terminal_punc);
    if (st.too_few_words) md["line-filter-too_few_words"] = std::to_string(st.too_few_words);
  }
  return rs;
}
inline std::string gq_outcome(const tb_gq_params& p, const tb_gq_stats& st, Meta& md) {  // gopher_quality.rs:198-315
  const uint64_t n = st.n_words;
  const double nw = (double)(n ? n : 1), nl = (double)(st.n_lines ? st.n_lines : 1);
  const double avg = n ? (double)st.sum_word_chars / (double)n : 0.0;
  std::vector<std::string> r;
  const uint32_t b = st.reason_bits;
  if (b & TB_GQ_REASON_SHORT_DOC) r.push_back("gopher_short_doc (" + std::to_string(n) + " non-symbol words, required " + std::to_string(p.min_doc_words) + ")");
  if (b & TB_GQ_REASON_LONG_DOC) r.push_back("gopher_long_doc (" + std::to_string(n) + " non-symbol words, max " + std::to_string(p.max_doc_words) + ")");
  if (b & TB_GQ_REASON_BELOW_AVG)
    r.push_back("gopher_below_avg_threshold (avg len " + f2(avg) + ", required " + f2(p.min_avg_word_length) +
                ((n == 0 && p.min_avg_word_length > 0.0) ? " - 0 non-symbol words" : "") + ")");
  if (b & TB_GQ_REASON_ABOVE_AVG) r.push_back("gopher_above_avg_threshold (avg len " + f2(avg) + ", max " + f2(p.max_avg_word_length) + ")");
  if (b & TB_GQ_REASON_TOO_MANY_HASHES) r.push_back("gopher_too_many_hashes (ratio " + f2((double)st.n_hash / nw) + ", max " + f2(p.max_symbol_word_ratio) + ")");
  if (b & TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS) r.push_back("gopher_too_many_ellipsis_units (ratio " + f2((double)st.n_ellipsis_units / nw) + ", max " + f2(p.max_symbol_word_ratio) + ")");
  if (b & TB_GQ_REASON_TOO_MANY_BULLETS) r.push_back("gopher_too_many_bullets (ratio " + f2((double)st.n_bullet_lines / nl) + ", max " + f2(p.max_bullet_lines_ratio) + ")");
  if (b & TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES) r.push_back("gopher_too_many_end_ellipsis_lines (ratio " + f2((double)st.n_ellipsis_lines / nl) + ", max " + f2(p.max_ellipsis_lines_ratio) + ")");
  if (b & TB_GQ_REASON_BELOW_ALPHA) r.push_back("gopher_below_alpha_threshold (alpha ratio " + f2((double)st.n_alpha_words / nw) + ", required min " + f2(p.max_non_alpha_words_ratio) + ")");
  if (b & TB_GQ_REASON_TOO_FEW_STOP_WORDS) r.push_back("gopher_too_few_stop_words (found " + std::to_string(st.n_stop_words) + ", required " + std::to_string(p.min_stop_words) + ")");
  if (r.empty()) { md["gopher_quality_filter_status"] = "passed"; return ""; }
  const std::string rs = join(r);
  md["gopher_quality_filter_status"] = "filtered"; md["gopher_quality_filter_reasons"] = rs;
  return rs;
}
inline std::string gr_outcome(const StepSpec& sp, const tb_gr_stats& st, Meta& md) {  // gopher_rep.rs:60-218
  const tb_gr_params& p = sp.cfg.u.gr;
  if (st.reason_bits & TB_GR_REASON_EMPTY) {  // note the singular key, gopher_rep.rs:65-68
    md["gopher_repetition_filter_status"] = "filtered"; md["gopher_repetition_filter_reason"] = "skipping empty content";
    return "skipping empty content";
  }
  const double tcl = (double)(st.char_len ? st.char_len : 1);
  std::vector<std::string> r;
  const uint32_t b = st.reason_bits;
  if (b & TB_GR_REASON_DUP_PARA_FRAC) r.push_back("dup_para_frac (ratio " + f2((double)st.para_dup_elems / (double)(st.n_paras ? st.n_paras : 1)) + ", max " + f2(p.dup_para_frac) + ")");
  if (b & TB_GR_REASON_DUP_PARA_CHAR_FRAC) r.push_back("dup_para_char_frac (ratio " + f2((double)st.para_dup_bytes / tcl) + ", max " + f2(p.dup_para_char_frac) + ")");
  if (b & TB_GR_REASON_DUP_LINE_FRAC) r.push_back("dup_line_frac (ratio " + f2((double)st.line_dup_elems / (double)(st.n_lines ? st.n_lines : 1)) + ", max " + f2(p.dup_line_frac) + ")");
  if (b & TB_GR_REASON_DUP_LINE_CHAR_FRAC) r.push_back("dup_line_char_frac (ratio " + f2((double)st.line_dup_bytes / tcl) + ", max " + f2(p.dup_line_char_frac) + ")");
  for (int i = 0; i < p.n_top; i++)
    if (b & (1u << (TB_GR_REASON_TOP_SHIFT + i))) r.push_back("top_" + std::to_string(p.top_n[i]) + "_gram (ratio " + f2((double)st.top_bytes[i] / tcl) + ", max " + f2(p.top_frac[i]) + ")");
  for (int i = 0; i < p.n_dup; i++)
    if (b & (1u << (TB_GR_REASON_DUP_SHIFT + i))) r.push_back("duplicated_" + std::to_string(p.dup_n[i]) + "_n_grams (ratio " + f2((double)st.dup_bytes[i] / tcl) + ", max " + f2(p.dup_frac[i]) + ")");
  if (r.empty()) { md["gopher_repetition_filter_status"] = "passed"; return ""; }
  const std::string rs = join(r);
  md["gopher_repetition_filter_status"] = "filtered"; md["gopher_repetition_filter_reasons"] = rs;
  return rs;
}
inline std::string fw_outcome(const tb_fw_params& p, const tb_fw_stats& st, Meta& md) {  // fineweb_quality.rs:79-225
  const uint32_t b = st.reason_bits;
  if (!b) return "";  // nothing is written on pass (:225)
  md["fineweb_filter_status"] = "filtered";
  if (b & TB_FW_REASON_EMPTY) { md["fineweb_filter_reason"] = "empty document"; return "empty"; }
  const double nl = (double)st.n_lines;
  std::string r;
  if (b & TB_FW_REASON_LINE_PUNCT)
    r = "line_punct_ratio: " + f4((double)st.n_stop_lines / nl) + " < threshold " + f4(p.line_punct_thr) + " (exclude_zero: " + (p.line_punct_exclude_zero ? "true" : "false") + ")";
  else if (b & TB_FW_REASON_SHORT_LINE) r = "short_line_ratio: " + f4((double)st.n_short_lines / nl) + " > threshold " + f4(p.short_line_thr);
  else if (b & TB_FW_REASON_CHAR_DUP) r = "char_dup_ratio: " + f4((double)st.dup_line_bytes / (double)st.n_chars_no_newline) + " > threshold " + f4(p.char_duplicates_ratio);
  else if (b & TB_FW_REASON_LIST_NO_WORDS) r = "list_ratio_no_words (newlines present but no words)";
  else r = "list_ratio: " + f4((double)st.n_newlines / (double)st.n_words) + " > threshold " + f4(p.new_line_ratio);
  md["fineweb_filter_reason"] = r;
  return r;
}
inline std::string step_outcome(const StepSpec& sp, const void* stats, int64_t i, Meta& md) {
  switch (sp.cfg.type) {
    case TB_STEP_C4_QUALITY: return c4_outcome(sp.cfg.u.c4, ((const tb_c4_stats*)stats)[i], md);
    case TB_STEP_GOPHER_QUALITY: return gq_outcome(sp.cfg.u.gq, ((const tb_gq_stats*)stats)[i], md);
    case TB_STEP_GOPHER_REPETITION: return gr_outcome(sp, ((const tb_gr_stats*)stats)[i], md);
    default: return fw_outcome(sp.cfg.u.fw, ((const tb_fw_stats*)stats)[i], md);
  }
}
}  // namespace detail

// ---- a chain of accelerated steps on one GPU (tb_pipeline) ------------------------------------------
using StepResult = std::variant<TextDocument, std::shared_ptr<StepError>>;  // Result<TextDocument>, executor.rs:30

class GpuPipeline {
 public:
  GpuPipeline(std::vector<StepSpec> specs, int device = 0) : specs_(std::move(specs)), ctx_(detail::ctx_for(device)) {
    std::vector<tb_step_cfg> cfgs;
    for (auto& s : specs_) { s.fix_pointers(); cfgs.push_back(s.cfg); }
    detail::check(tb_pipeline_create(ctx_->h, cfgs.data(), cfgs.size(), &h_));
  }
  ~GpuPipeline() { if (h_) tb_pipeline_destroy(h_); }
  GpuPipeline(const GpuPipeline&) = delete;
  GpuPipeline& operator=(const GpuPipeline&) = delete;
  // TB_OPT_*; e.g. TB_OPT_DECODE_HTML_ENTITIES = 1 for callers whose reader does not decode entities itself
  // (the reference's does, src/pipeline/readers/parquet_reader.rs:178-179)
  void set_option(int32_t option, int64_t value) { detail::check(tb_pipeline_set_option(h_, option, value)); }

  // One result per input document, in input order (run_batch_parallel_async, executor.rs:60-70).  `names`:
  // the step names used in StepError.
  std::vector<StepResult> run(const std::vector<TextDocument>& docs, const std::vector<std::string>& names) const {
    std::vector<StepResult> out;
    if (docs.empty()) return out;
    std::string bytes;
    std::vector<int64_t> off(docs.size() + 1, 0);
    for (size_t i = 0; i < docs.size(); i++) { bytes += docs[i].content; off[i + 1] = (int64_t)bytes.size(); }
    tb_batch_out bo;
    detail::check(tb_pipeline_run_batch(h_, (const uint8_t*)bytes.data(), off.data(), 1, (int64_t)docs.size(), &bo));
    for (size_t i = 0; i < docs.size(); i++) {
      TextDocument d = docs[i];
      const int fs = bo.failed_step[i];
      const int last = fs >= 0 ? fs : (int)specs_.size() - 1;
      std::string reason;
      for (int s = 0; s <= last; s++) {
        std::string r = detail::step_outcome(specs_[s], bo.step_stats[s], (int64_t)i, d.metadata);
        if (s == fs) reason = r;
      }
      if (bo.content_rewritten)
        d.content.assign((const char*)bo.content + bo.content_offsets[i], (size_t)(bo.content_offsets[i + 1] - bo.content_offsets[i]));
      if (fs < 0) out.emplace_back(std::move(d));
      else out.emplace_back(std::make_shared<StepError>(names[fs], std::make_shared<DocumentFiltered>(std::move(d), reason)));
    }
    tb_batch_out_release(h_, &bo);
    return out;
  }
  const std::vector<StepSpec>& specs() const { return specs_; }

 private:
  std::vector<StepSpec> specs_;
  std::shared_ptr<detail::Ctx> ctx_;
  tb_pipeline* h_ = nullptr;
};

// ---- trait ProcessingStep (src/executor.rs:8-15) ------------------------------------------------------
class ProcessingStep {
 public:
  virtual ~ProcessingStep() = default;
  virtual const char* name() const = 0;
  // Ok(document) or throws DocumentFiltered (the document, with the step's metadata, inside)
  virtual TextDocument process(TextDocument document) const = 0;
};

class GpuStep : public ProcessingStep {  // a one-step pipeline, driven with a batch of one
 public:
  TextDocument process(TextDocument document) const override {
    if (!pipe_) pipe_ = std::make_unique<GpuPipeline>(std::vector<StepSpec>{spec_}, device_);
    auto r = pipe_->run({std::move(document)}, {name()});
    if (auto* ok = std::get_if<TextDocument>(&r[0])) return std::move(*ok);
    throw *std::get<std::shared_ptr<StepError>>(r[0])->source;
  }
  const StepSpec& spec() const { return spec_; }
  int device() const { return device_; }

 protected:
  GpuStep(StepSpec spec, int device) : spec_(std::move(spec)), device_(device) {
    GpuPipeline probe({spec_}, device_);  // validates the parameters (ConfigValidationError) and the device (Unexpected)
  }
  StepSpec spec_;
  int device_;
  mutable std::unique_ptr<GpuPipeline> pipe_;
};

// ---- the four accelerated filters (constructor arguments in the reference's order) ---------------------
class C4QualityFilter : public GpuStep {  // C4QualityFilter::new, c4_filters.rs:112-138
 public:
  C4QualityFilter(bool split_paragraph, bool remove_citations, bool filter_no_terminal_punct, size_t min_num_sentences,
                  size_t min_words_per_line, size_t max_word_length, bool filter_lorem_ipsum, bool filter_javascript,
                  bool filter_curly_bracket, bool filter_policy, int device = 0)
      : GpuStep(make(split_paragraph, remove_citations, filter_no_terminal_punct, min_num_sentences, min_words_per_line,
                     max_word_length, filter_lorem_ipsum, filter_javascript, filter_curly_bracket, filter_policy), device) {}
  const char* name() const override { return "C4QualityFilter"; }

 private:
  static StepSpec make(bool sp, bool rc, bool tp, size_t ns, size_t wl, size_t mw, bool li, bool js, bool cb, bool po) {
    StepSpec s;
    s.cfg.type = TB_STEP_C4_QUALITY;
    tb_c4_params& p = s.cfg.u.c4;
    p.split_paragraph = sp; p.remove_citations = rc; p.filter_no_terminal_punct = tp; p.filter_lorem_ipsum = li;
    p.filter_javascript = js; p.filter_curly_bracket = cb; p.filter_policy = po;
    p.min_num_sentences = ns; p.min_words_per_line = wl; p.max_word_length = mw;
    return s;
  }
};

class GopherQualityFilter : public GpuStep {  // GopherQualityFilter::new, gopher_quality.rs:34-60 (None = std::nullopt)
 public:
  GopherQualityFilter(std::optional<size_t> min_doc_words, std::optional<size_t> max_doc_words,
                      std::optional<double> min_avg_word_length, std::optional<double> max_avg_word_length,
                      std::optional<double> max_symbol_word_ratio, std::optional<double> max_bullet_lines_ratio,
                      std::optional<double> max_ellipsis_lines_ratio, std::optional<double> max_non_alpha_words_ratio,
                      std::optional<size_t> min_stop_words, std::optional<std::vector<std::string>> stop_words, int device = 0)
      : GpuStep(make(min_doc_words, max_doc_words, min_avg_word_length, max_avg_word_length, max_symbol_word_ratio,
                     max_bullet_lines_ratio, max_ellipsis_lines_ratio, max_non_alpha_words_ratio, min_stop_words, stop_words),
                device) {}
  const char* name() const override { return "GopherQualityFilter"; }

 private:
  static StepSpec make(std::optional<size_t> a, std::optional<size_t> b, std::optional<double> c, std::optional<double> d,
                       std::optional<double> e, std::optional<double> f, std::optional<double> g, std::optional<double> h,
                       std::optional<size_t> i, const std::optional<std::vector<std::string>>& sw) {
    StepSpec s;
    s.cfg.type = TB_STEP_GOPHER_QUALITY;
    tb_gq_params& p = s.cfg.u.gq;
    if (a) { p.has |= TB_GQ_HAS_MIN_DOC_WORDS; p.min_doc_words = *a; }
    if (b) { p.has |= TB_GQ_HAS_MAX_DOC_WORDS; p.max_doc_words = *b; }
    if (c) { p.has |= TB_GQ_HAS_MIN_AVG_WORD_LENGTH; p.min_avg_word_length = *c; }
    if (d) { p.has |= TB_GQ_HAS_MAX_AVG_WORD_LENGTH; p.max_avg_word_length = *d; }
    if (e) { p.has |= TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO; p.max_symbol_word_ratio = *e; }
    if (f) { p.has |= TB_GQ_HAS_MAX_BULLET_LINES_RATIO; p.max_bullet_lines_ratio = *f; }
    if (g) { p.has |= TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO; p.max_ellipsis_lines_ratio = *g; }
    if (h) { p.has |= TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO; p.max_non_alpha_words_ratio = *h; }
    if (i) { p.has |= TB_GQ_HAS_MIN_STOP_WORDS; p.min_stop_words = *i; }
    if (sw) {
      p.has |= TB_GQ_HAS_STOP_WORDS;
      s.custom_stop_words = true;
      s.stop_offsets.push_back(0);
      for (const auto& w : *sw) { s.stop_bytes += w; s.stop_offsets.push_back((int32_t)s.stop_bytes.size()); }
      p.n_stop_words = (int32_t)sw->size();
    }
    return s;
  }
};

class GopherRepetitionFilter : public GpuStep {  // GopherRepetitionFilter::new, gopher_rep.rs:25-43
 public:
  using NGrams = std::vector<std::pair<size_t, double>>;
  GopherRepetitionFilter(std::optional<double> dup_line_frac, std::optional<double> dup_para_frac,
                         std::optional<double> dup_line_char_frac, std::optional<double> dup_para_char_frac,
                         const NGrams& top_n_grams, const NGrams& dup_n_grams, int device = 0)
      : GpuStep(make(dup_line_frac, dup_para_frac, dup_line_char_frac, dup_para_char_frac, top_n_grams, dup_n_grams), device) {}
  const char* name() const override { return "GopherRepetitionFilter"; }

 private:
  static StepSpec make(std::optional<double> a, std::optional<double> b, std::optional<double> c, std::optional<double> d,
                       const NGrams& top, const NGrams& dup) {
    if (top.size() > TB_MAX_NGRAMS || dup.size() > TB_MAX_NGRAMS)
      throw Unexpected("GopherRepetitionFilter: at most " + std::to_string(TB_MAX_NGRAMS) + " entries per n-gram list");
    StepSpec s;
    s.cfg.type = TB_STEP_GOPHER_REPETITION;
    tb_gr_params& p = s.cfg.u.gr;
    if (a) { p.has |= TB_GR_HAS_DUP_LINE_FRAC; p.dup_line_frac = *a; }
    if (b) { p.has |= TB_GR_HAS_DUP_PARA_FRAC; p.dup_para_frac = *b; }
    if (c) { p.has |= TB_GR_HAS_DUP_LINE_CHAR_FRAC; p.dup_line_char_frac = *c; }
    if (d) { p.has |= TB_GR_HAS_DUP_PARA_CHAR_FRAC; p.dup_para_char_frac = *d; }
    p.n_top = (int32_t)top.size(); p.n_dup = (int32_t)dup.size();
    for (size_t k = 0; k < top.size(); k++) { p.top_n[k] = (uint32_t)top[k].first; p.top_frac[k] = top[k].second; }
    for (size_t k = 0; k < dup.size(); k++) { p.dup_n[k] = (uint32_t)dup[k].first; p.dup_frac[k] = dup[k].second; }
    return s;
  }
};

class FineWebQualityFilter : public GpuStep {  // FineWebQualityFilter::new, fineweb_quality.rs:41-62
 public:
  FineWebQualityFilter(double line_punct_thr, bool line_punct_exclude_zero, double short_line_thr, size_t short_line_length,
                       double char_duplicates_ratio, double new_line_ratio,
                       std::optional<std::vector<char32_t>> stop_chars = std::nullopt, int device = 0)
      : GpuStep(make(line_punct_thr, line_punct_exclude_zero, short_line_thr, short_line_length, char_duplicates_ratio,
                     new_line_ratio, stop_chars), device) {}
  const char* name() const override { return "FineWebQualityFilter"; }

 private:
  static StepSpec make(double lp, bool ez, double sl, size_t sll, double cd, double nl,
                       const std::optional<std::vector<char32_t>>& sc) {
    StepSpec s;
    s.cfg.type = TB_STEP_FINEWEB_QUALITY;
    tb_fw_params& p = s.cfg.u.fw;
    p.line_punct_thr = lp; p.line_punct_exclude_zero = ez; p.short_line_thr = sl; p.short_line_length = sll;
    p.char_duplicates_ratio = cd; p.new_line_ratio = nl;
    p.n_stop_chars = -1;
    if (sc) {
      if (sc->size() > TB_FW_MAX_STOP_CHARS) throw Unexpected("FineWebQualityFilter: at most " + std::to_string(TB_FW_MAX_STOP_CHARS) + " stop_chars");
      p.n_stop_chars = (int32_t)sc->size();
      for (size_t k = 0; k < sc->size(); k++) p.stop_chars[k] = (uint32_t)(*sc)[k];
    }
    return s;
  }
};

// ---- PipelineExecutor (src/executor.rs:17-70) ------------------------------------------------------------
// All steps are accelerated ones here, so the chain is fused into one device pipeline: one round trip per batch,
// short-circuit at the first failing step, later steps' metadata absent (executor.rs:30-57).
class PipelineExecutor {
 public:
  explicit PipelineExecutor(std::vector<std::shared_ptr<GpuStep>> steps, int device = 0) : steps_(std::move(steps)) {
    std::vector<StepSpec> specs;
    for (auto& s : steps_) { specs.push_back(s->spec()); names_.push_back(s->name()); }
    pipe_ = std::make_unique<GpuPipeline>(std::move(specs), device);
  }
  // run_single_async: Ok(document) or throws StepError{step_name, DocumentFiltered}
  TextDocument run_single(TextDocument document) const {
    auto r = pipe_->run({std::move(document)}, names_);
    if (auto* ok = std::get_if<TextDocument>(&r[0])) return std::move(*ok);
    throw *std::get<std::shared_ptr<StepError>>(r[0]);
  }
  void set_option(int32_t option, int64_t value) { pipe_->set_option(option, value); }
  // run_batch_parallel_async: one Result per input, input order
  std::vector<StepResult> run_batch(const std::vector<TextDocument>& documents) const { return pipe_->run(documents, names_); }

 private:
  std::vector<std::shared_ptr<GpuStep>> steps_;
  std::vector<std::string> names_;
  std::unique_ptr<GpuPipeline> pipe_;
};

}  // namespace textblaster
