// The reference's own unit-test cases driven through the C++ host mirror (include/textblaster_b200.hpp).
// Built and run by tests/test_cpp_mirror.py.  Exit 0 = all passed; exit 3 with "no CUDA device" on stderr when
// the library refuses to start without a GPU (the CPU test expects exactly that); exit 1 = a check failed.
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>

#include "textblaster_b200.hpp"

using namespace textblaster;

static int g_fail = 0;
#define CHECK(cond)                                                            \
  do {                                                                         \
    if (!(cond)) { fprintf(stderr, "FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); g_fail++; } \
  } while (0)

static TextDocument doc(const std::string& id, const std::string& content) {
  TextDocument d;
  d.id = id; d.content = content; d.source = "test_source";
  return d;
}
static bool has(const std::string& s, const std::string& sub) { return s.find(sub) != std::string::npos; }

// runs the step and hands the DocumentFiltered error to `on_filtered`; a pass is a failure of the check
static void expect_filtered(const ProcessingStep& f, TextDocument d, const std::function<void(const DocumentFiltered&)>& on_filtered) {
  try {
    f.process(std::move(d));
    fprintf(stderr, "FAIL: %s passed a document it has to filter\n", f.name());
    g_fail++;
  } catch (const DocumentFiltered& e) {
    on_filtered(e);
  }
}

static void test_c4() {  // c4_filters.rs:594-790
  C4QualityFilter f(true, true, true, 5, 3, 1000, true, true, true, true);
  CHECK(std::string(f.name()) == "C4QualityFilter");
  const std::string good =
      "This is the first sentence. This is the second sentence. This is the third sentence. This is the fourth sentence. "
      "This is the fifth sentence.";
  TextDocument out = f.process(doc("pass", good));
  CHECK(out.content == good);
  CHECK(out.metadata["c4_filter_status"] == "passed");
  CHECK(out.source == "test_source" && out.id == "pass");
  expect_filtered(f, doc("fail_sentences", "One sentence. Two sentences. Three sentences. Four sentences."), [](const DocumentFiltered& e) {
    CHECK(has(e.reason, "too_few_sentences (found 4, required 5)"));
    CHECK(e.document.metadata.at("c4_filter_status") == "filtered");
    CHECK(e.document.id == "fail_sentences");
    CHECK(has(e.what(), "Document 'fail_sentences' filtered out: "));
  });
  expect_filtered(f, doc("lorem", "This is fine. Lorem ipsum dolor sit amet. This is also fine."), [](const DocumentFiltered& e) {
    CHECK(e.reason == "lorem_ipsum");
    CHECK(e.document.content.rfind("This is fine. Lorem", 0) == 0);
  });
  expect_filtered(f, doc("curly", "This is fine.\nSome code block {}.\nAnother good line."),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "curly_bracket")); });
  // line rules rewrite the content: the javascript and the policy line go, citations are removed
  C4QualityFilter g(true, true, true, 1, 3, 1000, true, true, true, true);
  out = g.process(doc("rewrite", "This line is kept[1] as it is.\nPlease enable javascript to view this.\nWe use cookies on this site.\nShort one."));
  CHECK(out.content == "This line is kept as it is.");
  // split_paragraph = false: the lines are the sentences (c4_filters.rs:152-155); "Yes." has too few words
  C4QualityFilter h(false, true, true, 1, 3, 1000, true, true, true, true);
  out = h.process(doc("sentences", "Stop right now please! Go there now? Yes."));
  CHECK(out.content == "Stop right now please!\nGo there now?");
}

static void test_gopher_quality() {  // gopher_quality.rs:360-829
  const auto none = std::nullopt;
  GopherQualityFilter f(3, none, none, none, none, none, none, none, none, none);
  CHECK(std::string(f.name()) == "GopherQualityFilter");
  CHECK(f.process(doc("a", "Hello world test . !")).metadata["gopher_quality_filter_status"] == "passed");
  expect_filtered(f, doc("b", "Hello world . !"),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "gopher_short_doc (2 non-symbol words, required 3)")); });
  GopherQualityFilter avg(none, none, 3.0, 5.0, none, none, none, none, none, none);
  expect_filtered(avg, doc("c", ". ! ."), [](const DocumentFiltered& e) {
    CHECK(has(e.reason, "gopher_below_avg_threshold (avg len 0.00, required 3.00 - 0 non-symbol words)"));
  });
  GopherQualityFilter sw(none, none, none, none, none, none, none, none, 1, std::vector<std::string>{"custom", "words"});
  sw.process(doc("d", "this is a custom test with other words"));
  expect_filtered(sw, doc("e", "this is a regular sentence"),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "gopher_too_few_stop_words (found 0, required 1)")); });
}

static void test_gopher_repetition() {  // gopher_rep.rs:412-642
  const auto none = std::nullopt;
  GopherRepetitionFilter f(0.3, none, none, none, {}, {});
  CHECK(std::string(f.name()) == "GopherRepetitionFilter");
  expect_filtered(f, doc("l", "This is line one.\nThis is line two.\nThis is line one."),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "dup_line_frac (ratio 0.33, max 0.30)")); });
  GopherRepetitionFilter top(none, none, none, none, {{2, 0.3}}, {});
  top.process(doc("p", "a b c d e f a b g h i j"));
  expect_filtered(top, doc("q", "a b c a b d a b e a b"),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "top_2_gram (ratio 0.57, max 0.30)")); });
  GopherRepetitionFilter dup(none, none, none, none, {}, {{2, 0.1}});
  expect_filtered(dup, doc("r", "a b c d e a b f g"),
                  [](const DocumentFiltered& e) { CHECK(has(e.reason, "duplicated_2_n_grams (ratio 0.12, max 0.10)")); });
  GopherRepetitionFilter empty(none, none, none, none, {}, {});
  expect_filtered(empty, doc("s", "   "), [](const DocumentFiltered& e) {
    CHECK(e.reason == "skipping empty content");
    CHECK(e.document.metadata.at("gopher_repetition_filter_reason") == "skipping empty content");
  });
}

static void test_fineweb() {  // fineweb_quality.rs:230-604
  FineWebQualityFilter f(0.12, false, 0.67, 30, 0.95, 0.3);
  CHECK(std::string(f.name()) == "FineWebQualityFilter");
  for (const char* t : {"", "   \n\t   \n ", "\n\n\n"})
    expect_filtered(f, doc("empty", t), [](const DocumentFiltered& e) {
      CHECK(e.reason == "empty");
      CHECK(e.document.metadata.at("fineweb_filter_reason") == "empty document");
    });
  expect_filtered(f, doc("punct", "Line one\nLine two\nLine three\nLine four\nLine five\nLine six\nLine seven\nLine eight\nLine nine\nLine ten."),
                  [](const DocumentFiltered& e) {
                    CHECK(e.reason.rfind("line_punct_ratio: 0.1000 < threshold 0.1200", 0) == 0);
                    CHECK(has(e.reason, "(exclude_zero: false)"));
                  });
  FineWebQualityFilter list(0.0, false, 1.0, 30, 0.95, 0.3);
  expect_filtered(list, doc("list", "word.\nword.\nword.\nword.\nword."),
                  [](const DocumentFiltered& e) { CHECK(e.reason.rfind("list_ratio: 0.8000 > threshold 0.3000", 0) == 0); });
  expect_filtered(list, doc("nowords", "!!!\n???"),
                  [](const DocumentFiltered& e) { CHECK(e.reason == "list_ratio_no_words (newlines present but no words)"); });
  TextDocument out = f.process(doc("ok", "Many words on a single line with no newlines effectively. This should pass easily."));
  CHECK(out.metadata.empty());  // nothing is written on a pass (:225)
}

static void test_executor() {  // executor.rs:30-70
  const auto none = std::nullopt;
  std::vector<std::shared_ptr<GpuStep>> steps = {
      std::make_shared<C4QualityFilter>(true, true, true, 5, 3, 1000, true, true, true, true),
      std::make_shared<GopherQualityFilter>(20, none, none, none, none, none, none, none, none, none),
      std::make_shared<GopherRepetitionFilter>(0.3, none, none, none, GopherRepetitionFilter::NGrams{}, GopherRepetitionFilter::NGrams{})};
  PipelineExecutor ex(steps);
  const std::string good =
      "This is the first sentence. This is the second sentence. This is the third sentence. This is the fourth sentence. "
      "This is the fifth sentence.";
  std::vector<TextDocument> docs = {doc("0", "Some code block {}."), doc("1", "One. Two. Three. Four. Five sentences are here now."),
                                    doc("2", good)};
  auto out = ex.run_batch(docs);
  CHECK(out.size() == 3);
  auto* e0 = std::get_if<std::shared_ptr<StepError>>(&out[0]);
  auto* e1 = std::get_if<std::shared_ptr<StepError>>(&out[1]);
  auto* ok2 = std::get_if<TextDocument>(&out[2]);
  CHECK(e0 && (*e0)->step_name == "C4QualityFilter");
  CHECK(e1 && (*e1)->step_name == "GopherQualityFilter");
  if (e0) CHECK(!(*e0)->source->document.metadata.count("gopher_quality_filter_status"));  // short-circuit
  if (e1) {
    CHECK((*e1)->source->document.metadata.at("c4_filter_status") == "passed");
    CHECK(!(*e1)->source->document.metadata.count("gopher_repetition_filter_status"));
    CHECK(has((*e1)->what(), "Error in processing step 'GopherQualityFilter': Document '1' filtered out: gopher_short_doc"));
  }
  CHECK(ok2 && ok2->id == "2" && ok2->metadata.size() == 3);
  // run_single: the document, or the StepError as the exception
  CHECK(ex.run_single(doc("2", good)).metadata.at("gopher_repetition_filter_status") == "passed");
  try { ex.run_single(doc("0", "Some code block {}.")); CHECK(!"run_single has to throw"); }
  catch (const StepError& e) { CHECK(e.step_name == "C4QualityFilter" && has(e.source->reason, "curly_bracket")); }
  // a reader that hands the raw column over lets the pipeline decode the entities first (parquet_reader.rs:178-179)
  ex.set_option(TB_OPT_DECODE_HTML_ENTITIES, 1);
  const std::string raw = "Tom &amp; Jerry wrote the first sentence. This is the second &#115;entence. This is the third sentence. "
                          "This is the fourth sentence. This is the fifth sentence.";
  TextDocument dec = ex.run_single(doc("3", raw));
  CHECK(dec.content.rfind("Tom & Jerry wrote", 0) == 0 && has(dec.content, "second sentence."));
}

static void test_config_errors() {  // src/config/pipeline.rs validate(): rejected before any document runs
  try { GopherRepetitionFilter f(1.5, std::nullopt, std::nullopt, std::nullopt, {}, {}); CHECK(!"1.5 is not a fraction"); }
  catch (const ConfigValidationError& e) { CHECK(has(e.message, "dup_line_frac must be between 0.0 and 1.0, got 1.5")); }
}

int main() {
  try {
    test_config_errors();
    test_c4();
    test_gopher_quality();
    test_gopher_repetition();
    test_fineweb();
    test_executor();
  } catch (const Unexpected& e) {
    fprintf(stderr, "Unexpected: %s\n", e.message.c_str());
    return 3;
  } catch (const std::exception& e) {
    fprintf(stderr, "exception: %s\n", e.what());
    return 1;
  }
  if (g_fail) { fprintf(stderr, "%d check(s) failed\n", g_fail); return 1; }
  printf("mirror: all checks passed\n");
  return 0;
}
