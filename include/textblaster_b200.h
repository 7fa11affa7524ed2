/*
 * textblaster_b200.h -- C ABI of the B200-native TextBlaster quality-filter hot path.
 *
 * The reference (kris927b/TextBlaster, Rust) has no FFI today: its boundary for this path is the
 * trait object `ProcessingStep` (src/executor.rs:8-15) built by `build_pipeline_from_config`
 * (src/worker_logic.rs:39-134) and driven by `PipelineExecutor::run_single_async`
 * (src/executor.rs:30-57).  This header is what a `-sys` crate over the CUDA library would bind;
 * every entry point cites the reference interface it stands in for.  Plain pointers and sizes only,
 * no torch / C++ types.  INTEGRATION.md shows the Rust (and ctypes) binding.
 *
 * Semantics mirrored from the reference:
 *   - a batch is one Arrow string column: contiguous UTF-8 `values` + n+1 offsets (i32 as in
 *     arrow::StringArray, src/pipeline/readers/parquet_reader.rs:108-112, or i64);
 *   - steps run in the configured order and a document stops at the first step that filters it
 *     (src/executor.rs:32-46); statistics of later steps are undefined for that document;
 *   - "filtered" is data (keep = 0, failed_step, reason bits + counters), not an error;
 *   - C4QualityFilter rewrites the content (src/pipeline/filters/c4_filters.rs:254-258); the
 *     content after the chain is returned in `content` / `content_offsets`.
 *
 * All integer counters are bit-exact against the reference algorithm; ratios are single IEEE-754
 * double divisions of those counters (helpers at the end of this header).
 */
#ifndef TEXTBLASTER_B200_H
#define TEXTBLASTER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TB_ABI_VERSION 3
#define TB_MAX_STEPS 8
#define TB_MAX_NGRAMS 12 /* entries per top_n_grams / dup_n_grams list */
#define TB_MAX_NGRAM_N 32 /* largest n-gram order accepted */

/* ---- status codes (tb_last_error() has the text) ---------------------------------------- */
enum {
  TB_OK = 0,
  TB_ERR_INVALID_ARGUMENT = 1,
  TB_ERR_CONFIG_VALIDATION = 2, /* PipelineError::ConfigValidationError, src/error.rs:62-63 */
  TB_ERR_CUDA = 3,
  TB_ERR_OUT_OF_MEMORY = 4,
  TB_ERR_UNSUPPORTED = 5,
  TB_ERR_INVALID_UTF8 = 6 /* the reference only ever sees Rust `String`s (valid UTF-8) */
};

/* ---- step configuration: src/config/pipeline.rs:26-37 (enum StepConfig, the three in-scope
 *      variants) ------------------------------------------------------------------------- */
typedef enum {
  TB_STEP_C4_QUALITY = 1,        /* StepConfig::C4QualityFilter        */
  TB_STEP_GOPHER_QUALITY = 2,    /* StepConfig::GopherQualityFilter    */
  TB_STEP_GOPHER_REPETITION = 3, /* StepConfig::GopherRepetitionFilter */
  TB_STEP_FINEWEB_QUALITY = 4    /* StepConfig::FineWebQualityFilter */
} tb_step_type;

/* C4QualityParams, src/config/pipeline.rs:66-79; field meaning src/pipeline/filters/c4_filters.rs:98-139 */
typedef struct tb_c4_params {
  int32_t split_paragraph;          /* 1: lines = str::lines(); 0: lines = split_into_sentences() (c4_filters.rs:150-155) */
  int32_t remove_citations;
  int32_t filter_no_terminal_punct;
  int32_t filter_lorem_ipsum;
  int32_t filter_javascript;
  int32_t filter_curly_bracket;
  int32_t filter_policy;
  int32_t _pad;
  uint64_t min_num_sentences;
  uint64_t min_words_per_line;
  uint64_t max_word_length;
} tb_c4_params;

/* GopherQualityParams, src/config/pipeline.rs:161-176.  `has` bit i set <=> the i-th Option is Some. */
enum {
  TB_GQ_HAS_MIN_DOC_WORDS = 1 << 0,
  TB_GQ_HAS_MAX_DOC_WORDS = 1 << 1,
  TB_GQ_HAS_MIN_AVG_WORD_LENGTH = 1 << 2,
  TB_GQ_HAS_MAX_AVG_WORD_LENGTH = 1 << 3,
  TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO = 1 << 4,
  TB_GQ_HAS_MAX_BULLET_LINES_RATIO = 1 << 5,
  TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO = 1 << 6,
  TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO = 1 << 7,
  TB_GQ_HAS_MIN_STOP_WORDS = 1 << 8,
  TB_GQ_HAS_STOP_WORDS = 1 << 9 /* else DEFAULT_STOP_WORDS, gopher_quality.rs:10 */
};
typedef struct tb_gq_params {
  uint32_t has;
  int32_t n_stop_words;
  uint64_t min_doc_words;
  uint64_t max_doc_words;
  double min_avg_word_length;
  double max_avg_word_length;
  double max_symbol_word_ratio;
  double max_bullet_lines_ratio;
  double max_ellipsis_lines_ratio;
  double max_non_alpha_words_ratio;
  uint64_t min_stop_words;
  const uint8_t* stop_words_utf8;     /* concatenated entries, used as given (not lower-cased) */
  const int32_t* stop_words_offsets;  /* n_stop_words + 1 */
} tb_gq_params;

/* GopherRepetitionParams, src/config/pipeline.rs:102-115 */
enum {
  TB_GR_HAS_DUP_LINE_FRAC = 1 << 0,
  TB_GR_HAS_DUP_PARA_FRAC = 1 << 1,
  TB_GR_HAS_DUP_LINE_CHAR_FRAC = 1 << 2,
  TB_GR_HAS_DUP_PARA_CHAR_FRAC = 1 << 3
};
typedef struct tb_gr_params {
  uint32_t has;
  int32_t n_top;
  int32_t n_dup;
  int32_t _pad;
  double dup_line_frac;
  double dup_para_frac;
  double dup_line_char_frac;
  double dup_para_char_frac;
  uint32_t top_n[TB_MAX_NGRAMS];
  double top_frac[TB_MAX_NGRAMS];
  uint32_t dup_n[TB_MAX_NGRAMS];
  double dup_frac[TB_MAX_NGRAMS];
} tb_gr_params;

/* FineWebQualityFilterParams, src/config/pipeline.rs:311-322; use src/pipeline/filters/fineweb_quality.rs:41-62 */
#define TB_FW_MAX_STOP_CHARS 32
typedef struct tb_fw_params {
  double line_punct_thr;
  int32_t line_punct_exclude_zero;
  int32_t n_stop_chars; /* < 0: stop_chars = None -> END_PUNCTUATION . ! ? " ' U+201D (fineweb_quality.rs:26,51) */
  uint32_t stop_chars[TB_FW_MAX_STOP_CHARS]; /* code points */
  double short_line_thr;
  uint64_t short_line_length;
  double char_duplicates_ratio;
  double new_line_ratio;
} tb_fw_params;

typedef struct tb_step_cfg {
  int32_t type; /* tb_step_type */
  int32_t _pad;
  union {
    tb_c4_params c4;
    tb_gq_params gq;
    tb_gr_params gr;
    tb_fw_params fw;
  } u;
} tb_step_cfg;

/* ---- per-document results ------------------------------------------------------------------
 * One array of the matching struct per configured step.  reason bits are in the order the
 * reference pushes reasons, so "; ".join over set bits reproduces its reason string. */

/* c4_filters.rs:166-173 (doc-level), :207-240 (line counters -> metadata keys line-filter-*), :261-269 */
enum {
  TB_C4_REASON_LOREM_IPSUM = 1 << 0,
  TB_C4_REASON_CURLY_BRACKET = 1 << 1,
  TB_C4_REASON_TOO_FEW_SENTENCES = 1 << 2
};
typedef struct tb_c4_stats {
  uint32_t reason_bits;
  uint32_t n_lines;              /* lines examined (0 on the early doc-level exit) */
  uint32_t too_long_word;        /* "line-filter-too_long_word"    */
  uint32_t no_terminal_punc;     /* "line-filter-no_terminal_punc" */
  uint32_t too_few_words;        /* "line-filter-too_few_words"    */
  uint32_t n_kept_lines;
  uint32_t n_sentences;          /* split_into_sentences(rewritten content).len() */
  uint32_t rewritten;            /* 1 if content was rewritten (line filtering ran) */
} tb_c4_stats;

/* gopher_quality.rs:198-295, in push order */
enum {
  TB_GQ_REASON_SHORT_DOC = 1 << 0,
  TB_GQ_REASON_LONG_DOC = 1 << 1,
  TB_GQ_REASON_BELOW_AVG = 1 << 2,
  TB_GQ_REASON_ABOVE_AVG = 1 << 3,
  TB_GQ_REASON_TOO_MANY_HASHES = 1 << 4,
  TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS = 1 << 5,
  TB_GQ_REASON_TOO_MANY_BULLETS = 1 << 6,
  TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES = 1 << 7,
  TB_GQ_REASON_BELOW_ALPHA = 1 << 8,
  TB_GQ_REASON_TOO_FEW_STOP_WORDS = 1 << 9
};
typedef struct tb_gq_stats {
  uint32_t reason_bits;
  uint32_t n_words;             /* n_total_words == n_non_symbol_words, gopher_quality.rs:74,85 */
  uint64_t sum_word_chars;      /* :92 */
  uint32_t n_hash;              /* :104 */
  uint32_t n_ellipsis_units;    /* :115 */
  uint32_t n_lines;             /* :127 */
  uint32_t n_bullet_lines;      /* :134-141 */
  uint32_t n_ellipsis_lines;    /* :152-158 */
  uint32_t n_alpha_words;       /* :169-172 */
  uint32_t n_stop_words;        /* :184-187 */
  uint32_t flags;               /* TB_DOCFLAG_* */
} tb_gq_stats;

/* gopher_rep.rs:95-195, in push order: 4 fixed bits, then top_n_grams[i], then dup_n_grams[i] */
enum {
  TB_GR_REASON_DUP_PARA_FRAC = 1 << 0,
  TB_GR_REASON_DUP_PARA_CHAR_FRAC = 1 << 1,
  TB_GR_REASON_DUP_LINE_FRAC = 1 << 2,
  TB_GR_REASON_DUP_LINE_CHAR_FRAC = 1 << 3,
  TB_GR_REASON_TOP_SHIFT = 4,                   /* bit 4+i  : top_n_grams[i] */
  TB_GR_REASON_DUP_SHIFT = 4 + TB_MAX_NGRAMS,   /* bit 16+i : dup_n_grams[i] */
  TB_GR_REASON_EMPTY = 1u << 31                 /* "skipping empty content", gopher_rep.rs:60-74 */
};
typedef struct tb_gr_stats {
  uint32_t reason_bits;
  uint32_t flags;               /* TB_DOCFLAG_* */
  uint64_t char_len;            /* trimmed.chars().count(), gopher_rep.rs:58 (before .max(1)) */
  uint32_t n_paras;
  uint32_t para_dup_elems;
  uint64_t para_dup_bytes;
  uint32_t n_lines;
  uint32_t line_dup_elems;
  uint64_t line_dup_bytes;
  uint64_t n_words;
  uint64_t top_bytes[TB_MAX_NGRAMS];  /* find_top_duplicate, text.rs:211-238 */
  uint64_t dup_bytes[TB_MAX_NGRAMS];  /* find_all_duplicate, text.rs:241-259 */
} tb_gr_stats;

/* fineweb_quality.rs:79-222: the checks return at the first failure, so exactly one bit is set */
enum {
  TB_FW_REASON_EMPTY = 1 << 0,          /* no non-blank line, :79-90 (reason "empty") */
  TB_FW_REASON_LINE_PUNCT = 1 << 1,     /* :106-123 */
  TB_FW_REASON_SHORT_LINE = 1 << 2,     /* :126-146 */
  TB_FW_REASON_CHAR_DUP = 1 << 3,       /* :148-185 */
  TB_FW_REASON_LIST_NO_WORDS = 1 << 4,  /* :191-204 */
  TB_FW_REASON_LIST_RATIO = 1 << 5      /* :205-222 */
};
typedef struct tb_fw_stats {
  uint32_t reason_bits;
  uint32_t flags;               /* TB_DOCFLAG_* */
  uint32_t n_lines;             /* non-blank lines of content.lines(), :74-77 */
  uint32_t n_stop_lines;        /* :93-104 */
  uint32_t n_short_lines;       /* :126-129 */
  uint32_t n_words;             /* split_into_words(content).len(), :188 */
  uint32_t n_newlines;          /* content.matches('\n').count(), :189 */
  uint32_t _pad;
  uint64_t dup_line_bytes;      /* find_duplicates(non-blank lines).1, :162 */
  uint64_t n_chars_no_newline;  /* :158-159 */
} tb_fw_stats;

/* Document flags (information, never affects decisions) */
enum {
  TB_DOCFLAG_COMPLEX_SCRIPT = 1 << 0, /* has Han/Kana/Thai/Lao/Khmer/Burmese text: ICU4X new_auto()
                                         (text.rs:107) would re-segment it with dictionary/LSTM data;
                                         we apply plain UAX #29 -- parity unpinned for such documents */
  TB_DOCFLAG_BAD_OFFSETS = 1 << 1     /* tb_gopher_repetition_noalloc only: the document's offsets were decreasing
                                         or outside [0, n_bytes]; it was not processed (keep = 0, counters zero) */
};

/* One run of a document's content: the bytes [begin, end) of the document's own input text (TB_OPT_RETURN_CONTENT = 2) */
typedef struct tb_run {
  uint32_t begin, end;
} tb_run;

typedef struct tb_batch_out {
  int64_t n_docs;
  int32_t n_steps;
  int32_t content_rewritten;       /* 1 if content/content_offsets are valid (a C4 step or the entity pre-pass ran, content returned as bytes) */
  uint8_t* keep;                   /* [n_docs] 1 = passed every step */
  int8_t* failed_step;             /* [n_docs] index into the pipeline, -1 = passed */
  void* step_stats[TB_MAX_STEPS];  /* [n_docs] of tb_c4_stats / tb_gq_stats / tb_gr_stats */
  uint8_t* content;                /* content after the chain (all docs), or NULL */
  int64_t* content_offsets;        /* [n_docs + 1] */
  int64_t content_bytes;
  double gpu_ms;                   /* device time of the kernels of this call (CUDA events) */
  int32_t gpu_launches;            /* kernels launched by this call */
  int32_t content_kind;            /* 0: no content returned, 1: content / content_offsets, 2: runs / run_offsets */
  void* _owner;                    /* internal */
  /* TB_OPT_RETURN_CONTENT = 2: the content of document d is the concatenation of the runs
   * runs[run_offsets[d] .. run_offsets[d + 1]) of its input text (tb_gather_content copies them out) */
  const tb_run* runs;
  const int64_t* run_offsets;      /* [n_docs + 1] */
  int64_t n_runs;
  /* TB_OPT_GATHER_KEPT: the content of the documents that passed every step, back to back in input order, and its
   * n_kept + 1 offsets -- the text column of the output file (src/producer_logic.rs:157-167), built by host threads
   * from the runs and the caller's input while the device works on the following chunks */
  const uint8_t* kept_content;
  const int64_t* kept_offsets;
  int64_t n_kept;
} tb_batch_out;

typedef struct tb_ctx tb_ctx;
typedef struct tb_pipeline tb_pipeline;

/* Creates the per-device context (streams, Unicode tables in HBM).  Replaces nothing in the
 * reference (it has no device); lifetime = the worker's `Arc<PipelineExecutor>` (src/bin/worker.rs:149).
 * Threads: the reference shares one executor between all worker tasks (ProcessingStep: Send + Sync,
 * src/bin/worker.rs:149, src/worker_logic.rs:260-264).  Here every entry point may be called from any host thread;
 * calls on one tb_pipeline are serialised by the pipeline (it owns one set of device buffers), pipelines of one
 * context may run at the same time (their kernels share the context's streams), and results do not depend on
 * either. */
int tb_ctx_create(int device, tb_ctx** out);
void tb_ctx_destroy(tb_ctx* ctx);
/* Thread-local text of the last failure on this thread (ctx may be NULL). */
const char* tb_last_error(const tb_ctx* ctx);
int tb_abi_version(void);

/* build_pipeline_from_config (src/worker_logic.rs:39-134) + StepConfig::validate
 * (src/config/pipeline.rs:53-63): validates with the reference's rules and messages, uploads
 * parameters.  The handle is immutable afterwards (ProcessingStep: Send + Sync). */
int tb_pipeline_create(tb_ctx* ctx, const tb_step_cfg* steps, size_t n_steps, tb_pipeline** out);
void tb_pipeline_destroy(tb_pipeline* p);
/* ProcessingStep::name (c4_filters.rs:143-145, gopher_quality.rs:65-67, gopher_rep.rs:48-50) */
const char* tb_step_name(int32_t step_type);

/* PipelineExecutor::run_batch_parallel_async (src/executor.rs:60-70) for a whole Arrow string
 * column held in HOST memory: pinned double-buffered H2D on a side stream, kernels, D2H of results.
 * offsets: n_docs + 1 entries of int32 (offsets_are_i64 == 0) or int64.  Results stay valid until
 * tb_batch_out_release.  Result order == input order (tests/executor_test.rs:284-291). */
int tb_pipeline_run_batch(tb_pipeline* p, const uint8_t* utf8, const void* offsets,
                          int offsets_are_i64, int64_t n_docs, tb_batch_out* out);

/* Same, inputs already resident in device memory (utf8_dev, int64 offsets_dev); results are copied
 * to host.  Used by bench.py's device-resident `value` and by callers that decode Parquet on GPU. */
int tb_pipeline_run_batch_device(tb_pipeline* p, const uint8_t* utf8_dev, const int64_t* offsets_dev,
                                 int64_t n_docs, int64_t n_bytes, tb_batch_out* out);

void tb_batch_out_release(tb_pipeline* p, tb_batch_out* out);

/* The content of the selected documents of a batch, gathered on the host from the runs of a TB_OPT_RETURN_CONTENT = 2
 * result and the caller's input column (the same utf8 / offsets that went into tb_pipeline_run_batch): what the
 * reference's aggregator does when it pushes `document.content` into the output batch (src/producer_logic.rs:157-167,
 * src/pipeline/writers/parquet_writer.rs:69-157).  select[d] != 0 picks document d (NULL: every document); the
 * picked documents' contents go to dst back to back, dst_offsets gets their n_selected + 1 offsets.  dst may be
 * NULL to learn the sizes only (*dst_bytes, dst_offsets).  nthreads <= 0: one thread per 64 MiB, at most 8.
 * Pure host code: no CUDA call, any thread. */
int tb_gather_content(const tb_batch_out* out, const uint8_t* utf8, const void* offsets, int offsets_are_i64, const uint8_t* select,
                      uint8_t* dst, int64_t dst_capacity, int64_t* dst_offsets, int64_t* dst_bytes, int nthreads);

/* Options (set before running; default in brackets) */
enum {
  TB_OPT_RETURN_CONTENT = 1, /* [1] how the content after the chain comes back.  1: as bytes (content / content_offsets).  0: not at all.
                                2: as runs of the caller's own text (runs / run_offsets) -- C4's output is a subsequence of its input
                                (c4_filters.rs:192-258: lines trimmed, citations cut, kept lines joined by the '\n' in front of them), so
                                a few (begin, end) pairs per document replace the bytes on the device->host link, and the caller (or
                                tb_gather_content) copies the runs of the documents it writes out straight from its input column.  Falls
                                back to 1 where the text C4 reads is not the caller's (entity pre-pass, split_paragraph = false, a chain with
                                several C4 steps) */
  TB_OPT_VALIDATE_UTF8 = 2,  /* [0] check inputs on the device, TB_ERR_INVALID_UTF8 on failure */
  TB_OPT_CHUNK_BYTES = 3,    /* [1 GiB] largest chunk the batch is processed in (device scratch: ~45 bytes per byte of it) */
  TB_OPT_GENERAL_WORD_PATH = 4, /* [0] testing: run the segment-walking word passes for every document (same results, slower) */
  TB_OPT_DECODE_HTML_ENTITIES = 5, /* [0] decode HTML entities before the first step, as the reference's reader does with
                                     html_escape::decode_html_entities (src/pipeline/readers/parquet_reader.rs:178-179): pass the
                                     Arrow buffers of the text column as they are.  The decoded text is the document's content
                                     (content / content_offsets are then returned even without a C4 step).  Parity with the
                                     html-escape crate is unpinned, see textblaster_b200/csrc/tb_entities.cuh */
  TB_OPT_CHUNK_RAMP = 7,        /* [1] chunk sizes ramp up at the start of a batch and down at its end, which hides the first copy to the
                                     device and the last copy back behind kernels (one batch at a time).  0: every chunk has the full
                                     size -- for callers that keep several pipelines of one context busy with a stream of batches
                                     (textblaster_b200/runner.py), where another batch's kernels cover the copies */
  TB_OPT_STAGE_THREADS = 8,     /* [4] text in pageable host memory (what Arrow allocates, src/pipeline/readers/parquet_reader.rs:103-112) is
                                     staged by the library: this many host threads copy it piece by piece into a ring of pinned slots
                                     (3 x 64 MiB) from which it goes to the device, instead of the driver's slow path for pageable
                                     copies.  Pinned or registered memory is detected and copied from where it is.  0: never stage */
  TB_OPT_BULK_STAGE_TILES = 9,  /* [0] documents of at least this many 1 KiB tiles have their text brought into shared memory by the copy engine
                                     (cp.async.bulk + mbarrier, a two-tile ring per warp) instead of per-lane vector loads in the
                                     GopherQuality, repetition-scan and C4 A / B passes.  0: off -- measured slower on B200 for this chain
                                     (the passes re-read bytes of the tile through L1, which a bulk copy bypasses) */
  TB_OPT_GATHER_KEPT = 6        /* [0] with TB_OPT_RETURN_CONTENT = 2 and host input: number of host threads that copy the content of the
                                     kept documents out of the caller's input column (kept_content / kept_offsets of tb_batch_out), chunk
                                     by chunk behind the device; 0: the caller gathers what it needs itself (tb_gather_content) */
};
int tb_pipeline_set_option(tb_pipeline* p, int32_t option, int64_t value);

/* Name of the dominant kernel and its accumulated device time since pipeline creation; used by
 * bench.py for the roofline line. */
int tb_pipeline_kernel_times(tb_pipeline* p, int32_t max_entries, const char** names, double* ms,
                             int64_t* launches);
void tb_pipeline_reset_kernel_times(tb_pipeline* p);

/* GopherRepetitionFilter for one small batch WITHOUT any allocation -- no counterpart in the reference.
 *
 * For embedders that may not enter the CUDA allocator on the calling thread: processes whose device
 * memory is handed out once by a pool that another component owns (fixed per-tenant quota, a host
 * runtime that pre-partitions the GPU), or latency-critical sections where cudaMalloc (which can
 * synchronise the device) is forbidden.  tb_pipeline_run_batch* grow their buffers with cudaMalloc
 * on demand; this call takes every intermediate from a fixed 24 MiB array inside the library image
 * (resident on every GPU that loads the library) and therefore accepts at most TB_NOALLOC_MAX_DOCS
 * documents and TB_NOALLOC_MAX_BYTES bytes of text per call (TB_ERR_INVALID_ARGUMENT beyond that).
 *
 * All pointers are DEVICE pointers owned by the caller: utf8_dev / offsets_dev (n_docs + 1 int64,
 * non-decreasing, 0 <= offsets <= n_bytes; the offsets cannot be checked on the host without a
 * synchronisation, so a document whose offsets break this is skipped on the device: keep = 0 and
 * flags = TB_DOCFLAG_BAD_OFFSETS) in; stats_dev[n_docs] and keep_dev[n_docs] (1 = passed) out.  Work
 * is enqueued on cuda_stream (a cudaStream_t, NULL = default stream) and not synchronised.  The first
 * call on a device performs a blocking set-up (address lookup of the array, table upload): make it at
 * start-up, before any stream capture.  All calls share the array: calls whose GPU work overlaps race.
 * Results equal those of a one-step GopherRepetitionFilter pipeline.  err/err_len (optional) receive
 * the message of a failure. */
#define TB_NOALLOC_MAX_DOCS 2048
#define TB_NOALLOC_MAX_BYTES (512 * 1024)
int tb_gopher_repetition_noalloc(const uint8_t* utf8_dev, const int64_t* offsets_dev, int64_t n_docs, int64_t n_bytes,
                                 const tb_gr_params* params, tb_gr_stats* stats_dev, uint8_t* keep_dev, void* cuda_stream,
                                 char* err, size_t err_len);

#ifdef __cplusplus
}
#endif
#endif /* TEXTBLASTER_B200_H */
