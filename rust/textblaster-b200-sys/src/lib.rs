//! Raw bindings of `include/textblaster_b200.h` (TB_ABI_VERSION 3), one item per declaration of the header.
//!
//! NOT COMPILED IN THIS REPOSITORY: the build image has no Rust toolchain.  The struct layouts are
//! asserted below against the sizes `tests/test_abi.py` measures for the C structs and the ctypes mirror
//! (`textblaster_b200/_abi.py`), so a drift in the header shows up as a compile error here.
#![allow(non_camel_case_types, non_upper_case_globals)]

use std::os::raw::{c_char, c_int, c_void};

pub const TB_ABI_VERSION: c_int = 3;
pub const TB_MAX_STEPS: usize = 8;
pub const TB_MAX_NGRAMS: usize = 12;
pub const TB_MAX_NGRAM_N: u32 = 32;
pub const TB_FW_MAX_STOP_CHARS: usize = 32;

// ---- status codes
pub const TB_OK: c_int = 0;
pub const TB_ERR_INVALID_ARGUMENT: c_int = 1;
pub const TB_ERR_CONFIG_VALIDATION: c_int = 2;
pub const TB_ERR_CUDA: c_int = 3;
pub const TB_ERR_OUT_OF_MEMORY: c_int = 4;
pub const TB_ERR_UNSUPPORTED: c_int = 5;
pub const TB_ERR_INVALID_UTF8: c_int = 6;

// ---- tb_step_type
pub const TB_STEP_C4_QUALITY: i32 = 1;
pub const TB_STEP_GOPHER_QUALITY: i32 = 2;
pub const TB_STEP_GOPHER_REPETITION: i32 = 3;
pub const TB_STEP_FINEWEB_QUALITY: i32 = 4;

// ---- `has` bits of tb_gq_params / tb_gr_params
pub const TB_GQ_HAS_MIN_DOC_WORDS: u32 = 1 << 0;
pub const TB_GQ_HAS_MAX_DOC_WORDS: u32 = 1 << 1;
pub const TB_GQ_HAS_MIN_AVG_WORD_LENGTH: u32 = 1 << 2;
pub const TB_GQ_HAS_MAX_AVG_WORD_LENGTH: u32 = 1 << 3;
pub const TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO: u32 = 1 << 4;
pub const TB_GQ_HAS_MAX_BULLET_LINES_RATIO: u32 = 1 << 5;
pub const TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO: u32 = 1 << 6;
pub const TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO: u32 = 1 << 7;
pub const TB_GQ_HAS_MIN_STOP_WORDS: u32 = 1 << 8;
pub const TB_GQ_HAS_STOP_WORDS: u32 = 1 << 9;
pub const TB_GR_HAS_DUP_LINE_FRAC: u32 = 1 << 0;
pub const TB_GR_HAS_DUP_PARA_FRAC: u32 = 1 << 1;
pub const TB_GR_HAS_DUP_LINE_CHAR_FRAC: u32 = 1 << 2;
pub const TB_GR_HAS_DUP_PARA_CHAR_FRAC: u32 = 1 << 3;

// ---- reason bits (in the order the reference pushes its reasons)
pub const TB_C4_REASON_LOREM_IPSUM: u32 = 1 << 0;
pub const TB_C4_REASON_CURLY_BRACKET: u32 = 1 << 1;
pub const TB_C4_REASON_TOO_FEW_SENTENCES: u32 = 1 << 2;
pub const TB_GQ_REASON_SHORT_DOC: u32 = 1 << 0;
pub const TB_GQ_REASON_LONG_DOC: u32 = 1 << 1;
pub const TB_GQ_REASON_BELOW_AVG: u32 = 1 << 2;
pub const TB_GQ_REASON_ABOVE_AVG: u32 = 1 << 3;
pub const TB_GQ_REASON_TOO_MANY_HASHES: u32 = 1 << 4;
pub const TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS: u32 = 1 << 5;
pub const TB_GQ_REASON_TOO_MANY_BULLETS: u32 = 1 << 6;
pub const TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES: u32 = 1 << 7;
pub const TB_GQ_REASON_BELOW_ALPHA: u32 = 1 << 8;
pub const TB_GQ_REASON_TOO_FEW_STOP_WORDS: u32 = 1 << 9;
pub const TB_GR_REASON_DUP_PARA_FRAC: u32 = 1 << 0;
pub const TB_GR_REASON_DUP_PARA_CHAR_FRAC: u32 = 1 << 1;
pub const TB_GR_REASON_DUP_LINE_FRAC: u32 = 1 << 2;
pub const TB_GR_REASON_DUP_LINE_CHAR_FRAC: u32 = 1 << 3;
pub const TB_GR_REASON_TOP_SHIFT: u32 = 4;
pub const TB_GR_REASON_DUP_SHIFT: u32 = 4 + TB_MAX_NGRAMS as u32;
pub const TB_GR_REASON_EMPTY: u32 = 1 << 31;
pub const TB_FW_REASON_EMPTY: u32 = 1 << 0;
pub const TB_FW_REASON_LINE_PUNCT: u32 = 1 << 1;
pub const TB_FW_REASON_SHORT_LINE: u32 = 1 << 2;
pub const TB_FW_REASON_CHAR_DUP: u32 = 1 << 3;
pub const TB_FW_REASON_LIST_NO_WORDS: u32 = 1 << 4;
pub const TB_FW_REASON_LIST_RATIO: u32 = 1 << 5;

pub const TB_DOCFLAG_COMPLEX_SCRIPT: u32 = 1 << 0;
pub const TB_DOCFLAG_BAD_OFFSETS: u32 = 1 << 1;

// ---- options of tb_pipeline_set_option
pub const TB_OPT_RETURN_CONTENT: i32 = 1;
pub const TB_OPT_VALIDATE_UTF8: i32 = 2;
pub const TB_OPT_CHUNK_BYTES: i32 = 3;
pub const TB_OPT_GENERAL_WORD_PATH: i32 = 4;
pub const TB_OPT_DECODE_HTML_ENTITIES: i32 = 5;
pub const TB_OPT_GATHER_KEPT: i32 = 6;
pub const TB_OPT_CHUNK_RAMP: i32 = 7;
pub const TB_OPT_STAGE_THREADS: i32 = 8;
pub const TB_OPT_BULK_STAGE_TILES: i32 = 9;

pub const TB_NOALLOC_MAX_DOCS: i64 = 2048;
pub const TB_NOALLOC_MAX_BYTES: i64 = 512 * 1024;

/// C4QualityParams, src/config/pipeline.rs:66-79
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_c4_params {
    pub split_paragraph: i32,
    pub remove_citations: i32,
    pub filter_no_terminal_punct: i32,
    pub filter_lorem_ipsum: i32,
    pub filter_javascript: i32,
    pub filter_curly_bracket: i32,
    pub filter_policy: i32,
    pub _pad: i32,
    pub min_num_sentences: u64,
    pub min_words_per_line: u64,
    pub max_word_length: u64,
}

/// GopherQualityParams, src/config/pipeline.rs:161-176.  `has` bit i set <=> the i-th Option is Some.
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct tb_gq_params {
    pub has: u32,
    pub n_stop_words: i32,
    pub min_doc_words: u64,
    pub max_doc_words: u64,
    pub min_avg_word_length: f64,
    pub max_avg_word_length: f64,
    pub max_symbol_word_ratio: f64,
    pub max_bullet_lines_ratio: f64,
    pub max_ellipsis_lines_ratio: f64,
    pub max_non_alpha_words_ratio: f64,
    pub min_stop_words: u64,
    pub stop_words_utf8: *const u8,
    pub stop_words_offsets: *const i32,
}

/// GopherRepetitionParams, src/config/pipeline.rs:102-115
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct tb_gr_params {
    pub has: u32,
    pub n_top: i32,
    pub n_dup: i32,
    pub _pad: i32,
    pub dup_line_frac: f64,
    pub dup_para_frac: f64,
    pub dup_line_char_frac: f64,
    pub dup_para_char_frac: f64,
    pub top_n: [u32; TB_MAX_NGRAMS],
    pub top_frac: [f64; TB_MAX_NGRAMS],
    pub dup_n: [u32; TB_MAX_NGRAMS],
    pub dup_frac: [f64; TB_MAX_NGRAMS],
}

/// FineWebQualityFilterParams, src/config/pipeline.rs:311-322
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct tb_fw_params {
    pub line_punct_thr: f64,
    pub line_punct_exclude_zero: i32,
    pub n_stop_chars: i32,
    pub stop_chars: [u32; TB_FW_MAX_STOP_CHARS],
    pub short_line_thr: f64,
    pub short_line_length: u64,
    pub char_duplicates_ratio: f64,
    pub new_line_ratio: f64,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub union tb_step_params {
    pub c4: tb_c4_params,
    pub gq: tb_gq_params,
    pub gr: tb_gr_params,
    pub fw: tb_fw_params,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct tb_step_cfg {
    pub type_: i32,
    pub _pad: i32,
    pub u: tb_step_params,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_c4_stats {
    pub reason_bits: u32,
    pub n_lines: u32,
    pub too_long_word: u32,
    pub no_terminal_punc: u32,
    pub too_few_words: u32,
    pub n_kept_lines: u32,
    pub n_sentences: u32,
    pub rewritten: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_gq_stats {
    pub reason_bits: u32,
    pub n_words: u32,
    pub sum_word_chars: u64,
    pub n_hash: u32,
    pub n_ellipsis_units: u32,
    pub n_lines: u32,
    pub n_bullet_lines: u32,
    pub n_ellipsis_lines: u32,
    pub n_alpha_words: u32,
    pub n_stop_words: u32,
    pub flags: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_gr_stats {
    pub reason_bits: u32,
    pub flags: u32,
    pub char_len: u64,
    pub n_paras: u32,
    pub para_dup_elems: u32,
    pub para_dup_bytes: u64,
    pub n_lines: u32,
    pub line_dup_elems: u32,
    pub line_dup_bytes: u64,
    pub n_words: u64,
    pub top_bytes: [u64; TB_MAX_NGRAMS],
    pub dup_bytes: [u64; TB_MAX_NGRAMS],
}

#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_fw_stats {
    pub reason_bits: u32,
    pub flags: u32,
    pub n_lines: u32,
    pub n_stop_lines: u32,
    pub n_short_lines: u32,
    pub n_words: u32,
    pub n_newlines: u32,
    pub _pad: u32,
    pub dup_line_bytes: u64,
    pub n_chars_no_newline: u64,
}

/// One run of a document's content: bytes [begin, end) of the document's own input text (TB_OPT_RETURN_CONTENT = 2).
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct tb_run {
    pub begin: u32,
    pub end: u32,
}

#[repr(C)]
#[derive(Debug)]
pub struct tb_batch_out {
    pub n_docs: i64,
    pub n_steps: i32,
    pub content_rewritten: i32,
    pub keep: *mut u8,
    pub failed_step: *mut i8,
    pub step_stats: [*mut c_void; TB_MAX_STEPS],
    pub content: *mut u8,
    pub content_offsets: *mut i64,
    pub content_bytes: i64,
    pub gpu_ms: f64,
    pub gpu_launches: i32,
    pub content_kind: i32,
    pub _owner: *mut c_void,
    pub runs: *const tb_run,
    pub run_offsets: *const i64,
    pub n_runs: i64,
    pub kept_content: *const u8,
    pub kept_offsets: *const i64,
    pub n_kept: i64,
}

impl Default for tb_batch_out {
    fn default() -> Self {
        // all-zero is the header's "empty" value (tb_pipeline_run_batch memsets the struct first)
        unsafe { std::mem::zeroed() }
    }
}

// Layout checks: sizeof() of the C structs on x86-64 / aarch64 Linux as printed by tests/test_abi.py.
const _: () = {
    assert!(std::mem::size_of::<tb_c4_params>() == 56);
    assert!(std::mem::size_of::<tb_gq_params>() == 96);
    assert!(std::mem::size_of::<tb_gr_params>() == 336);
    assert!(std::mem::size_of::<tb_fw_params>() == 176);
    assert!(std::mem::size_of::<tb_step_cfg>() == 344);
    assert!(std::mem::size_of::<tb_c4_stats>() == 32);
    assert!(std::mem::size_of::<tb_gq_stats>() == 48);
    assert!(std::mem::size_of::<tb_gr_stats>() == 248);
    assert!(std::mem::size_of::<tb_fw_stats>() == 48);
    assert!(std::mem::size_of::<tb_run>() == 8);
    assert!(std::mem::size_of::<tb_batch_out>() == 192);
};

#[repr(C)]
pub struct tb_ctx {
    _p: [u8; 0],
}
#[repr(C)]
pub struct tb_pipeline {
    _p: [u8; 0],
}

extern "C" {
    pub fn tb_abi_version() -> c_int;
    pub fn tb_last_error(ctx: *const tb_ctx) -> *const c_char;
    pub fn tb_step_name(step_type: i32) -> *const c_char;
    pub fn tb_ctx_create(device: c_int, out: *mut *mut tb_ctx) -> c_int;
    pub fn tb_ctx_destroy(ctx: *mut tb_ctx);
    pub fn tb_pipeline_create(ctx: *mut tb_ctx, steps: *const tb_step_cfg, n_steps: usize, out: *mut *mut tb_pipeline) -> c_int;
    pub fn tb_pipeline_destroy(p: *mut tb_pipeline);
    pub fn tb_pipeline_run_batch(p: *mut tb_pipeline, utf8: *const u8, offsets: *const c_void, offsets_are_i64: c_int, n_docs: i64,
                                 out: *mut tb_batch_out) -> c_int;
    pub fn tb_pipeline_run_batch_device(p: *mut tb_pipeline, utf8_dev: *const u8, offsets_dev: *const i64, n_docs: i64, n_bytes: i64,
                                        out: *mut tb_batch_out) -> c_int;
    pub fn tb_batch_out_release(p: *mut tb_pipeline, out: *mut tb_batch_out);
    pub fn tb_gather_content(out: *const tb_batch_out, utf8: *const u8, offsets: *const c_void, offsets_are_i64: c_int, select: *const u8,
                             dst: *mut u8, dst_capacity: i64, dst_offsets: *mut i64, dst_bytes: *mut i64, nthreads: c_int) -> c_int;
    pub fn tb_pipeline_set_option(p: *mut tb_pipeline, option: i32, value: i64) -> c_int;
    pub fn tb_pipeline_kernel_times(p: *mut tb_pipeline, max_entries: i32, names: *mut *const c_char, ms: *mut f64, launches: *mut i64) -> c_int;
    pub fn tb_pipeline_reset_kernel_times(p: *mut tb_pipeline);
    pub fn tb_gopher_repetition_noalloc(utf8_dev: *const u8, offsets_dev: *const i64, n_docs: i64, n_bytes: i64, params: *const tb_gr_params,
                                        stats_dev: *mut tb_gr_stats, keep_dev: *mut u8, cuda_stream: *mut c_void, err: *mut c_char,
                                        err_len: usize) -> c_int;
}
