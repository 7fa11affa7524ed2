//! GPU implementations of TextBlaster's quality filters behind the reference's own types.
//!
//! NOT COMPILED IN THIS REPOSITORY (no Rust toolchain in the build image).  The same C ABI is driven from
//! Python (`textblaster_b200/executor.py`, which this file mirrors item by item) and from C++ (`include/textblaster_b200.hpp`)
//! by the test-suite; the reason strings below are the ones `textblaster_b200/reasons.py` is tested with.
//!
//! What it provides (reference file:line each item stands in for):
//!   * `GpuPipeline`                      -- owner of a `tb_ctx` + `tb_pipeline`: `build_pipeline_from_config`, src/worker_logic.rs:39-134
//!   * `step_cfg_from(&StepConfig)`       -- `StepConfig` -> `tb_step_cfg`, src/config/pipeline.rs:26-37 (None for step types without a GPU form)
//!   * `GpuPipelineExecutor::run_batch`   -- `PipelineExecutor::run_batch_parallel_async`, src/executor.rs:60-70, one device round trip per batch
//!   * `GpuC4QualityFilter`, `GpuGopherQualityFilter`, `GpuGopherRepetitionFilter`, `GpuFineWebQualityFilter`
//!                                        -- `ProcessingStep` adapters with the reference's `name()` strings, for mixed pipelines
//!   * `reasons`                          -- metadata keys and reason strings from the integer statistics
//!                                           (c4_filters.rs:167-293, gopher_quality.rs:198-315, gopher_rep.rs:61-218, fineweb_quality.rs:79-225)
use std::collections::HashMap;
use std::ffi::CStr;
use std::os::raw::c_void;
use std::sync::Arc;

use async_trait::async_trait;
use textblaster_b200_sys as sys;
use TextBlaster::config::pipeline::{
    C4QualityParams, FineWebQualityFilterParams, GopherQualityParams, GopherRepetitionParams, PipelineConfig, StepConfig,
};
use TextBlaster::data_model::TextDocument;
use TextBlaster::error::{PipelineError, Result};
use TextBlaster::executor::ProcessingStep;

fn last_error() -> String {
    unsafe { CStr::from_ptr(sys::tb_last_error(std::ptr::null())).to_string_lossy().into_owned() }
}

fn check(rc: i32) -> Result<()> {
    match rc {
        sys::TB_OK => Ok(()),
        sys::TB_ERR_CONFIG_VALIDATION => Err(PipelineError::ConfigValidationError(last_error())),
        _ => Err(PipelineError::Unexpected(format!("textblaster_b200 native error {rc}: {}", last_error()))),
    }
}

// ---------------------------------------------------------------------------------------------
// StepConfig -> tb_step_cfg
// ---------------------------------------------------------------------------------------------

/// A `tb_step_cfg` together with the buffers its pointers refer to (the stop-word list).
pub struct OwnedStepCfg {
    pub cfg: sys::tb_step_cfg,
    pub params: StepParams,
    _stop_utf8: Vec<u8>,
    _stop_offsets: Vec<i32>,
}

/// The parameters of an accelerated step, kept for the reason strings.
#[derive(Clone, Debug)]
pub enum StepParams {
    C4(C4QualityParams),
    GopherQuality(GopherQualityParams),
    GopherRepetition(GopherRepetitionParams),
    FineWeb(FineWebQualityFilterParams),
}

impl StepParams {
    /// `ProcessingStep::name` of the reference step (c4_filters.rs:143-145, gopher_quality.rs:65-67, gopher_rep.rs:48-50)
    pub fn name(&self) -> &'static str {
        match self {
            StepParams::C4(_) => "C4QualityFilter",
            StepParams::GopherQuality(_) => "GopherQualityFilter",
            StepParams::GopherRepetition(_) => "GopherRepetitionFilter",
            StepParams::FineWeb(_) => "FineWebQualityFilter",
        }
    }
}

fn zeroed_cfg(step_type: i32) -> sys::tb_step_cfg {
    let mut c: sys::tb_step_cfg = unsafe { std::mem::zeroed() };
    c.type_ = step_type;
    c
}

pub fn c4_cfg(p: &C4QualityParams) -> OwnedStepCfg {
    let mut cfg = zeroed_cfg(sys::TB_STEP_C4_QUALITY);
    cfg.u.c4 = sys::tb_c4_params {
        split_paragraph: p.split_paragraph as i32,
        remove_citations: p.remove_citations as i32,
        filter_no_terminal_punct: p.filter_no_terminal_punct as i32,
        filter_lorem_ipsum: p.filter_lorem_ipsum as i32,
        filter_javascript: p.filter_javascript as i32,
        filter_curly_bracket: p.filter_curly_bracket as i32,
        filter_policy: p.filter_policy as i32,
        _pad: 0,
        min_num_sentences: p.min_num_sentences as u64,
        min_words_per_line: p.min_words_per_line as u64,
        max_word_length: p.max_word_length as u64,
    };
    OwnedStepCfg { cfg, params: StepParams::C4(p.clone()), _stop_utf8: Vec::new(), _stop_offsets: Vec::new() }
}

pub fn gq_cfg(p: &GopherQualityParams) -> OwnedStepCfg {
    let mut has = 0u32;
    let mut set = |bit: u32, present: bool| if present { has |= bit };
    set(sys::TB_GQ_HAS_MIN_DOC_WORDS, p.min_doc_words.is_some());
    set(sys::TB_GQ_HAS_MAX_DOC_WORDS, p.max_doc_words.is_some());
    set(sys::TB_GQ_HAS_MIN_AVG_WORD_LENGTH, p.min_avg_word_length.is_some());
    set(sys::TB_GQ_HAS_MAX_AVG_WORD_LENGTH, p.max_avg_word_length.is_some());
    set(sys::TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO, p.max_symbol_word_ratio.is_some());
    set(sys::TB_GQ_HAS_MAX_BULLET_LINES_RATIO, p.max_bullet_lines_ratio.is_some());
    set(sys::TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO, p.max_ellipsis_lines_ratio.is_some());
    set(sys::TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO, p.max_non_alpha_words_ratio.is_some());
    set(sys::TB_GQ_HAS_MIN_STOP_WORDS, p.min_stop_words.is_some());
    set(sys::TB_GQ_HAS_STOP_WORDS, p.stop_words.is_some());
    // the list is passed as given: the reference does not lower-case it (gopher_quality.rs:46-58), only the word it looks up
    let mut utf8 = Vec::new();
    let mut offsets = vec![0i32];
    for w in p.stop_words.iter().flatten() {
        utf8.extend_from_slice(w.as_bytes());
        offsets.push(utf8.len() as i32);
    }
    let mut cfg = zeroed_cfg(sys::TB_STEP_GOPHER_QUALITY);
    cfg.u.gq = sys::tb_gq_params {
        has,
        n_stop_words: p.stop_words.as_ref().map_or(0, |v| v.len() as i32),
        min_doc_words: p.min_doc_words.unwrap_or(0) as u64,
        max_doc_words: p.max_doc_words.unwrap_or(0) as u64,
        min_avg_word_length: p.min_avg_word_length.unwrap_or(0.0),
        max_avg_word_length: p.max_avg_word_length.unwrap_or(0.0),
        max_symbol_word_ratio: p.max_symbol_word_ratio.unwrap_or(0.0),
        max_bullet_lines_ratio: p.max_bullet_lines_ratio.unwrap_or(0.0),
        max_ellipsis_lines_ratio: p.max_ellipsis_lines_ratio.unwrap_or(0.0),
        max_non_alpha_words_ratio: p.max_non_alpha_words_ratio.unwrap_or(0.0),
        min_stop_words: p.min_stop_words.unwrap_or(0) as u64,
        stop_words_utf8: std::ptr::null(),
        stop_words_offsets: std::ptr::null(),
    };
    let mut owned = OwnedStepCfg { cfg, params: StepParams::GopherQuality(p.clone()), _stop_utf8: utf8, _stop_offsets: offsets };
    // the vectors do not move again: their heap buffers stay where they are when `owned` is moved
    unsafe {
        owned.cfg.u.gq.stop_words_utf8 = owned._stop_utf8.as_ptr();
        owned.cfg.u.gq.stop_words_offsets = owned._stop_offsets.as_ptr();
    }
    owned
}

pub fn gr_cfg(p: &GopherRepetitionParams) -> Result<OwnedStepCfg> {
    if p.top_n_grams.len() > sys::TB_MAX_NGRAMS || p.dup_n_grams.len() > sys::TB_MAX_NGRAMS {
        return Err(PipelineError::ConfigError(format!("GopherRepetitionFilter on the GPU: at most {} entries per n-gram list", sys::TB_MAX_NGRAMS)));
    }
    let mut g: sys::tb_gr_params = unsafe { std::mem::zeroed() };
    for (bit, v, slot) in [
        (sys::TB_GR_HAS_DUP_LINE_FRAC, p.dup_line_frac, &mut g.dup_line_frac as *mut f64),
        (sys::TB_GR_HAS_DUP_PARA_FRAC, p.dup_para_frac, &mut g.dup_para_frac as *mut f64),
        (sys::TB_GR_HAS_DUP_LINE_CHAR_FRAC, p.dup_line_char_frac, &mut g.dup_line_char_frac as *mut f64),
        (sys::TB_GR_HAS_DUP_PARA_CHAR_FRAC, p.dup_para_char_frac, &mut g.dup_para_char_frac as *mut f64),
    ] {
        if let Some(x) = v {
            g.has |= bit;
            unsafe { *slot = x };
        }
    }
    g.n_top = p.top_n_grams.len() as i32;
    g.n_dup = p.dup_n_grams.len() as i32;
    for (i, (n, f)) in p.top_n_grams.iter().enumerate() {
        g.top_n[i] = *n as u32;
        g.top_frac[i] = *f;
    }
    for (i, (n, f)) in p.dup_n_grams.iter().enumerate() {
        g.dup_n[i] = *n as u32;
        g.dup_frac[i] = *f;
    }
    let mut cfg = zeroed_cfg(sys::TB_STEP_GOPHER_REPETITION);
    cfg.u.gr = g;
    Ok(OwnedStepCfg { cfg, params: StepParams::GopherRepetition(p.clone()), _stop_utf8: Vec::new(), _stop_offsets: Vec::new() })
}

pub fn fw_cfg(p: &FineWebQualityFilterParams) -> Result<OwnedStepCfg> {
    let mut f: sys::tb_fw_params = unsafe { std::mem::zeroed() };
    f.line_punct_thr = p.line_punct_thr;
    f.line_punct_exclude_zero = p.line_punct_exclude_zero as i32;
    f.short_line_thr = p.short_line_thr;
    f.short_line_length = p.short_line_length as u64;
    f.char_duplicates_ratio = p.char_duplicates_ratio;
    f.new_line_ratio = p.new_line_ratio;
    match &p.stop_chars {
        None => f.n_stop_chars = -1, // END_PUNCTUATION, fineweb_quality.rs:26,51
        Some(set) => {
            if set.len() > sys::TB_FW_MAX_STOP_CHARS {
                return Err(PipelineError::ConfigError(format!("FineWebQualityFilter on the GPU: at most {} stop_chars", sys::TB_FW_MAX_STOP_CHARS)));
            }
            f.n_stop_chars = set.len() as i32;
            for (i, c) in set.iter().enumerate() {
                f.stop_chars[i] = *c as u32;
            }
        }
    }
    let mut cfg = zeroed_cfg(sys::TB_STEP_FINEWEB_QUALITY);
    cfg.u.fw = f;
    Ok(OwnedStepCfg { cfg, params: StepParams::FineWeb(p.clone()), _stop_utf8: Vec::new(), _stop_offsets: Vec::new() })
}

/// `None`: the step type has no GPU implementation (language detection, bad words, token counter).
pub fn step_cfg_from(step: &StepConfig) -> Result<Option<OwnedStepCfg>> {
    Ok(match step {
        StepConfig::C4QualityFilter(p) => Some(c4_cfg(p)),
        StepConfig::GopherQualityFilter(p) => Some(gq_cfg(p)),
        StepConfig::GopherRepetitionFilter(p) => Some(gr_cfg(p)?),
        StepConfig::FineWebQualityFilter(p) => Some(fw_cfg(p)?),
        _ => None,
    })
}

// ---------------------------------------------------------------------------------------------
// reasons: metadata keys + reason strings from the statistics
// ---------------------------------------------------------------------------------------------
pub mod reasons {
    use super::*;

    type Md = Vec<(&'static str, String)>;

    /// c4_filters.rs:166-187, 264-293
    pub fn c4(p: &C4QualityParams, st: &sys::tb_c4_stats) -> (Md, String) {
        let mut r: Vec<String> = Vec::new();
        if st.reason_bits & sys::TB_C4_REASON_LOREM_IPSUM != 0 { r.push("lorem_ipsum".into()); }
        if st.reason_bits & sys::TB_C4_REASON_CURLY_BRACKET != 0 { r.push("curly_bracket".into()); }
        if st.reason_bits & sys::TB_C4_REASON_TOO_FEW_SENTENCES != 0 {
            r.push(format!("too_few_sentences (found {}, required {})", st.n_sentences, p.min_num_sentences));
        }
        if r.is_empty() {
            return (vec![("c4_filter_status", "passed".into())], String::new());
        }
        let rs = r.join("; ");
        let mut md: Md = vec![("c4_filter_status", "filtered".into()), ("c4_filter_reasons", rs.clone())];
        if st.rewritten != 0 {
            // late exit: only the counters that are present in the HashMap (c4_filters.rs:281-283)
            for (key, v) in [("line-filter-too_long_word", st.too_long_word), ("line-filter-no_terminal_punc", st.no_terminal_punc),
                             ("line-filter-too_few_words", st.too_few_words)] {
                if v != 0 { md.push((key, v.to_string())); }
            }
        }
        (md, rs)
    }

    /// gopher_quality.rs:198-315
    pub fn gopher_quality(p: &GopherQualityParams, st: &sys::tb_gq_stats) -> (Md, String) {
        let n = st.n_words as u64;
        let nw = n.max(1) as f64;
        let nl = (st.n_lines as u64).max(1) as f64;
        let avg = if n > 0 { st.sum_word_chars as f64 / n as f64 } else { 0.0 };
        let b = st.reason_bits;
        let mut r: Vec<String> = Vec::new();
        if b & sys::TB_GQ_REASON_SHORT_DOC != 0 { r.push(format!("gopher_short_doc ({} non-symbol words, required {})", n, p.min_doc_words.unwrap_or(0))); }
        if b & sys::TB_GQ_REASON_LONG_DOC != 0 { r.push(format!("gopher_long_doc ({} non-symbol words, max {})", n, p.max_doc_words.unwrap_or(0))); }
        if b & sys::TB_GQ_REASON_BELOW_AVG != 0 {
            let min = p.min_avg_word_length.unwrap_or(0.0);
            let tail = if n == 0 && min > 0.0 { " - 0 non-symbol words" } else { "" };
            r.push(format!("gopher_below_avg_threshold (avg len {:.2}, required {:.2}{})", avg, min, tail));
        }
        if b & sys::TB_GQ_REASON_ABOVE_AVG != 0 { r.push(format!("gopher_above_avg_threshold (avg len {:.2}, max {:.2})", avg, p.max_avg_word_length.unwrap_or(0.0))); }
        let sym = p.max_symbol_word_ratio.unwrap_or(0.0);
        if b & sys::TB_GQ_REASON_TOO_MANY_HASHES != 0 { r.push(format!("gopher_too_many_hashes (ratio {:.2}, max {:.2})", st.n_hash as f64 / nw, sym)); }
        if b & sys::TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS != 0 { r.push(format!("gopher_too_many_ellipsis_units (ratio {:.2}, max {:.2})", st.n_ellipsis_units as f64 / nw, sym)); }
        if b & sys::TB_GQ_REASON_TOO_MANY_BULLETS != 0 { r.push(format!("gopher_too_many_bullets (ratio {:.2}, max {:.2})", st.n_bullet_lines as f64 / nl, p.max_bullet_lines_ratio.unwrap_or(0.0))); }
        if b & sys::TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES != 0 { r.push(format!("gopher_too_many_end_ellipsis_lines (ratio {:.2}, max {:.2})", st.n_ellipsis_lines as f64 / nl, p.max_ellipsis_lines_ratio.unwrap_or(0.0))); }
        if b & sys::TB_GQ_REASON_BELOW_ALPHA != 0 { r.push(format!("gopher_below_alpha_threshold (alpha ratio {:.2}, required min {:.2})", st.n_alpha_words as f64 / nw, p.max_non_alpha_words_ratio.unwrap_or(0.0))); }
        if b & sys::TB_GQ_REASON_TOO_FEW_STOP_WORDS != 0 { r.push(format!("gopher_too_few_stop_words (found {}, required {})", st.n_stop_words, p.min_stop_words.unwrap_or(0))); }
        if r.is_empty() {
            return (vec![("gopher_quality_filter_status", "passed".into())], String::new());
        }
        let rs = r.join("; ");
        (vec![("gopher_quality_filter_status", "filtered".into()), ("gopher_quality_filter_reasons", rs.clone())], rs)
    }

    /// gopher_rep.rs:60-74 (empty content: the singular key), :95-218
    pub fn gopher_repetition(p: &GopherRepetitionParams, st: &sys::tb_gr_stats) -> (Md, String) {
        let b = st.reason_bits;
        if b & sys::TB_GR_REASON_EMPTY != 0 {
            let rs = "skipping empty content".to_string();
            return (vec![("gopher_repetition_filter_status", "filtered".into()), ("gopher_repetition_filter_reason", rs.clone())], rs);
        }
        let tcl = st.char_len.max(1) as f64;
        let mut r: Vec<String> = Vec::new();
        if b & sys::TB_GR_REASON_DUP_PARA_FRAC != 0 { r.push(format!("dup_para_frac (ratio {:.2}, max {:.2})", st.para_dup_elems as f64 / (st.n_paras.max(1) as f64), p.dup_para_frac.unwrap_or(0.0))); }
        if b & sys::TB_GR_REASON_DUP_PARA_CHAR_FRAC != 0 { r.push(format!("dup_para_char_frac (ratio {:.2}, max {:.2})", st.para_dup_bytes as f64 / tcl, p.dup_para_char_frac.unwrap_or(0.0))); }
        if b & sys::TB_GR_REASON_DUP_LINE_FRAC != 0 { r.push(format!("dup_line_frac (ratio {:.2}, max {:.2})", st.line_dup_elems as f64 / (st.n_lines.max(1) as f64), p.dup_line_frac.unwrap_or(0.0))); }
        if b & sys::TB_GR_REASON_DUP_LINE_CHAR_FRAC != 0 { r.push(format!("dup_line_char_frac (ratio {:.2}, max {:.2})", st.line_dup_bytes as f64 / tcl, p.dup_line_char_frac.unwrap_or(0.0))); }
        for (i, (n, frac)) in p.top_n_grams.iter().enumerate() {
            if b & (1u32 << (sys::TB_GR_REASON_TOP_SHIFT + i as u32)) != 0 { r.push(format!("top_{}_gram (ratio {:.2}, max {:.2})", n, st.top_bytes[i] as f64 / tcl, frac)); }
        }
        for (i, (n, frac)) in p.dup_n_grams.iter().enumerate() {
            if b & (1u32 << (sys::TB_GR_REASON_DUP_SHIFT + i as u32)) != 0 { r.push(format!("duplicated_{}_n_grams (ratio {:.2}, max {:.2})", n, st.dup_bytes[i] as f64 / tcl, frac)); }
        }
        if r.is_empty() {
            return (vec![("gopher_repetition_filter_status", "passed".into())], String::new());
        }
        let rs = r.join("; ");
        (vec![("gopher_repetition_filter_status", "filtered".into()), ("gopher_repetition_filter_reasons", rs.clone())], rs)
    }

    /// fineweb_quality.rs:79-225: the first failing check returns; nothing is written on pass
    pub fn fineweb(p: &FineWebQualityFilterParams, st: &sys::tb_fw_stats) -> (Md, String) {
        let b = st.reason_bits;
        if b == 0 {
            return (Vec::new(), String::new());
        }
        if b & sys::TB_FW_REASON_EMPTY != 0 {
            // the metadata value differs from the reason of the error (:82-89)
            return (vec![("fineweb_filter_status", "filtered".into()), ("fineweb_filter_reason", "empty document".into())], "empty".into());
        }
        let nl = st.n_lines as f64;
        let r = if b & sys::TB_FW_REASON_LINE_PUNCT != 0 {
            format!("line_punct_ratio: {:.4} < threshold {:.4} (exclude_zero: {})", st.n_stop_lines as f64 / nl, p.line_punct_thr, p.line_punct_exclude_zero)
        } else if b & sys::TB_FW_REASON_SHORT_LINE != 0 {
            format!("short_line_ratio: {:.4} > threshold {:.4}", st.n_short_lines as f64 / nl, p.short_line_thr)
        } else if b & sys::TB_FW_REASON_CHAR_DUP != 0 {
            format!("char_dup_ratio: {:.4} > threshold {:.4}", st.dup_line_bytes as f64 / st.n_chars_no_newline as f64, p.char_duplicates_ratio)
        } else if b & sys::TB_FW_REASON_LIST_NO_WORDS != 0 {
            "list_ratio_no_words (newlines present but no words)".to_string()
        } else {
            format!("list_ratio: {:.4} > threshold {:.4}", st.n_newlines as f64 / st.n_words as f64, p.new_line_ratio)
        };
        (vec![("fineweb_filter_status", "filtered".into()), ("fineweb_filter_reason", r.clone())], r)
    }
}

// ---------------------------------------------------------------------------------------------
// the pipeline handle
// ---------------------------------------------------------------------------------------------
struct Handles {
    ctx: *mut sys::tb_ctx,
    pipe: *mut sys::tb_pipeline,
}
// the library locks a pipeline internally and keeps no thread affinity (include/textblaster_b200.h, tb_ctx_create)
unsafe impl Send for Handles {}
unsafe impl Sync for Handles {}
impl Drop for Handles {
    fn drop(&mut self) {
        unsafe {
            sys::tb_pipeline_destroy(self.pipe);
            sys::tb_ctx_destroy(self.ctx);
        }
    }
}

/// An immutable chain of accelerated steps on one GPU.  `Send + Sync` like the reference's steps (src/executor.rs:9).
pub struct GpuPipeline {
    h: Handles,
    steps: Vec<StepParams>,
    _cfgs: Vec<OwnedStepCfg>,
}

impl GpuPipeline {
    /// `build_pipeline_from_config` for a run of accelerated steps; validation messages are the reference's
    /// (`ConfigValidationError`, src/config/pipeline.rs:82-257).
    pub fn new(device: i32, cfgs: Vec<OwnedStepCfg>) -> Result<Self> {
        let mut ctx: *mut sys::tb_ctx = std::ptr::null_mut();
        check(unsafe { sys::tb_ctx_create(device, &mut ctx) })?;
        let raw: Vec<sys::tb_step_cfg> = cfgs.iter().map(|c| c.cfg).collect();
        let mut pipe: *mut sys::tb_pipeline = std::ptr::null_mut();
        let rc = unsafe { sys::tb_pipeline_create(ctx, raw.as_ptr(), raw.len(), &mut pipe) };
        if rc != sys::TB_OK {
            let e = check(rc);
            unsafe { sys::tb_ctx_destroy(ctx) };
            e?;
        }
        let steps = cfgs.iter().map(|c| c.params.clone()).collect();
        Ok(GpuPipeline { h: Handles { ctx, pipe }, steps, _cfgs: cfgs })
    }

    pub fn set_option(&self, option: i32, value: i64) -> Result<()> {
        check(unsafe { sys::tb_pipeline_set_option(self.h.pipe, option, value) })
    }

    pub fn step_names(&self) -> Vec<&'static str> {
        self.steps.iter().map(|s| s.name()).collect()
    }

    /// One Arrow string column (contiguous UTF-8 values + i32 offsets, src/pipeline/readers/parquet_reader.rs:108-112)
    /// through the chain.  `f` sees the results while they are valid; they are released when it returns.
    pub fn run_column<T>(&self, values: &[u8], offsets: &[i32], f: impl FnOnce(&BatchView) -> T) -> Result<T> {
        assert!(!offsets.is_empty());
        let n = (offsets.len() - 1) as i64;
        let mut out = sys::tb_batch_out::default();
        check(unsafe { sys::tb_pipeline_run_batch(self.h.pipe, values.as_ptr(), offsets.as_ptr() as *const c_void, 0, n, &mut out) })?;
        let view = BatchView { out: &out, steps: &self.steps };
        let r = f(&view);
        unsafe { sys::tb_batch_out_release(self.h.pipe, &mut out) };
        Ok(r)
    }
}

/// Results of one batch, borrowed from the library's pinned buffers.
pub struct BatchView<'a> {
    out: &'a sys::tb_batch_out,
    steps: &'a [StepParams],
}

impl<'a> BatchView<'a> {
    pub fn n_docs(&self) -> usize { self.out.n_docs as usize }
    pub fn keep(&self) -> &[u8] { unsafe { std::slice::from_raw_parts(self.out.keep, self.n_docs()) } }
    pub fn failed_step(&self) -> &[i8] { unsafe { std::slice::from_raw_parts(self.out.failed_step, self.n_docs()) } }

    /// content of document i after the chain, if the chain rewrote content and returned it as bytes
    pub fn content(&self, i: usize) -> Option<&str> {
        if self.out.content_rewritten == 0 { return None; }
        unsafe {
            let o = std::slice::from_raw_parts(self.out.content_offsets, self.n_docs() + 1);
            let bytes = std::slice::from_raw_parts(self.out.content.add(o[i] as usize), (o[i + 1] - o[i]) as usize);
            Some(std::str::from_utf8_unchecked(bytes)) // valid UTF-8 in, whole characters kept: valid UTF-8 out
        }
    }

    /// TB_OPT_RETURN_CONTENT = 2: the content of document i as runs of its input text
    pub fn content_from_runs(&self, i: usize, input: &str) -> Option<String> {
        if self.out.content_kind != 2 { return None; }
        unsafe {
            let ro = std::slice::from_raw_parts(self.out.run_offsets, self.n_docs() + 1);
            let runs = std::slice::from_raw_parts(self.out.runs.add(ro[i] as usize), (ro[i + 1] - ro[i]) as usize);
            let mut s = String::with_capacity(runs.iter().map(|r| (r.end - r.begin) as usize).sum());
            for r in runs { s.push_str(&input[r.begin as usize..r.end as usize]); }
            Some(s)
        }
    }

    /// Metadata written by every step that ran for document i (in step order), and the failing step's reason.
    pub fn metadata_and_reason(&self, i: usize) -> (Vec<(&'static str, String)>, String) {
        let fs = self.failed_step()[i];
        let last = if fs >= 0 { fs as usize } else { self.steps.len().saturating_sub(1) };
        let mut md = Vec::new();
        let mut reason = String::new();
        for s in 0..=last.min(self.steps.len().saturating_sub(1)) {
            if self.steps.is_empty() { break; }
            let base = self.out.step_stats[s];
            let (m, r) = unsafe {
                match &self.steps[s] {
                    StepParams::C4(p) => reasons::c4(p, &*(base as *const sys::tb_c4_stats).add(i)),
                    StepParams::GopherQuality(p) => reasons::gopher_quality(p, &*(base as *const sys::tb_gq_stats).add(i)),
                    StepParams::GopherRepetition(p) => reasons::gopher_repetition(p, &*(base as *const sys::tb_gr_stats).add(i)),
                    StepParams::FineWeb(p) => reasons::fineweb(p, &*(base as *const sys::tb_fw_stats).add(i)),
                }
            };
            md.extend(m);
            if fs >= 0 && s == fs as usize { reason = r; }
        }
        (md, reason)
    }
}

// ---------------------------------------------------------------------------------------------
// the batch executor: PipelineExecutor::run_batch_parallel_async for a chain of accelerated steps
// ---------------------------------------------------------------------------------------------
pub struct GpuPipelineExecutor {
    pipe: Arc<GpuPipeline>,
}

impl GpuPipelineExecutor {
    /// `Ok(None)` when the configuration holds a step without a GPU form: the caller keeps the stock `PipelineExecutor`
    /// (src/bin/worker.rs:147-149).
    pub fn from_config(config: &PipelineConfig, device: i32) -> Result<Option<Self>> {
        let mut cfgs = Vec::new();
        for step in &config.pipeline {
            match step_cfg_from(step)? {
                Some(c) => cfgs.push(c),
                None => return Ok(None),
            }
        }
        Ok(Some(GpuPipelineExecutor { pipe: Arc::new(GpuPipeline::new(device, cfgs)?) }))
    }

    /// Same contract as `PipelineExecutor::run_batch_parallel_async` (src/executor.rs:60-70) + `run_single_async`'s error
    /// wrapping (:37-46): one `Result` per input document, in input order (tests/executor_test.rs:284-291); a filtered
    /// document comes back as `StepError { step_name, source: DocumentFiltered { document, reason } }` with the document
    /// as the failing step left it (content rewritten by C4 if C4 ran, metadata of every step that ran).
    pub fn run_batch(&self, documents: Vec<TextDocument>) -> Vec<Result<TextDocument>> {
        if documents.is_empty() {
            return Vec::new();
        }
        // pack: contiguous UTF-8 + i32 offsets (a RecordBatch's StringArray hands over value_data() / value_offsets() instead)
        let total: usize = documents.iter().map(|d| d.content.len()).sum();
        if total > i32::MAX as usize {
            // a batch beyond 2 GiB goes in two halves (or with i64 offsets through tb_pipeline_run_batch(.., 1, ..))
            let mut docs = documents;
            let tail = docs.split_off(docs.len() / 2);
            let mut r = self.run_batch(docs);
            r.extend(self.run_batch(tail));
            return r;
        }
        let mut values = Vec::with_capacity(total);
        let mut offsets = Vec::with_capacity(documents.len() + 1);
        offsets.push(0i32);
        for d in &documents {
            values.extend_from_slice(d.content.as_bytes());
            offsets.push(values.len() as i32);
        }
        let names = self.pipe.step_names();
        let run = self.pipe.run_column(&values, &offsets, |view| {
            let mut results = Vec::with_capacity(documents.len());
            for (i, mut doc) in documents.into_iter().enumerate() {
                let (md, reason) = view.metadata_and_reason(i);
                if let Some(c) = view.content(i) {
                    if c != doc.content { doc.content = c.to_string(); }
                } else if let Some(c) = view.content_from_runs(i, &doc.content) {
                    doc.content = c;
                }
                for (k, v) in md { doc.metadata.insert(k.to_string(), v); }
                let fs = view.failed_step()[i];
                if fs < 0 {
                    results.push(Ok(doc));
                } else {
                    results.push(Err(PipelineError::StepError {
                        step_name: names[fs as usize].to_string(),
                        source: Box::new(PipelineError::DocumentFiltered { document: Box::new(doc), reason }),
                    }));
                }
            }
            results
        });
        match run {
            Ok(r) => r,
            Err(e) => {
                // a device failure is every document's error (the reference would report it per document too)
                let msg = e.to_string();
                (0..offsets.len() - 1).map(|_| Err(PipelineError::Unexpected(msg.clone()))).collect()
            }
        }
    }

    /// A whole Arrow string column without building `TextDocument`s: keep mask + failing step, as the aggregator needs
    /// them to route rows to the output / excluded files (src/producer_logic.rs:157-167).
    pub fn run_string_array(&self, column: &arrow::array::StringArray) -> Result<(Vec<bool>, Vec<i8>)> {
        let first = column.value_offsets()[0];
        let _ = first; // offsets may start anywhere: the library reads offsets[0] as the base
        self.pipe.run_column(column.value_data(), column.value_offsets(), |v| {
            (v.keep().iter().map(|&k| k != 0).collect(), v.failed_step().to_vec())
        })
    }
}

// ---------------------------------------------------------------------------------------------
// ProcessingStep adapters (for pipelines that mix accelerated and stock steps)
// ---------------------------------------------------------------------------------------------
macro_rules! gpu_step {
    ($ty:ident, $name:literal) => {
        pub struct $ty {
            pipe: GpuPipeline,
        }
        #[async_trait]
        impl ProcessingStep for $ty {
            fn name(&self) -> &'static str {
                $name
            }
            /// A batch of one: correct, but the device only pays off on whole record batches (GpuPipelineExecutor).
            async fn process(&self, mut document: TextDocument) -> Result<TextDocument> {
                let offsets = [0i32, document.content.len() as i32];
                let content = std::mem::take(&mut document.content);
                let (md, reason, keep, new_content) = self.pipe.run_column(content.as_bytes(), &offsets, |v| {
                    let (md, reason) = v.metadata_and_reason(0);
                    (md, reason, v.keep()[0] != 0, v.content(0).map(|s| s.to_string()))
                })?;
                document.content = new_content.unwrap_or(content);
                for (k, v) in md {
                    document.metadata.insert(k.to_string(), v);
                }
                if keep {
                    Ok(document)
                } else {
                    Err(PipelineError::DocumentFiltered { document: Box::new(document), reason })
                }
            }
        }
    };
}
gpu_step!(GpuC4QualityFilter, "C4QualityFilter");
gpu_step!(GpuGopherQualityFilter, "GopherQualityFilter");
gpu_step!(GpuGopherRepetitionFilter, "GopherRepetitionFilter");
gpu_step!(GpuFineWebQualityFilter, "FineWebQualityFilter");

impl GpuC4QualityFilter {
    pub fn new(p: &C4QualityParams, device: i32) -> Result<Self> { Ok(Self { pipe: GpuPipeline::new(device, vec![c4_cfg(p)])? }) }
}
impl GpuGopherQualityFilter {
    pub fn new(p: &GopherQualityParams, device: i32) -> Result<Self> { Ok(Self { pipe: GpuPipeline::new(device, vec![gq_cfg(p)])? }) }
}
impl GpuGopherRepetitionFilter {
    pub fn new(p: &GopherRepetitionParams, device: i32) -> Result<Self> { Ok(Self { pipe: GpuPipeline::new(device, vec![gr_cfg(p)?])? }) }
}
impl GpuFineWebQualityFilter {
    pub fn new(p: &FineWebQualityFilterParams, device: i32) -> Result<Self> { Ok(Self { pipe: GpuPipeline::new(device, vec![fw_cfg(p)?])? }) }
}

/// `build_pipeline_from_config` (src/worker_logic.rs:39-134) with the accelerated step types replaced one by one; every
/// other step type is built by `stock`, the reference's own factory for that step.
pub fn build_pipeline_from_config(
    config: &PipelineConfig,
    device: i32,
    stock: impl Fn(&StepConfig) -> Result<Box<dyn ProcessingStep>>,
) -> Result<Vec<Box<dyn ProcessingStep>>> {
    let mut steps: Vec<Box<dyn ProcessingStep>> = Vec::new();
    for step in &config.pipeline {
        steps.push(match step {
            StepConfig::C4QualityFilter(p) => Box::new(GpuC4QualityFilter::new(p, device)?),
            StepConfig::GopherQualityFilter(p) => Box::new(GpuGopherQualityFilter::new(p, device)?),
            StepConfig::GopherRepetitionFilter(p) => Box::new(GpuGopherRepetitionFilter::new(p, device)?),
            StepConfig::FineWebQualityFilter(p) => Box::new(GpuFineWebQualityFilter::new(p, device)?),
            other => stock(other)?,
        });
    }
    Ok(steps)
}

#[allow(dead_code)]
fn _metadata_type_check(doc: &TextDocument) -> &HashMap<String, String> {
    &doc.metadata
}
