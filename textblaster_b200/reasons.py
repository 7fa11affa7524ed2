"""Rebuild the reference's reason strings and metadata keys from the integer statistics the GPU
returns.  Ratios are recomputed on the host as the same single IEEE-754 division the reference does;
``{:.2}`` in Rust and ``%.2f`` here both round the exact binary value, so the text is identical.

Format strings: c4_filters.rs:167,172,265-268; gopher_quality.rs:200-293; gopher_rep.rs:66-68,97-194.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

from . import _abi
from .config import C4QualityParams, FineWebQualityParams, GopherQualityParams, GopherRepetitionParams


def _f2(v: float) -> str:
    return "%.2f" % v


def c4_reasons(p: C4QualityParams, st) -> List[str]:
    r = []
    bits = int(st["reason_bits"])
    if bits & _abi.TB_C4_REASON_LOREM_IPSUM:
        r.append("lorem_ipsum")
    if bits & _abi.TB_C4_REASON_CURLY_BRACKET:
        r.append("curly_bracket")
    if bits & _abi.TB_C4_REASON_TOO_FEW_SENTENCES:
        r.append(f"too_few_sentences (found {int(st['n_sentences'])}, required {p.min_num_sentences})")
    return r


def c4_metadata(p: C4QualityParams, st) -> Tuple[Dict[str, str], str]:
    """(metadata written by the step, reason string or '')  -- c4_filters.rs:175-187, 272-293"""
    reasons = c4_reasons(p, st)
    if not reasons:
        return {"c4_filter_status": "passed"}, ""
    rs = "; ".join(reasons)
    md = {"c4_filter_status": "filtered", "c4_filter_reasons": rs}
    if int(st["rewritten"]):  # late exit: the line counters that are present in the HashMap
        for key, field in (("line-filter-too_long_word", "too_long_word"), ("line-filter-no_terminal_punc", "no_terminal_punc"),
                           ("line-filter-too_few_words", "too_few_words")):
            if int(st[field]):
                md[key] = str(int(st[field]))
    return md, rs


def gq_reasons(p: GopherQualityParams, st) -> List[str]:
    bits = int(st["reason_bits"])
    n = int(st["n_words"])
    nw = float(max(n, 1))
    nl = float(max(int(st["n_lines"]), 1))
    avg = (float(int(st["sum_word_chars"])) / float(n)) if n > 0 else 0.0
    r = []
    if bits & _abi.TB_GQ_REASON_SHORT_DOC:
        r.append(f"gopher_short_doc ({n} non-symbol words, required {p.min_doc_words})")
    if bits & _abi.TB_GQ_REASON_LONG_DOC:
        r.append(f"gopher_long_doc ({n} non-symbol words, max {p.max_doc_words})")
    if bits & _abi.TB_GQ_REASON_BELOW_AVG:
        tail = " - 0 non-symbol words" if (n == 0 and p.min_avg_word_length > 0.0) else ""
        r.append(f"gopher_below_avg_threshold (avg len {_f2(avg)}, required {_f2(p.min_avg_word_length)}{tail})")
    if bits & _abi.TB_GQ_REASON_ABOVE_AVG:
        r.append(f"gopher_above_avg_threshold (avg len {_f2(avg)}, max {_f2(p.max_avg_word_length)})")
    if bits & _abi.TB_GQ_REASON_TOO_MANY_HASHES:
        r.append(f"gopher_too_many_hashes (ratio {_f2(float(int(st['n_hash'])) / nw)}, max {_f2(p.max_symbol_word_ratio)})")
    if bits & _abi.TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS:
        r.append(f"gopher_too_many_ellipsis_units (ratio {_f2(float(int(st['n_ellipsis_units'])) / nw)}, max {_f2(p.max_symbol_word_ratio)})")
    if bits & _abi.TB_GQ_REASON_TOO_MANY_BULLETS:
        r.append(f"gopher_too_many_bullets (ratio {_f2(float(int(st['n_bullet_lines'])) / nl)}, max {_f2(p.max_bullet_lines_ratio)})")
    if bits & _abi.TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES:
        r.append(f"gopher_too_many_end_ellipsis_lines (ratio {_f2(float(int(st['n_ellipsis_lines'])) / nl)}, max {_f2(p.max_ellipsis_lines_ratio)})")
    if bits & _abi.TB_GQ_REASON_BELOW_ALPHA:
        r.append(f"gopher_below_alpha_threshold (alpha ratio {_f2(float(int(st['n_alpha_words'])) / nw)}, required min {_f2(p.max_non_alpha_words_ratio)})")
    if bits & _abi.TB_GQ_REASON_TOO_FEW_STOP_WORDS:
        r.append(f"gopher_too_few_stop_words (found {int(st['n_stop_words'])}, required {p.min_stop_words})")
    return r


def gq_metadata(p: GopherQualityParams, st) -> Tuple[Dict[str, str], str]:
    reasons = gq_reasons(p, st)
    if not reasons:
        return {"gopher_quality_filter_status": "passed"}, ""
    rs = "; ".join(reasons)
    return {"gopher_quality_filter_status": "filtered", "gopher_quality_filter_reasons": rs}, rs


def gr_reasons(p: GopherRepetitionParams, st) -> List[str]:
    bits = int(st["reason_bits"])
    if bits & _abi.TB_GR_REASON_EMPTY:
        return ["skipping empty content"]
    tcl = float(max(int(st["char_len"]), 1))
    r = []
    if bits & _abi.TB_GR_REASON_DUP_PARA_FRAC:
        r.append(f"dup_para_frac (ratio {_f2(float(int(st['para_dup_elems'])) / float(max(int(st['n_paras']), 1)))}, max {_f2(p.dup_para_frac)})")
    if bits & _abi.TB_GR_REASON_DUP_PARA_CHAR_FRAC:
        r.append(f"dup_para_char_frac (ratio {_f2(float(int(st['para_dup_bytes'])) / tcl)}, max {_f2(p.dup_para_char_frac)})")
    if bits & _abi.TB_GR_REASON_DUP_LINE_FRAC:
        r.append(f"dup_line_frac (ratio {_f2(float(int(st['line_dup_elems'])) / float(max(int(st['n_lines']), 1)))}, max {_f2(p.dup_line_frac)})")
    if bits & _abi.TB_GR_REASON_DUP_LINE_CHAR_FRAC:
        r.append(f"dup_line_char_frac (ratio {_f2(float(int(st['line_dup_bytes'])) / tcl)}, max {_f2(p.dup_line_char_frac)})")
    for i, (n, frac) in enumerate(p.top_n_grams):
        if bits & (1 << (_abi.TB_GR_REASON_TOP_SHIFT + i)):
            r.append(f"top_{n}_gram (ratio {_f2(float(int(st['top_bytes'][i])) / tcl)}, max {_f2(frac)})")
    for i, (n, frac) in enumerate(p.dup_n_grams):
        if bits & (1 << (_abi.TB_GR_REASON_DUP_SHIFT + i)):
            r.append(f"duplicated_{n}_n_grams (ratio {_f2(float(int(st['dup_bytes'][i])) / tcl)}, max {_f2(frac)})")
    return r


def gr_metadata(p: GopherRepetitionParams, st) -> Tuple[Dict[str, str], str]:
    bits = int(st["reason_bits"])
    if bits & _abi.TB_GR_REASON_EMPTY:  # note the singular key, gopher_rep.rs:65-68
        return {"gopher_repetition_filter_status": "filtered", "gopher_repetition_filter_reason": "skipping empty content"}, "skipping empty content"
    reasons = gr_reasons(p, st)
    if not reasons:
        return {"gopher_repetition_filter_status": "passed"}, ""
    rs = "; ".join(reasons)
    return {"gopher_repetition_filter_status": "filtered", "gopher_repetition_filter_reasons": rs}, rs


def _f4(v: float) -> str:
    return "%.4f" % v


def fw_metadata(p: FineWebQualityParams, st) -> Tuple[Dict[str, str], str]:
    """fineweb_quality.rs:79-225: the first failing check returns; nothing is written on pass.
    (metadata, reason): the metadata value of the empty case differs from its reason (:82-89)."""
    bits = int(st["reason_bits"])
    if not bits:
        return {}, ""
    nl = float(int(st["n_lines"]))
    if bits & _abi.TB_FW_REASON_EMPTY:
        return {"fineweb_filter_status": "filtered", "fineweb_filter_reason": "empty document"}, "empty"
    if bits & _abi.TB_FW_REASON_LINE_PUNCT:
        r = (f"line_punct_ratio: {_f4(float(int(st['n_stop_lines'])) / nl)} < threshold {_f4(p.line_punct_thr)} "
             f"(exclude_zero: {'true' if p.line_punct_exclude_zero else 'false'})")
    elif bits & _abi.TB_FW_REASON_SHORT_LINE:
        r = f"short_line_ratio: {_f4(float(int(st['n_short_lines'])) / nl)} > threshold {_f4(p.short_line_thr)}"
    elif bits & _abi.TB_FW_REASON_CHAR_DUP:
        r = f"char_dup_ratio: {_f4(float(int(st['dup_line_bytes'])) / float(int(st['n_chars_no_newline'])))} > threshold {_f4(p.char_duplicates_ratio)}"
    elif bits & _abi.TB_FW_REASON_LIST_NO_WORDS:
        r = "list_ratio_no_words (newlines present but no words)"
    else:
        r = f"list_ratio: {_f4(float(int(st['n_newlines'])) / float(int(st['n_words'])))} > threshold {_f4(p.new_line_ratio)}"
    return {"fineweb_filter_status": "filtered", "fineweb_filter_reason": r}, r


def step_metadata(params, st) -> Tuple[Dict[str, str], str]:
    if isinstance(params, C4QualityParams):
        return c4_metadata(params, st)
    if isinstance(params, GopherQualityParams):
        return gq_metadata(params, st)
    if isinstance(params, GopherRepetitionParams):
        return gr_metadata(params, st)
    if isinstance(params, FineWebQualityParams):
        return fw_metadata(params, st)
    raise TypeError(type(params))
