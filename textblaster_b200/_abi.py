"""ctypes mirror of include/textblaster_b200.h (struct layouts and constants only).

Kept free of any loading logic so that both the product binding (``_native.py``) and the test
oracle's loader can describe the same C structs.
"""
import ctypes as C

TB_ABI_VERSION = 3
TB_MAX_STEPS = 8
TB_MAX_NGRAMS = 12
TB_MAX_NGRAM_N = 32

TB_OK = 0
TB_ERR_INVALID_ARGUMENT = 1
TB_ERR_CONFIG_VALIDATION = 2
TB_ERR_CUDA = 3
TB_ERR_OUT_OF_MEMORY = 4
TB_ERR_UNSUPPORTED = 5
TB_ERR_INVALID_UTF8 = 6

TB_STEP_C4_QUALITY = 1
TB_STEP_GOPHER_QUALITY = 2
TB_STEP_GOPHER_REPETITION = 3
TB_STEP_FINEWEB_QUALITY = 4
TB_FW_MAX_STOP_CHARS = 32

TB_GQ_HAS_MIN_DOC_WORDS = 1 << 0
TB_GQ_HAS_MAX_DOC_WORDS = 1 << 1
TB_GQ_HAS_MIN_AVG_WORD_LENGTH = 1 << 2
TB_GQ_HAS_MAX_AVG_WORD_LENGTH = 1 << 3
TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO = 1 << 4
TB_GQ_HAS_MAX_BULLET_LINES_RATIO = 1 << 5
TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO = 1 << 6
TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO = 1 << 7
TB_GQ_HAS_MIN_STOP_WORDS = 1 << 8
TB_GQ_HAS_STOP_WORDS = 1 << 9

TB_GR_HAS_DUP_LINE_FRAC = 1 << 0
TB_GR_HAS_DUP_PARA_FRAC = 1 << 1
TB_GR_HAS_DUP_LINE_CHAR_FRAC = 1 << 2
TB_GR_HAS_DUP_PARA_CHAR_FRAC = 1 << 3

TB_C4_REASON_LOREM_IPSUM = 1 << 0
TB_C4_REASON_CURLY_BRACKET = 1 << 1
TB_C4_REASON_TOO_FEW_SENTENCES = 1 << 2

TB_GQ_REASON_SHORT_DOC = 1 << 0
TB_GQ_REASON_LONG_DOC = 1 << 1
TB_GQ_REASON_BELOW_AVG = 1 << 2
TB_GQ_REASON_ABOVE_AVG = 1 << 3
TB_GQ_REASON_TOO_MANY_HASHES = 1 << 4
TB_GQ_REASON_TOO_MANY_ELLIPSIS_UNITS = 1 << 5
TB_GQ_REASON_TOO_MANY_BULLETS = 1 << 6
TB_GQ_REASON_TOO_MANY_END_ELLIPSIS_LINES = 1 << 7
TB_GQ_REASON_BELOW_ALPHA = 1 << 8
TB_GQ_REASON_TOO_FEW_STOP_WORDS = 1 << 9

TB_GR_REASON_DUP_PARA_FRAC = 1 << 0
TB_GR_REASON_DUP_PARA_CHAR_FRAC = 1 << 1
TB_GR_REASON_DUP_LINE_FRAC = 1 << 2
TB_GR_REASON_DUP_LINE_CHAR_FRAC = 1 << 3
TB_GR_REASON_TOP_SHIFT = 4
TB_GR_REASON_DUP_SHIFT = 4 + TB_MAX_NGRAMS
TB_GR_REASON_EMPTY = 1 << 31

TB_FW_REASON_EMPTY = 1 << 0
TB_FW_REASON_LINE_PUNCT = 1 << 1
TB_FW_REASON_SHORT_LINE = 1 << 2
TB_FW_REASON_CHAR_DUP = 1 << 3
TB_FW_REASON_LIST_NO_WORDS = 1 << 4
TB_FW_REASON_LIST_RATIO = 1 << 5

TB_DOCFLAG_COMPLEX_SCRIPT = 1 << 0
TB_DOCFLAG_BAD_OFFSETS = 1 << 1

TB_OPT_RETURN_CONTENT = 1
TB_OPT_VALIDATE_UTF8 = 2
TB_OPT_CHUNK_BYTES = 3
TB_OPT_GENERAL_WORD_PATH = 4
TB_OPT_DECODE_HTML_ENTITIES = 5
TB_OPT_GATHER_KEPT = 6
TB_OPT_CHUNK_RAMP = 7
TB_OPT_STAGE_THREADS = 8
TB_OPT_BULK_STAGE_TILES = 9


class tb_c4_params(C.Structure):
    _fields_ = [
        ("split_paragraph", C.c_int32),
        ("remove_citations", C.c_int32),
        ("filter_no_terminal_punct", C.c_int32),
        ("filter_lorem_ipsum", C.c_int32),
        ("filter_javascript", C.c_int32),
        ("filter_curly_bracket", C.c_int32),
        ("filter_policy", C.c_int32),
        ("_pad", C.c_int32),
        ("min_num_sentences", C.c_uint64),
        ("min_words_per_line", C.c_uint64),
        ("max_word_length", C.c_uint64),
    ]


class tb_gq_params(C.Structure):
    _fields_ = [
        ("has", C.c_uint32),
        ("n_stop_words", C.c_int32),
        ("min_doc_words", C.c_uint64),
        ("max_doc_words", C.c_uint64),
        ("min_avg_word_length", C.c_double),
        ("max_avg_word_length", C.c_double),
        ("max_symbol_word_ratio", C.c_double),
        ("max_bullet_lines_ratio", C.c_double),
        ("max_ellipsis_lines_ratio", C.c_double),
        ("max_non_alpha_words_ratio", C.c_double),
        ("min_stop_words", C.c_uint64),
        ("stop_words_utf8", C.c_void_p),
        ("stop_words_offsets", C.c_void_p),
    ]


class tb_gr_params(C.Structure):
    _fields_ = [
        ("has", C.c_uint32),
        ("n_top", C.c_int32),
        ("n_dup", C.c_int32),
        ("_pad", C.c_int32),
        ("dup_line_frac", C.c_double),
        ("dup_para_frac", C.c_double),
        ("dup_line_char_frac", C.c_double),
        ("dup_para_char_frac", C.c_double),
        ("top_n", C.c_uint32 * TB_MAX_NGRAMS),
        ("top_frac", C.c_double * TB_MAX_NGRAMS),
        ("dup_n", C.c_uint32 * TB_MAX_NGRAMS),
        ("dup_frac", C.c_double * TB_MAX_NGRAMS),
    ]


class tb_fw_params(C.Structure):
    _fields_ = [
        ("line_punct_thr", C.c_double),
        ("line_punct_exclude_zero", C.c_int32),
        ("n_stop_chars", C.c_int32),
        ("stop_chars", C.c_uint32 * TB_FW_MAX_STOP_CHARS),
        ("short_line_thr", C.c_double),
        ("short_line_length", C.c_uint64),
        ("char_duplicates_ratio", C.c_double),
        ("new_line_ratio", C.c_double),
    ]


class _tb_step_union(C.Union):
    _fields_ = [("c4", tb_c4_params), ("gq", tb_gq_params), ("gr", tb_gr_params), ("fw", tb_fw_params)]


class tb_step_cfg(C.Structure):
    _fields_ = [("type", C.c_int32), ("_pad", C.c_int32), ("u", _tb_step_union)]


class tb_c4_stats(C.Structure):
    _fields_ = [
        ("reason_bits", C.c_uint32),
        ("n_lines", C.c_uint32),
        ("too_long_word", C.c_uint32),
        ("no_terminal_punc", C.c_uint32),
        ("too_few_words", C.c_uint32),
        ("n_kept_lines", C.c_uint32),
        ("n_sentences", C.c_uint32),
        ("rewritten", C.c_uint32),
    ]


class tb_gq_stats(C.Structure):
    _fields_ = [
        ("reason_bits", C.c_uint32),
        ("n_words", C.c_uint32),
        ("sum_word_chars", C.c_uint64),
        ("n_hash", C.c_uint32),
        ("n_ellipsis_units", C.c_uint32),
        ("n_lines", C.c_uint32),
        ("n_bullet_lines", C.c_uint32),
        ("n_ellipsis_lines", C.c_uint32),
        ("n_alpha_words", C.c_uint32),
        ("n_stop_words", C.c_uint32),
        ("flags", C.c_uint32),
    ]


class tb_gr_stats(C.Structure):
    _fields_ = [
        ("reason_bits", C.c_uint32),
        ("flags", C.c_uint32),
        ("char_len", C.c_uint64),
        ("n_paras", C.c_uint32),
        ("para_dup_elems", C.c_uint32),
        ("para_dup_bytes", C.c_uint64),
        ("n_lines", C.c_uint32),
        ("line_dup_elems", C.c_uint32),
        ("line_dup_bytes", C.c_uint64),
        ("n_words", C.c_uint64),
        ("top_bytes", C.c_uint64 * TB_MAX_NGRAMS),
        ("dup_bytes", C.c_uint64 * TB_MAX_NGRAMS),
    ]


class tb_fw_stats(C.Structure):
    _fields_ = [
        ("reason_bits", C.c_uint32),
        ("flags", C.c_uint32),
        ("n_lines", C.c_uint32),
        ("n_stop_lines", C.c_uint32),
        ("n_short_lines", C.c_uint32),
        ("n_words", C.c_uint32),
        ("n_newlines", C.c_uint32),
        ("_pad", C.c_uint32),
        ("dup_line_bytes", C.c_uint64),
        ("n_chars_no_newline", C.c_uint64),
    ]


class tb_run(C.Structure):
    _fields_ = [("begin", C.c_uint32), ("end", C.c_uint32)]


class tb_batch_out(C.Structure):
    _fields_ = [
        ("n_docs", C.c_int64),
        ("n_steps", C.c_int32),
        ("content_rewritten", C.c_int32),
        ("keep", C.POINTER(C.c_uint8)),
        ("failed_step", C.POINTER(C.c_int8)),
        ("step_stats", C.c_void_p * TB_MAX_STEPS),
        ("content", C.POINTER(C.c_uint8)),
        ("content_offsets", C.POINTER(C.c_int64)),
        ("content_bytes", C.c_int64),
        ("gpu_ms", C.c_double),
        ("gpu_launches", C.c_int32),
        ("content_kind", C.c_int32),
        ("_owner", C.c_void_p),
        ("runs", C.POINTER(tb_run)),
        ("run_offsets", C.POINTER(C.c_int64)),
        ("n_runs", C.c_int64),
        ("kept_content", C.POINTER(C.c_uint8)),
        ("kept_offsets", C.POINTER(C.c_int64)),
        ("n_kept", C.c_int64),
    ]


STATS_STRUCT = {
    TB_STEP_C4_QUALITY: tb_c4_stats,
    TB_STEP_GOPHER_QUALITY: tb_gq_stats,
    TB_STEP_GOPHER_REPETITION: tb_gr_stats,
    TB_STEP_FINEWEB_QUALITY: tb_fw_stats,
}

STEP_NAMES = {
    TB_STEP_C4_QUALITY: "C4QualityFilter",
    TB_STEP_GOPHER_QUALITY: "GopherQualityFilter",
    TB_STEP_GOPHER_REPETITION: "GopherRepetitionFilter",
    TB_STEP_FINEWEB_QUALITY: "FineWebQualityFilter",
}


def stats_dtype(step_type):
    """numpy structured dtype matching the C stats struct of ``step_type``."""
    import numpy as np

    return np.dtype(STATS_STRUCT[step_type])
