"""Step configurations shared by the emulation (CPU) and CUDA (GPU) parity tests."""
from textblaster_b200.config import C4QualityParams, FineWebQualityParams, GopherQualityParams, GopherRepetitionParams

# permissive: (almost) nothing is filtered, so every statistic of every step is compared
GQ_OPEN = GopherQualityParams(min_doc_words=1, max_doc_words=1000000, min_avg_word_length=0.01, max_avg_word_length=5000.0,
                              max_symbol_word_ratio=1000.0, max_bullet_lines_ratio=1.0, max_ellipsis_lines_ratio=1.0,
                              max_non_alpha_words_ratio=0.0, min_stop_words=0)
# shipped thresholds (config/pipeline_config.yaml:23-32) + a stop list exercising case mapping
GQ_SHIPPED = GopherQualityParams(min_doc_words=5, max_doc_words=200, min_avg_word_length=3.0, max_avg_word_length=10.0,
                                 max_symbol_word_ratio=0.1, max_bullet_lines_ratio=0.9, max_ellipsis_lines_ratio=0.3,
                                 max_non_alpha_words_ratio=0.8, min_stop_words=2,
                                 stop_words=["the", "of", "café", "σοφος", "i̇stanbul", "kelvin", "a", "i",
                                             "thisisaverylongstopwordxx", "og", "på"])
# a short stop list without one-byte entries (the second-byte prefilter and the mask-domain length test of the tile path
# are live): mixed case, non-ASCII second bytes, digits
GQ_SHORT_SW = GopherQualityParams(min_doc_words=1, max_doc_words=1000000, min_avg_word_length=0.01, max_avg_word_length=5000.0,
                                  max_symbol_word_ratio=1000.0, max_bullet_lines_ratio=1.0, max_ellipsis_lines_ratio=1.0,
                                  max_non_alpha_words_ratio=0.0, min_stop_words=0,
                                  stop_words=["the", "Of", "på", "og", "to", "it's", "x9", "9x", "http", "foo_bar", "café"])
GQ_DEFAULT_SW = GopherQualityParams(min_doc_words=50, max_doc_words=100000, min_avg_word_length=3.0, max_avg_word_length=10.0,
                                    max_symbol_word_ratio=0.1, max_bullet_lines_ratio=0.9, max_ellipsis_lines_ratio=0.3,
                                    max_non_alpha_words_ratio=0.8, min_stop_words=2)
GR_OPEN = GopherRepetitionParams(top_n_grams=[(2, 1.0), (3, 1.0), (4, 1.0), (1, 1.0)],
                                 dup_n_grams=[(n, 1.0) for n in (5, 6, 7, 8, 9, 10, 2, 1)])
# shipped thresholds (config/pipeline_config.yaml:6-21)
GR_SHIPPED = GopherRepetitionParams(dup_line_frac=0.3, dup_para_frac=0.3, dup_line_char_frac=0.2, dup_para_char_frac=0.2,
                                    top_n_grams=[(2, 0.2), (3, 0.18), (4, 0.16)],
                                    dup_n_grams=[(5, 0.15), (6, 0.14), (7, 0.13), (8, 0.12), (9, 0.11), (10, 0.10)])
C4_SHIPPED = C4QualityParams(True, True, True, 5, 3, 1000, True, True, True, True)  # pipeline_config.yaml:56-66
C4_ONE_SENT = C4QualityParams(True, True, True, 1, 3, 1000, True, True, True, True)
C4_LAX = C4QualityParams(True, False, False, 1, 1, 50, False, False, False, False)
C4_NOPUNCT = C4QualityParams(True, True, False, 1, 2, 1000, False, True, False, True)
# split_paragraph = false: lines are the pieces of split_into_sentences (c4_filters.rs:152-155)
C4_SENT = C4QualityParams(False, True, True, 2, 3, 1000, True, True, True, True)
C4_SENT_LAX = C4QualityParams(False, False, False, 1, 1, 50, False, False, False, False)

# FineWebQualityFilter: shipped thresholds (config/pipeline_config.yaml:74-80), a permissive
# variant that lets every counter be compared, and a custom stop-character set with non-ASCII members
FW_SHIPPED = FineWebQualityParams(0.12, False, 0.67, 30, 0.01, 0.3)
FW_OPEN = FineWebQualityParams(0.0, False, 1.0, 30, 1.0, 1.0)
FW_CUSTOM = FineWebQualityParams(0.5, True, 0.9, 12, 0.5, 0.8, stop_chars=["\u3002", "a", "\u2026", ".", "\U0001f600"])
CHAINS = {
    "fw_shipped": [FW_SHIPPED],
    "fw_open": [FW_OPEN],
    "fw_custom": [FW_CUSTOM],
    "shipped_with_fineweb": [GR_SHIPPED, GQ_SHIPPED, C4_SHIPPED, FW_SHIPPED],  # the accelerated steps of pipeline_config.yaml, in its order
    "gq_open": [GQ_OPEN],
    "gq_shipped": [GQ_SHIPPED],
    "gq_short_sw": [GQ_SHORT_SW],
    "gr_open": [GR_OPEN],
    "gr_shipped": [GR_SHIPPED],
    "c4_one_sentence": [C4_ONE_SENT],
    "c4_lax": [C4_LAX],
    "c4_nopunct": [C4_NOPUNCT],
    "c4_shipped": [C4_SHIPPED],
    "c4_sentences": [C4_SENT],
    "c4_sentences_lax": [C4_SENT_LAX],
    "sentences_then_open": [C4_SENT_LAX, GQ_OPEN, GR_OPEN],
    "baseline_order": [C4_SHIPPED, GQ_DEFAULT_SW, GR_SHIPPED],      # BASELINE.json config 2 order
    "shipped_order": [GR_SHIPPED, GQ_SHIPPED, C4_SHIPPED],          # config/pipeline_config.yaml order
    "c4_then_open": [C4_ONE_SENT, GQ_OPEN, GR_OPEN],
    # several C4 steps in one chain (the reference takes any list of steps, worker_logic.rs:39-134): each reads the
    # output of the one before it; documents that drop out in between keep the content they had at that point
    "c4_twice": [C4_LAX, GQ_SHIPPED, C4_ONE_SENT],
    "c4_thrice": [C4_LAX, C4_NOPUNCT, GQ_SHIPPED, C4_SHIPPED, GR_OPEN],
}


def step_types(steps):
    return [s.to_c([]).type for s in steps]
