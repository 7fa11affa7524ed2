"""pipeline_config.yaml parsing and validation -- mirrors the reference's tests/config_tests.rs
(cited per test) for the three accelerated step types and the loader."""
import os

import pytest

from textblaster_b200.config import (C4QualityParams, GopherQualityParams, GopherRepetitionParams, PipelineConfig,
                                     load_pipeline_config, rust_f64_display)
from textblaster_b200.error import ConfigError, ConfigValidationError

VALID = """
pipeline:
  - type: C4QualityFilter
    split_paragraph: false
    remove_citations: true
    filter_no_terminal_punct: true
    min_num_sentences: 5
    min_words_per_line: 3
    max_word_length: 15
    filter_lorem_ipsum: true
    filter_javascript: true
    filter_curly_bracket: true
    filter_policy: true
  - type: GopherRepetitionFilter
    dup_line_frac: 0.20
    top_n_grams: [[2, 0.2], [3, 0.18]]
"""


def _write(tmp_path, text):
    p = tmp_path / "cfg.yaml"
    p.write_text(text, encoding="utf-8")
    return str(p)


def test_load_valid_config(tmp_path):  # config_tests.rs:17-58
    cfg = load_pipeline_config(_write(tmp_path, VALID))
    assert len(cfg.pipeline) == 2
    assert isinstance(cfg.pipeline[0].params, C4QualityParams) and cfg.pipeline[0].params.min_num_sentences == 5
    p = cfg.pipeline[1].params
    assert isinstance(p, GopherRepetitionParams) and p.dup_line_frac == 0.20 and len(p.top_n_grams) == 2
    assert cfg.pipeline[0].name() == "C4QualityFilter" and cfg.pipeline[1].name() == "GopherRepetitionFilter"


def test_file_not_found():  # :61-71
    with pytest.raises(ConfigError) as e:
        load_pipeline_config("non_existent_config.yaml")
    assert "Failed to read pipeline config file" in str(e.value) and "non_existent_config.yaml" in str(e.value)


def test_invalid_yaml_syntax(tmp_path):  # :74-93
    with pytest.raises(ConfigError) as e:
        load_pipeline_config(_write(tmp_path, "pipeline:\n  - type: C4QualityFilter\n    split_paragraph: [unclosed\n"))
    assert "Failed to parse pipeline config YAML" in str(e.value)


def test_unknown_step_type(tmp_path):  # :96-114
    with pytest.raises(ConfigError) as e:
        load_pipeline_config(_write(tmp_path, "pipeline:\n  - type: UnknownFilterType\n    x: 1\n"))
    assert "Failed to parse pipeline config YAML" in str(e.value)
    assert "UnknownFilterType" in str(e.value) or "unknown variant" in str(e.value)


def test_missing_pipeline_field(tmp_path):  # :118-138 (the README's `steps:` layout is rejected)
    with pytest.raises(ConfigError) as e:
        load_pipeline_config(_write(tmp_path, "steps:\n  - type: C4QualityFilter\n"))
    assert "Failed to parse pipeline config YAML" in str(e.value) and "missing field `pipeline`" in str(e.value)


def test_empty_pipeline_is_valid(tmp_path):  # :141-153
    assert load_pipeline_config(_write(tmp_path, "pipeline: []\n")).pipeline == []


def c4(**kw):
    d = dict(split_paragraph=True, remove_citations=True, filter_no_terminal_punct=True, min_num_sentences=5,
             min_words_per_line=3, max_word_length=1000, filter_lorem_ipsum=True, filter_javascript=True,
             filter_curly_bracket=True, filter_policy=True)
    d.update(kw)
    return C4QualityParams(**d)


def test_c4_params():  # :213-243
    c4().validate()
    for field in ("min_num_sentences", "min_words_per_line", "max_word_length"):
        with pytest.raises(ConfigValidationError) as e:
            c4(**{field: 0}).validate()
        assert field in str(e.value) and "must be greater than 0" in str(e.value)


def test_gopher_repetition_params():  # :258-286
    GopherRepetitionParams(0.3, 0.3, 0.2, 0.2, [(2, 0.2)], [(5, 0.15)]).validate()
    with pytest.raises(ConfigValidationError) as e:
        GopherRepetitionParams(dup_line_frac=1.5).validate()
    assert str(e.value) == "Configuration validation error: GopherRepetitionParams: dup_line_frac must be between 0.0 and 1.0, got 1.5"
    with pytest.raises(ConfigValidationError) as e:
        GopherRepetitionParams(dup_para_frac=-0.1).validate()
    assert "dup_para_frac" in str(e.value) and "got -0.1" in str(e.value)
    with pytest.raises(ConfigValidationError) as e:
        GopherRepetitionParams(top_n_grams=[(0, 0.2)]).validate()
    assert "n-gram size in top_n_grams at index 0 must be greater than 0" in str(e.value)
    with pytest.raises(ConfigValidationError) as e:
        GopherRepetitionParams(dup_n_grams=[(5, 0.1), (6, 2.0)]).validate()
    assert "n-gram fraction in dup_n_grams at index 1 must be between 0.0 and 1.0, got 2" in str(e.value)


def test_gopher_quality_params():  # :305-382
    GopherQualityParams(50, 100000, 3.0, 10.0, 0.1, 0.9, 0.3, 0.8, 2, None).validate()
    cases = [(dict(min_doc_words=0), "min_doc_words must be greater than 0"),
             (dict(max_doc_words=0), "max_doc_words must be greater than 0"),
             (dict(min_doc_words=10, max_doc_words=5), "min_doc_words (10) cannot be greater than max_doc_words (5)"),
             (dict(min_avg_word_length=0.0), "min_avg_word_length must be greater than 0.0"),
             (dict(max_avg_word_length=0.0), "max_avg_word_length must be greater than 0.0"),
             (dict(min_avg_word_length=10.0, max_avg_word_length=5.0), "min_avg_word_length (10) cannot be greater than max_avg_word_length (5)"),
             (dict(max_symbol_word_ratio=-0.1), "max_symbol_word_ratio must be non-negative, got -0.1")]
    for kw, msg in cases:
        with pytest.raises(ConfigValidationError) as e:
            GopherQualityParams(**kw).validate()
        assert msg in str(e.value)


def test_load_rejects_invalid_step(tmp_path):  # :538-556
    bad = VALID.replace("min_num_sentences: 5", "min_num_sentences: 0")
    with pytest.raises(ConfigValidationError) as e:
        load_pipeline_config(_write(tmp_path, bad))
    assert "min_num_sentences" in str(e.value)


def test_shipped_config_layout(tmp_path):
    """The layout of the reference's config/pipeline_config.yaml: all seven step types parse and validate; the
    out-of-scope ones are flagged (restated here, not read from the reference tree)."""
    text = """
pipeline:
  - type: LanguageDetectionFilter
    min_confidence: 0.65
    allowed_languages: [ "dan" ]
  - type: GopherRepetitionFilter
    dup_line_frac: 0.3
    dup_para_frac: 0.3
    dup_line_char_frac: 0.2
    dup_para_char_frac: 0.2
    top_n_grams:
      - [2, 0.2]
      - [3, 0.18]
      - [4, 0.16]
    dup_n_grams:
      - [5, 0.15]
      - [10, 0.10]
  - type: GopherQualityFilter
    min_doc_words: 50
    max_doc_words: 100000
    min_avg_word_length: 3.0
    max_avg_word_length: 10.0
    max_symbol_word_ratio: 0.1
    max_bullet_lines_ratio: 0.9
    max_ellipsis_lines_ratio: 0.3
    max_non_alpha_words_ratio: 0.8
    min_stop_words: 2
    stop_words: ["ad", "af", "på", "fordi", "fordi"]
  - type: C4QualityFilter
    split_paragraph: true
    remove_citations: true
    filter_no_terminal_punct: true
    min_num_sentences: 5
    min_words_per_line: 3
    max_word_length: 1000
    filter_lorem_ipsum: true
    filter_javascript: true
    filter_curly_bracket: true
    filter_policy: true
  - type: FineWebQualityFilter
    line_punct_thr: 0.12
    line_punct_exclude_zero: false
    short_line_thr: 0.67
    short_line_length: 30
    char_duplicates_ratio: 0.01
    new_line_ratio: 0.3
  - type: TokenCounter
    tokenizer_name: "gpt2"
"""
    cfg = load_pipeline_config(_write(tmp_path, text))
    assert [s.name() for s in cfg.pipeline] == ["LanguageDetectionFilter", "GopherRepetitionFilter", "GopherQualityFilter",
                                                "C4QualityFilter", "FineWebQualityFilter", "TokenCounter"]
    assert [s.in_scope for s in cfg.pipeline] == [False, True, True, True, True, False]
    gq = cfg.pipeline[2].params
    assert gq.stop_words == ["ad", "af", "på", "fordi", "fordi"] and gq.min_stop_words == 2
    assert cfg.pipeline[1].params.dup_n_grams == [(5, 0.15), (10, 0.10)]


def test_out_of_scope_validation(tmp_path):  # :443-466, :489-512, :529-533, :559-580
    with pytest.raises(ConfigValidationError) as e:
        load_pipeline_config(_write(tmp_path, "pipeline:\n  - type: LanguageDetectionFilter\n    min_confidence: 1.5\n    allowed_languages: [dan]\n"))
    assert "min_confidence" in str(e.value)
    with pytest.raises(ConfigValidationError) as e:
        load_pipeline_config(_write(tmp_path, "pipeline:\n  - type: TokenCounter\n    tokenizer_name: \"\"\n"))
    assert "tokenizer_name" in str(e.value)


def test_rust_f64_display():
    assert rust_f64_display(1.0) == "1" and rust_f64_display(1.5) == "1.5" and rust_f64_display(-0.1) == "-0.1"
    assert rust_f64_display(0.30000000000000004) == "0.30000000000000004" and rust_f64_display(1e21) == "1000000000000000000000"
    assert rust_f64_display(1e-7) == "0.0000001"


def test_fineweb_params_validate_and_marshal():  # src/config/pipeline.rs:311-348
    from textblaster_b200 import _abi
    from textblaster_b200.config import FineWebQualityParams
    from textblaster_b200.error import ConfigValidationError
    ok = FineWebQualityParams.from_dict(dict(line_punct_thr=0.12, line_punct_exclude_zero=False, short_line_thr=0.67,
                                             short_line_length=30, char_duplicates_ratio=0.01, new_line_ratio=0.3))
    ok.validate()
    c = ok.to_c([])
    assert c.type == _abi.TB_STEP_FINEWEB_QUALITY and c.u.fw.n_stop_chars == -1 and c.u.fw.short_line_length == 30
    with pytest.raises(ConfigValidationError, match="FineWebQualityFilterParams: new_line_ratio must be between 0.0 and 1.0, got 1.5"):
        FineWebQualityParams(0.1, False, 0.5, 30, 0.1, 1.5).validate()
    with pytest.raises(ConfigValidationError, match="short_line_length must be greater than 0"):
        FineWebQualityParams(0.1, False, 0.5, 0, 0.1, 0.5).validate()
    custom = FineWebQualityParams(0.1, True, 0.5, 10, 0.1, 0.5, stop_chars=["\u3002", ".", "."]).to_c([])
    assert custom.u.fw.n_stop_chars == 2 and sorted(custom.u.fw.stop_chars[:2]) == [ord("."), 0x3002]
