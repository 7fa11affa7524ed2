"""Pipeline configuration: the reference's ``pipeline_config.yaml`` format, unchanged.

Mirrors ``src/config/pipeline.rs`` of the reference:
  * ``PipelineConfig { pipeline: Vec<StepConfig> }``                       (:10-22)
  * ``#[serde(tag = "type")] enum StepConfig`` with flattened params      (:26-37)
  * ``C4QualityParams`` / ``GopherRepetitionParams`` / ``GopherQualityParams`` and their
    ``validate()`` rules and messages                                      (:66-100, :102-159, :161-258)
  * ``load_pipeline_config``                                               (:372-393)

The four step types outside the accelerated path (LanguageDetectionFilter, C4BadWordsFilter,
TokenCounter) still parse and validate with the reference's rules so the
shipped ``config/pipeline_config.yaml`` loads; building a GPU pipeline from them is refused
unless the caller asks to skip them.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

from . import _abi
from .error import ConfigError, ConfigValidationError


def rust_f64_display(v: float) -> str:
    """``format!("{}", v)`` for an f64: shortest round-trip digits, never an exponent, and no
    trailing ``.0`` for integral values (Rust prints ``1`` for ``1.0``)."""
    if v != v:
        return "NaN"
    if v in (float("inf"), float("-inf")):
        return "inf" if v > 0 else "-inf"
    r = repr(float(v))
    if "e" in r or "E" in r:
        from decimal import Decimal

        r = format(Decimal(r), "f")
    if r.endswith(".0"):
        r = r[:-2]
    return r


def _req(d: dict, key: str, typ, ctx: str):
    if key not in d:
        raise ConfigError(f"missing field `{key}`")
    return _conv(d[key], typ, key, ctx)


def _conv(v, typ, key, ctx):
    if typ is bool:
        if not isinstance(v, bool):
            raise ConfigError(f"{ctx}: invalid type for `{key}`: expected a boolean")
        return v
    if typ is int:  # usize
        if isinstance(v, bool) or not isinstance(v, int) or v < 0:
            raise ConfigError(f"{ctx}: invalid type for `{key}`: expected an unsigned integer")
        return v
    if typ is float:
        if isinstance(v, bool) or not isinstance(v, (int, float)):
            raise ConfigError(f"{ctx}: invalid type for `{key}`: expected a float")
        return float(v)
    if typ is str:
        if not isinstance(v, str):
            raise ConfigError(f"{ctx}: invalid type for `{key}`: expected a string")
        return v
    raise AssertionError(typ)


def _opt(d: dict, key: str, typ, ctx: str):
    v = d.get(key)
    if v is None:
        return None
    return _conv(v, typ, key, ctx)


@dataclass
class C4QualityParams:
    """src/config/pipeline.rs:66-79"""

    split_paragraph: bool
    remove_citations: bool
    filter_no_terminal_punct: bool
    min_num_sentences: int
    min_words_per_line: int
    max_word_length: int
    filter_lorem_ipsum: bool
    filter_javascript: bool
    filter_curly_bracket: bool
    filter_policy: bool

    TYPE = "C4QualityFilter"
    CTYPE = _abi.TB_STEP_C4_QUALITY

    @classmethod
    def from_dict(cls, d: dict) -> "C4QualityParams":
        ctx = cls.TYPE
        return cls(
            _req(d, "split_paragraph", bool, ctx), _req(d, "remove_citations", bool, ctx),
            _req(d, "filter_no_terminal_punct", bool, ctx), _req(d, "min_num_sentences", int, ctx),
            _req(d, "min_words_per_line", int, ctx), _req(d, "max_word_length", int, ctx),
            _req(d, "filter_lorem_ipsum", bool, ctx), _req(d, "filter_javascript", bool, ctx),
            _req(d, "filter_curly_bracket", bool, ctx), _req(d, "filter_policy", bool, ctx))

    def validate(self) -> None:
        """src/config/pipeline.rs:82-99"""
        if self.min_num_sentences == 0:
            raise ConfigValidationError("C4QualityParams: min_num_sentences must be greater than 0")
        if self.min_words_per_line == 0:
            raise ConfigValidationError("C4QualityParams: min_words_per_line must be greater than 0")
        if self.max_word_length == 0:
            raise ConfigValidationError("C4QualityParams: max_word_length must be greater than 0")

    def to_c(self, keepalive: list) -> _abi.tb_step_cfg:
        s = _abi.tb_step_cfg()
        s.type = _abi.TB_STEP_C4_QUALITY
        p = s.u.c4
        for k in ("split_paragraph", "remove_citations", "filter_no_terminal_punct", "filter_lorem_ipsum",
                  "filter_javascript", "filter_curly_bracket", "filter_policy"):
            setattr(p, k, 1 if getattr(self, k) else 0)
        p.min_num_sentences = self.min_num_sentences
        p.min_words_per_line = self.min_words_per_line
        p.max_word_length = self.max_word_length
        return s


@dataclass
class GopherRepetitionParams:
    """src/config/pipeline.rs:102-115"""

    dup_line_frac: Optional[float] = None
    dup_para_frac: Optional[float] = None
    dup_line_char_frac: Optional[float] = None
    dup_para_char_frac: Optional[float] = None
    top_n_grams: List[Tuple[int, float]] = field(default_factory=list)
    dup_n_grams: List[Tuple[int, float]] = field(default_factory=list)

    TYPE = "GopherRepetitionFilter"
    CTYPE = _abi.TB_STEP_GOPHER_REPETITION

    @classmethod
    def from_dict(cls, d: dict) -> "GopherRepetitionParams":
        ctx = cls.TYPE

        def grams(key):
            v = d.get(key)
            if v is None:
                return []
            if not isinstance(v, list):
                raise ConfigError(f"{ctx}: invalid type for `{key}`: expected a sequence")
            out = []
            for item in v:
                if not isinstance(item, (list, tuple)) or len(item) != 2:
                    raise ConfigError(f"{ctx}: invalid type for `{key}`: expected a tuple of size 2")
                out.append((_conv(item[0], int, key, ctx), _conv(item[1], float, key, ctx)))
            return out

        return cls(_opt(d, "dup_line_frac", float, ctx), _opt(d, "dup_para_frac", float, ctx),
                   _opt(d, "dup_line_char_frac", float, ctx), _opt(d, "dup_para_char_frac", float, ctx),
                   grams("top_n_grams"), grams("dup_n_grams"))

    def validate(self) -> None:
        """src/config/pipeline.rs:118-158"""
        for name in ("dup_line_frac", "dup_para_frac", "dup_line_char_frac", "dup_para_char_frac"):
            v = getattr(self, name)
            if v is not None and not (0.0 <= v <= 1.0):
                raise ConfigValidationError(
                    f"GopherRepetitionParams: {name} must be between 0.0 and 1.0, got {rust_f64_display(v)}")
        for name in ("top_n_grams", "dup_n_grams"):
            for idx, (size, fraction) in enumerate(getattr(self, name)):
                if size == 0:
                    raise ConfigValidationError(
                        f"GopherRepetitionParams: n-gram size in {name} at index {idx} must be greater than 0")
                if not (0.0 <= fraction <= 1.0):
                    raise ConfigValidationError(
                        f"GopherRepetitionParams: n-gram fraction in {name} at index {idx} must be between "
                        f"0.0 and 1.0, got {rust_f64_display(fraction)}")

    def to_c(self, keepalive: list) -> _abi.tb_step_cfg:
        s = _abi.tb_step_cfg()
        s.type = _abi.TB_STEP_GOPHER_REPETITION
        p = s.u.gr
        has = 0
        for bit, name in ((_abi.TB_GR_HAS_DUP_LINE_FRAC, "dup_line_frac"), (_abi.TB_GR_HAS_DUP_PARA_FRAC, "dup_para_frac"),
                          (_abi.TB_GR_HAS_DUP_LINE_CHAR_FRAC, "dup_line_char_frac"),
                          (_abi.TB_GR_HAS_DUP_PARA_CHAR_FRAC, "dup_para_char_frac")):
            v = getattr(self, name)
            if v is not None:
                has |= bit
                setattr(p, name, v)
        p.has = has
        if len(self.top_n_grams) > _abi.TB_MAX_NGRAMS or len(self.dup_n_grams) > _abi.TB_MAX_NGRAMS:
            raise ConfigError(f"GopherRepetitionFilter: at most {_abi.TB_MAX_NGRAMS} entries per n-gram list are supported")
        p.n_top = len(self.top_n_grams)
        p.n_dup = len(self.dup_n_grams)
        for i, (n, f) in enumerate(self.top_n_grams):
            p.top_n[i] = n
            p.top_frac[i] = f
        for i, (n, f) in enumerate(self.dup_n_grams):
            p.dup_n[i] = n
            p.dup_frac[i] = f
        return s


@dataclass
class GopherQualityParams:
    """src/config/pipeline.rs:161-176"""

    min_doc_words: Optional[int] = None
    max_doc_words: Optional[int] = None
    min_avg_word_length: Optional[float] = None
    max_avg_word_length: Optional[float] = None
    max_symbol_word_ratio: Optional[float] = None
    max_bullet_lines_ratio: Optional[float] = None
    max_ellipsis_lines_ratio: Optional[float] = None
    max_non_alpha_words_ratio: Optional[float] = None
    min_stop_words: Optional[int] = None
    stop_words: Optional[List[str]] = None

    TYPE = "GopherQualityFilter"
    CTYPE = _abi.TB_STEP_GOPHER_QUALITY

    @classmethod
    def from_dict(cls, d: dict) -> "GopherQualityParams":
        ctx = cls.TYPE
        sw = d.get("stop_words")
        if sw is not None:
            if not isinstance(sw, list) or not all(isinstance(x, str) for x in sw):
                raise ConfigError(f"{ctx}: invalid type for `stop_words`: expected a sequence of strings")
            sw = list(sw)
        return cls(_opt(d, "min_doc_words", int, ctx), _opt(d, "max_doc_words", int, ctx),
                   _opt(d, "min_avg_word_length", float, ctx), _opt(d, "max_avg_word_length", float, ctx),
                   _opt(d, "max_symbol_word_ratio", float, ctx), _opt(d, "max_bullet_lines_ratio", float, ctx),
                   _opt(d, "max_ellipsis_lines_ratio", float, ctx), _opt(d, "max_non_alpha_words_ratio", float, ctx),
                   _opt(d, "min_stop_words", int, ctx), sw)

    def validate(self) -> None:
        """src/config/pipeline.rs:179-257"""
        if self.min_doc_words is not None and self.min_doc_words == 0:
            raise ConfigValidationError("GopherQualityParams: min_doc_words must be greater than 0")
        if self.max_doc_words is not None and self.max_doc_words == 0:
            raise ConfigValidationError("GopherQualityParams: max_doc_words must be greater than 0")
        if self.min_doc_words is not None and self.max_doc_words is not None and self.min_doc_words > self.max_doc_words:
            raise ConfigValidationError(
                f"GopherQualityParams: min_doc_words ({self.min_doc_words}) cannot be greater than "
                f"max_doc_words ({self.max_doc_words})")
        if self.min_avg_word_length is not None and self.min_avg_word_length <= 0.0:
            raise ConfigValidationError("GopherQualityParams: min_avg_word_length must be greater than 0.0")
        if self.max_avg_word_length is not None and self.max_avg_word_length <= 0.0:
            raise ConfigValidationError("GopherQualityParams: max_avg_word_length must be greater than 0.0")
        if (self.min_avg_word_length is not None and self.max_avg_word_length is not None
                and self.min_avg_word_length > self.max_avg_word_length):
            raise ConfigValidationError(
                f"GopherQualityParams: min_avg_word_length ({rust_f64_display(self.min_avg_word_length)}) cannot be "
                f"greater than max_avg_word_length ({rust_f64_display(self.max_avg_word_length)})")
        for name in ("max_symbol_word_ratio", "max_bullet_lines_ratio", "max_ellipsis_lines_ratio",
                     "max_non_alpha_words_ratio"):
            v = getattr(self, name)
            if v is not None and v < 0.0:
                raise ConfigValidationError(
                    f"GopherQualityParams: {name} must be non-negative, got {rust_f64_display(v)}")

    def to_c(self, keepalive: list) -> _abi.tb_step_cfg:
        s = _abi.tb_step_cfg()
        s.type = _abi.TB_STEP_GOPHER_QUALITY
        p = s.u.gq
        has = 0
        for bit, name in ((_abi.TB_GQ_HAS_MIN_DOC_WORDS, "min_doc_words"), (_abi.TB_GQ_HAS_MAX_DOC_WORDS, "max_doc_words"),
                          (_abi.TB_GQ_HAS_MIN_AVG_WORD_LENGTH, "min_avg_word_length"),
                          (_abi.TB_GQ_HAS_MAX_AVG_WORD_LENGTH, "max_avg_word_length"),
                          (_abi.TB_GQ_HAS_MAX_SYMBOL_WORD_RATIO, "max_symbol_word_ratio"),
                          (_abi.TB_GQ_HAS_MAX_BULLET_LINES_RATIO, "max_bullet_lines_ratio"),
                          (_abi.TB_GQ_HAS_MAX_ELLIPSIS_LINES_RATIO, "max_ellipsis_lines_ratio"),
                          (_abi.TB_GQ_HAS_MAX_NON_ALPHA_WORDS_RATIO, "max_non_alpha_words_ratio"),
                          (_abi.TB_GQ_HAS_MIN_STOP_WORDS, "min_stop_words")):
            v = getattr(self, name)
            if v is not None:
                has |= bit
                setattr(p, name, v)
        if self.stop_words is not None:
            has |= _abi.TB_GQ_HAS_STOP_WORDS
            enc = [w.encode("utf-8") for w in self.stop_words]
            blob = b"".join(enc)
            offs = [0]
            for e in enc:
                offs.append(offs[-1] + len(e))
            buf = C.create_string_buffer(blob, max(len(blob), 1))
            arr = (C.c_int32 * len(offs))(*offs)
            keepalive.extend([buf, arr])
            p.n_stop_words = len(enc)
            p.stop_words_utf8 = C.cast(buf, C.c_void_p).value
            p.stop_words_offsets = C.cast(arr, C.c_void_p).value
        p.has = has
        return s


@dataclass
class FineWebQualityParams:
    """FineWebQualityFilterParams, src/config/pipeline.rs:311-322 (stop_chars: Option<HashSet<char>>)"""

    line_punct_thr: float
    line_punct_exclude_zero: bool
    short_line_thr: float
    short_line_length: int
    char_duplicates_ratio: float
    new_line_ratio: float
    stop_chars: Optional[List[str]] = None

    TYPE = "FineWebQualityFilter"
    CTYPE = _abi.TB_STEP_FINEWEB_QUALITY

    @classmethod
    def from_dict(cls, d: dict) -> "FineWebQualityParams":
        ctx = cls.TYPE
        sc = d.get("stop_chars")
        if sc is not None:
            if not isinstance(sc, (list, tuple)) or any(not isinstance(c, str) or len(c) != 1 for c in sc):
                raise ConfigError(f"{ctx}: invalid type for `stop_chars`: expected a sequence of single characters")
            sc = list(sc)
        return cls(_req(d, "line_punct_thr", float, ctx), _req(d, "line_punct_exclude_zero", bool, ctx),
                   _req(d, "short_line_thr", float, ctx), _req(d, "short_line_length", int, ctx),
                   _req(d, "char_duplicates_ratio", float, ctx), _req(d, "new_line_ratio", float, ctx), sc)

    def validate(self) -> None:
        """src/config/pipeline.rs:324-348"""
        for name in ("line_punct_thr", "short_line_thr", "char_duplicates_ratio", "new_line_ratio"):
            v = getattr(self, name)
            if not (0.0 <= v <= 1.0):
                raise ConfigValidationError(
                    f"FineWebQualityFilterParams: {name} must be between 0.0 and 1.0, got {rust_f64_display(v)}")
        if self.short_line_length == 0:
            raise ConfigValidationError("FineWebQualityFilterParams: short_line_length must be greater than 0")

    def to_c(self, keepalive: list) -> _abi.tb_step_cfg:
        s = _abi.tb_step_cfg()
        s.type = _abi.TB_STEP_FINEWEB_QUALITY
        p = s.u.fw
        p.line_punct_thr = self.line_punct_thr
        p.line_punct_exclude_zero = 1 if self.line_punct_exclude_zero else 0
        if self.stop_chars is None:
            p.n_stop_chars = -1
        else:
            chars = sorted(set(self.stop_chars))
            if len(chars) > _abi.TB_FW_MAX_STOP_CHARS:
                raise ConfigError(f"FineWebQualityFilter: at most {_abi.TB_FW_MAX_STOP_CHARS} stop_chars are supported")
            p.n_stop_chars = len(chars)
            for i, c in enumerate(chars):
                p.stop_chars[i] = ord(c)
        p.short_line_thr = self.short_line_thr
        p.short_line_length = self.short_line_length
        p.char_duplicates_ratio = self.char_duplicates_ratio
        p.new_line_ratio = self.new_line_ratio
        return s


@dataclass
class OutOfScopeParams:
    """A reference step type that parses and validates but has no GPU implementation here
    (LanguageDetectionFilter, C4BadWordsFilter, TokenCounter)."""

    type_name: str
    params: dict

    def validate(self) -> None:
        d, t = self.params, self.type_name
        if t == "LanguageDetectionFilter":  # src/config/pipeline.rs:294-308
            mc = d.get("min_confidence")
            if not (0.0 <= mc <= 1.0):
                raise ConfigValidationError(
                    f"LanguageDetectionParams: min_confidence must be between 0.0 and 1.0, got {rust_f64_display(mc)}")
            if not d.get("allowed_languages"):
                raise ConfigValidationError("LanguageDetectionParams: allowed_languages cannot be empty")
        elif t == "C4BadWordsFilter":  # :270-284
            kf = d.get("keep_fraction")
            if not (0.0 <= kf <= 1.0):
                raise ConfigValidationError(
                    f"C4BadWordsParams: keep_fraction must be between 0.0 and 1.0, got {rust_f64_display(kf)}")
            if not d.get("default_language"):
                raise ConfigValidationError("C4BadWordsParams: default_language cannot be empty")
        elif t == "TokenCounter":  # :357-367
            if not d.get("tokenizer_name"):
                raise ConfigValidationError("TokenCounterParams: tokenizer_name cannot be empty")


_OUT_OF_SCOPE_REQUIRED = {
    "LanguageDetectionFilter": ("min_confidence", "allowed_languages"),
    "C4BadWordsFilter": ("keep_fraction", "fail_on_missing_language", "default_language"),
    "TokenCounter": ("tokenizer_name",),
}
_IN_SCOPE = {c.TYPE: c for c in (C4QualityParams, GopherRepetitionParams, GopherQualityParams, FineWebQualityParams)}
_VARIANTS = ["C4QualityFilter", "GopherRepetitionFilter", "GopherQualityFilter", "C4BadWordsFilter",
             "LanguageDetectionFilter", "FineWebQualityFilter", "TokenCounter"]


@dataclass
class StepConfig:
    """One entry of ``pipeline:`` -- ``enum StepConfig`` (src/config/pipeline.rs:26-37)."""

    params: object

    def name(self) -> str:
        """StepConfig::name, src/config/pipeline.rs:41-51"""
        return getattr(self.params, "TYPE", None) or self.params.type_name

    def validate(self) -> None:
        self.params.validate()

    @property
    def in_scope(self) -> bool:
        return not isinstance(self.params, OutOfScopeParams)

    @classmethod
    def from_dict(cls, d) -> "StepConfig":
        if not isinstance(d, dict):
            raise ConfigError("invalid type: expected a map with a `type` tag")
        if "type" not in d:
            raise ConfigError("missing field `type`")
        t = d["type"]
        rest = {k: v for k, v in d.items() if k != "type"}
        if t in _IN_SCOPE:
            return cls(_IN_SCOPE[t].from_dict(rest))
        if t in _OUT_OF_SCOPE_REQUIRED:
            for k in _OUT_OF_SCOPE_REQUIRED[t]:
                if k not in rest:
                    raise ConfigError(f"missing field `{k}`")
            return cls(OutOfScopeParams(t, rest))
        raise ConfigError(f"unknown variant `{t}`, expected one of " + ", ".join(f"`{v}`" for v in _VARIANTS))


@dataclass
class PipelineConfig:
    """src/config/pipeline.rs:10-22"""

    pipeline: List[StepConfig]

    def validate(self) -> None:
        for s in self.pipeline:
            s.validate()

    @classmethod
    def from_dict(cls, d) -> "PipelineConfig":
        if not isinstance(d, dict) or "pipeline" not in d:
            raise ConfigError("missing field `pipeline`")
        steps = d["pipeline"]
        if steps is None:
            steps = []
        if not isinstance(steps, list):
            raise ConfigError("invalid type for `pipeline`: expected a sequence")
        return cls([StepConfig.from_dict(s) for s in steps])

    @classmethod
    def from_yaml(cls, text: str) -> "PipelineConfig":
        import yaml

        try:
            d = yaml.safe_load(text)
        except yaml.YAMLError as e:
            raise ConfigError(str(e)) from e
        return cls.from_dict(d)


def load_pipeline_config(path) -> PipelineConfig:
    """load_pipeline_config, src/config/pipeline.rs:372-393: read, parse, validate."""
    try:
        with open(path, "r", encoding="utf-8") as f:
            content = f.read()
    except OSError as e:
        raise ConfigError(f"Failed to read pipeline config file '{path}': {e}") from e
    try:
        cfg = PipelineConfig.from_yaml(content)
    except ConfigError as e:
        raise ConfigError(f"Failed to parse pipeline config YAML from '{path}': {e.message}") from e
    cfg.validate()
    return cfg
