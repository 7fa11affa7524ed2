"""Host-side mirror of the reference's step / executor interface for the accelerated path.

  * ``ProcessingStep``          -- trait ProcessingStep, src/executor.rs:8-15 (``name``, ``process``)
  * ``C4QualityFilter`` / ``GopherQualityFilter`` / ``GopherRepetitionFilter``
                                -- same constructor arguments as the Rust ``new`` functions
                                   (c4_filters.rs:112-138, gopher_quality.rs:34-60, gopher_rep.rs:25-43)
  * ``PipelineExecutor``        -- src/executor.rs:17-70 (``run_single``, ``run_batch``)
  * ``build_pipeline_from_config`` -- src/worker_logic.rs:39-134
  * ``GpuBatchPipeline``        -- the batch executor: a whole Arrow string column per call through the
                                   C ABI (``tb_pipeline_run_batch``); the only place the GPU is driven.

``process`` returns the document or raises ``DocumentFiltered`` exactly like the Rust ``Result``;
``PipelineExecutor`` wraps a step's error in ``StepError`` (src/executor.rs:37-46).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Union

import numpy as np

from . import _abi, _native
from .config import (C4QualityParams, FineWebQualityParams, GopherQualityParams, GopherRepetitionParams, PipelineConfig,
                     StepConfig)
from .data_model import ProcessingOutcome, TextDocument
from .error import ConfigError, DocumentFiltered, PipelineError, StepError, Unexpected
from .reasons import step_metadata

_ctx_cache = {}


def _ctx(device: int):
    """One tb_ctx per device per process."""
    h = _ctx_cache.get(device)
    if h is None:
        p = C.c_void_p()
        _native.check(_native.lib().tb_ctx_create(device, C.byref(p)))
        h = p.value
        _ctx_cache[device] = h
    return h


class BatchResult:
    """Results of one batch, copied out of the library's pinned buffers."""

    def __init__(self, params, out: _abi.tb_batch_out, copy_content: bool = True, source=None):
        """source: (utf8, offsets) of the input column; with TB_OPT_RETURN_CONTENT = 2 the content is gathered from it"""
        n = out.n_docs
        self.n_docs = n
        self.params = list(params)
        self.keep = np.ctypeslib.as_array(out.keep, (n,)).copy() if n else np.zeros(0, np.uint8)
        self.failed_step = np.ctypeslib.as_array(out.failed_step, (n,)).copy() if n else np.zeros(0, np.int8)
        self.stats = []
        for s, p in enumerate(self.params):
            dt = _abi.stats_dtype(p.CTYPE)
            if n:
                raw = C.string_at(out.step_stats[s], n * dt.itemsize)
                self.stats.append(np.frombuffer(raw, dtype=dt).copy())
            else:
                self.stats.append(np.zeros(0, dt))
        self.content = None
        self.content_offsets = None
        if out.content_rewritten and copy_content:
            self.content_offsets = np.ctypeslib.as_array(out.content_offsets, (n + 1,)).copy()
            nb = int(out.content_bytes)
            self.content = np.ctypeslib.as_array(out.content, (nb,)).copy() if nb else np.zeros(0, np.uint8)
        self.runs = None
        self.run_offsets = None
        if out.content_kind == 2:
            # runs of the input text: kept as they came, and (unless copy_content is off) gathered into the same
            # content / content_offsets a TB_OPT_RETURN_CONTENT = 1 call returns
            self.run_offsets = np.ctypeslib.as_array(out.run_offsets, (n + 1,)).copy()
            nr = int(out.n_runs)
            self.runs = (np.frombuffer(C.string_at(out.runs, nr * 8), dtype=np.uint32).reshape(nr, 2).copy() if nr
                         else np.zeros((0, 2), np.uint32))
            if copy_content and source is not None:
                self.content, self.content_offsets = gather_content(out, source[0], source[1])
        self.gpu_ms = out.gpu_ms
        self.gpu_launches = out.gpu_launches

    def content_of(self, i: int) -> Optional[str]:
        if self.content is None:
            return None
        a, b = int(self.content_offsets[i]), int(self.content_offsets[i + 1])
        return bytes(self.content[a:b]).decode("utf-8")

    def metadata_and_reason(self, i: int):
        """Metadata every executed step wrote for document i, and the failing step's reason string."""
        md = {}
        reason = ""
        fs = int(self.failed_step[i])
        last = fs if fs >= 0 else len(self.params) - 1
        for s in range(last + 1):
            m, r = step_metadata(self.params[s], self.stats[s][i])
            md.update(m)
            if s == fs:
                reason = r
        return md, reason


def gather_content(out: _abi.tb_batch_out, utf8: np.ndarray, offsets: np.ndarray, select: Optional[np.ndarray] = None,
                   nthreads: int = 0):
    """tb_gather_content: the contents of the selected documents (all if ``select`` is None) of a TB_OPT_RETURN_CONTENT = 2
    result, copied out of the input column.  Returns (content u8, offsets i64)."""
    lib = _native.lib()
    n = int(out.n_docs)
    sel = None if select is None else np.ascontiguousarray(select, dtype=np.uint8)
    nsel = n if sel is None else int(np.count_nonzero(sel))
    doff = np.zeros(nsel + 1, np.int64)
    total = C.c_int64()
    args = (C.byref(out), utf8.ctypes.data if utf8.size else None, offsets.ctypes.data, 1 if offsets.dtype == np.int64 else 0,
            None if sel is None else sel.ctypes.data)
    _native.check(lib.tb_gather_content(*args, None, 0, doff.ctypes.data, C.byref(total), nthreads))
    dst = np.empty(max(int(total.value), 1), np.uint8)
    _native.check(lib.tb_gather_content(*args, dst.ctypes.data, int(total.value), doff.ctypes.data, C.byref(total), nthreads))
    return dst[: int(total.value)], doff


class GpuBatchPipeline:
    """An immutable chain of C4 / Gopher steps on one GPU (tb_pipeline).  Thread-safe."""

    def __init__(self, params: Sequence[object], device: int = 0, validate: bool = True, decode_html_entities: bool = False):
        """decode_html_entities: decode entities on the device before the first step, as the reference's reader does
        (parquet_reader.rs:178-179); the decoded text is then the content of every document."""
        self.params = list(params)
        for p in self.params:
            if not isinstance(p, (C4QualityParams, GopherQualityParams, GopherRepetitionParams, FineWebQualityParams)):
                raise ConfigError(f"step type {getattr(p, 'type_name', type(p).__name__)} has no GPU implementation")
            if validate:
                p.validate()
        self._keep = []
        arr = (_abi.tb_step_cfg * max(len(self.params), 1))()
        for i, p in enumerate(self.params):
            arr[i] = p.to_c(self._keep)
        self._arr = arr
        h = C.c_void_p()
        self._lib = _native.lib()
        _native.check(self._lib.tb_pipeline_create(_ctx(device), C.addressof(arr), len(self.params), C.byref(h)))
        self._h = h.value
        self.device = device
        if decode_html_entities:
            self.set_option(_abi.TB_OPT_DECODE_HTML_ENTITIES, 1)

    def set_option(self, option: int, value: int):
        _native.check(self._lib.tb_pipeline_set_option(self._h, option, value))

    def run_packed(self, utf8: np.ndarray, offsets: np.ndarray, copy_content: bool = True) -> BatchResult:
        """utf8: uint8 values buffer; offsets: int32 or int64, n+1 entries (an Arrow string column)."""
        if offsets.dtype not in (np.int32, np.int64):
            raise TypeError("offsets must be int32 or int64")
        utf8 = np.ascontiguousarray(utf8, dtype=np.uint8)
        offsets = np.ascontiguousarray(offsets)
        n = len(offsets) - 1
        out = _abi.tb_batch_out()
        _native.check(self._lib.tb_pipeline_run_batch(self._h, utf8.ctypes.data if utf8.size else None, offsets.ctypes.data,
                                                      1 if offsets.dtype == np.int64 else 0, n, C.byref(out)))
        try:
            return BatchResult(self.params, out, copy_content, source=(utf8, offsets))
        finally:
            self._lib.tb_batch_out_release(self._h, C.byref(out))

    def run_raw(self, utf8_ptr: int, offsets_ptr: int, offsets_are_i64: bool, n: int, out: _abi.tb_batch_out):
        """No copies, no numpy: the caller owns ``out`` and must call ``release``. Used by bench.py."""
        _native.check(self._lib.tb_pipeline_run_batch(self._h, utf8_ptr, offsets_ptr, 1 if offsets_are_i64 else 0, n, C.byref(out)))

    def run_device_raw(self, utf8_dev: int, offsets_dev: int, n: int, n_bytes: int, out: _abi.tb_batch_out):
        _native.check(self._lib.tb_pipeline_run_batch_device(self._h, utf8_dev, offsets_dev, n, n_bytes, C.byref(out)))

    def release(self, out: _abi.tb_batch_out):
        self._lib.tb_batch_out_release(self._h, C.byref(out))

    def run_texts(self, texts: Sequence[str]) -> BatchResult:
        enc = [t.encode("utf-8") for t in texts]
        off = np.zeros(len(enc) + 1, np.int64)
        if enc:
            off[1:] = np.cumsum([len(e) for e in enc])
        return self.run_packed(np.frombuffer(b"".join(enc), dtype=np.uint8), off)

    def kernel_times(self):
        names = (C.c_char_p * 32)()
        ms = (C.c_double * 32)()
        launches = (C.c_int64 * 32)()
        k = self._lib.tb_pipeline_kernel_times(self._h, 32, names, ms, launches)
        return {names[i].decode(): (ms[i], launches[i]) for i in range(min(k, 32))}

    def reset_kernel_times(self):
        self._lib.tb_pipeline_reset_kernel_times(self._h)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.tb_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------
# ProcessingStep mirror
# ---------------------------------------------------------------------------------------------
class ProcessingStep:
    """trait ProcessingStep (src/executor.rs:8-15)."""

    def name(self) -> str:
        raise NotImplementedError

    def process(self, document: TextDocument) -> TextDocument:
        raise NotImplementedError


class _GpuStep(ProcessingStep):
    NAME = ""

    def __init__(self, params, device: int = 0):
        self.params = params
        self.device = device
        self._pipe = None

    def name(self) -> str:
        return self.NAME

    def _pipeline(self) -> GpuBatchPipeline:
        if self._pipe is None:
            self._pipe = GpuBatchPipeline([self.params], self.device, validate=False)
        return self._pipe

    def process(self, document: TextDocument) -> TextDocument:
        res = self._pipeline().run_texts([document.content])
        return _apply_result(document, res, 0, raise_step_error=False, names=[self.NAME])


class C4QualityFilter(_GpuStep):
    """C4QualityFilter::new, c4_filters.rs:112-138 (same argument order)."""
    NAME = "C4QualityFilter"

    def __init__(self, split_paragraph, remove_citations, filter_no_terminal_punct, min_num_sentences, min_words_per_line,
                 max_word_length, filter_lorem_ipsum, filter_javascript, filter_curly_bracket, filter_policy, device: int = 0):
        super().__init__(C4QualityParams(split_paragraph, remove_citations, filter_no_terminal_punct, min_num_sentences,
                                         min_words_per_line, max_word_length, filter_lorem_ipsum, filter_javascript,
                                         filter_curly_bracket, filter_policy), device)


class GopherQualityFilter(_GpuStep):
    """GopherQualityFilter::new, gopher_quality.rs:34-60."""
    NAME = "GopherQualityFilter"

    def __init__(self, min_doc_words=None, max_doc_words=None, min_avg_word_length=None, max_avg_word_length=None,
                 max_symbol_word_ratio=None, max_bullet_lines_ratio=None, max_ellipsis_lines_ratio=None,
                 max_non_alpha_words_ratio=None, min_stop_words=None, stop_words=None, device: int = 0):
        super().__init__(GopherQualityParams(min_doc_words, max_doc_words, min_avg_word_length, max_avg_word_length,
                                             max_symbol_word_ratio, max_bullet_lines_ratio, max_ellipsis_lines_ratio,
                                             max_non_alpha_words_ratio, min_stop_words, stop_words), device)


class GopherRepetitionFilter(_GpuStep):
    """GopherRepetitionFilter::new, gopher_rep.rs:25-43."""
    NAME = "GopherRepetitionFilter"

    def __init__(self, dup_line_frac=None, dup_para_frac=None, dup_line_char_frac=None, dup_para_char_frac=None,
                 top_n_grams=(), dup_n_grams=(), device: int = 0):
        super().__init__(GopherRepetitionParams(dup_line_frac, dup_para_frac, dup_line_char_frac, dup_para_char_frac,
                                                list(top_n_grams), list(dup_n_grams)), device)


class FineWebQualityFilter(_GpuStep):
    """FineWebQualityFilter::new, fineweb_quality.rs:41-62 (same argument order)."""
    NAME = "FineWebQualityFilter"

    def __init__(self, line_punct_thr, line_punct_exclude_zero, short_line_thr, short_line_length, char_duplicates_ratio,
                 new_line_ratio, stop_chars=None, device: int = 0):
        super().__init__(FineWebQualityParams(line_punct_thr, line_punct_exclude_zero, short_line_thr, short_line_length,
                                              char_duplicates_ratio, new_line_ratio,
                                              None if stop_chars is None else list(stop_chars)), device)


def _apply_result(doc: TextDocument, res: BatchResult, i: int, raise_step_error: bool, names: List[str]) -> TextDocument:
    """Turn row i of a batch result into the Rust ``Result<TextDocument>``: returns the document
    (content + metadata updated) or raises DocumentFiltered / StepError{DocumentFiltered}."""
    md, reason = res.metadata_and_reason(i)
    content = res.content_of(i)
    out = TextDocument(id=doc.id, content=doc.content if content is None else content, source=doc.source, added=doc.added,
                       created=doc.created, metadata=dict(doc.metadata))
    out.metadata.update(md)
    fs = int(res.failed_step[i])
    if fs < 0:
        return out
    err = DocumentFiltered(out, reason)
    if raise_step_error:
        raise StepError(names[fs], err)
    raise err


class PipelineExecutor:
    """PipelineExecutor, src/executor.rs:17-70.  When every step is one of the three GPU steps the
    whole chain is fused into one ``GpuBatchPipeline`` (one device round trip per batch); otherwise
    steps run one after the other like the reference."""

    def __init__(self, steps: Sequence[ProcessingStep], device: int = 0):
        self.steps = list(steps)
        self.device = device
        self._fused = None
        if self.steps and all(isinstance(s, _GpuStep) for s in self.steps):
            self._fused = GpuBatchPipeline([s.params for s in self.steps], device, validate=False)

    def run_single(self, document: TextDocument) -> TextDocument:
        """run_single_async (src/executor.rs:30-57): raises StepError on the first failing step."""
        r = self.run_batch([document])[0]
        if isinstance(r, Exception):
            raise r
        return r

    def run_batch(self, documents: Sequence[TextDocument]) -> List[Union[TextDocument, PipelineError]]:
        """run_batch_parallel_async (src/executor.rs:60-70): one entry per input, input order."""
        docs = list(documents)
        if not docs:
            return []
        if self._fused is not None:
            res = self._fused.run_texts([d.content for d in docs])
            names = [s.name() for s in self.steps]
            out = []
            for i, d in enumerate(docs):
                try:
                    out.append(_apply_result(d, res, i, True, names))
                except PipelineError as e:
                    out.append(e)
            return out
        results: List[Union[TextDocument, PipelineError]] = []
        for d in docs:
            cur = d
            try:
                for s in self.steps:
                    try:
                        cur = s.process(cur)
                    except PipelineError as e:
                        raise StepError(s.name(), e)
                results.append(cur)
            except PipelineError as e:
                results.append(e)
        return results

    def run_outcomes(self, documents: Sequence[TextDocument]) -> List[Optional[ProcessingOutcome]]:
        """execute_processing_pipeline's mapping (src/worker_logic.rs:155-180): Ok -> Success,
        StepError{DocumentFiltered} -> Filtered, anything else -> None."""
        out = []
        for r in self.run_batch(documents):
            if isinstance(r, TextDocument):
                out.append(ProcessingOutcome.success(r))
            elif isinstance(r, StepError) and isinstance(r.source, DocumentFiltered):
                out.append(ProcessingOutcome.filtered(r.source.document, r.source.reason))
            else:
                out.append(None)
        return out


def build_pipeline_from_config(config: PipelineConfig, device: int = 0, skip_out_of_scope: bool = False) -> List[ProcessingStep]:
    """build_pipeline_from_config (src/worker_logic.rs:39-134) for the three accelerated step types.
    The reference's other step types raise ConfigError unless ``skip_out_of_scope``."""
    steps: List[ProcessingStep] = []
    for sc in config.pipeline:
        p = sc.params
        if isinstance(p, C4QualityParams):
            st = C4QualityFilter.__new__(C4QualityFilter)
        elif isinstance(p, GopherQualityParams):
            st = GopherQualityFilter.__new__(GopherQualityFilter)
        elif isinstance(p, GopherRepetitionParams):
            st = GopherRepetitionFilter.__new__(GopherRepetitionFilter)
        elif isinstance(p, FineWebQualityParams):
            st = FineWebQualityFilter.__new__(FineWebQualityFilter)
        else:
            if skip_out_of_scope:
                continue
            raise ConfigError(f"step type {sc.name()} is outside the GPU hot path (C4QualityFilter, GopherQualityFilter, "
                              "GopherRepetitionFilter, FineWebQualityFilter)")
        _GpuStep.__init__(st, p, device)
        steps.append(st)
    return steps
