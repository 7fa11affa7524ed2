"""Parquet reader / outcome writer around the pipeline: schema and routing of
producer_logic.rs:129-196 + parquet_writer.rs:17-157, reader checks of parquet_reader.rs:67-166."""
import json

import numpy as np
import pyarrow as pa
import pyarrow.parquet as pq
import pytest

import oracle
from parity_cases import CHAINS
from textblaster_b200 import parquet_io
from textblaster_b200.error import ConfigError, Unexpected

import random

import fuzzgen

_rng = random.Random(77)
DOCS = [fuzzgen.english_doc(_rng, 1800), "Some code block {}.",
        "Tom &amp; Jerry said &#72;i&#x21; and that was it.\n" + fuzzgen.english_doc(_rng, 2500), "short"]


class OracleAsPipeline:
    """Stands in for GpuBatchPipeline on a machine without a GPU (tests only)."""

    def __init__(self, steps):
        self.pl = oracle.OraclePipeline(steps)
        self.decode = False

    def set_option(self, option, value):
        from textblaster_b200 import _abi
        if option == _abi.TB_OPT_DECODE_HTML_ENTITIES:
            self.decode = bool(value)

    def run_packed(self, values, off):
        values, off = np.asarray(values), np.asarray(off, dtype=np.int64)
        if self.decode:  # what the device pre-pass does, by the oracle's restatement of the crate
            texts = [oracle.decode_html_entities(bytes(values[off[i]:off[i + 1]]).decode("utf-8")) for i in range(len(off) - 1)]
            r = self.pl.run_texts(texts)
        else:
            r = self.pl.run_packed(values, off)

        class R:
            keep = r.keep
            content = True
            def metadata_and_reason(self, i): return r.metadata(i), r.reason(i)
            def content_of(self, i): return r.content(i)
        return R()


def _write_input(path, large=False, with_null=False):
    t = pa.large_string() if large else pa.string()
    texts = list(DOCS)
    if with_null:
        texts[1] = None
    tab = pa.table({"id": pa.array([f"d{i}" for i in range(len(DOCS))]), "text": pa.array(texts, t),
                    "source": pa.array(["s0", None, "s2", "s3"]), "added": pa.array([738000, None, 738002, 738003], pa.int32()).cast(pa.date32()),
                    "metadata": pa.array(['{"lang":"en"}', None, "not json", '{"a": 1, "b": "x"}'])})
    pq.write_table(tab, path)


def _run(tmp_path, pipeline, **kw):
    tmp_path.mkdir(parents=True, exist_ok=True)
    src = str(tmp_path / "in.parquet")
    _write_input(src, **kw)
    kept, exc = str(tmp_path / "out" / "kept.parquet"), str(tmp_path / "out" / "excluded.parquet")
    counts = parquet_io.run_parquet(pipeline, parquet_io.ParquetInputConfig(src), kept, exc)
    return counts, pq.read_table(kept), pq.read_table(exc), src


def test_outcome_files_schema_and_routing(tmp_path):
    counts, kept, exc, src = _run(tmp_path, OracleAsPipeline(CHAINS["baseline_order"]))
    assert counts == (4, kept.num_rows, exc.num_rows) and kept.num_rows + exc.num_rows == 4
    assert kept.schema.equals(parquet_io.output_schema()) and exc.schema.equals(parquet_io.output_schema())
    k, e = kept.to_pydict(), exc.to_pydict()
    ref = oracle.OraclePipeline(CHAINS["baseline_order"]).run_texts([oracle.decode_html_entities(d) for d in DOCS])
    assert list(ref.keep) == [1, 0, 1, 0], [ref.reason(i) for i in range(4)]
    assert k["id"] == ["d0", "d2"] and e["id"] == ["d1", "d3"]
    assert "&amp;" not in k["text"][1] and "Tom & Jerry said Hi! and that was it." in k["text"][1]   # entities decoded before the filters
    assert e["source"][0] == src and k["source"] == ["s0", "s2"]                    # missing source -> the input path
    md0 = json.loads(k["metadata"][0])
    assert md0["lang"] == "en" and md0["c4_filter_status"] == "passed" and md0["gopher_repetition_filter_status"] == "passed"
    md1 = json.loads(e["metadata"][0])
    assert md1 == {"c4_filter_status": "filtered", "c4_filter_reasons": "curly_bracket"}   # later steps never ran
    # a metadata object with a non-string value does not deserialise into HashMap<String, String>: the reader
    # falls back to an empty map (parquet_reader.rs:224-229), so only the step's keys are written
    md3 = json.loads(e["metadata"][1])
    assert "a" not in md3 and "b" not in md3 and md3["c4_filter_status"] == "filtered"
    assert kept.column("added").cast(pa.int32()).to_pylist() == [738000, 738002]          # days pass through unchanged


def test_large_utf8_and_buffers_zero_copy(tmp_path):
    src = str(tmp_path / "l.parquet")
    _write_input(src, large=True)
    batch = next(parquet_io.read_batches(parquet_io.ParquetInputConfig(src)))
    values, off = parquet_io.text_buffers(batch.column("text"))
    assert off.dtype == np.int64 and bytes(values[off[3]:off[4]]) == b"short"
    sliced = batch.column("text").slice(2, 2)
    v2, o2 = parquet_io.text_buffers(sliced)
    assert len(o2) == 3 and bytes(v2[o2[1]:o2[2]]) == b"short"


def test_reader_errors(tmp_path):
    src = str(tmp_path / "in.parquet")
    _write_input(src)
    with pytest.raises(ConfigError, match="Required column 'body' not found in schema."):
        next(parquet_io.read_batches(parquet_io.ParquetInputConfig(src, text_column="body")))
    with pytest.raises(ConfigError, match="must be Utf8 or LargeUtf8"):
        next(parquet_io.read_batches(parquet_io.ParquetInputConfig(src, text_column="added")))
    with pytest.raises(Unexpected, match="Row 1 has null text column 'text'"):
        _run(tmp_path, OracleAsPipeline(CHAINS["gq_open"]), with_null=True)


@pytest.mark.gpu
def test_gpu_pipeline_writes_the_same_files(tmp_path):
    from textblaster_b200.executor import GpuBatchPipeline
    a = _run(tmp_path / "gpu", GpuBatchPipeline(CHAINS["shipped_with_fineweb"]))
    b = _run(tmp_path / "cpu", OracleAsPipeline(CHAINS["shipped_with_fineweb"]))
    assert a[0] == b[0]
    for x, y in ((a[1], b[1]), (a[2], b[2])):
        dx, dy = x.to_pydict(), y.to_pydict()
        dx.pop("source"); dy.pop("source")   # the default source is the (different) input path
        assert dx == dy
