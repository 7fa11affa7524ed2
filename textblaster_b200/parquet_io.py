"""Parquet in / out around the GPU pipeline -- SURVEY 8(f) rows 1, 2 and 4, host side.

  * ``read_batches``   mirrors ParquetReader::read_documents (src/pipeline/readers/parquet_reader.rs:67-215):
    required text / id columns, Utf8 or LargeUtf8, null rows are errors, ``source`` defaults to the path.
    The text column's Arrow buffers (values + offsets) go to ``GpuBatchPipeline.run_packed`` as they are.
  * ``run_parquet``    replaces the producer -> RabbitMQ -> worker -> aggregator round trip
    (src/producer_logic.rs:30-196, src/worker_logic.rs:140-239) for a single box: batches through the
    pipeline, outcomes into the kept / excluded files with the writer's schema
    (src/pipeline/writers/parquet_writer.rs:17-44): id, source, text, added, created{start,end}, metadata (JSON).

HTML entities: the reference decodes them in the reader with the `html-escape 0.2.13` crate
(parquet_reader.rs:178-179).  Here the Arrow buffers go to the device as they are and the pipeline decodes
them in a pre-pass (TB_OPT_DECODE_HTML_ENTITIES, csrc/tb_entities.cuh); the crate is not part of the
reference tree, so parity of this one step is unpinned.
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np

from .error import ConfigError, Unexpected

DAYS_CE_TO_EPOCH = 719163  # chrono num_days_from_ce() of 1970-01-01


@dataclass
class ParquetInputConfig:
    """src/config/parquet.rs ParquetInputConfig"""
    path: str
    text_column: str = "text"
    id_column: str = "id"
    batch_size: Optional[int] = None


def output_schema():
    """create_schema(), parquet_writer.rs:17-44"""
    import pyarrow as pa
    ts = pa.timestamp("us")
    return pa.schema([
        pa.field("id", pa.string(), nullable=False), pa.field("source", pa.string(), nullable=False),
        pa.field("text", pa.string(), nullable=False), pa.field("added", pa.date32(), nullable=True),
        pa.field("created", pa.struct([pa.field("start", ts), pa.field("end", ts)]), nullable=True),
        pa.field("metadata", pa.string(), nullable=True)])


def read_batches(cfg: ParquetInputConfig) -> Iterator["object"]:
    """Record batches of the input file after the reader's schema checks (parquet_reader.rs:67-95)."""
    import pyarrow as pa
    import pyarrow.parquet as pq

    pf = pq.ParquetFile(cfg.path)
    schema = pf.schema_arrow
    for name in (cfg.text_column, cfg.id_column):
        if schema.get_field_index(name) < 0:
            raise ConfigError(f"Required column '{name}' not found in schema.")
    t = schema.field(cfg.text_column).type
    if not (pa.types.is_string(t) or pa.types.is_large_string(t)):
        raise ConfigError(f"Column '{cfg.text_column}' must be Utf8 or LargeUtf8, found: {t}")
    kw = {"batch_size": cfg.batch_size} if cfg.batch_size else {}
    for batch in pf.iter_batches(**kw):
        yield batch


def text_buffers(arr) -> Tuple[np.ndarray, np.ndarray]:
    """(values u8, offsets i32/i64 starting anywhere) of a string array, zero-copy."""
    import pyarrow as pa

    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    large = pa.types.is_large_string(arr.type)
    bufs = arr.buffers()
    odt = np.int64 if large else np.int32
    off = np.frombuffer(bufs[1], dtype=odt, count=len(arr) + 1, offset=arr.offset * odt().itemsize)
    values = np.frombuffer(bufs[2], dtype=np.uint8) if bufs[2] is not None else np.zeros(0, np.uint8)
    return values, off


def _added_days_ce(col, n: int) -> List[Optional[int]]:
    """`added` as the writer stores it: chrono's num_days_from_ce() in a Date32 slot (parquet_writer.rs:88-93);
    the reader undoes exactly that (parquet_reader.rs:44-52), so a Date32 input passes through unchanged."""
    import pyarrow as pa

    if col is None:
        return [None] * n
    if pa.types.is_date32(col.type):
        return col.cast(pa.int32()).to_pylist()
    if pa.types.is_timestamp(col.type):
        us = col.cast(pa.timestamp("us")).cast(pa.int64()).to_pylist()
        return [None if v is None else v // 86_400_000_000 + DAYS_CE_TO_EPOCH for v in us]
    return [None] * n


class OutcomeWriter:
    """Two ParquetWriter instances (kept / excluded), src/producer_logic.rs:129-196."""

    def __init__(self, kept_path: str, excluded_path: str):
        import os
        import pyarrow.parquet as pq

        for p in (kept_path, excluded_path):
            d = os.path.dirname(p)
            if d:
                os.makedirs(d, exist_ok=True)
        self.schema = output_schema()
        self.kept = pq.ParquetWriter(kept_path, self.schema)
        self.excluded = pq.ParquetWriter(excluded_path, self.schema)
        self.n_kept = self.n_excluded = 0

    def _table(self, ids, sources, texts, added, created, metadata):
        import pyarrow as pa

        ts = pa.timestamp("us")
        if created is None:
            created = pa.array([None] * len(ids), type=pa.struct([pa.field("start", ts), pa.field("end", ts)]))
        cols = [pa.array(ids, pa.string()), pa.array(sources, pa.string()), pa.array(texts, pa.string()),
                pa.array(added, pa.int32()).cast(pa.date32()), created,
                pa.array([None if not m else json.dumps(m, ensure_ascii=False, separators=(",", ":"), sort_keys=True)
                          for m in metadata], pa.string())]
        return pa.Table.from_arrays(cols, schema=self.schema)

    def write(self, keep: Sequence[bool], ids, sources, texts, added, created, metadata):
        import pyarrow as pa

        keep = np.asarray(keep, dtype=bool)
        for sel, writer in ((keep, self.kept), (~keep, self.excluded)):
            idx = np.flatnonzero(sel)
            if not len(idx):
                continue
            cr = None if created is None else created.take(pa.array(idx))
            writer.write_table(self._table([ids[i] for i in idx], [sources[i] for i in idx], [texts[i] for i in idx],
                                           [added[i] for i in idx], cr, [metadata[i] for i in idx]))
        self.n_kept += int(keep.sum())
        self.n_excluded += int((~keep).sum())

    def close(self):
        self.kept.close()
        self.excluded.close()


def run_parquet(pipeline, cfg: ParquetInputConfig, kept_path: str, excluded_path: str, decode_entities: bool = True):
    """Read -> filter on the GPU -> write.  ``pipeline``: a GpuBatchPipeline.  Returns (received, kept, excluded),
    the triple aggregate_results_from_stream returns (producer_logic.rs:195).  ``decode_entities`` switches the
    pipeline's entity pre-pass on or off (the reference always decodes)."""
    import pyarrow as pa

    from . import _abi

    pipeline.set_option(_abi.TB_OPT_DECODE_HTML_ENTITIES, 1 if decode_entities else 0)
    out = OutcomeWriter(kept_path, excluded_path)
    received = 0
    for batch in read_batches(cfg):
        n = batch.num_rows
        names = batch.schema.names
        text_arr, id_arr = batch.column(cfg.text_column), batch.column(cfg.id_column)
        for arr, nm in ((text_arr, cfg.text_column), (id_arr, cfg.id_column)):
            if arr.null_count:
                i = int(np.flatnonzero(~np.asarray(arr.is_valid()))[0])
                kind = "text" if nm == cfg.text_column else "id"
                raise Unexpected(f"Row {i} has null {kind} column '{nm}'")
        values, off = text_buffers(text_arr)
        res = pipeline.run_packed(values, off)
        ids = id_arr.to_pylist()
        sources = batch.column("source").to_pylist() if "source" in names else [None] * n
        sources = [cfg.path if s is None else s for s in sources]
        added = _added_days_ce(batch.column("added") if "added" in names else None, n)
        created = batch.column("created") if "created" in names and pa.types.is_struct(batch.column("created").type) else None
        if created is not None:
            ts = pa.timestamp("us")
            created = created.cast(pa.struct([pa.field("start", ts), pa.field("end", ts)]))
        in_md = batch.column("metadata").to_pylist() if "metadata" in names and (
            pa.types.is_string(batch.column("metadata").type) or pa.types.is_large_string(batch.column("metadata").type)) else [None] * n
        texts, metadata = [], []
        for i in range(n):
            md = {}
            if in_md[i]:
                # serde_json::from_str::<HashMap<String, String>>: anything but an object whose values are all
                # strings fails to parse, and the reader then falls back to an empty map (parquet_reader.rs:224-229)
                try:
                    obj = json.loads(in_md[i])
                    md = dict(obj) if isinstance(obj, dict) and all(isinstance(v, str) for v in obj.values()) else {}
                except ValueError:
                    md = {}
            step_md, _reason = res.metadata_and_reason(i)
            md.update(step_md)
            metadata.append(md)
            texts.append(res.content_of(i) if res.content is not None else
                         bytes(values[int(off[i]):int(off[i + 1])]).decode("utf-8"))
        out.write(res.keep, ids, sources, texts, added, created, metadata)
        received += n
    out.close()
    return received, out.n_kept, out.n_excluded
