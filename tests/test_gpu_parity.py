"""GPU parity tests: the CUDA kernels, called through the C ABI (tb_pipeline_run_batch), against the
CPU oracle on the same inputs.  Bar: every integer counter, keep / failed_step decision, reason bit
and the rewritten content bit-exact; the f64 ratios are single divisions of those integers, so the
reason strings are compared verbatim (tolerance: 0).  Nothing here reads /root/reference."""
import ctypes as C
import os
import random

import numpy as np
import pytest

import fuzzgen
import multilingual
import oracle
from parity_cases import CHAINS, GQ_OPEN, GR_OPEN, step_types
from parity_util import compare, executed
from test_emul_parity import (level_task_docs, GOLDEN_DOCS, SENTENCE_LINE_DOCS, boundary_sentence_docs, entity_docs, fineweb_docs,
                              signature_pruning_docs, split_ambiguity_docs)
from test_oracle_golden import C4_LINE_CASES, c4_default
from textblaster_b200 import _abi, synth
from textblaster_b200.config import GopherRepetitionParams
from textblaster_b200.data_model import TextDocument
from textblaster_b200.error import DocumentFiltered, StepError
from textblaster_b200.executor import (C4QualityFilter, GopherQualityFilter, GopherRepetitionFilter, GpuBatchPipeline,
                                       PipelineExecutor)

pytestmark = pytest.mark.gpu
NCPU = os.cpu_count() or 1
_pipes = {}


def pipe(name):
    if name not in _pipes:
        _pipes[name] = GpuBatchPipeline(CHAINS[name])
    return _pipes[name]


def check_texts(name, texts, **kw):
    steps = CHAINS[name]
    res = pipe(name).run_texts(texts)
    ores = oracle.OraclePipeline(steps).run_texts(texts, nthreads=NCPU)
    contents = [res.content_of(i) for i in range(res.n_docs)] if res.content is not None else None
    compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name)
    for i in range(0, len(texts), max(1, len(texts) // 50)):  # reason strings + metadata, sampled
        md, reason = res.metadata_and_reason(i)
        assert reason == ores.reason(i) and md == ores.metadata(i)
    return res


def check_packed(name, utf8, off, p=None):
    steps = CHAINS[name]
    res = (p or pipe(name)).run_packed(utf8, off)
    ores = oracle.OraclePipeline(steps).run_packed(utf8, off.astype(np.int64), nthreads=NCPU)
    compare(ores, res.keep, res.failed_step, res.stats, None, step_types(steps), None, label=name)
    if res.content is not None:
        ob, oo = ores.content_packed()
        assert np.array_equal(res.content_offsets, oo) and np.array_equal(res.content, ob)
    return res, ores


def _same_stats(r, ref):
    """statistics agree on every row whose step executed (later steps are undefined after a failure)"""
    ran = ref.failed_step.astype(np.int64)
    for s_i, (x, y) in enumerate(zip(r.stats, ref.stats)):
        m = (ran < 0) | (ran >= s_i)
        assert np.array_equal(x[m], y[m])


@pytest.mark.parametrize("name", sorted(CHAINS))
def test_golden_documents(name):
    check_texts(name, GOLDEN_DOCS)


@pytest.mark.parametrize("name", sorted(CHAINS))
def test_fuzzed_documents(name):
    check_texts(name, fuzzgen.FuzzGen(101).docs(1500))


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "c4_then_open", "gr_open", "gq_open"])
def test_multi_tile_documents(name):
    rng = random.Random(5)
    texts = [fuzzgen.english_doc(rng, sz) for sz in (200, 900, 1023, 1024, 1025, 1100, 3000, 8000, 20000, 70000)]
    texts += fuzzgen.FuzzGen(6).docs(60, max_lines=120)
    texts += ["start " + "a" * 5000 + " end. Another sentence here. And a third one.", "a" * 3000, "é" * 1500 + " x"]
    check_texts(name, texts)


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "c4_then_open"])
def test_synthetic_c4_corpus(name):
    """BASELINE.json config 2 shape (2 KB median, log-normal), 20k documents."""
    utf8, off = synth.generate(20000, seed=synth.BASE_SEED + 2)
    res, ores = check_packed(name, utf8, off)
    assert 0 < int(res.keep.sum()) < res.n_docs or name == "shipped_order"


@pytest.mark.parametrize("name", ["gq_open", "gq_shipped", "gr_open", "c4_lax", "c4_then_open", "baseline_order"])
def test_ascii_fast_path(name):
    """pure-ASCII text dense in Mid*/quote/underscore/digit patterns: the bit-parallel word pass"""
    rng = random.Random(19)
    texts = [fuzzgen.ascii_doc(rng, rng.choice([5, 40, 400, 1500, 6000])) for _ in range(600)]
    texts += ["a" * 31 + " " + "b" * 40, "x" * 100, "ab " * 200, "a.b" * 30, "1,2" * 40, "it's " * 50, "a\r\nb c\r\n" * 20]
    check_texts(name, texts)


@pytest.mark.parametrize("name", ["c4_lax", "c4_shipped", "gq_open", "gr_open", "baseline_order", "shipped_order", "fw_open", "fw_custom"])
def test_boundary_documents(name):
    """White space, citations, multi-byte characters and MidLetter joins sliding across the 32-byte stretch and
    1 KiB tile boundaries of the mask-based passes."""
    check_texts(name, fuzzgen.boundary_docs())


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "gq_open", "gr_open", "c4_lax"])
def test_general_word_path_matches(name):
    """TB_OPT_GENERAL_WORD_PATH: the segment-walking kernels bodies, which take over for documents the mask form
    cannot express, give the same results on every document."""
    rng = random.Random(5)
    texts = fuzzgen.FuzzGen(21).docs(150) + [fuzzgen.english_doc(rng, sz) for sz in (900, 2500, 9000)] + fuzzgen.boundary_docs()[:60]
    p = GpuBatchPipeline(CHAINS[name])
    p.set_option(_abi.TB_OPT_GENERAL_WORD_PATH, 1)
    steps = CHAINS[name]
    res = p.run_texts(texts)
    ores = oracle.OraclePipeline(steps).run_texts(texts, nthreads=NCPU)
    contents = [res.content_of(i) for i in range(res.n_docs)] if res.content is not None else None
    compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name + "/general")
    p.close()


@pytest.mark.parametrize("name", ["fw_open", "fw_shipped", "fw_custom", "shipped_with_fineweb"])
def test_fineweb_documents(name):
    rng = random.Random(9)
    check_texts(name, fineweb_docs() + [fuzzgen.english_doc(rng, sz) for sz in (300, 2000, 9000, 40000)] + multilingual.corpus(5, 200))


@pytest.mark.parametrize("name", ["c4_sentences", "c4_sentences_lax", "sentences_then_open"])
def test_sentence_wise_lines(name):  # split_paragraph = false, c4_filters.rs:152-155
    rng = random.Random(23)
    texts = SENTENCE_LINE_DOCS + boundary_sentence_docs() + fuzzgen.boundary_docs()[:80]
    texts += [fuzzgen.english_doc(rng, sz) for sz in (300, 1024, 2500, 9000, 40000, 150000)] + multilingual.corpus(7, 150)
    check_texts(name, texts)


def test_sentence_wise_lines_corpus_and_chunks():
    """BASELINE config 2 shape with sentences as lines; the rewritten text can outgrow a chunk's input."""
    utf8, off = synth.generate(8000, seed=synth.BASE_SEED + 6)
    p = GpuBatchPipeline(CHAINS["sentences_then_open"])
    ref, _ = check_packed("sentences_then_open", utf8, off, p)
    p.set_option(_abi.TB_OPT_CHUNK_BYTES, 1 << 18)
    r = p.run_packed(utf8, off)
    assert np.array_equal(r.keep, ref.keep) and np.array_equal(r.failed_step, ref.failed_step)
    _same_stats(r, ref)
    assert np.array_equal(r.content, ref.content) and np.array_equal(r.content_offsets, ref.content_offsets)
    grow = ["?".join("ab" for _ in range(4000))] * 50  # every break adds a byte
    rg = p.run_texts(grow)
    assert rg.content_of(0) == grow[0].replace("?", "?\n") and int(rg.content_offsets[-1]) > sum(len(t) for t in grow)
    p.close()


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "gq_open", "gr_open", "c4_lax", "sentences_then_open", "fw_open"])
def test_html_entities_decoded_first(name):  # parquet_reader.rs:178-179; parity with html-escape unpinned
    rng = random.Random(29)
    texts = entity_docs() + [fuzzgen.english_doc(rng, sz).replace(" and ", " &amp; ").replace("the ", "the&nbsp;") for sz in (500, 3000, 20000, 90000)]
    steps = CHAINS[name]
    decoded = [oracle.decode_html_entities(t) for t in texts]
    p = GpuBatchPipeline(steps, decode_html_entities=True)
    for chunk in (None, 1 << 16):  # one chunk, then many
        if chunk:
            p.set_option(_abi.TB_OPT_CHUNK_BYTES, chunk)
        res = p.run_texts(texts)
        ores = oracle.OraclePipeline(steps).run_texts(decoded, nthreads=NCPU)
        assert res.content is not None  # the decoded text is the content, with or without a C4 step
        contents = [res.content_of(i) for i in range(res.n_docs)]
        compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name + "/entities")
    # without entities the pre-pass is the identity
    plain = [t for t in GOLDEN_DOCS if "&" not in t]
    r1, r0 = p.run_texts(plain), pipe(name).run_texts(plain)
    assert np.array_equal(r1.keep, r0.keep) and np.array_equal(r1.failed_step, r0.failed_step)
    _same_stats(r1, r0)
    p.close()


def test_ngram_split_ambiguity():
    """Concatenated n-gram keys that agree although the word sequences differ ("no where" / "now here")."""
    check_texts("gr_open", split_ambiguity_docs())
    check_texts("gr_shipped", split_ambiguity_docs())


def test_int32_offsets_and_chunking():
    """Arrow StringArray offsets are i32 (parquet_reader.rs:108-112); tiny chunks force the
    double-buffered multi-chunk path and must not change any result."""
    utf8, off = synth.generate(3000, seed=11, median_bytes=1200)
    p = GpuBatchPipeline(CHAINS["baseline_order"])
    ref, _ = check_packed("baseline_order", utf8, off, p)
    r32 = p.run_packed(utf8, off.astype(np.int32))
    p.set_option(_abi.TB_OPT_CHUNK_BYTES, 1 << 16)
    rch = p.run_packed(utf8, off)
    for r in (r32, rch):
        assert np.array_equal(r.keep, ref.keep) and np.array_equal(r.failed_step, ref.failed_step)
        _same_stats(r, ref)
        assert np.array_equal(r.content, ref.content) and np.array_equal(r.content_offsets, ref.content_offsets)
    # a sliced column: offsets[0] != 0
    k = 500
    rs = p.run_packed(utf8, off[k:])
    assert np.array_equal(rs.keep, ref.keep[k:]) and np.array_equal(rs.failed_step, ref.failed_step[k:])
    p.close()


def test_device_resident_entry_point():
    import torch

    utf8, off = synth.generate(4000, seed=12)
    p = GpuBatchPipeline(CHAINS["baseline_order"])
    host = p.run_packed(utf8, off)
    tu = torch.from_numpy(utf8.copy()).cuda()
    to = torch.from_numpy(off.copy()).cuda()
    out = _abi.tb_batch_out()
    p.run_device_raw(tu.data_ptr(), to.data_ptr(), len(off) - 1, int(off[-1]), out)
    keep = np.ctypeslib.as_array(out.keep, (out.n_docs,)).copy()
    failed = np.ctypeslib.as_array(out.failed_step, (out.n_docs,)).copy()
    nb = int(out.content_bytes)
    content = np.ctypeslib.as_array(out.content, (nb,)).copy()
    assert out.gpu_launches > 0
    p.release(out)
    assert np.array_equal(keep, host.keep) and np.array_equal(failed, host.failed_step) and np.array_equal(content, host.content)
    p.close()


def test_concurrent_callers():
    """ProcessingStep is Send + Sync and one executor is shared by all worker tasks (bin/worker.rs:149,
    worker_logic.rs:260-264): calls on one pipeline from several host threads, and on several pipelines of one
    context at once, give the results of the same calls made one after the other."""
    import threading
    utf8, off = synth.generate(4000, seed=77, median_bytes=1500)
    shared = GpuBatchPipeline(CHAINS["baseline_order"])
    own = [GpuBatchPipeline(CHAINS[name]) for name in ("shipped_order", "c4_then_open", "gr_shipped")]
    jobs = [(shared, 0, 1500), (shared, 1000, 3000), (shared, 2500, 4000)] + [(p, 500 * k, 500 * k + 2500) for k, p in enumerate(own)]
    serial = [p.run_packed(utf8, off[a:b + 1]) for p, a, b in jobs]
    got = [None] * len(jobs)

    def work(k):
        p, a, b = jobs[k]
        for _ in range(3):
            got[k] = p.run_packed(utf8, off[a:b + 1])

    th = [threading.Thread(target=work, args=(k,)) for k in range(len(jobs))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for r, ref in zip(got, serial):
        assert r is not None and np.array_equal(r.keep, ref.keep) and np.array_equal(r.failed_step, ref.failed_step)
        _same_stats(r, ref)
        assert np.array_equal(r.content, ref.content) and np.array_equal(r.content_offsets, ref.content_offsets)
    for p in [shared] + own:
        p.close()


def test_empty_and_degenerate_batches():
    p = pipe("baseline_order")
    r = p.run_packed(np.zeros(0, np.uint8), np.zeros(1, np.int64))
    assert r.n_docs == 0 and len(r.keep) == 0
    check_texts("baseline_order", ["", "", ""])
    check_texts("gr_open", ["", " ", "\n"])
    check_texts("gq_open", [""])


def test_long_documents_ngram_sweep():
    """BASELINE.json config 3 in miniature: 10-200 KB documents, n = 2..10 for both n-gram lists;
    the per-document tables exceed shared memory and spill to the global work area."""
    sweep = GopherRepetitionParams(dup_line_frac=0.3, dup_para_frac=0.3, dup_line_char_frac=0.2, dup_para_char_frac=0.2,
                                   top_n_grams=[(n, 0.2) for n in range(2, 11)], dup_n_grams=[(n, 0.15) for n in range(2, 11)])
    rng = random.Random(9)
    sizes = [int(10_000 * (20 ** rng.random())) for _ in range(60)]
    texts = [fuzzgen.english_doc(rng, sz) for sz in sizes]
    p = GpuBatchPipeline([sweep])
    res = p.run_texts(texts)
    ores = oracle.OraclePipeline([sweep]).run_texts(texts, nthreads=NCPU)
    compare(ores, res.keep, res.failed_step, res.stats, None, [_abi.TB_STEP_GOPHER_REPETITION], texts, label="sweep")
    assert res.stats[0]["n_words"].max() > 10000
    p.close()


@pytest.mark.parametrize("name", ["gq_open", "gq_shipped", "gr_open", "c4_lax", "c4_then_open", "shipped_order"])
def test_multilingual_stress(name):
    """BASELINE.json config 5: bit-exact decisions on non-ASCII-heavy text (rule-based scripts)."""
    check_texts(name, multilingual.corpus(31, 700))


def test_complex_script_subset_is_flagged():
    """Han / Kana / Thai: plain UAX #29 on both sides (bit-exact vs the oracle) and flagged, because
    ICU4X new_auto() would use dictionary / LSTM data we do not have (declared divergence)."""
    texts = multilingual.corpus(32, 90, langs=["han", "kana", "thai"])
    for name in ("gq_open", "gr_open"):
        res = check_texts(name, texts)
        assert (res.stats[0]["flags"] & _abi.TB_DOCFLAG_COMPLEX_SCRIPT).all()


# ---- the reference's own unit tests, driven through the ProcessingStep mirror -------------------
def _doc(i, c):
    return TextDocument(id=i, content=c, source="test_source")


def test_c4_step_reference_cases():  # c4_filters.rs:594-790
    f = C4QualityFilter(True, True, True, 5, 3, 1000, True, True, True, True)
    assert f.name() == "C4QualityFilter"
    for text, expected in C4_LINE_CASES:
        out = f.process(_doc("d", text))
        assert out.content.strip() == expected and out.metadata["c4_filter_status"] == "passed"
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("fail_sentences", "One sentence. Two sentences. Three sentences. Four sentences."))
    assert "too_few_sentences (found 4, required 5)" in e.value.reason
    assert e.value.document.metadata["c4_filter_status"] == "filtered"
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("lorem", "This is fine. Lorem ipsum dolor sit amet. This is also fine."))
    assert e.value.reason == "lorem_ipsum" and e.value.document.content.startswith("This is fine. Lorem")
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("curly", "This is fine.\nSome code block {}.\nAnother good line."))
    assert "curly_bracket" in e.value.reason


def test_gopher_quality_step_reference_cases():  # gopher_quality.rs:360-829
    f = GopherQualityFilter(min_doc_words=3)
    assert f.name() == "GopherQualityFilter"
    assert f.process(_doc("a", "Hello world test . !")).metadata["gopher_quality_filter_status"] == "passed"
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("b", "Hello world . !"))
    assert "gopher_short_doc (2 non-symbol words, required 3)" in e.value.reason
    f = GopherQualityFilter(min_avg_word_length=3.0, max_avg_word_length=5.0)
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("c", ". ! ."))
    assert "gopher_below_avg_threshold (avg len 0.00, required 3.00 - 0 non-symbol words)" in e.value.reason
    f = GopherQualityFilter(min_stop_words=1, stop_words=["custom", "words"])
    f.process(_doc("d", "this is a custom test with other words"))
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("e", "this is a regular sentence"))
    assert "gopher_too_few_stop_words (found 0, required 1)" in e.value.reason


def test_gopher_repetition_step_reference_cases():  # gopher_rep.rs:412-642
    f = GopherRepetitionFilter(dup_line_frac=0.3)
    assert f.name() == "GopherRepetitionFilter"
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("l", "This is line one.\nThis is line two.\nThis is line one."))
    assert "dup_line_frac (ratio 0.33, max 0.30)" in e.value.reason
    f = GopherRepetitionFilter(top_n_grams=[(2, 0.3)])
    f.process(_doc("p", "a b c d e f a b g h i j"))
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("q", "a b c a b d a b e a b"))
    assert "top_2_gram (ratio 0.57, max 0.30)" in e.value.reason
    f = GopherRepetitionFilter(dup_n_grams=[(2, 0.1)])
    with pytest.raises(DocumentFiltered) as e:
        f.process(_doc("r", "a b c d e a b f g"))
    assert "duplicated_2_n_grams (ratio 0.12, max 0.10)" in e.value.reason
    with pytest.raises(DocumentFiltered) as e:
        GopherRepetitionFilter().process(_doc("s", "   "))
    assert e.value.reason == "skipping empty content"
    assert e.value.document.metadata["gopher_repetition_filter_reason"] == "skipping empty content"


def test_fused_executor_short_circuit_and_order():  # executor.rs:30-70
    steps = [C4QualityFilter(True, True, True, 5, 3, 1000, True, True, True, True), GopherQualityFilter(min_doc_words=20),
             GopherRepetitionFilter(dup_line_frac=0.3)]
    ex = PipelineExecutor(steps)
    good = "This is the first sentence. This is the second sentence. This is the third sentence. This is the fourth sentence. This is the fifth sentence."
    docs = [_doc("0", "Some code block {}."), _doc("1", "One. Two. Three. Four. Five sentences are here now."), _doc("2", good)]
    out = ex.run_batch(docs)
    assert isinstance(out[0], StepError) and out[0].step_name == "C4QualityFilter"
    assert isinstance(out[1], StepError) and out[1].step_name == "GopherQualityFilter"
    assert "gopher_quality_filter_status" not in out[0].source.document.metadata
    assert out[1].source.document.metadata["c4_filter_status"] == "passed"
    assert "gopher_repetition_filter_status" not in out[1].source.document.metadata
    assert isinstance(out[2], TextDocument) and out[2].id == "2"
    assert out[2].metadata == {"c4_filter_status": "passed", "gopher_quality_filter_status": "passed",
                               "gopher_repetition_filter_status": "passed"}
    oc = ex.run_outcomes(docs)
    assert [o.kind for o in oc] == ["Filtered", "Filtered", "Success"]


def test_size_independent_properties_at_scale():
    """Properties that hold at any size (no oracle): determinism, batch-splitting invariance, and
    idempotence of C4's rewrite on this corpus (a kept document run again is kept unchanged; this
    holds because the generator never starts a line with a citation)."""
    utf8, off = synth.generate(60000, seed=synth.BASE_SEED + 4)
    p = GpuBatchPipeline(CHAINS["baseline_order"])
    r1 = p.run_packed(utf8, off)
    r2 = p.run_packed(utf8, off)
    assert np.array_equal(r1.keep, r2.keep) and np.array_equal(r1.content, r2.content)
    _same_stats(r1, r2)
    # halves == whole
    h = 30000
    ra = p.run_packed(utf8[: off[h]], off[: h + 1])
    rb = p.run_packed(utf8, off[h:])
    assert np.array_equal(np.concatenate([ra.keep, rb.keep]), r1.keep)
    assert np.array_equal(np.concatenate([ra.failed_step, rb.failed_step]), r1.failed_step)
    # idempotence: C4 applied to its own output changes nothing and keeps every kept document
    c4 = GpuBatchPipeline(CHAINS["c4_shipped"])
    first = c4.run_packed(utf8, off)
    kept = np.nonzero(first.keep)[0]
    lens = np.diff(first.content_offsets)[kept]
    off2 = np.zeros(len(kept) + 1, np.int64)
    off2[1:] = np.cumsum(lens)
    buf2 = np.concatenate([first.content[first.content_offsets[i]:first.content_offsets[i + 1]] for i in kept[:5000]]) if len(kept) else np.zeros(0, np.uint8)
    off2 = off2[:min(len(kept), 5000) + 1]
    second = c4.run_packed(buf2, off2)
    assert second.keep.all()
    assert np.array_equal(second.content, buf2) and np.array_equal(second.content_offsets, off2)
    p.close()
    c4.close()


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "c4_shipped", "c4_lax", "c4_nopunct", "c4_then_open"])
def test_content_as_runs(name):
    """TB_OPT_RETURN_CONTENT = 2: the content comes back as runs of the caller's own text (C4's output is a subsequence of
    its input, c4_filters.rs:192-258).  Gathered from the input column (tb_gather_content) it must equal the oracle's
    content byte for byte; statistics and decisions are those of the byte mode.  Fuzzed, boundary-sliding, multi-tile
    and synthetic documents; one chunk and many."""
    steps = CHAINS[name]
    rng = random.Random(41)
    texts = (GOLDEN_DOCS + fuzzgen.FuzzGen(77).docs(400) + fuzzgen.boundary_docs() +
             [fuzzgen.english_doc(rng, sz) for sz in (900, 1024, 3000, 20000, 70000)] + multilingual.corpus(4, 60))
    p = GpuBatchPipeline(steps)
    p.set_option(_abi.TB_OPT_RETURN_CONTENT, 2)
    for chunk in (1 << 30, 1 << 16):
        p.set_option(_abi.TB_OPT_CHUNK_BYTES, chunk)
        res = p.run_texts(texts)
        assert res.runs is not None and res.content is not None
        ores = oracle.OraclePipeline(steps).run_texts(texts, nthreads=NCPU)
        contents = [res.content_of(i) for i in range(res.n_docs)]
        compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name + "/runs")
        # runs are maximal, ordered and inside their document
        lens = np.array([len(t.encode("utf-8")) for t in texts], np.int64)
        for d in range(res.n_docs):
            r = res.runs[int(res.run_offsets[d]):int(res.run_offsets[d + 1])].astype(np.int64)
            assert (r[:, 1] > r[:, 0]).all() and (r[:, 1] <= lens[d]).all() and (r[1:, 0] > r[:-1, 1]).all()
    utf8, off = synth.generate(20000, seed=synth.BASE_SEED + 5)
    p.set_option(_abi.TB_OPT_CHUNK_BYTES, 8 << 20)
    res = p.run_packed(utf8, off)
    ores = oracle.OraclePipeline(steps).run_packed(utf8, off.astype(np.int64), nthreads=NCPU)
    compare(ores, res.keep, res.failed_step, res.stats, None, step_types(steps), None, label=name + "/runs/synthetic")
    ob, oo = ores.content_packed()
    assert np.array_equal(res.content_offsets, oo) and np.array_equal(res.content, ob)
    # far fewer bytes cross the link than the content holds
    assert res.runs.nbytes < 0.2 * max(len(ob), 1)
    p.close()


def test_content_as_runs_select_and_fallback():
    """tb_gather_content with a selection (the kept documents, as the writer of the output file wants them), and the
    fall-back to bytes where the text C4 reads is not the caller's (sentences as lines, entity pre-pass)."""
    from textblaster_b200.executor import gather_content
    from textblaster_b200 import _native
    steps = CHAINS["baseline_order"]
    utf8, off = synth.generate(3000, seed=synth.BASE_SEED + 6)
    p = GpuBatchPipeline(steps)
    p.set_option(_abi.TB_OPT_RETURN_CONTENT, 2)
    out = _abi.tb_batch_out()
    p.run_raw(utf8.ctypes.data, off.ctypes.data, True, len(off) - 1, out)
    assert out.content_kind == 2 and not out.content
    keep = np.ctypeslib.as_array(out.keep, (len(off) - 1,)).copy()
    got, goff = gather_content(out, utf8, off, select=keep, nthreads=3)
    p.release(out)
    ores = oracle.OraclePipeline(steps).run_packed(utf8, off.astype(np.int64), nthreads=NCPU)
    assert np.array_equal(keep, ores.keep)
    want = [ores.content(i) for i in np.flatnonzero(keep)]
    assert len(goff) == len(want) + 1
    for k, w in enumerate(want[:400]):
        assert bytes(got[goff[k]:goff[k + 1]]).decode("utf-8") == w
    assert int(goff[-1]) == sum(len(w.encode("utf-8")) for w in want)
    # TB_OPT_GATHER_KEPT: the library gathers the kept documents itself, chunk by chunk behind the device
    p.set_option(_abi.TB_OPT_GATHER_KEPT, 3)
    for chunk in (1 << 30, 1 << 20):
        p.set_option(_abi.TB_OPT_CHUNK_BYTES, chunk)
        p.run_raw(utf8.ctypes.data, off.ctypes.data, True, len(off) - 1, out)
        nk = int(out.n_kept)
        assert nk == int(keep.sum())
        koff = np.ctypeslib.as_array(out.kept_offsets, (nk + 1,)).copy()
        kbytes = np.ctypeslib.as_array(out.kept_content, (int(koff[-1]),)).copy() if koff[-1] else np.zeros(0, np.uint8)
        p.release(out)
        assert np.array_equal(koff, goff) and np.array_equal(kbytes, got)
    p.close()
    for chain, opt in (("c4_sentences", None), ("c4_shipped", _abi.TB_OPT_DECODE_HTML_ENTITIES)):
        q = GpuBatchPipeline(CHAINS[chain])
        q.set_option(_abi.TB_OPT_RETURN_CONTENT, 2)
        if opt:
            q.set_option(opt, 1)
        r = q.run_texts(["One sentence here. And a second one follows it. A third &amp; last one ends it. Four. Five is it."])
        assert r.runs is None and r.content is not None
        q.close()


def test_pageable_input_is_staged_by_the_library():
    """Text in ordinary (pageable) host memory -- what Arrow allocates -- goes through the library's pinned ring
    (TB_OPT_STAGE_THREADS); the results are those of the direct copy, with one chunk and with many."""
    utf8, off = synth.generate(30000, seed=synth.BASE_SEED + 7)   # ~100 MB: several slots of the ring
    steps = CHAINS["baseline_order"]
    p = GpuBatchPipeline(steps)
    p.set_option(_abi.TB_OPT_STAGE_THREADS, 0)
    ref = p.run_packed(utf8, off)
    for threads, chunk in ((4, 1 << 30), (1, 24 << 20), (7, 5 << 20)):
        p.set_option(_abi.TB_OPT_STAGE_THREADS, threads)
        p.set_option(_abi.TB_OPT_CHUNK_BYTES, chunk)
        r = p.run_packed(utf8, off)
        assert np.array_equal(r.keep, ref.keep) and np.array_equal(r.failed_step, ref.failed_step)
        _same_stats(r, ref)
        assert np.array_equal(r.content_offsets, ref.content_offsets) and np.array_equal(r.content, ref.content)
    ores = oracle.OraclePipeline(steps).run_packed(utf8, off.astype(np.int64), nthreads=NCPU)
    compare(ores, ref.keep, ref.failed_step, ref.stats, None, step_types(steps), None, label="staged")
    p.close()


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "c4_then_open", "gq_shipped"])
def test_bulk_copy_staging_of_long_documents(name):
    """TB_OPT_BULK_STAGE_TILES: the tiles of long documents arrive in shared memory by cp.async.bulk (tb_feed.cuh) instead of
    per-lane loads.  Same results on every document: long and short ones mixed, documents the mask path hands to the
    walking path in the middle (the ring must be drained), the first and last tiles of the buffer (not staged)."""
    steps = CHAINS[name]
    rng = random.Random(61)
    texts = [fuzzgen.english_doc(rng, sz) for sz in (300, 2100, 4096, 5000, 9000, 20000, 70000, 150000)]
    texts += fuzzgen.FuzzGen(62).docs(40, max_lines=200) + multilingual.corpus(6, 40)
    texts += ["plain words " * 900 + "a\u202fb mixed white space in a word here " + "tail words " * 700]   # exotic segment deep inside a long document
    texts += [fuzzgen.english_doc(rng, 30000)]
    p = GpuBatchPipeline(steps)
    for tiles in (2, 8):
        p.set_option(_abi.TB_OPT_BULK_STAGE_TILES, tiles)
        res = p.run_texts(texts)
        ores = oracle.OraclePipeline(steps).run_texts(texts, nthreads=NCPU)
        contents = [res.content_of(i) for i in range(res.n_docs)] if res.content is not None else None
        compare(ores, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=name + "/bulk%d" % tiles)
    p.close()


def repetition_pass_docs():
    """what the second repetition pass is sensitive to: every tier size, few and many candidates (register and table
    paths), split ambiguity at the pruned levels, the densest word pattern"""
    rng = random.Random(8)
    base = fuzzgen.english_doc(rng, 1500)
    docs = [fuzzgen.english_doc(rng, 2500), fuzzgen.english_doc(rng, 12000), fuzzgen.english_doc(rng, 36000), fuzzgen.english_doc(rng, 90000),
            (base + "\n") * 3, (base + "\n") * 9, (base + "\n") * 30, "a b c d e f g " * 700, "x y " * 5000]
    return docs + signature_pruning_docs() + split_ambiguity_docs() + fuzzgen.dense_word_docs() + level_task_docs()


@pytest.mark.parametrize("name", ["gr_open", "gr_shipped", "shipped_order"])
def test_repetition_pass_documents(name):
    check_texts(name, repetition_pass_docs())


def test_utf8_validation_option():
    """TB_OPT_VALIDATE_UTF8: the reference only ever sees Rust Strings; a caller that cannot vouch for its bytes has them
    checked on the device (the count of malformed documents comes back through a pinned word, chunk by chunk)."""
    from textblaster_b200.error import PipelineError
    good = ["plain text. more of it here.", "café naïve 中文 \U0001f600 fine."] * 50
    p = GpuBatchPipeline(CHAINS["gq_open"])
    p.set_option(_abi.TB_OPT_VALIDATE_UTF8, 1)
    assert p.run_texts(good).n_docs == 100
    enc = [t.encode("utf-8") for t in good]
    enc[37] = b"broken \xc3 lead without continuation"
    enc[80] = b"overlong \xc0\xaf and a lone \x80 byte"
    off = np.zeros(len(enc) + 1, np.int64)
    off[1:] = np.cumsum([len(e) for e in enc])
    blob = np.frombuffer(b"".join(enc), np.uint8)
    for chunk in (1 << 30, 1 << 16):
        p.set_option(_abi.TB_OPT_CHUNK_BYTES, chunk)
        with pytest.raises(PipelineError, match="invalid UTF-8"):
            p.run_packed(blob, off)
    p.set_option(_abi.TB_OPT_VALIDATE_UTF8, 0)
    p.close()


def test_repeated_runs_are_identical():
    """compute-sanitizer's racecheck is closed on the GPU pool this was developed on; what stands in for it: the
    documents that drive every table, register and walk path of the repetition pass and the C4 bitmap passes go through
    the chain twenty times, cut into chunks of different sizes (which changes which documents share an SM and in which
    order the persistent CTAs pick them up), and every run must reproduce the oracle's result in every field.  A barrier
    argued away wrongly shows up as a counter that differs once in a while."""
    rng = random.Random(23)
    texts = repetition_pass_docs() + [fuzzgen.english_doc(rng, rng.choice([300, 800, 1500, 3000, 6000])) for _ in range(300)] + fuzzgen.FuzzGen(31).docs(200)
    for name, mode in (("shipped_order", 2), ("gr_open", 1)):
        steps = CHAINS[name]
        want = oracle.OraclePipeline(steps).run_texts(texts, nthreads=os.cpu_count() or 1)
        p = GpuBatchPipeline(steps)
        p.set_option(_abi.TB_OPT_RETURN_CONTENT, mode)
        first = None
        for it in range(20):
            p.set_option(_abi.TB_OPT_CHUNK_BYTES, [1 << 30, 1 << 20, 1 << 18, 3 << 16][it % 4])
            res = p.run_texts(texts)
            if it < 2:
                contents = [res.content_of(i) for i in range(res.n_docs)] if res.content is not None else None
                compare(want, res.keep, res.failed_step, res.stats, contents, step_types(steps), texts, label=f"{name}/run{it}")
                if first is None:
                    first = (res.keep.copy(), res.failed_step.copy(), [a.copy() for a in res.stats])
            else:
                assert np.array_equal(res.keep, first[0]) and np.array_equal(res.failed_step, first[1]), f"{name}: run {it} differs"
                for si, (a, b) in enumerate(zip(res.stats, first[2])):
                    ran = executed(want.failed_step, si)  # (the statistics of a step that did not run for a document are not defined)
                    for f in a.dtype.names:
                        assert np.array_equal(a[f][ran], b[f][ran]), f"{name}: run {it}, step {si}, field {f} differs"
        p.close()


def test_streaming_runner_matches_oracle_and_single_call():
    """textblaster_b200/runner.py on the GPU: row groups given as views of one column (absolute offsets) go through three
    pipelines of one context, the kept documents' content rebuilt on the host; the summaries arrive in row-group order
    and equal the oracle's result (and so the single call's) document by document, content included."""
    from textblaster_b200.runner import RowGroup, StreamingRunner
    utf8, off = synth.generate(24000, seed=77)
    off = off.astype(np.int64)
    steps = CHAINS["baseline_order"]
    want = oracle.OraclePipeline(steps).run_packed(utf8, off, nthreads=os.cpu_count() or 1)
    n = len(off) - 1
    G = 7
    bounds = [(n * k // G, n * (k + 1) // G) for k in range(G)]
    runner = StreamingRunner(lambda: GpuBatchPipeline(steps), pipelines=3, copy_stats=True, gather_kept=2)
    seen = []

    def sink(index, s):
        a, b = bounds[index % G]
        seen.append(index)
        assert np.array_equal(s.keep, want.keep[a:b]) and np.array_equal(s.failed_step, want.failed_step[a:b]), f"group {index}"
        kept = np.nonzero(want.keep[a:b])[0]
        assert len(s.kept_offsets) == len(kept) + 1
        for j, d in enumerate(kept[:: max(1, len(kept) // 200)]):  # a sample of the kept documents, byte for byte
            jj = int(np.searchsorted(kept, d))
            got = bytes(s.kept_content[int(s.kept_offsets[jj]):int(s.kept_offsets[jj + 1])])
            c = want.content(a + int(d))
            assert got == (c.encode("utf-8") if isinstance(c, str) else bytes(c)), f"group {index}, document {a + int(d)}"
        for si, t in enumerate(step_types(steps)):
            ran = executed(want.failed_step[a:b], si)
            for f in s.stats[si].dtype.names:
                assert np.array_equal(s.stats[si][f][ran], want.stats[si][f][a:b][ran]), f"group {index}, step {si}, field {f}"

    tot = runner.run((RowGroup(i, utf8, off[bounds[i % G][0]:bounds[i % G][1] + 1]) for i in range(2 * G)), sink)
    runner.close()
    assert seen == list(range(2 * G)) and tot.docs == 2 * n and tot.kept == 2 * int(want.keep.sum())
