"""PipelineExecutor semantics with plain Python steps (no GPU) -- mirrors the reference's
tests/executor_test.rs and tests/worker_tests.rs (cited per test)."""
import pytest

from textblaster_b200.data_model import TextDocument
from textblaster_b200.error import DocumentFiltered, PipelineError, StepError, Unexpected
from textblaster_b200.executor import PipelineExecutor, ProcessingStep


class MockStep(ProcessingStep):  # executor_test.rs:19-67
    def __init__(self, name, fn):
        self._name, self._fn = name, fn

    def name(self):
        return self._name

    def process(self, document):
        return self._fn(document)


def append(s):
    def f(doc):
        doc.content += s
        return doc
    return f


def doc(i, c):
    return TextDocument(id=i, content=c, source="test_source")


def test_empty_pipeline_passes_through():  # :115-124, :229-247
    ex = PipelineExecutor([])
    assert ex.run_single(doc("doc1", "initial content")).content == "initial content"
    out = ex.run_batch([doc("doc1", "content1"), doc("doc2", "content2")])
    assert [d.content for d in out] == ["content1", "content2"]
    assert ex.run_batch([]) == []  # :213-226


def test_steps_apply_in_order():  # :127-186
    ex = PipelineExecutor([MockStep("step1_append", append(" + step1")), MockStep("step2_append", append(" + step2"))])
    assert ex.run_single(doc("doc1", "initial")).content == "initial + step1 + step2"


def test_error_is_wrapped_with_step_name_and_short_circuits():  # :189-215
    ran = []

    def boom(d):
        raise Unexpected("Simulated error")

    def never(d):
        ran.append(1)
        return d

    ex = PipelineExecutor([MockStep("ok_step", append(" + ok")), MockStep("error_step", boom), MockStep("step3", never)])
    with pytest.raises(StepError) as e:
        ex.run_single(doc("doc1", "initial"))
    assert e.value.step_name == "error_step" and not ran
    assert str(e.value).startswith("Error in processing step 'error_step': ")


def test_batch_results_keep_input_order_and_mix_errors():  # :250-291, :294-349
    def maybe(d):
        if d.id == "doc2_error":
            raise DocumentFiltered(d, "Specific error for doc2")
        return append(" + ok")(d)

    ex = PipelineExecutor([MockStep("step_ok", maybe)])
    out = ex.run_batch([doc("doc1", "a"), doc("doc2_error", "b"), doc("doc3", "c")])
    assert out[0].content == "a + ok" and out[2].content == "c + ok"
    assert isinstance(out[1], StepError) and out[1].step_name == "step_ok"
    assert isinstance(out[1].source, DocumentFiltered) and out[1].source.reason == "Specific error for doc2"
    assert str(out[1].source) == "Document 'doc2_error' filtered out: Specific error for doc2"


def test_outcome_mapping():  # worker_tests.rs:60-160 / worker_logic.rs:155-180
    def filt(d):
        d.metadata["k"] = "v"
        raise DocumentFiltered(d, "Filtered by FilteringStep")

    def fail(d):
        raise Unexpected("boom")

    ok = PipelineExecutor([MockStep("IdentityStep", lambda d: d)]).run_outcomes([doc("1", "x")])[0]
    assert ok.kind == "Success" and ok.document.content == "x"
    fl = PipelineExecutor([MockStep("FilteringStep", filt)]).run_outcomes([doc("1", "x")])[0]
    assert fl.kind == "Filtered" and fl.reason == "Filtered by FilteringStep" and fl.document.metadata == {"k": "v"}
    assert PipelineExecutor([MockStep("FailingStep", fail)]).run_outcomes([doc("1", "x")])[0] is None


def test_chunk_plan_ramps():
    """Chunk sizes of the host executor (tb_chunk_targets): they cover the batch exactly, never exceed the configured
    size, about double on the way up (the next chunk's copy fits behind the current chunk's kernels), shrink by about
    thirds on the way down, and a small batch is one chunk."""
    import _emul
    MiB = 1 << 20
    for total in (0, 1, 5 * MiB, 64 * MiB, 200 * MiB, 1000 * MiB, 3534533291, 40 * 1024 * MiB):
        for C in (1 << 16, 64 * MiB, 256 * MiB, 1024 * MiB):
            for dev in (False, True):
                if total // C > 3000:
                    continue
                t = _emul.chunk_targets(total, C, dev)
                assert (sum(t) == total or total == 0) and all(0 < x <= C for x in t), (total, C, dev, t)
                peak = t.index(max(t))
                for a, b in zip(t[:peak], t[1:peak + 1]):
                    assert b <= 3 * a or a >= C // 2, (total, C, dev, t)   # up: doubling (a remainder may stretch one step)
                last = len(t) - 1 - t[::-1].index(max(t))
                for a, b in zip(t[last:], t[last + 1:]):
                    assert 4 * b >= a or b >= C // 3, (total, C, dev, t)   # down: thirds (a remainder may stretch one step)
    assert _emul.chunk_targets(5 * MiB, 1024 * MiB) == [5 * MiB]
    t = _emul.chunk_targets(3534533291, 1024 * MiB)
    assert t[0] == 64 * MiB and t[-1] == 48 * MiB and max(t) <= 1024 * MiB and len(t) == 9
    assert _emul.chunk_targets(3534533291, 1024 * MiB, True)[0] == 256 * MiB
