"""The host-side formatter (textblaster_b200/reasons.py) must rebuild the reference's reason strings
and metadata keys from the integer statistics alone.  Checked against the oracle, which builds the
strings the way the reference does (format! with {:.2})."""
import pytest

import fuzzgen
import oracle
from parity_cases import CHAINS
from test_emul_parity import GOLDEN_DOCS
from textblaster_b200.reasons import step_metadata


@pytest.mark.parametrize("name", sorted(CHAINS))
def test_reason_strings_and_metadata(name):
    steps = CHAINS[name]
    texts = GOLDEN_DOCS + fuzzgen.FuzzGen(21).docs(150)
    r = oracle.OraclePipeline(steps).run_texts(texts)
    for i in range(len(texts)):
        fs = int(r.failed_step[i])
        last = fs if fs >= 0 else len(steps) - 1
        md, reason = {}, ""
        for s in range(last + 1):
            m, rs = step_metadata(steps[s], r.stats[s][i])
            md.update(m)
            if s == fs:
                reason = rs
        assert reason == r.reason(i), (i, texts[i][:200])
        assert md == r.metadata(i), (i, texts[i][:200])
