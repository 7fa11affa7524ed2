"""Fuzz the oracle's UAX #29 restatement against ICU4C 71 (bundled with Nsight Compute).

ICU4C 71 is NOT the reference (that is ICU4X 1.5, absent here): it is a proxy that implements the
same published rules, so agreement on rule-based scripts is evidence the restatement is right.
Han/Kana/Thai are excluded (ICU4C re-segments them with a dictionary).  Skipped when the library
is not present.

The colon: ICU4C 71 (CLDR 41) still has U+003A in Word_Break=MidLetter; the tables here follow the CLDR >= 42 root
tailoring that takes it out (tools/gen_unicode_tables.py --colon cldr-root, the best-evidence setting for ICU4X 1.5).
For the word rules the proxy is therefore asked about the text with every ':' replaced by '+' (Word_Break=Other,
like the untailored colon); the sentence rules see the text as it is.
"""
import os
import random
import sys

import pytest

import oracle

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))
import icu4c_probe as icu  # noqa: E402

pytestmark = pytest.mark.skipif(not icu.available(), reason="ICU4C 71 not found")

ALPHABET = list("abcXYZ019 .,;:!?'\"_-()[]\n\r\t") + [
    # Latin-1, Cyrillic, Greek, Hebrew, Extend, ZWJ, Format (SHY), emoji, regional indicators, NBSP, NNBSP,
    # LS, ideographic space, Arabic-Indic digit, quotes, ellipsis, middle dot, heart + VS16, NEL, VT, em dash,
    # gershayim, Arabic comma / decimal separator, fullwidth full stop, one dot leader, ogham space, figure space
    "\xe6", "\xc5", "\u0416", "\u0436", "\u03a3", "\u03c3", "\u05d0", "\u05d1", "\u0301", "\u200d",
    "\xad", "\U0001f600", "\U0001f1e9", "\U0001f1f0", "\xa0", "\u202f", "\u2028", "\u3000", "\u0661", "\u2019",
    "\u201c", "\u201d", "\u2026", "\xb7", "\u2764", "\ufe0f", "\x85", "\x0b", "\u2014", "\u05f4",
    "\u060c", "\u066b", "\uff0e", "\u2024", "\x27", "\u1680", "\u2007", "\xdf",
]


def _cp_offsets(text, byte_offsets):
    m, k = {}, 0
    for i, ch in enumerate(text):
        m[k] = i
        k += len(ch.encode("utf-8"))
    m[k] = len(text)
    return [m[o] for o in byte_offsets]


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_word_and_sentence_boundaries_match_icu4c(seed):
    rng = random.Random(seed)
    bad = []
    for _ in range(3000):
        t = "".join(rng.choice(ALPHABET) for _ in range(rng.randint(1, 14)))
        for kind, fn in (("word", oracle.word_boundaries), ("sentence", oracle.sentence_boundaries)):
            a = _cp_offsets(t, fn(t))
            b = icu.breaks(t.replace(":", "+") if kind == "word" else t, kind)
            if a != b:
                bad.append((kind, [hex(ord(c)) for c in t], a, b))
    assert not bad, bad[:10]


def test_known_segmentations():
    assert icu.segments("mid...dle") == ["mid", ".", ".", ".", "dle"]
    for t in ["can't 3.14 foo_bar a.b 1,000 a;b", "Hello. How are you? Fine! ok", "e.g. this. (Really?) \"Yes.\" he said."]:
        assert _cp_offsets(t, oracle.word_boundaries(t)) == icu.breaks(t, "word")
        assert _cp_offsets(t, oracle.sentence_boundaries(t)) == icu.breaks(t, "sentence")


def test_colon_follows_the_cldr_root_tailoring():
    """U+003A / U+FE55 / U+FF1A are not MidLetter in the root word-break rules since CLDR 42 (only fi / sv keep them):
    a colon between letters separates two words.  The proxy (CLDR 41) still joins them -- the difference this setting is about."""
    assert oracle.split_into_words("foo:bar") == ["foo", "bar"]
    assert oracle.split_into_words("a:b:c") == ["a", "b", "c"]
    assert oracle.split_into_words("12:30 http://x.y key:value C:\\dir") == ["12", "30", "http", "x.y", "key", "value", "C", "dir"]
    # (U+FF1A is in the reference's PUNCTUATION set, U+FE55 is not: on its own it counts as a word, text.rs:139-146)
    assert oracle.split_into_words("a\uff1ab a\ufe55b") == ["a", "b", "a", "\ufe55", "b"]
    assert icu.segments("foo:bar") == ["foo:bar"]  # ICU4C 71: one segment
    assert oracle.split_into_words("don't a.b x\u00b7y") == ["don't", "a.b", "x\u00b7y"]  # the other Mid* characters still join
