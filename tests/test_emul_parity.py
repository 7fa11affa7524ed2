"""CPU check of the kernel logic: the exact warp bodies of the CUDA kernels, run by the host emulation
harness (tests/host_emul), against the oracle -- bit-exact counters, decisions and rewritten content.
The real kernels are checked the same way on the GPU in test_gpu_parity.py."""
import random

import pytest

import _emul
import fuzzgen
import oracle
from parity_cases import CHAINS, step_types
from parity_util import compare
from test_oracle_golden import C4_LINE_CASES

GOLDEN_DOCS = [d for d, _ in C4_LINE_CASES] + [
    "This is fine. Lorem ipsum dolor sit amet.", "code {}.", "", "   \n   ",
    "One sentence. Two sentences. Three sentences. Four sentences.",
    "Foo bar baz. [1]\nThis line stays here.", "Keep this line please.\nno punct here\nab.\n",
    "a b c d e f a b g h i j", "a b c a b d a b e a b", "a b c d e a b f g",
    "word1 ... word2 … word3 word4 word5 word6 word7 word8", "- item 1\n- item 2\n- item 3\nnormal line",
    "Line one...\nLine two…\nLine three...\nNormal line", "the quick brown fox and the lazy dog", "Hello world . !",
    "This is line one.\nThis is line two.\nThis is line one.", "P one.\n\nP two.\n\nP one.", "\n", "\n\n", "a\n", "a\n\n",
    "x", "#", ". ! ?", "word 123 456 !!!",
]


def _check(name, texts):
    steps = CHAINS[name]
    o = oracle.OraclePipeline(steps).run_texts(texts)
    e = _emul.run_chain(steps, texts)
    compare(o, e.keep, e.failed_step, e.stats, e.contents, step_types(steps), texts, label=name)


@pytest.mark.parametrize("name", sorted(CHAINS))
def test_golden_documents(name):
    _check(name, GOLDEN_DOCS)


@pytest.mark.parametrize("name", sorted(CHAINS))
def test_fuzzed_documents(name):
    _check(name, fuzzgen.FuzzGen(11).docs(120))


@pytest.mark.parametrize("name", ["baseline_order", "shipped_order", "c4_then_open", "gr_open", "gq_open", "fw_open", "fw_custom"])
def test_multi_tile_documents(name):
    rng = random.Random(3)
    texts = [fuzzgen.english_doc(rng, sz) for sz in (200, 900, 1023, 1024, 1025, 1100, 3000, 8000, 20000)]
    texts += [fuzzgen.FuzzGen(4).doc(max_lines=80) for _ in range(10)]
    _check(name, texts)


@pytest.mark.parametrize("name", ["gq_open", "gq_shipped", "gr_open", "c4_lax", "c4_then_open", "baseline_order"])
def test_ascii_fast_path(name):
    rng = random.Random(17)
    texts = [fuzzgen.ascii_doc(rng, rng.choice([5, 40, 400, 1500])) for _ in range(60)]
    texts += ["a" * 31 + " " + "b" * 40, "x" * 100, "ab " * 200, "a.b" * 30, "1,2" * 40, "it's " * 50, "a\r\nb c\r\n" * 20]
    _check(name, texts)


@pytest.mark.parametrize("name", ["gr_shipped", "gr_open", "gq_shipped", "gq_open", "baseline_order", "shipped_order", "fw_open"])
def test_dense_word_documents(name):  # two words per three bytes: the bound the word slots are sized by
    _check(name, fuzzgen.dense_word_docs())


def test_repetition_tiers():
    """tb_rep.cuh picks shared-memory (warp / CTA) or global work areas by document size; a document
    with many repeated n-grams exercises the candidate walk in each tier."""
    rng = random.Random(8)
    base = fuzzgen.english_doc(rng, 1500)
    texts = [fuzzgen.english_doc(rng, 2500), fuzzgen.english_doc(rng, 12000), fuzzgen.english_doc(rng, 36000),
             (base + "\n") * 3, (base + "\n") * 9, (base + "\n") * 30, "a b c d e f g " * 700, "x y " * 5000]
    _check("gr_open", texts)
    _check("gr_shipped", texts)


def test_word_spanning_many_tiles():
    # a single segment longer than several tiles is walked by the lane that owns its first byte
    texts = ["start " + "a" * 5000 + " end. Another sentence here. And a third one.", "a" * 3000, "é" * 1500 + " x"]
    for name in ("gq_open", "gr_open", "c4_lax"):
        _check(name, texts)


@pytest.mark.parametrize("name", ["c4_lax", "c4_shipped", "gq_open", "gr_open", "baseline_order", "fw_open", "fw_custom", "fw_shipped"])
def test_boundary_documents(name):
    _check(name, fuzzgen.boundary_docs())


def test_general_word_path_matches(monkeypatch):
    """The segment-walking bodies (the general path behind the mask-based passes) give the same results."""
    monkeypatch.setenv("TBE_FORCE_WALK", "1")
    rng = random.Random(5)
    texts = fuzzgen.FuzzGen(21).docs(40) + [fuzzgen.english_doc(rng, 2500)] + fuzzgen.boundary_docs()[:30]
    for name in ("baseline_order", "gq_open", "gr_open", "fw_open"):
        _check(name, texts)


SENTENCE_LINE_DOCS = [
    "a?b?c?d", "?a?a?a", "Hello.World and more", "Stop! Go now? Yes.", "foo.\n\nbar.", "x one two.\r\ny three four.", "  \n  ",
    "One two three.  Four five six.\nSeven eight nine!\n\n\n  Ten eleven twelve?  ", "a.\rb c d.\r\re f g.",
    "He said \"stop.\" Then left (quickly.) And more words here.", "e.g. this is not a break. But this one is.",
    "Dr. Smith went home. 3.5 apples fell.", "Line\u2028sep here now.\u2029Para sep here now.\u0085Next line here now.",
    "Wait...really? Yes!!! No?! Sure.", "\u3053\u308c\u306f\u6587\u3067\u3059\u3002\u6b21\u306e\u6587\u3067\u3059\u3002final one here.",
    "ends with spaces and newline.   \n", "\n\n  leading newlines here. And more text.", "no terminator at all",
    "This has a citation [1]. And another [2, 3] one here. Please enable javascript now. We use cookies here.",
    "a.\u00a0b c d. e f g.\u3000h i j.", "x" * 40 + "? " + "y" * 1100 + ". z z z.",
    " ".join("Sentence number %d is here." % i for i in range(120)),
    "\n".join("w%d? next%d" % (i, i) for i in range(300)),
]


def boundary_sentence_docs():
    """terminators and separators around the 32-byte stretch and 1 KiB tile edges"""
    out = []
    for pad in (28, 29, 30, 31, 32, 33, 1020, 1021, 1022, 1023, 1024, 1025):
        for sep in ("? ", "?", ".\n", ". \n ", "!\r\n", ".\u2029", "?\u00a0", ".) "):
            out.append("a" * pad + sep + "Bcd efg hij." + sep + "k l m.")
    return out


@pytest.mark.parametrize("name", ["c4_sentences", "c4_sentences_lax", "sentences_then_open"])
def test_sentence_wise_lines(name):  # split_paragraph = false, c4_filters.rs:152-155
    _check(name, SENTENCE_LINE_DOCS + boundary_sentence_docs())


def entity_docs():
    """HTML entities of every kind, well-formed and not, also sliding across the 32-byte stretch and 1 KiB tile edges"""
    import html.entities
    rng = random.Random(41)
    names = sorted(k for k in html.entities.html5 if k.endswith(";"))
    docs = [
        "AT&T said &amp; co. Fish &amp; chips &lt;b&gt;bold&lt;/b&gt; &quot;quoted&quot; &apos;x&apos; &nbsp;done.",
        "&#65;&#x41;&#X1F600; &nLt;&nGt; &bogus; &amp &#; &#x; &#xD800; &#1114112; &#0;x &#00000000000000000065; &#99999999999; &;&&amp;&CounterClockwiseContourIntegral;",
        "&", "&&&&", "&amp", "&amp;", "x&amp;", "&amp;&amp;&amp;&amp;&amp;&amp;&amp;&amp;&amp;&amp;", "&nLt;&nGt;&nLt;&nGt;&nLt;&nGt;&nLt;",
        "&#x000000000000000000000000000000000041;", "&#x0000000000000000000000000000000041;", "&AMP;&aMp;&Amp;", "&lt;&#10;&gt;\n&#13;&#10;line two.&#x2028;three",
        "Caf&eacute; na&iuml;ve &euro;5 &mdash; &hellip; &ldquo;So it goes.&rdquo; He said &#8220;stop.&#8221; Then left. And more words here now.",
        "", "no entities at all. Just text. And more text.",
    ]
    for pad in (26, 27, 28, 29, 30, 31, 32, 33, 1015, 1018, 1020, 1022, 1023, 1024, 1025):
        for e in ("&amp;", "&nLt;", "&#x1F600;", "&CounterClockwiseContourIntegral;", "&bogus;", "&#12"):
            docs.append("a" * pad + e + " Bcd efg hij." + e + e + " k l m.")
    for _ in range(60):  # fuzzed documents with entities sprinkled in
        words = fuzzgen.FuzzGen(rng.randrange(1 << 30)).doc(max_lines=10).split(" ")
        for _ in range(rng.randrange(1, 12)):
            k = rng.randrange(len(words) + 1)
            kind = rng.randrange(6)
            e = ("&" + rng.choice(names) if kind < 2 else "&#%d;" % rng.choice([38, 60, 233, 8212, 128512, 55296, 1114112, 0]) if kind == 2
                 else "&#x%X;" % rng.randrange(0x30000) if kind == 3 else "&" + rng.choice(names)[:-1] if kind == 4 else "&" + rng.choice(["", "#", "#x", "zz;", " amp;"]))
            words.insert(k, e)
        docs.append(" ".join(words))
    return docs


@pytest.mark.parametrize("name", ["baseline_order", "gq_open", "gr_open", "c4_lax", "sentences_then_open", "fw_open"])
def test_html_entities_decoded_first(name):  # parquet_reader.rs:178-179, parity with html-escape unpinned
    steps = CHAINS[name]
    texts = entity_docs()
    decoded = [oracle.decode_html_entities(t) for t in texts]
    o = oracle.OraclePipeline(steps).run_texts(decoded)
    e = _emul.run_chain(steps, texts, decode_entities=True)
    compare(o, e.keep, e.failed_step, e.stats, e.contents, step_types(steps), texts, label=name + "/entities")


def split_ambiguity_docs():
    """find_all_duplicate keys are bare concatenations (text.rs:250): "no where" and "now here" give the same
    key although the word sequences differ, and such a pair need not share any shorter repeated n-gram.
    Tiny alphabets make these collisions frequent."""
    rng = random.Random(12)
    toks = ["a", "b", "ab", "ba", "aa", "aba", "bab", "no", "now", "where", "here", "w", "he", "re"]
    docs = ["is no where to be seen today but is now here to be seen today ok",
            "a ba b ab a ba b ab x a ba", "ab a ab a a ba a ba ab a"]
    for _ in range(40):
        docs.append(" ".join(rng.choice(toks) for _ in range(rng.choice([8, 30, 120, 700]))))
    return docs


def signature_pruning_docs():
    """The duplicate n-gram levels only visit word ends whose suffix signature (last k key bytes, four bits of each)
    repeats (tb_rep.cuh).  Documents that stress the filter: characters that differ only in the bits the signature
    drops, shortest 5-grams around the switch-over lengths (k < 6: no pruning; 6..16; > 16), words longer than the
    16 bytes a signature holds, repeats that overlap themselves, equal concatenations with different word splits at
    the pruned levels, and every repeat pattern at each tier size."""
    rng = random.Random(31)
    docs = []
    same_low = ["a", "q", "1", "A", "Q", "\u00e1", "b", "r", "2", "B"]  # 0x61 0x71 0x31 0x41 0x51 ... : equal low nibbles
    for nw in (9, 40, 300, 700, 2500, 9000):
        docs.append(" ".join(rng.choice(same_low) * rng.randint(1, 3) for _ in range(nw)))
    for wl in (1, 2, 3, 4, 7, 17, 40):  # word length sets the shortest 5-gram: 5, 10, 15, 20, 35, 85, 200 bytes
        vocab = ["".join(rng.choice("abcdefgh") for _ in range(wl)) for _ in range(12)]
        for nw in (30, 600, 3000):
            docs.append(" ".join(rng.choice(vocab) for _ in range(nw)))
    phrase = "one two three four five six seven eight nine ten eleven twelve"
    filler = lambda n: " ".join("".join(rng.choice("klmnopqrstuvwxyz") for _ in range(rng.randint(2, 9))) for _ in range(n))
    for reps, gap in ((2, 3), (2, 40), (5, 7), (30, 1), (3, 400), (12, 90)):
        docs.append(" ".join(phrase + " " + filler(gap) for _ in range(reps)))
    docs.append(("x y z " * 4 + "no where to be seen at all today ") + filler(20) + (" x y z " * 2 + "now here to be seen at all today"))
    docs.append("ab c d e f g h i j k " + filler(15) + " a bc d e f g h i j k " + filler(5) + " a b cd e f g h i j k")
    docs.append("the end of the story is the end of the story is the end of the story " * 3)
    docs.append(" ".join(["w" * 40, "v" * 18] * 30 + ["u" * 16] * 12))
    docs.append("a b c d e f " * 200 + filler(50) + " a b c d e f" * 7)
    docs.append("\u00e9t\u00e9 \u00fcber na\u00efve caf\u00e9 \u4e2d\u6587 " * 25 + filler(30))
    for _ in range(12):
        toks = ["no", "now", "where", "here", "w", "he", "re", "to", "be", "seen", "a", "ll", "all", "al", "l"]
        docs.append(" ".join(rng.choice(toks) for _ in range(rng.choice([40, 400, 2400]))))
    return docs


def level_task_docs():
    """Documents of the CTA tiers (more than 512 words) whose pruned duplicate n-gram levels keep few word ends in play:
    there a warp works a whole level off alone (tb_rep.cuh) -- up to 32 ends in registers, up to 256 with a table in its
    share of the CTA's arrays, more than that the CTA does level by level.  Repeated blocks of several sizes at each tier."""
    rng = random.Random(77)
    docs = []
    for nw in (700, 1500, 3000, 6000, 12000):
        for block, reps in ((8, 2), (14, 3), (40, 2), (60, 3), (120, 2), (300, 2)):
            body = ["".join(rng.choice("klmnopqrstuvwxyz") for _ in range(rng.randint(2, 9))) for _ in range(nw)]  # (hardly any chance repeats)
            rep = body[10:10 + block]
            for r in range(reps):
                at = rng.randrange(len(body))
                body[at:at] = rep
            docs.append(" ".join(body))
    return docs


def test_signature_pruning():
    for name in ("gr_open", "gr_shipped"):
        _check(name, signature_pruning_docs())
    _emul.rep_task_stats()
    _check("gr_shipped", level_task_docs())
    small, table = _emul.rep_task_stats()
    assert small > 20 and table > 20, (small, table)
    # and it does prune: on C4-shaped prose only a few percent of the word ends stay in play
    from textblaster_b200 import synth
    utf8, off = synth.generate(300, seed=9)
    texts = [bytes(utf8[off[i]:off[i + 1]]).decode("utf-8") for i in range(300)]
    _emul.rep_stats()
    _check("gr_shipped", texts)
    words, items, docs = _emul.rep_stats()
    assert docs > 250 and items < 0.1 * words, (words, items, docs)


def test_ngram_split_ambiguity():
    _check("gr_open", split_ambiguity_docs())
    _check("gr_shipped", split_ambiguity_docs())


def fineweb_docs():
    """Line shapes FineWebQualityFilter is sensitive to: CR LF and bare CR, blank lines of odd white space,
    repeated lines (more than the 512 the shared-memory table of pass 2 holds), lines around the short-line
    length, stop characters behind trailing white space, very long lines, no final newline / only newlines."""
    rng = random.Random(23)
    docs = ["a\r\nb\r\na\r\n", "a\rb\n\r\n \u00a0\n\u3000 \nlast\r", "x" * 30 + ".\n" + "y" * 31 + ".  \n" + "z" * 29 + "\u201d \t\n",
            "same line here.\n" * 700 + "other.\n" * 3, ("l" * 5000 + ". \n") * 3, "\n\n\n", "word", "\u3002\n\u3002\n!!!\n",
            "Ends with an emoji \U0001f600\nEnds with a dot.\u2003\nEnds with an a\nEnds with ellipsis\u2026\n"]
    for _ in range(30):
        n = rng.choice([3, 20, 200])
        pool = [fuzzgen.english_doc(rng, rng.choice([10, 40, 120])).replace("\n", " ") for _ in range(max(2, n // 3))]
        eol = rng.choice(["\n", "\r\n", "\n\n", " \n"])
        docs.append(eol.join(rng.choice(pool) for _ in range(n)) + rng.choice(["", "\n", "\r"]))
    return docs


@pytest.mark.parametrize("name", ["fw_open", "fw_shipped", "fw_custom", "shipped_with_fineweb"])
def test_fineweb_documents(name):
    _check(name, fineweb_docs())
