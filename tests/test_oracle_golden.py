"""Pins the CPU oracle against every known-answer test the reference holds for the hot path
(SURVEY.md 8c).  Each test cites the reference test it restates (paths relative to the reference
repo).  Inputs and expected values are the reference's own golden vectors."""
import pytest

import oracle
from textblaster_b200.config import C4QualityParams, FineWebQualityParams, GopherQualityParams, GopherRepetitionParams


# ---------------------------------------------------------------------------------------------
# src/utils/text.rs tests
# ---------------------------------------------------------------------------------------------
def test_split_sentences_empty_and_simple():  # text.rs:266-305
    s = oracle.split_into_sentences
    assert s("") == []
    assert s("   ") == []
    assert s("Hello world.") == ["Hello world."]
    assert s("  Hello world.  ") == ["Hello world."]
    assert s("Dette er en sætning.") == ["Dette er en sætning."]
    assert s("SingleWord") == ["SingleWord"]
    assert s("  SingleWord  ") == ["SingleWord"]


def test_split_sentences_multiple():  # text.rs:308-344
    s = oracle.split_into_sentences
    exp = ["Første sætning.", "Anden sætning!", "Tredje sætning?"]
    assert s("Første sætning. Anden sætning! Tredje sætning?") == exp
    assert s("  Første sætning.   Anden sætning!  Tredje sætning?  ") == exp
    assert s(" Hello. How are you? Fine! ") == ["Hello.", "How are you?", "Fine!"]
    assert s("This is a sentence. This is another") == ["This is a sentence.", "This is another"]
    assert s("  This is a sentence.   This is another  ") == ["This is a sentence.", "This is another"]


def test_split_words_empty_and_simple():  # text.rs:347-351
    w = oracle.split_into_words
    assert w("") == []
    assert w("hello") == ["hello"]
    assert w("hello world") == ["hello", "world"]


def test_split_words_with_punctuation():  # text.rs:354-361, 429
    w = oracle.split_into_words
    assert w("hello, world!") == ["hello", "world"]
    assert w("first. second; third?") == ["first", "second", "third"]
    assert w("...leading") == ["leading"]
    assert w("trailing...") == ["trailing"]
    assert w("mid...dle") == ["mid", "dle"]


def test_split_words_danish():  # text.rs:439-442
    w = oracle.split_into_words
    assert w("hej med dig") == ["hej", "med", "dig"]
    assert w("en, to, tre!") == ["en", "to", "tre"]


def test_punctuation_set_contents():  # text.rs:445-457 (+ the 129-member count, text.rs:28-37)
    p = oracle.punctuation_contains
    for ch in ".,!?\"":
        assert p(ch)
    assert p(chr(0)) and p(chr(0x1F))
    for ch in "aA5":
        assert not p(ch)
    assert sum(1 for cp in range(0x110000) if oracle.cinfo(cp) & (1 << 10)) == 129
    assert not p("\t") and not p("\n") and not p(" ")
    assert p("１") and p("”") and p("…")


# ---------------------------------------------------------------------------------------------
# gopher_rep.rs helper tests
# ---------------------------------------------------------------------------------------------
def test_get_n_grams_logic():  # gopher_rep.rs:247-270
    words = ["a", "b", "c", "d"]
    assert oracle.get_n_grams(words, 2) == ["a b", "b c", "c d"]
    assert oracle.get_n_grams(words, 1) == ["a", "b", "c", "d"]
    assert oracle.get_n_grams(words, 4) == ["a b c d"]
    assert oracle.get_n_grams(words, 5) == []
    assert oracle.get_n_grams([], 2) == []
    assert oracle.get_n_grams(words, 0) == []


def test_find_duplicates_logic():  # gopher_rep.rs:273-297
    fd = oracle.find_duplicates
    assert fd(["a", "b", "c"]) == (0, 0)
    assert fd(["a", "b", "a"]) == (1, 1)
    assert fd(["ab", "cd", "ab", "ef", "cd"]) == (2, 4)
    assert fd(["a", "a", "a"]) == (2, 2)
    assert fd([]) == (0, 0)


def test_find_top_duplicate_logic():  # gopher_rep.rs:300-355
    ft = oracle.find_top_duplicate
    assert ft(["a", "a"]) == 2
    assert ft(["a", "a", "b", "b"]) == 2
    assert ft(["a b", "c d", "a b"]) == 6
    assert ft(["a", "b", "c"]) == 0
    assert ft(["aa", "aa", "b", "b"]) == 4
    assert ft(["a", "a", "a"]) == 3
    assert ft([]) == 0
    assert ft(["unique"]) == 0


def test_find_all_duplicate():  # gopher_rep.rs:358-408
    fa = oracle.find_all_duplicate
    nd = ["a", "b", "c", "d", "e"]
    assert fa(nd, 2) == 0 and fa(nd, 3) == 0
    assert fa(["a", "b", "c", "a", "b", "d"], 2) == 2
    assert fa(["a", "b", "a", "b", "a", "b"], 2) == 4
    assert fa(["a", "a", "a", "a", "a"], 2) == 4
    assert fa([], 2) == 0
    assert fa(nd, 0) == 0
    assert fa(nd, 6) == 0


# ---------------------------------------------------------------------------------------------
# C4QualityFilter, c4_filters.rs:578-840
# ---------------------------------------------------------------------------------------------
def c4_default(**kw):  # c4_filters.rs:578-591
    d = dict(split_paragraph=True, remove_citations=True, filter_no_terminal_punct=True, min_num_sentences=5,
             min_words_per_line=3, max_word_length=1000, filter_lorem_ipsum=True, filter_javascript=True,
             filter_curly_bracket=True, filter_policy=True)
    d.update(kw)
    return C4QualityParams(**d)


def run1(params, text):
    r = oracle.OraclePipeline([params]).run_texts([text])
    return bool(r.keep[0]), r.reason(0), r.content(0), r.metadata(0), r


def test_c4_document_passes():  # :594-607
    t = ("This is the first sentence. This is the second sentence. This is the third sentence. "
         "This is the fourth sentence. This is the fifth sentence.")
    ok, reason, content, md, _ = run1(c4_default(), t)
    assert ok and md["c4_filter_status"] == "passed"
    assert content.strip() == t.strip()


def test_c4_too_few_sentences():  # :610-621
    ok, reason, *_ = run1(c4_default(), "One sentence. Two sentences. Three sentences. Four sentences.")
    assert not ok and "too_few_sentences (found 4, required 5)" in reason


TAIL = " This is the fourth sentence. And the fifth sentence. Here is the sixth."

C4_LINE_CASES = [
    # :624-641 line too few words
    ("This line is fine.\nTwo words.\nAnother good line." + TAIL, "This line is fine.\nAnother good line." + TAIL),
    # :644-661 missing terminal punctuation
    ("This line is fine.\nThis one is not\nAnd this is okay. Here is another sentence. And a fifth one. This is the sixth sentence.",
     "This line is fine.\nAnd this is okay. Here is another sentence. And a fifth one. This is the sixth sentence."),
    # :664-681 ellipsis
    ("This line is fine.\nThis one ends with ellipsis...\nAnd this is okay." + TAIL, "This line is fine.\nAnd this is okay." + TAIL),
    # :684-702 word too long
    ("This line is fine.\nA line with a verylongword " + "a" * 1001 + ".\nAnother good line." + TAIL,
     "This line is fine.\nAnother good line." + TAIL),
    # :719-736 javascript
    ("This is fine.\nSome javascript code here.\nAnother good line." + TAIL, "This is fine.\nAnother good line." + TAIL),
    # :753-770 policy
    ("This is fine.\nRead our privacy policy.\nAnother good line." + TAIL, "This is fine.\nAnother good line." + TAIL),
    # :773-790 citations
    ("This is text [1]. Another sentence [2, 3]. Final text [45]. Here is the fourth sentence. And the fifth sentence. This is the sixth sentence.",
     "This is text . Another sentence . Final text . Here is the fourth sentence. And the fifth sentence. This is the sixth sentence."),
]


@pytest.mark.parametrize("doc,expected", C4_LINE_CASES)
def test_c4_line_filters(doc, expected):
    ok, reason, content, md, _ = run1(c4_default(), doc)
    assert ok, reason
    assert content.strip() == expected
    assert md["c4_filter_status"] == "passed"


def test_c4_lorem_ipsum():  # :705-716
    ok, reason, content, md, r = run1(c4_default(), "This is fine. Lorem ipsum dolor sit amet. This is also fine.")
    assert not ok and "lorem_ipsum" in reason
    assert md == {"c4_filter_status": "filtered", "c4_filter_reasons": "lorem_ipsum"}
    assert content == "This is fine. Lorem ipsum dolor sit amet. This is also fine."  # untouched on the early exit


def test_c4_curly_bracket():  # :739-750
    ok, reason, *_ = run1(c4_default(), "This is fine.\nSome code block {}.\nAnother good line.")
    assert not ok and "curly_bracket" in reason


def test_c4_empty_and_spaces():  # :793-817
    for t in ("", "   \n   "):
        ok, reason, *_ = run1(c4_default(), t)
        assert not ok and "too_few_sentences (found 0, required 5)" in reason


def test_c4_zero_min_values():  # :820-840
    p = C4QualityParams(True, False, False, 0, 0, 0, False, False, False, False)
    ok, *_ = run1(p, "Ok.")
    assert ok


def test_c4_quirks():  # SURVEY 8a a8 (i)-(viii), derived from c4_filters.rs:193-199,219-226,281-283
    # (i) citation removed but not re-trimmed -> "Foo bar baz. " -> last char ' ' -> dropped
    ok, reason, content, md, r = run1(c4_default(min_num_sentences=1), "Foo bar baz. [1]\nThis line stays here.")
    assert ok and content == "This line stays here."
    assert r.stats[0]["no_terminal_punc"][0] == 1
    # (viii) late filtered path: rewritten content + line-filter counters in metadata
    ok, reason, content, md, r = run1(c4_default(), "Keep this line please.\nno punct here\nab.\n")
    assert not ok and reason == "too_few_sentences (found 1, required 5)"
    assert content == "Keep this line please."
    assert md["line-filter-no_terminal_punc"] == "1" and md["line-filter-too_few_words"] == "1"
    assert "line-filter-too_long_word" not in md


# ---------------------------------------------------------------------------------------------
# GopherQualityFilter, gopher_quality.rs:339-829
# ---------------------------------------------------------------------------------------------
def gq(**kw):
    return GopherQualityParams(**kw)


def test_gq_permissive():  # :343-356
    assert run1(gq(), "This is a perfectly normal document with the and of words.")[0]


def test_gq_min_doc_words():  # :360-389
    p = gq(min_doc_words=3)
    assert run1(p, "Hello world test . !")[0]
    ok, reason, *_ = run1(p, "Hello world . !")
    assert not ok and "gopher_short_doc (2 non-symbol words, required 3)" in reason
    ok, reason, *_ = run1(p, ". ! ?")
    assert not ok and "gopher_short_doc (0 non-symbol words, required 3)" in reason


def test_gq_max_doc_words():  # :392-410
    p = gq(max_doc_words=3)
    assert run1(p, "One two three .")[0]
    ok, reason, *_ = run1(p, "One two three four .")
    assert not ok and "gopher_long_doc (4 non-symbol words, max 3)" in reason


def test_gq_avg_word_length():  # :414-472
    p = gq(min_avg_word_length=3.0, max_avg_word_length=5.0)
    assert run1(p, "cat words test .")[0]
    ok, reason, *_ = run1(p, "a it .")
    assert not ok and "gopher_below_avg_threshold (avg len 1.50, required 3.00)" in reason
    ok, reason, *_ = run1(p, "testing another .")
    assert not ok and "gopher_above_avg_threshold (avg len 7.00, max 5.00)" in reason
    ok, reason, *_ = run1(p, ". ! .")
    assert not ok and "gopher_below_avg_threshold (avg len 0.00, required 3.00 - 0 non-symbol words)" in reason


def test_gq_hashes():  # :476-533
    p = gq(max_symbol_word_ratio=0.1)
    assert run1(p, "word1 word2 # word3 word4 word5 word6 word7 word8 word9 word10")[0]
    ok, reason, *_ = run1(p, "word1 # word2 # word3 word4 word5 word6 word7 word8")
    assert not ok and "gopher_too_many_hashes (ratio 0.25, max 0.10)" in reason
    assert run1(p, "")[0]
    ok, reason, *_ = run1(p, "#")
    assert not ok and "gopher_too_many_hashes (ratio 1.00, max 0.10)" in reason


def test_gq_ellipsis_units():  # :536-568
    p = gq(max_symbol_word_ratio=0.1)
    assert run1(p, "word1 word2 ... word3 word4 word5 word6 word7 word8 word9 word10")[0]
    ok, reason, *_ = run1(p, "word1 ... word2 … word3 word4 word5 word6 word7 word8")
    assert not ok and "gopher_too_many_ellipsis_units (ratio 0.25, max 0.10)" in reason


def test_gq_bullets():  # :572-615
    p = gq(max_bullet_lines_ratio=0.5)
    assert run1(p, "- item 1\n- item 2\nnormal line\nanother normal line")[0]
    ok, reason, *_ = run1(p, "- item 1\n- item 2\n- item 3\nnormal line")
    assert not ok and "gopher_too_many_bullets (ratio 0.75, max 0.50)" in reason
    assert run1(p, "")[0]
    ok, reason, *_ = run1(p, "- all bullets")
    assert not ok and "gopher_too_many_bullets (ratio 1.00, max 0.50)" in reason


def test_gq_ellipsis_lines():  # :618-644
    p = gq(max_ellipsis_lines_ratio=0.5)
    assert run1(p, "Line one...\nLine two…\nNormal line\nAnother normal")[0]
    ok, reason, *_ = run1(p, "Line one...\nLine two…\nLine three...\nNormal line")
    assert not ok and "gopher_too_many_end_ellipsis_lines (ratio 0.75, max 0.50)" in reason


def test_gq_alpha_ratio():  # :648-756
    p = gq(max_non_alpha_words_ratio=0.5)
    assert run1(p, "word 123 word !!!")[0]
    ok, reason, *_ = run1(p, "word 123 456 !!!")
    assert not ok and "gopher_below_alpha_threshold (alpha ratio 0.33, required min 0.50)" in reason
    ok, reason, *_ = run1(p, "123 456 789 !!!")
    assert not ok and "gopher_below_alpha_threshold (alpha ratio 0.00, required min 0.50)" in reason
    ok, reason, *_ = run1(p, "")
    assert not ok and "gopher_below_alpha_threshold (alpha ratio 0.00, required min 0.50)" in reason


def test_gq_stop_words():  # :761-829
    p = gq(min_stop_words=2)
    assert run1(p, "the quick brown fox and the lazy dog")[0]
    ok, reason, *_ = run1(p, "a quick brown fox is lazy")
    assert not ok and "gopher_too_few_stop_words (found 0, required 2)" in reason
    p = gq(min_stop_words=1, stop_words=["custom", "words"])
    assert run1(p, "this is a custom test with other words")[0]
    ok, reason, *_ = run1(p, "this is a regular sentence")
    assert not ok and "gopher_too_few_stop_words (found 0, required 1)" in reason
    assert run1(gq(min_stop_words=0), "no stop words here")[0]
    assert run1(gq(), "no stop words here")[0]


# ---------------------------------------------------------------------------------------------
# GopherRepetitionFilter, gopher_rep.rs:412-642
# ---------------------------------------------------------------------------------------------
def gr(**kw):
    return GopherRepetitionParams(**kw)


def test_gr_permissive():  # :412-419
    assert run1(gr(), "This is a normal document.\nIt has multiple lines.\n\nAnd multiple paragraphs.")[0]


def test_gr_duplicate_paragraphs():  # :423-475
    p1, p2 = "This is the first paragraph.", "This is the second paragraph."
    c_pass = f"{p1}\n\n{p2}\n\nAnother unique."
    c_fail = f"{p1}\n\n{p2}\n\n{p1}"
    c_fail_char = f"{p1}\n\n{p1}\n\n{p1}"
    f = gr(dup_para_frac=0.3)
    assert run1(f, c_pass)[0]
    ok, reason, *_ = run1(f, c_fail)
    assert not ok and "dup_para_frac (ratio 0.33, max 0.30)" in reason
    ratio = (2 * len(p1)) / len(c_fail_char)
    f = gr(dup_para_char_frac=ratio - 0.01)
    assert run1(f, c_pass)[0]
    ok, reason, *_ = run1(f, c_fail_char)
    assert not ok and "dup_para_char_frac" in reason


def test_gr_duplicate_lines():  # :479-563
    l1, l2 = "This is line one.", "This is line two."
    c_pass = f"{l1}\n{l2}\nUnique line"
    c_fail = f"{l1}\n{l2}\n{l1}"
    c_fail_char = f"{l1}\n{l1}\n{l1}"
    f = gr(dup_line_frac=0.3)
    assert run1(f, c_pass)[0]
    ok, reason, *_ = run1(f, c_fail)
    assert not ok and "dup_line_frac (ratio 0.33, max 0.30)" in reason
    ratio = (2 * len(l1)) / max(len(c_fail_char), 1)
    thr = ratio - 0.01
    f = gr(dup_line_char_frac=thr)
    assert run1(f, c_pass)[0]
    ok, reason, *_ = run1(f, c_fail_char)
    assert not ok and reason.startswith("dup_line_char_frac (ratio") and ("max %.2f" % thr) in reason


def test_gr_top_n_grams():  # :567-603
    f = gr(top_n_grams=[(2, 0.3)])
    ok, _, _, _, r = run1(f, "a b c d e f a b g h i j")
    assert ok and r.stats[0]["top_bytes"][0][0] == 6 and r.stats[0]["char_len"][0] == 23
    ok, reason, _, _, r = run1(f, "a b c a b d a b e a b")
    assert not ok and "top_2_gram" in reason
    assert r.stats[0]["top_bytes"][0][0] == 12 and r.stats[0]["char_len"][0] == 21


def test_gr_duplicate_n_grams():  # :606-642
    f = gr(dup_n_grams=[(2, 0.1)])
    ok, reason, _, _, r = run1(f, "a b c d e a b f g")
    assert not ok and "duplicated_2_n_grams" in reason
    assert r.stats[0]["dup_bytes"][0][0] == 2 and r.stats[0]["char_len"][0] == 17
    assert run1(f, "a b c d e f g h i")[0]


def test_gr_empty():  # gopher_rep.rs:60-74
    ok, reason, _, md, _ = run1(gr(), "  \n ")
    assert not ok and reason == "skipping empty content"
    assert md == {"gopher_repetition_filter_status": "filtered", "gopher_repetition_filter_reason": "skipping empty content"}


# ---------------------------------------------------------------------------------------------
# FineWebQualityFilter: src/pipeline/filters/fineweb_quality.rs:230-604
# ---------------------------------------------------------------------------------------------
def fw(**kw):  # default_filter(), fineweb_quality.rs:240-250
    d = dict(line_punct_thr=0.12, line_punct_exclude_zero=False, short_line_thr=0.67, short_line_length=30,
             char_duplicates_ratio=0.95, new_line_ratio=0.3, stop_chars=None)
    d.update(kw)
    return FineWebQualityParams(**d)


def test_fw_empty_and_whitespace():  # :263-285, :475-488
    for t in ("", "   \n\t   \n ", "\n\n\n"):
        ok, reason, _, md, _ = run1(fw(), t)
        assert not ok and reason == "empty"
        assert md == {"fineweb_filter_status": "filtered", "fineweb_filter_reason": "empty document"}


def test_fw_line_punct_ratio():  # :288-344
    t = "Line one\nLine two\nLine three\nLine four\nLine five\nLine six\nLine seven\nLine eight\nLine nine\nLine ten."
    ok, reason, *_ = run1(fw(), t)
    assert not ok and reason.startswith("line_punct_ratio: 0.1000 < threshold 0.1200") and reason.endswith("(exclude_zero: false)")
    t = ("Line one is long enough and ends with a period.\nLine two is also long enough and ends with a question mark?\n"
         "Line three is also very long indeed and ends with an exclamation mark!")
    assert run1(fw(short_line_thr=1.0), t)[0]
    t = ("Looooooooong line one, no punctuation here\nLooooooooong line two, also no punctuation\n"
         "Looooooooong line three, definitely no punctuation")
    assert run1(fw(line_punct_exclude_zero=True, short_line_thr=1.0), t)[0]
    ok, reason, *_ = run1(fw(), "Line one\nLine two\nLine three")
    assert not ok and reason.startswith("line_punct_ratio: 0.0000 < threshold 0.1200")


def test_fw_short_line_ratio():  # :347-369
    t = ("Short line.\nThis is another short one.\nWay too short.\n"
         "This line is definitely longer than thirty characters to provide some balance.")
    ok, reason, *_ = run1(fw(), t)
    assert not ok and reason.startswith("short_line_ratio: 0.7500 > threshold 0.6700")
    t = ("This line is adequately long and should pass.\nSo is this one, it meets the criteria perfectly.\n"
         "And another one just to be sure it's fine.")
    assert run1(fw(), t)[0]


def test_fw_char_dup_ratio():  # :374-425
    assert run1(fw(line_punct_thr=0.0, short_line_thr=1.0, new_line_ratio=1.0), "abcdefghijklmnopqrstuvwxyz.\n1234567890.")[0]
    assert run1(fw(line_punct_thr=0.0, short_line_thr=1.0), "abcde fghij klmno pqrst uvwxyz.")[0]
    ok, reason, _, _, r = run1(fw(line_punct_thr=0.0, short_line_thr=1.0, new_line_ratio=1.0, char_duplicates_ratio=0.66),
                               "Hello World\nHello World\nHello World")
    assert not ok and reason.startswith("char_dup_ratio: 0.6667 > threshold 0.6600")
    assert r.stats[0]["dup_line_bytes"][0] == 22 and r.stats[0]["n_chars_no_newline"][0] == 33


def test_fw_new_line_ratio():  # :430-472, :491-508
    ok, reason, *_ = run1(fw(line_punct_thr=0.0, short_line_thr=1.0), "word.\nword.\nword.\nword.\nword.")
    assert not ok and reason.startswith("list_ratio: 0.8000 > threshold 0.3000")
    assert run1(fw(), "Many words on a single line with no newlines effectively. This should pass easily.")[0]
    t = ("Word one is long enough and ends with a period.\nWord two is also quite long and ends with a period.\n"
         "Word three is suitably lengthy and ends with a period.\nWord four and five and six are here and it ends with a period.")
    assert run1(fw(), t)[0]
    ok, reason, *_ = run1(fw(), "... --- !!!")
    assert not ok and reason.startswith("short_line_ratio: 1.0000 > threshold 0.6700")


def test_fw_passing_document():  # :511-527 (no metadata is written on pass, :225)
    t = ("This is a good line that ends with a period.\nAnother good line also ends with a question mark?\n"
         "Short lines are not too frequent here, which is great!\n"
         "Character duplication is hopefully not too high in this example text.\n"
         "And the ratio of newlines to words should be reasonable as well.")
    ok, reason, _, md, _ = run1(fw(), t)
    assert ok and reason == "" and md == {}


def test_fw_no_words_but_newlines():  # fineweb_quality.rs:191-204 (no reference test reaches it)
    ok, reason, *_ = run1(fw(line_punct_thr=0.0, short_line_thr=1.0), "!!!\n???")
    assert not ok and reason == "list_ratio_no_words (newlines present but no words)"


# ---------------------------------------------------------------------------------------------
# executor short-circuit, src/executor.rs:30-57 (tests/executor_test.rs:207-215 semantics)
# ---------------------------------------------------------------------------------------------
def test_chain_short_circuit():
    steps = [c4_default(), gq(min_doc_words=20), gr(dup_line_frac=0.3)]
    pl = oracle.OraclePipeline(steps)
    r = pl.run_texts(["Some code block {}.", "One. Two. Three. Four. Five sentences are here now.",
                      "This is the first sentence. This is the second sentence. This is the third sentence. "
                      "This is the fourth sentence. This is the fifth sentence."])
    assert list(r.failed_step) == [0, 1, -1]
    assert r.metadata(1)["c4_filter_status"] == "passed" and "gopher_repetition_filter_status" not in r.metadata(1)
    assert "gopher_quality_filter_status" not in r.metadata(0)  # later steps never ran
    md = r.metadata(2)
    assert md["c4_filter_status"] == md["gopher_quality_filter_status"] == md["gopher_repetition_filter_status"] == "passed"


# ---------------------------------------------------------------------------------------------
# html_escape::decode_html_entities (parquet_reader.rs:178-179), SURVEY 8(f) row 1.  The crate is not in the
# reference tree and the reference has no test with an entity: these pin the restatement to the published
# meaning of the references (WHATWG named list, decimal, hexadecimal), not to the crate -- parity unpinned.
# ---------------------------------------------------------------------------------------------
def test_decode_html_entities_published_forms():
    d = oracle.decode_html_entities
    assert d("Tom &amp; Jerry &lt;3 &gt; &quot;x&quot; &apos;y&apos;") == "Tom & Jerry <3 > \"x\" 'y'"
    assert d("&#72;i&#x21; &#X41; &eacute;&Eacute; &nbsp;|&euro;") == "Hi! A \u00e9\u00c9 \u00a0|\u20ac"
    assert d("&nLt;&#128512;&CounterClockwiseContourIntegral;") == "\u226a\u20d2\U0001f600\u2233"
    # not references: no semicolon, unknown or wrongly cased name, no digits, not a scalar value, does not fit u32
    for t in ("&amp", "&bogus;", "&AMp;", "&#;", "&#x;", "&#xD800;", "&#1114112;", "&#99999999999;", "&;", "& amp;", "AT&T"):
        assert d(t) == t
    assert d("AT&T &amp; co") == "AT&T & co" and d("&&amp;;") == "&&;" and d("&amp;amp;") == "&amp;"


def test_decode_html_entities_agrees_with_python_on_well_formed_input():
    import html
    import html.entities
    names = sorted(k for k in html.entities.html5 if k.endswith(";"))
    text = " ".join("&%s x" % n for n in names) + " " + " ".join("&#%d; &#x%X;" % (c, c) for c in (9, 10, 38, 65, 233, 0x20AC, 0x1F600, 0x10000))
    assert oracle.decode_html_entities(text) == html.unescape(text)
