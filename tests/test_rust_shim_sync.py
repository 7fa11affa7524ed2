"""The Rust side of the boundary (rust/textblaster-b200-sys, rust/textblaster-b200) cannot be compiled here (no Rust
toolchain in the image).  What can be checked without one: that the -sys crate declares every entry point of the
header, carries the same constants, and asserts the struct sizes the C compiler and the ctypes mirror agree on."""
import ctypes as C
import os
import re

from textblaster_b200 import _abi

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
HEADER = open(os.path.join(ROOT, "include", "textblaster_b200.h")).read()
SYS = open(os.path.join(ROOT, "rust", "textblaster-b200-sys", "src", "lib.rs")).read()
SHIM = open(os.path.join(ROOT, "rust", "textblaster-b200", "src", "lib.rs")).read()


def test_every_entry_point_is_declared():
    exported = set(re.findall(r"^(?:int|void|const char\*)\s+(tb_\w+)\s*\(", HEADER, re.M))
    declared = set(re.findall(r"pub fn (tb_\w+)\s*\(", SYS))
    assert exported and exported == declared, (sorted(exported - declared), sorted(declared - exported))


def test_struct_sizes_match_the_c_layout():
    sizes = dict(re.findall(r"size_of::<(tb_\w+)>\(\) == (\d+)", SYS))
    assert len(sizes) >= 11
    for name, size in sizes.items():
        assert C.sizeof(getattr(_abi, name)) == int(size), name
    # field order of the result struct: the names appear in the header's order
    fields = re.findall(r"pub (\w+):", SYS[SYS.index("pub struct tb_batch_out"):SYS.index("impl Default for tb_batch_out")])
    assert fields == [f[0] for f in _abi.tb_batch_out._fields_]


def test_constants_match_the_header():
    for name, value in re.findall(r"pub const (TB_\w+): \w+ = ([^;]+);", SYS):
        if not hasattr(_abi, name):
            continue
        v = value.replace("TB_MAX_NGRAMS as u32", str(_abi.TB_MAX_NGRAMS)).replace("TB_MAX_NGRAMS", str(_abi.TB_MAX_NGRAMS))
        assert eval(v) == getattr(_abi, name), name
    assert "TB_ABI_VERSION: c_int = %d" % _abi.TB_ABI_VERSION in SYS


def test_step_names_and_reason_formats_are_the_reference_strings():
    for s in ('"C4QualityFilter"', '"GopherQualityFilter"', '"GopherRepetitionFilter"', '"FineWebQualityFilter"',
              "too_few_sentences (found {}, required {})", "gopher_short_doc ({} non-symbol words, required {})",
              "gopher_below_alpha_threshold (alpha ratio {:.2}, required min {:.2})", "duplicated_{}_n_grams (ratio {:.2}, max {:.2})",
              "top_{}_gram (ratio {:.2}, max {:.2})", "gopher_repetition_filter_reason", "line-filter-no_terminal_punc"):
        assert s in SHIM, s
