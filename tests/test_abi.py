"""The C-ABI shared library loads without a GPU, exports every entry point include/textblaster_b200.h
declares, the ctypes mirror has the C struct sizes, and the product refuses to run without CUDA."""
import ctypes as C
import os
import re
import subprocess

import pytest

from textblaster_b200 import _abi, _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "textblaster_b200.h")


def declared_functions():
    src = open(HEADER, encoding="utf-8").read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported():
    names = declared_functions()
    assert len(names) >= 13 and "tb_pipeline_run_batch" in names
    lib = _native.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_native.EXPORTS) == sorted(n for n in names if n in _native.EXPORTS)
    assert lib.tb_abi_version() == _abi.TB_ABI_VERSION


def test_struct_layout_matches_ctypes(tmp_path):
    structs = ["tb_c4_params", "tb_gq_params", "tb_gr_params", "tb_step_cfg", "tb_c4_stats", "tb_gq_stats", "tb_gr_stats",
               "tb_batch_out"]
    prog = "#include <stdio.h>\n#include \"textblaster_b200.h\"\nint main(void){" + "".join(
        f'printf("{s} %zu\\n", sizeof({s}));' for s in structs) + "return 0;}\n"
    c = tmp_path / "sz.c"
    c.write_text(prog)
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)], check=True)  # header is plain C
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    sizes = dict(l.split() for l in out.strip().splitlines())
    for s in structs:
        assert int(sizes[s]) == C.sizeof(getattr(_abi, s)), s


def test_step_names():
    lib = _native.lib()
    assert lib.tb_step_name(_abi.TB_STEP_C4_QUALITY) == b"C4QualityFilter"        # c4_filters.rs:143-145
    assert lib.tb_step_name(_abi.TB_STEP_GOPHER_QUALITY) == b"GopherQualityFilter"  # gopher_quality.rs:65-67
    assert lib.tb_step_name(_abi.TB_STEP_GOPHER_REPETITION) == b"GopherRepetitionFilter"  # gopher_rep.rs:48-50


def test_no_cpu_fallback():
    """Without a CUDA device the product fails loudly instead of computing on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from textblaster_b200.error import PipelineError
    from textblaster_b200.executor import GpuBatchPipeline
    from parity_cases import GQ_OPEN

    with pytest.raises(PipelineError) as e:
        GpuBatchPipeline([GQ_OPEN]).run_texts(["hello world"])
    assert "no CUDA device" in str(e.value)


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "textblaster_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), encoding="utf-8", errors="replace").read()
                assert "import oracle" not in src and "tb_oracle" not in src and "libtb_oracle" not in src, f
