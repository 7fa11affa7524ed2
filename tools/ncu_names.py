"""Kernel names as ncu prints them -> short names: 'void <unnamed>::k_gr_rep_c<(int)2>(KBatch, ...)' -> 'k_gr_rep_c<2>'."""
import re


def kname(full):
    m = re.search(r"(k_\w+)(<[^>]*>)?", full)
    if not m:
        return full.split("(")[0].split("::")[-1]
    return m.group(1) + (m.group(2) or "").replace("(int)", "").replace("(bool)", "").replace(" ", "")
