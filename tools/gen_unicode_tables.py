#!/usr/bin/env python3
"""Generate textblaster_b200/csrc/unicode_tables.inc (committed; re-run only to regenerate).

Sources (offline, this container): perl 5.38 Unicode::UCD (Unicode 15.0.0) for
Word_Break, Sentence_Break, White_Space, Alphabetic, Extended_Pictographic, Gc=Nd,
Cased, Case_Ignorable, Lowercase_Mapping, Script, Line_Break=SA; patched by hand to
Unicode 15.1 (the version ICU4X 1.5's compiled data uses): CJK
Extension I U+2EBF0..2EE5D (Lo, Alphabetic, Ideographic, Script=Han => WB=Other,
SB=OLetter).  U+2FFC..2FFF and U+31EF (new IDCs, So) carry none of these properties.

Per code point we pack one u16 ("cinfo"):
  bits 0..4  Word_Break class      (UAX #29 table 3; tb_wb_* constants)
  bits 5..8  Sentence_Break class  (UAX #29 table 4; tb_sb_* constants)
  bit  9     White_Space           (Rust char::is_whitespace, str::trim)
  bit 10     member of the reference's PUNCTUATION set (src/utils/text.rs:28-57)
  bit 11     Alphabetic            (Rust char::is_alphabetic)
  bit 12     Extended_Pictographic (WB3c)
  bit 13     General_Category=Nd   (regex crate \\d, c4_filters.rs:33)
  bit 14     "complex script": Han/Hiragana/Katakana or Line_Break=SA -- ICU4X new_auto()
             re-segments these with dictionary/LSTM data we do not have (declared
             divergence class)
  bit 15     has a lowercase mapping different from itself (str::to_lowercase)

Layout: 2-stage trie, stage1[cp >> 8] -> block index, stage2[block * 256 + (cp & 255)].

--colon {cldr-root,uax29}: Word_Break of U+003A COLON, U+FE55 SMALL COLON and U+FF1A FULLWIDTH COLON.
  uax29      MidLetter, as UAX #29's WordBreakProperty.txt lists them ("a:b" is one word).
  cldr-root  Other (the default here).  Since CLDR 42 / ICU 72 the ROOT word-break rules take these three out of MidLetter
             ("$MidLetter = [\p{Word_Break = MidLetter} - [\: \uFE55 \uFF1A]];" in ICU's brkitr/rules/word.txt) and only the
             Finnish and Swedish tailorings keep them; ICU4X carries exactly those two locale overrides for its word
             segmenter, which only makes sense if its root data excludes the colon.  The reference calls
             WordSegmenter::new_auto() without a locale (src/utils/text.rs:107), i.e. root: "foo:bar" is foo | : | bar, two
             words.  Neither ICU4X nor CLDR data is in this image, so this is the best-evidence setting, not a pinned one
             (DESIGN.md section 2); the ICU4C 71 cross-check (Unicode 14, before the change) cannot see it either way.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "textblaster_b200", "csrc", "unicode_tables.inc")

WB = ["Other", "CR", "LF", "Newline", "Extend", "ZWJ", "Regional_Indicator", "Format",
      "Katakana", "Hebrew_Letter", "ALetter", "Single_Quote", "Double_Quote", "MidNumLet",
      "MidLetter", "MidNum", "Numeric", "ExtendNumLet", "WSegSpace"]
SB = ["Other", "CR", "LF", "Sep", "Extend", "Format", "Sp", "Lower", "Upper", "OLetter",
      "Numeric", "ATerm", "SContinue", "STerm", "Close"]

# src/utils/text.rs:28-29 PUNCTUATION_LIT (66 code points) + :32-37 ranges [0,9) [11,13) [13,32) [127,160)
PUNCT_LIT = """0021 0022 0023 0024 0025 0026 0027 0028 0029 002A 002B 002C 002D 002E 002F 003A 003B
003C 003D 003E 003F 0040 005B 005C 005D 005E 005F 0060 007B 007C 007D 007E 00AB 00B4 00BB 2013 2014
2019 201C 201D 201E 2026 2236 2501 25BA 3001 3002 3008 3009 300A 300B 300C 300D 3010 3011 FF01 FF05
FF08 FF09 FF0C FF0E FF11 FF1A FF1B FF1F FF5E"""
PUNCT_RANGES = [(0, 9), (11, 13), (13, 32), (127, 160)]

PERL = r'''
use Unicode::UCD qw(prop_invlist prop_invmap);
print "VERSION ", Unicode::UCD::UnicodeVersion(), "\n";
for my $p ("Word_Break", "Sentence_Break") {
  my ($list, $map, $fmt, $def) = prop_invmap($p);
  for my $i (0..$#$list) {
    my $end = $i < $#$list ? $list->[$i+1] : 0x110000;
    print "MAP $p $list->[$i] $end $map->[$i]\n";
  }
}
for my $p ("White_Space", "Alphabetic", "Extended_Pictographic", "General_Category=Nd", "Cased",
           "Case_Ignorable", "Script=Han", "Script=Hiragana", "Script=Katakana", "Line_Break=SA") {
  my @l = prop_invlist($p);
  push @l, 0x110000 if @l % 2;
  for (my $i = 0; $i < @l; $i += 2) { print "LIST $p $l[$i] $l[$i+1]\n"; }
}
my ($list, $map, $fmt, $def) = prop_invmap("Lowercase_Mapping");
for my $i (0..$#$list) {
  my $end = $i < $#$list ? $list->[$i+1] : 0x110000;
  my $m = $map->[$i];
  next if !ref($m) && $m == 0;
  if (ref($m)) { print "LOWER $list->[$i] $end S ", join(" ", @$m), "\n"; }
  else { print "LOWER $list->[$i] $end A $m\n"; }   # format "al": adjusts by offset within range
}
'''


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--colon", default="cldr-root", choices=["cldr-root", "uax29"])
    args = ap.parse_args()
    out = subprocess.run(["perl", "-e", PERL], capture_output=True, text=True, check=True).stdout
    N = 0x110000
    wb = [0] * N
    sb = [0] * N
    flags = {k: bytearray(N) for k in ("ws", "alpha", "extpict", "nd", "cased", "ci", "complex")}
    lower = {}
    special_lower = {}
    version = None
    for line in out.splitlines():
        t = line.split()
        if t[0] == "VERSION":
            version = t[1]
        elif t[0] == "MAP":
            prop, a, b, v = t[1], int(t[2]), int(t[3]), t[4]
            if prop == "Word_Break":
                ep = False
                if v == "ExtPict_LE":
                    v, ep = "ALetter", True
                elif v == "ExtPict_XX":
                    v, ep = "Other", True
                if v == "Perl_Tailored_HSpace":
                    # perl tailors \b{wb}; restore the UAX #29 values
                    for cp in range(a, b):
                        if cp in (0x09, 0xA0, 0x2007):
                            wb[cp] = WB.index("Other")
                        else:
                            wb[cp] = WB.index("WSegSpace")
                    continue
                k = WB.index(v)
                for cp in range(a, b):
                    wb[cp] = k
            else:
                k = SB.index(v)
                for cp in range(a, b):
                    sb[cp] = k
        elif t[0] == "LIST":
            prop, a, b = t[1], int(t[2]), int(t[3])
            key = {"White_Space": "ws", "Alphabetic": "alpha", "Extended_Pictographic": "extpict",
                   "General_Category=Nd": "nd", "Cased": "cased", "Case_Ignorable": "ci"}.get(prop, "complex")
            f = flags[key]
            for cp in range(a, b):
                f[cp] = 1
        elif t[0] == "LOWER":
            a, b, kind = int(t[1]), int(t[2]), t[3]
            if kind == "A":
                base = int(t[4])
                for cp in range(a, b):
                    lower[cp] = base + (cp - a)
            else:
                seq = [int(x) for x in t[4:]]
                assert b == a + 1
                special_lower[a] = seq
    assert version == "15.0.0", version
    assert special_lower == {0x130: [0x69, 0x307]}, special_lower
    # ---- Unicode 15.1 patch: CJK Unified Ideographs Extension I
    for cp in range(0x2EBF0, 0x2EE5E):
        wb[cp] = WB.index("Other")
        sb[cp] = SB.index("OLetter")
        flags["alpha"][cp] = 1
        flags["complex"][cp] = 1
    assert all(wb[cp] == WB.index("MidLetter") for cp in (0x3A, 0xFE55, 0xFF1A))
    if args.colon == "cldr-root":
        for cp in (0x3A, 0xFE55, 0xFF1A):
            wb[cp] = WB.index("Other")
    # sanity anchors from SURVEY.md Appendix B
    assert sum(1 for c in range(N) if flags["ws"][c]) == 25
    assert sum(1 for c in range(N) if wb[c] == WB.index("WSegSpace")) == 14
    assert wb[0x202F] == WB.index("ExtendNumLet")
    assert sum(1 for c in range(N) if sb[c] == SB.index("ATerm")) == 4

    punct = set(int(x, 16) for x in PUNCT_LIT.split())
    assert len(punct) == 66
    for a, b in PUNCT_RANGES:
        punct.update(range(a, b))
    assert len(punct) == 129

    cinfo = [0] * N
    for cp in range(N):
        v = wb[cp] | (sb[cp] << 5)
        if flags["ws"][cp]:
            v |= 1 << 9
        if cp in punct:
            v |= 1 << 10
        if flags["alpha"][cp]:
            v |= 1 << 11
        if flags["extpict"][cp]:
            v |= 1 << 12
        if flags["nd"][cp]:
            v |= 1 << 13
        if flags["complex"][cp]:
            v |= 1 << 14
        if (cp in lower and lower[cp] != cp) or cp in special_lower:
            v |= 1 << 15
        cinfo[cp] = v
    # surrogates are not scalar values; leave as computed (never reached from valid UTF-8)

    blocks = {}
    stage1 = []
    stage2 = []
    for hi in range(N >> 8):
        blk = tuple(cinfo[hi << 8:(hi + 1) << 8])
        if blk not in blocks:
            blocks[blk] = len(blocks)
            stage2.extend(blk)
        stage1.append(blocks[blk])

    lower_pairs = sorted((cp, lo) for cp, lo in lower.items() if lo != cp)

    def ranges(f):
        r = []
        cp = 0
        while cp < N:
            if f[cp]:
                s = cp
                while cp < N and f[cp]:
                    cp += 1
                r.append((s, cp))
            else:
                cp += 1
        return r

    cased_r = ranges(flags["cased"])
    ci_r = ranges(flags["ci"])

    with open(OUT, "w") as fo:
        w = fo.write
        w("// GENERATED by tools/gen_unicode_tables.py -- do not edit.\n")
        w("// Unicode 15.0.0 (perl Unicode::UCD) + 15.1 CJK Ext-I patch. See the generator for the bit layout.\n")
        w("#define TB_UC_VERSION \"15.1 (15.0.0 + ExtI patch)\"\n")
        w("#define TB_UC_COLON_MIDLETTER %d  // 0: CLDR root tailoring (colon is Word_Break=Other), 1: plain UAX #29\n" % (0 if args.colon == "cldr-root" else 1))
        for i, n in enumerate(WB):
            w(f"#define TB_WB_{n.upper()} {i}\n")
        for i, n in enumerate(SB):
            w(f"#define TB_SB_{n.upper()} {i}\n")
        w(f"#define TB_UC_NBLOCKS {len(blocks)}\n")
        w(f"#define TB_UC_NLOWER {len(lower_pairs)}\n")
        w(f"#define TB_UC_NCASED {len(cased_r)}\n")
        w(f"#define TB_UC_NCASEIGN {len(ci_r)}\n")
        w(f"static const unsigned short tb_uc_stage1[{len(stage1)}] = {{\n")
        for i in range(0, len(stage1), 32):
            w(",".join(str(x) for x in stage1[i:i + 32]) + ",\n")
        w("};\n")
        w(f"static const unsigned short tb_uc_stage2[{len(stage2)}] = {{\n")
        for i in range(0, len(stage2), 32):
            w(",".join(str(x) for x in stage2[i:i + 32]) + ",\n")
        w("};\n")
        w(f"static const unsigned int tb_uc_lower[{len(lower_pairs)}][2] = {{\n")
        for i in range(0, len(lower_pairs), 8):
            w(",".join("{%d,%d}" % p for p in lower_pairs[i:i + 8]) + ",\n")
        w("};\n")
        for name, r in (("tb_uc_cased", cased_r), ("tb_uc_caseign", ci_r)):
            w(f"static const unsigned int {name}[{len(r)}][2] = {{\n")
            for i in range(0, len(r), 8):
                w(",".join("{%d,%d}" % p for p in r[i:i + 8]) + ",\n")
            w("};\n")
    print(f"wrote {OUT}: {len(blocks)} blocks, {len(lower_pairs)} lower pairs, "
          f"{len(cased_r)} cased ranges, {len(ci_r)} case-ignorable ranges", file=sys.stderr)


if __name__ == "__main__":
    main()
