"""Adversarial document generator for the parity tests (seeded; pure Python).

Mixes ASCII prose with the characters the filters are sensitive to: all White_Space variants, the
PUNCTUATION members, UAX #29 Mid*/quote/Extend/ZWJ/Regional_Indicator classes, case-mapping
specials (U+0130, U+03A3 final sigma, U+212A), citations, policy phrases, long words, duplicate
lines / paragraphs / n-grams, and every newline flavour.
"""
import random

SPECIAL = [
    "\xe6", "\xc5", "Ж", "ж", "Σ", "σ", "ς", "א", "ב", "́", "‍", "\xad",
    "\U0001f600", "\U0001f1e9", "\U0001f1f0", "\xa0", " ", " ", "　", "١", "’", "“", "”",
    "…", "\xb7", "❤", "️", "\x85", "\x0b", "—", "״", "،", "٫", "．", "․",
    " ", " ", "\xdf", "İ", "K", "中", "文", "ア", "カ", "ก", "•", "１",
    "\xa7", "\r", "\t", "{", "#", "[", "]", ",", "\x00", "\x1f", "\x7f", " ", "\x0c",
]
WORDS = [
    "the", "be", "to", "of", "and", "that", "have", "with", "lorem", "ipsum", "javascript", "privacy", "policy", "cookie",
    "cookies", "uses", "use", "terms", "hello", "world", "a", "I", "x1", "3.14", "don't", "foo_bar", "U.S.A.", "e.g.",
    "Mr.", "10,000", "...", "…", "-", "•", "#", "[1]", "[2, 3]", "[4,5,6]", "[7,", "[a]", "[١٢]",
    "(see)", "\"quoted\"", "end.", "stop!", "why?", "ok”", "caf\xe9", "na\xefve", "Жук",
    "ΣΟΦΟΣ", "İstanbul", "Kelvin", "TERMS OF USE", "Use Of Cookies", "JavaScript",
    "LOREM IPSUM", "cooKie policy", "a:b", "foo:bar", "a:b:c", "12:30", "http://x", "key:value", "1.5.2", "it's", "rock'n'roll", "\U0001f1e9\U0001f1f0\U0001f1e9",
]


class FuzzGen:
    def __init__(self, seed):
        self.rng = random.Random(seed)

    def word(self):
        rng = self.rng
        r = rng.random()
        if r < 0.7:
            return rng.choice(WORDS)
        if r < 0.85:
            return "".join(rng.choice("abcdefghijklmnopqrstuvwxyzABCDEFGH0123456789") for _ in range(rng.randint(1, 12)))
        if r < 0.97:
            return "".join(rng.choice(SPECIAL + list("ab1.',:;")) for _ in range(rng.randint(1, 6)))
        return rng.choice("wZ9") * rng.randint(30, 1200)

    def line(self):
        rng = self.rng
        n = rng.randint(0, 30)
        s = "".join(self.word() + rng.choice([" ", " ", " ", "  ", "\t", ",", ", ", ". ", "\xa0"]) for _ in range(n))
        if rng.random() < 0.6:
            s = s.rstrip() + rng.choice([".", "!", "?", "...", "…", "\"", "'", "”", " .", "."])
        if rng.random() < 0.1:
            s = "  " + s + "  "
        return s

    def doc(self, max_lines=12):
        rng = self.rng
        lines = []
        for _ in range(rng.randint(0, max_lines)):
            lines.append(self.line())
            if rng.random() < 0.2:
                lines.append(rng.choice(lines))
        doc = ""
        for ln in lines:
            doc += ln + rng.choice(["\n", "\n", "\n\n", "\r\n", "\n\n\n", " \n"])
        if rng.random() < 0.5:
            doc = doc.rstrip("\n")
        if rng.random() < 0.1:
            doc = rng.choice(["\n", " ", "\n\n"]) + doc
        return doc

    def docs(self, n, max_lines=12):
        return [self.doc(max_lines) for _ in range(n)]


def english_doc(rng, target_bytes):
    """C4-shaped ASCII prose with repeated boilerplate, for multi-tile (> 1 KiB) coverage."""
    vocab = ["alpha", "beta", "gamma", "delta", "the", "of", "and", "to", "with", "that", "have", "be", "system",
             "data", "value", "people", "time", "year", "work", "government", "company", "number", "a", "in", "is"]
    out = []
    size = 0
    boiler = "Click here to read more about our products and services today."
    while size < target_bytes:
        nsent = rng.randint(1, 5)
        line = []
        for _ in range(nsent):
            nw = rng.randint(3, 20)
            ws = [rng.choice(vocab) for _ in range(nw)]
            ws[0] = ws[0].capitalize()
            line.append(" ".join(ws) + rng.choice([".", ".", ".", "?", "!", ""]))
        text = " ".join(line)
        if rng.random() < 0.15:
            text = boiler
        if rng.random() < 0.05:
            text += " [%d]" % rng.randint(1, 99)
        out.append(text)
        size += len(text) + 1
        if rng.random() < 0.3:
            out.append("")
    return "\n".join(out)


def ascii_doc(rng, n_tokens=400):
    """Pure-ASCII text dense in the characters the UAX #29 word rules care about (Mid*, quotes, '_',
    digits, runs of spaces, tabs, VT/FF) -- exercises the bit-parallel fast path of the word pass."""
    atoms = ["a", "b", "Z", "0", "7", ".", ",", ";", ":", "'", "\"", "_", "-", " ", " ", " ", "  ", "\t", "\n", "\n\n",
             "\x0b", "\x0c", "(", ")", "#", "...", "don't", "3.14", "1,000", "a.b.c", "x_y", "_x", "9_", "a:b", "a'", "'a", "1.",
             ".1", "a..b", "1,,2", "a'b'c", "U.S.A.", "the", "of", "and", "javascript", "privacy policy", "[1]", "[2, 3]", "{", "!", "?"]
    out = []
    for _ in range(n_tokens):
        out.append(rng.choice(atoms))
    return "".join(out)


def boundary_docs():
    """Documents whose interesting spots (white space at line ends, citations, multi-byte characters, words
    joined by MidLetter / MidNum, CR LF) slide across the 32-byte stretch and 1 KiB tile boundaries of the
    kernels, at several alignments of the document start."""
    ws = [" ", "\t", "\u00a0", "\u3000", "\u2003 ", " \u202f", "\r", " \r", "\u0085", "\u2028"]
    mids = ["don't", "3.14", "a_b", "x:y", "1,000", "café", "naïve", "über", "été", "\U0001f600ok"]
    docs = []
    for pad in list(range(0, 40)) + list(range(1000, 1040, 3)):
        lead = "w" * pad
        k = pad % len(ws)
        m = mids[pad % len(mids)]
        docs.append(lead + ws[k] + "\n" + ws[(k + 1) % len(ws)] + "Second line here [12, 3] ends." + ws[(k + 2) % len(ws)]
                    + "\n\n" + m + " " + m + ". Tail [7] words" + ws[(k + 3) % len(ws)] * 3 + "and more [1, 2] text.\n")
        docs.append(lead + " " + (m + " ") * 8 + "\r\n" + ws[k] * 2 + m + ".\r\n")
        docs.append(("ab " * (pad // 3)) + "x" + ws[k] * (1 + pad % 5) + "[1" + "," * (pad % 2) + " 2]" + ws[(k + 1) % len(ws)] + "y z. Done here now.")
    # newline runs longer than a stretch, kept lines on both sides
    docs.append("First kept line here.\n" + "\n" * 70 + "Second kept line here.\n" + "\n" * 33 + "x\n" * 40 + "Third kept line is here.")
    docs.append("\n" * 100 + "Only line, it ends well." + "\n" * 64)
    return docs


def dense_word_docs():
    """Documents with the most words a text can hold per byte: a one-byte letter or digit alternating with a
    two-byte symbol that is a word of its own (not in PUNCTUATION, Word_Break = Other) gives two words per
    three bytes.  Each is followed by an ordinary document so that a write past the first one's word slots shows
    up in its neighbour."""
    docs = []
    for sym, k in (("\u00a9", 4), ("\u00a9", 40), ("\u00d7", 300), ("\u00a7", 700), ("\u00b0", 3000), ("\u00a9", 11)):
        for lead in ("", "zz ", "\n"):
            docs.append(lead + ("a" + sym) * k + "a")
            docs.append("An ordinary line of text follows here. It has two sentences.\nAnd a second line that repeats a b c d e f a b c d e f.")
    docs.append("".join("x\u00a9"[i % 2] for i in range(5000)) + "\n" + "7\u00d7" * 100)
    docs.append("\u00a9" * 600)
    docs.append("a\u00a9 " * 400)
    return docs
