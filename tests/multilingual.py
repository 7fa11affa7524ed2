"""BASELINE.json config 5 in miniature: non-ASCII-heavy documents (Cyrillic, Greek with final sigma,
Turkish dotted/dotless i, German sharp s, Danish with the shipped stop words, Hangul, Arabic with
Arabic-Indic citations, emoji / ZWJ sequences / flags, mixed White_Space, combining marks) plus a
separately reported Han / Kana / Thai subset (declared divergence class vs ICU4X)."""
import random

LANGS = {
    "russian": "и в не на я быть он с что а по это она этот к но они мы как из у который то за свой весь год от так о для ты же все тот мочь вы человек такой его сказать только или ещё бы себя один как уже до время если сам когда другой вот говорить наш мой знать стать при чтобы дело жизнь кто первый очень два день её новый рука даже во со раз где там под можно ну какой после их работа без самый потом надо хотеть ли слово идти большой должен место иметь ничто".split(),
    "greek": "ΚΑΙ ΤΟ ΣΟΦΟΣ ΟΔΟΣ λόγος ΛΟΓΟΣ ἄνθρωπος ΑΝΘΡΩΠΟΣ της και το να είναι σε που με για από τον την ένα τους δεν θα στο ΣΑΣ ΕΙΣ ΠΟΛΙΣ πόλις ΟΔΥΣΣΕΥΣ Σ ΑΣ".split(),
    "turkish": "İstanbul ISPARTA ılık Iğdır için bir ve bu da de ile çok daha gibi kadar İYİ iyi ŞİİR şiir DİYARBAKIR ağaç gün".split(),
    "german": "der die das und ist nicht STRASSE Straße groß GROSS weiß Fuß über müssen können für auf mit sich des ein eine auch als an nach wie im".split(),
    "danish": "og i at det er en til på de med af for ikke der var han jeg så men et har om vi havde ham hun nu over da fra du ud sin dem os op man hans hvor eller hvad skal selv her alle vil blev kunne ind når være dog noget ville jo deres efter ned skulle denne end dette mit også under have dig anden hende mine alt meget sit sine vor mod disse hvis din nogle hos blive mange ad bliver hendes været thi jer sådan ÅR Øl Æble".split(),
    "hangul": "그리고 나는 한국어 문장 입니다 사람 시간 오늘 내일 학교 에서 를 이 가 은 는 하다 있다 없다 좋다 서울 대한민국".split(),
    "arabic": "في من على أن إلى هذا التي الذي كان عن مع هو ما لا بين كل بعد قد ذلك عند لم أو حتى إذا كما له فيه".split(),
}
EMOJI = ["\U0001f600", "\U0001f468‍\U0001f469‍\U0001f467", "\U0001f1e9\U0001f1f0", "\U0001f1fa\U0001f1f8\U0001f1e9", "❤️",
         "\U0001f44d\U0001f3fd", "☺", "©", "™", "→", "•", "\U0001f3f3️‍\U0001f308"]
SPACES = [" ", " ", " ", "\u00a0", "\u2009", "\u202f", "\u2007", "\u3000", "\u1680", "\t", "\u2003", "\u205f", "\u200a"]
BREAKS = ["\n", "\n", "\n\n", "\r\n", "\r", "\u2028", "\u2029", "\u0085", "\x0b", "\x0c", "\n\n\n"]
COMBINING = ["́", "̈", "̧", "⃝", "゙"]
CJK = {
    "han": "我 爱 北京 天安门 中华人民共和国 今天 天气 很 好 我们 去 公园 散步 这 是 一个 测试 文档".split(),
    "kana": "これ は テスト です ひらがな カタカナ コンピュータ わたし の なまえ ありがとう ございます データ".split(),
    "thai": "ภาษาไทย สวัสดี ครับ ประเทศไทย วันนี้ อากาศ ดี มาก เรา ไป เดิน เล่น ที่ สวน".split(),
}
PUNCT = [".", ".", ".", "!", "?", "…", "...", "。", "؟", "·", ";", ":", ",", "'", "\"", "”", "»"]


def _sentence(rng, words, nmin=3, nmax=18, join=None):
    n = rng.randint(nmin, nmax)
    parts = []
    for i in range(n):
        w = rng.choice(words)
        r = rng.random()
        if r < 0.04:
            w += rng.choice(COMBINING) * rng.randint(1, 3)
        elif r < 0.07:
            w = rng.choice(EMOJI)
        elif r < 0.10:
            w = str(rng.randint(0, 99999)) if rng.random() < 0.5 else "".join(rng.choice("٠١٢٣٤٥٦٧٨٩") for _ in range(rng.randint(1, 4)))
        elif r < 0.12:
            w = w + rng.choice(["'", "’", ".", ":", "_", "-"]) + rng.choice(words)
        elif r < 0.13:
            w = "[" + rng.choice(["1", "٣", "12, ٤", "7,8"]) + "]"
        parts.append(w)
    sep = join if join is not None else None
    out = ""
    for i, w in enumerate(parts):
        out += w
        if i + 1 < len(parts):
            out += sep if sep is not None else (rng.choice(SPACES) if rng.random() < 0.25 else " ")
    return out + rng.choice(PUNCT)


def doc(rng, lang, target=1500):
    words = (LANGS.get(lang) or CJK[lang])
    join = "" if lang in ("han", "kana", "thai") and rng.random() < 0.7 else None
    out = ""
    while len(out.encode("utf-8")) < target:
        line = " ".join(_sentence(rng, words, join=join) for _ in range(rng.randint(1, 4)))
        if rng.random() < 0.1:
            line = rng.choice(["- ", "• ", "  ", "\xa0"]) + line
        out += line + rng.choice(BREAKS)
        if rng.random() < 0.1:
            out += line + "\n"  # duplicate line
    return out


def corpus(seed, n, langs=None, target=1500):
    rng = random.Random(seed)
    langs = list(langs or LANGS)
    return [doc(rng, langs[i % len(langs)], rng.choice([target // 4, target, target * 3])) for i in range(n)]
