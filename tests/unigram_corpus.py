"""Test sentences of tests/test_ingest_unigram.py (invisible characters are written as escapes)."""

CORPUS = [
    "The quick brown fox jumps over the lazy dog. Sentence embeddings are computed in batches of thirty-two!",
    "Is tokenisation really the bottleneck? Naïve café façade déjà vu; Ångström, Zürich, Škoda, İstanbul, ǅ, straße.",
    "Ελληνικά κείμενα: Αθήνα. Русский текст: Москва, ёлка.",
    "中文分词测试。日本語のテキスト、カタカナとひらがな。한국어 문장도. 豈 更 車",
    "ｆｕｌｌ\u3000ｗｉｄｔｈ ＡＢＣ１２３ ﬁ ﬂ ½ ¾ ™ Ⅷ ㍿ ｶﾞ ｷﾞ ㈱ ① ② ⑳ ㌔ ℃ Å K",
    "ש\u05b8\u05c1לו\u05b9ם עו\u05b9ל\u05b8ם م\u064eر\u0652ح\u064eب\u064bا ب\u0650ال\u0652ع\u064eال\u064eم नमस\u094dत\u0947 द\u0941निया สว\u0e31สด\u0e35 ক\u09cdষ ত\u09cdর",
    "e\u0301 a\u0308\u0304 o\u0302\u0301 n\u0303 A\u030a 각 ക\u0d4dഷ \u0600١ क\u094dष กำ",
    "Emoji \U0001f600 \U0001f680 \U0001f1e9\U0001f1ea \U0001f468\u200d\U0001f469\u200d\U0001f467 ©\u200d® ❤\ufe0f \U0001f44d\U0001f3fd and symbols © ® € £ ¥ § ¶ † ‡ • … — –",
    "tabs\tand\nnewlines\r\nand nbsp\u00a0thin\u2009hair\u200aideographic\u3000zero\u200bwidth\u2060joiner\u200dline\u2028sep\u2029para\u0085nel end",
    "x▁y ▁z  a▁ ▁▁b ▁", "<s> hi </s> <mask> there<unk><pad>", "", "   ", "a" * 700, "word " * 300,
    "\u0001\u0002 nul \u007f ctl � repl \ufeff bom \u00ad shy \u034f cgj \U000e0001 tag \U000e0100 vs",
]

FUZZ_EXTRAS = [" ", "  ", "\u200d", "\u0301", "\u3000", "\n"]
