"""Native WordPiece tokenizer (lsc_tokenizer_*, host code, no GPU needed) against the HuggingFace `tokenizers`
BERT pipeline - the library DJL's HuggingFaceTokenizer wraps in the reference: committed golden ids
(tests/golden/tokenizer_golden.json, made by make_tokenizer_golden.py) and, where transformers is importable,
a live comparison on a larger corpus and a random-Unicode fuzz.  Bar: identical ids."""
import json
import os
import random
import sys
from pathlib import Path

import numpy as np
import pytest

sys.path.insert(0, str(Path(__file__).parent))
import tokenizer_corpus as tc  # noqa: E402

GOLDEN = Path(__file__).parent / "golden"


@pytest.fixture(scope="module")
def vocab_file(tmp_path_factory):
    p = tmp_path_factory.mktemp("vocab") / "vocab.txt"
    p.write_text(tc.vocab_text(), encoding="utf-8")
    return str(p)


@pytest.mark.parametrize("name,lower", [("uncased", True), ("cased", False)])
def test_golden_ids(native_lib, vocab_file, name, lower):
    from langstream_b200._native import WordPieceTokenizer
    fx = json.loads((GOLDEN / "tokenizer_golden.json").read_text())
    texts = tc.texts(n_random=fx["n_random"], seed=fx["seed"])
    tok = WordPieceTokenizer.from_file(vocab_file, do_lower_case=lower, max_length=fx["max_length"], n_threads=3)
    got = tok(texts)
    assert len(got) == len(fx[name])
    bad = [i for i, (g, e) in enumerate(zip(got, fx[name])) if g != e]
    assert not bad, (ascii(texts[bad[0]]), got[bad[0]], fx[name][bad[0]])
    # the known answers a reader can check by eye (ids = positions in tokenizer_corpus.VOCAB)
    v = {t: i for i, t in enumerate(tc.VOCAB)}
    if lower:
        i = texts.index("Hello, World!")
        assert got[i] == [v["[CLS]"], v["hello"], v[","], v["world"], v["!"], v["[SEP]"]]
        i = texts.index("unable unablely reranking")
        assert got[i] == [v["[CLS]"], v["un"], v["##able"], v["un"], v["##able"], v["##ly"], v["[UNK]"], v["[SEP]"]]
        # ("reranking": "re" matches, "##ranking" / "##rank..." do not -> the whole word is [UNK])
        assert got[texts.index("")] == [v["[CLS]"], v["[SEP]"]]
        assert got[texts.index("x" * 101)] == [v["[CLS]"], v["[UNK]"], v["[SEP]"]]   # more than 100 characters


def test_packed_layout_truncation_and_threads(native_lib, vocab_file):
    from langstream_b200._native import WordPieceTokenizer
    texts = tc.texts(n_random=500, seed=3)
    one = WordPieceTokenizer.from_file(vocab_file, max_length=16, n_threads=1)
    many = WordPieceTokenizer.from_file(vocab_file, max_length=16, n_threads=8)
    ids1, cu1 = one.encode_packed(texts)
    ids8, cu8 = many.encode_packed(texts)
    assert np.array_equal(ids1, ids8) and np.array_equal(cu1, cu8)          # thread count never changes the result
    assert cu1.dtype == np.int32 and cu1[0] == 0 and len(cu1) == len(texts) + 1 and cu1[-1] == ids1.size
    lens = np.diff(cu1)
    assert lens.min() >= 2 and lens.max() == 16                             # [CLS] ... [SEP], truncated to max_length
    v = {t: i for i, t in enumerate(tc.VOCAB)}
    assert all(ids1[cu1[i]] == v["[CLS]"] and ids1[cu1[i + 1] - 1] == v["[SEP]"] for i in range(len(texts)))
    full = WordPieceTokenizer.from_file(vocab_file, max_length=512)(texts)
    assert all(a[1:-1] == b[1:-1][:14] for a, b in zip(one(texts), full))        # truncation keeps the first tokens
    assert one([]) == [] and one.encode_packed([])[1].tolist() == [0]


def test_errors(native_lib, vocab_file):
    from langstream_b200._native import NativeError, WordPieceTokenizer
    with pytest.raises(NativeError, match=r"\[UNK\]"):
        WordPieceTokenizer("hello\nworld\n")
    with pytest.raises(NativeError):
        WordPieceTokenizer(tc.vocab_text(), max_length=2)
    assert WordPieceTokenizer.from_file(vocab_file).vocab_size == len(tc.VOCAB)


def test_live_against_huggingface(native_lib, vocab_file):
    transformers = pytest.importorskip("transformers")
    from langstream_b200._native import WordPieceTokenizer
    texts = tc.texts(n_random=3000, seed=11)
    for lower in (True, False):
        hf = transformers.BertTokenizer(vocab_file, do_lower_case=lower)
        mine = WordPieceTokenizer.from_file(vocab_file, do_lower_case=lower, max_length=48)
        assert mine(texts) == hf(texts, truncation=True, max_length=48)["input_ids"]


def test_mpnet_template_against_huggingface(native_lib, tmp_path):
    """all-mpnet-base-v2's tokenizer: the BERT pipeline with the template <s> A </s> (MPNetTokenizer)."""
    transformers = pytest.importorskip("transformers")
    from langstream_b200._native import NativeError, WordPieceTokenizer
    bert = [t for t in tc.vocab_text().split("\n") if t]
    vp = tmp_path / "vocab.txt"
    vp.write_text("\n".join(["<s>", "<pad>", "</s>", "<unk>"] + bert + ["<mask>"]) + "\n", encoding="utf-8")
    texts = tc.texts(n_random=1500, seed=5)
    hf = transformers.MPNetTokenizer(str(vp), do_lower_case=True)
    mine = WordPieceTokenizer.from_file(str(vp), do_lower_case=True, max_length=64, cls_token="<s>", sep_token="</s>")
    assert mine(texts) == hf(texts, truncation=True, max_length=64)["input_ids"]
    with pytest.raises(NativeError, match="no token"):
        WordPieceTokenizer.from_file(str(vp), cls_token="<bos>")


def test_fuzz_ascii_bytes_against_huggingface(native_lib, vocab_file):
    """Pure-ASCII texts take a byte-level path of their own: every ASCII byte incl. the control characters (removed
    without splitting a word), tabs / line ends (blanks), punctuation runs, words of more than 100 characters."""
    transformers = pytest.importorskip("transformers")
    from langstream_b200._native import WordPieceTokenizer
    rng = random.Random(3)
    alphabet = [chr(c) for c in range(1, 128)] + list("  \t\n\r aeiou the of and THE Of") * 3
    texts = ["".join(rng.choice(alphabet) for _ in range(rng.randint(0, 60))) for _ in range(3000)]
    texts += ["a" * 101, "a" * 100, "x\x01y\x7fz", "\x0b\x0c", "ab\x00cd", "Hello, World!!", "the" + "\x02" * 5 + "of", ""]
    for lower in (True, False):
        hf = transformers.BertTokenizer(vocab_file, do_lower_case=lower)
        mine = WordPieceTokenizer.from_file(vocab_file, do_lower_case=lower, max_length=32)
        exp, got = hf(texts, truncation=True, max_length=32)["input_ids"], mine(texts)
        bad = [i for i in range(len(texts)) if exp[i] != got[i]]
        assert not bad, (repr(texts[bad[0]]), exp[bad[0]], got[bad[0]])


def test_fuzz_random_unicode_against_huggingface(native_lib, tmp_path):
    """Random code points from every plane; the vocabulary is built from the library's own normalised pre-tokens,
    so any difference in cleaning, accent stripping, lower-casing or splitting changes the ids."""
    transformers = pytest.importorskip("transformers")
    from langstream_b200._native import WordPieceTokenizer
    rng = random.Random(7)
    special = [0x300, 0x301, 0x327, 0x308, 0x323, 0x307, 0x3099, 0x1161, 0x11A8, 0x9, 0xA, 0x85, 0x2028, 0xAD, 0x600,
               0x61C, 0x180E, 0xFFF9, 0xAFF, 0x1AC3, 0xC3C]

    def rand_cp():
        r = rng.random()
        if r < 0.25:
            return rng.randint(0x20, 0x7E)
        if r < 0.45:
            return rng.randint(0x80, 0x24FF)
        if r < 0.55:
            return rng.choice([rng.randint(0x2500, 0xD7FF), rng.randint(0xE000, 0xFFFF)])
        if r < 0.60:
            return rng.randint(0x10000, 0x2FFFF)
        if r < 0.62:
            return rng.randint(0xE0000, 0x10FFFF)
        if r < 0.80:
            return 0x20
        return rng.choice(special + [0xAC00 + rng.randint(0, 11171)])

    texts = ["".join(chr(rand_cp()) for _ in range(rng.randint(1, 30))) for _ in range(2500)]
    bt = transformers.BertTokenizer(None, do_lower_case=True).backend_tokenizer
    words = set()
    for t in texts:
        for w, _ in bt.pre_tokenizer.pre_tokenize_str(bt.normalizer.normalize_str(t)):
            words.add(w)
            if len(w) > 2:
                words.update((w[:2], "##" + w[2:]))
    vp = tmp_path / "vocab.txt"
    vp.write_text("\n".join(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + sorted(words)) + "\n", encoding="utf-8")
    for lower in (True, False):
        hf = transformers.BertTokenizer(str(vp), do_lower_case=lower)
        mine = WordPieceTokenizer.from_file(str(vp), do_lower_case=lower, max_length=128)
        exp, got = hf(texts, truncation=True, max_length=128)["input_ids"], mine(texts)
        bad = [i for i in range(len(texts)) if exp[i] != got[i]]
        assert not bad, (lower, ascii(texts[bad[0]]), exp[bad[0]], got[bad[0]])
