"""Stress corpus and toy vocabulary for the WordPiece tokenizer tests (shared by the golden-vector generator
tests/golden/make_tokenizer_golden.py and tests/test_tokenizer.py).  All literals are ASCII escapes."""
import random

BASE_WORDS = [
    'the', 'a', 'of', 'and', 'to', 'in', 'is', 'that', 'it', 'stream', 'lang', 'embed', 'vector', 'data',
    'base', 'query', 'un', 'able', 're', 'rank', 'token', 'model', 'cafe', 'naive', 'uber', 'strasse',
    'hello', 'world', 'i', '\u0307', '\u4e2d', '\u6587', '\u65e5', '\u672c', '\u1112', '\u1161', '\u11ab',
    '\u03c3', '\u03c2', '\xe9', '\xdf', 'ss', 'ffi', 'x', 'y', 'z', '1', '2', '3', '10', '000',
]
SUB_WORDS = [
    '##s', '##ing', '##ed', '##ly', '##able', '##er', '##e', '##n', '##t', '##a', '##o', '##1', '##2', '##x',
    '##ss', '##\u0307', '##\u1161', '##\u11ab', '##\u4e2d',
]
PUNCT = [
    '!', '"', '#', '$', '%', '&', "'", '(', ')', '*', '+', ',', '-', '.', '/', ':', ';', '<', '=', '>', '?',
    '@', '[', '\\', ']', '^', '_', '`', '{', '|', '}', '~', '\u2014', '\u2026', '\xab', '\xbb', '\u3002',
    '\uff0c', '\xbf', '\u3001',
]
VOCAB = list(dict.fromkeys(["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + BASE_WORDS + SUB_WORDS + PUNCT))

# characters that exercise every rule of the normaliser: case pairs with multi-character lowercase, precomposed
# and combining accents, ligatures, CJK, Hangul (algorithmic NFD), all kinds of white space, controls, format
# characters, the replacement character, private use, unassigned, symbols that are ASCII punctuation only, ...
CHARS = [
    'a', 'b', 'e', 's', 't', 'x', 'A', 'B', '\xc9', '\xe9', '\xfc', '\xdc', '\xdf', '\u0130', 'I', '\u0131',
    '\u03a3', '\u03c3', '\u03c2', '\u01c4', '\u01c5', '\ufb03', '\xc5', '\u212b', '\xf1', 'n\u0303', '\u0301',
    '\u0327\u0301', '\u0301\u0327', '\u4e2d', '\u6587', '\u65e5', '\u672c', '\ud55c', '\uae00', '\uac01',
    '\u1112', '\uac00', ' ', '  ', '\t', '\n', '\r', '\xa0', '\u2009', '\u3000', '\x85', '\u200b', '\u200d',
    '\ufeff', '\x00', '\x07', '\x1f', '\ufffd', '\ue000', '\u0378', '!', ',', '.', '-', '\u2014', '\u2026',
    '\xab', '$', '^', '`', '~', '|', '\u3002', '\uff11', '\uff21', '\U0001d49c', '\U0001f600', '\U0001f1fa',
    '\U000e0041', '\U00020000', '\u0e01\u0e34', '\u0915\u094d', 'a\u0300\u0316', '\u1e69', 's\u0323\u0307',
    's\u0307\u0323',
]
FIXED_TEXTS = [
    '', ' ', 'Hello, World!', 'unable unablely reranking', 'Caf\xe9 na\xefve \xdcBER Stra\xdfe',
    '\u0130stanbul I\u0131', '\u03a3\u038a\u03a3\u03a5\u03a6\u039f\u03a3 \u03c3\u03c2',
    '\u4e2d\u6587\u65e5\u672c mixed\u4e2dtext', '\ud55c\uae00 \uac01',
    'xxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxx',
    'xxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxxx',
    'tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens tokens ',
    '\ufb03 \u01c4 \u01c5', 'a\u0327\u0301 b\u0301\u0327', '\u200bzero\u200dwidth\ufeff',
    'tab\tnew\nline\r\n', '\uff11\uff12\uff13 \uff21\uff22\uff23', 'something to embed',
]


def texts(n_random: int = 3000, seed: int = 0):
    rng = random.Random(seed)
    out = list(FIXED_TEXTS)
    pool = CHARS + BASE_WORDS + [" "] * 6
    for _ in range(n_random):
        out.append("".join(rng.choice(pool) for _ in range(rng.randint(0, 40))))
    return out


def vocab_text() -> str:
    return "\n".join(VOCAB) + "\n"
