"""Writes tests/golden/tokenizer_golden.json: token ids produced by the HuggingFace `tokenizers` BERT pipeline
(through transformers.BertTokenizer, the library DJL's HuggingFaceTokenizer wraps in the reference) for the
stress corpus of tests/tokenizer_corpus.py with its toy vocabulary.  Run here, where transformers is installed;
the fixture travels."""
import json
import os
import sys
import tempfile
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import tokenizer_corpus as tc  # noqa: E402
from transformers import BertTokenizer  # noqa: E402

MAX_LENGTH = 64


def main():
    d = tempfile.mkdtemp()
    vp = os.path.join(d, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write(tc.vocab_text())
    texts = tc.texts(n_random=400, seed=0)
    out = {"source": "transformers.BertTokenizer (tokenizers BertNormalizer + BertPreTokenizer + WordPiece), "
                     "truncation=True, max_length=%d" % MAX_LENGTH,
           "max_length": MAX_LENGTH, "n_random": 400, "seed": 0}
    for name, lower in (("uncased", True), ("cased", False)):
        tok = BertTokenizer(vp, do_lower_case=lower)
        out[name] = tok(texts, truncation=True, max_length=MAX_LENGTH)["input_ids"]
    Path(__file__).with_name("tokenizer_golden.json").write_text(json.dumps(out, separators=(",", ":")))
    print(len(texts), "texts")


if __name__ == "__main__":
    main()
