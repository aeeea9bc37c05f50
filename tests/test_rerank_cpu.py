"""Host side of the re-rank mirror (no GPU): Java's split semantics and the term statistics handed to the kernel."""
import numpy as np


def test_java_split_whitespace():
    from langstream_b200.rerank import java_split_whitespace as js
    assert js("a b  c") == ["a", "b", "c"]
    assert js("a b ") == ["a", "b"]            # trailing empty strings are removed
    assert js(" a b") == ["", "a", "b"]        # a leading one is kept
    assert js("") == [""]                      # no match: the input itself
    assert js("   ") == []                     # only separators
    assert js("a\tb\nc") == ["a", "b", "c"]


def test_term_statistics_match_oracle_tokeniser():
    from langstream_b200.rerank import term_statistics
    from oracle import rerank_oracle

    texts = ["this is a text", " leading space text text", "other words ", "text"]
    query = "text is text missing"
    tf, doc_len = term_statistics(texts, query)
    assert tf.shape == (4, 4) and tf.dtype == np.int32
    assert doc_len.tolist() == [len(rerank_oracle._tokenise(t)) for t in texts] == [4, 5, 2, 1]
    assert tf[:, 0].tolist() == [1, 2, 0, 1] and tf[:, 2].tolist() == tf[:, 0].tolist()  # duplicate query token
    assert tf[:, 1].tolist() == [1, 0, 0, 0] and tf[:, 3].tolist() == [0, 0, 0, 0]


def test_native_term_statistics_match_the_python_restatement(native_lib):
    """lsc_rerank_term_statistics (host code of the C ABI) against the Python restatement of ReRankAgent.tokenise on
    texts with every white-space shape Java's split("\\s+") distinguishes; U+00A0 and U+2003 are NOT separators."""
    import random
    from langstream_b200._native import NativeError, rerank_term_statistics
    from langstream_b200.rerank import term_statistics
    import pytest

    rng = random.Random(4)
    alphabet = ["a", "b", "text", "is", "\u00e9t\u00e9", "x\u00a0y", "\u2003", " ", "  ", "\t", "\n", "\x0b", "\f", "\r"]
    texts = ["".join(rng.choice(alphabet) for _ in range(rng.randint(0, 12))) for _ in range(200)]
    texts += ["", "   ", " a", "a ", "a\x0bb", "x\u00a0y x\u00a0y"]
    for query in ["text is text missing", " a b", "", "x\u00a0y", "\t"]:
        tf_py, len_py = term_statistics(texts, query)
        tf_c, len_c = rerank_term_statistics(texts, query)
        assert len_c.tolist() == len_py.tolist()
        assert tf_c.shape == tf_py.shape and (tf_c == tf_py).all()
    with pytest.raises(NativeError, match="more terms"):
        rerank_term_statistics(["a"], "a b c", max_terms=2)
    tf0, len0 = rerank_term_statistics([], "a b")
    assert tf0.shape == (0, 2) and len0.shape == (0,)


def test_agent_configuration_defaults():
    import pytest
    from langstream_b200.rerank import CudaReRankAgent

    a = CudaReRankAgent()
    with pytest.raises(ValueError, match="field"):
        a.init({"output-field": "value.x"})
    a.init({"field": "value.docs", "output-field": "value.docs"})
    # ReRankAgent.java:80-90: algorithm none, max 100, lambda 0.5, k1 1.5, b 0.75
    assert (a.algorithm, a.max, a.mmr_lambda, a.bm25_k1, a.bm25_b) == ("none", 100, 0.5, 1.5, 0.75)
    docs = [{"text": "x"}]
    assert a.rerank(docs, "anything") == docs   # algorithm none never touches the GPU
