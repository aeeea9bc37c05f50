"""The oracle against every golden vector / known-answer test the reference holds for this path,
and against the committed fixtures.  CPU only."""
import json
from pathlib import Path

import numpy as np
import pytest

GOLDEN = Path(__file__).parent / "golden"


def test_rerank_it_known_answer():
    """RerankAgentRunnerIT.java:62-92 — sink 10 rows, cosine query (LIMIT 20), MMR picks 8 in an
    exact order.  BM25 is 0 for every document here, so the order is decided by the cosine ranking
    and the cosine-based diversity: it pins oracle cosine + ordering."""
    from oracle import rerank_oracle, topk_oracle

    fx = json.loads((GOLDEN / "rerank_it.json").read_text())
    rows = np.array([r["embeddings_vector"] for r in fx["rows"]], np.float32)
    q = np.array(fx["query_embeddings"], np.float32)
    ids, scores = topk_oracle.search_scalar(rows, q[None], fx["limit"])
    got = [i for i in ids[0] if i >= 0]
    assert len(got) == 10 and got[0] == 1            # LIMIT 20 over 10 rows returns all, best first
    assert (ids[0, 10:] == -1).all()
    assert (np.diff(scores[0, :10]) <= 0).all()
    docs = [(fx["rows"][i]["text"], rows[i]) for i in got]
    m = fx["mmr"]
    order = rerank_oracle.rerank_mmr(docs, int(m["max"]), fx["query_text"], m["lambda"], m["k1"], m["b"])
    assert [docs[i][0] for i in order] == [d["text"] for d in fx["expected_related_documents"]]
    for i, d in zip(order, fx["expected_related_documents"]):
        assert docs[i][1].tolist() == d["embeddings_vector"]


def test_cosine_semantics_of_rerank_agent():
    """ReRankAgent.cosineSimilarity (ReRankAgent.java:262-273): 0 when a norm is 0."""
    from oracle import topk_oracle

    z = np.zeros(5, np.float32)
    a = np.array([1, 2, 3, 4, 5], np.float32)
    assert topk_oracle.cosine_scalar(z, a) == 0.0
    assert topk_oracle.cosine_scalar(a, z) == 0.0
    assert abs(topk_oracle.cosine_scalar(a, a) - 1.0) < 1e-6
    assert abs(topk_oracle.cosine_scalar(a, -a) + 1.0) < 1e-6


def test_e5_golden_vector_pins_pooling_and_normalisation():
    """ComputeEmbeddingsIT.java:205-210: the DJL output is 384-d with unit L2 norm (both platform
    variants, which agree to ~1e-7) — the evidence that the path mean-pools and L2-normalises.
    The oracle's outputs have the same dimension and norm; VALUES are unpinned (weights absent)."""
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import bert_numpy

    fx = json.loads((GOLDEN / "e5_small_golden.json").read_text())
    v = np.array(fx["vectors"], np.float64)
    assert v.shape == (2, 384)
    assert np.abs(np.linalg.norm(v, axis=1) - 1.0).max() < 1e-6
    assert np.abs(v[0] - v[1]).max() < 1e-6
    spec = W.EncoderSpec(layers=1, hidden=384, heads=12, ffn=256, vocab=500, max_position=32)
    ids, cu = W.synthetic_token_ids(2, 9, vocab=spec.vocab)
    out = bert_numpy(spec, W.random_init(spec, 0, True), ids, cu)
    assert out.shape == (2, 384)
    assert np.abs(np.linalg.norm(out.astype(np.float64), axis=1) - 1.0).max() < 1e-6


def test_numpy_restatement_matches_hf_graph_and_fixture():
    """bert_numpy == HuggingFace BertModel (the graph DJL traces) == committed fixture."""
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import HFOracle, bert_numpy, to_list_of_double

    fx = np.load(GOLDEN / "encoder_tiny.npz")
    blob = W.random_init(W.TINY, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    ids, cu = fx["token_ids"], fx["cu_seqlens"]
    a = bert_numpy(W.TINY, blob, ids, cu)
    assert np.abs(a - fx["expected"]).max() < 2e-6
    hf = HFOracle(W.TINY, blob)
    b = hf.embed_one_by_one(ids, cu)
    c = hf.embed_batched(ids, cu, batch=4)
    assert np.abs(b - fx["expected"]).max() < 2e-6
    assert np.abs(c - b).max() < 2e-6          # padding + mask == one text per forward
    empty = [i for i in range(len(cu) - 1) if cu[i] == cu[i + 1]]
    assert empty and (b[empty] == 0).all()
    d = to_list_of_double(b[:1])                # float -> Double widening is exact
    assert isinstance(d[0][0], float) and np.float32(d[0][3]) == b[0, 3]


def test_minilm_fixture_is_reproducible_subset():
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import HFOracle

    fx = np.load(GOLDEN / "encoder_minilm.npz")
    blob = W.random_init(W.MINILM_L6, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    ids, cu = fx["token_ids"], fx["cu_seqlens"]
    sel = [5, 7]  # two ragged sequences: keeps the CPU suite fast
    sub_ids = np.concatenate([ids[cu[i]:cu[i + 1]] for i in sel])
    sub_cu = np.concatenate([[0], np.cumsum([cu[i + 1] - cu[i] for i in sel])]).astype(np.int32)
    got = HFOracle(W.MINILM_L6, blob).embed_one_by_one(sub_ids, sub_cu)
    assert np.abs(got - fx["expected"][sel]).max() < 2e-6


def test_topk_fixture_and_fast_form():
    """Scalar (ReRankAgent-form) oracle reproduces the fixture; the fast bf16 matmul form agrees with
    it up to the bf16 storage rounding."""
    from oracle import topk_oracle

    fx = np.load(GOLDEN / "topk_small.npz")
    rows = np.random.default_rng(int(fx["corpus_seed"])).standard_normal((int(fx["n"]), int(fx["dim"])), dtype=np.float32)
    q = np.random.default_rng(int(fx["query_seed"])).standard_normal((int(fx["B"]), int(fx["dim"])), dtype=np.float32)
    ids, scores = topk_oracle.search_scalar(rows, q, int(fx["k"]))
    assert (ids == fx["expected_ids"]).all() and np.array_equal(scores, fx["expected_scores"])
    ids2, scores2 = topk_oracle.search_numpy(rows, q, int(fx["k"]))
    assert np.abs(scores2 - scores).max() < 4e-3
    # bf16 storage can swap near-equal neighbours; every id of the fast form must still be a
    # near-top-k row under the scalar fp32 cosine
    for b in range(ids.shape[0]):
        for i in ids2[b]:
            assert topk_oracle.cosine_scalar(rows[i], q[b]) >= scores[b, -1] - 4e-3


def test_topk_edge_cases():
    from oracle import topk_oracle

    rows = np.array([[1, 0], [1, 0], [0, 1], [0, 0]], np.float32)
    ids, scores = topk_oracle.search_scalar(rows, np.array([[1, 0]], np.float32), 6)
    assert ids[0].tolist() == [0, 1, 2, 3, -1, -1]          # ties by lower row index; zero row scores 0
    assert scores[0, :4].tolist() == [1.0, 1.0, 0.0, 0.0]
    ids, _ = topk_oracle.search_numpy(rows, np.array([[1, 0]], np.float32), 6)
    assert ids[0].tolist() == [0, 1, 2, 3, -1, -1]
    ids, scores = topk_oracle.search_scalar(np.zeros((0, 2), np.float32), np.array([[1, 0]], np.float32), 2)
    assert (ids == -1).all()
    x = np.array([1.0, 1.00390625, 3.1415927, -2.5e-8], np.float32)
    r = topk_oracle.to_bf16_rounded(x)
    assert r[0] == 1.0 and r[1] == 1.0 and abs(r[2] - 3.140625) < 1e-6  # round to nearest even
    with pytest.raises(AssertionError):
        topk_oracle.assert_topk_matches(np.array([[1]]), np.array([[0.5]], np.float32),
                                        np.array([[0.9, 0.5]], np.float32), 1)


def test_mpnet_numpy_restatement_matches_hf():
    """The NumPy restatement of the MPNet graph (bucketed relative-position bias, positions from 2, no token types)
    against transformers.MPNetModel on seeded random-init weights; lengths reach the saturated buckets."""
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import HFOracle, bert_numpy

    spec = W.MPNET_TINY
    blob = W.random_init(spec, seed=3, perturb=True)
    lens = [1, 9, 40, 150, 298]
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    ids = np.random.default_rng(2).integers(4, spec.vocab, size=int(cu[-1])).astype(np.int32)
    a = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    b = bert_numpy(spec, blob, ids, cu)
    assert np.abs(a - b).max() < 5e-6


@pytest.mark.parametrize("buckets,max_distance", [(32, 128), (16, 128), (32, 64), (64, 100)])
def test_mpnet_bias_table_of_the_native_library(native_lib, buckets, max_distance):
    """The table lsc_encoder_create builds for an MPNet model (host code, no GPU): for every relative position in
    [-191, 191] the bucket of transformers' MPNetEncoder.relative_position_bucket (evaluated by torch in float32,
    including the exact boundaries 16 / 32 / 64), times log2(e); beyond +-max_distance the entries are saturated, which
    is what lets the attention kernel clamp its first index."""
    import ctypes as C
    import torch
    from transformers.models.mpnet.modeling_mpnet import MPNetEncoder
    from langstream_b200 import _native

    heads = 12
    rng = np.random.default_rng(buckets + max_distance)
    w = rng.standard_normal((buckets, heads)).astype(np.float32)
    out = np.zeros((heads, 384), np.float32)
    _native.check(native_lib.lsc_debug_mpnet_bias_table(w.ctypes.data, buckets, max_distance, heads, out.ctypes.data))
    rel = torch.arange(-191, 192, dtype=torch.long)
    if max_distance == 128:
        bucket = MPNetEncoder.relative_position_bucket(rel, num_buckets=buckets).numpy()
    else:
        bucket = MPNetEncoder.relative_position_bucket(rel, num_buckets=buckets, max_distance=max_distance).numpy()
    want = (w[bucket].T * np.float32(1.4426950408889634)).astype(np.float32)   # [heads, 383]
    assert np.array_equal(out[:, :383], want)
    sat = 191 - max_distance
    assert (out[:, :sat + 1] == out[:, :1]).all() and (out[:, 382 - sat:383] == out[:, 382:383]).all()


def test_mpnet_golden_fixture_matches_the_numpy_restatement():
    """tests/golden/encoder_mpnet_tiny.npz (written by transformers.MPNetModel in the build container) against the NumPy
    restatement: pins the fixture the GPU test uses to a second, independent evaluation of the MPNet graph."""
    from langstream_b200 import weights as W
    from oracle.encoder_oracle import bert_numpy

    fx = np.load(GOLDEN / "encoder_mpnet_tiny.npz")
    blob = W.random_init(W.MPNET_TINY, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    got = bert_numpy(W.MPNET_TINY, blob, fx["token_ids"], fx["cu_seqlens"])
    assert np.abs(got - fx["expected"]).max() < 5e-6
    assert np.abs(fx["expected"][9]).max() == 0.0   # the empty sequence
