"""CPU oracle for the `compute-ai-embeddings` local-model path.  TEST INFRASTRUCTURE ONLY.

What the reference executes (third-party, NOT in the reference tree): DJL 0.24.0
`TextEmbeddingTranslatorFactory` + PyTorch engine 2.0.1 CPU run a TorchScript trace of the
HuggingFace BERT model in fp32, then mean-pool over the attention mask and L2-normalise
(call sites: langstream-agents/langstream-ai-agents/src/main/java/com/datastax/oss/streaming/ai/
embeddings/AbstractHuggingFaceEmbeddingService.java:45-49 engine, :132-161 model load,
:183-207 one `predictor.predict(text)` per text; float[] -> List<Double> widening at
.../HuggingFaceEmbeddingService.java:52-63).

Two restatements live here:
  * `HFOracle`  — the same HuggingFace graph (`transformers.BertModel`, eager attention, fp32, CPU),
                  i.e. the model DJL traces; one forward per text like the reference, or batched.
  * `bert_numpy` — an independent NumPy restatement of that graph (embeddings+LN, post-LN encoder
                  layers with exact-erf GELU, additive mask) used to cross-check the HF oracle and as
                  a dependency-free checker.

PARITY PIN STATUS: the only value-level golden vector in the reference
(langstream-runtime/langstream-runtime-impl/src/test/java/ai/langstream/kafka/ComputeEmbeddingsIT.java:205-210,
384 doubles for "something to embed", model intfloat/multilingual-e5-small) needs weights that are
not in the tree and cannot be downloaded here, so ENCODER VALUES ARE "parity unpinned" against the
reference's own tests.  What IS pinned: output dimension and unit L2 norm of that vector (the
evidence for mean-pool + normalise), and `bert_numpy` == `HFOracle` on seeded random-init weights
(tests/test_oracle.py, fixtures in tests/golden/).
"""
from __future__ import annotations

import math
from typing import List

import numpy as np

from langstream_b200.weights import EncoderSpec, unpack


# ---------------------------------------------------------------------------------------------
# HuggingFace graph (what DJL traces)
# ---------------------------------------------------------------------------------------------
class HFOracle:
    def __init__(self, spec: EncoderSpec, blob: np.ndarray, threads: int | None = None):
        import torch
        from transformers import BertConfig, BertModel

        if threads:
            torch.set_num_threads(threads)
        if spec.arch == "mpnet":
            self._init_mpnet(spec, blob)
            return
        cfg = BertConfig(
            vocab_size=spec.vocab, hidden_size=spec.hidden, num_hidden_layers=spec.layers,
            num_attention_heads=spec.heads, intermediate_size=spec.ffn, hidden_act="gelu",
            hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
            max_position_embeddings=spec.max_position, type_vocab_size=spec.type_vocab,
            layer_norm_eps=spec.ln_eps, pad_token_id=0,
        )
        try:
            cfg._attn_implementation = "eager"
        except Exception:  # pragma: no cover
            pass
        self.torch = torch
        self.spec = spec
        self.model = BertModel(cfg, add_pooling_layer=False).eval()
        state = {k: torch.from_numpy(np.array(v)) for k, v in unpack(spec, blob).items()}
        missing, unexpected = self.model.load_state_dict(state, strict=False)
        bad = [m for m in missing if "position_ids" not in m and "token_type_ids" not in m]
        if bad or unexpected:
            raise RuntimeError(f"weight blob does not match BertModel: missing={bad} unexpected={unexpected}")

    def _init_mpnet(self, spec: EncoderSpec, blob: np.ndarray):
        """transformers.MPNetModel: the graph DJL traces for all-mpnet-base-v2 (relative-position bias from
        MPNetEncoder.compute_position_bias, position ids from padding_idx + 1)."""
        import torch
        from transformers import MPNetConfig, MPNetModel

        cfg = MPNetConfig(
            vocab_size=spec.vocab, hidden_size=spec.hidden, num_hidden_layers=spec.layers,
            num_attention_heads=spec.heads, intermediate_size=spec.ffn, hidden_act="gelu",
            hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0,
            max_position_embeddings=spec.max_position, layer_norm_eps=spec.ln_eps,
            relative_attention_num_buckets=spec.rel_buckets, pad_token_id=1,
        )
        if spec.rel_max_distance != 128:
            raise ValueError("MPNetEncoder.relative_position_bucket hard-codes max_distance=128")
        try:
            cfg._attn_implementation = "eager"
        except Exception:  # pragma: no cover
            pass
        self.torch = torch
        self.spec = spec
        self.model = MPNetModel(cfg, add_pooling_layer=False).eval()
        state = {k: torch.from_numpy(np.array(v)) for k, v in unpack(spec, blob).items()}
        missing, unexpected = self.model.load_state_dict(state, strict=False)
        bad = [m for m in missing if "position_ids" not in m]
        if bad or unexpected:
            raise RuntimeError(f"weight blob does not match MPNetModel: missing={bad} unexpected={unexpected}")

    def _pool(self, hidden, mask):
        torch = self.torch
        m = mask.unsqueeze(-1).to(hidden.dtype)
        pooled = (hidden * m).sum(1) / m.sum(1).clamp(min=1e-9)
        return torch.nn.functional.normalize(pooled, p=2, dim=1, eps=1e-12)

    def embed_one_by_one(self, token_ids: np.ndarray, cu_seqlens: np.ndarray) -> np.ndarray:
        """Reference shape: one text per forward pass
        (AbstractHuggingFaceEmbeddingService.java:191-206)."""
        torch = self.torch
        out = np.zeros((len(cu_seqlens) - 1, self.spec.hidden), np.float32)
        with torch.no_grad():
            for i in range(len(cu_seqlens) - 1):
                ids = token_ids[cu_seqlens[i]:cu_seqlens[i + 1]]
                if len(ids) == 0:
                    continue
                t = torch.from_numpy(np.asarray(ids, np.int64))[None, :]
                mask = torch.ones_like(t)
                h = self.model(input_ids=t, attention_mask=mask).last_hidden_state
                out[i] = self._pool(h, mask)[0].numpy()
        return out

    def embed_batched(self, token_ids: np.ndarray, cu_seqlens: np.ndarray, batch: int = 64) -> np.ndarray:
        """Same math, padded batches with an attention mask (charitable to the CPU)."""
        torch = self.torch
        n = len(cu_seqlens) - 1
        out = np.zeros((n, self.spec.hidden), np.float32)
        with torch.no_grad():
            for b0 in range(0, n, batch):
                b1 = min(n, b0 + batch)
                lens = [int(cu_seqlens[i + 1] - cu_seqlens[i]) for i in range(b0, b1)]
                L = max(max(lens), 1)
                ids = np.zeros((b1 - b0, L), np.int64)
                mask = np.zeros((b1 - b0, L), np.int64)
                for j, i in enumerate(range(b0, b1)):
                    ids[j, :lens[j]] = token_ids[cu_seqlens[i]:cu_seqlens[i + 1]]
                    mask[j, :lens[j]] = 1
                rows = [j for j in range(b1 - b0) if lens[j] > 0]
                if not rows:
                    continue
                t, m = torch.from_numpy(ids[rows]), torch.from_numpy(mask[rows])
                h = self.model(input_ids=t, attention_mask=m).last_hidden_state
                res = self._pool(h, m).numpy()
                for r, j in enumerate(rows):
                    out[b0 + j] = res[r]
        return out


def to_list_of_double(emb: np.ndarray) -> List[List[float]]:
    """float[] -> List<Double> widening, as HuggingFaceEmbeddingService.convertOutput does
    (…/embeddings/HuggingFaceEmbeddingService.java:52-63)."""
    return [[float(np.float64(v)) for v in row] for row in np.asarray(emb, np.float32)]


# ---------------------------------------------------------------------------------------------
# Independent NumPy restatement
# ---------------------------------------------------------------------------------------------
_erf = np.vectorize(math.erf, otypes=[np.float64])


def _layer_norm(x, g, b, eps):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)
    return (x - mu) / np.sqrt(var + eps) * g + b


def _mpnet_bias(w: np.ndarray, s: int, buckets: int, max_distance: int) -> np.ndarray:
    """[heads, s, s] relative-position bias: bias[h, q, k] = w[bucket(k - q), h]
    (transformers MPNetEncoder.compute_position_bias / relative_position_bucket: half of the buckets per direction,
    the first half of those exact distances, the rest logarithmic up to max_distance, evaluated in float32)."""
    rel = np.arange(s)[None, :] - np.arange(s)[:, None]
    n = -rel
    nb = buckets // 2
    ret = np.where(n < 0, nb, 0)
    n = np.abs(n)
    max_exact = nb // 2
    with np.errstate(divide="ignore"):
        v = (np.log(n.astype(np.float32) / np.float32(max_exact)) / np.float32(math.log(max_distance / max_exact))
             * np.float32(nb - max_exact))
    large = np.minimum(max_exact + np.where(n >= max_exact, v, 0).astype(np.int64), nb - 1)
    bucket = ret + np.where(n < max_exact, n, large)
    return np.transpose(w[bucket], (2, 0, 1))


def bert_numpy(spec: EncoderSpec, blob: np.ndarray, token_ids: np.ndarray, cu_seqlens: np.ndarray,
               dtype=np.float64) -> np.ndarray:
    """BERT forward + masked mean-pool + L2-normalise, one sequence at a time, in `dtype`."""
    W = {k: v.astype(dtype) for k, v in unpack(spec, blob).items()}
    mpnet = spec.arch == "mpnet"
    if mpnet:  # same graph under MPNetModel's names, plus what differs: see _mpnet_bias and the embeddings below
        ren = {"attention.attn.q.": "attention.self.query.", "attention.attn.k.": "attention.self.key.",
               "attention.attn.v.": "attention.self.value.", "attention.attn.o.": "attention.output.dense.",
               "attention.LayerNorm.": "attention.output.LayerNorm."}
        for k in list(W):
            for a, b in ren.items():
                if a in k:
                    W[k.replace(a, b)] = W.pop(k)
    H, nh = spec.hidden, spec.heads
    d = H // nh
    out = np.zeros((len(cu_seqlens) - 1, H), np.float32)
    for i in range(len(cu_seqlens) - 1):
        ids = np.asarray(token_ids[cu_seqlens[i]:cu_seqlens[i + 1]], np.int64)
        s = len(ids)
        if s == 0:
            continue
        if mpnet:  # MPNetEmbeddings: no token types; positions count from padding_idx + 1 = 2
            x = W["embeddings.word_embeddings.weight"][ids] + W["embeddings.position_embeddings.weight"][2:s + 2]
            bias = _mpnet_bias(W["encoder.relative_attention_bias.weight"], s, spec.rel_buckets, spec.rel_max_distance)
        else:
            x = (W["embeddings.word_embeddings.weight"][ids]
                 + W["embeddings.position_embeddings.weight"][:s]
                 + W["embeddings.token_type_embeddings.weight"][0])
        x = _layer_norm(x, W["embeddings.LayerNorm.weight"], W["embeddings.LayerNorm.bias"], spec.ln_eps)
        for l in range(spec.layers):
            p = f"encoder.layer.{l}."
            q = x @ W[p + "attention.self.query.weight"].T + W[p + "attention.self.query.bias"]
            k = x @ W[p + "attention.self.key.weight"].T + W[p + "attention.self.key.bias"]
            v = x @ W[p + "attention.self.value.weight"].T + W[p + "attention.self.value.bias"]
            q = q.reshape(s, nh, d).transpose(1, 0, 2)
            k = k.reshape(s, nh, d).transpose(1, 0, 2)
            v = v.reshape(s, nh, d).transpose(1, 0, 2)
            sc = q @ k.transpose(0, 2, 1) / math.sqrt(d)
            if mpnet:
                sc = sc + bias
            sc = sc - sc.max(-1, keepdims=True)
            pr = np.exp(sc)
            pr = pr / pr.sum(-1, keepdims=True)
            ctx = (pr @ v).transpose(1, 0, 2).reshape(s, H)
            a = ctx @ W[p + "attention.output.dense.weight"].T + W[p + "attention.output.dense.bias"]
            x = _layer_norm(x + a, W[p + "attention.output.LayerNorm.weight"],
                            W[p + "attention.output.LayerNorm.bias"], spec.ln_eps)
            f = x @ W[p + "intermediate.dense.weight"].T + W[p + "intermediate.dense.bias"]
            f = 0.5 * f * (1.0 + _erf(f / math.sqrt(2.0)).astype(dtype))
            f = f @ W[p + "output.dense.weight"].T + W[p + "output.dense.bias"]
            x = _layer_norm(x + f, W[p + "output.LayerNorm.weight"], W[p + "output.LayerNorm.bias"], spec.ln_eps)
        pooled = x.sum(0) / max(float(s), 1e-9)
        out[i] = (pooled / max(float(np.sqrt((pooled ** 2).sum())), 1e-12)).astype(np.float32)
    return out
