"""Encoder configuration and the fp32 weight-blob format of `lsc_encoder_create`.

The blob is the HuggingFace `BertModel` state_dict (the graph DJL traces and runs, reference:
langstream-agents/langstream-ai-agents/src/main/java/com/datastax/oss/streaming/ai/embeddings/
AbstractHuggingFaceEmbeddingService.java:132-161) flattened in the fixed order documented in
include/langstream_cuda.h.  `random_init` reproduces HuggingFace's default initialisation
(normal(0, 0.02) matrices and embeddings, zero biases, unit LayerNorm) from a NumPy seed so that the
CUDA path, the oracle and the benchmark all see the same random-init weights without any
checkpoint on disk.
"""
from __future__ import annotations

from dataclasses import dataclass, asdict
from typing import Dict, Iterator, List, Tuple

import numpy as np


@dataclass(frozen=True)
class EncoderSpec:
    layers: int = 6
    hidden: int = 384
    heads: int = 12
    ffn: int = 1536
    vocab: int = 30522
    max_position: int = 512
    type_vocab: int = 2
    ln_eps: float = 1e-12
    # "bert" = HuggingFace BertModel; "mpnet" = MPNetModel (all-mpnet-base-v2): no token-type embeddings, position ids
    # from 2, relative-position attention bias shared by all layers (include/langstream_cuda.h: lsc_encoder_config.arch)
    arch: str = "bert"
    rel_buckets: int = 32
    rel_max_distance: int = 128

    @property
    def arch_id(self) -> int:
        return {"bert": 0, "mpnet": 1}[self.arch]

    @property
    def max_len(self) -> int:
        """Longest sequence the position table serves (MPNet spends rows 0 and 1 on padding_idx)."""
        return self.max_position - (2 if self.arch == "mpnet" else 0)

    def to_dict(self) -> dict:
        return asdict(self)


MINILM_L6 = EncoderSpec()  # all-MiniLM-L6-v2 architecture
BERT_BASE = EncoderSpec(layers=12, hidden=768, heads=12, ffn=3072)  # BGE-base architecture
TINY = EncoderSpec(layers=2, hidden=128, heads=4, ffn=256, vocab=1000, max_position=64)
MPNET_BASE = EncoderSpec(layers=12, hidden=768, heads=12, ffn=3072, vocab=30527, max_position=514, type_vocab=0,
                         ln_eps=1e-5, arch="mpnet")  # all-mpnet-base-v2 architecture
MPNET_TINY = EncoderSpec(layers=2, hidden=256, heads=4, ffn=512, vocab=1000, max_position=300, type_vocab=0,
                         ln_eps=1e-5, arch="mpnet")


def tensor_order(spec: EncoderSpec) -> List[Tuple[str, Tuple[int, ...]]]:
    """(state_dict key, shape) in blob order."""
    H, F = spec.hidden, spec.ffn
    if spec.arch == "mpnet":
        return _tensor_order_mpnet(spec)
    out: List[Tuple[str, Tuple[int, ...]]] = [
        ("embeddings.word_embeddings.weight", (spec.vocab, H)),
        ("embeddings.position_embeddings.weight", (spec.max_position, H)),
        ("embeddings.token_type_embeddings.weight", (spec.type_vocab, H)),
        ("embeddings.LayerNorm.weight", (H,)),
        ("embeddings.LayerNorm.bias", (H,)),
    ]
    for l in range(spec.layers):
        p = f"encoder.layer.{l}."
        out += [
            (p + "attention.self.query.weight", (H, H)),
            (p + "attention.self.query.bias", (H,)),
            (p + "attention.self.key.weight", (H, H)),
            (p + "attention.self.key.bias", (H,)),
            (p + "attention.self.value.weight", (H, H)),
            (p + "attention.self.value.bias", (H,)),
            (p + "attention.output.dense.weight", (H, H)),
            (p + "attention.output.dense.bias", (H,)),
            (p + "attention.output.LayerNorm.weight", (H,)),
            (p + "attention.output.LayerNorm.bias", (H,)),
            (p + "intermediate.dense.weight", (F, H)),
            (p + "intermediate.dense.bias", (F,)),
            (p + "output.dense.weight", (H, F)),
            (p + "output.dense.bias", (H,)),
            (p + "output.LayerNorm.weight", (H,)),
            (p + "output.LayerNorm.bias", (H,)),
        ]
    return out


def _tensor_order_mpnet(spec: EncoderSpec) -> List[Tuple[str, Tuple[int, ...]]]:
    """MPNetModel state_dict in blob order: the BERT order with the relative-position bias behind the embeddings."""
    H, F = spec.hidden, spec.ffn
    out: List[Tuple[str, Tuple[int, ...]]] = [
        ("embeddings.word_embeddings.weight", (spec.vocab, H)),
        ("embeddings.position_embeddings.weight", (spec.max_position, H)),
        ("embeddings.LayerNorm.weight", (H,)),
        ("embeddings.LayerNorm.bias", (H,)),
        ("encoder.relative_attention_bias.weight", (spec.rel_buckets, spec.heads)),
    ]
    for l in range(spec.layers):
        p = f"encoder.layer.{l}."
        for n in ("q", "k", "v", "o"):
            out += [(p + f"attention.attn.{n}.weight", (H, H)), (p + f"attention.attn.{n}.bias", (H,))]
        out += [
            (p + "attention.LayerNorm.weight", (H,)),
            (p + "attention.LayerNorm.bias", (H,)),
            (p + "intermediate.dense.weight", (F, H)),
            (p + "intermediate.dense.bias", (F,)),
            (p + "output.dense.weight", (H, F)),
            (p + "output.dense.bias", (H,)),
            (p + "output.LayerNorm.weight", (H,)),
            (p + "output.LayerNorm.bias", (H,)),
        ]
    return out


def num_params(spec: EncoderSpec) -> int:
    return sum(int(np.prod(s)) for _, s in tensor_order(spec))


def pack(spec: EncoderSpec, state: Dict[str, np.ndarray]) -> np.ndarray:
    """state_dict (numpy fp32 arrays) -> flat fp32 blob."""
    parts = []
    for name, shape in tensor_order(spec):
        a = np.asarray(state[name], dtype=np.float32)
        if tuple(a.shape) != tuple(shape):
            raise ValueError(f"{name}: expected shape {shape}, got {a.shape}")
        parts.append(a.reshape(-1))
    return np.ascontiguousarray(np.concatenate(parts))


def unpack(spec: EncoderSpec, blob: np.ndarray) -> Dict[str, np.ndarray]:
    """flat fp32 blob -> state_dict of numpy views."""
    blob = np.asarray(blob, dtype=np.float32).reshape(-1)
    if blob.size != num_params(spec):
        raise ValueError(f"blob has {blob.size} values, spec needs {num_params(spec)}")
    out: Dict[str, np.ndarray] = {}
    o = 0
    for name, shape in tensor_order(spec):
        n = int(np.prod(shape))
        out[name] = blob[o:o + n].reshape(shape)
        o += n
    return out


def random_init(spec: EncoderSpec, seed: int = 0, perturb: bool = False) -> np.ndarray:
    """HuggingFace-default random initialisation as a blob.

    perturb=True additionally randomises biases and LayerNorm parameters (HF initialises them to
    0 / 1, which would hide bias and affine bugs from a parity test).
    """
    rng = np.random.default_rng(seed)
    state: Dict[str, np.ndarray] = {}
    for name, shape in tensor_order(spec):
        if name.endswith("LayerNorm.weight"):
            a = np.ones(shape, np.float32)
            if perturb:
                a = a + 0.1 * rng.standard_normal(shape, dtype=np.float32)
        elif name.endswith(".bias"):
            a = np.zeros(shape, np.float32)
            if perturb:
                a = 0.05 * rng.standard_normal(shape, dtype=np.float32)
        else:
            a = 0.02 * rng.standard_normal(shape, dtype=np.float32)
            if name == "embeddings.word_embeddings.weight":
                a[1 if spec.arch == "mpnet" else 0] = 0.0  # padding_idx row, as nn.Embedding(padding_idx=...) initialises it
            if name == "encoder.relative_attention_bias.weight":
                a = a * 100.0  # trained tables span several units; 0.02 would hide the bias from a parity test
        state[name] = a.astype(np.float32)
    return pack(spec, state)


def synthetic_token_ids(n_seq: int, max_len: int, seed: int = 1234, min_len: int | None = None,
                        vocab: int = 30522) -> Tuple[np.ndarray, np.ndarray]:
    """Synthetic packed token ids: [CLS]=101 first, [SEP]=102 last, random ids between.

    min_len=None gives fixed-length sequences; otherwise lengths ~ UniformInt[min_len, max_len].
    Returns (token_ids int32 [total], cu_seqlens int32 [n_seq+1]).
    """
    rng = np.random.default_rng(seed)
    lo, hi = min(1000, vocab - 2), vocab
    if min_len is None:
        lens = np.full(n_seq, max_len, np.int64)
    else:
        lens = rng.integers(min_len, max_len + 1, size=n_seq)
    cu = np.zeros(n_seq + 1, np.int32)
    cu[1:] = np.cumsum(lens)
    ids = rng.integers(lo, hi, size=int(cu[-1])).astype(np.int32)
    cls_id, sep_id = min(101, vocab - 1), min(102, vocab - 1)
    for i in range(n_seq):
        if lens[i] >= 1:
            ids[cu[i]] = cls_id
        if lens[i] >= 2:
            ids[cu[i + 1] - 1] = sep_id
    return ids, cu
