"""Test doubles for the CPU suite: a NumPy index / encoder with the interface of the native handles,
built on the ORACLE (test infrastructure only).  They let the host-side logic (providers, SQL parsing,
micro-batching, sharding) run without a GPU; the -m gpu tests run the same flows on the CUDA path."""
from __future__ import annotations

import numpy as np

from oracle import topk_oracle
from oracle.encoder_oracle import bert_numpy


class FakeIndex:
    def __init__(self, dim, device=0):
        self.dim = dim
        self.ids: list = []
        self.rows: list = []

    def __len__(self):
        return len(self.ids)

    def upsert(self, ids, vecs):
        vecs = np.asarray(vecs, np.float32).reshape(len(ids), self.dim)
        for i, v in zip(np.asarray(ids, np.int64), vecs):
            if int(i) in self.ids:
                self.rows[self.ids.index(int(i))] = v.copy()
            else:
                self.ids.append(int(i))
                self.rows.append(v.copy())

    def delete(self, ids):
        for i in np.asarray(ids, np.int64):
            if int(i) in self.ids:
                j = self.ids.index(int(i))
                self.ids[j], self.rows[j] = self.ids[-1], self.rows[-1]  # swap-with-last like the native index
                self.ids.pop()
                self.rows.pop()

    def search(self, queries, k):
        q = np.asarray(queries, np.float32).reshape(-1, self.dim)
        if not self.ids:
            return np.full((len(q), k), -1, np.int64), np.full((len(q), k), -np.inf, np.float32)
        slots, scores = topk_oracle.search_numpy(np.stack(self.rows), q, k)
        ids = np.where(slots >= 0, np.asarray(self.ids, np.int64)[np.maximum(slots, 0)], -1)
        return ids, scores

    def save(self, path):
        np.savez(path + ".npz", dim=self.dim, ids=np.asarray(self.ids, np.int64),
                 rows=np.stack(self.rows) if self.rows else np.zeros((0, self.dim), np.float32))
        open(path, "wb").close()  # the native file name exists too

    @classmethod
    def load(cls, path, device=0, dim=None):
        z = np.load(path + ".npz")
        ix = cls(int(z["dim"]), device)
        ix.ids = [int(i) for i in z["ids"]]
        ix.rows = [r.copy() for r in z["rows"]]
        return ix

    def close(self):
        pass


class FakeEncoder:
    def __init__(self, spec, blob):
        self.spec, self.blob, self.calls = spec, blob, []

    def embed(self, token_ids, cu_seqlens):
        self.calls.append(len(cu_seqlens) - 1)
        for i in range(len(cu_seqlens) - 1):
            if cu_seqlens[i + 1] - cu_seqlens[i] > self.spec.max_position:
                raise RuntimeError("Illegal input, maybe the number of tokens is too high")
        return bert_numpy(self.spec, self.blob, token_ids, cu_seqlens, dtype=np.float32)

    def close(self):
        pass
