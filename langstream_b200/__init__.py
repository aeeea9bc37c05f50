"""langstream_b200 — B200-native replacement for LangStream's embeddings + vector top-k hot path.

Only what the hot path needs lives here: the CUDA kernels and C ABI (`csrc/`, built into
`_lib/liblangstream_cuda.so`), the ctypes binding (`_native`), the weight-blob format (`weights`)
and host-side mirrors of the reference plugin interfaces (`embeddings`, `vector`).
There is no CPU fallback: loading the native library is mandatory.
"""
__all__ = ["build", "_native", "weights", "embeddings", "vector", "distributed"]
__version__ = "0.1"
