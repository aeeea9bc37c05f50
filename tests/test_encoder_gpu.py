"""Encoder forward through the C ABI against the fp32 CPU oracle.
Tolerance (BASELINE.json north_star): per-vector cosine >= 0.999 and max-abs error <= 2e-3."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).parent / "golden"
COS_MIN = 0.999
MAX_ABS = 2e-3


def _margins(got, want):
    """(cos_min, max_abs) of a batch, the two quantities BASELINE.json's tolerance is stated in."""
    num = (got * want).sum(1)
    den = np.linalg.norm(got, axis=1) * np.linalg.norm(want, axis=1)
    ok = den > 0
    cos_min = float((num[ok] / den[ok]).min()) if ok.any() else 1.0
    return cos_min, float(np.abs(got - want).max())


def _report(name, got, want):
    """Print the parity margin of a headline configuration and append it to gpurun_out/parity_margins.jsonl (the
    figures quoted in DESIGN.md / profiles/ come from there)."""
    import json

    cos_min, max_abs = _margins(got, want)
    rec = {"case": name, "rows": int(got.shape[0]), "cos_min": cos_min, "max_abs": max_abs,
           "cos_budget": COS_MIN, "max_abs_budget": MAX_ABS}
    print("PARITY", json.dumps(rec))
    out = Path(__file__).resolve().parents[1] / "gpurun_out"
    try:
        out.mkdir(exist_ok=True)
        with open(out / "parity_margins.jsonl", "a") as f:
            f.write(json.dumps(rec) + "\n")
    except OSError:
        pass
    return cos_min, max_abs


def _check(got, want):
    assert got.shape == want.shape
    assert np.isfinite(got).all()
    for i in range(want.shape[0]):
        nw = np.linalg.norm(want[i])
        if nw == 0:
            assert np.abs(got[i]).max() == 0
            continue
        cos = float(got[i] @ want[i] / (np.linalg.norm(got[i]) * nw))
        assert cos >= COS_MIN, f"row {i}: cosine {cos}"
        assert abs(np.linalg.norm(got[i]) - 1.0) < 1e-3
    assert np.abs(got - want).max() <= MAX_ABS, f"max-abs {np.abs(got - want).max()}"


@pytest.mark.parametrize("gemm_impl", [1, 0])
def test_tiny_golden(native_lib, cuda_device, gemm_impl):
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    fx = np.load(GOLDEN / "encoder_tiny.npz")
    blob = W.random_init(W.TINY, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    enc = Encoder(W.TINY, blob, gemm_impl=gemm_impl)
    got = enc.embed(fx["token_ids"], fx["cu_seqlens"])
    _check(got, fx["expected"])
    enc.close()


@pytest.mark.parametrize("gemm_impl", [1, 0])
def test_minilm_golden(native_lib, cuda_device, gemm_impl):
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    fx = np.load(GOLDEN / "encoder_minilm.npz")
    blob = W.random_init(W.MINILM_L6, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    enc = Encoder(W.MINILM_L6, blob, gemm_impl=gemm_impl)
    got = enc.embed(fx["token_ids"], fx["cu_seqlens"])
    _check(got, fx["expected"])
    enc.close()


def test_minilm_config1_live_oracle(native_lib, cuda_device):
    """BASELINE config 1 shape: 32 x 128 tokens, perturbed weights, live HuggingFace oracle run one
    text per forward like the reference."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=3, perturb=True)
    ids, cu = W.synthetic_token_ids(32, 128, seed=1234)
    want = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    _check(got, want)
    enc.close()


def test_minilm_ragged_and_batch_invariance(native_lib, cuda_device):
    """Packed variable-length batch: each sequence's embedding does not depend on its neighbours
    (the reference embeds texts independently), checked against the oracle on a subset."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=5, perturb=True)
    ids, cu = W.synthetic_token_ids(300, 128, seed=77, min_len=1)
    enc = Encoder(spec, blob)
    all_out = enc.embed(ids, cu)
    pick = [0, 1, 17, 150, 299]
    sub_ids = np.concatenate([ids[cu[i]:cu[i + 1]] for i in pick])
    sub_cu = np.concatenate([[0], np.cumsum([cu[i + 1] - cu[i] for i in pick])]).astype(np.int32)
    sub_out = enc.embed(sub_ids, sub_cu)
    assert np.abs(sub_out - all_out[pick]).max() <= 1e-6  # same kernels, same per-sequence math
    want = HFOracle(spec, blob).embed_one_by_one(sub_ids, sub_cu)
    _check(sub_out, want)
    enc.close()


def test_base_shape_subset(native_lib, cuda_device):
    """12L/768-d (BGE-base shape), 512-token and ragged sequences, oracle on 6 sequences."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.BERT_BASE
    blob = W.random_init(spec, seed=1, perturb=True)
    ids_a, cu_a = W.synthetic_token_ids(2, 512, seed=1)
    ids_b, cu_b = W.synthetic_token_ids(4, 512, seed=2, min_len=64)
    ids = np.concatenate([ids_a, ids_b])
    cu = np.concatenate([cu_a, cu_a[-1] + cu_b[1:]]).astype(np.int32)
    want = HFOracle(spec, blob).embed_batched(ids, cu, batch=2)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    _check(got, want)
    enc.close()


def test_edge_cases(native_lib, cuda_device):
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder, NativeError, LSC_E_TOO_LONG

    spec = W.TINY
    enc = Encoder(spec, W.random_init(spec, 0, True))
    # empty batch, and a batch of empty sequences
    assert enc.embed(np.zeros(0, np.int32), np.zeros(1, np.int32)).shape == (0, spec.hidden)
    out = enc.embed(np.zeros(0, np.int32), np.zeros(4, np.int32))
    assert out.shape == (3, spec.hidden) and (out == 0).all()
    # maximum length is accepted, one more token is the reference's "too many tokens" error
    ids, cu = W.synthetic_token_ids(1, spec.max_position, vocab=spec.vocab)
    assert np.isfinite(enc.embed(ids, cu)).all()
    ids, cu = W.synthetic_token_ids(1, spec.max_position + 1, vocab=spec.vocab)
    with pytest.raises(NativeError) as ei:
        enc.embed(ids, cu)
    assert ei.value.code == LSC_E_TOO_LONG and "number of tokens is too high" in ei.value.message
    with pytest.raises(NativeError):
        enc.embed(np.array([spec.vocab], np.int32), np.array([0, 1], np.int32))  # id out of range
    enc.close()


def test_device_api_matches_host_api(native_lib, cuda_device):
    import torch
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    ids, cu = W.synthetic_token_ids(64, 128, seed=4, min_len=16)
    host = enc.embed(ids, cu)
    d_ids = torch.from_numpy(ids).to(cuda_device)
    d_cu = torch.from_numpy(cu).to(cuda_device)
    d_out = torch.empty(64, spec.hidden, device=cuda_device)
    enc.embed_dev(d_ids, d_cu, 64, int(cu[-1]), int(np.diff(cu).max()), d_out,
                  torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy(), host)
    assert enc.launch_count > 0
    enc.close()


def test_submit_wait_pipeline_matches_blocking_call(native_lib, cuda_device):
    """lsc_encoder_submit / lsc_encoder_wait: the same embeddings as lsc_encoder_embed, in submission order, with
    two batches in flight, out-of-order waits, an empty batch and a repeated wait."""
    import torch
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder, NativeError

    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    batches = []
    for i, (n, L) in enumerate([(64, 128), (17, 40), (200, 64), (1, 1), (96, 128), (33, 90)]):
        ids, cu = W.synthetic_token_ids(n, L, seed=100 + i, min_len=1)
        h_ids, h_cu = torch.from_numpy(ids).pin_memory(), torch.from_numpy(cu).pin_memory()
        h_out = torch.full((n, spec.hidden), float("nan")).pin_memory()
        batches.append((h_ids.numpy(), h_cu.numpy(), h_out.numpy(), enc.embed(ids, cu).copy()))
    tickets = [enc.submit(b[0], b[1], b[2]) for b in batches[:2]]
    for i in range(2, len(batches)):            # keep two in flight
        enc.wait(tickets[i - 2])
        tickets.append(enc.submit(batches[i][0], batches[i][1], batches[i][2]))
    enc.wait(tickets[-1])                       # newest first, then the older one, then again
    enc.wait(tickets[-2])
    enc.wait(tickets[-2])
    assert tickets == list(range(len(batches)))
    for _, _, out, ref in batches:
        assert np.array_equal(out, ref)         # same kernels, same order of operations: identical bits
    empty_out = np.zeros((0, spec.hidden), np.float32)
    t = enc.submit(np.zeros(0, np.int32), np.zeros(1, np.int32), empty_out)
    enc.wait(t)
    with pytest.raises(NativeError):
        enc.wait(10_000)
    with pytest.raises(NativeError, match="vocabulary"):
        enc.submit(np.array([spec.vocab], np.int32), np.array([0, 1], np.int32), np.zeros((1, spec.hidden), np.float32))
    enc.close()


@pytest.mark.parametrize("ragged", [False, True])
def test_headline_config_all_1024_rows(native_lib, cuda_device, ragged):
    """BASELINE configs[1] at its full size: 1024 sequences of 128 tokens, fixed length
    and UniformInt[16,128] packed, EVERY row against the HuggingFace fp32 oracle run one text per forward like the
    reference (AbstractHuggingFaceEmbeddingService.java:191-206), on the weights bench.py uses (seed 0)."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=0)
    ids, cu = W.synthetic_token_ids(1024, 128, seed=1234, min_len=16 if ragged else None)
    want = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    enc.close()
    _report("minilm_1024x128_" + ("ragged16-128" if ragged else "fixed"), got, want)
    _check(got, want)


def test_headline_config_perturbed_weights_1024_rows(native_lib, cuda_device):
    """Same size with randomised biases / LayerNorm parameters (HF's default init leaves them at 0 / 1, which
    would hide an error in the LayerNorm folding)."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=3, perturb=True)
    ids, cu = W.synthetic_token_ids(1024, 128, seed=4321, min_len=16)
    want = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    enc.close()
    _report("minilm_1024x128_ragged_perturbed", got, want)
    _check(got, want)


def test_base_shape_64_sequences(native_lib, cuda_device):
    """SURVEY.md 8(d) config 3's parity subset: 12L/768-d, 64 sequences - 16 of exactly 512 tokens, 48 drawn from
    UniformInt[64,512] - against the fp32 oracle."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.BERT_BASE
    blob = W.random_init(spec, seed=0)
    ids_a, cu_a = W.synthetic_token_ids(16, 512, seed=99)
    ids_b, cu_b = W.synthetic_token_ids(48, 512, seed=98, min_len=64)
    ids = np.concatenate([ids_a, ids_b])
    cu = np.concatenate([cu_a, cu_a[-1] + cu_b[1:]]).astype(np.int32)
    want = HFOracle(spec, blob).embed_batched(ids, cu, batch=4)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    enc.close()
    _report("base_64seq_512tok", got, want)
    _check(got, want)


def test_two_threads_submitting_keep_their_tickets_apart(native_lib, cuda_device):
    """The handle promises thread safety: two threads submit at the same time, one of them a batch of more than
    262,144 tokens (served in several passes inside submit).  Every ticket is distinct and every batch comes back
    as the blocking call computes it."""
    import threading

    import torch
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    shapes = {0: [(2100, 128), (40, 64), (64, 128)], 1: [(48, 100), (300, 128), (7, 30), (128, 128)]}
    work = {}
    for t, lst in shapes.items():
        work[t] = []
        for i, (n, L) in enumerate(lst):
            ids, cu = W.synthetic_token_ids(n, L, seed=500 + 10 * t + i, min_len=L if n > 1000 else 1)
            ref = enc.embed(ids, cu).copy()
            h_ids, h_cu = torch.from_numpy(ids).pin_memory(), torch.from_numpy(cu).pin_memory()
            h_out = torch.full((n, spec.hidden), float("nan")).pin_memory()
            work[t].append((h_ids, h_cu, h_out, ref))
    assert int(work[0][0][1][-1]) > 262_144
    tickets, errors = {0: [], 1: []}, []
    start = threading.Barrier(2)

    def run(t):
        try:
            start.wait()
            for rep in range(3):
                for h_ids, h_cu, h_out, _ in work[t]:
                    tk = enc.submit(h_ids.numpy(), h_cu.numpy(), h_out.numpy())
                    tickets[t].append(tk)
                    enc.wait(tk)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=run, args=(t,)) for t in (0, 1)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    every = tickets[0] + tickets[1]
    assert len(set(every)) == len(every) == 3 * (len(shapes[0]) + len(shapes[1]))
    for t in (0, 1):
        for _, _, h_out, ref in work[t]:
            assert np.array_equal(h_out.numpy(), ref)
    enc.close()


def test_passes_on_different_streams_do_not_overlap(native_lib, cuda_device):
    """All passes of a handle share one activation workspace; lsc_encoder_embed_dev on two caller streams (and the
    handle's own stream through the host call) must still give each batch its own result."""
    import torch
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    spec = W.MINILM_L6
    enc = Encoder(spec, W.random_init(spec, 0))
    batches = []
    for i in range(4):
        ids, cu = W.synthetic_token_ids(256, 128, seed=900 + i, min_len=8)
        batches.append((ids, cu, enc.embed(ids, cu).copy()))
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = []
    dev = [(torch.from_numpy(i).to(cuda_device), torch.from_numpy(c).to(cuda_device)) for i, c, _ in batches]
    torch.cuda.synchronize()
    for rep in range(3):
        outs = []
        for j, ((ids, cu, _), (d_ids, d_cu)) in enumerate(zip(batches, dev)):
            d_out = torch.empty(256, spec.hidden, device=cuda_device)
            enc.embed_dev(d_ids, d_cu, 256, int(cu[-1]), int(np.diff(cu).max()), d_out, streams[j & 1].cuda_stream)
            outs.append(d_out)
        host = enc.embed(batches[0][0], batches[0][1])  # the handle's own stream, behind the four passes
        torch.cuda.synchronize()
        assert np.array_equal(host, batches[0][2])
        for d_out, (_, _, ref) in zip(outs, batches):
            assert np.array_equal(d_out.cpu().numpy(), ref)
    enc.close()


def test_short_sequences_take_the_packed_attention_tiles(native_lib, cuda_device):
    """Texts of a few dozen tokens (average below 48): several sequences share one 128-row attention tile with a
    block-diagonal mask.  Every row against the fp32 oracle, and each embedding independent of its neighbours."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder
    from oracle.encoder_oracle import HFOracle

    spec = W.MINILM_L6
    blob = W.random_init(spec, seed=5, perturb=True)
    ids, cu = W.synthetic_token_ids(600, 64, seed=78, min_len=1)
    assert cu[-1] < 48 * 600
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    want = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    _report("minilm_600_short_sequences_packed_tiles", got, want)
    _check(got, want)
    pick = [0, 3, 4, 5, 250, 599]   # other neighbours, other positions inside their tiles
    sub_ids = np.concatenate([ids[cu[i]:cu[i + 1]] for i in pick])
    sub_cu = np.concatenate([[0], np.cumsum([cu[i + 1] - cu[i] for i in pick])]).astype(np.int32)
    sub = enc.embed(sub_ids, sub_cu)
    # same math per sequence; the fp32 summation order inside a tile depends on the position, and one flipped bf16
    # rounding of an attention output travels through the remaining layers
    assert np.abs(sub - got[pick]).max() <= 2e-4
    enc.close()


def test_mpnet_golden(native_lib, cuda_device):
    """The committed MPNet fixture (transformers.MPNetModel, tests/golden/make_golden.py) through the C ABI."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder

    fx = np.load(GOLDEN / "encoder_mpnet_tiny.npz")
    blob = W.random_init(W.MPNET_TINY, seed=int(fx["seed"]), perturb=bool(fx["perturb"]))
    enc = Encoder(W.MPNET_TINY, blob)
    got = enc.embed(fx["token_ids"], fx["cu_seqlens"])
    _check(got, fx["expected"])
    enc.close()


@pytest.mark.parametrize("which", ["tiny", "base"])
def test_mpnet_relative_position_bias(native_lib, cuda_device, which):
    """MPNet (BASELINE configs[2] names all-mpnet-base-v2): no token types, positions from 2, and the bucketed
    relative-position bias of MPNetEncoder.compute_position_bias added to the scores inside the d_head 64 attention
    kernel - against transformers.MPNetModel on the same random-init weights (bias table scaled to trained magnitudes).
    Lengths cover both directions of every bucket, the saturated range (|distance| >= 128) and several key tiles."""
    from langstream_b200 import weights as W
    from langstream_b200._native import Encoder, NativeError
    from oracle.encoder_oracle import HFOracle

    if which == "tiny":
        spec, lens = W.MPNET_TINY, [1, 2, 7, 9, 17, 63, 64, 65, 128, 129, 130, 200, 257, 298, 0, 33]
    else:
        spec, lens = W.MPNET_BASE, [512, 384, 100, 17, 300, 129]
    blob = W.random_init(spec, seed=5, perturb=True)
    rng = np.random.default_rng(9)
    cu = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    ids = rng.integers(4, spec.vocab, size=int(cu[-1])).astype(np.int32)  # never <pad> = 1
    want = HFOracle(spec, blob).embed_one_by_one(ids, cu)
    enc = Encoder(spec, blob)
    got = enc.embed(ids, cu)
    _report("mpnet_" + which, got, want)
    _check(got, want)
    # the bias matters: the same weights with a zero table must NOT pass
    state = W.unpack(spec, blob.copy())
    state = {k: np.array(v) for k, v in state.items()}
    state["encoder.relative_attention_bias.weight"][:] = 0.0
    enc0 = Encoder(spec, W.pack(spec, state))
    got0 = enc0.embed(ids, cu)
    enc0.close()
    assert np.abs(got0 - want).max() > 2 * MAX_ABS
    # positions: max_position - 2 tokens fit, one more does not (the reference's "Illegal input" message)
    too_long = np.concatenate([[0], [spec.max_len + 1]]).astype(np.int32)
    with pytest.raises(NativeError, match="Illegal input"):
        enc.embed(np.full(spec.max_len + 1, 5, np.int32), too_long)
    enc.close()
