"""N>1 host logic on CPU: world-size-2 gloo process group, utterance sharding with no data-path
collective, off-path gather of the waveforms.  The per-rank "synthesis" is the oracle (tests may use it);
the GPU path of the same logic is bench.py under torchrun."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from neuraltexttospeech_b200.sharding import gather_waveforms, group_by_length, shard_utterances


def test_shard_utterances_partitions_and_balances():
    lengths = [5, 9, 3, 7, 7, 2, 8]
    for world in (1, 2, 3, 4, 8):
        for balance in (False, True):
            parts = [shard_utterances(lengths, world, r, balance) for r in range(world)]
            flat = sorted(i for p in parts for i in p)
            assert flat == list(range(len(lengths)))                    # disjoint cover
            assert max(map(len, parts)) - min(map(len, parts)) <= 1
    a, b = (shard_utterances(lengths, 2, r, True) for r in range(2))
    fa, fb = sum(lengths[i] for i in a), sum(lengths[i] for i in b)
    assert abs(fa - fb) <= max(lengths)                                  # frames balanced within one utterance
    assert shard_utterances([], 2, 1) == []
    with pytest.raises(ValueError):
        shard_utterances(lengths, 2, 2)
    assert group_by_length(lengths, [1, 3, 4, 6]) == [[1], [6], [3, 4]]


def test_call_ragged_groups_equal_lengths_and_keeps_input_order():
    # host logic only: the dense-batch call is replaced by a recorder (the GPU test runs the real one)
    from neuraltexttospeech_b200 import Generator, HParams
    g = Generator.__new__(Generator)
    g.hparams = HParams.builtin("1")
    batches = []

    def fake_call(batch):
        batches.append(batch.shape)
        return np.repeat(batch[:, :, :1], 2, axis=1)          # [B, 2T, 1], tagged by the utterance's first mel bin

    g.call = fake_call
    lengths = [5, 2, 5, 0, 2, 7]
    mels = [np.full((t, 80), float(i), np.float32) for i, t in enumerate(lengths)]
    outs = g.call_ragged(mels)
    assert batches == [(1, 7, 80), (2, 5, 80), (2, 2, 80), (1, 0, 80)]      # longest first, one batch per length
    assert [o.shape for o in outs] == [(2 * t, 1) for t in lengths]
    assert all((o == i).all() for i, o in enumerate(outs))
    assert g.call_ragged([]) == []
    with pytest.raises(ValueError):
        g.call_ragged([np.zeros((4, 79), np.float32)])
    with pytest.raises(ValueError):
        g.call_ragged([np.zeros((1, 4, 80), np.float32)])


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from neuraltexttospeech_b200.hparams import HParams
        from neuraltexttospeech_b200.synthetic import synthetic_mel, synthetic_weights
        from oracle import hifigan_oracle as O
        h = HParams.builtin("2")
        w = synthetic_weights(h, "B")
        B, T = 5, 6
        mel = synthetic_mel(B, T, h.num_mels)
        mine = shard_utterances([T] * B, world, rank)
        local = O.generator_forward(h, w, mel[mine], "torch32") if mine else np.zeros((0, T * 256, 1), np.float32)
        full = gather_waveforms(local, mine, B, dst=0)
        if rank == 0:
            ref = O.generator_forward(h, w, mel, "torch32")
            q.put(float(np.abs(full - ref).max()))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_synthesis_matches_single_process():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
        assert p.exitcode == 0
    # batch items are independent: sharded == unsharded up to fp32 conv algorithm choice per batch size
    assert q.get(timeout=5) <= 1e-5
