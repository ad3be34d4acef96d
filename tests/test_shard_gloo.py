"""N > 1 host path on CPU: world_size-2 gloo process group, clip sharding + host gather (no GPU, so the engine
is replaced by a deterministic per-window function — what is under test is partitioning, batching and ordering)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from whisper_tflite_b200 import shard


def fake_transcribe(pcm):
    """ids depend only on the window's content (like independent 30 s windows do)."""
    n = pcm.shape[0]
    ids = np.full((n, shard.MAX_TOKENS), -1, dtype=np.int32)
    lens = np.zeros(n, dtype=np.int32)
    for i in range(n):
        key = int(np.abs(pcm[i, :64]).sum() * 1000) % 50000
        ln = 5 + key % 27
        ids[i, :ln] = (key + np.arange(ln)) % 51865
        lens[i] = ln
    return ids, lens


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_windows, max_batch, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rs = np.random.RandomState(7)
    audio = rs.uniform(-1, 1, size=n_windows * 3000 - 1234).astype(np.float32)
    windows = shard.chunk_windows(audio, window=3000)
    out = shard.transcribe_sharded(fake_transcribe, windows, rank, world, dist, max_batch=max_batch)
    if rank == 0:
        q.put((out[0], out[1]))
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_windows,max_batch", [(7, 2), (8, 64), (1, 4)])
def test_two_rank_gloo_matches_single_rank(n_windows, max_batch):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_windows, max_batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    ids, lens = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    rs = np.random.RandomState(7)
    audio = rs.uniform(-1, 1, size=n_windows * 3000 - 1234).astype(np.float32)
    windows = shard.chunk_windows(audio, window=3000)
    ref_ids, ref_lens = shard.transcribe_sharded(fake_transcribe, windows, 0, 1, None, max_batch=3)
    np.testing.assert_array_equal(ids, ref_ids)
    np.testing.assert_array_equal(lens, ref_lens)


def test_shard_range_and_chunking():
    for n in (0, 1, 5, 120, 121):
        for w in (1, 2, 4, 8):
            spans = [shard.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            assert max(e - b for b, e in spans) == -(-n // w)
    assert shard.shard_range(120, 3, 8) == (45, 60)  # 1 h of audio over 8 GPUs: 15 windows each
    w = shard.chunk_windows(np.ones(57_600_000 // 100 + 7, np.float32), window=4800)
    assert w.shape == (121, 4800) and w[-1, :7].sum() == 7 and w[-1, 7:].sum() == 0
    assert shard.chunk_windows(np.zeros(0, np.float32)).shape == (0, shard.WINDOW)
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)
