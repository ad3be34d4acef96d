"""Clip sharding across the GPUs of one box: every 30 s window is an independent unit
(the reference truncates to one window per call and carries no state across calls: whisper.cpp:698, :753),
so rank r of G takes a contiguous block of ceil(N/G) windows, runs its own engine replica, and the token ids
(<= 31 int32 per window) are gathered on the host in window order. There is NO collective on the data path;
`gather_tokens` is the only communication and it moves N*32*4 bytes once per job."""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

WINDOW = 480000  # kSampleRate * kChunkSize (whisper.h:34,38)
MAX_TOKENS = 31


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition: [begin, end) of rank's items; the last ranks may get fewer (or none)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    per = -(-n_items // world)
    b = min(n_items, rank * per)
    return b, min(n_items, b + per)


def chunk_windows(pcm: np.ndarray, window: int = WINDOW) -> np.ndarray:
    """Long mono f32 audio -> [n_windows][window], last window zero-padded (BASELINE.json configs[2]: 1 h of
    audio = 120 windows). An empty input yields zero windows."""
    pcm = np.ascontiguousarray(pcm, dtype=np.float32).reshape(-1)
    n = -(-pcm.shape[0] // window)
    out = np.zeros((n, window), dtype=np.float32)
    if n:
        out.reshape(-1)[:pcm.shape[0]] = pcm
    return out


def gather_tokens(ids: np.ndarray, lens: np.ndarray, n_total: int, rank: int, world: int,
                  dist=None) -> Optional[Tuple[np.ndarray, np.ndarray]]:
    """Rank-local results -> (ids [n_total][MAX_TOKENS], lens [n_total]) on rank 0 (None elsewhere), in window
    order. `dist` is torch.distributed (gloo or nccl process group already initialised); world == 1 needs none."""
    ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1, MAX_TOKENS)
    lens = np.ascontiguousarray(lens, dtype=np.int32).reshape(-1)
    if world == 1:
        assert ids.shape[0] == n_total
        return ids, lens
    import torch
    per = -(-n_total // world)
    pad_ids = np.full((per, MAX_TOKENS), -1, dtype=np.int32)
    pad_lens = np.zeros((per,), dtype=np.int32)
    pad_ids[:ids.shape[0]] = ids
    pad_lens[:lens.shape[0]] = lens
    device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.from_numpy(np.concatenate([pad_ids, pad_lens[:, None]], axis=1)).to(device)
    bucket: Optional[List] = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    if dist.get_backend() == "nccl":
        allb = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allb, t)
        bucket = allb if rank == 0 else None
    else:
        dist.gather(t, bucket, dst=0)
    if rank != 0:
        return None
    full = torch.cat(bucket, dim=0).cpu().numpy()[:n_total]
    return np.ascontiguousarray(full[:, :MAX_TOKENS]), np.ascontiguousarray(full[:, MAX_TOKENS])


def transcribe_sharded(transcribe_fn, windows: np.ndarray, rank: int, world: int, dist=None, max_batch: int = 64):
    """Runs `transcribe_fn(pcm [n][WINDOW]) -> (ids [n][MAX_TOKENS], lens [n])` (an Engine.transcribe_tokens bound
    to this rank's device) over this rank's block in batches of max_batch, then gathers on rank 0."""
    b, e = shard_range(windows.shape[0], rank, world)
    ids = np.full((e - b, MAX_TOKENS), -1, dtype=np.int32)
    lens = np.zeros((e - b,), dtype=np.int32)
    for s in range(b, e, max_batch):
        n = min(max_batch, e - s)
        i, l = transcribe_fn(np.ascontiguousarray(windows[s:s + n]))
        ids[s - b:s - b + n] = np.asarray(i)[:, :MAX_TOKENS]
        lens[s - b:s - b + n] = np.asarray(l)
    return gather_tokens(ids, lens, windows.shape[0], rank, world, dist)
