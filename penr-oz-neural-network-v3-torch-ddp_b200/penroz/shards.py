"""Per-rank strided reader over tokenised uint16 `.npy` shards.

Same window arithmetic as the reference's loaders.Loader (loaders.py:43-87), which the training step's data
parallelism is defined by: rank r of W starts at token B*T*r and advances by B*T*W per micro-step; inputs are
tokens[i : i+BT], targets tokens[i+1 : i+BT+1]; when the window (plus one stride of look-ahead) runs past the loaded
tokens, the unread tail is carried over in front of the next shard (wrapping to shard 0 after the last).
The reference module stays in use inside the reference service; this restatement is what runs where the reference
checkout (and its `datasets` dependency) is absent, and tests/test_host_logic.py pins it to golden indices recorded
from the reference.
"""
import logging
import os

import numpy as np

import ddp

log = logging.getLogger(__name__)
DATA_FOLDER = "data"


class Loader:
    def __init__(self, dataset_id: str, begin_shard=0, begin_idx=0, buffer_size=0, idx_offset=0):
        names = os.listdir(DATA_FOLDER) if os.path.isdir(DATA_FOLDER) else []
        self.shards = sorted(n for n in names if dataset_id in n)
        if ddp.master_proc():
            log.info(f"Found {len(self.shards)} shard(s) for {dataset_id}")
        self.shard_idx = begin_shard
        self.token_idx = begin_idx
        self.buffer_size = buffer_size
        self.idx_offset = idx_offset
        self.tokens = np.empty((0,))

    def list(self) -> list[str]:
        return self.shards

    def delete(self):
        for name in self.shards:
            os.remove(os.path.join(DATA_FOLDER, name))

    def _load(self) -> np.ndarray:
        path = os.path.join(DATA_FOLDER, self.shards[self.shard_idx])
        return np.load(path).astype(np.int32)

    def _ensure(self, needed_past_cursor: int):
        """Roll shards until `needed_past_cursor` tokens exist after the cursor (at most one lap)."""
        laps = 0
        while len(self.tokens) < self.token_idx + needed_past_cursor and laps < len(self.shards):
            self.shard_idx = (self.shard_idx + 1) % len(self.shards)
            carry = self.tokens[self.token_idx:]
            self.tokens = np.concatenate((carry, self._load()))
            self.token_idx = 0
            laps += 1

    def next_batch(self, target_offset=1):
        if len(self.tokens) == 0:
            self.tokens = self._load()
        self._ensure(self.idx_offset + target_offset)
        lo, hi = self.token_idx, self.token_idx + self.buffer_size
        inputs = self.tokens[lo:hi]
        targets = self.tokens[lo + target_offset:hi + target_offset] if target_offset > 0 else None
        self.token_idx += self.idx_offset
        return inputs, targets


def write_synthetic_shards(dataset_id: str, vocab: int, n_shards: int = 1, tokens_per_shard: int = 2 ** 24,
                           folder: str = DATA_FOLDER):
    """Synthetic token shards in the format Downloader._save writes (loaders.py:22-25): uint16 .npy,
    `{dataset_id}_{k:06d}.npy`, tokens drawn from default_rng(1234 + k) (synthetic inputs)."""
    os.makedirs(folder, exist_ok=True)
    for k in range(n_shards):
        arr = np.random.default_rng(1234 + k).integers(0, vocab, size=tokens_per_shard).astype(np.uint16)
        np.save(os.path.join(folder, f"{dataset_id}_{k:06d}"), arr)
