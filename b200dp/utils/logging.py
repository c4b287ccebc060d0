"""Rank-0 logging: the reference's two human-readable lines plus a JSONL stream.

Format strings are the reference's, character for character
(train.py:53-55 and train.py:74-76), including its quirks: the sample counter
is ``batch_idx * len(data)`` (under-counts on a ragged last batch) and the
denominator is the *shard* size.
"""
from __future__ import annotations

import json
import time
from typing import Any, Dict, Optional, TextIO

TRAIN_FMT = "Train Epoch: {} [{}/{} ({:.0f}%)]\tLoss: {:.6f}"
TEST_FMT = "\nTest set on {}: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n"


def train_line(epoch: int, batch_idx: int, batch_len: int, shard_len: int, n_batches: int, loss: float) -> str:
    return TRAIN_FMT.format(epoch, batch_idx * batch_len, shard_len, 100.0 * batch_idx / n_batches, loss)


def test_line(device, test_loss: float, correct: int, shard_len: int) -> str:
    return TEST_FMT.format(device, test_loss, correct, shard_len, 100.0 * correct / shard_len)


class RankLogger:
    """Prints on the main rank only (reference gates on ``device == 0``)."""

    def __init__(self, is_main: bool, jsonl_path: Optional[str] = None, stream: Optional[TextIO] = None):
        self.is_main = is_main
        self.stream = stream
        self._jsonl = open(jsonl_path, "a") if (is_main and jsonl_path) else None

    def print(self, msg: str) -> None:
        if self.is_main:
            print(msg, file=self.stream, flush=True)

    def event(self, **fields: Any) -> None:
        if self._jsonl is not None:
            rec: Dict[str, Any] = {"ts": time.time()}
            rec.update(fields)
            self._jsonl.write(json.dumps(rec) + "\n")
            self._jsonl.flush()

    def close(self) -> None:
        if self._jsonl is not None:
            self._jsonl.close()
            self._jsonl = None
