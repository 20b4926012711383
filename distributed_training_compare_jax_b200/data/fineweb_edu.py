"""Data pipeline (component C3; reference: data/fineweb_edu.py:8-39).

Same two entry points as the reference — ``get_tokenizer()`` and ``get_batch_iterator(batch_size, seq_len)``
yielding ``np.int32 [batch, seq_len]`` chunks of a contiguous token stream (the trainer asks for
``seq_len = max_seq_len + 1`` and splits x = row[:-1], y = row[1:], reference train/train.py:55,66-67).

The GPU boxes have no network, so by default the stream is synthetic but FineWeb-edu-*shaped*: GPT-2 vocabulary
(50257 ids + the `<pad>` id 50257 that is never emitted, reference data/fineweb_edu.py:10-11), a Zipf unigram
distribution and a learnable bigram component so that the loss decreases like on text.  If a local HF cache
holds the GPT-2 tokenizer and the dataset, ``synthetic=False`` streams the real thing exactly like the reference.

Differences by design: batches are produced by a background thread into a bounded
queue (the reference tokenises synchronously inside the timed loop) and each data-parallel rank can take its
own slice of the global batch (``rank`` / ``world``).
"""
from __future__ import annotations

import queue
import threading
from typing import Any, Iterator

import numpy as np

GPT2_VOCAB = 50257
PAD_ID = GPT2_VOCAB          # `<pad>` added on top of GPT-2's vocabulary
VOCAB_SIZE = GPT2_VOCAB + 1  # len(tokenizer) in the reference (main.py:17-18) = 50258


class OfflineTokenizer:
    """Stand-in used when the GPT-2 tokenizer files are not available locally: it only knows its size
    (which is all ``main.py`` needs) and a byte-level fallback encoding for smoke tests."""

    pad_token = "<pad>"
    pad_token_id = PAD_ID

    def __len__(self) -> int:
        return VOCAB_SIZE

    def encode(self, text: str) -> list[int]:
        return [b % GPT2_VOCAB for b in text.encode("utf-8")]


def get_tokenizer() -> Any:
    """GPT-2 tokenizer + dedicated ``<pad>`` token when cached locally, else :class:`OfflineTokenizer`."""
    try:
        import os

        os.environ.setdefault("HF_HUB_OFFLINE", "1")
        from transformers import AutoTokenizer

        tok = AutoTokenizer.from_pretrained("gpt2", local_files_only=True)
        tok.add_special_tokens({"pad_token": "<pad>"})
        return tok
    except Exception:
        return OfflineTokenizer()


class SyntheticTokenStream:
    """Deterministic FineWeb-edu-shaped token stream.

    token[t] = bigram_map[base[t-1]] with probability ``bigram_p`` else base[t], where base[] are i.i.d. draws
    from a Zipf law over the 50257 real GPT-2 ids mapped through a fixed permutation.  Fully vectorised."""

    def __init__(self, seed: int = 0, vocab: int = GPT2_VOCAB, zipf_a: float = 1.05, bigram_p: float = 0.5):
        self.rng = np.random.default_rng(seed)
        setup = np.random.default_rng(12345)  # stream structure is independent of the data seed
        ranks = np.arange(vocab, dtype=np.float64)
        p = 1.0 / np.power(ranks + 2.7, zipf_a)
        self.cdf = np.cumsum(p / p.sum())
        self.rank_to_id = setup.permutation(vocab).astype(np.int32)
        self.bigram_map = setup.permutation(vocab).astype(np.int32)
        self.bigram_p = bigram_p
        self._prev_base = np.int32(0)

    def take(self, n: int) -> np.ndarray:
        u = self.rng.random(n)
        base = self.rank_to_id[np.minimum(np.searchsorted(self.cdf, u), len(self.cdf) - 1)]
        follow = self.rng.random(n) < self.bigram_p
        prev = np.empty(n, dtype=np.int32)
        prev[0] = self._prev_base
        prev[1:] = base[:-1]
        self._prev_base = base[-1]
        return np.where(follow, self.bigram_map[prev], base).astype(np.int32)


def _synthetic_batches(batch_size: int, seq_len: int, seed: int, rank: int, world: int) -> Iterator[np.ndarray]:
    stream = SyntheticTokenStream(seed)
    per_rank = batch_size // world
    while True:
        chunk = stream.take(batch_size * seq_len).reshape(batch_size, seq_len)
        yield np.ascontiguousarray(chunk[rank * per_rank:(rank + 1) * per_rank]) if world > 1 else chunk


def _hf_batches(batch_size: int, seq_len: int, rank: int, world: int) -> Iterator[np.ndarray]:
    from datasets import load_dataset  # needs network or a populated cache

    tokenizer = get_tokenizer()
    ds = load_dataset("HuggingFaceFW/fineweb-edu", split="train", streaming=True)
    chunk = batch_size * seq_len
    buf = np.empty(0, dtype=np.int32)
    per_rank = batch_size // world
    for item in ds:
        buf = np.concatenate([buf, np.asarray(tokenizer.encode(item["text"]), dtype=np.int32)])
        while buf.size >= chunk:
            out, buf = buf[:chunk].reshape(batch_size, seq_len), buf[chunk:]
            yield np.ascontiguousarray(out[rank * per_rank:(rank + 1) * per_rank]) if world > 1 else out


class _Prefetcher:
    """Runs a batch generator in a daemon thread with a bounded queue (keeps data work off the step loop)."""

    def __init__(self, gen: Iterator[np.ndarray], depth: int = 4):
        self._q: queue.Queue = queue.Queue(maxsize=depth)
        self._gen = gen
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def _run(self) -> None:
        try:
            for b in self._gen:
                self._q.put(b)
            self._q.put(StopIteration)
        except BaseException as e:  # surface loader failures on the consumer side
            self._q.put(e)

    def __iter__(self):
        return self

    def __next__(self) -> np.ndarray:
        item = self._q.get()
        if item is StopIteration:
            raise StopIteration
        if isinstance(item, BaseException):
            raise item
        return item


def _skip(gen: Iterator[np.ndarray], n: int) -> Iterator[np.ndarray]:
    for _ in range(n):
        next(gen)
    yield from gen


def get_batch_iterator(batch_size: int, seq_len: int, *, seed: int = 0, rank: int = 0, world: int = 1,
                       synthetic: bool = True, prefetch: int = 4, skip_batches: int = 0) -> Any:
    """Infinite iterator of ``np.int32 [batch_size // world, seq_len]`` batches (``[batch_size, seq_len]`` when
    ``world == 1``).  All ranks walk the same global stream, so DP/TP/PP see identical global batches.
    ``skip_batches``: global batches already consumed by the run being resumed (the stream is deterministic, so the
    resumed run continues with the batch the interrupted one would have seen next)."""
    if world > 1 and batch_size % world:
        raise ValueError(f"batch {batch_size} not divisible by {world} data-parallel ranks")
    gen = _synthetic_batches(batch_size, seq_len, seed, rank, world) if synthetic \
        else _hf_batches(batch_size, seq_len, rank, world)
    if skip_batches > 0:
        gen = _skip(gen, skip_batches)
    return _Prefetcher(gen, prefetch) if prefetch > 0 else gen
