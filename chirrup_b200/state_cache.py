"""Prefix state cache with the semantics of Chirrup's ``SimpleStateCache`` (``/root/reference/chirrup/utils/state_cache.py:51-215``):
a token trie maps a prompt to its longest cached prefix, entries are LRU-evicted, values are single-request RWKV states
((L,2,1,C), (L,1,H,64,64), (1,)).  B200 redesign: entries live in a GPU-resident tier (device tensors, optionally on another GPU:
admission is then one peer copy, ``state_move.inject_slot``) with an optional host tier below it; the reference keeps every
entry on the host.  Integer logic only (keys are token tuples): lookups must be exact."""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Tuple

import torch


class _Node:
    __slots__ = ("children", "key")

    def __init__(self):
        self.children: Dict[int, "_Node"] = {}
        self.key: Optional[Tuple[int, ...]] = None   # set when a cached state ends exactly here


class StateCache:
    def __init__(self, max_gpu_entries: int = 64, max_host_entries: int = 0):
        self.root = _Node()
        self.gpu: "OrderedDict[Tuple[int, ...], List[torch.Tensor]]" = OrderedDict()
        self.host: "OrderedDict[Tuple[int, ...], List[torch.Tensor]]" = OrderedDict()
        self.max_gpu_entries, self.max_host_entries = max_gpu_entries, max_host_entries
        self.hits = self.misses = 0

    # ---------------------------------------------------------------- trie
    def _insert_key(self, key: Tuple[int, ...]) -> None:
        n = self.root
        for t in key:
            n = n.children.setdefault(t, _Node())
        n.key = key

    def _remove_key(self, key: Tuple[int, ...]) -> None:
        path = [self.root]
        for t in key:
            nxt = path[-1].children.get(t)
            if nxt is None:
                return
            path.append(nxt)
        path[-1].key = None
        for depth in range(len(key), 0, -1):          # prune empty branches
            node = path[depth]
            if node.children or node.key is not None:
                break
            del path[depth - 1].children[key[depth - 1]]

    def longest_prefix(self, tokens: Sequence[int], max_len: Optional[int] = None) -> Optional[Tuple[int, ...]]:
        """Longest cached key that is a prefix of ``tokens`` (and no longer than ``max_len``: a prompt must keep at least one
        token to run, so callers pass len(tokens) - 1)."""
        n, best = self.root, None
        limit = len(tokens) if max_len is None else min(max_len, len(tokens))
        for i in range(limit):
            n = n.children.get(tokens[i])
            if n is None:
                break
            if n.key is not None:
                best = n.key
        return best

    # ---------------------------------------------------------------- entries
    def put(self, tokens: Sequence[int], state: List[torch.Tensor]) -> None:
        key = tuple(int(t) for t in tokens)
        if key in self.host:
            del self.host[key]
        self.gpu[key] = state
        self.gpu.move_to_end(key)
        self._insert_key(key)
        while len(self.gpu) > self.max_gpu_entries:   # LRU: demote to the host tier or drop
            old, st = self.gpu.popitem(last=False)
            if self.max_host_entries > 0:
                self.host[old] = [t.to("cpu") for t in st]
                while len(self.host) > self.max_host_entries:
                    dropped, _ = self.host.popitem(last=False)
                    self._remove_key(dropped)
            else:
                self._remove_key(old)

    def get(self, tokens: Sequence[int], max_len: Optional[int] = None):
        """Returns (prefix_len, state or None).  A hit refreshes the entry's LRU position."""
        key = self.longest_prefix(tokens, max_len)
        if key is None:
            self.misses += 1
            return 0, None
        self.hits += 1
        if key in self.gpu:
            self.gpu.move_to_end(key)
            return len(key), self.gpu[key]
        st = self.host[key]
        self.host.move_to_end(key)
        return len(key), st

    def __len__(self) -> int:
        return len(self.gpu) + len(self.host)


class PrefixStateCache:
    """Drop-in for the web service's ``SimpleStateCache`` (``/root/reference/chirrup/utils/state_cache.py:51-215``) with the same
    method names and observable behaviour — what ``chirrup/web_service/app.py:187-190, :293-298`` calls around
    ``engine_core.completion``:

    * ``check(tokens)`` -> ``(remaining_tokens, state | None, cached_len)``: the longest cached key that is a STRICT prefix of
      ``tokens`` (``:58-84``: a state is only noticed while a further token remains, so a request always keeps one token to run);
      a hit refreshes the entry's LRU position (``:80``);
    * ``cache(tokens, state)``: register under the token tuple; a key already present is refreshed and KEEPS its old value
      (``LRUCache.put`` ``:17-27``); beyond ``max_size`` entries the least recently used one is dropped (``:122-147``);
    * ``remove(tokens)`` (``:152-178``);
    * ``check_and_wait_prefill(tokens, padding)`` / ``awake_hang_up_prefills`` (``:86-136``): identical concurrent prefills are
      parked until the first one has cached its state, then served from the cache.

    Different mechanics: the reference walks a trie with per-node dependency counters; integer keys make that equivalent to
    hashing the prefixes of the few distinct cached LENGTHS, longest first, which is what this class does (no node bookkeeping
    to go stale when a key is re-cached).  Values are whatever the caller caches — here GPU-resident slot states
    (``state_move.extract_slot``), possibly on another GPU than the worker that will admit the request (one peer copy at
    admission); the reference keeps them on the host.  ``tests/test_state_cache_cpu.py`` replays a seeded operation sequence
    against the imported reference class: same hits, same evictions.  One deliberate deviation: the pending-prefill table is
    keyed by the tokens that WILL be cached (``tokens[:len - max(padding, 1)]``, the worker's trigger point,
    ``chirrup/worker.py:411-435``), not by the trie node the reference's walk happens to end on (which is a different node
    when another cached key shares a longer path, so its waiters are never woken)."""

    def __init__(self, max_size: int = 100):
        if max_size <= 0:
            raise ValueError("capacity must be positive")
        self.max_size = max_size
        self.entries: "OrderedDict[Tuple[int, ...], object]" = OrderedDict()
        self._lengths: Dict[int, int] = {}                     # key length -> number of entries of that length
        self._pending: Dict[Tuple[int, ...], "asyncio.Condition"] = {}
        self._lock = None
        self.hits = self.misses = 0

    # ---------------------------------------------------------------- integer part
    def _drop(self, key: Tuple[int, ...]) -> None:
        del self.entries[key]
        n = self._lengths[len(key)] - 1
        if n:
            self._lengths[len(key)] = n
        else:
            del self._lengths[len(key)]

    def check(self, tokens: Sequence[int]):
        tokens = list(tokens)
        for n in sorted((n for n in self._lengths if n < len(tokens)), reverse=True):
            key = tuple(tokens[:n])
            if key in self.entries:
                self.entries.move_to_end(key)
                self.hits += 1
                return tokens[n:], self.entries[key], n
        self.misses += 1
        return tokens, None, 0

    def cache(self, tokens: Sequence[int], state: object) -> Optional[Tuple[int, ...]]:
        """Returns the key (what the reference returns as its trie node: the handle ``awake_hang_up_prefills`` takes)."""
        key = tuple(int(t) for t in tokens)
        if not key:
            return None
        if key in self.entries:
            self.entries.move_to_end(key)
            return key
        self.entries[key] = state
        self._lengths[len(key)] = self._lengths.get(len(key), 0) + 1
        if len(self.entries) > self.max_size:
            self._drop(next(iter(self.entries)))
        return key

    def remove(self, tokens: Sequence[int]) -> None:
        key = tuple(int(t) for t in tokens)
        if key in self.entries and self.entries[key]:           # the reference tests the VALUE's truthiness (:154)
            self._drop(key)

    def __len__(self) -> int:
        return len(self.entries)

    def __contains__(self, tokens) -> bool:
        return tuple(tokens) in self.entries

    # ---------------------------------------------------------------- concurrent identical prefills
    async def check_and_wait_prefill(self, tokens: Sequence[int], cache_prefill_padding: int):
        import asyncio

        tokens = list(tokens)
        if self._lock is None:
            self._lock = asyncio.Lock()
        keep = max(int(cache_prefill_padding), 1)
        async with self._lock:
            rest, state, n = self.check(tokens)
            if n + keep == len(tokens) or len(tokens) <= keep:
                return rest, state, n                           # everything cacheable is cached already
            target = tuple(tokens[:len(tokens) - keep])
            cond = self._pending.get(target)
            if cond is None:                                    # first of its kind: this request does the prefill
                self._pending[target] = asyncio.Condition()
                return rest, state, n
        async with cond:
            await cond.wait()
        if target in self.entries:
            self.entries.move_to_end(target)
            return tokens[len(target):], self.entries[target], len(target)
        return self.check(tokens)                               # the prefill it waited for failed: proceed on its own

    async def awake_hang_up_prefills(self, key) -> bool:
        cond = self._pending.pop(tuple(key), None) if key is not None else None
        if cond is None:
            return False
        async with cond:
            cond.notify_all()
        return True
