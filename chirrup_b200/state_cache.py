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
