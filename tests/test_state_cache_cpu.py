"""CPU tier: trie / LRU semantics of the prefix state cache (integer keys, exact lookups) — the behaviours of the reference's
SimpleStateCache (chirrup/utils/state_cache.py:58-136): longest cached prefix, LRU eviction, refresh on hit."""
import torch

from chirrup_b200.state_cache import StateCache


def st(v):
    return [torch.full((1,), float(v))]


def test_longest_prefix_and_max_len():
    c = StateCache(max_gpu_entries=8)
    c.put([1, 2, 3], st(3))
    c.put([1, 2, 3, 4, 5], st(5))
    c.put([9], st(9))
    assert c.get([1, 2, 3, 4, 5, 6])[0] == 5
    assert c.get([1, 2, 3, 4])[0] == 3
    assert c.get([1, 2, 3, 4, 5], max_len=4)[0] == 3          # must leave at least one token to run
    assert c.get([2, 2, 3]) == (0, None)
    assert c.get([9, 9])[1][0].item() == 9.0
    assert c.hits == 4 and c.misses == 1


def test_lru_eviction_refresh_and_trie_pruning():
    c = StateCache(max_gpu_entries=2)
    c.put([1, 1], st(1))
    c.put([2, 2], st(2))
    assert c.get([1, 1, 7])[0] == 2                            # refreshes [1,1]
    c.put([3, 3], st(3))                                       # evicts [2,2] (least recently used)
    assert c.get([2, 2, 7]) == (0, None)
    assert c.get([1, 1, 7])[0] == 2 and c.get([3, 3, 7])[0] == 2
    assert 2 not in c.root.children and len(c) == 2            # branch pruned with the entry


def test_host_tier_demotion():
    c = StateCache(max_gpu_entries=1, max_host_entries=1)
    c.put([1], st(1))
    c.put([2], st(2))                                          # [1] demoted to host
    assert c.get([1, 5])[0] == 1 and len(c) == 2
    c.put([3], st(3))                                          # [2] demoted, [1] dropped from host
    assert c.get([1, 5]) == (0, None) and c.get([2, 5])[0] == 1


# ---------------------------------------------------------------------------------------------------------------------------
# PrefixStateCache: the web service's SimpleStateCache semantics, pinned against the UNMODIFIED reference class where it can be
# imported (build container: /root/reference; it is pure Python), and against a recording of the same seeded run elsewhere.
import asyncio
import importlib.util
import json
import os
import random

import pytest

from chirrup_b200.state_cache import PrefixStateCache

_REF = "/root/reference/chirrup/utils/state_cache.py"
_TRACE = os.path.join(os.path.dirname(__file__), "golden", "state_cache_trace.json")


def _ops(seed, n_ops=1500):
    rng = random.Random(seed)
    pool = [tuple(rng.randrange(1, 5) for _ in range(rng.randrange(1, 9))) for _ in range(60)]
    for _ in range(n_ops):
        r = rng.random()
        if r < 0.45:
            yield ("cache", rng.choice(pool))
        elif r < 0.9:
            base = list(rng.choice(pool))
            yield ("check", tuple(base[:rng.randrange(1, len(base) + 1)] + [rng.randrange(1, 5) for _ in range(rng.randrange(0, 4))]))
        else:
            yield ("remove", rng.choice(pool))


def _replay(cache, seed):
    """Observable trace of a seeded op sequence: every check's (remaining, value, cached_len) and the LRU order after each op."""
    trace = []
    for i, (op, key) in enumerate(_ops(seed)):
        if op == "cache":
            cache.cache(key, f"s{i}")
        elif op == "check":
            rest, st, n = cache.check(list(key))
            trace.append([list(rest), st, n])
        else:
            cache.remove(list(key))
        od = cache.LRU_cache.od if hasattr(cache, "LRU_cache") else cache.entries
        trace.append([list(k) for k in od.keys()])
    return trace


@pytest.mark.parametrize("seed,cap", [(0, 3), (1, 8), (2, 25)])
def test_prefix_cache_same_hits_and_evictions_as_reference(seed, cap):
    mine = _replay(PrefixStateCache(cap), seed)
    if os.path.exists(_REF):
        spec = importlib.util.spec_from_file_location("ref_state_cache", _REF)
        R = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(R)
        ref = _replay(R.SimpleStateCache(cap), seed)
        assert mine == ref
    rec = json.load(open(_TRACE))[f"{seed}_{cap}"]        # recorded from the reference by tests/golden/make_state_cache_trace.py
    import hashlib
    assert hashlib.sha256(json.dumps(mine).encode()).hexdigest() == rec["sha256"] and len(mine) == rec["len"]


def test_prefix_cache_strict_prefix_and_value_kept_on_recache():
    c = PrefixStateCache(4)
    c.cache([1, 2, 3, 4], "a")
    assert c.check([1, 2, 3, 4]) == ([1, 2, 3, 4], None, 0)            # an exact match is NOT a hit: one token must remain
    assert c.check([1, 2, 3, 4, 5]) == ([5], "a", 4)
    c.cache([1, 2, 3, 4], "b")
    assert c.check([1, 2, 3, 4, 9])[1] == "a"                          # LRUCache.put keeps the old value (:17-21)
    c.cache((), "x")
    assert len(c) == 1


def test_concurrent_identical_prefills_are_deduplicated():
    """chirrup/utils/state_cache.py:86-136 + app.py:187-190, :293-298: the second identical request waits for the first one's
    ("cache_prefill", ...) event and is then served tokens[-padding:] with the cached state."""
    async def main():
        c = PrefixStateCache(8)
        toks, pad = [5, 6, 7, 8, 9, 10], 2
        first = await c.check_and_wait_prefill(toks, pad)
        assert first == (toks, None, 0)
        waiter = asyncio.ensure_future(c.check_and_wait_prefill(toks, pad))
        other = await c.check_and_wait_prefill([5, 6, 1, 1, 1], pad)   # a different prompt is not held up
        assert other == ([5, 6, 1, 1, 1], None, 0)
        await asyncio.sleep(0.01)
        assert not waiter.done()
        key = c.cache(tuple(toks[:-pad]), "STATE")                      # what the worker's cache_prefill event carries
        assert await c.awake_hang_up_prefills(key) is True
        assert await waiter == (toks[-pad:], "STATE", len(toks) - pad)
        assert await c.check_and_wait_prefill(toks, pad) == (toks[-pad:], "STATE", len(toks) - pad)   # all-hit fast path
        assert await c.awake_hang_up_prefills(key) is False
    asyncio.run(main())
