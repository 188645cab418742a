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
