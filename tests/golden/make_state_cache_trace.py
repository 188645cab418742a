"""Records tests/golden/state_cache_trace.json: the observable trace (every check() result and the LRU key order after every
op) of seeded operation sequences replayed on the UNMODIFIED reference SimpleStateCache (chirrup/utils/state_cache.py, loaded by
path from /root/reference; pure Python).  Build container only; tests/test_state_cache_cpu.py compares PrefixStateCache to it."""
import hashlib, importlib.util, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_state_cache_cpu import _replay
spec = importlib.util.spec_from_file_location("ref_state_cache", "/root/reference/chirrup/utils/state_cache.py")
R = importlib.util.module_from_spec(spec); spec.loader.exec_module(R)
out = {}
for seed, cap in [(0, 3), (1, 8), (2, 25)]:
    t = _replay(R.SimpleStateCache(cap), seed)
    out[f"{seed}_{cap}"] = {"sha256": hashlib.sha256(json.dumps(t).encode()).hexdigest(), "len": len(t)}
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "state_cache_trace.json"), "w"), indent=1)
print(out)
