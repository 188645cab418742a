"""Debug aid: host / device timeline of consecutive worker steps in the steady state (all slots decoding)."""
import os, queue, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from chirrup_b200.engine import ModelLoadConfig, Task, Worker
from chirrup_b200.rwkv7 import RWKV_x070, make_args
from chirrup_b200.synth import GEOMETRIES, make_checkpoint

geom = GEOMETRIES[sys.argv[1] if len(sys.argv) > 1 else "7.2B"]
dev = torch.device("cuda")
m = RWKV_x070(make_args(), state_dict=make_checkpoint(geom, seed=0, recipe="trained_like", device=dev, dtype=torch.float16))
class Sink:
    def put_nowait(self, x): pass
w = Worker("w0", [0], ModelLoadConfig(vocab_size=geom.vocab_size), queue.Queue(), batch_size=257, model=m)
w.max_prefill_count = 300
w._init_worker()
for i in range(256):
    w.task_queue.put(Task(prefill_tokens=[i + 1, 5, 9], output_queue=Sink(), top_k=1, presence_penalty=0., frequency_penalty=0.,
                          max_tokens=10000, stop_tokens=[]))
for _ in range(20): w.step()
torch.cuda.synchronize()
base = torch.cuda.Event(enable_timing=True); base.record(); torch.cuda.synchronize(); t_base = time.perf_counter()
log = []
orig_step, orig_launch, orig_finish = w.model._step, w._launch_decode, w._finish_decode
def mstep(*a, **k):
    e = torch.cuda.Event(enable_timing=True); e.record(); cur["g0"] = e
    cur["h_step0"] = time.perf_counter(); r = orig_step(*a, **k); cur["h_step1"] = time.perf_counter(); return r
def launch(rows, feed):
    cur["h_launch0"] = time.perf_counter(); rec = orig_launch(rows, feed); cur["h_launch1"] = time.perf_counter()
    e = torch.cuda.Event(enable_timing=True); e.record(); rec["g"] = (cur["g0"], e); return rec
def finish(rec):
    cur["h_wait0"] = time.perf_counter()
    if rec["event"] is not None: rec["event"].synchronize()
    cur["h_wait1"] = time.perf_counter(); r = orig_finish(rec); cur["h_fin1"] = time.perf_counter()
    cur["gpu"] = rec.get("g"); return r
w.model._step, w._launch_decode, w._finish_decode = mstep, launch, finish
for i in range(12):
    cur = {"h_call0": time.perf_counter()}
    w.step()
    cur["h_call1"] = time.perf_counter()
    log.append(cur)
torch.cuda.synchronize()
for c in log:
    f = lambda k: f"{(c[k] - t_base) * 1e3:8.2f}" if k in c else "     n/a"
    if not c.get("gpu"):
        continue
    g0, g1 = c["gpu"]
    print("call", f("h_call0"), "launch", f("h_launch0"), "_step", f("h_step0"), f("h_step1"), "launched", f("h_launch1"), "wait", f("h_wait0"),
          f("h_wait1"), "fin", f("h_fin1"), "ret", f("h_call1"), "| gpu span of the FINISHED step", f"{base.elapsed_time(g0):8.2f} {base.elapsed_time(g1):8.2f}")
