"""decode on two GPUs (13.3B geometry, 512 slots each) while cached state slots (31.5 MiB each) migrate
between the two workers' slot tables by peer-to-peer DMA (chirrup_b200.state_move.migrate_slot -> b200_slot_copy).  The
reference moves a slot through host memory (chirrup/worker.py:427-429, :591).  Reports decode tok/s per GPU alone and under
migration, and the achieved inter-GPU GB/s (alone and under decode).

Usage (2-GPU box):  python tools/bench_migration.py [--geom 13.3B] [--slots 512] [--spare 16] [--steps 30] [--per-step 8]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from chirrup_b200 import ops, state_move


BATCHED = os.environ.get("B200_MIG_UNBATCHED") is None      # state_move.migrate_slots (one C call per slot) vs migrate_slot
N_STREAMS = int(os.environ.get("B200_MIG_STREAMS", "1"))   # slot copies of one direction spread over this many streams


class Replica:
    def __init__(self, dev: int, geom_name: str, slots: int, spare: int):
        self.dev = torch.device(f"cuda:{dev}")
        with torch.cuda.device(self.dev):
            self.geom, self.m = bench.build_model(geom_name, self.dev)
            V = self.geom.vocab_size
            self.B = slots
            self.state = self.m.generate_zero_state(slots + spare)
            g = torch.Generator().manual_seed(7 + dev)
            self.state[1].normal_(0, 0.05)
            self.idx = torch.arange(slots, dtype=torch.int32, device=self.dev)
            self.el = torch.zeros((slots,), dtype=torch.int32, device=self.dev)
            self.tok = torch.randint(1, V, (slots,), generator=g).to(torch.int32).to(self.dev)
            self.logits = torch.empty((slots, V), dtype=torch.float16, device=self.dev)
            self.nxt = torch.empty((slots,), dtype=torch.int32, device=self.dev)
            self.stream = torch.cuda.Stream(self.dev)
            self.mig = torch.cuda.Stream(self.dev)
            self.migs = [self.mig] + [torch.cuda.Stream(self.dev) for _ in range(max(0, N_STREAMS - 1))]
            self.m.graph_decode = False      # this tool captures the whole step (forward + argmax + feedback) itself
            with torch.cuda.stream(self.stream):
                for _ in range(2):
                    self._step()
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph, stream=self.stream):
                    self._step()
            torch.cuda.synchronize(self.dev)

    def _step(self):
        self.m._step(self.tok, self.B, 1, [self.state[0], self.state[1], self.el], False, slot_idx=self.idx, out=self.logits)
        ops.argmax(self.logits, out=self.nxt)
        self.tok.copy_(self.nxt)

    def replay(self):
        with torch.cuda.device(self.dev), torch.cuda.stream(self.stream):
            self.graph.replay()


def sync_all(reps):
    for r in reps:
        torch.cuda.synchronize(r.dev)


def decode_phase(reps, steps, per_step, slots, spare):
    """`steps` decode steps on every replica, the host waiting for each step's ids like a worker loop does; per_step > 0: that
    many slot migrations per direction are enqueued right after each step's launch and run under it.  Returns per-replica
    decode ms, per-replica device ms spent in the migration batches, bytes moved per direction."""
    ev, mig_ms = [], [0.0 for _ in reps]
    for r in reps:
        with torch.cuda.device(r.dev):
            ev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
    sync_all(reps)
    for r, (e0, _) in zip(reps, ev):
        with torch.cuda.device(r.dev):
            e0.record(r.stream)
    moved = 0
    k = 0
    for _ in range(steps):
        for r in reps:
            r.replay()
        batch = []
        if per_step:
            for d, r in enumerate(reps):            # destination d pulls spare slots from the other replica
                src = reps[1 - d]
                with torch.cuda.device(r.dev):
                    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    m0.record(r.mig)
                    for st in r.migs[1:]:
                        st.wait_event(m0)
                    for si, st in enumerate(r.migs):
                        idx = list(range(si, per_step, len(r.migs)))
                        with torch.cuda.stream(st):
                            if BATCHED:
                                b = state_move.migrate_slots(src.state, [slots + ((k + j) % spare) for j in idx], r.state,
                                                             [slots + ((k + j + 1) % spare) for j in idx])
                            else:
                                b = sum(state_move.migrate_slot(src.state, slots + ((k + j) % spare), r.state,
                                                                slots + ((k + j + 1) % spare)) for j in idx)
                        moved += b if d == 0 else 0
                    for st in r.migs[1:]:
                        r.mig.wait_stream(st)
                    m1.record(r.mig)
                    batch.append((m0, m1))
            k += per_step
        for r in reps:
            r.stream.synchronize()                  # the worker reads this step's ids before it launches the next one
        for d, (m0, m1) in enumerate(batch):
            m1.synchronize()
            mig_ms[d] += m0.elapsed_time(m1)
    for r, (_, e1) in zip(reps, ev):
        with torch.cuda.device(r.dev):
            e1.record(r.stream)
    sync_all(reps)
    ms = [e0.elapsed_time(e1) for e0, e1 in ev]
    return ms, mig_ms, moved


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--geom", default="13.3B")
    ap.add_argument("--slots", type=int, default=512)
    ap.add_argument("--spare", type=int, default=16)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--per-step", type=int, default=8)
    a = ap.parse_args()
    if torch.cuda.device_count() < 2:
        print(json.dumps({"unavailable": "needs two GPUs"}))
        return
    reps = [Replica(d, a.geom, a.slots, a.spare) for d in (0, 1)]
    slot_b = state_move.slot_bytes(reps[0].state)
    # migration alone (both directions at once)
    n = 64
    decode_phase(reps, 0, 0, a.slots, a.spare)
    state_move.migrate_slot(reps[0].state, a.slots, reps[1].state, a.slots)      # enables peer access
    state_move.migrate_slot(reps[1].state, a.slots, reps[0].state, a.slots)
    sync_all(reps)
    mev = []
    t0 = time.perf_counter()
    for d, r in enumerate(reps):
        with torch.cuda.device(r.dev):
            m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            m0.record(r.mig)
            for st in r.migs[1:]:
                st.wait_event(m0)
            for si, st in enumerate(r.migs):
                idx = list(range(si, n, len(r.migs)))
                with torch.cuda.stream(st):
                    if BATCHED:
                        state_move.migrate_slots(reps[1 - d].state, [a.slots + i % a.spare for i in idx], r.state,
                                                 [a.slots + (i + 1) % a.spare for i in idx])
                    else:
                        for i in idx:
                            state_move.migrate_slot(reps[1 - d].state, a.slots + i % a.spare, r.state, a.slots + (i + 1) % a.spare)
            for st in r.migs[1:]:
                r.mig.wait_stream(st)
            m1.record(r.mig)
            mev.append((m0, m1))
    t_enq = time.perf_counter() - t0
    sync_all(reps)
    alone_ms = [m0.elapsed_time(m1) for m0, m1 in mev]
    alone_gbps = [n * slot_b / (m * 1e-3) * 1e-9 for m in alone_ms]
    # the link's own ceiling for this message size: contiguous 32 MiB peer copies (copy engine), one way and both ways at once
    def ceiling(both):
        bufs = []
        for d in (0, 1):
            with torch.cuda.device(reps[d].dev):
                bufs.append((torch.empty(32 << 20, dtype=torch.uint8, device=reps[d].dev), torch.empty(32 << 20, dtype=torch.uint8, device=reps[d].dev)))
        sync_all(reps)
        evs = []
        for d in ((0, 1) if both else (0,)):
            r = reps[d]
            with torch.cuda.device(r.dev), torch.cuda.stream(r.mig):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(r.mig)
                for _ in range(32):
                    bufs[d][0].copy_(bufs[1 - d][1], non_blocking=True)      # destination pulls from the peer
                e1.record(r.mig)
                evs.append((e0, e1))
        sync_all(reps)
        return [32 * (32 << 20) / (e0.elapsed_time(e1) * 1e-3) * 1e-9 for e0, e1 in evs]

    ceil_one, ceil_both = ceiling(False), ceiling(True)
    # decode alone, then decode under migration
    decode_phase(reps, 5, 0, a.slots, a.spare)
    ms0, _, _ = decode_phase(reps, a.steps, 0, a.slots, a.spare)
    ms1, mms, moved = decode_phase(reps, a.steps, a.per_step, a.slots, a.spare)
    out = {
        "config": f"{a.geom} geometry, {a.slots} decode slots per GPU (+{a.spare} spare), 2 GPUs, greedy decode from a CUDA graph per GPU; "
                  f"migration = {a.per_step} slots per decode step per direction, {slot_b / 2**20:.1f} MiB each, b200_slot_copy (peer DMA)",
        "p2p_access": bool(torch.cuda.can_device_access_peer(0, 1)), "streams_per_direction": N_STREAMS, "host_path": "migrate_slots (b200_slot_move)" if BATCHED else "migrate_slot",
        "slot_copy_mode": "copy engine (cudaMemcpy2DAsync)",
        "slot_MiB": slot_b / 2**20,
        "migration_alone_GBps_per_direction": alone_gbps,
        "link_ceiling_contiguous_32MiB_copies_GBps": {"one_way": ceil_one, "both_ways_at_once_per_direction": ceil_both},
        "migration_alone_us_per_slot_device": [m / n * 1e3 for m in alone_ms],
        "migration_host_enqueue_us_per_slot": t_enq / (2 * n) * 1e6,
        "decode_alone_ms_per_step": [m / a.steps for m in ms0],
        "decode_alone_tok_per_s_per_gpu": [a.slots * a.steps / (m * 1e-3) for m in ms0],
        "decode_under_migration_ms_per_step": [m / a.steps for m in ms1],
        "decode_under_migration_tok_per_s_per_gpu": [a.slots * a.steps / (m * 1e-3) for m in ms1],
        "migration_under_decode_GBps_per_direction": [moved / (m * 1e-3) * 1e-9 for m in mms],
        "slots_migrated_per_direction": a.steps * a.per_step,
    }
    print(json.dumps(out))


if __name__ == "__main__":
    main()
