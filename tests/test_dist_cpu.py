"""CPU tier, world_size 2 over gloo: the N > 1 plumbing of the request-sharded path (chirrup_b200/dist_util.py) — sharding covers
every request once, timing is the max over ranks, throughput aggregates over ranks, and no data-path collective is needed:
each rank steps its own shard with the CPU oracle and the union equals the single-rank result."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from chirrup_b200.dist_util import max_over_ranks, shard_requests, sum_over_ranks


def test_shard_requests_partition():
    for n in (0, 1, 7, 8, 256, 257):
        for world in (1, 2, 3, 8):
            ids = [i for r in range(world) for i in shard_requests(n, r, world)]
            assert ids == list(range(n))
            sizes = [len(shard_requests(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from chirrup_b200.synth import GEOMETRIES, make_checkpoint
    from oracle import rwkv7_oracle as O

    geom = GEOMETRIES["tiny"]
    z = O.transform_checkpoint(make_checkpoint(geom, seed=0, recipe="trained_like"))
    og = O.Geometry(z)
    prompts = torch.randint(1, geom.vocab_size, (6, 4), generator=torch.Generator().manual_seed(3)).tolist()
    mine = list(shard_requests(len(prompts), rank, world))
    toks, _, _ = O.greedy_decode(z, og, [prompts[i] for i in mine], 5)
    t_ms = 10.0 + 5.0 * rank                                   # pretend device time; rank 1 is the slow one
    (t_max,) = max_over_ranks([t_ms])
    (n_tok,) = sum_over_ranks([toks.numel()])
    torch.save({"ids": mine, "toks": toks, "t_max": t_max, "n_tok": n_tok}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_replicas_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert r0["t_max"] == r1["t_max"] == 15.0                   # max over ranks, identical on both
    assert r0["n_tok"] == r1["n_tok"] == 30.0                   # whole-job token count
    assert r0["ids"] + r1["ids"] == list(range(6))
    # union of the shards == the unsharded run (requests are independent: no collective on the data path)
    from chirrup_b200.synth import GEOMETRIES, make_checkpoint
    from oracle import rwkv7_oracle as O

    geom = GEOMETRIES["tiny"]
    z = O.transform_checkpoint(make_checkpoint(geom, seed=0, recipe="trained_like"))
    prompts = torch.randint(1, geom.vocab_size, (6, 4), generator=torch.Generator().manual_seed(3)).tolist()
    full, _, _ = O.greedy_decode(z, O.Geometry(z), prompts, 5)
    assert torch.equal(torch.cat([r0["toks"], r1["toks"]]), full)
