"""Host-side multi-GPU logic on CPU: partition properties and a world_size-2 gloo gather."""
import os
import random

import torch.distributed as dist
import torch.multiprocessing as mp

from replisense_rfq_b200.shard import parse_sharded, partition


def test_partition_properties():
    r = random.Random(0)
    for n in (0, 1, 7, 1000):
        sizes = [r.randint(0, 10 ** 6) for _ in range(n)]
        for w in (1, 2, 4, 8):
            sh = partition(sizes, w)
            assert sorted(i for s in sh for i in s) == list(range(n))
            assert all(s == sorted(s) for s in sh)
            loads = [sum(sizes[i] for i in s) for s in sh]
            if n >= 8 * w:
                assert max(loads) - min(loads) <= max(sizes)   # LPT bound
            assert partition(sizes, w) == sh                     # deterministic


def _fake_parse(docs, kinds):
    return ["%s:%d:%s" % (k, len(d), d[:4].decode()) for d, k in zip(docs, kinds)]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    r = random.Random(1)
    docs = [bytes(r.choice(b"abcd") for _ in range(r.randint(0, 300))) for _ in range(57)]
    kinds = [r.choice(["txt", "csv", "xlsx"]) for _ in docs]
    out = parse_sharded(docs, kinds, _fake_parse, rank, world)
    q.put((rank, out == _fake_parse(docs, kinds)))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_gloo_gather():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + random.randint(0, 2000)
    procs = [ctx.Process(target=_worker, args=(rk, 2, port, q)) for rk in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(0, True), (1, True)]
