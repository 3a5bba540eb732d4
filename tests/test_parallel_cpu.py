"""The N > 1 host logic on CPU: two `gloo` ranks run the data-parallel plumbing of dualhyp_b200/parallel.py --
the flat fp32 gradient bucket, its single all-reduce per step, the re-homed `.grad` views -- and the utterance sharding
of the inference path (rank r takes utterances r, r + world, ...; no communication, predictions merged by index)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dualhyp_b200.parallel import FlatTrainables, shard_indices


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(7)                                   # same parameters on every rank, like the replicated PEFT set
        shapes = [(10, 64), (1, 1, 32, 1), (16, 64), (64, 16), (3,)]
        params = [torch.nn.Parameter(torch.randn(s)) for s in shapes]
        frozen = torch.nn.Parameter(torch.randn(4, 4), requires_grad=False)
        before = [p.detach().clone() for p in params]
        bucket = FlatTrainables(params + [frozen])
        assert bucket.numel == sum(p.numel() for p in params), "the frozen tensor must stay out of the bucket"
        for p, b in zip(params, before):                       # re-homing keeps the values and makes data / grad views of the flats
            assert torch.equal(p.detach(), b)
            assert p.data.untyped_storage().data_ptr() == bucket.flat.untyped_storage().data_ptr()
            assert p.grad.untyped_storage().data_ptr() == bucket.grad.untyped_storage().data_ptr()
        # a backward that ACCUMULATES into the views (what autograd does) with rank-dependent gradients
        bucket.zero_grad()
        loss = sum(((rank + 1) * (i + 1)) * p.sum() for i, p in enumerate(params))
        loss.backward()
        for i, p in enumerate(params):
            assert p.grad.untyped_storage().data_ptr() == bucket.grad.untyped_storage().data_ptr(), "autograd replaced a .grad view"
            assert torch.all(p.grad == (rank + 1) * (i + 1))
        bucket.allreduce()                                     # ONE collective for the whole trainable set
        total = sum(r + 1 for r in range(world))
        for i, p in enumerate(params):
            assert torch.all(p.grad == total * (i + 1)), (rank, i)
        # a second step must start from zero (zero_grad clears the bucket and keeps the views)
        bucket.zero_grad()
        assert float(bucket.grad.abs().sum()) == 0.0 and all(float(p.grad.abs().sum()) == 0.0 for p in params)
        # inference: utterances sharded by index, predictions gathered on rank 0 by index
        n_utts = 11
        mine = shard_indices(n_utts, rank, world)
        preds = {i: i * i for i in mine}
        gathered = [None] * world
        dist.all_gather_object(gathered, preds)
        merged = {}
        for g in gathered:
            assert not (merged.keys() & g.keys()), "two ranks decoded the same utterance"
            merged.update(g)
        assert sorted(merged) == list(range(n_utts)) and all(merged[i] == i * i for i in merged)
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_flat_bucket_allreduce_and_utterance_sharding_two_gloo_ranks(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_shard_indices_partition():
    for n in (0, 1, 7, 64, 513):
        for world in (1, 2, 4, 8):
            parts = [shard_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
