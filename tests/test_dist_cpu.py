"""world_size-2 gloo test of the N>1 evaluation path (host logic; no GPU)."""

import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepfakedetection_b200.evaluate import gather_records, shard_range, unpack_records


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, n_total: int, out_dir: str) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        full = torch.rand(n_total, 4, generator=g)  # the records a 1-process run would produce
        lo, hi = shard_range(n_total, rank, world)
        got = gather_records(full[lo:hi].clone())
        assert got.shape == full.shape and torch.equal(got, full), "gather must restore dataset order"
        torch.save(got, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_gather_records_world2_ragged(tmp_path):
    n_total = 11  # ragged: shards of 6 and 5
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_total, str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(a, b)
    probs, preds, targets = unpack_records(a)
    assert probs.shape == (n_total, 2) and preds.dtype == torch.int64 and targets.dtype == torch.int64


def _worker_empty(rank: int, world: int, port: int) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(4.0).view(1, 4)
        lo, hi = shard_range(1, rank, world)  # rank 1 gets nothing
        got = gather_records(full[lo:hi].clone())
        assert torch.equal(got, full)
    finally:
        dist.destroy_process_group()


def test_gather_records_with_an_empty_shard():
    mp.spawn(_worker_empty, args=(2, _free_port()), nprocs=2, join=True)
