"""Row sharding / gather logic of the launcher layer, world_size 2 over gloo on CPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import ROOT


def test_shard_ranges_partition_rows(pkg):
    sr = pkg.shard.shard_range
    for n in (0, 1, 3, 4, 5, 1023, 1024, 1025, 10_000_001):
        for world in (1, 2, 3, 4, 8):
            spans = [sr(n, g, world) for g in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all(b % 4 == 0 for b, _ in spans if b < n)
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 4 or n < 4 * world
    with pytest.raises(ValueError):
        sr(10, 2, 2)


def test_rank_cpu_sets_partition_the_local_cores(pkg):
    rc = pkg.shard.rank_cpus
    allowed = set(range(32))
    sets = [rc(r, 8, allowed=allowed) for r in range(8)]
    assert all(len(s) == 4 for s in sets) and set().union(*sets) == allowed          # one NUMA node: split evenly
    assert sum(len(s) for s in sets) == 32
    near = set(range(16, 32))
    sets = [rc(r, 4, allowed=allowed, gpu_local_cpus=near) for r in range(4)]
    assert set().union(*sets) == near and all(len(s) == 4 for s in sets)             # only the GPU's own cores
    assert rc(0, 1, allowed={3, 5}) == {3, 5}
    assert rc(5, 8, allowed={0, 1}) <= {0, 1} and rc(5, 8, allowed={0, 1})            # more ranks than cores: still a core
    assert pkg.shard._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}


def _worker(rank, world, port, n_total, tmp):
    sys.path.insert(0, ROOT)
    import importlib
    pkg = importlib.import_module("machine-learning-replications_b200")
    from oracle import stack_oracle as so            # stands in for the GPU scorer in this CPU-only test
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    r, lr, w = pkg.shard.init_process_group(backend="gloo")
    assert (r, w) == (rank, world)
    blob = open(os.path.join(ROOT, "tests", "golden", "published_model.blob"), "rb").read()
    model = pkg.export.read_blob(blob)
    specs = pkg.synth.schema_specs()
    b, e = pkg.shard.shard_range(n_total, rank, world)
    X = pkg.synth.host_rows(specs, 2020, b, e - b).astype(np.float64)
    local = torch.from_numpy(so.predict_proba(model, X).astype(np.float32))
    full = pkg.shard.gather_rows(local, n_total, world, rank)
    t = pkg.shard.max_over_ranks(float(rank + 1), world)
    assert t == float(world)
    np.save(os.path.join(tmp, f"gathered_{rank}.npy"), full.numpy())
    torch.distributed.destroy_process_group()


def test_two_rank_sharded_scoring_matches_single_rank(pkg, published, tmp_path):
    from oracle import stack_oracle as so
    n_total = 1001
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, n_total, str(tmp_path)), nprocs=2, join=True)
    X = pkg.synth.host_rows(published["specs"], 2020, 0, n_total).astype(np.float64)
    want = so.predict_proba(published["arrays"], X).astype(np.float32)
    for rank in range(2):
        got = np.load(tmp_path / f"gathered_{rank}.npy")
        assert got.shape == (n_total, 2) and np.array_equal(got, want)      # byte-identical for any shard count
