"""Row sharding across the GPUs of one box: one process per GPU, no data-path collective.

Rows of the patient matrix are independent, so rank ``g`` of ``G`` scores the contiguous range
``shard_range(n, g, G)`` with its own copy of the (tiny) model and nothing is exchanged while scoring.  The only
collective is the optional gather of the ``[n_g, 2]`` float32 outputs (``gather_rows``), which runs over NCCL on GPUs
and over gloo in the CPU tests.  Synthetic rows are a pure function of the *global* row index, so any shard count
produces byte-identical results for the same rows.
"""
from __future__ import annotations

import os


def shard_range(n_total: int, rank: int, world: int):
    """Contiguous ``[begin, end)`` of rank's rows; sizes differ by at most 4 and every begin is a multiple of 4 (so a
    shard of a larger column-major matrix keeps the 16-byte alignment the kernels need)."""
    if world < 1 or not (0 <= rank < world) or n_total < 0:
        raise ValueError("bad shard request")
    quads = (n_total + 3) // 4
    q0 = (quads * rank) // world
    q1 = (quads * (rank + 1)) // world
    return min(4 * q0, n_total), min(4 * q1, n_total)


def dist_env():
    """(rank, local_rank, world) from the torchrun environment; (0, 0, 1) when launched plainly."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def rank_cpus(local_rank: int, local_world: int, allowed=None, gpu_local_cpus=None):
    """CPU set for this rank's host threads (the chunk pipeline of ``hfe_predict_host`` narrows rows on host threads and
    copies out of pinned memory, so eight ranks must not all sit on the same cores): the CPUs local to the rank's GPU
    (its PCIe root / NUMA node, from sysfs) intersected with what the process may use; ranks that share such a set split
    it evenly.  Pure function of its arguments (tested on CPU)."""
    allowed = sorted(allowed if allowed is not None else os.sched_getaffinity(0))
    pool = sorted(set(allowed) & set(gpu_local_cpus)) if gpu_local_cpus else allowed
    if not pool:
        pool = allowed
    per = max(len(pool) // max(local_world, 1), 1)
    start = (local_rank * per) % len(pool)
    mine = pool[start:start + per]
    return set(mine) if mine else set(pool)


def bind_rank_to_gpu_cpus(local_rank: int, local_world: int):
    """Set this process's CPU affinity with ``rank_cpus``; returns the set (or None where sysfs / affinity calls are not
    available).  Pinned buffers allocated afterwards are first touched from these cores."""
    try:
        import torch
        local = None
        if torch.cuda.is_available():
            bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
            dom = getattr(torch.cuda.get_device_properties(local_rank), "pci_domain_id", 0)
            dev = getattr(torch.cuda.get_device_properties(local_rank), "pci_device_id", 0)
            path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/local_cpulist"
            if os.path.exists(path):
                local = _parse_cpulist(open(path).read())
        cpus = rank_cpus(local_rank, local_world, gpu_local_cpus=local)
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def init_process_group(backend=None, bind_cpus=True):
    """Join the torchrun rendezvous (MASTER_ADDR/MASTER_PORT from the environment).  Returns (rank, local_rank, world).
    With more than one rank each process is bound to its share of the CPUs local to its GPU (``bind_rank_to_gpu_cpus``)."""
    import torch
    import torch.distributed as dist

    rank, local_rank, world = dist_env()
    if world > 1 and bind_cpus:
        bind_rank_to_gpu_cpus(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            # NCCL writes its banner ("NCCL version ...") to stdout when NCCL_DEBUG is set; stdout belongs to the caller
            # (bench.py prints exactly one JSON line there), so send NCCL's own output to stderr unless told otherwise
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            torch.cuda.set_device(local_rank)
            dist.init_process_group(backend=backend, rank=rank, world_size=world,
                                    device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local_rank, world


def gather_rows(local, n_total: int, world: int, rank: int):
    """All-gather per-rank row blocks ``[n_g, k]`` (ragged by at most 4 rows) into ``[n_total, k]`` on every rank."""
    import torch
    import torch.distributed as dist

    if world == 1:
        return local
    sizes = [shard_range(n_total, g, world) for g in range(world)]
    longest = max(e - b for b, e in sizes)
    padded = torch.zeros((longest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded)
    return torch.cat([parts[g][: e - b] for g, (b, e) in enumerate(sizes)], dim=0)


def max_over_ranks(value: float, world: int, device="cpu") -> float:
    import torch
    import torch.distributed as dist

    if world == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def score_shard(clf, specs, seed: int, n_total: int, rank: int, world: int, out=None):
    """Generate this rank's rows on its GPU and score them (device-resident in, device-resident out).
    Returns ``(x, out, (begin, end))``."""
    from .facade import synth_rows_device

    begin, end = shard_range(n_total, rank, world)
    x = synth_rows_device(specs, seed, begin, end - begin, device=clf.device)
    out = clf.predict_proba_device(x, n=end - begin, out=out)
    return x, out, (begin, end)
