"""Multi-GPU partition of the hot path: one process per GPU, studies sharded over ranks, no data-path collective.

The reference processes its inputs one after the other in one process (reference ts2d/main.py:80-83).  Studies are
independent, so rank r of W takes studies r, r+W, r+2W, ... (all five group models are resident on every GPU); the only
communication is the bookkeeping below (a max of device times, a sum of counts, and an ordered gather of small
per-study records), which works over NCCL on the GPU box and over gloo in the CPU tests.
"""
from __future__ import annotations

import os
from typing import Any, List, Sequence, Tuple


def rank_world() -> Tuple[int, int, int]:
    """(rank, world size, local rank) from the torchrun environment; (0, 1, 0) when launched plainly"""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """indices of the studies rank `rank` processes: i with i % world == rank (round robin keeps ranks balanced to +-1)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"invalid rank {rank} of world size {world}")
    return list(range(rank, n_items, world))


def shard(items: Sequence[Any], rank: int, world: int) -> List[Any]:
    return [items[i] for i in shard_indices(len(items), rank, world)]


def _parse_cpulist(text: str) -> List[int]:
    """'0-3,8,10-11' -> [0, 1, 2, 3, 8, 10, 11]"""
    cpus: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def gpu_numa_cpus(pci_bus_id: str, sysfs: str = "/sys") -> List[int]:
    """CPUs of the NUMA node the GPU with this PCI address ('0000:1b:00.0') hangs off; [] when sysfs does not say"""
    try:
        bdf = pci_bus_id.lower()
        if len(bdf.split(":")) == 2:
            bdf = "0000:" + bdf
        with open(os.path.join(sysfs, "bus/pci/devices", bdf, "numa_node")) as f:
            node = int(f.read().strip())
        if node < 0:
            return []
        with open(os.path.join(sysfs, "devices/system/node", f"node{node}", "cpulist")) as f:
            return _parse_cpulist(f.read())
    except (OSError, ValueError):
        return []


def bind_to_gpu(local_rank: int, n_local: int, pci_bus_id: str = "", sysfs: str = "/sys") -> List[int]:
    """Pin this process (and the pinned host buffers it allocates afterwards: first touch) to the CPUs next to its GPU.
    With 8 ranks streaming 524 MB volumes each, buffers that all sit on one socket make the other socket's GPUs pull their
    input over the inter-socket link.  Uses the GPU's NUMA node from sysfs; when that is unknown, an even split of the
    allowed CPUs by local rank (GPUs are enumerated socket by socket on the usual 8-GPU boards).  Returns the CPU list in
    effect (empty = left alone).  TS2D_BIND=0 disables it."""
    if os.environ.get("TS2D_BIND", "1") == "0" or n_local <= 1 or not hasattr(os, "sched_setaffinity"):
        return []
    try:
        allowed = sorted(os.sched_getaffinity(0))
        cpus = [c for c in gpu_numa_cpus(pci_bus_id, sysfs) if c in allowed] if pci_bus_id else []
        if not cpus:
            per = len(allowed) // n_local
            if per < 1:
                return []
            cpus = allowed[local_rank * per:(local_rank + 1) * per]
        os.sched_setaffinity(0, cpus)
        return cpus
    except OSError:
        return []


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def _device():
    import torch
    d = _dist()
    if d is not None and d.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_max(x: float) -> float:
    """max over ranks (device time of a step = the slowest rank)"""
    import torch
    d = _dist()
    if d is None:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    d.all_reduce(t, op=d.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(x: float) -> float:
    import torch
    d = _dist()
    if d is None:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    d.all_reduce(t, op=d.ReduceOp.SUM)
    return float(t.item())


def gather_in_order(local_records: Sequence[Any], n_items: int) -> List[Any]:
    """inverse of `shard`: every rank contributes the records of its studies, every rank gets the list in study order"""
    d = _dist()
    if d is None:
        return list(local_records)
    world, rank = d.get_world_size(), d.get_rank()
    if len(local_records) != len(shard_indices(n_items, rank, world)):
        raise ValueError("rank {} holds {} records for {} studies".format(rank, len(local_records), len(shard_indices(n_items, rank, world))))
    parts: List[Any] = [None] * world
    d.all_gather_object(parts, list(local_records))
    out: List[Any] = [None] * n_items
    for r in range(world):
        for k, i in enumerate(shard_indices(n_items, r, world)):
            out[i] = parts[r][k]
    return out


# ------------------------------------------------------------------------------------------------------------------
# Latency mode: ONE study spread over the ranks.  The group models of a study are independent (reference
# ts2d/tool.py:110-112) and share only the projected image, so the home rank projects, broadcasts the 4 MB image, every
# rank runs the group models it owns, and the per-model label masks are gathered on the home rank (the only exchange
# step of the path; NCCL over NVLink on the GPU box, gloo in the CPU tests).
# ------------------------------------------------------------------------------------------------------------------
def model_owner(k: int, world: int, home: int = 0) -> int:
    """rank that runs group model k of a study whose volume lives on rank `home` (round robin starting at home)"""
    if world < 1 or not (0 <= home < world) or k < 0:
        raise ValueError(f"invalid model {k} / home {home} / world {world}")
    return (home + k) % world


def owned_models(n_models: int, rank: int, world: int, home: int = 0) -> List[int]:
    return [k for k in range(n_models) if model_owner(k, world, home) == rank]


def broadcast_image(img, src: int, meta: Sequence[float] = (), header_len: int = 8):
    """rank `src` passes `img` (C, H, W) float32 and up to 5 numbers `meta` (e.g. the pixel spacing); every rank returns
    (img, meta).  Two broadcasts: a fixed-size float64 header (shape + meta), then the pixels."""
    import torch
    d = _dist()
    if d is None:
        return img, [float(v) for v in meta]
    dev = _device()
    rank = d.get_rank()
    n_meta = header_len - 3
    hdr = torch.zeros(header_len, dtype=torch.float64)
    if rank == src:
        if len(meta) > n_meta:
            raise ValueError("too many meta values")
        hdr[:3] = torch.tensor([float(v) for v in img.shape], dtype=torch.float64)
        hdr[3:3 + len(meta)] = torch.tensor([float(v) for v in meta], dtype=torch.float64)
    hdr = hdr.to(dev)
    d.broadcast(hdr, src=src)
    vals = hdr.tolist()
    C, H, W = (int(v) for v in vals[:3])
    if rank != src:
        img = torch.empty((C, H, W), dtype=torch.float32, device=dev)
    else:
        img = img.contiguous()
    d.broadcast(img, src=src)
    return img, vals[3:]


def gather_windows(local: dict, n_models: int, widths: Sequence[int], lead_shape: Sequence[int], home: int, out=None, dtype=None):
    """local: {model index k: tensor [*lead_shape, widths[k]]} for the models this rank owns.  On the home rank returns
    `out` [*lead_shape, sum(widths)] with every model's window in place (model-id order); other ranks return None.
    One batch of point-to-point transfers: sizes are known on every rank, nothing is padded."""
    import torch
    d = _dist()
    offs = [0]
    for w in widths:
        offs.append(offs[-1] + int(w))
    if d is None:
        world, rank = 1, 0
    else:
        world, rank = d.get_world_size(), d.get_rank()
    mine = owned_models(n_models, rank, world, home)
    if sorted(local.keys()) != mine:
        raise ValueError(f"rank {rank} holds windows {sorted(local.keys())}, expected {mine}")
    if rank == home:
        any_t = next(iter(local.values())) if local else None
        if out is None:
            out = torch.zeros((*lead_shape, offs[-1]), dtype=dtype or any_t.dtype, device=any_t.device if any_t is not None else _device())
        ops, bufs = [], {}
        for k in range(n_models):
            o = model_owner(k, world, home)
            if o == home:
                out[..., offs[k]:offs[k + 1]].copy_(local[k])
            else:
                bufs[k] = torch.empty((*lead_shape, int(widths[k])), dtype=out.dtype, device=out.device)
                ops.append(d.P2POp(d.irecv, bufs[k], o))
        if ops:
            for r in d.batch_isend_irecv(ops):
                r.wait()
        for k, b in bufs.items():
            out[..., offs[k]:offs[k + 1]].copy_(b)
        return out
    ops = [d.P2POp(d.isend, local[k].contiguous(), home) for k in mine]
    if ops:
        for r in d.batch_isend_irecv(ops):
            r.wait()
    return None
