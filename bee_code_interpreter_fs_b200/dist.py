"""torch.distributed plumbing for multi-GPU runs: rendezvous, barrier, max-over-ranks,
and gathering per-rank digests.  The data path has no collective (objects are
independent); these helpers only synchronise timing and collect 32-byte results.
Works with backend "nccl" (one rank per GPU) and "gloo" (CPU tests)."""
from __future__ import annotations

import os

import numpy as np


def env_world() -> tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment, (0, 0, 1) without it."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)))


_ORDER_FILE = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "h2d_gpu_order.json")


def gpu_order(device_count: int) -> list[int]:
    """The order in which the GPUs of this node are handed to ranks (and to MultiGpuHasher).

    Objects are independent, so ranks never talk to each other and any GPU serves any rank; what
    differs between GPUs is the host uplink they hang off.  With N < all GPUs in use, taking them
    in an order measured to spread over distinct uplinks (tools/h2d_matrix.cu, "greedy" section,
    kept in profiles/h2d_gpu_order.json) gives the N concurrent host->device streams more of the
    host's bandwidth than GPUs 0..N-1 would.  ``B200CAS_GPU_ORDER=0,4,1,5,...`` overrides it; a
    map measured on a node with a different GPU count is ignored (identity order)."""
    order = None
    env = os.environ.get("B200CAS_GPU_ORDER")
    try:
        if env:
            order = [int(x) for x in env.split(",") if x.strip()]
        else:
            import json

            with open(_ORDER_FILE) as f:
                order = [int(x) for x in json.load(f)["order"]]
    except (OSError, ValueError, KeyError):
        order = None
    if not order or sorted(order) != list(range(device_count)):
        return list(range(device_count))
    return order


def h2d_gate_for(device: int, device_count: int) -> tuple[str, int] | None:
    """(gate name, slots) for the host->device gate of ``device``'s uplink group, or None.

    Ranks of one torchrun launch (same parent process) whose GPUs hang off the same group of the
    measured map share a gate, so that no more than ``gate_slots_per_group`` of them copy from
    host memory at the same time (GpuHasher.set_h2d_gate).  Only meaningful when more ranks than
    slots share a group."""
    try:
        import json

        with open(_ORDER_FILE) as f:
            info = json.load(f)
        groups, slots = info["uplink_groups"], int(info["gate_slots_per_group"])
        if sorted(g for grp in groups for g in grp) != list(range(device_count)):
            return None
        for k, grp in enumerate(groups):
            if device in grp:
                return f"/b200cas-h2d-{os.getppid()}-g{k}", slots
    except (OSError, ValueError, KeyError):
        pass
    return None


def remove_h2d_gate(name: str) -> None:
    """sem_unlink for a gate name (Linux keeps named semaphores as /dev/shm/sem.<name>)."""
    try:
        os.unlink("/dev/shm/sem." + name.lstrip("/"))
    except OSError:
        pass


def gpu_for_rank(local_rank: int, device_count: int) -> int:
    order = gpu_order(device_count)
    return order[local_rank] if local_rank < len(order) else local_rank


def init(backend: str | None = None) -> tuple[int, int, int]:
    """Rendezvous.  Returns (rank, device index for this rank, world size); the device index is
    ``gpu_for_rank(local_rank)`` with the nccl backend and the local rank otherwise."""
    import torch
    import torch.distributed as dist

    rank, local_rank, world = env_world()
    device = local_rank
    if torch.cuda.is_available():
        device = gpu_for_rank(local_rank, torch.cuda.device_count())
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29531")
        backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
        if backend == "nccl":
            torch.cuda.set_device(device)
            dist.init_process_group(backend, device_id=torch.device("cuda", device))
        else:
            dist.init_process_group(backend)
    return rank, device, world


def _device():
    import torch
    import torch.distributed as dist

    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def barrier() -> None:
    import torch.distributed as dist

    if dist.is_initialized():
        dist.barrier()


def max_over_ranks(value: float) -> float:
    import torch
    import torch.distributed as dist

    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float) -> float:
    import torch
    import torch.distributed as dist

    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_digests(local: np.ndarray, counts: list[int]) -> np.ndarray:
    """Concatenate per-rank (k_r, 32) digest arrays in rank order on every rank."""
    import torch
    import torch.distributed as dist

    local = np.ascontiguousarray(local, dtype=np.uint8).reshape(-1, 32)
    if not dist.is_initialized():
        return local
    world = dist.get_world_size()
    if len(counts) != world or counts[dist.get_rank()] != local.shape[0]:
        raise ValueError("counts do not describe the shards")
    dev = _device()
    width = max(counts) if counts else 0
    pad = torch.zeros((max(width, 1), 32), dtype=torch.uint8, device=dev)
    pad[: local.shape[0]] = torch.from_numpy(local).to(dev)
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    return np.concatenate([p[:c].cpu().numpy() for p, c in zip(parts, counts)], axis=0)


def bind_to_gpu_numa_node(local_rank: int) -> dict:
    """Pins this process to the CPUs of the NUMA node its GPU hangs off, so that pinned
    staging is allocated next to the GPU's PCIe root.  No-op (and says so) when the
    platform hides the topology, as most virtual machines do."""
    import torch

    info = {"numa_node": None, "cpus": None}
    try:
        p = torch.cuda.get_device_properties(local_rank)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        info["numa_node"] = node
        if node < 0:
            return info
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["cpus"] = len(allowed)
    except Exception:  # noqa: BLE001 - topology is a hint; any failure means "leave the process alone"
        pass
    return info
