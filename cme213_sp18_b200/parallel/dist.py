"""Process bootstrap and peer-mapped (symmetric) device memory.

The reference initialises MPI, binds rank -> GPU with cudaSetDevice(rank) and runs every
collective on host buffers (ref fpcode/main.cpp:37-50, fpcode/neural_network.cpp:407-411).
Here one process drives one GPU (torchrun), `torch.distributed` is used only for the bootstrap
and for the NCCL baseline path, and the hot-path buffers are symmetric allocations whose peer
addresses (and NVSwitch multicast address, when available) are handed to this repo's kernels.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class DistInfo:
    rank: int = 0
    world: int = 1
    local_rank: int = 0
    backend: str = "none"
    device: torch.device = torch.device("cpu")

    @property
    def is_main(self) -> bool:
        return self.rank == 0


def env_world() -> tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def init_distributed(device: str | None = None) -> DistInfo:
    """Reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* from the environment (torchrun).  With a
    GPU: binds cuda:LOCAL_RANK and uses NCCL; without: gloo on the CPU (the multi-process tests)."""
    rank, world, local = env_world()
    use_cuda = torch.cuda.is_available() if device is None else device.startswith("cuda")
    if use_cuda:
        ngpu = torch.cuda.device_count()
        if ngpu < 1:
            raise RuntimeError("CUDA requested but no device is visible")
        if local >= ngpu:
            # the reference only warned here (ref fpcode/main.cpp:45-48); sharing a GPU between
            # ranks that spin on each other's flags can deadlock, so refuse
            raise RuntimeError(f"rank {rank}: LOCAL_RANK {local} >= visible GPUs {ngpu}")
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
    else:
        dev = torch.device("cpu")
    backend = "none"
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        backend = "nccl" if use_cuda else "gloo"
        if not dist.is_initialized():
            kw = {"device_id": dev} if use_cuda else {}
            dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return DistInfo(rank, world, local, backend, dev)


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def barrier(info: DistInfo) -> None:
    if info.world > 1:
        if info.backend == "nccl":
            dist.barrier(device_ids=[info.local_rank])
        else:
            dist.barrier()


class SymmetricTensor:
    """A device tensor allocated identically on every rank with the peers' addresses known.

    world == 1: a plain tensor.  world > 1: `torch.distributed._symmetric_memory` allocation +
    rendezvous, which yields the peer-mapped addresses (`peer_ptrs`) and, where the fabric
    supports it, the multicast address (`multicast_ptr`, 0 if unavailable)."""

    def __init__(self, numel: int, dtype: torch.dtype, info: DistInfo, zero: bool = True):
        self.info = info
        self.numel = numel
        self.dtype = dtype
        self.handle = None
        if info.world == 1:
            self.tensor = torch.empty(numel, dtype=dtype, device=info.device)
            self.peer_ptrs = [int(self.tensor.data_ptr())]
            self.multicast_ptr = 0
        else:
            import torch.distributed._symmetric_memory as symm_mem

            self.tensor = symm_mem.empty(numel, dtype=dtype, device=info.device)
            self.handle = symm_mem.rendezvous(self.tensor, dist.group.WORLD)
            self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
            try:
                self.multicast_ptr = int(self.handle.multicast_ptr or 0)
            except Exception:  # noqa: BLE001 - older builds raise when multicast is unsupported
                self.multicast_ptr = 0
        if zero:
            self.tensor.zero_()

    def ptr_table_dev(self) -> torch.Tensor:
        """int64 device tensor holding the peer addresses (for kernels that take a device array)."""
        return torch.tensor(self.peer_ptrs, dtype=torch.int64, device=self.info.device)
