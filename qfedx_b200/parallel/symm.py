"""Symmetric heap over CUDA IPC: the same buffer on every rank, every peer's copy mapped locally.

Bootstrap is the only place a library collective is used (an object all-gather of 64-byte IPC
handles through torch.distributed); after it, kernels address peer memory directly over NVLink.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch
import torch.distributed as dist

from ..ops import _lib


class _DevicePtr:
    """Wraps a raw device pointer so torch can view it (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class SymmetricBuffer:
    def __init__(self, nbytes: int, rank: int, world: int, device: torch.device, group=None):
        self.nbytes, self.rank, self.world, self.device = nbytes, rank, world, device
        lib = _lib.lib()
        local = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        with torch.cuda.device(device):
            _lib.call("qfx_symm_alloc", ctypes.c_size_t(nbytes), ctypes.byref(local), handle, launches=0)
        self.local_ptr = local.value
        handles: List[bytes] = [None] * world
        if world > 1:
            dist.all_gather_object(handles, bytes(handle), group=group)
        else:
            handles[0] = bytes(handle)
        self.peer_ptrs: List[int] = []
        for r in range(world):
            if r == rank:
                self.peer_ptrs.append(self.local_ptr)
                continue
            p = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(handles[r])
            with torch.cuda.device(device):
                _lib.call("qfx_symm_open", buf, ctypes.byref(p), launches=0)
            self.peer_ptrs.append(p.value)
        self._holder = _DevicePtr(self.local_ptr, nbytes)
        self.local = torch.as_tensor(self._holder, device=device)  # uint8 view of the local copy
        # device-side table of peer base pointers, for kernels that take `T* const*`
        self.ptr_table = torch.tensor(self.peer_ptrs, dtype=torch.int64, device=device)
        if world > 1:
            dist.barrier(group=group)

    def view(self, dtype, offset_bytes: int = 0, numel: int | None = None) -> torch.Tensor:
        t = self.local[offset_bytes:]
        t = t[: numel * torch.empty((), dtype=dtype).element_size()] if numel is not None else t
        return t.view(dtype)

    def peer_table(self, offset_bytes: int = 0) -> torch.Tensor:
        return (self.ptr_table + offset_bytes).contiguous()


class MulticastBuffer:
    """A buffer of ``nbytes`` on every rank bound to ONE multicast object over the job's GPUs (NVLS): ``mc_ptr`` names
    the same offset of all copies at once - ``multimem.ld_reduce`` on it returns the sum over the ranks computed in the
    NVSwitch, ``multimem.st`` writes all copies.  Allocation, handle exchange and the multicast binding are torch's
    symmetric-memory rendezvous (plumbing only); the kernels that use the addresses are ours (csrc/fedavg_fused.cu).
    ``available`` is False where the driver or fabric has no multicast support - callers fall back to peer stores."""

    def __init__(self, nbytes: int, device: torch.device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.nbytes, self.device = nbytes, device
        group = group if group is not None else dist.group.WORLD
        self.local = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.handle = symm_mem.rendezvous(self.local, group)
        self.local_ptr = self.local.data_ptr()
        self.mc_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        ok = torch.tensor([1 if self.mc_ptr else 0], device=device, dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        self.available = bool(int(ok))

    @staticmethod
    def supported(device: torch.device) -> bool:
        try:
            from torch._C._distributed_c10d import _SymmetricMemory
            from torch._C._autograd import DeviceType
            return bool(_SymmetricMemory.has_multicast_support(DeviceType.CUDA, device.index or 0))
        except Exception:
            return False
