"""In-stream peer-memory exchange of the adapter-gradient bucket (csrc/p2p.cuh) for one NVSwitch box.

The reference leaves its one exchange step -- averaging the LoRA gradients across data-parallel
ranks -- to torch DDP over gloo (finetune_lora_distilgpt2.py:142; launch at
sbatch/run_finetune_distilgpt2.sbatch).  Here it is one kernel in the compute stream: every rank
copies its bucket into a buffer all peers have mapped (CUDA IPC), flags travel over NVLink, and each
rank sums the peers' copies in rank order, so all replicas hold bit-identical averages.

`PeerBucketExchange.allreduce_avg_` is the primitive; `ddp_hook` plugs it into torch DDP
(`register_comm_hook`) in place of the NCCL all-reduce.  torch.distributed is used only to hand the
64-byte IPC handles around at start-up.  No CPU path: CUDA tensors only.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.distributed as dist

from . import binding

MAX_WORLD = 8


class PeerBucketExchange:
    def __init__(self, capacity_floats: int, group: Optional[dist.ProcessGroup] = None):
        if not dist.is_initialized():
            raise RuntimeError("PeerBucketExchange needs an initialised process group (one process per GPU)")
        if not torch.cuda.is_available():
            raise RuntimeError("PeerBucketExchange needs CUDA devices (no CPU fallback)")
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > MAX_WORLD:
            raise RuntimeError(f"PeerBucketExchange spans one box: at most {MAX_WORLD} ranks, got {self.world}")
        self.capacity = int(capacity_floats)
        self._lib = binding.lib()
        self._own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        binding._check(self._lib.lora_p2p_alloc(self.capacity, ctypes.byref(self._own), handle), "lora_p2p_alloc")
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw, group=group)
        self._syms = (ctypes.c_void_p * self.world)()
        self._opened = []
        for q in range(self.world):
            if q == self.rank:
                self._syms[q] = self._own
            else:
                peer = ctypes.c_void_p()
                binding._check(self._lib.lora_p2p_open(handles[q], ctypes.byref(peer)), "lora_p2p_open")
                self._syms[q] = peer
                self._opened.append(peer)
        dist.barrier(group=group)  # nobody frees or reuses anything before everybody has mapped everything

    def allreduce_avg_(self, t: torch.Tensor, mode: int = 0) -> torch.Tensor:
        """t <- mean over ranks, in place, enqueued on the current stream (fp32, contiguous, CUDA)."""
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError("allreduce_avg_ needs a contiguous fp32 CUDA tensor")
        rc = self._lib.lora_bucket_allreduce_avg(ctypes.c_void_p(t.data_ptr()), t.numel(), self._syms, self.rank, self.world,
                                                 self.capacity, mode,
                                                 ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        binding._check(rc, "lora_bucket_allreduce_avg")
        return t

    def ddp_hook(self, state, bucket):
        """torch DDP communication hook: averages the bucket in the backward's own stream."""
        buf = bucket.buffer()
        self.allreduce_avg_(buf)
        fut = torch.futures.Future()
        fut.set_result(buf)
        return fut

    def close(self):
        if self._own is None:
            return
        torch.cuda.synchronize()
        dist.barrier(group=self.group)  # every rank's last exchange has finished reading
        for p in self._opened:
            self._lib.lora_p2p_close(p)
        self._lib.lora_p2p_free(self._own)
        self._own, self._opened = None, []
