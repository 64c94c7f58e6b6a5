"""In-stream peer-memory exchange of the adapter-gradient bucket (csrc/p2p.cuh) for one NVSwitch box.

The reference leaves its one exchange step -- averaging the LoRA gradients across data-parallel
ranks -- to torch DDP over gloo (finetune_lora_distilgpt2.py:142; launched by
"Cluster/Project 1 - Fine Tuning Distilgpt2/submit_distilgpt2_lora.sbatch").  Here it is one kernel in the compute stream: every rank
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
        # Set-up is collective and must fail on ALL ranks or on none: local steps record their error, the
        # collectives that follow always run, and everybody learns whether anybody failed.
        self._own = ctypes.c_void_p()
        self._opened = []
        handle = ctypes.create_string_buffer(64)
        err = None
        if self._lib.lora_p2p_alloc(self.capacity, ctypes.byref(self._own), handle) != 0:
            err = "lora_p2p_alloc: " + binding.last_error()
            self._own = None
        handles = [None] * self.world
        dist.all_gather_object(handles, handle.raw if err is None else None, group=group)
        self._syms = (ctypes.c_void_p * self.world)()
        if err is None and all(h is not None for h in handles):
            for q in range(self.world):
                if q == self.rank:
                    self._syms[q] = self._own
                    continue
                peer = ctypes.c_void_p()
                if self._lib.lora_p2p_open(handles[q], ctypes.byref(peer)) != 0:
                    err = f"lora_p2p_open(rank {q}): " + binding.last_error()
                    break
                self._syms[q] = peer
                self._opened.append(peer)
        elif err is None:
            err = "a peer rank could not allocate its exchange buffer"
        ok = torch.tensor([1 if err is None else 0], device="cuda", dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)  # also: nobody proceeds before everybody has mapped
        if int(ok.item()) == 0:
            for p in self._opened:
                self._lib.lora_p2p_close(p)
            if self._own is not None:
                self._lib.lora_p2p_free(self._own)
            self._own, self._opened = None, []
            raise RuntimeError("peer-memory exchange set-up failed: " + (err or "on another rank"))

    def allreduce_avg_(self, t: torch.Tensor, mode: int = 0) -> torch.Tensor:
        """t <- mean over ranks, in place, enqueued on the current stream (fp32, contiguous, CUDA)."""
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise RuntimeError("allreduce_avg_ needs a contiguous fp32 CUDA tensor")
        with binding.device_guard(t):
            rc = self._lib.lora_bucket_allreduce_avg(ctypes.c_void_p(t.data_ptr()), t.numel(), self._syms, self.rank,
                                                     self.world, self.capacity, mode,
                                                     ctypes.c_void_p(binding._stream(t.device)))
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


def try_peer_exchange(capacity_floats: int, group: Optional[dist.ProcessGroup] = None) -> Optional[PeerBucketExchange]:
    """PeerBucketExchange if EVERY rank could set it up (CUDA IPC allowed, peers P2P-capable), else None on every
    rank -- the caller then keeps the NCCL all-reduce.  Set-up fails collectively, so ranks cannot diverge."""
    import sys

    try:
        return PeerBucketExchange(capacity_floats, group)
    except RuntimeError as e:
        print(f"[lora_b200] rank {dist.get_rank(group) if dist.is_initialized() else 0}: {e}; using NCCL", file=sys.stderr)
        return None
