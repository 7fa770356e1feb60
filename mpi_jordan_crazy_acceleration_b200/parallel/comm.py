"""Process-group plumbing: one process per GPU, ``torch.distributed`` for bootstrap and the few
control-plane collectives the reference does outside its hot loop (allocation / file-error votes,
``main.cpp:371,237,277``; max of the timings, ``:455``; max of the residual, ``:495``).  The per-step
exchanges (pivot min-loc ``:1074``, pivot-row broadcast ``:1097``, row swap ``:1120-1130``) never come
through here on GPUs: they are peer-memory loads/stores issued by the kernels.

Control messages use a gloo group (CPU) so they never serialise with the CUDA stream; the NCCL
group is kept for the in-tree NCCL+cuBLAS baseline and for symmetric-memory rendezvous.
"""
from __future__ import annotations

import os
import pickle
from typing import Any

import torch
import torch.distributed as dist


class Comm:
    def __init__(self, group: Any = None):
        self.enabled = dist.is_available() and dist.is_initialized()
        if self.enabled:
            self.rank = dist.get_rank()
            self.world = dist.get_world_size()
            self.backend = dist.get_backend()
            if self.backend == "gloo":
                self.ctrl = dist.group.WORLD
            else:
                self.ctrl = _gloo_group()
        else:
            self.rank, self.world, self.backend, self.ctrl = 0, 1, "none", None

    # ------------------------------------------------------------ bootstrap helpers
    @staticmethod
    def init_from_env(prefer_cuda: bool = True) -> "Comm":
        """Initialise torch.distributed from RANK/WORLD_SIZE/MASTER_* if launched by torchrun."""
        if "RANK" in os.environ and "WORLD_SIZE" in os.environ and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29533")
            use_cuda = prefer_cuda and torch.cuda.is_available()
            if use_cuda:
                torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
                dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
            else:
                dist.init_process_group("gloo")
        return Comm()

    # ------------------------------------------------------------ control-plane collectives
    def barrier(self) -> None:
        if self.enabled and self.world > 1:
            dist.barrier(group=self.ctrl)

    def allgather_bytes(self, payload: bytes) -> list[bytes]:
        if not self.enabled or self.world == 1:
            return [payload]
        out: list[Any] = [None] * self.world
        dist.all_gather_object(out, payload, group=self.ctrl)
        return out

    def allgather_obj(self, obj: Any) -> list[Any]:
        return [pickle.loads(b) for b in self.allgather_bytes(pickle.dumps(obj))]

    def allreduce_max(self, x: float) -> float:
        if not self.enabled or self.world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.ctrl)
        return float(t.item())

    def allreduce_sum_int(self, x: int) -> int:
        if not self.enabled or self.world == 1:
            return int(x)
        t = torch.tensor([x], dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.ctrl)
        return int(t.item())

    def bcast_obj(self, obj: Any, src: int = 0) -> Any:
        if not self.enabled or self.world == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=src, group=self.ctrl)
        return box[0]

    # ------------------------------------------------------------ file descriptors (same host)
    def _fd_sock_name(self, rank: int, tag: str) -> str:
        port = os.environ.get("MASTER_PORT", "0")
        return f"\0jb_fd_{os.getuid()}_{port}_{tag}_{rank}"   # abstract namespace: nothing to unlink

    def exchange_fds(self, fd: int, tag: str = "x") -> list[int]:
        """All-to-all of one open file descriptor per rank (SCM_RIGHTS over abstract Unix sockets):
        the channel that carries CUDA VMM allocation handles between the rank processes of one
        host (csrc/comm/symm_heap.cpp).  Returns the received descriptors, own slot = ``fd``."""
        import socket

        if not self.enabled or self.world == 1:
            return [fd]
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(self._fd_sock_name(self.rank, tag))
        srv.listen(self.world)
        self.barrier()                      # every rank listens before anybody connects
        for r in range(self.world):
            if r == self.rank:
                continue
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(self._fd_sock_name(r, tag))
            socket.send_fds(c, [self.rank.to_bytes(4, "little")], [fd])
            c.close()
        out = [-1] * self.world
        out[self.rank] = fd
        for _ in range(self.world - 1):
            conn, _addr = srv.accept()
            msg, fds, _flags, _a = socket.recv_fds(conn, 4, 1)
            out[int.from_bytes(msg, "little")] = fds[0]
            conn.close()
        srv.close()
        self.barrier()
        return out

    def bcast_fd(self, fd: int, src: int = 0, tag: str = "b") -> int:
        """One descriptor from ``src`` to every rank (the multicast object's handle)."""
        import socket

        if not self.enabled or self.world == 1:
            return fd
        if self.rank != src:
            srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            srv.bind(self._fd_sock_name(self.rank, tag))
            srv.listen(1)
        self.barrier()
        got = fd
        if self.rank == src:
            for r in range(self.world):
                if r == src:
                    continue
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                c.connect(self._fd_sock_name(r, tag))
                socket.send_fds(c, [b"m"], [fd])
                c.close()
        else:
            conn, _addr = srv.accept()
            _msg, fds, _flags, _a = socket.recv_fds(conn, 1, 1)
            got = fds[0]
            conn.close()
            srv.close()
        self.barrier()
        return got

    def send_tensor(self, t: torch.Tensor, dst: int) -> None:
        dist.send(t, dst=dst, group=self.ctrl)

    def recv_tensor(self, t: torch.Tensor, src: int) -> None:
        dist.recv(t, src=src, group=self.ctrl)


_GLOO = None


def _gloo_group():
    global _GLOO
    if _GLOO is None:
        _GLOO = dist.new_group(backend="gloo")
    return _GLOO
