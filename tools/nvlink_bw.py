"""Measured NVLink / NVLS roofs for the fused scale+broadcast path (profiles/nvlink_bw_N*.json).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/nvlink_bw.py [MiB]

Per rank, over the native symmetric heap (csrc/comm/symm_heap.cpp):
  local_store   16-byte stores into the own region               (HBM write roof of the kernel)
  peer_store    16-byte stores into rank+1's region over NVLink  (what unicast SCALE does per peer)
  peer_load     16-byte loads from rank+1's region               (what the residual's X pull does)
  mc_store      multimem.st into the NVLS multicast alias, rank 0 only while the others idle
                (what SCALE does with NVLS: one store, replicated by the switch into all N copies)
  memcpy_peer   cudaMemcpyPeer-style copy (torch copy_ between the two mappings)
GB/s are payload bytes / CUDA-event time, best of the timed launches' average."""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mpi_jordan_crazy_acceleration_b200 import _native  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.parallel.comm import Comm  # noqa: E402
from mpi_jordan_crazy_acceleration_b200.parallel.symm import NativeSymmHeap  # noqa: E402


def main():
    mib = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    comm = Comm.init_from_env()
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    lib = _native.cuda()
    lib.jb_store_bandwidth.restype = C.c_int
    lib.jb_store_bandwidth.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_float)]
    nbytes = mib << 20
    heap = NativeSymmHeap.create(comm, nbytes, dev)
    assert heap is not None, NativeSymmHeap.last_error
    p, k = comm.world, comm.rank

    def run(ptr, mode, iters=10):
        ms = C.c_float()
        _native.check(lib.jb_store_bandwidth(C.c_void_p(ptr), nbytes, mode, iters, C.byref(ms)))
        return nbytes / (ms.value * 1e-3) / 1e9

    res = {"rank": k, "mib": mib}
    comm.barrier()
    res["local_store_gbs"] = run(heap.ptrs[k], 0)
    comm.barrier()
    if p > 1:
        res["peer_store_gbs"] = run(heap.ptrs[(k + 1) % p], 0)       # every rank at once: ring pattern
        comm.barrier()
        res["peer_load_gbs"] = run(heap.ptrs[(k + 1) % p], 2)
        comm.barrier()
        if k == 0:                                                      # one sender, others idle
            res["peer_store_alone_gbs"] = run(heap.ptrs[1], 0)
        comm.barrier()
        if heap.mc_ptr:
            if k == 0:
                res["mc_store_gbs_payload"] = run(heap.mc_ptr, 1)       # payload rate; the switch delivers p copies
            comm.barrier()
    torch.cuda.synchronize()
    allr = comm.allgather_obj(res)
    if k == 0:
        print(json.dumps({"n_gpus": p, "multicast": bool(heap.mc_ptr), "per_rank": allr}))
    comm.barrier()
    heap.close()


if __name__ == "__main__":
    main()
