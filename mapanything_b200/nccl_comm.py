"""An NCCL communicator for the view-sharded forward inside libmab200 (`mab_set_comm`, include/mab200.h).

torch.distributed does not hand out its ncclComm_t, so the ranks that share a scene create one of their own
on the NCCL library torch itself has loaded (ctypes; ncclGetUniqueId on rank 0, the 128 id bytes broadcast
through torch.distributed, ncclCommInitRank everywhere) and give it to the library together with that
library's path.  No reference counterpart: the reference has no multi-GPU inference path.
"""
import ctypes as C
import os

import torch
import torch.distributed as dist


class _UniqueId(C.Structure):
    _fields_ = [("internal", C.c_byte * 128)]


def nccl_library_path():
    """The libnccl this process has mapped (torch's), else the wheel's copy next to torch."""
    try:
        with open("/proc/self/maps") as f:
            for line in f:
                if "libnccl.so" in line:
                    return line.split()[-1]
    except OSError:
        pass
    import importlib.util
    spec = importlib.util.find_spec("nvidia.nccl")
    if spec and spec.submodule_search_locations:
        p = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
        if os.path.exists(p):
            return p
    raise RuntimeError("libnccl.so.2 not found (torch's NCCL is neither mapped nor installed as nvidia.nccl)")


def create(device, group=None):
    """Collective over `group`: returns (comm pointer, rank, world, library path).  The communicator lives
    until the process exits (it is shared by every model instance on this device)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    path = nccl_library_path()
    lib = C.CDLL(path)
    lib.ncclGetErrorString.restype = C.c_char_p
    uid = _UniqueId()
    if rank == 0:
        rc = lib.ncclGetUniqueId(C.byref(uid))
        if rc != 0:
            raise RuntimeError("ncclGetUniqueId: " + lib.ncclGetErrorString(rc).decode())
    t = torch.frombuffer(bytearray(bytes(uid)), dtype=torch.uint8).to(device)
    src = dist.get_global_rank(group, 0) if group is not None else 0
    dist.broadcast(t, src=src, group=group)
    C.memmove(C.byref(uid), bytes(t.cpu().numpy().tobytes()), 128)
    comm = C.c_void_p()
    lib.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _UniqueId, C.c_int]
    with torch.cuda.device(device):
        rc = lib.ncclCommInitRank(C.byref(comm), world, uid, rank)
    if rc != 0:
        raise RuntimeError("ncclCommInitRank: " + lib.ncclGetErrorString(rc).decode())
    return comm, rank, world, path
