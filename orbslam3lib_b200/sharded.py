"""Map-scale Hamming kNN partitioned over the GPUs of one box (BASELINE configs[4]): the partitioned
form of cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, 2) (Frame.cc:45,1227).  One process per GPU; query rows are
sharded, the train set is broadcast and the results are exchanged with NCCL inside liborbfe.so (orbfe_knn2_sharded);
torch.distributed is used only to hand the NCCL unique id to the other ranks."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import OrbfeError


def shard_range(n, world, rank):
    """rows [begin, begin + count) of rank `rank` (orbfe_shard_range: [g n / G, (g + 1) n / G))."""
    b, c = C.c_int64(), C.c_int64()
    rc = _lib.lib().orbfe_shard_range(int(n), int(world), int(rank), C.byref(b), C.byref(c))
    if rc != _lib.OK:
        raise OrbfeError(rc, "orbfe_shard_range")
    return b.value, c.value


class ShardedMatcher:
    """knnMatch(k=2) of nq_total query rows against nt train rows over `world` ranks.  Tensors are torch CUDA tensors
    of this rank's device."""

    def __init__(self, matcher, rank=0, world=1, device=0, dist=None):
        self.m, self.rank, self.world, self.device, self._comm = matcher, int(rank), int(world), int(device), None
        if self.world > 1:
            import torch
            L = _lib.lib()
            uid = torch.zeros(128, dtype=torch.uint8)
            if self.rank == 0:
                buf = (C.c_char * 128)()
                rc = L.orbfe_nccl_unique_id(buf)
                if rc != _lib.OK:
                    raise OrbfeError(rc, L.orbfe_nccl_last_error().decode())
                uid = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
            # the id travels over the caller's process group (NCCL groups move CUDA tensors)
            backend = dist.get_backend()
            t = uid.cuda(self.device) if backend == "nccl" else uid
            dist.broadcast(t, src=0)
            raw = bytes(t.cpu().numpy().tobytes())
            comm = C.c_void_p()
            rc = L.orbfe_nccl_comm_create(raw, self.rank, self.world, self.device, C.byref(comm))
            if rc != _lib.OK:
                raise OrbfeError(rc, L.orbfe_nccl_last_error().decode())
            self._comm = comm

    def close(self):
        if self._comm is not None:
            _lib.lib().orbfe_nccl_comm_destroy(self._comm)
            self._comm = None

    def knn2(self, q_shard, nq_total, train, nt, idx_all, dist_all, root=0):
        """q_shard: this rank's rows (shard_range) uint8 (count, 32); train: uint8 (nt, 32), valid on `root`;
        idx_all / dist_all: int32 (nq_total, 2) on every rank.  Returns the device time in ms."""
        ms = C.c_float()
        rc = _lib.lib().orbfe_knn2_sharded(self.m._h, self._comm, self.rank, self.world, _lib.ptr(q_shard),
                                           int(nq_total), _lib.ptr(train), int(nt), int(root), _lib.ptr(idx_all),
                                           _lib.ptr(dist_all), C.byref(ms))
        if rc != _lib.OK:
            raise OrbfeError(rc, _lib.lib().orbfe_matcher_last_error(self.m._h).decode())
        return ms.value

    def last_times(self):
        """device ms of the last knn2: whole call, kNN launches, train broadcast (own stream), result exchange."""
        out = (C.c_float * 4)()
        _lib.lib().orbfe_knn2_sharded_times(self.m._h, out)
        return {"total_ms": out[0], "kernel_ms": out[1], "bcast_ms": out[2], "gather_ms": out[3]}
