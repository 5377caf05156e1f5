"""Host <-> device copy ceiling of the box with N ranks copying at the same time, no kernels.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/h2d_ceiling.py [--mb 370] [--out-mb 85] [--iters 10]

Every rank copies `--mb` MB of pinned host memory to its GPU and `--out-mb` MB back (two streams, as the e2e leg of
bench.py does: 1024 images 752x480 in, keypoints + descriptors + matches out), all ranks between the same two barriers.
Variants: torch pin_memory() (cudaHostRegister of malloc'ed pages), cudaHostAlloc default, cudaHostAllocWriteCombined
for the input, and each of them with the rank's threads bound to a disjoint core set.  Rank 0 prints one JSON line:
per-rank and aggregate GB/s per variant = what the e2e number of bench.py --gpus N can reach at most on this host."""
import argparse
import ctypes
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=370.0)
    ap.add_argument("--out-mb", type=float, default=85.0)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nin, nout = int(a.mb * 1e6), int(a.out_mb * 1e6)
    d_in = torch.empty(nin, dtype=torch.uint8, device=dev)
    d_out = torch.zeros(nout, dtype=torch.uint8, device=dev)
    rt = ctypes.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else ctypes.CDLL("libcudart.so")
    rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(h_in_ptr, h_out_ptr):
        def once():
            rt.cudaMemcpyAsync(d_in.data_ptr(), h_in_ptr, nin, 1, s_in.cuda_stream)
            rt.cudaMemcpyAsync(h_out_ptr, d_out.data_ptr(), nout, 2, s_out.cuda_stream)
        for _ in range(2):
            once()
        barrier()
        t0 = time.perf_counter()
        for _ in range(a.iters):
            once()
        torch.cuda.synchronize()
        mine = time.perf_counter() - t0
        barrier()
        t = torch.tensor([mine], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item()) / a.iters
        return {"ms_per_step": sec * 1e3, "h2d_gbs_per_rank": nin / sec * 1e-9, "h2d_gbs_all": world * nin / sec * 1e-9,
                "d2h_gbs_all": world * nout / sec * 1e-9}

    def host_alloc(n, flags):
        p = ctypes.c_void_p()
        rc = rt.cudaHostAlloc(ctypes.byref(p), n, flags)
        assert rc == 0, rc
        ctypes.memset(p, 1, n)
        return p

    res = {"ranks": world, "in_mb": a.mb, "out_mb": a.out_mb, "cpus": os.cpu_count()}
    ncpu = os.cpu_count() or 1
    for bind in (False, True):
        if bind:
            per = max(1, ncpu // world)
            try:
                os.sched_setaffinity(0, set(range(rank * per, min(ncpu, (rank + 1) * per))))
            except OSError:
                pass
        tag = "_bound" if bind else ""
        t_in = torch.ones(nin, dtype=torch.uint8).pin_memory()
        t_out = torch.zeros(nout, dtype=torch.uint8).pin_memory()
        res["torch_pin_memory" + tag] = run(t_in.data_ptr(), t_out.data_ptr())
        del t_in, t_out
        for name, fl in (("cudaHostAlloc_default", 0), ("cudaHostAlloc_write_combined_input", 4)):
            pi, po = host_alloc(nin, fl), host_alloc(nout, 0)
            res[name + tag] = run(pi, po)
            rt.cudaFreeHost(pi)
            rt.cudaFreeHost(po)
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
