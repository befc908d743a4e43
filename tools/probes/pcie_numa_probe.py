"""Analysis aid (not part of the product path): how fast do N GPUs read back to host memory AT THE SAME TIME,
and does it depend on where the pages live?   torchrun --nproc-per-node N tools/probes/pcie_numa_probe.py

A  pinned buffer allocated by the rank as it happens to be scheduled
B  the same after binding the process to the GPU's NUMA node (NVML cpu affinity), fresh buffer
C  one /dev/shm image shared by the ranks, every rank first-touching and reading back its own share (B's binding)
D  as C, but rank 0 touches every page first (all pages on one node)
"""
import mmap
import os
import subprocess
import time

import numpy as np
import torch
import torch.distributed as dist

MB = 64


def d2h_rate(dst_ptr_tensor_or_ptr, src, nbytes, reps=20):
    from cuda.bindings import runtime as cr

    s = torch.cuda.current_stream()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if isinstance(dst_ptr_tensor_or_ptr, torch.Tensor):
            dst_ptr_tensor_or_ptr.copy_(src, non_blocking=True)
        else:
            (err,) = cr.cudaMemcpyAsync(dst_ptr_tensor_or_ptr, src.data_ptr(), nbytes,
                                        cr.cudaMemcpyKind.cudaMemcpyDeviceToHost, s.cuda_stream)
            assert int(err) == 0, err
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    dist.barrier()
    return nbytes * reps / dt / 1e9


def gather(x):
    t = torch.tensor([x], dtype=torch.float64)
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [round(float(o.item()), 1) for o in out]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    torch.cuda.set_device(rank)
    n = MB << 20
    src = torch.empty(n, dtype=torch.uint8, device=f"cuda:{rank}")
    if rank == 0:
        print("allowed cpus:", len(os.sched_getaffinity(0)), sorted(os.sched_getaffinity(0))[:4], "...")
        for node in sorted(os.listdir("/sys/devices/system/node")) if os.path.isdir("/sys/devices/system/node") else []:
            if node.startswith("node"):
                print(node, open(f"/sys/devices/system/node/{node}/cpulist").read().strip())
        r = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True)
        print(r.stdout)
    a = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    a.fill_(1)
    ra = gather(d2h_rate(a, src, n))
    info = "no nvml"
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(rank)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        info = f"bound to {len(os.sched_getaffinity(0))} cpus from {min(os.sched_getaffinity(0))}"
    except Exception as e:  # noqa: BLE001
        info = f"affinity failed: {e!r}"
    infos = [None] * world
    dist.all_gather_object(infos, info)
    b = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    b.fill_(1)
    rb = gather(d2h_rate(b, src, n))
    res = {}
    for tag, touch_all in (("C", False), ("D", True)):
        path = f"/dev/shm/pcie_probe_{tag}"
        total = n * world
        if rank == 0:
            with open(path, "wb") as f:
                f.truncate(total)
        dist.barrier()
        fd = os.open(path, os.O_RDWR)
        m = mmap.mmap(fd, total)
        os.close(fd)
        arr = np.frombuffer(m, dtype=np.uint8)
        if touch_all:
            if rank == 0:
                arr[:] = 1
            dist.barrier()
        else:
            arr[rank * n:(rank + 1) * n] = 1
            dist.barrier()
        rtl = torch.cuda.cudart()
        err = rtl.cudaHostRegister(arr.ctypes.data, total, 0)
        assert int(err) == 0, err
        dist.barrier()
        res[tag] = gather(d2h_rate(arr.ctypes.data + rank * n, src, n))
        rtl.cudaHostUnregister(arr.ctypes.data)
        del arr
        dist.barrier()
        try:
            m.close()
        except BufferError:
            pass
        dist.barrier()
        if rank == 0:
            os.unlink(path)
    if rank == 0:
        print("affinity:", infos)
        for tag, r in (("A unbound private", ra), ("B bound private", rb), ("C shm own-share touch", res["C"]),
                       ("D shm rank0 touch", res["D"])):
            print(f"{tag:24s} per-rank GB/s {r}  total {sum(r):.0f}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
