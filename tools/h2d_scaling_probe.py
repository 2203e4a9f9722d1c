"""What feeds N GPUs from host memory at once?  (e2e efficiency 0.53 at N = 4, 0.43 at N = 8.)

torchrun --nproc-per-node N tools/h2d_scaling_probe.py            # every rank streams 8 MB pinned blocks to its GPU, concurrently
Modes: default allocation; NUMA-local (the rank pins itself to the CPUs next to its GPU before it allocates and touches the
pinned pool: first-touch places the pages on that node); and write-combined pinned memory.  Prints aggregate GB/s per mode.
"""
import ctypes as C
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
GB = 4                      # pinned pool per rank
BLOCK = 8 * 1000 * 1000     # one 1000 x 1000 FFMatrixBlock of doubles
n_blocks = GB * (1 << 30) // BLOCK
dev = torch.empty(n_blocks * BLOCK, dtype=torch.uint8, device="cuda")
rt = C.CDLL("libcudart.so.12")


def local_cpus():
    """CPUs next to this rank's GPU (sysfs local_cpulist of its PCI device)."""
    bus = torch.cuda.get_device_properties(local).pci_bus_id if hasattr(torch.cuda.get_device_properties(local), "pci_bus_id") else None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = (bus.decode() if isinstance(bus, bytes) else bus).lower()
        bus = bus[-12:] if len(bus) > 12 else bus
        txt = open(f"/sys/bus/pci/devices/{bus}/local_cpulist").read().strip()
        node = open(f"/sys/bus/pci/devices/{bus}/numa_node").read().strip()
    except Exception as e:  # noqa: BLE001
        return None, f"unknown ({e!r})"
    cpus = set()
    for part in txt.split(","):
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus, node


def run(mode):
    if mode == "numa":
        cpus, node = local_cpus()
        if cpus:
            os.sched_setaffinity(0, cpus)
    if mode == "wc":
        p = C.c_void_p()
        assert rt.cudaHostAlloc(C.byref(p), C.c_size_t(n_blocks * BLOCK), C.c_uint(4)) == 0     # cudaHostAllocWriteCombined
        host = None
        C.memset(p, 1, n_blocks * BLOCK)
        hp = p.value
    else:
        host = torch.empty(n_blocks * BLOCK, dtype=torch.uint8)
        host.fill_(1)                                       # first touch on the node this rank runs on
        assert rt.cudaHostRegister(C.c_void_p(host.data_ptr()), C.c_size_t(host.numel()), C.c_uint(1)) == 0
        hp = host.data_ptr()
    s = torch.cuda.current_stream().cuda_stream

    def sweep():
        for i in range(n_blocks):
            rt.cudaMemcpyAsync(C.c_void_p(dev.data_ptr() + i * BLOCK), C.c_void_p(hp + i * BLOCK), C.c_size_t(BLOCK), C.c_int(1), C.c_void_p(s))

    sweep(); torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        sweep()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if mode == "wc":
        rt.cudaFreeHost(p)
    else:
        rt.cudaHostUnregister(C.c_void_p(host.data_ptr()))
    agg = world * 3 * n_blocks * BLOCK / float(t.item()) / 1e9
    if rank == 0:
        print(f"N={world} mode={mode:8s} aggregate H2D {agg:7.1f} GB/s ({agg / world:5.1f} per GPU), {BLOCK / 1e6:.0f} MB copies", flush=True)


cpus, node = local_cpus()
print(f"rank {rank}: GPU {local} numa node {node}, {len(cpus) if cpus else '?'} local CPUs, process may run on {len(os.sched_getaffinity(0))} CPUs", flush=True)
for mode in ("default", "numa", "wc"):
    run(mode)
if world > 1:
    dist.destroy_process_group()
