"""Per-rank PCIe bandwidth with all ranks copying at once (run under torchrun, one rank per GPU):
H2D alone, D2H alone, both directions together; pinned host memory, 512 MiB per copy, CUDA
events, all ranks start behind a barrier.  Shows whether the end-to-end leg of bench.py at
N = 8 is bound by one host memory system feeding eight links.
Also H2D from a write-combined pinned buffer (cudaHostAllocWriteCombined through libcudart) and
the host's NUMA layout, the two remedies VERDICT r1 item 10 names.
One JSON line from rank 0: per-rank GB/s min / mean over ranks."""
import ctypes, glob, json, os
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 512 << 20
h_a = torch.empty(n, dtype=torch.uint8).pin_memory()
h_b = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device="cuda")
d_b = torch.empty(n, dtype=torch.uint8, device="cuda")
s2 = torch.cuda.Stream()


def timed(fn, reps=6):
    best = 1e9
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        s2.synchronize() if False else None
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return n / best / 1e6  # GB/s


def duplex():
    d_a.copy_(h_a, non_blocking=True)
    with torch.cuda.stream(s2):
        h_b.copy_(d_b, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s2)


def wc_h2d():
    """GB/s of H2D from write-combined pinned memory (0.0 when libcudart cannot be loaded)."""
    rt = None
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = ctypes.CDLL(name)
            break
        except OSError:
            pass
    if rt is None:
        import nvidia.cuda_runtime
        rt = ctypes.CDLL(glob.glob(os.path.join(os.path.dirname(nvidia.cuda_runtime.__file__), "lib", "libcudart.so*"))[0])
    ptr = ctypes.c_void_p()
    if rt.cudaHostAlloc(ctypes.byref(ptr), ctypes.c_size_t(n), ctypes.c_uint(4)) != 0:  # 4 = cudaHostAllocWriteCombined
        return 0.0
    ctypes.memset(ptr, 1, n)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    gbs = timed(lambda: rt.cudaMemcpyAsync(ctypes.c_void_p(d_a.data_ptr()), ptr, ctypes.c_size_t(n), ctypes.c_int(1), stream))
    rt.cudaFreeHost(ptr)
    return gbs


res = [timed(lambda: d_a.copy_(h_a, non_blocking=True)), timed(lambda: h_b.copy_(d_b, non_blocking=True)), timed(duplex), wc_h2d()]
t = torch.tensor(res, dtype=torch.float64, device="cuda")
if world > 1:
    allr = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(allr, t)
    allr = torch.stack(allr).cpu()
else:
    allr = t.cpu()[None]
if rank == 0:
    names = ["h2d_gbs", "d2h_gbs", "duplex_each_way_gbs", "h2d_write_combined_gbs"]
    print(json.dumps({"n_gpus": world, **{k: {"min": float(allr[:, i].min()), "mean": float(allr[:, i].mean())} for i, k in enumerate(names)},
                      "aggregate_h2d_gbs": float(allr[:, 0].sum()),
                      "numa_nodes": len(glob.glob("/sys/devices/system/node/node[0-9]*")),
                      "host_cpus": len(os.sched_getaffinity(0))}), flush=True)
if world > 1:
    dist.destroy_process_group()
