"""Why does the end-to-end figure stop scaling past 2 GPUs?  (VERDICT r1 weak #9.)  Under torchrun every rank uploads a
pinned host buffer to its GPU, first alone (the others wait), then all at once; plus the NUMA node of each GPU and of the
pinned allocation.  If the concurrent per-rank rate is a fraction of the solo rate, the uploads share something on the
host side (PCIe switch uplink, one socket's memory controllers), which no kernel can fix.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 scripts/diag_h2d.py"""
import os
import subprocess
import time

import torch
import torch.distributed as dist


def rate(dst, src, reps=4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    return src.numel() * src.element_size() * reps / (time.perf_counter() - t0) / 1e9


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = 1 << 28   # 1 GiB of int32
    src = torch.empty(n, dtype=torch.int32).pin_memory()
    src.fill_(rank)
    dst = torch.empty(n, dtype=torch.int32, device="cuda")
    rate(dst, src, 1)
    solo = torch.zeros(world, dtype=torch.float64, device="cuda")
    for r in range(world):
        dist.barrier()
        if r == rank:
            solo[r] = rate(dst, src)
        dist.barrier()
    dist.all_reduce(solo)
    dist.barrier()
    both = torch.zeros(world, dtype=torch.float64, device="cuda")
    both[rank] = rate(dst, src)
    dist.all_reduce(both)
    topo = ""
    if rank == 0:
        try:
            topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=30).stdout
        except Exception as e:  # noqa: BLE001
            topo = repr(e)
        print("pinned host -> device, 1 GiB per copy, GB/s per rank")
        print("rank  alone  all %d at once" % world)
        for r in range(world):
            print("%4d  %5.1f  %5.1f" % (r, solo[r].item(), both[r].item()))
        print("sum   %5.1f  %5.1f   (host cores: %d)" % (solo.sum().item(), both.sum().item(), os.cpu_count()))
        print(topo)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
