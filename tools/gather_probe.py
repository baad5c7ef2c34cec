"""Developer tool (torchrun, N >= 2): where the time of dist.gather_varlen goes for a payload of the eval-set size.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/gather_probe.py
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    n = 153_000_000 // world  # floats per rank (613 MB in total)
    pay = torch.randn(n, device=dev)

    def t(label, fn, reps=1):
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            out = fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        if rank == 0:
            print(f"{label}: {dt * 1e3:.2f} ms", flush=True)
        return out

    for rnd in range(3):
        if rank == 0:
            print(f"-- round {rnd}", flush=True)
        allp = t("torch.empty(world, n)", lambda: torch.empty(world, n, device=dev))
        t("zeros + copy (pad)", lambda: torch.zeros(n, device=dev).copy_(pay))
        if rank == 0:
            t("dist.gather", lambda: dist.gather(pay, list(allp.unbind(0)), dst=0))
        else:
            t("dist.gather", lambda: dist.gather(pay, None, dst=0))
        flat = torch.empty(world * n, device=dev)
        t("all_gather_into_tensor", lambda: dist.all_gather_into_tensor(flat, pay))
        t("cat", lambda: torch.cat([allp[r] for r in range(world)]))
        if rnd == 0:  # every rank calls t() for every label: t() holds a barrier
            host = t("pinned alloc", lambda: torch.empty(world * n, pin_memory=True) if rank == 0 else None)
        t("D2H pinned", lambda: host.copy_(flat, non_blocking=True) if rank == 0 else None)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
