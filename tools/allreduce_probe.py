"""Time the mapper's one collective alone: all_reduce(sum) of the fused fp32 [16, P] buffer over the ranks of one box.
torchrun --nproc-per-node N tools/allreduce_probe.py [P]   (prints device ms per all-reduce, max over ranks)"""
import os
import sys

import torch
import torch.distributed as dist


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_200_000
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl")
    for rows in (16, 8, 4):
        buf = torch.ones(rows * P, device="cuda")
        for _ in range(5):
            dist.all_reduce(buf)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 20
        for _ in range(n):
            dist.all_reduce(buf)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            mb = rows * P * 4 / 1e6
            print("world %d  %d rows x %d: %.1f MB  %.3f ms  algbw %.0f GB/s  busbw %.0f GB/s" % (
                world, rows, P, mb, float(ms), mb / float(ms), mb / float(ms) * 2 * (world - 1) / world), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
