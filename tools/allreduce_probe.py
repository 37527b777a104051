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
    # the same exchange over symmetric (peer-mapped) memory: torch's symm_mem kernels; multimem = in-switch reduction
    # through NVLink multicast (NVLS), two_shot = reduce-scatter + all-gather over peer loads/stores
    try:
        import torch.distributed._symmetric_memory as symm_mem
        name = dist.group.WORLD.group_name
        t = symm_mem.empty(16 * P, dtype=torch.float32, device="cuda")
        symm_mem.rendezvous(t, dist.group.WORLD)
        for op_name in ("multimem_all_reduce_", "two_shot_all_reduce_"):
            op = getattr(torch.ops.symm_mem, op_name)
            t.fill_(1.0)
            torch.cuda.synchronize()
            dist.barrier()
            op(t, "sum", name)
            torch.cuda.synchronize()
            ok = bool((t[:1000] == world).all()) and bool((t[-1000:] == world).all())
            t.fill_(0.0)
            for _ in range(3):
                op(t, "sum", name)
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            n = 20
            for _ in range(n):
                op(t, "sum", name)
            e1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            if rank == 0:
                mb = 16 * P * 4 / 1e6
                print("world %d  symm_mem.%s: %.1f MB  %.3f ms  algbw %.0f GB/s  correct=%s" % (
                    world, op_name, mb, float(ms), mb / float(ms), ok), flush=True)
    except Exception as e:  # noqa: BLE001 -- a probe: report and go on
        if rank == 0:
            print("symm_mem probe failed: %r" % (e,), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
