"""Probe (torchrun, N GPUs): cost of the sharded step's exchanges over NCCL point-to-point vs peer
memory (torch symmetric memory), and the host-side time of issuing them.  Scratch tool."""
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mvs_gaussian_splatting_b200.sharded import ShardPlan, all_to_all_rows  # noqa: E402


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    host = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, 1e3 * host


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    P = 3_000_000
    plan = ShardPlan(P, world)
    Ps, stride = plan.sizes[rank], plan.slot_stride(rank)
    widths = [1, 2, 4, 4, 1, 1]
    send = [torch.zeros(world * stride, k, device=dev) for k in widths]
    full = [torch.zeros(P, k, device=dev) for k in widths]
    pairs = [([s[v * stride:v * stride + Ps] for v in range(world)], [f[plan.rows(j)] for j in range(world)])
             for s, f in zip(send, full)]
    acc_full, acc_send = torch.zeros(P, 12, device=dev), torch.zeros(world * stride, 12, device=dev)
    bpairs = [([acc_full[plan.rows(j)] for j in range(world)], [acc_send[v * stride:v * stride + Ps] for v in range(world)])]
    out = {}
    out["nccl_p2p_fwd_6arrays"] = timed(lambda: all_to_all_rows(pairs))
    out["nccl_p2p_bwd_1array"] = timed(lambda: all_to_all_rows(bpairs))
    flat_in = torch.zeros(world, Ps * 13, device=dev)
    flat_out = torch.zeros(world, Ps * 13, device=dev)
    if len(set(plan.sizes)) == 1:
        out["nccl_all_to_all_single_52B"] = timed(lambda: dist.all_to_all_single(flat_out, flat_in))
    # ---- symmetric memory ----
    try:
        import torch.distributed._symmetric_memory as symm_mem
        nbytes = P * 52
        t = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
        try:
            hdl = symm_mem.rendezvous(t, dist.group.WORLD)
        except Exception as e1:
            print(f"[{rank}] rendezvous(group) failed: {e1!r}; trying group name", flush=True)
            hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
        if rank == 0:
            print("symm_mem handle attrs:", [a for a in dir(hdl) if not a.startswith("_")], flush=True)
        ptrs = list(hdl.buffer_ptrs)
        src = torch.zeros(Ps * 52, dtype=torch.uint8, device=dev)

        def push():
            hdl.barrier(channel=0)
            for j in range(world):
                peer = hdl.get_buffer(j, (nbytes,), torch.uint8, 0)
                peer[plan.starts[rank] * 52: plan.starts[rank] * 52 + Ps * 52].copy_(src)
            hdl.barrier(channel=1)
        out["symm_push_52B_copy_engine"] = timed(push)
        out["symm_barrier_only"] = timed(lambda: hdl.barrier(channel=0))
        if rank == 0:
            print("buffer_ptrs", [hex(p) for p in ptrs], flush=True)
    except Exception as e:
        print(f"[{rank}] symmetric memory unavailable: {e!r}", flush=True)
    if rank == 0:
        for k, (gpu_ms, host_ms) in out.items():
            print(f"{k:32s} gpu {gpu_ms:.3f} ms   host {host_ms:.3f} ms", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
