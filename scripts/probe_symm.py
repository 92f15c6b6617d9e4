"""Probe (2+ GPUs, torchrun): is symmetric memory with peer pointers / multicast available on this box, and how long does a
20 MB fp32 all-reduce take through NCCL and through torch's symmetric-memory kernels (reference points for our own)?"""
import os, sys, time
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl")
n = 5_010_401
def timeit(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
x = torch.randn(n, device="cuda")
res = {"nccl_us": timeit(lambda: dist.all_reduce(x, op=dist.ReduceOp.AVG))}
try:
    t = sm.empty(n, dtype=torch.float32, device="cuda")
    h = sm.rendezvous(t, dist.group.WORLD.group_name)
    res["multicast"] = bool(h.has_multicast_support)
    res["mc_ptr"] = int(h.multicast_ptr) != 0
    res["signal_pad_bytes"] = int(h.signal_pad_size)
    t.copy_(x)
    g = dist.group.WORLD.group_name
    for name in ("two_shot_all_reduce_", "multimem_all_reduce_"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            res[name + "_us"] = timeit(lambda: op(t, "sum", g))
        except Exception as e:
            res[name] = "failed: " + str(e)[:200]
    for name in ("one_shot_all_reduce", "multimem_one_shot_all_reduce"):
        try:
            op = getattr(torch.ops.symm_mem, name)
            res[name + "_us"] = timeit(lambda: op(t, "sum", g))
        except Exception as e:
            res[name] = "failed: " + str(e)[:200]
except Exception as e:
    res["symm_mem"] = "failed: " + str(e)[:300]
if rank == 0:
    print(res, flush=True)
dist.destroy_process_group()
