"""Probe (torchrun, 2+ GPUs): time of the pieces of the NVLS gradient averaging -- the cross-GPU barrier, our multimem
kernel at several grid sizes, the whole sequence -- against NCCL's all-reduce on the same 20 MB."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
from gnn_b200 import _lib
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl")
n = 5_010_432
buf = symm.empty(n, dtype=torch.float32, device="cuda")
h = symm.rendezvous(buf, dist.group.WORLD.group_name)
buf.normal_()
mc = ctypes.c_void_p(int(h.multicast_ptr))
st = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def timeit(fn, it=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / it * 1e3, 1)
res = {"world": world}
x = torch.randn(n, device="cuda")
res["nccl_avg_us"] = timeit(lambda: dist.all_reduce(x, op=dist.ReduceOp.AVG))
res["barrier_us"] = timeit(lambda: h.barrier(channel=0))
peers = (ctypes.c_uint64 * world)(*[int(q) for q in h.buffer_ptrs])
for ctas in (16, 32, 64, 148, 296, 592):
    def seqp(c=ctas):
        h.barrier(channel=0)
        _lib.call("gnnb_allreduce_mean_p2p", peers, n, rank, world, c, st())
        h.barrier(channel=1)
    res[f"p2p_seq_{ctas}_us"] = timeit(seqp)
    res[f"p2p_kernel_{ctas}_us"] = timeit(lambda c=ctas: _lib.call("gnnb_allreduce_mean_p2p", peers, n, rank, world, c, st()))
for ctas in (16, 64, 148):
    def seq(c=ctas):
        h.barrier(channel=0)
        _lib.call("gnnb_allreduce_mean_multimem", mc, n, rank, world, c, st())
        h.barrier(channel=1)
    res[f"seq_{ctas}_us"] = timeit(seq)
    res[f"kernel_{ctas}_us"] = timeit(lambda c=ctas: _lib.call("gnnb_allreduce_mean_multimem", mc, n, rank, world, c, st()))
if rank == 0:
    print(res, flush=True)
dist.destroy_process_group()
