"""Row-sharded neighbour graph of ONE 65,536-agent env over the ranks of a torchrun job:
every rank searches its share of the queries, one all-reduce assembles the graph; checked against the
single-GPU graph and timed.  torchrun --nproc-per-node N scripts/sharded_graph_check.py [agents] [k]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "macbf-gnn_b200")]
import torch
import torch.distributed as dist
from macbf_b200 import ops
from oracle import macbf_oracle as O

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
k = int(sys.argv[2]) if len(sys.argv) > 2 else 12
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
s, _ = O.synthetic_swarm(1, N, 5)
s = s.cuda()
full = ops.neighbor_topk(s, k)
got = ops.neighbor_topk_sharded(s, k)
ok = torch.equal(full, got)


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


t_full = timeit(lambda: ops.neighbor_topk(s, k))
t_shard = timeit(lambda: ops.neighbor_topk_sharded(s, k))
if rank == 0:
    print(f"agents={N} k={k} ranks={world}: sharded == single-GPU graph: {ok}; single GPU {t_full:.1f} us, "
          f"sharded incl. all-reduce {t_shard:.1f} us")
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if ok else 1)
