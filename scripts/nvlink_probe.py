"""How fast can one B200 store into another one's memory?  2 processes (torchrun), rank 0 stores
into rank 1's mailbox (NVLink peer mapping from analyzer_b200.sharded) and, for comparison, into
its own.  Build first: nvcc -shared -Xcompiler -fPIC -gencode arch=compute_100a,code=sm_100a
-o scripts/libnvprobe.so scripts/nvlink_probe.cu"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from analyzer_b200 import engine, sharded

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
r = engine.Rater(engine.Env(), rank)
u = 10_000_000
lad = sharded.ShardedLadder(r, u, 1024, 6, rank, world)
lib = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "libnvprobe.so"))
lib.probe.restype = C.c_float
lib.probe.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint64, C.c_uint64, C.c_int, C.c_int]
names = {0: "random 32 B rows (one 256-bit store)", 1: "random 4 B words", 2: "consecutive 8 B words", 3: "random 128 B chunks (4 x 32 B)",
         4: "random 16 B", 5: "consecutive 32 B rows"}
size = {0: 32, 1: 4, 2: 8, 3: 32, 4: 16, 5: 32}


def run(where, ranks_to, modes, count=100_000_000):
    arr = (C.c_void_p * len(ranks_to))(*[lad.bases[g] + lad.layouts[g][0][0] for g in ranks_to])
    for mode in modes:
        c = 20_000_000 if mode == 1 and where != "local" else count
        ms = lib.probe(arr, len(ranks_to), u, c, mode, 148 * 8)
        print("rank %d -> %-14s %-40s %8.3f ms for %.0fM stores: %6.2f G stores/s, %7.1f GB/s" % (
            rank, where, names[mode], ms, c / 1e6, c / ms / 1e6, c * size[mode] / ms / 1e6), flush=True)


dist.barrier()
peers = [g for g in range(world) if g != rank]
if rank == 0:
    run("1 peer", peers[:1], (0, 4, 1, 3, 2, 5))
    if world > 2:
        run("%d peers" % min(3, len(peers)), peers[:3], (0, 3))
        run("%d peers" % len(peers), peers, (0, 3, 1, 5, 2))
    run("local", [0], (0, 1, 3))
torch.cuda.synchronize()
dist.barrier()
if world > 2:                       # every rank stores into all its peers at the same time
    run("all peers, all ranks at once", peers, (5,), 100_000_000)
    torch.cuda.synchronize()
    dist.barrier()
    run("all peers, all ranks at once", peers, (0,), 20_000_000)
torch.cuda.synchronize()
dist.barrier()
lad.close()
dist.destroy_process_group()
