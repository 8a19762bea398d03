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
lib.probe.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int]
names = {0: "random 32 B rows (one 256-bit store)", 1: "random 4 B words", 2: "consecutive 8 B words", 3: "random 128 B chunks (4 x 32 B)",
         4: "random 16 B", 5: "consecutive 32 B rows"}
size = {0: 32, 1: 4, 2: 8, 3: 32, 4: 16, 5: 32}
dist.barrier()
if rank == 0:
    for where, g in (("peer", 1), ("local", 0)):
        dst = lad.bases[g] + lad.layouts[g][0][0]
        for mode in (0, 4, 1, 3, 2, 5):
            count = 20_000_000 if mode == 1 and where == "peer" else 100_000_000
            ms = lib.probe(dst, u, count, mode, 148 * 8)
            print("%-5s %-40s %8.3f ms for %.0fM stores: %6.2f G stores/s, %7.1f GB/s" % (
                where, names[mode], ms, count / 1e6, count / ms / 1e6, count * size[mode] / ms / 1e6), flush=True)
torch.cuda.synchronize()
dist.barrier()
lad.close()
dist.destroy_process_group()
