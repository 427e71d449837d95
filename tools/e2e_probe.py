"""Where does the time of one sharded host-buffer product go?  torchrun --nproc-per-node N tools/e2e_probe.py
Prints the wall time of each dist.mul_host call (8192^3 f32, A and C row-sharded) and, for the last call, the library's own
timeline on rank 0 (LA_DEBUG_HOSTGEMM)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as td
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
if rank == 0 and "--debug" in sys.argv:
    os.environ["LA_DEBUG_HOSTGEMM"] = "1"
torch.cuda.set_device(local)
td.init_process_group("nccl", device_id=torch.device("cuda", local))
if "--bind" in sys.argv:
    import bench
    print("rank %d: %s" % (rank, bench.bind_near_gpu(local)), flush=True)
import linear_algebra_b200 as la
from linear_algebra_b200 import dist as ladist
la.init(local)
comm = ladist.Comm()
n = 8192; rows = n // world
hA = la.pinned(rows * n, "f32"); hB = la.pinned(n * n, "f32"); hC = la.pinned(rows * n, "f32")
hA[:] = np.random.default_rng(rank).uniform(-1, 1, rows * n).astype(np.float32)
hB[:] = np.random.default_rng(99).uniform(-1, 1, n * n).astype(np.float32)
for it in range(6):
    la.synchronize(); comm.barrier()
    t = time.perf_counter()
    ladist.mul_host(comm, hA, hB, n, n, n, out=hC)
    dt = (time.perf_counter() - t) * 1e3
    print("rank %d call %d: %.3f ms" % (rank, it, dt), flush=True)
comm.barrier()
td.destroy_process_group()
