#!/usr/bin/env python3
"""Development tool (run under torch.distributed.run, one rank per GPU): device time of the four phases of the MPIAIJ
MatMult on the bench matrix (7-pt 256^3 per GPU), both transports."""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, lrank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lrank)
dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
pb = importlib.import_module("petsc-dev_b200"); par = importlib.import_module("petsc-dev_b200.parallel"); P = importlib.import_module("workloads.problems")
pb.init(lrank); par.comm_init(rank, world, dist, device="cuda")
if (sys.argv[1] if len(sys.argv) > 1 else "p2p") == "p2p":
    par.enable_p2p()
nx = 256
N = nx ** 3 * world
ranges = par.split_ownership(N, world)
ai, aj, aa = P.laplace3d_7pt(nx, nx, nx * world, r0=int(ranges[rank]), r1=int(ranges[rank + 1]))
host = par.setup_mpiaij_host(rank, world, ranges, ai - ai[0], aj, aa, N, dist=dist, device="cuda")
A = par.create_mpiaij(host); m = host["m"]
x, y = pb.Vec(m, P.random_vector(m, 2, r0=rank * m)), pb.Vec(m)
ms = (C.c_double * 4)()
acc = np.zeros(4)
for it in range(30):
    pb._chk(pb.lib.b200_mat_mult_phases(A.h, x.p, y.p, ms), "phases")
    if it >= 10:
        acc += np.array(list(ms))
acc /= 20
# back-to-back loop for comparison
for _ in range(5): A.mult(x, y)
pb.synchronize(); dist.barrier(); pb.timer_start()
for _ in range(50): A.mult(x, y)
t = pb.timer_stop() / 50
print("rank %d %s: begin %.1f us | diag %.1f | end %.1f | offdiag %.1f | sum %.1f ; back-to-back %.1f us" % (rank, sys.argv[1] if len(sys.argv) > 1 else "p2p", *(acc * 1e3), acc.sum() * 1e3, t * 1e3), flush=True)
dist.barrier(); dist.destroy_process_group()
