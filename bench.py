#!/usr/bin/env python3
"""bench.py — the headline measurement: AIJ MatMult GB/s (and KSPCG iterations/s) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--nx 256] [--cg-iters 100]

A step is one MatMult of BASELINE.json configs[1]: the 3-D 7-point Laplacian on a 256^3 grid per GPU (weak scaling:
N GPUs hold a 256 x 256 x 256N grid, row-partitioned into z-slabs exactly as PetscSplitOwnership does, ghost planes
exchanged GPU to GPU).  `value` = algorithmic bytes of the MatMult (12 B/nnz + 4 B/row + 8 B/row y +
8 B/col x, plus the ghost pack/receive traffic) x K / device time, aggregated over ranks.  Inputs are 1.7 GB per GPU,
far larger than the 126 MB L2, so no flush is needed between steps.  `e2e` is the same MatMult through the public API
with HOST vectors (pinned H2D of x, D2H of y inside the timed region).  The KSPCG/Jacobi rate on the same operator is
reported beside it (`kspcg`).  `--impl reference` times the reference's own CPU MatMult_SeqAIJ (oracle/_ref) on the
same workload on this box's host cores.
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "AIJ MatMult GB/s (7-pt Laplacian 256^3 per GPU); KSPCG iter/s beside it"


def spmv_bytes(nz, m, n):
    return 12.0 * nz + 4.0 * (m + 1) + 8.0 * n + 8.0 * m


class ClockSampler:
    """Samples SM clocks and throttle reasons during the timed region (nvidia-smi, 200 ms)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        import subprocess
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.proc.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(2)
        sm, mx, reasons = [], None, set()
        for l in self.lines:
            f = [s.strip() for s in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def env_rank():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def _ref_slab_worker(rank, nprocs, nx, reps, warm, barrier, q):
    """One host core: the reference's MatMult_SeqAIJ (oracle/_ref, unmodified aij.c:1225) on this core's row slab of the
    7-pt matrix (rows split like PetscSplitOwnership; global column ids, so x is the whole vector as lvec+local would
    be).  All cores start together; returns the average seconds per MatMult on this core."""
    try:
        from oracle import refpetsc as R
        P = importlib.import_module("workloads.problems")
        N = nx ** 3
        r0 = rank * (N // nprocs) + min(rank, N % nprocs)
        r1 = r0 + N // nprocs + (1 if rank < N % nprocs else 0)
        ai, aj, aa = P.laplace3d_7pt(nx, r0=r0, r1=r1)
        A = R.RefMat(r1 - r0, N, ai - ai[0], aj, aa)
        x = P.random_vector(N, 2)
        if warm:
            A.mult(x, reps=warm)
        barrier.wait()
        t0 = time.perf_counter()
        A.mult(x, reps=reps)
        dt = (time.perf_counter() - t0) / reps
        barrier.wait()
        q.put((rank, dt, int(len(aa)), r1 - r0))
    except Exception as e:                                  # report, never hang the other cores
        try:
            barrier.abort()
        except Exception:
            pass
        q.put((rank, None, repr(e), 0))


def reference_all_cores(nx, reps, warm=1, nprocs=None):
    """The reference CPU MatMult on all host cores of this box: one process per core, each with the unmodified reference
    kernel on its row slab (what MPIAIJ's ranks would each do, minus the ghost exchange this box has no MPI for).
    Returns (GB/s aggregate, cores, seconds per MatMult = slowest core)."""
    import multiprocessing as mp
    nprocs = nprocs or len(os.sched_getaffinity(0))
    ctx = mp.get_context("spawn")            # the parent may hold a CUDA context: never fork it
    barrier, q = ctx.Barrier(nprocs), ctx.Queue()
    procs = [ctx.Process(target=_ref_slab_worker, args=(r, nprocs, nx, reps, warm, barrier, q)) for r in range(nprocs)]
    [p.start() for p in procs]
    res = [q.get(timeout=900) for _ in procs]
    [p.join(60) for p in procs]
    bad = [r for r in res if r[1] is None]
    if bad:
        raise RuntimeError("reference worker failed: %s" % (bad[0][2],))
    N = nx ** 3
    nz = sum(r[2] for r in res)
    tmax = max(r[1] for r in res)
    return spmv_bytes(nz, N, N) / tmax / 1e9, nprocs, tmax


def run_reference(args):
    """The reference's own CPU implementation (unmodified MatMult_SeqAIJ, src/mat/impls/aij/seq/aij.c:1225, built in
    oracle/_ref) on the host cores; one MPIUNI rank = one core (no MPI exists on this box)."""
    rank, _, world = env_rank()
    if rank != 0:
        return
    from oracle import refpetsc as R
    P = importlib.import_module("workloads.problems")
    nx = args.nx
    N = nx ** 3
    # all host cores first (forked before this process holds the big matrix)
    gbs_all, cores, t_all = reference_all_cores(nx, reps=max(args.steps, 1), warm=max(args.warmup, 1))
    ip, j, a = P.laplace3d_7pt(nx)
    A = R.RefMat(N, N, ip, j, a)
    x = P.random_vector(N, 2)
    A.mult(x)
    t0 = time.perf_counter()
    reps1 = min(args.steps, 5)
    A.mult(x, reps=reps1)
    dt1 = (time.perf_counter() - t0) / reps1
    B = spmv_bytes(len(a), N, N)
    b = A.mult(np.ones(N))
    cg_its = min(args.cg_iters, 20)
    r = A.ksp_solve(b, "cg", "jacobi", rtol=1e-50, maxit=cg_its)
    line = {"impl": "reference", "metric": METRIC, "value": gbs_all, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_all * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: 3D 7-point Laplacian %d^3, AIJ MatMult (+KSPCG/Jacobi)" % nx, "rows": N,
                       "nnz": int(len(a)), "l2_policy": "inputs (1.7 GB) larger than cache"},
            "cpu_baseline": {"value": gbs_all, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "full %d^3 matrix, %d MatMult reps, one process per host core each running the "
                                       "unmodified reference MatMult_SeqAIJ on its row slab" % (nx, args.steps),
                             "one_core": {"value": B / dt1 / 1e9, "ms_per_step": dt1 * 1e3}},
            "kspcg": {"iter_per_s": r["its"] / r["seconds"], "iterations": r["its"], "cores": 1,
                      "note": "reference KSPSolve_CG + PCJACOBI, one MPIUNI rank (no MPI on this box)"},
            "e2e": {"value": gbs_all, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def run_b200(args):
    import ctypes as C
    rank, local_rank, world = env_rank()
    if world != args.gpus and world > 1:
        args.gpus = world
    pb = importlib.import_module("petsc-dev_b200")
    par = importlib.import_module("petsc-dev_b200.parallel")
    P = importlib.import_module("workloads.problems")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pb.init(local_rank)
    nx = args.nx
    Nglob = nx * nx * nx * world
    t_setup = time.perf_counter()
    if world == 1:
        ip, j, a = P.laplace3d_7pt(nx)
        A = pb.Mat.from_csr(Nglob, Nglob, ip, j, a)
        nz, m = len(a), Nglob
        B_alg = spmv_bytes(nz, m, m)
        del ip, j, a
    else:
        import torch
        par.comm_init(rank, world, dist, device="cuda")
        if args.transport == "p2p" and not par.enable_p2p(strict=False):
            args.transport = "nccl"          # agreed on by all ranks inside the library; reported in config.transport
        ranges = par.split_ownership(Nglob, world)
        ai, aj, aa = P.laplace3d_7pt(nx, nx, nx * world, r0=int(ranges[rank]), r1=int(ranges[rank + 1]))
        host = par.setup_mpiaij_host(rank, world, ranges, ai - ai[0], aj, aa, Nglob, dist=dist, device="cuda")
        A = par.create_mpiaij(host)
        m = host["m"]
        nz = len(aa)
        B_alg = A.bytes()
        del ai, aj, aa, host
    t_setup = time.perf_counter() - t_setup
    xh = P.random_vector(m, 2, r0=rank * m)
    x, y = pb.Vec(m, xh), pb.Vec(m)

    def barrier():
        pb.synchronize()
        if dist is not None:
            dist.barrier()

    def maxr(v):
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- MatMult (device-resident) ----------------
    for _ in range(max(args.warmup, 3)):
        A.mult(x, y)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()               # BEFORE the barrier: spawning the sampler costs rank 0 ~1 ms the other ranks would wait for
    barrier()
    l0 = pb.launch_count()
    pb.timer_start()
    for _ in range(args.steps):
        A.mult(x, y)
    ms = pb.timer_stop()
    launches = pb.launch_count() - l0
    barrier()
    ms = maxr(ms)
    ms_step = ms / args.steps
    # total algorithmic bytes over ranks
    Btot = B_alg
    if dist is not None:
        import torch
        t = torch.tensor([B_alg], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        Btot = float(t.item())
    value = Btot / (ms_step * 1e-3) / 1e9

    # ---------------- KSPCG + Jacobi, fixed iteration count ----------------
    one, b, sol = pb.Vec(m), pb.Vec(m), pb.Vec(m)
    one.set(1.0)
    A.mult(one, b)
    ksp = {}
    for label, fused in (("fused", 1), ("op_by_op", 0)):
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=5, fused=fused, histmax=0)
        barrier()
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=args.cg_iters, fused=fused,
                         check_every=args.cg_iters, histmax=0)
        t = maxr(r["solve_ms"])
        ksp[label] = {"iter_per_s": r["its"] / (t * 1e-3), "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"]}
    # CG roofline against B_cg,min = B_spmv + 88 n
    Bcg = Btot + 88.0 * m * world
    ksp["fused"]["gbs_vs_Bcg_min"] = Bcg / (ksp["fused"]["ms_per_iter"] * 1e-3) / 1e9
    clk = clocks.stop() if rank == 0 else None

    # ---------------- end to end: host vectors through the public API ----------------
    hx, hy = pb.pinned_array(m), pb.pinned_array(m)
    hx[:] = xh
    for _ in range(2):
        x.set_array(hx); A.mult(x, y); y.get_array(hy)
    barrier()
    t0 = time.perf_counter()
    pb.timer_start()
    for _ in range(args.steps):
        x.set_array(hx); A.mult(x, y); y.get_array(hy)
    ms_e2e = pb.timer_stop()
    wall_e2e = (time.perf_counter() - t0) * 1e3
    ms_e2e = maxr(max(ms_e2e, wall_e2e)) / args.steps
    e2e_val = Btot / (ms_e2e * 1e-3) / 1e9
    checksum = float(np.abs(hy).sum())
    e2e_mode = "serial: H2D x, MatMult, D2H y on one stream"
    e2e_serial = {"value": e2e_val, "ms_per_step": ms_e2e}
    if world == 1:
        # the same calls double-buffered on two caller-owned streams (b200_stream_create / b200_set_stream): every step still
        # uploads its x and reads back its y, but step k+1's upload overlaps step k's download (PCIe is full duplex)
        S = [pb.Stream(), pb.Stream()]
        xd, yd, hyb = [x, pb.Vec(m)], [y, pb.Vec(m)], [hy, pb.pinned_array(m)]

        def piped(nsteps):
            for k in range(nsteps):
                q = k & 1
                S[q].use()
                xd[q].set_array_async(hx); A.mult(xd[q], yd[q]); yd[q].get_array_async(hyb[q])
            S[0].synchronize(); S[1].synchronize()
        piped(4)
        t0 = time.perf_counter()
        piped(args.steps)
        ms_pipe = (time.perf_counter() - t0) * 1e3 / args.steps
        pb.use_default_stream()
        assert float(np.abs(hyb[(args.steps - 1) & 1]).sum()) == checksum, "double-buffered result differs"
        e2e_val, ms_e2e = Btot / (ms_pipe * 1e-3) / 1e9, ms_pipe
        e2e_mode = "double-buffered on two streams: every step uploads x and downloads y, consecutive steps overlap (wall clock)"
        del xd, yd
        [st.destroy() for st in S]

    # ---------------- optional: KSPCG strong scaling on a FIXED global grid (north_star: 512^3 7-pt, 1 -> 8 GPUs) ----------------
    strong = None
    if args.strong_nx:
        del A, x, y, one, b, sol, hx, hy
        snx = args.strong_nx
        SN = snx ** 3
        if world == 1:
            ip, j, a = P.laplace3d_7pt(snx)
            SA = pb.Mat.from_csr(SN, SN, ip, j, a)
            sm, sB = SN, spmv_bytes(len(a), SN, SN)
            del ip, j, a
        else:
            sr = par.split_ownership(SN, world)
            ai, aj, aa = P.laplace3d_7pt(snx, r0=int(sr[rank]), r1=int(sr[rank + 1]))
            sh = par.setup_mpiaij_host(rank, world, sr, ai - ai[0], aj, aa, SN, dist=dist, device="cuda")
            SA = par.create_mpiaij(sh)
            sm, sB = sh["m"], SA.bytes()
            del ai, aj, aa, sh
        so, sb, ss = pb.Vec(sm), pb.Vec(sm), pb.Vec(sm)
        so.set(1.0)
        SA.mult(so, sb)
        pb.ksp_solve(SA, sb, ss, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=5, fused=1, histmax=0)
        barrier()
        r = pb.ksp_solve(SA, sb, ss, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=args.cg_iters, fused=1,
                         check_every=args.cg_iters, histmax=0)
        t = maxr(r["solve_ms"])
        for _ in range(3):
            SA.mult(so, sb)
        barrier()
        pb.timer_start()
        for _ in range(20):
            SA.mult(so, sb)
        tm = maxr(pb.timer_stop()) / 20
        sBt = sB
        if dist is not None:
            import torch
            tt = torch.tensor([sB], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt)
            sBt = float(tt.item())
        strong = {"grid": "%d^3 7-pt, fixed total size" % snx, "rows": SN, "iter_per_s": r["its"] / (t * 1e-3),
                  "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"], "matmult_ms": tm,
                  "matmult_gbs_total": sBt / (tm * 1e-3) / 1e9,
                  "cg_gbs_vs_Bcg_min": (sBt + 88.0 * SN) / (t / max(r["its"], 1) * 1e-3) / 1e9}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None                       # dram bytes of the dominant kernel from the committed `ncu --set full` capture
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["spmv_tile_kernel_lap7_256"]["dram_bytes"]
    except Exception:
        pass
    per_gpu = (B_alg / (ms_step * 1e-3) / 1e9)
    if rank != 0:
        if dist is not None:
            dist.barrier(); dist.destroy_process_group()
        return
    line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": "configs[1]: 3D 7-point Laplacian %d^3 per GPU, AIJ MatMult (+KSPCG/Jacobi)" % nx,
                       "rows_per_gpu": m, "nnz_per_gpu": int(nz), "parallelism": "row slabs x%d (MPIAIJ diag+offdiag)" % world,
                       "transport": "none (1 rank)" if world == 1 else ("NVLink peer stores, own kernels" if args.transport == "p2p" else "NCCL send/recv + allreduce"),
                       "l2_policy": "inputs (1.7 GB per GPU) larger than the 126 MB L2; no flush"},
            "gflops": (2.0 * nz * world - m * world) / (ms_step * 1e-3) / 1e9,
            "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s", "frac": per_gpu / peak,
                         "traffic": traffic, "kernel": "spmv_tile_kernel<G=1,PARK=0,ADD=0,DOT=0,256,2048,1>",
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "algorithmic_bytes_per_launch": B_alg},
            "kspcg": ksp,
            **({"kspcg_strong": strong} if strong else {}),
            "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": 8 * m * world, "d2h_bytes_per_step": 8 * m * world,
                    "ms_per_step": ms_e2e, "checksum": checksum, "mode": e2e_mode, "serial": e2e_serial},
            "gpu_launches": int(launches), "clocks": clk, "setup_s": t_setup}
    # CPU baseline on rank 0 at N = 1 only: the reference library on a bounded sample of the same workload
    if world == 1 and not args.no_cpu_baseline:
        try:
            gbs_all, cores, t_all = reference_all_cores(nx, reps=5, warm=1)
            line["cpu_baseline"] = {"value": gbs_all, "unit": "GB/s", "cores": cores, "kind": "reference", "ms_per_step": t_all * 1e3,
                                    "sample": "unmodified reference MatMult_SeqAIJ (oracle/_ref), full %d^3 7-pt matrix row-split "
                                              "over %d host processes (one per core), 5 reps after 1 warm-up" % (nx, cores)}
        except Exception as e:                       # the checker is absent: report it, never substitute
            line["cpu_baseline"] = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % e}
    print(json.dumps(line))
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--nx", type=int, default=256)
    ap.add_argument("--cg-iters", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strong-nx", type=int, default=0,
                    help="also time KSPCG on a FIXED nx^3 7-pt grid split over the ranks (strong scaling; north_star uses 512)")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: ghost exchange + scalar allreduces over NVLink peer memory (own kernels) or NCCL")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
