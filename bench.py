#!/usr/bin/env python3
"""bench.py — the headline measurement: AIJ MatMult GB/s (and KSPCG iterations/s) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--nx 256] [--cg-iters 100] [--strong-nx 512]

A step is one MatMult of BASELINE.json configs[1]: the 3-D 7-point Laplacian with 256^3 rows per GPU (weak scaling: N GPUs
hold a 256 x 256 x 256N grid, row-partitioned into z-slabs exactly as PetscSplitOwnership does, ghost planes exchanged
GPU to GPU).  `value` = algorithmic bytes of the global MatMult (12 B/nnz + 4 B/row + 8 B/row y + 8 B/col
x) x K / device time.  Inputs are 1.7 GB per GPU, far larger than the 126 MB L2, so no flush is needed between steps.
Beside it, in the same JSON line:
  parity      every rank's y against the UNMODIFIED reference kernels on the same arrays (MatMult_SeqAIJ on the diagonal
              block, then MatMultAdd_SeqAIJ on the off-diagonal block, mpiaij.c:1111-1114) — bit-exact — and, at N = 1, the
              first 20 CG residual norms against the reference's KSPSolve_CG; a mismatch ends the run with a non-zero status
  kspcg       KSPCG + PCJACOBI iterations/s on the same operator (fused device driver, and op by op)
  kspcg_strong  the same on a FIXED 512^3 grid split over the N GPUs (north_star's scaling target)
  e2e         the MatMult through the public API with HOST vectors (pinned H2D of x, D2H of y inside the timed region)
  e2e_ksp     host b -> KSPSolve(rtol 1e-8) -> host x
`--impl reference` runs the reference's own CPU implementation of the same global problem: the unmodified reference
library (oracle/_ref_mpi) driven by oracle/ref_bench.c on P = all host cores under the b200mpi launcher — MatMult_MPIAIJ
with its real ghost exchange and KSPSolve_CG.  That arm imports nothing of petsc-dev_b200 and loads none of its libraries.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
METRIC = "AIJ MatMult GB/s (7-pt Laplacian 256^3 per GPU); KSPCG iter/s beside it"
MPIEXEC = os.path.join(ROOT, "petsc-dev_b200", "lib", "b200mpiexec")
REF_BENCH = os.path.join(ROOT, "oracle", "_ref_mpi", "bin", "b200_ref_bench")
PLUGIN_MPI = os.path.join(ROOT, "petsc-dev_b200", "lib", "mpi", "libpetscb200plugin.so")


def spmv_bytes(nz, m, n):
    return 12.0 * nz + 4.0 * (m + 1) + 8.0 * n + 8.0 * m


def lap7_nnz(nx, ny, nz):
    return 7 * nx * ny * nz - 2 * (nx * ny + ny * nz + nx * nz)


def workload_config(nx, gpus):
    """The SAME dictionary in both arms: what is multiplied, not how."""
    g = [nx, nx, nx * gpus]
    N = g[0] * g[1] * g[2]
    return {"workload": "configs[1]: 3D 7-point Laplacian, %d^3 rows per GPU, AIJ MatMult (+KSPCG/Jacobi)" % nx,
            "grid": g, "rows": N, "nnz": lap7_nnz(*g), "gpus": gpus,
            "l2_policy": "inputs (1.7 GB per 256^3 rows) larger than the 126 MB L2; no flush"}


class ClockSampler:
    """Samples SM clocks and throttle reasons during the timed region (nvidia-smi, 200 ms)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw"

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
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


def host_cores():
    return len(os.sched_getaffinity(0))


# ---------------------------------------------------------------------------------------------------- reference arm
def run_ref_bench(grid, ranks, reps, warm, cg_its, solve_rtol=0.0, extra=(), timeout=3000):
    """The reference's own code on `ranks` host processes: oracle/_ref_mpi (unmodified sources built against the b200mpi
    subset) driven by oracle/ref_bench.c.  Returns the JSON it prints."""
    if not (os.path.exists(MPIEXEC) and os.path.exists(REF_BENCH)):
        raise RuntimeError("oracle/_ref_mpi is not built (oracle/build_ref.py --mpi, petsc-dev_b200/plugin/build_plugin.py)")
    cmd = [MPIEXEC, "-n", str(ranks), REF_BENCH, "-nx", str(grid[0]), "-ny", str(grid[1]), "-nz", str(grid[2]),
           "-reps", str(reps), "-warm", str(warm), "-cg_its", str(cg_its)]
    if solve_rtol:
        cmd += ["-solve_rtol", repr(solve_rtol)]
    r = subprocess.run(cmd + list(extra), capture_output=True, text=True, timeout=timeout)
    if r.returncode:
        raise RuntimeError("reference bench failed (rc %d): %s %s" % (r.returncode, r.stdout[-1500:], r.stderr[-1500:]))
    for line in r.stdout.splitlines():
        if line.startswith("{"):
            return json.loads(line)
    raise RuntimeError("reference bench printed no result: %s" % r.stdout[-1500:])


def run_reference(args):
    rank, _, world = env_rank()
    if rank != 0:
        return
    gpus = max(args.gpus, world)
    cfg = workload_config(args.nx, gpus)
    cores = host_cores()
    N, nnz = cfg["rows"], cfg["nnz"]
    B = spmv_bytes(nnz, N, N)
    # KSP to rtol 1e-8 on the all-cores arm only for the 1-GPU problem (at N > 1 the global grid needs thousands of
    # iterations of ~0.2 s each on the host cores); the fixed-iteration rate is reported at every N
    d = run_ref_bench(cfg["grid"], cores, reps=max(args.steps, 1), warm=max(args.warmup, 1), cg_its=20,
                      solve_rtol=1e-8 if (gpus == 1 and not args.no_solve) else 0.0)
    one = run_ref_bench([args.nx, args.nx, args.nx], 1, reps=3, warm=1, cg_its=0)
    gbs = B / d["matmult_mean_s"] / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": d["matmult_mean_s"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "per_launch": {"min_ms": d["matmult_min_s"] * 1e3, "median_ms": d["matmult_median_s"] * 1e3},
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "the whole %dx%dx%d problem: unmodified reference MatMult_MPIAIJ (ghost exchange included) on %d "
                                       "MPI ranks (b200mpi, one per host core), %d reps after %d warm-ups"
                                       % (cfg["grid"][0], cfg["grid"][1], cfg["grid"][2], cores, args.steps, args.warmup),
                             "one_core": {"value": spmv_bytes(one["nnz"], one["rows"], one["rows"]) / one["matmult_mean_s"] / 1e9,
                                          "ms_per_step": one["matmult_mean_s"] * 1e3, "rows": one["rows"]}},
            "kspcg": {"iter_per_s": d["cg_its"] / d["cg_s"] if d["cg_s"] else None, "iterations": d["cg_its"], "cores": cores,
                      "note": "reference KSPSolve_CG + PCJACOBI on %d ranks" % cores},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "y_norm1": d["y_norm1"], "assembly_s": d["assembly_s"]}
    if d.get("solve_its"):
        line["e2e_ksp"] = {"rtol": 1e-8, "its": d["solve_its"], "seconds": d["solve_s"], "iter_per_s": d["solve_its"] / d["solve_s"],
                           "h2d_bytes": 0, "d2h_bytes": 0}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------- B200 arm
class Ranks:
    """torch.distributed plumbing of the B200 arm: NCCL for device-side maxima, a gloo group for CPU-side barriers (an
    NCCL barrier would keep a spinning kernel on every GPU while rank 0 runs host-side checks)."""

    def __init__(self, rank, local_rank, world):
        self.rank, self.world, self.dist, self.cpu = rank, world, None, None
        if world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            self.dist, self.torch = dist, torch
            self.cpu = dist.new_group(backend="gloo")

    def barrier(self, pb):
        pb.synchronize()
        if self.dist is not None:
            self.dist.barrier()

    def cpu_barrier(self):
        if self.dist is not None:
            self.dist.barrier(group=self.cpu)

    def reduce(self, v, op="max"):
        """v: float or 1-D array -> the same, reduced over ranks."""
        if self.dist is None:
            return v
        t = self.torch.tensor(np.atleast_1d(np.asarray(v, np.float64)), dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op={"max": self.dist.ReduceOp.MAX, "min": self.dist.ReduceOp.MIN, "sum": self.dist.ReduceOp.SUM}[op])
        out = t.cpu().numpy()
        return float(out[0]) if np.isscalar(v) or np.ndim(v) == 0 else out

    def close(self):
        if self.dist is not None:
            self.dist.barrier()
            self.dist.destroy_process_group()


def build_operator(pb, par, P, R, nx, ny, nz, keep_host=False):
    """The 7-point Laplacian of an nx x ny x nz grid, rows split over the ranks like PetscSplitOwnership.
    Returns (A, m, nnz_local, bytes_per_rank, host) — host = what the parity check needs (None unless keep_host)."""
    N = nx * ny * nz
    if R.world == 1:
        ip, j, a = P.laplace3d_7pt(nx, ny, nz)
        A = pb.Mat.from_csr(N, N, ip, j, a)
        return A, N, len(a), spmv_bytes(len(a), N, N), ((ip, j, a) if keep_host else None)
    ranges = par.split_ownership(N, R.world)
    ai, aj, aa = P.laplace3d_7pt(nx, ny, nz, r0=int(ranges[R.rank]), r1=int(ranges[R.rank + 1]))
    host = par.setup_mpiaij_host(R.rank, R.world, ranges, ai - ai[0], aj, aa, N, dist=R.dist, device="cuda")
    A = par.create_mpiaij(host)
    return A, host["m"], len(aa), A.bytes(), (host if keep_host else None)


def reference_product(P, m, rank_r0, xh, seq_csr=None, host=None, seed=2):
    """y of this rank from the UNMODIFIED reference kernels (oracle/_ref): MatMult_SeqAIJ, resp. MatMult_SeqAIJ on the
    diagonal block followed by MatMultAdd_SeqAIJ on the off-diagonal block with the ghost values (mpiaij.c:1111-1114)."""
    from oracle import refpetsc as ORA
    if seq_csr is not None:
        ip, j, a = seq_csr
        return ORA.RefMat(m, m, ip, j, a, no_inode=True).mult(xh)
    sp = host["split"]
    yd = ORA.RefMat(m, m, sp["Ai"], sp["Aj"], sp["Aa"], no_inode=True).mult(xh)
    g = sp["garray"]
    if not len(g):
        return yd
    lvec = P.random_vector_at(g, seed)
    return ORA.RefMat(m, len(g), sp["Bi"], sp["Bj"], sp["Ba"], no_inode=True).multadd(lvec, yd)


def hist_dev(h, hr):
    k = min(len(h), len(hr))
    h, hr = np.asarray(h[:k]), np.asarray(hr[:k])
    return float(np.max(np.abs(h - hr) / hr)) if k else None


def plugin_on_ranks(n, grid, cg_its=20, ksp_type="cg", timeout=240):
    """The reference's C driver (oracle/ref_bench.c, PETSc API only) on n ranks, one GPU each, with the B200 Mat/Vec types
    loaded through -dll_prepend: the drop-in boundary exercised from plain C on more than one rank."""
    return run_ref_bench(grid, n, reps=5, warm=2, cg_its=cg_its, timeout=timeout,
                         extra=["-dll_prepend", PLUGIN_MPI, "-mat_type", "aijb200", "-vec_type", "b200", "-ksp_type", ksp_type])


def run_b200(args):
    rank, local_rank, world = env_rank()
    if world != args.gpus and world > 1:
        args.gpus = world
    pb = importlib.import_module("petsc-dev_b200")
    par = importlib.import_module("petsc-dev_b200.parallel")
    P = importlib.import_module("workloads.problems")
    R = Ranks(rank, local_rank, world)
    pb.init(local_rank)
    nx = args.nx
    cfg = workload_config(nx, world)
    transport = "none (1 rank)"
    if world > 1:
        par.comm_init(rank, world, R.dist, device="cuda")
        if args.transport == "p2p" and not par.enable_p2p(strict=False):
            args.transport = "nccl"          # agreed on by all ranks inside the library
        transport = "NVLink peer stores, own kernels" if args.transport == "p2p" else "NCCL send/recv + allreduce"
    t_setup = time.perf_counter()
    A, m, nz, B_rank, host = build_operator(pb, par, P, R, nx, nx, nx * world, keep_host=True)
    t_setup = time.perf_counter() - t_setup
    r0 = int(host["rstart"]) if world > 1 else 0
    xh = P.random_vector(m, 2, r0=r0)
    x, y = pb.Vec(m, xh), pb.Vec(m)
    B_glob = spmv_bytes(cfg["nnz"], cfg["rows"], cfg["rows"])

    # ---------------- MatMult (device-resident): W warm-ups, then EXACTLY K steps between barriers ----------------
    W = max(args.warmup, 3)
    for _ in range(W):
        A.mult(x, y)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()               # BEFORE the barrier: spawning the sampler costs rank 0 ~1 ms the other ranks would wait for
    R.barrier(pb)
    l0 = pb.launch_count()
    pb.timer_start()
    for _ in range(args.steps):
        A.mult(x, y)
    ms = pb.timer_stop()
    launches = pb.launch_count() - l0
    R.barrier(pb)
    ms_step = R.reduce(ms, "max") / args.steps
    value = B_glob / (ms_step * 1e-3) / 1e9
    # every launch on its own (device time between two events around ONE MatMult): min and median over the K steps
    per = np.zeros(args.steps)
    for k in range(args.steps):
        pb.timer_start(); A.mult(x, y); per[k] = pb.timer_stop()
    per = np.atleast_1d(R.reduce(per, "max"))
    clk = clocks.stop() if rank == 0 else None

    # ---------------- parity, outside every timed region ----------------
    A.mult(x, y)
    got = y.get_array()
    want = reference_product(P, m, r0, xh, seq_csr=host if world == 1 else None, host=host if world > 1 else None)
    exact = bool(np.array_equal(got, want))
    err = float(np.max(np.abs(got - want))) if m else 0.0
    parity = {"matmult_bit_exact": bool(R.reduce(1.0 if exact else 0.0, "min") == 1.0), "max_abs_err": R.reduce(err, "max"),
              "ranks_checked": world,
              "oracle": "unmodified reference MatMult_SeqAIJ (+ MatMultAdd_SeqAIJ on the off-diagonal block) on each rank's own arrays"}
    y_norm1 = R.reduce(float(np.abs(got).sum()), "sum")
    del got, want
    host = None

    # ---------------- KSPCG + Jacobi, fixed iteration count ----------------
    one, b, sol = pb.Vec(m), pb.Vec(m), pb.Vec(m)
    one.set(1.0)
    A.mult(one, b)
    ksp = {}
    for label, fused in (("fused", 1), ("op_by_op", 0)):
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=5, fused=fused, histmax=0)
        R.barrier(pb)
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=args.cg_iters, fused=fused,
                         check_every=args.cg_iters, histmax=0)
        t = R.reduce(r["solve_ms"], "max")
        ksp[label] = {"iter_per_s": r["its"] / (t * 1e-3), "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"]}
    Bcg = B_glob + 88.0 * cfg["rows"]                        # B_cg,min = B_spmv + 88 n
    ksp["fused"]["gbs_vs_Bcg_min"] = Bcg / (ksp["fused"]["ms_per_iter"] * 1e-3) / 1e9
    hist20 = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=20, fused=1)["history"]

    # ---------------- end to end: host vectors through the public API, double-buffered on two caller-owned streams ----------------
    hx, hy = pb.pinned_array(m), pb.pinned_array(m)
    hx[:] = xh
    for _ in range(2):
        x.set_array(hx); A.mult(x, y); y.get_array(hy)
    R.barrier(pb)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        x.set_array(hx); A.mult(x, y); y.get_array(hy)
    pb.synchronize()
    ms_serial = R.reduce((time.perf_counter() - t0) * 1e3, "max") / args.steps
    checksum = float(np.abs(hy).sum())
    S = [pb.Stream(), pb.Stream()]
    xd, yd, hyb = [x, pb.Vec(m)], [y, pb.Vec(m)], [hy, pb.pinned_array(m)]

    def piped(nsteps):
        # every step uploads its x and reads back its y; step k+1's upload overlaps step k's download (PCIe is full duplex)
        for k in range(nsteps):
            q = k & 1
            S[q].use()
            xd[q].set_array_async(hx); A.mult(xd[q], yd[q]); yd[q].get_array_async(hyb[q])
        S[0].synchronize(); S[1].synchronize()
    piped(4)
    R.barrier(pb)
    t0 = time.perf_counter()
    piped(args.steps)
    ms_e2e = R.reduce((time.perf_counter() - t0) * 1e3, "max") / args.steps
    pb.use_default_stream()
    e2e_ok = float(np.abs(hyb[(args.steps - 1) & 1]).sum()) == checksum
    e2e_val = B_glob / (ms_e2e * 1e-3) / 1e9
    del xd, yd
    [st.destroy() for st in S]

    # ---------------- end to end at the solve level: host b in, KSPSolve to rtol 1e-8, host x out ----------------
    e2e_ksp = None
    if not args.no_solve:
        hb = pb.pinned_array(m)
        b.get_array(hb)
        R.barrier(pb)
        t0 = time.perf_counter()
        b.set_array(hb)
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-8, max_it=100000, fused=1, check_every=32, histmax=0)
        sol.get_array(hx)
        secs = R.reduce(time.perf_counter() - t0, "max")
        xerr = R.reduce(float(np.max(np.abs(hx - 1.0))), "max")             # b = A * 1
        e2e_ksp = {"rtol": 1e-8, "its": r["its"], "reason": r["reason"], "seconds": secs, "iter_per_s": r["its"] / secs,
                   "h2d_bytes": 8 * cfg["rows"], "d2h_bytes": 8 * cfg["rows"], "max_abs_error_of_x": xerr}
    del A, x, y, one, b, sol, hx, hy

    # ---------------- KSPCG strong scaling on a FIXED global grid (north_star: 512^3 7-pt, 1 -> 8 GPUs) ----------------
    strong = None
    if args.strong_nx:
        snx = args.strong_nx
        SN = snx ** 3
        SA, sm, _, _, _ = build_operator(pb, par, P, R, snx, snx, snx)
        so, sb, ss = pb.Vec(sm), pb.Vec(sm), pb.Vec(sm)
        so.set(1.0)
        SA.mult(so, sb)
        pb.ksp_solve(SA, sb, ss, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=5, fused=1, histmax=0)
        R.barrier(pb)
        r = pb.ksp_solve(SA, sb, ss, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=args.cg_iters, fused=1,
                         check_every=args.cg_iters, histmax=0)
        t = R.reduce(r["solve_ms"], "max")
        for _ in range(3):
            SA.mult(so, sb)
        R.barrier(pb)
        pb.timer_start()
        for _ in range(20):
            SA.mult(so, sb)
        tm = R.reduce(pb.timer_stop(), "max") / 20
        sB = spmv_bytes(lap7_nnz(snx, snx, snx), SN, SN)
        strong = {"grid": "%d^3 7-pt, fixed total size" % snx, "rows": SN, "iter_per_s": r["its"] / (t * 1e-3),
                  "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"], "matmult_ms": tm,
                  "matmult_gbs_total": sB / (tm * 1e-3) / 1e9,
                  "cg_gbs_vs_Bcg_min": (sB + 88.0 * SN) / (t / max(r["its"], 1) * 1e-3) / 1e9}
        del SA, so, sb, ss

    # ---------------- CG history against the reference's KSPSolve_CG ----------------
    # N = 1: the full 256^3 problem, reference = the all-cores CPU baseline run below.  N > 1: a 64 x 64 x 64N grid on the N
    # GPUs against the reference on N host ranks (the full-size global solve on the host cores belongs to --impl reference).
    cpu_baseline, plugin_np = None, None
    if world > 1:
        g = [64, 64, 64 * world]
        SA, sm, _, _, _ = build_operator(pb, par, P, R, *g)
        so, sb, ss = pb.Vec(sm), pb.Vec(sm), pb.Vec(sm)
        so.set(1.0); SA.mult(so, sb)
        hsmall = pb.ksp_solve(SA, sb, ss, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=20, fused=1)["history"]
        del SA, so, sb, ss
        pb.synchronize()
        R.cpu_barrier()
        if rank == 0:
            try:
                ref = run_ref_bench(g, world, reps=2, warm=1, cg_its=20, timeout=600)
                parity["cg_hist_dev"] = hist_dev(hsmall, ref["hist"])
                parity["cg_problem"] = "%dx%dx%d on %d GPUs vs reference KSPSolve_CG on %d host ranks" % (g[0], g[1], g[2], world, world)
                # the same C program on the B200 types, one rank per GPU, through -dll_prepend (the other bench ranks are idle)
                if os.path.exists(PLUGIN_MPI) and not args.no_plugin:
                    plugin_np = {}
                    for kt in ("cg", "cgb200"):
                        try:
                            pr = plugin_on_ranks(world, g, ksp_type=kt)
                            plugin_np[kt] = {"mat_type": pr["mat_type"], "ranks": pr["ranks"], "cg_its": pr["cg_its"],
                                             "hist_dev_vs_reference": hist_dev(pr["hist"], ref["hist"]),
                                             "y_norm1_rel_diff": abs(pr["y_norm1"] - ref["y_norm1"]) / ref["y_norm1"],
                                             "matmult_median_ms": pr["matmult_median_s"] * 1e3, "iter_per_s": pr["cg_its"] / pr["cg_s"]}
                        except Exception as e:
                            plugin_np[kt] = {"error": str(e)[-800:]}
            except Exception as e:
                parity["cg_hist_dev"] = None
                parity["cg_error"] = str(e)[-800:]
        R.cpu_barrier()
    elif not args.no_cpu_baseline:
        # CPU baseline on rank 0 at N = 1 only: the reference library on all host cores, the whole 256^3 problem, 5 reps
        try:
            cores = host_cores()
            ref = run_ref_bench(cfg["grid"], cores, reps=5, warm=1, cg_its=20)
            cpu_baseline = {"value": B_glob / ref["matmult_mean_s"] / 1e9, "unit": "GB/s", "cores": cores, "kind": "reference",
                            "ms_per_step": ref["matmult_mean_s"] * 1e3, "kspcg_iter_per_s": ref["cg_its"] / ref["cg_s"],
                            "sample": "unmodified reference MatMult_MPIAIJ / KSPSolve_CG (oracle/_ref_mpi), the whole %d^3 7-pt problem on %d "
                                      "MPI ranks (b200mpi, one per host core), 5 MatMult reps after 1 warm-up, 20 CG iterations" % (nx, cores)}
            parity["cg_hist_dev"] = hist_dev(hist20, ref["hist"])
            parity["cg_problem"] = "%d^3 on 1 GPU vs reference KSPSolve_CG on %d host ranks" % (nx, cores)
            parity["y_norm1_rel_diff_vs_reference_mpiaij"] = abs(y_norm1 - ref["y_norm1"]) / ref["y_norm1"]
        except Exception as e:                       # the checker is absent: report it, never substitute
            cpu_baseline = {"value": None, "unit": "GB/s", "cores": 0, "kind": "reference", "sample": "unavailable: %s" % str(e)[-600:]}
    if "cg_hist_dev" not in parity and world == 1:
        from oracle import refpetsc as ORA           # no all-cores run: the 1-rank reference library
        ip, j, a = P.laplace3d_7pt(nx)
        rm = ORA.RefMat(cfg["rows"], cfg["rows"], ip, j, a, no_inode=True)
        parity["cg_hist_dev"] = hist_dev(hist20, rm.ksp_solve(rm.mult(np.ones(cfg["rows"])), "cg", "jacobi", rtol=1e-50, maxit=20)["history"])
        parity["cg_problem"] = "%d^3 on 1 GPU vs reference KSPSolve_CG on 1 rank" % nx
    parity["ok"] = bool(parity["matmult_bit_exact"] and e2e_ok and (parity.get("cg_hist_dev") is None or parity["cg_hist_dev"] <= 1e-10)
                        and not (world > 1 and rank == 0 and parity.get("cg_hist_dev") is None))

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic, traffic_src = None, None      # dram bytes of the dominant kernel from the committed `ncu --set full` capture
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))["spmv_tile_kernel_lap7_256"]
        traffic, traffic_src = tj["dram_bytes"], tj.get("source", "profiles/ncu_traffic.json")
    except Exception:
        pass
    per_gpu = B_rank / (float(np.median(per)) * 1e-3) / 1e9
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": W,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": cfg,
                "impl_detail": {"rows_per_gpu": m, "nnz_per_gpu": int(nz), "parallelism": "row slabs x%d (MPIAIJ diag+offdiag)" % world,
                                "transport": transport},
                "per_launch": {"min_ms": float(per.min()), "median_ms": float(np.median(per)), "max_ms": float(per.max()), "n": int(per.size)},
                "gflops": (2.0 * cfg["nnz"] - cfg["rows"]) / (ms_step * 1e-3) / 1e9,
                "roofline": {"bound": "hbm", "achieved": per_gpu, "peak": peak, "unit": "GB/s", "frac": per_gpu / peak,
                             "traffic": traffic, "traffic_source": traffic_src,
                             "kernel": ("spmv_tile_kernel<G=1,PARK=0,ADD=0,DOT=0,256,2048,1>" if (world == 1 or not transport.startswith("NVLink")) else
                                        "spmv_mpiaij_kernel<G=1,ADD=0,DOT=0,256,2048> (MatMult_MPIAIJ in one launch: ghost exchange, diagonal block, boundary rows)"),
                             "frac_of_nominal_8TBs": per_gpu / 8000.0,
                             "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                             "algorithmic_bytes_per_launch": B_rank, "launch_ms": float(np.median(per)),
                             "how": "per-GPU algorithmic bytes / median device time of one MatMult (CUDA events on the compute stream)"},
                "parity": parity, "kspcg": ksp,
                **({"kspcg_strong": strong} if strong else {}),
                "e2e": {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": 8 * cfg["rows"], "d2h_bytes_per_step": 8 * cfg["rows"],
                        "ms_per_step": ms_e2e, "checksum": checksum, "result_identical": bool(e2e_ok),
                        "mode": "double-buffered on two caller-owned streams: every step uploads x and downloads y, consecutive steps overlap (wall clock, max over ranks)",
                        "serial": {"value": B_glob / (ms_serial * 1e-3) / 1e9, "ms_per_step": ms_serial}},
                **({"e2e_ksp": e2e_ksp} if e2e_ksp else {}),
                **({"plugin_on_ranks": plugin_np} if plugin_np else {}),
                "y_norm1": y_norm1, "gpu_launches": int(launches), "clocks": clk, "setup_s": t_setup}
        if cpu_baseline:
            line["cpu_baseline"] = cpu_baseline
        print(json.dumps(line))
        sys.stdout.flush()
    ok = R.reduce(1.0 if (rank != 0 or parity["ok"]) else 0.0, "min") == 1.0
    R.close()
    if not ok:
        sys.stderr.write("bench.py: PARITY FAILURE: %s\n" % json.dumps(parity))
        sys.exit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--nx", type=int, default=256)
    ap.add_argument("--cg-iters", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-solve", action="store_true", help="skip the solve-level end-to-end legs (KSPSolve to rtol 1e-8)")
    ap.add_argument("--no-plugin", action="store_true", help="N > 1: skip the C-program-on-N-ranks check of the plugin")
    ap.add_argument("--strong-nx", type=int, default=512,
                    help="KSPCG on a FIXED nx^3 7-pt grid split over the ranks (strong scaling; north_star uses 512); 0 = skip")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: ghost exchange + scalar allreduces over NVLink peer memory (own kernels) or NCCL")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
