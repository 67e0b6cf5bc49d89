#!/bin/bash
# development tool: outputs of the failing multi-rank plugin cases, kept under gpurun_out/
o=gpurun_out; mkdir -p $o
X=petsc-dev_b200/lib/b200mpiexec; B=oracle/_ref_mpi/bin; P=petsc-dev_b200/lib/mpi/libpetscb200plugin.so; H=petsc-dev_b200/lib/mpi/libpetscb200scatterhook.so
A="-nx 24 -ny 20 -nz 8 -reps 2 -warm 1 -cg_its 20"
$X -n 1 $B/b200_ref_bench $A -dll_prepend $P -mat_type aijb200 -vec_type b200 -info > $o/dbg_refbench_np1.txt 2>&1; echo "refbench np1 rc=$?"
$X -n 2 $B/b200_ref_bench -nx 24 -ny 20 -nz 15 -reps 2 -warm 1 -cg_its 20 -dll_prepend $P -mat_type aijb200 -vec_type b200 -info > $o/dbg_refbench_np2.txt 2>&1; echo "refbench np2 rc=$?"
$X -n 2 $B/b200_zerorows_refresh -n 15 -dll_prepend $P -mat_type aijb200 -vec_type b200 -info > $o/dbg_zerorows_np2.txt 2>&1; echo "zerorows np2 rc=$?"
LD_PRELOAD=$H $X -n 2 $B/b200_scatter_cases -dll_prepend $P -vec_type b200 -info > $o/dbg_scatter_np2.txt 2>&1; echo "scatter np2 rc=$?"
LD_PRELOAD=$H $X -n 2 $B/b200_scatter_cases -dll_prepend $P -vec_type b200 -b200_p2p 0 > $o/dbg_scatter_np2_nccl.txt 2>&1; echo "scatter np2 nccl rc=$?"
$X -n 2 $B/b200_scatter_cases > $o/dbg_scatter_np2_ref.txt 2>&1
grep -c . $o/dbg_*.txt
