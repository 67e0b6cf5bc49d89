"""The reference's own example programs (compiled UNMODIFIED into oracle/_ref/bin by petsc-dev_b200/plugin/build_plugin.py)
and the option sets the reference's makefiles run them with (file:line of each `runex*` target), used three ways:
  reference arm : <prog> args + ref_args            -> the reference's CPU Mat/Vec types
  golden        : the reference's committed output file, where one exists
  B200 arm      : <prog> args + -dll_prepend libpetscb200plugin.so + b200_args -> the new types, nothing else changed
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
PLUGIN = os.path.join(ROOT, "petsc-dev_b200", "lib", "libpetscb200plugin.so")
B200 = ["-mat_type", "aijb200", "-vec_type", "b200"]

CASES = {
    # src/ksp/ksp/examples/tutorials/makefile:317 (GMRES + ILU(0): the factorisation and MatSolve stay host code and
    # reach the device vectors through VecGetArray -> exercises the coherence seam)
    "ksp_ex2_1": dict(prog="ksp_ex2", args=["-ksp_monitor_short", "-m", "5", "-n", "5", "-ksp_gmres_cgs_refinement_type", "refine_always"],
                      b200_args=B200, golden="src/ksp/ksp/examples/tutorials/output/ex2_1.out"),
    # BASELINE.json configs[0] at a size the CPU finishes at once: ex2 with KSPCG + PCJACOBI
    "ksp_ex2_cg_jacobi": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "40", "-n", "37", "-ksp_type", "cg", "-pc_type", "jacobi"],
                              b200_args=B200, history=True),
    "ksp_ex2_gmres_jacobi": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "33", "-n", "41", "-ksp_type", "gmres", "-pc_type", "jacobi", "-ksp_gmres_restart", "30"],
                                 b200_args=B200, history=True),
    "ksp_ex2_cg_random": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "25", "-n", "25", "-ksp_type", "cg", "-pc_type", "jacobi", "-random_exact_sol"],
                              b200_args=B200, history=True),
    "ksp_ex2_mpiaij": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "30", "-n", "30", "-ksp_type", "cg", "-pc_type", "jacobi"],
                           ref_args=["-mat_type", "mpiaij", "-vec_type", "mpi"], b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200"], history=True),
    # the device-resident Krylov drivers as new KSP types (-ksp_type cgb200 / gmresb200) against the reference's KSPCG / KSPGMRES
    "ksp_ex2_cgb200": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "40", "-n", "37", "-pc_type", "jacobi"],
                           ref_args=["-ksp_type", "cg"], b200_args=B200 + ["-ksp_type", "cgb200"], history=True),
    "ksp_ex2_cgb200_opbyop": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "31", "-n", "29", "-pc_type", "jacobi", "-random_exact_sol"],
                                  ref_args=["-ksp_type", "cg"], b200_args=B200 + ["-ksp_type", "cgb200", "-ksp_b200_fused", "0"], history=True),
    "ksp_ex2_cgb200_mpiaij": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "30", "-n", "30", "-pc_type", "none"],
                                  ref_args=["-ksp_type", "cg", "-mat_type", "mpiaij", "-vec_type", "mpi"],
                                  b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200", "-ksp_type", "cgb200"], history=True),
    "ksp_ex2_gmresb200": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "33", "-n", "41", "-pc_type", "jacobi", "-ksp_gmres_restart", "30"],
                              ref_args=["-ksp_type", "gmres"], b200_args=B200 + ["-ksp_type", "gmresb200"], history=True),
    "ksp_ex2_gmresb200_mgs": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "21", "-n", "23", "-pc_type", "jacobi", "-ksp_gmres_modifiedgramschmidt"],
                                  ref_args=["-ksp_type", "gmres"], b200_args=B200 + ["-ksp_type", "gmresb200"], history=True),
    "ksp_ex2_gmresb200_refine": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "19", "-n", "27", "-pc_type", "jacobi", "-ksp_gmres_cgs_refinement_type", "refine_always"],
                                     ref_args=["-ksp_type", "gmres"], b200_args=B200 + ["-ksp_type", "gmresb200"], history=True),
    # PCJACOBI variants (jacobi.c:125-206): the device drivers take the inverse diagonal from the PC itself
    "ksp_ex2_cgb200_rowsum": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "23", "-n", "19", "-pc_type", "jacobi", "-pc_jacobi_rowsum"],
                                  ref_args=["-ksp_type", "cg"], b200_args=B200 + ["-ksp_type", "cgb200"], history=True),
    "ksp_ex2_gmresb200_rowmax": dict(prog="ksp_ex2", args=["-ksp_monitor", "-m", "17", "-n", "22", "-pc_type", "jacobi", "-pc_jacobi_rowmax", "-pc_jacobi_abs"],
                                     ref_args=["-ksp_type", "gmres"], b200_args=B200 + ["-ksp_type", "gmresb200"], history=True),
    # src/ksp/ksp/examples/tutorials/makefile:410 — two systems, same nonzero structure, different values: the second
    # assembly reaches the device as a value-only refresh (no re-binning, no new exchange plan)
    "ksp_ex5_1": dict(prog="ksp_ex5", args=["-pc_type", "jacobi", "-ksp_monitor_short", "-ksp_gmres_cgs_refinement_type", "refine_always"],
                      b200_args=B200, golden="src/ksp/ksp/examples/tutorials/output/ex5_1.out"),
    "ksp_ex5_mpiaij": dict(prog="ksp_ex5", args=["-pc_type", "jacobi", "-ksp_monitor_short", "-ksp_gmres_cgs_refinement_type", "refine_always", "-m", "7"],
                           ref_args=["-mat_type", "mpiaij", "-vec_type", "mpi"], b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200"]),
    # MatSetValuesBatch on the parallel type (tests/c/batch_mpiaij.c: ex4's mesh through MatCreate/MatSetType(MATMPIAIJ))
    "batch_mpiaij": dict(prog="b200_batch_mpiaij", args=["-X", "6", "-Y", "5"], ref_args=["-mat_type", "mpiaij"],
                         b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200"]),
    # tests/c/zerorows_refresh.c: MatZeroRows rewrites column indices in place, row offsets and nz unchanged
    "zerorows_seqaij": dict(prog="b200_zerorows_refresh", args=[], ref_args=["-mat_type", "seqaij"], b200_args=["-mat_type", "seqaijb200", "-vec_type", "seqb200"]),
    # (the parallel type's MatZeroRows sends every row to its owner, itself included, which MPIUNI cannot do: that case runs
    # in the MPI flavour, tests/test_gpu_plugin_mpi.py)
    # src/mat/examples/tests/makefile:753,774,795
    "mat_ex5_11_A": dict(prog="mat_ex5", args=["-rectA"], ref_args=["-mat_type", "seqaij"], b200_args=["-mat_type", "seqaijb200", "-vec_type", "seqb200"],
                         golden="src/mat/examples/tests/output/ex5_11_A.out"),
    "mat_ex5_21": dict(prog="mat_ex5", args=[], ref_args=["-mat_type", "mpiaij"], b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200"],
                       golden="src/mat/examples/tests/output/ex5_21.out"),
    "mat_ex5_31": dict(prog="mat_ex5", args=["-test_diagonalscale"], ref_args=["-mat_type", "mpiaij"], b200_args=["-mat_type", "mpiaijb200", "-vec_type", "mpib200"],
                       golden="src/mat/examples/tests/output/ex5_31.out"),
    # src/vec/vec/examples/tutorials/makefile:142 (VecDot, VecMDot, VecScale, VecCopy, VecAXPY, VecAYPX, VecWAXPY,
    # VecPointwiseMult/Divide, VecMAXPY, VecNorm)
    "vec_tut_ex1": dict(prog="vec_tut_ex1", args=[], b200_args=["-vec_type", "b200"], golden="src/vec/vec/examples/tutorials/output/ex1_1.out"),
    # src/vec/vec/examples/tests/makefile:418 (VecMDot against VecDot)
    "vec_ex43": dict(prog="vec_ex43", args=[], b200_args=["-vec_type", "b200"], golden="src/vec/vec/examples/tests/output/ex43_1.out"),
    # strided ops and VecMax/VecMin on one rank (golden files are for 2 ranks -> compared with the reference arm only)
    "vec_ex21": dict(prog="vec_ex21", args=[], b200_args=["-vec_type", "b200"]),
}
