/* ref_bench.c — TEST/BENCH INFRASTRUCTURE: the CPU baseline of bench.py.  A plain PETSc program (public API only) that is
 * linked against the UNMODIFIED reference built in oracle/_ref_mpi and started with b200mpiexec -n P (P = host cores):
 * assembles the 7-point Laplacian of an nx x ny x nz grid as MATMPIAIJ with MatSetValues (rows split by
 * PetscSplitOwnership), times MatMult_MPIAIJ (mpiaij.c:1102-1116: real ghost exchange between the P ranks) and
 * KSPSolve_CG + PCJACOBI (cg.c:92-286), and prints ONE JSON line from rank 0.
 *
 *   b200mpiexec -n P b200_ref_bench -nx 256 -ny 256 -nz 256 -reps 20 -warm 2 -cg_its 20 [-solve_rtol 1e-8]
 *
 * x = the same hashed vector bench.py's GPU arm multiplies (workloads/problems.py: random_vector(N, 2)); b = A * 1.
 * With -mat_type aijb200 -vec_type b200 -dll_prepend <plugin> the same program runs on the B200 types (one rank per GPU). */
static char help[] = "CPU baseline: MatMult_MPIAIJ and KSPSolve_CG on the 7-point Laplacian.\n";

#include <petscksp.h>

static unsigned long long splitmix64(unsigned long long x)
{
  unsigned long long z = x + 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static double hashed_entry(unsigned long long seed, unsigned long long g)     /* workloads/problems.py: random_vector */
{
  unsigned long long h = splitmix64(seed * 0xD1342543DE82EF95ULL + g);
  h = splitmix64(h ^ (0ULL * 0xA0761D6478BD642FULL));
  return -1.0 + 2.0 * (((double)(h >> 11) + 0.5) * (1.0 / 9007199254740992.0));
}

static int cmp_double(const void *a, const void *b) { double x = *(const double *)a, y = *(const double *)b; return x < y ? -1 : (x > y); }

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc, char **args)
{
  Mat            A;
  Vec            x, y, b, u, one;
  KSP            ksp;
  PC             pc;
  PetscInt       nx = 64, ny = -1, nz = -1, N, rstart, rend, row, reps = 10, warm = 2, cg_its = 20, its = 0, k, nhist = 0, *dnz, *onz, m;
  PetscMPIInt    rank, size;
  PetscReal      solve_rtol = 0.0, hist[4096], ysum, ynrm;
  PetscScalar    *xa;
  double         *t, t0, tcg = 0.0, tsolve = 0.0, tasm;
  PetscInt       solve_its = 0;
  KSPConvergedReason reason;
  long long      nnz = 0, gnnz = 0;
  PetscErrorCode ierr;

  PetscInitialize(&argc, &args, (char *)0, help);
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD, &rank);CHKERRQ(ierr);
  ierr = MPI_Comm_size(PETSC_COMM_WORLD, &size);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-nx", &nx, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-ny", &ny, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-nz", &nz, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-reps", &reps, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-warm", &warm, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-cg_its", &cg_its, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetReal(PETSC_NULL, "-solve_rtol", &solve_rtol, PETSC_NULL);CHKERRQ(ierr);
  if (ny < 0) ny = nx;
  if (nz < 0) nz = nx;
  N = nx * ny * nz;

  t0 = MPI_Wtime();
  ierr = MatCreate(PETSC_COMM_WORLD, &A);CHKERRQ(ierr);
  m = PETSC_DECIDE;
  ierr = PetscSplitOwnership(PETSC_COMM_WORLD, &m, &N);CHKERRQ(ierr);           /* psplit.c:88 */
  ierr = MPI_Scan(&m, &rend, 1, MPIU_INT, MPI_SUM, PETSC_COMM_WORLD);CHKERRQ(ierr);
  rstart = rend - m;
  ierr = MatSetSizes(A, m, m, N, N);CHKERRQ(ierr);
  ierr = MatSetType(A, MATAIJ);CHKERRQ(ierr);
  ierr = MatSetFromOptions(A);CHKERRQ(ierr);
  ierr = PetscMalloc2(m, PetscInt, &dnz, m, PetscInt, &onz);CHKERRQ(ierr);
#define STENCIL(body) do { \
    const PetscInt ix = row % nx, iy = (row / nx) % ny, iz = row / (nx * ny); \
    PetscInt c; PetscScalar v; \
    if (iz > 0)      { c = row - nx * ny; v = -1.0; body; } \
    if (iy > 0)      { c = row - nx;      v = -1.0; body; } \
    if (ix > 0)      { c = row - 1;       v = -1.0; body; } \
                     { c = row;           v =  6.0; body; } \
    if (ix < nx - 1) { c = row + 1;       v = -1.0; body; } \
    if (iy < ny - 1) { c = row + nx;      v = -1.0; body; } \
    if (iz < nz - 1) { c = row + nx * ny; v = -1.0; body; } } while (0)
  for (row = rstart; row < rend; row++) {
    PetscInt d = 0, o = 0;
    STENCIL({ (void)v; if (c >= rstart && c < rend) d++; else o++; });
    dnz[row - rstart] = d; onz[row - rstart] = o; nnz += d + o;
  }
  ierr = MatMPIAIJSetPreallocation(A, 0, dnz, 0, onz);CHKERRQ(ierr);
  ierr = MatSeqAIJSetPreallocation(A, 0, dnz);CHKERRQ(ierr);
  ierr = PetscFree2(dnz, onz);CHKERRQ(ierr);
  for (row = rstart; row < rend; row++) {
    PetscInt    cols[7], n = 0;
    PetscScalar vals[7];
    STENCIL({ cols[n] = c; vals[n] = v; n++; });
    ierr = MatSetValues(A, 1, &row, n, cols, vals, INSERT_VALUES);CHKERRQ(ierr);
  }
  ierr = MatAssemblyBegin(A, MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatAssemblyEnd(A, MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatSetOption(A, MAT_SYMMETRIC, PETSC_TRUE);CHKERRQ(ierr);
  tasm = MPI_Wtime() - t0;
  ierr = MPI_Allreduce(&nnz, &gnnz, 1, MPI_LONG_LONG_INT, MPI_SUM, PETSC_COMM_WORLD);CHKERRQ(ierr);

  ierr = MatGetVecs(A, &x, &y);CHKERRQ(ierr);
  ierr = VecDuplicate(x, &b);CHKERRQ(ierr);
  ierr = VecDuplicate(x, &u);CHKERRQ(ierr);
  ierr = VecDuplicate(x, &one);CHKERRQ(ierr);
  ierr = VecGetArray(x, &xa);CHKERRQ(ierr);
  for (row = rstart; row < rend; row++) xa[row - rstart] = hashed_entry(2, (unsigned long long)row);
  ierr = VecRestoreArray(x, &xa);CHKERRQ(ierr);

  /* ---- MatMult: every repetition timed between barriers, slowest rank counts ---- */
  for (k = 0; k < warm; k++) { ierr = MatMult(A, x, y);CHKERRQ(ierr); }
  t = (double *)malloc(sizeof(double) * (size_t)(reps > 0 ? reps : 1));
  for (k = 0; k < reps; k++) {
    double t1, dt, dtmax;
    ierr = MPI_Barrier(PETSC_COMM_WORLD);CHKERRQ(ierr);
    t1 = MPI_Wtime();
    ierr = MatMult(A, x, y);CHKERRQ(ierr);
    dt = MPI_Wtime() - t1;
    ierr = MPI_Allreduce(&dt, &dtmax, 1, MPI_DOUBLE, MPI_MAX, PETSC_COMM_WORLD);CHKERRQ(ierr);
    t[k] = dtmax;
  }
  ierr = VecNorm(y, NORM_1, &ysum);CHKERRQ(ierr);
  ierr = VecNorm(y, NORM_2, &ynrm);CHKERRQ(ierr);

  /* ---- KSPCG + Jacobi, b = A * 1 ---- */
  ierr = VecSet(one, 1.0);CHKERRQ(ierr);
  ierr = MatMult(A, one, b);CHKERRQ(ierr);
  ierr = KSPCreate(PETSC_COMM_WORLD, &ksp);CHKERRQ(ierr);
  ierr = KSPSetOperators(ksp, A, A, SAME_NONZERO_PATTERN);CHKERRQ(ierr);
  ierr = KSPSetType(ksp, KSPCG);CHKERRQ(ierr);
  ierr = KSPGetPC(ksp, &pc);CHKERRQ(ierr);
  ierr = PCSetType(pc, PCJACOBI);CHKERRQ(ierr);
  ierr = KSPSetFromOptions(ksp);CHKERRQ(ierr);
  ierr = KSPSetResidualHistory(ksp, hist, 4096, PETSC_TRUE);CHKERRQ(ierr);
  if (cg_its > 0) {
    ierr = KSPSetTolerances(ksp, 1e-50, 1e-50, 1e50, cg_its);CHKERRQ(ierr);
    ierr = KSPSetUp(ksp);CHKERRQ(ierr);
    ierr = MPI_Barrier(PETSC_COMM_WORLD);CHKERRQ(ierr);
    t0 = MPI_Wtime();
    ierr = KSPSolve(ksp, b, u);CHKERRQ(ierr);
    ierr = MPI_Barrier(PETSC_COMM_WORLD);CHKERRQ(ierr);
    tcg = MPI_Wtime() - t0;
    ierr = KSPGetIterationNumber(ksp, &its);CHKERRQ(ierr);
    ierr = KSPGetResidualHistory(ksp, PETSC_NULL, &nhist);CHKERRQ(ierr);
  }
  if (solve_rtol > 0.0) {
    PetscReal h2[4];
    ierr = KSPSetResidualHistory(ksp, h2, 4, PETSC_TRUE);CHKERRQ(ierr);
    ierr = KSPSetTolerances(ksp, solve_rtol, 1e-50, 1e5, 100000);CHKERRQ(ierr);
    ierr = MPI_Barrier(PETSC_COMM_WORLD);CHKERRQ(ierr);
    t0 = MPI_Wtime();
    ierr = KSPSolve(ksp, b, u);CHKERRQ(ierr);
    ierr = MPI_Barrier(PETSC_COMM_WORLD);CHKERRQ(ierr);
    tsolve = MPI_Wtime() - t0;
    ierr = KSPGetIterationNumber(ksp, &solve_its);CHKERRQ(ierr);
    ierr = KSPGetConvergedReason(ksp, &reason);CHKERRQ(ierr);
  }
  if (!rank) {
    double tmin, tmed, tsum = 0.0;
    const char *mtype;
    ierr = MatGetType(A, &mtype);CHKERRQ(ierr);
    for (k = 0; k < reps; k++) tsum += t[k];
    qsort(t, (size_t)reps, sizeof(double), cmp_double);
    tmin = reps ? t[0] : 0.0; tmed = reps ? t[reps / 2] : 0.0;
    printf("{\"ranks\": %d, \"mat_type\": \"%s\", \"grid\": [%d, %d, %d], \"rows\": %d, \"nnz\": %lld, \"assembly_s\": %.6f, \"reps\": %d, "
           "\"matmult_mean_s\": %.9e, \"matmult_min_s\": %.9e, \"matmult_median_s\": %.9e, \"y_norm1\": %.17g, \"y_norm2\": %.17g, "
           "\"cg_its\": %d, \"cg_s\": %.9e, \"solve_rtol\": %g, \"solve_its\": %d, \"solve_s\": %.9e, \"hist\": [",
           (int)size, mtype, (int)nx, (int)ny, (int)nz, (int)N, gnnz, tasm, (int)reps, reps ? tsum / reps : 0.0, tmin, tmed,
           (double)ysum, (double)ynrm, (int)its, tcg, (double)solve_rtol, (int)solve_its, tsolve);
    for (k = 0; k < nhist && k < 64; k++) printf("%s%.17g", k ? ", " : "", (double)hist[k]);
    printf("]}\n");
    fflush(stdout);
  }
  free(t);
  ierr = KSPDestroy(&ksp);CHKERRQ(ierr);
  ierr = VecDestroy(&x);CHKERRQ(ierr); ierr = VecDestroy(&y);CHKERRQ(ierr); ierr = VecDestroy(&b);CHKERRQ(ierr);
  ierr = VecDestroy(&u);CHKERRQ(ierr); ierr = VecDestroy(&one);CHKERRQ(ierr);
  ierr = MatDestroy(&A);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
