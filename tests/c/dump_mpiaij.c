/* dump_mpiaij.c — TEST INFRASTRUCTURE.  Runs the REFERENCE's own parallel assembly on P ranks (under b200mpiexec) and writes
 * down, per rank, every integer structure the MPIAIJ hot path is built from, plus y = A x:
 *   MatSetValues_MPIAIJ / MatStash (mpiaij.c:476-580, matstash.c)   -> A (diag) and B (off-diag) CSR
 *   MatSetUpMultiply_MPIAIJ (mmaij.c:9-161)                          -> garray, B->j renumbered, the block/general IS choice
 *   MatCreateColmap_MPIAIJ_Private (mpiaij.c:357-377)                -> colmap (PetscTable under PETSC_USE_CTABLE)
 *   VecScatterCreate_PtoS (vpscat.c:1730-1924)                       -> to/from {n, procs, starts, indices, bs}, local part
 *   MatMult_MPIAIJ (mpiaij.c:1102-1116)                              -> y
 * tests/test_mpi_oracle.py compares these with libpetscb200's host restatements (b200_mpiaij_split, b200_scatter_plan_*).
 *
 *   b200mpiexec -n P dump_mpiaij -prefix /tmp/case [-stash] [-mat_type aijb200 -vec_type b200 -dll_prepend ...]
 * Rank r reads  <prefix>.<r>.in :  int32 m, N, bs, nz, rstart | int32 i[m+1] | int32 j[nz] (global ids) | double a[nz] | double x[m]
 * and writes    <prefix>.<r>.out:  lines "name count v0 v1 ..." (integers in decimal, doubles as C99 hex floats).
 * -stash: rank r inserts the rows of rank (r+1) mod P instead of its own, so every entry travels through the stash. */
static char help[] = "Dumps the MPIAIJ structures the reference builds on P ranks.\n";

#include <petscksp.h>
#include <../src/mat/impls/aij/mpi/mpiaij.h>
#include <petsc-private/vecimpl.h>

extern PetscErrorCode MatCreateColmap_MPIAIJ_Private(Mat);

static void put_ints(FILE *f, const char *name, PetscInt n, const PetscInt *v)
{
  PetscInt i;
  fprintf(f, "%s %d", name, (int)n);
  for (i = 0; i < n; i++) fprintf(f, " %d", (int)v[i]);
  fprintf(f, "\n");
}
static void put_mpiints(FILE *f, const char *name, PetscInt n, const PetscMPIInt *v)
{
  PetscInt i;
  fprintf(f, "%s %d", name, (int)n);
  for (i = 0; i < n; i++) fprintf(f, " %d", (int)v[i]);
  fprintf(f, "\n");
}
static void put_reals(FILE *f, const char *name, PetscInt n, const PetscScalar *v)
{
  PetscInt i;
  fprintf(f, "%s %d", name, (int)n);
  for (i = 0; i < n; i++) fprintf(f, " %a", (double)v[i]);
  fprintf(f, "\n");
}

typedef struct { int m, N, bs, nz, rstart; int *i, *j; double *a, *x; } Slab;

static PetscErrorCode read_slab(const char *prefix, int r, Slab *s)
{
  char  name[PETSC_MAX_PATH_LEN];
  FILE *f;
  int   hdr[5];
  PetscFunctionBegin;
  snprintf(name, sizeof(name), "%s.%d.in", prefix, r);
  f = fopen(name, "rb");
  if (!f) SETERRQ1(PETSC_COMM_SELF, PETSC_ERR_FILE_OPEN, "cannot open %s", name);
  if (fread(hdr, sizeof(int), 5, f) != 5) SETERRQ1(PETSC_COMM_SELF, PETSC_ERR_FILE_READ, "short header in %s", name);
  s->m = hdr[0]; s->N = hdr[1]; s->bs = hdr[2]; s->nz = hdr[3]; s->rstart = hdr[4];
  s->i = (int *)malloc(sizeof(int) * (s->m + 1)); s->j = (int *)malloc(sizeof(int) * (s->nz + 1));
  s->a = (double *)malloc(sizeof(double) * (s->nz + 1)); s->x = (double *)malloc(sizeof(double) * (s->m + 1));
  if (fread(s->i, sizeof(int), s->m + 1, f) != (size_t)(s->m + 1) || fread(s->j, sizeof(int), s->nz, f) != (size_t)s->nz ||
      fread(s->a, sizeof(double), s->nz, f) != (size_t)s->nz || fread(s->x, sizeof(double), s->m, f) != (size_t)s->m)
    SETERRQ1(PETSC_COMM_SELF, PETSC_ERR_FILE_READ, "short read in %s", name);
  fclose(f);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc, char **args)
{
  Mat            A;
  Vec            x, y;
  Slab           own, ins;
  PetscMPIInt    rank, size;
  PetscBool      flg, stash = PETSC_FALSE;
  char           prefix[PETSC_MAX_PATH_LEN], name[PETSC_MAX_PATH_LEN];
  PetscInt       r, *dnz, *onz, k, reps = 1;
  PetscScalar   *xa, *ya;
  FILE          *f;
  PetscErrorCode ierr;

  PetscInitialize(&argc, &args, (char *)0, help);
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD, &rank);CHKERRQ(ierr);
  ierr = MPI_Comm_size(PETSC_COMM_WORLD, &size);CHKERRQ(ierr);
  ierr = PetscOptionsGetString(PETSC_NULL, "-prefix", prefix, sizeof(prefix), &flg);CHKERRQ(ierr);
  if (!flg) SETERRQ(PETSC_COMM_WORLD, PETSC_ERR_USER, "need -prefix");
  ierr = PetscOptionsGetBool(PETSC_NULL, "-stash", &stash, PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL, "-reps", &reps, PETSC_NULL);CHKERRQ(ierr);
  ierr = read_slab(prefix, rank, &own);CHKERRQ(ierr);
  if (stash) { ierr = read_slab(prefix, (rank + 1) % size, &ins);CHKERRQ(ierr); } else ins = own;

  ierr = MatCreate(PETSC_COMM_WORLD, &A);CHKERRQ(ierr);
  ierr = MatSetSizes(A, own.m, own.m, own.N, own.N);CHKERRQ(ierr);
  ierr = MatSetType(A, MATAIJ);CHKERRQ(ierr);
  ierr = MatSetFromOptions(A);CHKERRQ(ierr);
  if (own.bs > 1) { ierr = MatSetBlockSize(A, own.bs);CHKERRQ(ierr); }
  /* preallocation from the rows this rank OWNS */
  ierr = PetscMalloc2(own.m, PetscInt, &dnz, own.m, PetscInt, &onz);CHKERRQ(ierr);
  for (r = 0; r < own.m; r++) {
    dnz[r] = onz[r] = 0;
    for (k = own.i[r]; k < own.i[r + 1]; k++) { if (own.j[k] >= own.rstart && own.j[k] < own.rstart + own.m) dnz[r]++; else onz[r]++; }
  }
  ierr = MatMPIAIJSetPreallocation(A, 0, dnz, 0, onz);CHKERRQ(ierr);
  ierr = MatSeqAIJSetPreallocation(A, 0, dnz);CHKERRQ(ierr);
  ierr = PetscFree2(dnz, onz);CHKERRQ(ierr);
  for (r = 0; r < ins.m; r++) {
    PetscInt row = ins.rstart + r;
    ierr = MatSetValues(A, 1, &row, ins.i[r + 1] - ins.i[r], ins.j + ins.i[r], ins.a + ins.i[r], INSERT_VALUES);CHKERRQ(ierr);
  }
  ierr = MatAssemblyBegin(A, MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatAssemblyEnd(A, MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);

  ierr = MatGetVecs(A, &x, &y);CHKERRQ(ierr);
  ierr = VecGetArray(x, &xa);CHKERRQ(ierr);
  for (r = 0; r < own.m; r++) xa[r] = own.x[r];
  ierr = VecRestoreArray(x, &xa);CHKERRQ(ierr);
  for (k = 0; k < reps; k++) { ierr = MatMult(A, x, y);CHKERRQ(ierr); }

  snprintf(name, sizeof(name), "%s.%d.out", prefix, (int)rank);
  f = fopen(name, "w");
  if (!f) SETERRQ1(PETSC_COMM_SELF, PETSC_ERR_FILE_OPEN, "cannot write %s", name);
  put_ints(f, "range", size + 1, A->rmap->range);
  if (size > 1) {
    Mat_MPIAIJ             *a  = (Mat_MPIAIJ *)A->data;
    Mat_SeqAIJ             *ad = (Mat_SeqAIJ *)a->A->data, *ao = (Mat_SeqAIJ *)a->B->data;
    VecScatter_MPI_General *to = (VecScatter_MPI_General *)a->Mvctx->todata, *from = (VecScatter_MPI_General *)a->Mvctx->fromdata;
    PetscInt                nghost = a->B->cmap->n, *cm, c;
    put_ints(f, "Ai", own.m + 1, ad->i); put_ints(f, "Aj", ad->nz, ad->j); put_reals(f, "Aa", ad->nz, ad->a);
    put_ints(f, "Bi", own.m + 1, ao->i); put_ints(f, "Bj", ao->nz, ao->j); put_reals(f, "Ba", ao->nz, ao->a);
    put_ints(f, "garray", nghost, a->garray);
    /* colmap: dense image of whatever representation the build uses (global column -> local ghost id + 1, else 0) */
    ierr = MatCreateColmap_MPIAIJ_Private(A);CHKERRQ(ierr);
    ierr = PetscMalloc((own.N + 1) * sizeof(PetscInt), &cm);CHKERRQ(ierr);
    for (c = 0; c < own.N; c++) {
#if defined(PETSC_USE_CTABLE)
      ierr = PetscTableFind(a->colmap, c + 1, &cm[c]);CHKERRQ(ierr);
#else
      cm[c] = a->colmap[c];
#endif
    }
    put_ints(f, "colmap", own.N, cm);
    ierr = PetscFree(cm);CHKERRQ(ierr);
    if (to->type != VEC_SCATTER_MPI_GENERAL || from->type != VEC_SCATTER_MPI_GENERAL) SETERRQ(PETSC_COMM_SELF, PETSC_ERR_PLIB, "unexpected scatter type");
    fprintf(f, "scatter_bs 1 %d\n", (int)to->bs);
    put_mpiints(f, "to_procs", to->n, to->procs);     put_ints(f, "to_starts", to->n + 1, to->starts);
    put_ints(f, "to_indices", to->starts[to->n], to->indices);
    put_mpiints(f, "from_procs", from->n, from->procs); put_ints(f, "from_starts", from->n + 1, from->starts);
    put_ints(f, "from_indices", from->starts[from->n], from->indices);
    put_ints(f, "to_local", to->local.n, to->local.vslots); put_ints(f, "from_local", from->local.n, from->local.vslots);
    fprintf(f, "from_contiq 1 %d\n", (int)from->contiq);
  }
  ierr = VecGetArray(y, &ya);CHKERRQ(ierr);
  put_reals(f, "y", own.m, ya);
  ierr = VecRestoreArray(y, &ya);CHKERRQ(ierr);
  fclose(f);
  ierr = VecDestroy(&x);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);
  ierr = MatDestroy(&A);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
