/* b200mat.c — MATSEQAIJB200 / MATMPIAIJB200 / "aijb200": AIJ matrices whose MatMult / MatMultAdd run on a B200.
 *
 * Mirrors MatCreate_SeqAIJCUSP (reference src/mat/impls/aij/seq/seqcusp/aijcusp.cu:657-681) and MatCreate_MPIAIJCUSP
 * (src/mat/impls/aij/mpi/mpicusp/mpiaijcusp.cu:192-226): inherit the complete host CSR type (MatCreate_SeqAIJ aij.c:3886,
 * MatCreate_MPIAIJ mpiaij.c:5617 — preallocation, MatSetValues, assembly, views, MatGetDiagonal ... all stay host code),
 * override mult / multadd / multtranspose / multtransposeadd / getvecs / destroy (+ setvaluesbatch for the sequential
 * type, aijcusp.cu:639), switch the inode path off (aijcusp.cu:665).
 *
 * Device copy: built lazily at the first MatMult after the matrix changed.  "Changed" is detected with the object state
 * every mutating Mat front end bumps (MatAssemblyEnd, MatScale, MatShift, MatZeroEntries, MatDiagonalScale, MatZeroRows:
 * PetscObjectStateIncrease in src/mat/interface/matrix.c) — the PETSC_HAVE_CUSP build marks the same places with
 * valid_GPU_matrix = PETSC_CUSP_CPU (matrix.c:1085-1089,4917-4921).  When the row offsets (and, for the parallel type, the
 * ghost columns) are unchanged only the values are re-sent (b200_mat_update_values[_mpiaij]) — no re-binning, no new
 * exchange plan — not the whole matrix as MatCUSPCopyToGPU does (aijcusp.cu:138-152); a new preallocation drops the
 * device copy, because it can keep the row lengths and change the columns. */
#include "petscb200plugin.h"
#include <../src/mat/impls/aij/seq/aij.h>
#include <../src/mat/impls/aij/mpi/mpiaij.h>
#include <petsc-private/vecimpl.h>
#include "petscb200.h"

extern PetscInt PetscB200MatUploads, PetscB200MatValueUpdates;

#define CHKB200(call) do { int b200rc_ = (call); if (b200rc_) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_LIB,"libpetscb200 error %d: %s",b200rc_,b200_last_error()); } while (0)

typedef unsigned long long B200Hash;

typedef struct {
  b200_mat  mat;          /* device CSR + row binning                           */
  PetscInt  state;        /* object state of the host matrix at the last upload */
  PetscInt  m,nz;
  PetscInt *i;            /* host copy of the uploaded row offsets              */
  B200Hash  jhash;        /* hash of the uploaded column indices                */
} Mat_SeqAIJB200;

typedef struct {
  b200_mat  mat;          /* diagonal block, off-diagonal block, ghost plan     */
  PetscInt  state,stateA,stateB;
  PetscInt  m,nzA,nzB,nghost;
  PetscInt *iA,*iB,*garray; /* host copies of what was uploaded: row offsets of both blocks, ghost columns */
  B200Hash  jhashA,jhashB;  /* hashes of the uploaded column indices of both blocks                          */
} Mat_MPIAIJB200;

/* Hash of a column-index array.  Equal row offsets do not imply equal columns: MatZeroRows with a diagonal value on a row
 * whose single entry was off the diagonal rewrites a->j in place (aij.c:1680-1683) and leaves i and nz alone.  Four
 * independent multiply-xor lanes keep it at memory speed (the check reads 4 bytes per nonzero; the value refresh it
 * guards moves 8 over PCIe). */
static B200Hash MatB200HashInts(const PetscInt *v,PetscInt n)
{
  const B200Hash K = 0x9E3779B97F4A7C15ull;
  B200Hash       h0 = 1,h1 = 2,h2 = 3,h3 = 4;
  PetscInt       k = 0;
  for (; k+4<=n; k+=4) {
    h0 = (h0 ^ (B200Hash)(unsigned)v[k])*K;   h1 = (h1 ^ (B200Hash)(unsigned)v[k+1])*K;
    h2 = (h2 ^ (B200Hash)(unsigned)v[k+2])*K; h3 = (h3 ^ (B200Hash)(unsigned)v[k+3])*K;
  }
  for (; k<n; k++) h0 = (h0 ^ (B200Hash)(unsigned)v[k])*K;
  h0 ^= h0>>29; h1 ^= h1>>31; h2 ^= h2>>33; h3 ^= h3>>27;
  return ((h0*K ^ h1)*K ^ h2)*K ^ h3;
}

/* ------------------------------------------------------------------------------------------------ sequential */
#undef __FUNCT__
#define __FUNCT__ "MatB200CopyToDevice_SeqAIJ"
static PetscErrorCode MatB200CopyToDevice_SeqAIJ(Mat A)                           /* MatCUSPCopyToGPU aijcusp.cu:126 */
{
  Mat_SeqAIJ     *a = (Mat_SeqAIJ*)A->data;
  Mat_SeqAIJB200 *d = (Mat_SeqAIJB200*)A->spptr;
  PetscInt       m = A->rmap->n;
  PetscBool      same = PETSC_FALSE;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (!A->assembled) SETERRQ(((PetscObject)A)->comm,PETSC_ERR_ARG_WRONGSTATE,"Not for unassembled matrix");
  if (d->mat && d->state == ((PetscObject)A)->state) PetscFunctionReturn(0);
  if (d->mat && d->m == m && d->nz == a->nz) {
    ierr = PetscMemcmp(d->i,a->i,(m+1)*sizeof(PetscInt),&same);CHKERRQ(ierr);
    if (same && MatB200HashInts(a->j,a->nz) != d->jhash) same = PETSC_FALSE;
  }
  if (same) {
    CHKB200(b200_mat_update_values(d->mat,a->a));
    PetscB200MatValueUpdates++;
    ierr = PetscInfo1(A,"B200 device copy: values refreshed (%D nonzeros), pattern unchanged\n",a->nz);CHKERRQ(ierr);
  } else {
    if (d->mat) { CHKB200(b200_mat_destroy(d->mat)); d->mat = 0; }
    ierr = PetscFree(d->i);CHKERRQ(ierr);
    CHKB200(b200_mat_create_seqaij(m,A->cmap->n,a->i,a->j,a->a,&d->mat));
    ierr = PetscMalloc((m+1)*sizeof(PetscInt),&d->i);CHKERRQ(ierr);
    ierr = PetscMemcpy(d->i,a->i,(m+1)*sizeof(PetscInt));CHKERRQ(ierr);
    d->m = m; d->nz = a->nz; d->jhash = MatB200HashInts(a->j,a->nz);
    PetscB200MatUploads++;
    ierr = PetscInfo1(A,"B200 device copy: full upload (%D nonzeros)\n",a->nz);CHKERRQ(ierr);
  }
  d->state = ((PetscObject)A)->state;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMult_SeqAIJB200"
static PetscErrorCode MatMult_SeqAIJB200(Mat A,Vec xx,Vec yy)                     /* MatMult_SeqAIJ aij.c:1225 */
{
  const PetscScalar *x;
  PetscScalar       *y;
  double            flops;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_SeqAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  ierr = VecB200GetDeviceWrite(yy,&y);CHKERRQ(ierr);
  CHKB200(b200_mat_mult(((Mat_SeqAIJB200*)A->spptr)->mat,x,y));
  CHKB200(b200_mat_flops(((Mat_SeqAIJB200*)A->spptr)->mat,&flops));               /* 2 nz - nonzerorow, aij.c:1281 */
  ierr = PetscLogFlops(flops);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultAdd_SeqAIJB200"
static PetscErrorCode MatMultAdd_SeqAIJB200(Mat A,Vec xx,Vec yy,Vec zz)           /* MatMultAdd_SeqAIJ aij.c:1291 */
{
  Mat_SeqAIJ        *a = (Mat_SeqAIJ*)A->data;
  const PetscScalar *x,*y;
  PetscScalar       *z;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_SeqAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  if (zz == yy) { ierr = VecB200GetDeviceReadWrite(zz,&z);CHKERRQ(ierr); y = z; }
  else {
    ierr = VecB200GetDeviceRead(yy,&y);CHKERRQ(ierr);
    ierr = VecB200GetDeviceWrite(zz,&z);CHKERRQ(ierr);
  }
  CHKB200(b200_mat_multadd(((Mat_SeqAIJB200*)A->spptr)->mat,x,y,z));
  ierr = PetscLogFlops(2.0*a->nz);CHKERRQ(ierr);                                  /* aij.c:1343 */
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultTranspose_SeqAIJB200"
static PetscErrorCode MatMultTranspose_SeqAIJB200(Mat A,Vec xx,Vec yy)            /* MatMultTranspose_SeqAIJ aij.c:1127 */
{
  Mat_SeqAIJ        *a = (Mat_SeqAIJ*)A->data;
  const PetscScalar *x;
  PetscScalar       *y;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_SeqAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  ierr = VecB200GetDeviceWrite(yy,&y);CHKERRQ(ierr);
  CHKB200(b200_mat_multtranspose(((Mat_SeqAIJB200*)A->spptr)->mat,x,y));
  ierr = PetscLogFlops(2.0*a->nz);CHKERRQ(ierr);                                  /* aij.c:1120 */
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultTransposeAdd_SeqAIJB200"
static PetscErrorCode MatMultTransposeAdd_SeqAIJB200(Mat A,Vec xx,Vec yy,Vec zz)  /* MatMultTransposeAdd_SeqAIJ aij.c:1078 */
{
  Mat_SeqAIJ        *a = (Mat_SeqAIJ*)A->data;
  const PetscScalar *x,*y;
  PetscScalar       *z;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_SeqAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  if (zz == yy) { ierr = VecB200GetDeviceReadWrite(zz,&z);CHKERRQ(ierr); y = z; }
  else {
    ierr = VecB200GetDeviceRead(yy,&y);CHKERRQ(ierr);
    ierr = VecB200GetDeviceWrite(zz,&z);CHKERRQ(ierr);
  }
  CHKB200(b200_mat_multtransposeadd(((Mat_SeqAIJB200*)A->spptr)->mat,x,y,z));
  ierr = PetscLogFlops(2.0*a->nz);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatGetVecs_SeqAIJB200"
static PetscErrorCode MatGetVecs_SeqAIJB200(Mat mat,Vec *right,Vec *left)         /* MatGetVecs_SeqAIJCUSP aijcusp.cu:324 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (right) {
    ierr = VecCreate(((PetscObject)mat)->comm,right);CHKERRQ(ierr);
    ierr = VecSetSizes(*right,mat->cmap->n,PETSC_DETERMINE);CHKERRQ(ierr);
    ierr = VecSetBlockSize(*right,mat->rmap->bs);CHKERRQ(ierr);
    ierr = VecSetType(*right,VECSEQB200);CHKERRQ(ierr);
    ierr = PetscLayoutReference(mat->cmap,&(*right)->map);CHKERRQ(ierr);
  }
  if (left) {
    ierr = VecCreate(((PetscObject)mat)->comm,left);CHKERRQ(ierr);
    ierr = VecSetSizes(*left,mat->rmap->n,PETSC_DETERMINE);CHKERRQ(ierr);
    ierr = VecSetBlockSize(*left,mat->rmap->bs);CHKERRQ(ierr);
    ierr = VecSetType(*left,VECSEQB200);CHKERRQ(ierr);
    ierr = PetscLayoutReference(mat->rmap,&(*left)->map);CHKERRQ(ierr);
  }
  PetscFunctionReturn(0);
}

/* The values-only refresh above trusts "same row offsets => same pattern".  That holds for everything that edits an
 * assembled matrix (a new nonzero lengthens its row), but NOT for a second preallocation, which throws the arrays away
 * and may be followed by different columns with the same row lengths.  So preallocation drops the device copy. */
EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "MatSeqAIJSetPreallocation_SeqAIJB200"
PetscErrorCode MatSeqAIJSetPreallocation_SeqAIJB200(Mat B,PetscInt nz,const PetscInt *nnz)
{
  Mat_SeqAIJB200 *d = (Mat_SeqAIJB200*)B->spptr;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (d && d->mat) { CHKB200(b200_mat_destroy(d->mat)); d->mat = 0; }
  ierr = MatSeqAIJSetPreallocation_SeqAIJ(B,nz,nnz);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
EXTERN_C_END

/* MatSetValuesBatch_SeqAIJCUSP (src/mat/impls/aij/seq/seqcusp/aijAssemble.cu:157-290): the element matrices are
 * assembled on the device (b200_mat_setvaluesbatch: per-row sort + ordered sums, bit-identical to the host loop of
 * matrix.c:1712-1716 on an empty matrix); like the reference's version it "creates the given matrix" — whatever
 * preallocated pattern J had is replaced — and the host copy is filled from the device result (the reference's
 * MatCUSPCopyFromGPU step, aijAssemble.cu:287).  The device matrix stays: the next MatMult only refreshes values. */
#undef __FUNCT__
#define __FUNCT__ "MatSetValuesBatch_SeqAIJB200"
static PetscErrorCode MatSetValuesBatch_SeqAIJB200(Mat J,PetscInt Ne,PetscInt Nl,PetscInt *elemRows,const PetscScalar *elemMats)
{
  Mat_SeqAIJB200 *d = (Mat_SeqAIJB200*)J->spptr;
  PetscInt       m = J->rmap->n,n = J->cmap->n,*ci,*cj;
  PetscScalar    *ca;
  long long      nz;
  b200_mat       dm;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscInfo3(J,"Assembling %D element matrices of %D x %D on the device\n",Ne,Nl,Nl);CHKERRQ(ierr);
  CHKB200(b200_mat_setvaluesbatch(m,n,(long long)Ne,Nl,elemRows,elemMats,0,0,&dm));
  CHKB200(b200_mat_get_size(dm,0,0,&nz));
  ierr = PetscMalloc3(m+1,PetscInt,&ci,(size_t)nz+1,PetscInt,&cj,(size_t)nz+1,PetscScalar,&ca);CHKERRQ(ierr);
  CHKB200(b200_mat_get_csr(dm,ci,cj,ca));
  ierr = MatSeqAIJSetPreallocationCSR(J,ci,cj,ca);CHKERRQ(ierr);                  /* sorted rows: straight copies, then assembles */
  if (d->mat) { CHKB200(b200_mat_destroy(d->mat)); }
  ierr = PetscFree(d->i);CHKERRQ(ierr);
  d->mat = dm; d->m = m; d->nz = (PetscInt)nz;
  ierr = PetscMalloc((m+1)*sizeof(PetscInt),&d->i);CHKERRQ(ierr);
  ierr = PetscMemcpy(d->i,ci,(m+1)*sizeof(PetscInt));CHKERRQ(ierr);
  d->state = ((PetscObject)J)->state;
  PetscB200MatUploads++;
  ierr = PetscFree3(ci,cj,ca);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatDestroy_SeqAIJB200"
static PetscErrorCode MatDestroy_SeqAIJB200(Mat A)                                /* MatDestroy_SeqAIJCUSP aijcusp.cu:566 */
{
  Mat_SeqAIJB200 *d = (Mat_SeqAIJB200*)A->spptr;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (d) {
    if (d->mat) { CHKB200(b200_mat_destroy(d->mat)); }
    ierr = PetscFree(d->i);CHKERRQ(ierr);
    ierr = PetscFree(d);CHKERRQ(ierr);
  }
  A->spptr = 0;                                                                   /* MatDestroy frees a non-null spptr */
  ierr = MatDestroy_SeqAIJ(A);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "MatCreate_SeqAIJB200"
PetscErrorCode MatCreate_SeqAIJB200(Mat B)
{
  Mat_SeqAIJB200 *d;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MatCreate_SeqAIJ(B);CHKERRQ(ierr);
  ((Mat_SeqAIJ*)B->data)->inode.use = PETSC_FALSE;                                /* aijcusp.cu:665 */
  ierr = PetscNewLog(B,Mat_SeqAIJB200,&d);CHKERRQ(ierr);
  B->spptr        = (void*)d;
  B->ops->mult    = MatMult_SeqAIJB200;
  B->ops->multadd = MatMultAdd_SeqAIJB200;
  B->ops->multtranspose    = MatMultTranspose_SeqAIJB200;
  B->ops->multtransposeadd = MatMultTransposeAdd_SeqAIJB200;
  B->ops->getvecs = MatGetVecs_SeqAIJB200;
  B->ops->destroy = MatDestroy_SeqAIJB200;
  B->ops->setvaluesbatch = MatSetValuesBatch_SeqAIJB200;                          /* aijcusp.cu:639 */
  ierr = PetscObjectComposeFunctionDynamic((PetscObject)B,"MatSeqAIJSetPreallocation_C","MatSeqAIJSetPreallocation_SeqAIJB200",
                                           MatSeqAIJSetPreallocation_SeqAIJB200);CHKERRQ(ierr);
  ierr = PetscObjectChangeTypeName((PetscObject)B,MATSEQAIJB200);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
EXTERN_C_END

/* ------------------------------------------------------------------------------------------------ parallel */
#undef __FUNCT__
#define __FUNCT__ "MatMPIAIJSetPreallocation_MPIAIJB200"
static PetscErrorCode MatMPIAIJSetPreallocation_MPIAIJB200(Mat B,PetscInt d_nz,const PetscInt d_nnz[],PetscInt o_nz,const PetscInt o_nnz[])
{                                                                                 /* mpiaijcusp.cu:9-55 */
  Mat_MPIAIJ     *b = (Mat_MPIAIJ*)B->data;
  Mat_MPIAIJB200 *d = (Mat_MPIAIJB200*)b->spptr;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (d && d->mat) { CHKB200(b200_mat_destroy(d->mat)); d->mat = 0; }              /* a new preallocation may keep the row lengths and change the columns */
  if (!B->preallocated) {
    /* the two sequential blocks are of our type, so every host op on them stays coherent with the device copy */
    ierr = PetscLayoutSetUp(B->rmap);CHKERRQ(ierr);
    ierr = PetscLayoutSetUp(B->cmap);CHKERRQ(ierr);
    ierr = MatCreate(PETSC_COMM_SELF,&b->A);CHKERRQ(ierr);
    ierr = MatSetSizes(b->A,B->rmap->n,B->cmap->n,B->rmap->n,B->cmap->n);CHKERRQ(ierr);
    ierr = MatSetBlockSizes(b->A,B->rmap->bs,B->cmap->bs);CHKERRQ(ierr);
    ierr = MatSetType(b->A,MATSEQAIJB200);CHKERRQ(ierr);
    ierr = PetscLogObjectParent(B,b->A);CHKERRQ(ierr);
    ierr = MatCreate(PETSC_COMM_SELF,&b->B);CHKERRQ(ierr);
    ierr = MatSetSizes(b->B,B->rmap->n,B->cmap->N,B->rmap->n,B->cmap->N);CHKERRQ(ierr);
    ierr = MatSetBlockSizes(b->B,B->rmap->bs,B->cmap->bs);CHKERRQ(ierr);
    ierr = MatSetType(b->B,MATSEQAIJB200);CHKERRQ(ierr);
    ierr = PetscLogObjectParent(B,b->B);CHKERRQ(ierr);
    B->preallocated = PETSC_TRUE;                                                 /* the base routine then only preallocates */
  }
  ierr = MatMPIAIJSetPreallocation_MPIAIJ(B,d_nz,d_nnz,o_nz,o_nnz);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* (re)build the device operator from the assembled host pieces: Mat_MPIAIJ::A, ::B (columns already compacted to
 * ghost ids by MatSetUpMultiply_MPIAIJ mmaij.c:56-63), ::garray, and the VecScatter_MPI_General send/receive lists of
 * ::Mvctx (vpscat.c:1827-1879) — what VecScatterInitializeForGPU digs out for the CUSP path (vscatcusp.c:29-133). */
#undef __FUNCT__
#define __FUNCT__ "MatB200CopyToDevice_MPIAIJ"
static PetscErrorCode MatB200CopyToDevice_MPIAIJ(Mat A)
{
  Mat_MPIAIJ     *a  = (Mat_MPIAIJ*)A->data;
  Mat_MPIAIJB200 *d  = (Mat_MPIAIJB200*)a->spptr;
  Mat_SeqAIJ     *ad = (Mat_SeqAIJ*)a->A->data,*ao = (Mat_SeqAIJ*)a->B->data;
  PetscInt       nghost = a->B->cmap->n,nto = 0,nfrom = 0,i,k,c;
  int            *to_procs = 0,*to_starts = 0,*to_indices = 0,*from_procs = 0,*from_starts = 0;
  PetscBool      same = PETSC_FALSE;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (!A->assembled) SETERRQ(((PetscObject)A)->comm,PETSC_ERR_ARG_WRONGSTATE,"Not for unassembled matrix");
  if (d->mat && d->state == ((PetscObject)A)->state && d->stateA == ((PetscObject)a->A)->state && d->stateB == ((PetscObject)a->B)->state) PetscFunctionReturn(0);
  if (d->mat && d->m == A->rmap->n && d->nzA == ad->nz && d->nzB == ao->nz && d->nghost == nghost) {
    /* same row offsets in both blocks and the same ghost columns: the pattern, the exchange plan and the peer buffers
     * stand, only the values travel (MatAssemblyEnd_MPIAIJ rebuilds B and garray whenever a new off-process column
     * appears, mpiaij.c:696-710, which changes these arrays).  The decision must be the same on every rank: a full
     * rebuild is collective. */
    PetscBool sA,sB,sG = PETSC_TRUE;
    ierr = PetscMemcmp(d->iA,ad->i,(d->m+1)*sizeof(PetscInt),&sA);CHKERRQ(ierr);
    ierr = PetscMemcmp(d->iB,ao->i,(d->m+1)*sizeof(PetscInt),&sB);CHKERRQ(ierr);
    if (nghost) { ierr = PetscMemcmp(d->garray,a->garray,nghost*sizeof(PetscInt),&sG);CHKERRQ(ierr); }
    same = (PetscBool)(sA && sB && sG);
    if (same && (MatB200HashInts(ad->j,ad->nz) != d->jhashA || MatB200HashInts(ao->j,ao->nz) != d->jhashB)) same = PETSC_FALSE;
    if (same) {                                                                   /* column-stripped blocks are not refreshed in place */
      int sa = 1,sb = 1;
      CHKB200(b200_mat_strips(d->mat,&sa,&sb));
      if (sa > 1 || sb > 1) same = PETSC_FALSE;
    }
  }
  {
    PetscInt mine = same ? 1 : 0,all = mine;
    if (a->size > 1) { ierr = MPI_Allreduce(&mine,&all,1,MPIU_INT,MPI_MIN,((PetscObject)A)->comm);CHKERRQ(ierr); }
    same = all ? PETSC_TRUE : PETSC_FALSE;
  }
  if (same) {
    CHKB200(b200_mat_update_values_mpiaij(d->mat,ad->a,ao->a));
    PetscB200MatValueUpdates++;
    ierr = PetscInfo2(A,"B200 device copy: values refreshed (%D + %D nonzeros), pattern and ghost exchange unchanged\n",ad->nz,ao->nz);CHKERRQ(ierr);
    d->state = ((PetscObject)A)->state; d->stateA = ((PetscObject)a->A)->state; d->stateB = ((PetscObject)a->B)->state;
    PetscFunctionReturn(0);
  }
  if (d->mat) { CHKB200(b200_mat_destroy(d->mat)); d->mat = 0; }
  if (a->size > 1) {
    VecScatter_MPI_General *to = (VecScatter_MPI_General*)a->Mvctx->todata,*from = (VecScatter_MPI_General*)a->Mvctx->fromdata;
    PetscInt               bs = to->bs;
    if (to->type != VEC_SCATTER_MPI_GENERAL || from->type != VEC_SCATTER_MPI_GENERAL) SETERRQ(((PetscObject)A)->comm,PETSC_ERR_SUP,"MatMult scatter is not the general parallel scatter");
    nto = to->n; nfrom = from->n;
    ierr = PetscMalloc5(nto+1,int,&to_procs,nto+1,int,&to_starts,bs*to->starts[nto]+1,int,&to_indices,nfrom+1,int,&from_procs,nfrom+1,int,&from_starts);CHKERRQ(ierr);
    for (i=0; i<nto; i++)    to_procs[i] = to->procs[i];
    for (i=0; i<=nto; i++)   to_starts[i] = bs*to->starts[i];
    for (k=0; k<to->starts[nto]; k++) for (c=0; c<bs; c++) to_indices[bs*k+c] = to->indices[k]+c;   /* Pack_bs, vpscat.c:674 */
    for (i=0; i<nfrom; i++)  from_procs[i] = from->procs[i];
    for (i=0; i<=nfrom; i++) from_starts[i] = bs*from->starts[i];
    /* receives land in lvec in order: MatSetUpMultiply's `to` index set is the stride 0..nghost-1 (mmaij.c:144) */
    for (k=0; k<from->starts[nfrom]; k++) if (from->indices[k] != bs*k) SETERRQ(((PetscObject)A)->comm,PETSC_ERR_PLIB,"Ghost receive positions are not contiguous");
  }
  CHKB200(b200_mat_create_mpiaij(A->rmap->n,A->cmap->n,A->cmap->N,ad->i,ad->j,ad->a,nghost,ao->i,ao->j,ao->a,a->garray,
                                 nto,to_procs,to_starts,to_indices,nfrom,from_procs,from_starts,&d->mat));
  if (a->size > 1) { ierr = PetscFree5(to_procs,to_starts,to_indices,from_procs,from_starts);CHKERRQ(ierr); }
  PetscB200MatUploads++;
  ierr = PetscInfo2(A,"B200 device copy: full upload (%D + %D nonzeros) and ghost exchange plan\n",ad->nz,ao->nz);CHKERRQ(ierr);
  ierr = PetscFree3(d->iA,d->iB,d->garray);CHKERRQ(ierr);
  d->m = A->rmap->n; d->nzA = ad->nz; d->nzB = ao->nz; d->nghost = nghost;
  d->jhashA = MatB200HashInts(ad->j,ad->nz); d->jhashB = MatB200HashInts(ao->j,ao->nz);
  ierr = PetscMalloc3(d->m+1,PetscInt,&d->iA,d->m+1,PetscInt,&d->iB,nghost+1,PetscInt,&d->garray);CHKERRQ(ierr);
  ierr = PetscMemcpy(d->iA,ad->i,(d->m+1)*sizeof(PetscInt));CHKERRQ(ierr);
  ierr = PetscMemcpy(d->iB,ao->i,(d->m+1)*sizeof(PetscInt));CHKERRQ(ierr);
  if (nghost) { ierr = PetscMemcpy(d->garray,a->garray,nghost*sizeof(PetscInt));CHKERRQ(ierr); }
  d->state = ((PetscObject)A)->state; d->stateA = ((PetscObject)a->A)->state; d->stateB = ((PetscObject)a->B)->state;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMult_MPIAIJB200"
static PetscErrorCode MatMult_MPIAIJB200(Mat A,Vec xx,Vec yy)                     /* MatMult_MPIAIJ mpiaij.c:1102 */
{
  Mat_MPIAIJ        *a = (Mat_MPIAIJ*)A->data;
  const PetscScalar *x;
  PetscScalar       *y;
  double            flops;
  PetscInt          nt;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = VecGetLocalSize(xx,&nt);CHKERRQ(ierr);
  if (nt != A->cmap->n) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_ARG_SIZ,"Incompatible partition of A (%D) and xx (%D)",A->cmap->n,nt);
  ierr = MatB200CopyToDevice_MPIAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  ierr = VecB200GetDeviceWrite(yy,&y);CHKERRQ(ierr);
  /* ghost exchange (GPU to GPU) overlapped with the diagonal block, then y += B lvec: all inside the library */
  CHKB200(b200_mat_mult(((Mat_MPIAIJB200*)a->spptr)->mat,x,y));
  CHKB200(b200_mat_flops(((Mat_MPIAIJB200*)a->spptr)->mat,&flops));
  ierr = PetscLogFlops(flops);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultAdd_MPIAIJB200"
static PetscErrorCode MatMultAdd_MPIAIJB200(Mat A,Vec xx,Vec yy,Vec zz)           /* MatMultAdd_MPIAIJ mpiaij.c:1132 */
{
  Mat_MPIAIJ        *a = (Mat_MPIAIJ*)A->data;
  const PetscScalar *x,*y;
  PetscScalar       *z;
  double            flops;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_MPIAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  if (zz == yy) { ierr = VecB200GetDeviceReadWrite(zz,&z);CHKERRQ(ierr); y = z; }
  else {
    ierr = VecB200GetDeviceRead(yy,&y);CHKERRQ(ierr);
    ierr = VecB200GetDeviceWrite(zz,&z);CHKERRQ(ierr);
  }
  CHKB200(b200_mat_multadd(((Mat_MPIAIJB200*)a->spptr)->mat,x,y,z));
  CHKB200(b200_mat_flops(((Mat_MPIAIJB200*)a->spptr)->mat,&flops));
  ierr = PetscLogFlops(flops + A->rmap->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultTranspose_MPIAIJB200"
static PetscErrorCode MatMultTranspose_MPIAIJB200(Mat A,Vec xx,Vec yy)            /* MatMultTranspose_MPIAIJ mpiaij.c:1147 */
{
  Mat_MPIAIJ        *a = (Mat_MPIAIJ*)A->data;
  Mat_SeqAIJ        *ad = (Mat_SeqAIJ*)a->A->data,*ao = (Mat_SeqAIJ*)a->B->data;
  const PetscScalar *x;
  PetscScalar       *y;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_MPIAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  ierr = VecB200GetDeviceWrite(yy,&y);CHKERRQ(ierr);
  /* B^T x, reverse ghost exchange (ADD) GPU to GPU overlapped with A^T x: all inside the library */
  CHKB200(b200_mat_multtranspose(((Mat_MPIAIJB200*)a->spptr)->mat,x,y));
  ierr = PetscLogFlops(2.0*ad->nz + 2.0*ao->nz);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatMultTransposeAdd_MPIAIJB200"
static PetscErrorCode MatMultTransposeAdd_MPIAIJB200(Mat A,Vec xx,Vec yy,Vec zz)  /* MatMultTransposeAdd_MPIAIJ mpiaij.c:1178 */
{
  Mat_MPIAIJ        *a = (Mat_MPIAIJ*)A->data;
  Mat_SeqAIJ        *ad = (Mat_SeqAIJ*)a->A->data,*ao = (Mat_SeqAIJ*)a->B->data;
  const PetscScalar *x,*y;
  PetscScalar       *z;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = MatB200CopyToDevice_MPIAIJ(A);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(xx,&x);CHKERRQ(ierr);
  if (zz == yy) { ierr = VecB200GetDeviceReadWrite(zz,&z);CHKERRQ(ierr); y = z; }
  else {
    ierr = VecB200GetDeviceRead(yy,&y);CHKERRQ(ierr);
    ierr = VecB200GetDeviceWrite(zz,&z);CHKERRQ(ierr);
  }
  CHKB200(b200_mat_multtransposeadd(((Mat_MPIAIJB200*)a->spptr)->mat,x,y,z));
  ierr = PetscLogFlops(2.0*ad->nz + 2.0*ao->nz);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatGetVecs_MPIAIJB200"
static PetscErrorCode MatGetVecs_MPIAIJB200(Mat mat,Vec *right,Vec *left)         /* MatGetVecs_MPIAIJCUSP mpiaijcusp.cu:59 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (right) {
    ierr = VecCreate(((PetscObject)mat)->comm,right);CHKERRQ(ierr);
    ierr = VecSetSizes(*right,mat->cmap->n,PETSC_DETERMINE);CHKERRQ(ierr);
    ierr = VecSetBlockSize(*right,mat->rmap->bs);CHKERRQ(ierr);
    ierr = VecSetType(*right,VECMPIB200);CHKERRQ(ierr);
    ierr = PetscLayoutReference(mat->cmap,&(*right)->map);CHKERRQ(ierr);
  }
  if (left) {
    ierr = VecCreate(((PetscObject)mat)->comm,left);CHKERRQ(ierr);
    ierr = VecSetSizes(*left,mat->rmap->n,PETSC_DETERMINE);CHKERRQ(ierr);
    ierr = VecSetBlockSize(*left,mat->rmap->bs);CHKERRQ(ierr);
    ierr = VecSetType(*left,VECMPIB200);CHKERRQ(ierr);
    ierr = PetscLayoutReference(mat->rmap,&(*left)->map);CHKERRQ(ierr);
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "MatDestroy_MPIAIJB200"
static PetscErrorCode MatDestroy_MPIAIJB200(Mat A)                                /* MatDestroy_MPIAIJCUSP mpiaijcusp.cu:176 */
{
  Mat_MPIAIJ     *a = (Mat_MPIAIJ*)A->data;
  Mat_MPIAIJB200 *d = (Mat_MPIAIJB200*)a->spptr;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (d) {
    if (d->mat) { CHKB200(b200_mat_destroy(d->mat)); }
    ierr = PetscFree3(d->iA,d->iB,d->garray);CHKERRQ(ierr);
    ierr = PetscFree(d);CHKERRQ(ierr);
    a->spptr = 0;
  }
  ierr = MatDestroy_MPIAIJ(A);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* MatSetValuesBatch_MPIAIJCUSP (src/mat/impls/aij/mpi/mpicusp/mpiaijAssemble.cu:200-460).  The reference ships the
 * (i,j,v) entries of rows this rank does not own to their owners (three MPI_Alltoallv, :326-352) and sorts diagonal and
 * off-diagonal COO lists; here every rank receives the whole ELEMENTS that touch its rows, the device assembles the
 * local rows from own + received elements (b200_mat_setvaluesbatch_rows passes over rows owned elsewhere and keeps
 * their ids as global columns), and the result enters the host matrix through MatMPIAIJSetPreallocationCSR, which
 * splits it into the two blocks and assembles.  The exchange is setup-time and kept simple: one MPI_Allgatherv of
 * the interface elements; who sends what and where a rank's share starts in a sender's buffer is decided by the
 * library (b200_batch_send_plan / b200_batch_recv_plan — the same two functions petsc-dev_b200/parallel.py uses on
 * real ranks and tests/test_assembly.py drives with virtual ranks), so only the two MPI calls below are specific to
 * this file. */
#undef __FUNCT__
#define __FUNCT__ "MatSetValuesBatch_MPIAIJB200"
static PetscErrorCode MatSetValuesBatch_MPIAIJB200(Mat J,PetscInt Ne,PetscInt Nl,PetscInt *elemRows,const PetscScalar *elemMats)
{
  MPI_Comm       comm = ((PetscObject)J)->comm;
  PetscMPIInt    size,rank;
  PetscInt       rstart,rend,m,N,ne = Ne,*rows = elemRows,*ci,*cj,p,s;
  PetscInt       *allrows = 0,*recvrows = 0,*sendrows = 0;
  PetscScalar    *allmats = 0,*recvmats = 0,*sendmats = 0,*ca;
  const PetscScalar *mats = elemMats;
  long long      nz;
  b200_mat       dm;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MPI_Comm_size(comm,&size);CHKERRQ(ierr);
  ierr = MPI_Comm_rank(comm,&rank);CHKERRQ(ierr);
  ierr = PetscLayoutSetUp(J->rmap);CHKERRQ(ierr);
  ierr = PetscLayoutSetUp(J->cmap);CHKERRQ(ierr);
  rstart = J->rmap->rstart; rend = J->rmap->rend; m = rend - rstart; N = J->cmap->N;
  if (size > 1) {
    PetscMPIInt *rc,*rd,*mc,*md;
    long long   *cnt,*allc,*order,*soff,*scnt,nsend = 0,total = 0,nrecv = 0,q;
    int         *irange;
    /* which of my elements touch rows of rank d != rank, grouped by destination (b200_batch_send_plan) */
    ierr = PetscMalloc5(size,long long,&cnt,size*size,long long,&allc,size,long long,&soff,size,long long,&scnt,size+1,int,&irange);CHKERRQ(ierr);
    for (s=0; s<=size; s++) irange[s] = (int)J->rmap->range[s];
    CHKB200(b200_batch_send_plan(size,rank,irange,(long long)Ne,(int)Nl,elemRows,cnt,0));
    for (s=0; s<size; s++) nsend += cnt[s];
    ierr = PetscMalloc((nsend+1)*sizeof(long long),&order);CHKERRQ(ierr);
    CHKB200(b200_batch_send_plan(size,rank,irange,(long long)Ne,(int)Nl,elemRows,cnt,order));
    ierr = MPI_Allgather(cnt,size,MPI_LONG_LONG_INT,allc,size,MPI_LONG_LONG_INT,comm);CHKERRQ(ierr);   /* allc[s*size+d] */
    ierr = PetscMalloc2((size_t)nsend*Nl+1,PetscInt,&sendrows,(size_t)nsend*Nl*Nl+1,PetscScalar,&sendmats);CHKERRQ(ierr);
    for (q=0; q<nsend; q++) {
      ierr = PetscMemcpy(sendrows+(size_t)q*Nl,elemRows+(size_t)order[q]*Nl,Nl*sizeof(PetscInt));CHKERRQ(ierr);
      ierr = PetscMemcpy(sendmats+(size_t)q*Nl*Nl,elemMats+(size_t)order[q]*Nl*Nl,Nl*Nl*sizeof(PetscScalar));CHKERRQ(ierr);
    }
    /* every rank publishes its one buffer (groups in destination order); where mine start: b200_batch_recv_plan */
    ierr = PetscMalloc4(size,PetscMPIInt,&rc,size,PetscMPIInt,&rd,size,PetscMPIInt,&mc,size,PetscMPIInt,&md);CHKERRQ(ierr);
    for (s=0; s<size; s++) {
      long long tot = 0;
      PetscInt  d;
      for (d=0; d<size; d++) tot += allc[(size_t)s*size+d];
      rc[s] = (PetscMPIInt)(tot*Nl); mc[s] = (PetscMPIInt)(tot*Nl*Nl);
      rd[s] = (PetscMPIInt)(total*Nl); md[s] = (PetscMPIInt)(total*Nl*Nl);
      total += tot;
    }
    CHKB200(b200_batch_recv_plan(size,rank,allc,soff,scnt));
    for (s=0; s<size; s++) nrecv += scnt[s];
    ierr = PetscMalloc2((size_t)total*Nl+1,PetscInt,&recvrows,(size_t)total*Nl*Nl+1,PetscScalar,&recvmats);CHKERRQ(ierr);
    ierr = MPI_Allgatherv(sendrows,(PetscMPIInt)(nsend*Nl),MPIU_INT,recvrows,rc,rd,MPIU_INT,comm);CHKERRQ(ierr);
    ierr = MPI_Allgatherv(sendmats,(PetscMPIInt)(nsend*Nl*Nl),MPIU_SCALAR,recvmats,mc,md,MPIU_SCALAR,comm);CHKERRQ(ierr);
    /* own elements first, then what ranks 0,1,... sent me, in their order */
    ne = Ne + (PetscInt)nrecv;
    ierr = PetscMalloc2((size_t)ne*Nl+1,PetscInt,&allrows,(size_t)ne*Nl*Nl+1,PetscScalar,&allmats);CHKERRQ(ierr);
    ierr = PetscMemcpy(allrows,elemRows,(size_t)Ne*Nl*sizeof(PetscInt));CHKERRQ(ierr);
    ierr = PetscMemcpy(allmats,elemMats,(size_t)Ne*Nl*Nl*sizeof(PetscScalar));CHKERRQ(ierr);
    for (s=0,p=Ne; s<size; s++) {
      if (!scnt[s]) continue;
      ierr = PetscMemcpy(allrows+(size_t)p*Nl,recvrows+rd[s]+(size_t)soff[s]*Nl,(size_t)scnt[s]*Nl*sizeof(PetscInt));CHKERRQ(ierr);
      ierr = PetscMemcpy(allmats+(size_t)p*Nl*Nl,recvmats+md[s]+(size_t)soff[s]*Nl*Nl,(size_t)scnt[s]*Nl*Nl*sizeof(PetscScalar));CHKERRQ(ierr);
      p += (PetscInt)scnt[s];
    }
    rows = allrows; mats = allmats;
    ierr = PetscFree4(rc,rd,mc,md);CHKERRQ(ierr);
    ierr = PetscFree2(recvrows,recvmats);CHKERRQ(ierr);
    ierr = PetscFree2(sendrows,sendmats);CHKERRQ(ierr);
    ierr = PetscFree(order);CHKERRQ(ierr);
    ierr = PetscFree5(cnt,allc,soff,scnt,irange);CHKERRQ(ierr);
  }
  ierr = PetscInfo3(J,"Assembling rows %D..%D from %D element matrices on the device\n",rstart,rend,ne);CHKERRQ(ierr);
  CHKB200(b200_mat_setvaluesbatch_rows(m,rstart,N,(long long)ne,Nl,rows,mats,0,0,&dm));
  CHKB200(b200_mat_get_size(dm,0,0,&nz));
  ierr = PetscMalloc3(m+1,PetscInt,&ci,(size_t)nz+1,PetscInt,&cj,(size_t)nz+1,PetscScalar,&ca);CHKERRQ(ierr);
  CHKB200(b200_mat_get_csr(dm,ci,cj,ca));
  CHKB200(b200_mat_destroy(dm));
  if (size > 1) { ierr = PetscFree2(allrows,allmats);CHKERRQ(ierr); }
  ierr = MatMPIAIJSetPreallocationCSR(J,ci,cj,ca);CHKERRQ(ierr);                  /* splits into the two blocks, assembles */
  ierr = PetscFree3(ci,cj,ca);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* the device operator behind a seqaijb200 / mpiaijb200 matrix (uploaded or refreshed first), for the KSP types */
#undef __FUNCT__
#define __FUNCT__ "MatB200GetHandle"
PetscErrorCode MatB200GetHandle(Mat A,void **handle)
{
  PetscBool      isseq,ismpi;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscObjectTypeCompare((PetscObject)A,MATSEQAIJB200,&isseq);CHKERRQ(ierr);
  ierr = PetscObjectTypeCompare((PetscObject)A,MATMPIAIJB200,&ismpi);CHKERRQ(ierr);
  if (isseq) {
    ierr = MatB200CopyToDevice_SeqAIJ(A);CHKERRQ(ierr);
    *handle = (void*)((Mat_SeqAIJB200*)A->spptr)->mat;
  } else if (ismpi) {
    ierr = MatB200CopyToDevice_MPIAIJ(A);CHKERRQ(ierr);
    *handle = (void*)((Mat_MPIAIJB200*)((Mat_MPIAIJ*)A->data)->spptr)->mat;
  } else SETERRQ1(((PetscObject)A)->comm,PETSC_ERR_SUP,"Matrix of type %s where a B200 matrix (aijb200) is needed",((PetscObject)A)->type_name);
  PetscFunctionReturn(0);
}

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "MatCreate_MPIAIJB200"
PetscErrorCode MatCreate_MPIAIJB200(Mat A)
{
  Mat_MPIAIJ     *a;
  Mat_MPIAIJB200 *d;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MatCreate_MPIAIJ(A);CHKERRQ(ierr);
  ierr = PetscObjectComposeFunctionDynamic((PetscObject)A,"MatMPIAIJSetPreallocation_C","MatMPIAIJSetPreallocation_MPIAIJB200",MatMPIAIJSetPreallocation_MPIAIJB200);CHKERRQ(ierr);
  a = (Mat_MPIAIJ*)A->data;
  ierr = PetscNewLog(A,Mat_MPIAIJB200,&d);CHKERRQ(ierr);
  a->spptr        = (void*)d;
  A->ops->mult    = MatMult_MPIAIJB200;
  A->ops->multadd = MatMultAdd_MPIAIJB200;
  A->ops->multtranspose    = MatMultTranspose_MPIAIJB200;
  A->ops->multtransposeadd = MatMultTransposeAdd_MPIAIJB200;
  A->ops->getvecs = MatGetVecs_MPIAIJB200;
  A->ops->destroy = MatDestroy_MPIAIJB200;
  A->ops->setvaluesbatch = MatSetValuesBatch_MPIAIJB200;                          /* mpiaijcusp.cu:217 */
  ierr = PetscObjectChangeTypeName((PetscObject)A,MATMPIAIJB200);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
EXTERN_C_END
