/* b200scatter.c — PETSc's own VecScatter object routed to the device exchange engine.
 *
 * VecScatterBegin/End dispatch through ctx->begin / ctx->end of struct _p_VecScatter
 * (include/petsc-private/vecimpl.h:539-554; src/vec/vec/utils/vscat.c:1627,1698).  The reference's GPU seam for this object is
 * VecScatterInitializeForGPU (src/vec/vec/utils/veccusp/vscatcusp.c:29-133), which hangs device index lists on ctx->spptr
 * and still messages through the host.  VecScatterB200Attach() uses the same slot: it keeps the scatter's original
 * function pointers in a hook stored in ctx->spptr and installs wrappers.  When BOTH vectors of a VecScatterBegin are
 * B200 vectors the wrappers run the scatter on the device — pack, GPU-to-GPU transfer, unpack through
 * b200_scatter_begin/end of the C ABI — and nothing touches the host; in every other case (host vectors, MAX_VALUES,
 * SCATTER_LOCAL, ToAll/ToZero scatters, x == y) they call the saved originals, so behaviour is unchanged.
 *
 * Covered layouts: the general parallel scatter (VEC_SCATTER_MPI_GENERAL on both sides: PtoS, StoP and PtoP,
 * vpscat.c:1730-2414, any block size) and the four sequential ones (general/stride index sets on either side,
 * vscat.c:344-640).  The device plan is built from todata/fromdata at the first device scatter.  VecScatterCopy of an
 * attached scatter yields an attached scatter.  After VecScatterRemap (which rewrites the lists in place) call
 * VecScatterB200Attach again to drop the stale plan. */
#include "petscb200plugin.h"
#include <petsc-private/vecimpl.h>
#include "petscb200.h"

#define CHKB200(call) do { int b200rc_ = (call); if (b200rc_) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_LIB,"libpetscb200 error %d: %s",b200rc_,b200_last_error()); } while (0)

typedef struct {
  PetscErrorCode (*begin)(VecScatter,Vec,Vec,InsertMode,ScatterMode);
  PetscErrorCode (*end)(VecScatter,Vec,Vec,InsertMode,ScatterMode);
  PetscErrorCode (*copy)(VecScatter,VecScatter);
  PetscErrorCode (*destroy)(VecScatter);
  b200_scatter   plan;           /* built at the first device scatter                                               */
  PetscInt       nx,ny;          /* local sizes the plan was built for (forward orientation: x source, y target)   */
  PetscBool      parallel;       /* VEC_SCATTER_MPI_GENERAL: messages travel; otherwise purely local                */
  PetscBool      supported;      /* layout the device engine covers                                                 */
  PetscBool      active;         /* the Begin in flight went to the device: End must too                            */
} VecScatterB200Hook;

static PetscInt b200_scatters_device = 0, b200_scatters_host = 0;

static PetscErrorCode VecScatterBegin_B200(VecScatter,Vec,Vec,InsertMode,ScatterMode);
static PetscErrorCode VecScatterEnd_B200(VecScatter,Vec,Vec,InsertMode,ScatterMode);
static PetscErrorCode VecScatterCopy_B200(VecScatter,VecScatter);
static PetscErrorCode VecScatterDestroy_B200(VecScatter);

#undef __FUNCT__
#define __FUNCT__ "PetscB200GetScatterCounters"
PetscErrorCode PetscB200GetScatterCounters(PetscInt *ondevice,PetscInt *onhost)
{
  PetscFunctionBegin;
  if (ondevice) *ondevice = b200_scatters_device;
  if (onhost)   *onhost   = b200_scatters_host;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200IsDeviceVec"
static PetscBool VecB200IsDeviceVec(Vec v)
{
  const char *t = ((PetscObject)v)->type_name;
  return (t && (!strcmp(t,VECSEQB200) || !strcmp(t,VECMPIB200))) ? PETSC_TRUE : PETSC_FALSE;
}

/* index list of one side of a sequential scatter: general (vslots) or stride (first + i*step) */
#undef __FUNCT__
#define __FUNCT__ "VecScatterB200SeqIndices"
static PetscErrorCode VecScatterB200SeqIndices(void *data,PetscInt *n,int **idx)
{
  VecScatterType type = ((VecScatter_Seq_General*)data)->type;
  PetscInt       i;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (type == VEC_SCATTER_SEQ_GENERAL) {
    VecScatter_Seq_General *g = (VecScatter_Seq_General*)data;
    *n = g->n;
    ierr = PetscMalloc((g->n+1)*sizeof(int),idx);CHKERRQ(ierr);
    for (i=0; i<g->n; i++) (*idx)[i] = (int)g->vslots[i];
  } else if (type == VEC_SCATTER_SEQ_STRIDE) {
    VecScatter_Seq_Stride *s = (VecScatter_Seq_Stride*)data;
    *n = s->n;
    ierr = PetscMalloc((s->n+1)*sizeof(int),idx);CHKERRQ(ierr);
    for (i=0; i<s->n; i++) (*idx)[i] = (int)(s->first + i*s->step);
  } else SETERRQ(PETSC_COMM_SELF,PETSC_ERR_PLIB,"Not a sequential scatter side");
  PetscFunctionReturn(0);
}

/* device plan from the scatter's own lists (forward orientation).  Parallel: what MatB200CopyToDevice_MPIAIJ reads for
 * the MatMult scatter, plus the receive positions and the self part; blocks expanded to scalar entries (Pack_bs /
 * UnPack_bs / Scatter_bs, vpscat.c:674-760).  Collective for a parallel scatter. */
#undef __FUNCT__
#define __FUNCT__ "VecScatterB200BuildPlan"
static PetscErrorCode VecScatterB200BuildPlan(VecScatter ctx,VecScatterB200Hook *h,PetscInt nx,PetscInt ny)
{
  PetscErrorCode ierr;
  PetscInt       i,k,c;

  PetscFunctionBegin;
  if (h->plan) { CHKB200(b200_scatter_destroy(h->plan)); h->plan = 0; }
  if (h->parallel) {
    VecScatter_MPI_General *to = (VecScatter_MPI_General*)ctx->todata,*from = (VecScatter_MPI_General*)ctx->fromdata;
    PetscInt               bs = to->bs,nto = to->n,nfrom = from->n,ns = to->starts[nto],nr = from->starts[nfrom],nl = to->local.n;
    int                    *to_procs,*to_starts,*to_indices,*from_procs,*from_starts,*from_indices,*to_local,*from_local;
    ierr = PetscMalloc4(nto+1,int,&to_procs,nto+1,int,&to_starts,nfrom+1,int,&from_procs,nfrom+1,int,&from_starts);CHKERRQ(ierr);
    ierr = PetscMalloc4(bs*ns+1,int,&to_indices,bs*nr+1,int,&from_indices,bs*nl+1,int,&to_local,bs*nl+1,int,&from_local);CHKERRQ(ierr);
    for (i=0; i<nto; i++)    to_procs[i]    = (int)to->procs[i];
    for (i=0; i<=nto; i++)   to_starts[i]   = (int)(bs*to->starts[i]);
    for (i=0; i<nfrom; i++)  from_procs[i]  = (int)from->procs[i];
    for (i=0; i<=nfrom; i++) from_starts[i] = (int)(bs*from->starts[i]);
    for (k=0; k<ns; k++) for (c=0; c<bs; c++) to_indices[bs*k+c]   = (int)(to->indices[k]+c);
    for (k=0; k<nr; k++) for (c=0; c<bs; c++) from_indices[bs*k+c] = (int)(from->indices[k]+c);
    for (k=0; k<nl; k++) for (c=0; c<bs; c++) { to_local[bs*k+c] = (int)(to->local.vslots[k]+c); from_local[bs*k+c] = (int)(from->local.vslots[k]+c); }
    CHKB200(b200_scatter_create((int)nx,(int)ny,(int)nto,to_procs,to_starts,to_indices,(int)nfrom,from_procs,from_starts,from_indices,
                                (int)(bs*nl),to_local,from_local,&h->plan));
    ierr = PetscInfo5(ctx,"B200 scatter: device plan, %D neighbours out (%D entries), %D in (%D entries), %D local entries\n",nto,bs*ns,nfrom,bs*nr,bs*nl);CHKERRQ(ierr);
    ierr = PetscFree4(to_indices,from_indices,to_local,from_local);CHKERRQ(ierr);
    ierr = PetscFree4(to_procs,to_starts,from_procs,from_starts);CHKERRQ(ierr);
  } else {
    PetscInt nf,nt;
    int      *src,*dst;
    ierr = VecScatterB200SeqIndices(ctx->fromdata,&nf,&src);CHKERRQ(ierr);
    ierr = VecScatterB200SeqIndices(ctx->todata,&nt,&dst);CHKERRQ(ierr);
    if (nf != nt) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_PLIB,"Sequential scatter sides differ in length: %D and %D",nf,nt);
    CHKB200(b200_scatter_create_local((int)nx,(int)ny,(int)nf,src,dst,&h->plan));
    ierr = PetscInfo1(ctx,"B200 scatter: device plan, %D local entries\n",nf);CHKERRQ(ierr);
    ierr = PetscFree(src);CHKERRQ(ierr);
    ierr = PetscFree(dst);CHKERRQ(ierr);
  }
  h->nx = nx; h->ny = ny;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecScatterBegin_B200"
static PetscErrorCode VecScatterBegin_B200(VecScatter ctx,Vec xin,Vec yin,InsertMode addv,ScatterMode mode)
{
  VecScatterB200Hook *h = (VecScatterB200Hook*)ctx->spptr;
  PetscErrorCode     ierr;

  PetscFunctionBegin;
  h->active = PETSC_FALSE;
  if (h->supported && xin != yin && (addv == INSERT_VALUES || addv == ADD_VALUES) && !(mode & SCATTER_LOCAL) &&
      VecB200IsDeviceVec(xin) && VecB200IsDeviceVec(yin)) {
    const PetscBool   reverse = (mode & SCATTER_REVERSE) ? PETSC_TRUE : PETSC_FALSE;
    Vec               X = reverse ? yin : xin,Y = reverse ? xin : yin;    /* forward orientation: X source side, Y target side */
    const PetscScalar *src;
    PetscScalar       *dst;
    if (!h->plan || h->nx != X->map->n || h->ny != Y->map->n) { ierr = VecScatterB200BuildPlan(ctx,h,X->map->n,Y->map->n);CHKERRQ(ierr); }
    ierr = VecB200GetDeviceRead(xin,&src);CHKERRQ(ierr);
    ierr = VecB200GetDeviceReadWrite(yin,&dst);CHKERRQ(ierr);          /* entries the scatter does not name keep their values */
    if (reverse) { CHKB200(b200_scatter_begin(h->plan,dst,(double*)src,addv == ADD_VALUES,1)); }
    else         { CHKB200(b200_scatter_begin(h->plan,(double*)src,dst,addv == ADD_VALUES,0)); }
    h->active = PETSC_TRUE;
    b200_scatters_device++;
    if (!h->parallel) { ierr = PetscObjectStateIncrease((PetscObject)yin);CHKERRQ(ierr); }   /* VecRestoreArray(yin) of the host loops */
    PetscFunctionReturn(0);
  }
  b200_scatters_host++;
  ierr = (*h->begin)(ctx,xin,yin,addv,mode);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecScatterEnd_B200"
static PetscErrorCode VecScatterEnd_B200(VecScatter ctx,Vec xin,Vec yin,InsertMode addv,ScatterMode mode)
{
  VecScatterB200Hook *h = (VecScatterB200Hook*)ctx->spptr;
  PetscErrorCode     ierr;

  PetscFunctionBegin;
  if (h->active) {
    const PetscBool   reverse = (mode & SCATTER_REVERSE) ? PETSC_TRUE : PETSC_FALSE;
    const PetscScalar *src;
    PetscScalar       *dst;
    h->active = PETSC_FALSE;
    ierr = VecB200GetDeviceRead(xin,&src);CHKERRQ(ierr);
    ierr = VecB200GetDeviceReadWrite(yin,&dst);CHKERRQ(ierr);
    if (reverse) { CHKB200(b200_scatter_end(h->plan,dst,(double*)src,addv == ADD_VALUES,1)); }
    else         { CHKB200(b200_scatter_end(h->plan,(double*)src,dst,addv == ADD_VALUES,0)); }
    ierr = PetscObjectStateIncrease((PetscObject)yin);CHKERRQ(ierr);
    PetscFunctionReturn(0);
  }
  if (h->end) { ierr = (*h->end)(ctx,xin,yin,addv,mode);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

/* the scatter's own copy routine copies the function pointers of `in` (vscat.c:643-660, vpscat.c:280-400): run it on
 * the originals, then attach the copy as well */
#undef __FUNCT__
#define __FUNCT__ "VecScatterCopy_B200"
static PetscErrorCode VecScatterCopy_B200(VecScatter in,VecScatter out)
{
  VecScatterB200Hook *h = (VecScatterB200Hook*)in->spptr;
  PetscErrorCode     ierr;

  PetscFunctionBegin;
  in->begin = h->begin; in->end = h->end; in->copy = h->copy; in->destroy = h->destroy; in->spptr = 0;
  ierr = (*h->copy)(in,out);
  in->begin = VecScatterBegin_B200; in->end = h->end ? VecScatterEnd_B200 : (h->parallel ? VecScatterEnd_B200 : 0);
  in->copy = VecScatterCopy_B200; in->destroy = VecScatterDestroy_B200; in->spptr = h;
  CHKERRQ(ierr);
  out->spptr = 0;
  ierr = VecScatterB200Attach(out);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecScatterDestroy_B200"
static PetscErrorCode VecScatterDestroy_B200(VecScatter ctx)
{
  VecScatterB200Hook *h = (VecScatterB200Hook*)ctx->spptr;
  PetscErrorCode     (*destroy)(VecScatter) = h->destroy;
  PetscErrorCode     ierr;

  PetscFunctionBegin;
  if (h->plan) { CHKB200(b200_scatter_destroy(h->plan)); }
  ctx->spptr = 0;
  ierr = PetscFree(h);CHKERRQ(ierr);
  ierr = (*destroy)(ctx);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* VecScatterB200Attach - route this scatter to the device whenever both vectors of a Begin/End are B200 vectors.
 * Harmless on any scatter (unsupported layouts and host vectors keep the reference's code); calling it again drops the
 * device plan (needed after VecScatterRemap).  Not collective; the plan itself is built collectively at the first
 * device scatter of a parallel scatter. */
#undef __FUNCT__
#define __FUNCT__ "VecScatterB200Attach"
PetscErrorCode VecScatterB200Attach(VecScatter ctx)
{
  VecScatterB200Hook *h;
  VecScatterType     tt,ft;
  PetscErrorCode     ierr;

  PetscFunctionBegin;
  PetscValidHeaderSpecific(ctx,VEC_SCATTER_CLASSID,1);
  if (ctx->begin == VecScatterBegin_B200) {                               /* attached already: forget the plan */
    h = (VecScatterB200Hook*)ctx->spptr;
    if (h->plan) { CHKB200(b200_scatter_destroy(h->plan)); h->plan = 0; }
    PetscFunctionReturn(0);
  }
  if (ctx->spptr || !ctx->begin || !ctx->todata || !ctx->fromdata) PetscFunctionReturn(0);
  ierr = PetscNew(VecScatterB200Hook,&h);CHKERRQ(ierr);
  h->begin = ctx->begin; h->end = ctx->end; h->copy = ctx->copy; h->destroy = ctx->destroy;
  tt = ((VecScatter_Seq_General*)ctx->todata)->type; ft = ((VecScatter_Seq_General*)ctx->fromdata)->type;   /* first member of every layout */
  if (tt == VEC_SCATTER_MPI_GENERAL && ft == VEC_SCATTER_MPI_GENERAL) {
    PetscMPIInt size,wsize;
    ierr = MPI_Comm_size(((PetscObject)ctx)->comm,&size);CHKERRQ(ierr);
    ierr = MPI_Comm_size(PETSC_COMM_WORLD,&wsize);CHKERRQ(ierr);
    h->parallel  = PETSC_TRUE;
    h->supported = (size == wsize) ? PETSC_TRUE : PETSC_FALSE;            /* the device transport spans PETSC_COMM_WORLD */
  } else if ((tt == VEC_SCATTER_SEQ_GENERAL || tt == VEC_SCATTER_SEQ_STRIDE) && (ft == VEC_SCATTER_SEQ_GENERAL || ft == VEC_SCATTER_SEQ_STRIDE)) {
    h->parallel = PETSC_FALSE; h->supported = PETSC_TRUE;
  } else h->supported = PETSC_FALSE;                                      /* ToAll / ToZero: the reference's code */
  ctx->spptr   = h;
  ctx->begin   = VecScatterBegin_B200;
  ctx->end     = (h->end || h->parallel) ? VecScatterEnd_B200 : 0;
  ctx->copy    = h->copy ? VecScatterCopy_B200 : 0;
  ctx->destroy = VecScatterDestroy_B200;
  PetscFunctionReturn(0);
}
