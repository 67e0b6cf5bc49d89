/* b200vec.c — VECSEQB200 / VECMPIB200 / VECB200: PETSc vectors whose data lives on a B200.
 *
 * Mirrors VecCreate_SeqCUSP / VecCreate_MPICUSP / VecCreate_CUSP (reference
 * src/vec/vec/impls/seq/seqcusp/veccusp.cu:1903-1947, src/vec/vec/impls/mpi/mpicusp/mpicusp.cu:187-244): inherit the
 * host type (VecCreate_Seq_Private bvec2.c:1252, VecCreate_MPI_Private pbvec.c:175), override the ops table, keep the
 * data on the device.  Every device op is one call of the C ABI (include/petscb200.h); flop logging follows the
 * reference counts so -log_summary stays comparable (SURVEY 8(b) "Logging obligations").
 *
 * Coherence: `where` says which copy is current.  Host access goes through ops->getarray (petscnative is FALSE), which
 * copies back when the device is newer.  VecGetArrayRead and VecGetArray reach the same op (vecimpl.h:382,411), but
 * only VecRestoreArray bumps the object state (vecimpl.h:430-432 against :395-401): getarray records the state and
 * leaves both copies valid; the next access of any kind compares — an unchanged state means the host only read, and
 * the device copy is still current (no re-upload after a VecGetArrayRead, the read/write distinction the CUSP seam has
 * through valid_GPU_array, vecimpl.h:375-379,410-414).  Inherited host ops that read `->array` directly (VecSetValues_Seq, VecView,
 * VecMax ...) are wrapped so that they see current data.  The host array is pinned and allocated lazily (KSP work
 * vectors never get one), as VecCUSPAllocateCheckHost does (veccusp.cu:34-60). */
#include "petscb200plugin.h"
#include <petsc-private/vecimpl.h>
#include <../src/vec/vec/impls/dvecimpl.h>
#include <../src/vec/vec/impls/mpi/pvecimpl.h>
#include "petscb200.h"

typedef struct {
  PetscScalar  *d;         /* device array, map->n entries                                  */
  PetscScalar  *pinned;    /* the lazily allocated pinned host array (also VECHEADER.array) */
  B200VecWhere  where;
  PetscBool     hostpending; /* the host array was handed out while both copies were current ...               */
  PetscInt      hoststate;   /* ... at this object state: a changed state since means it was written (below)  */
} B200VecData;

typedef struct { Vec_Seq base; B200VecData dev; } Vec_SeqB200;
typedef struct { Vec_MPI base; B200VecData dev; } Vec_MPIB200;

static PetscInt b200_h2d_count = 0, b200_d2h_count = 0;
PetscInt        PetscB200MatUploads = 0, PetscB200MatValueUpdates = 0;

#define CHKB200(call) do { int b200rc_ = (call); if (b200rc_) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_LIB,"libpetscb200 error %d: %s",b200rc_,b200_last_error()); } while (0)

static PetscErrorCode VecDestroy_SeqB200(Vec);
static PetscErrorCode VecDestroy_MPIB200(Vec);

#undef __FUNCT__
#define __FUNCT__ "VecB200Data"
PETSC_STATIC_INLINE PetscErrorCode VecB200Data(Vec v,B200VecData **dev)
{
  PetscFunctionBegin;
  if (v->ops->destroy == VecDestroy_SeqB200)      *dev = &((Vec_SeqB200*)v->data)->dev;
  else if (v->ops->destroy == VecDestroy_MPIB200) *dev = &((Vec_MPIB200*)v->data)->dev;
  else SETERRQ1(((PetscObject)v)->comm,PETSC_ERR_ARG_WRONG,"Vector of type %s where a B200 vector is needed",((PetscObject)v)->type_name);
  PetscFunctionReturn(0);
}

/* settle a pending host hand-out: state moved on => someone called VecRestoreArray (or changed the vector through the
 * host some other way) => the host copy is the newer one.  Called first thing by every access. */
PETSC_STATIC_INLINE void VecB200ResolveHostAccess(Vec v,B200VecData *dev)
{
  if (dev->hostpending) {
    dev->hostpending = PETSC_FALSE;
    if (dev->where == B200_VEC_BOTH && ((PetscObject)v)->state != dev->hoststate) dev->where = B200_VEC_HOST;
  }
}

/* ---- allocation and the two copies ---------------------------------------------------------------------------- */
#undef __FUNCT__
#define __FUNCT__ "VecB200AllocateCheck"
static PetscErrorCode VecB200AllocateCheck(Vec v,B200VecData *dev)      /* VecCUSPAllocateCheck, veccusp.cu:68 */
{
  PetscFunctionBegin;
  if (!dev->d) {
    CHKB200(b200_malloc(sizeof(PetscScalar)*(size_t)(v->map->n > 0 ? v->map->n : 1),(void**)&dev->d));
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200AllocateCheckHost"
static PetscErrorCode VecB200AllocateCheckHost(Vec v,B200VecData *dev)  /* VecCUSPAllocateCheckHost, veccusp.cu:34 */
{
  Vec_Seq *s = (Vec_Seq*)v->data;                                       /* VECHEADER is the first member of both */

  PetscFunctionBegin;
  if (!s->array) {
    PetscInt n = v->map->n;
    if (v->ops->destroy == VecDestroy_MPIB200) n += ((Vec_MPI*)v->data)->nghost;
    CHKB200(b200_host_alloc(sizeof(PetscScalar)*(size_t)(n > 0 ? n : 1),(void**)&dev->pinned));
    s->array = dev->pinned;                                             /* array_allocated stays 0: we free it */
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200CopyToDevice"
static PetscErrorCode VecB200CopyToDevice(Vec v,B200VecData *dev)       /* VecCUSPCopyToGPU, veccusp.cu:109 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200AllocateCheck(v,dev);CHKERRQ(ierr);
  if (dev->where == B200_VEC_HOST) {
    CHKB200(b200_memcpy_h2d(dev->d,((Vec_Seq*)v->data)->array,sizeof(PetscScalar)*(size_t)v->map->n));
    b200_h2d_count++;
    dev->where = B200_VEC_BOTH;
    ierr = PetscInfo1(v,"B200 vector upload (%D entries)\n",v->map->n);CHKERRQ(ierr);
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200CopyFromDevice"
static PetscErrorCode VecB200CopyFromDevice(Vec v,B200VecData *dev)     /* VecCUSPCopyFromGPU, veccusp.cu:173 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200AllocateCheckHost(v,dev);CHKERRQ(ierr);
  if (dev->where == B200_VEC_DEVICE) {
    CHKB200(b200_memcpy_d2h(((Vec_Seq*)v->data)->array,dev->d,sizeof(PetscScalar)*(size_t)v->map->n));
    b200_d2h_count++;
    dev->where = B200_VEC_BOTH;
    ierr = PetscInfo1(v,"B200 vector download (%D entries)\n",v->map->n);CHKERRQ(ierr);
  } else if (dev->where == B200_VEC_UNALLOCATED) {
    ierr = PetscMemzero(((Vec_Seq*)v->data)->array,sizeof(PetscScalar)*(size_t)v->map->n);CHKERRQ(ierr);
    dev->where = B200_VEC_HOST;
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200GetDeviceRead"
PetscErrorCode VecB200GetDeviceRead(Vec v,const PetscScalar **a)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  VecB200ResolveHostAccess(v,dev);
  ierr = VecB200CopyToDevice(v,dev);CHKERRQ(ierr);
  if (dev->where == B200_VEC_UNALLOCATED) {
    CHKB200(b200_vec_set(v->map->n,dev->d,0.0));
    dev->where = B200_VEC_DEVICE;
  }
  *a = dev->d;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200GetDeviceWrite"
PetscErrorCode VecB200GetDeviceWrite(Vec v,PetscScalar **a)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  dev->hostpending = PETSC_FALSE;
  ierr = VecB200AllocateCheck(v,dev);CHKERRQ(ierr);
  dev->where = B200_VEC_DEVICE;
  /* what VecRestoreArray does for the host loops (vecimpl.h:430) and VecCUSPRestoreArrayWrite for the CUSP type: callers
     outside the Vec front ends (MatMult in 3.3 does not touch y's state, matrix.c:2154) must still invalidate cached norms */
  ierr = PetscObjectStateIncrease((PetscObject)v);CHKERRQ(ierr);
  *a = dev->d;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200GetDeviceReadWrite"
PetscErrorCode VecB200GetDeviceReadWrite(Vec v,PetscScalar **a)
{
  const PetscScalar *r;
  B200VecData       *dev;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = VecB200GetDeviceRead(v,&r);CHKERRQ(ierr);
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  dev->where = B200_VEC_DEVICE;
  ierr = PetscObjectStateIncrease((PetscObject)v);CHKERRQ(ierr);
  *a = (PetscScalar*)r;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecB200GetWhere"
PetscErrorCode VecB200GetWhere(Vec v,B200VecWhere *w)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  VecB200ResolveHostAccess(v,dev);
  *w = dev->where;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "PetscB200GetCounters"
PetscErrorCode PetscB200GetCounters(PetscInt *h2d,PetscInt *d2h,PetscInt *up,PetscInt *vup)
{
  PetscFunctionBegin;
  if (h2d) *h2d = b200_h2d_count;
  if (d2h) *d2h = b200_d2h_count;
  if (up)  *up  = PetscB200MatUploads;
  if (vup) *vup = PetscB200MatValueUpdates;
  PetscFunctionReturn(0);
}

/* host access: current data on the host, host becomes the newest copy */
#undef __FUNCT__
#define __FUNCT__ "VecB200HostReadWrite"
static PetscErrorCode VecB200HostReadWrite(Vec v)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  VecB200ResolveHostAccess(v,dev);
  ierr = VecB200CopyFromDevice(v,dev);CHKERRQ(ierr);
  dev->where = B200_VEC_HOST;
  PetscFunctionReturn(0);
}

/* host access that does not modify: both copies stay valid */
#undef __FUNCT__
#define __FUNCT__ "VecB200HostRead"
static PetscErrorCode VecB200HostRead(Vec v)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  VecB200ResolveHostAccess(v,dev);
  ierr = VecB200CopyFromDevice(v,dev);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* VecGetArray AND VecGetArrayRead (vecimpl.h:382,411): current data on the host; whether the caller writes shows in the
 * object state at the next access (VecRestoreArray increases it, VecRestoreArrayRead does not) */
#undef __FUNCT__
#define __FUNCT__ "VecGetArray_B200"
static PetscErrorCode VecGetArray_B200(Vec v,PetscScalar **a)
{
  B200VecData    *dev;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200Data(v,&dev);CHKERRQ(ierr);
  VecB200ResolveHostAccess(v,dev);
  ierr = VecB200CopyFromDevice(v,dev);CHKERRQ(ierr);
  if (dev->where == B200_VEC_BOTH) { dev->hostpending = PETSC_TRUE; dev->hoststate = ((PetscObject)v)->state; }
  else dev->where = B200_VEC_HOST;
  *a = ((Vec_Seq*)v->data)->array;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecRestoreArray_B200"
static PetscErrorCode VecRestoreArray_B200(Vec v,PetscScalar **a)
{
  PetscFunctionBegin;
  PetscFunctionReturn(0);
}

/* ---- device ops (each shadows the VECSEQ op named; the C ABI carries the reference's special cases) ------------ */
#define DEVR(v,p)  do { ierr = VecB200GetDeviceRead(v,&(p));CHKERRQ(ierr); } while (0)
#define DEVW(v,p)  do { ierr = VecB200GetDeviceWrite(v,&(p));CHKERRQ(ierr); } while (0)
#define DEVRW(v,p) do { ierr = VecB200GetDeviceReadWrite(v,&(p));CHKERRQ(ierr); } while (0)

#undef __FUNCT__
#define __FUNCT__ "VecSet_SeqB200"
static PetscErrorCode VecSet_SeqB200(Vec x,PetscScalar alpha)                    /* VecSet_Seq dvec2.c:722 */
{
  PetscScalar    *xd;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  DEVW(x,xd);
  CHKB200(b200_vec_set(x->map->n,xd,alpha));
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecCopy_SeqB200"
static PetscErrorCode VecCopy_SeqB200(Vec x,Vec y)                               /* VecCopy_Seq bvec2.c:464 */
{
  const PetscScalar *xd;
  PetscScalar       *yd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  if (x != y) {
    if (y->ops->copy != VecCopy_SeqB200) {                                       /* into a host vector */
      PetscScalar *ya;
      ierr = VecB200HostRead(x);CHKERRQ(ierr);
      ierr = VecGetArray(y,&ya);CHKERRQ(ierr);
      ierr = PetscMemcpy(ya,((Vec_Seq*)x->data)->array,x->map->n*sizeof(PetscScalar));CHKERRQ(ierr);
      ierr = VecRestoreArray(y,&ya);CHKERRQ(ierr);
      PetscFunctionReturn(0);
    }
    DEVR(x,xd); DEVW(y,yd);
    CHKB200(b200_vec_copy(x->map->n,xd,yd));
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecSwap_SeqB200"
static PetscErrorCode VecSwap_SeqB200(Vec x,Vec y)                               /* VecSwap_Seq bvec2.c:519 */
{
  PetscScalar    *xd,*yd;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (x != y) {
    DEVRW(x,xd); DEVRW(y,yd);
    CHKB200(b200_vec_swap(x->map->n,xd,yd));
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecScale_SeqB200"
static PetscErrorCode VecScale_SeqB200(Vec x,PetscScalar alpha)                  /* VecScale_Seq bvec1.c:183 */
{
  PetscScalar    *xd;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (alpha == 0.0) { ierr = VecSet_SeqB200(x,0.0);CHKERRQ(ierr); }
  else if (alpha != 1.0) {
    DEVRW(x,xd);
    CHKB200(b200_vec_scale(x->map->n,xd,alpha));
    ierr = PetscLogFlops(x->map->n);CHKERRQ(ierr);
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecAXPY_SeqB200"
static PetscErrorCode VecAXPY_SeqB200(Vec y,PetscScalar alpha,Vec x)             /* VecAXPY_Seq bvec1.c:244 */
{
  const PetscScalar *xd;
  PetscScalar       *yd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  if (alpha != 0.0) {                                                            /* bvec1.c:253 */
    DEVR(x,xd); DEVRW(y,yd);
    CHKB200(b200_vec_axpy(y->map->n,yd,alpha,xd));
    ierr = PetscLogFlops(2.0*y->map->n);CHKERRQ(ierr);
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecAYPX_SeqB200"
static PetscErrorCode VecAYPX_SeqB200(Vec y,PetscScalar alpha,Vec x)             /* VecAYPX_Seq dvec2.c:971 */
{
  const PetscScalar *xd;
  PetscScalar       *yd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd);
  if (alpha == 0.0) DEVW(y,yd); else DEVRW(y,yd);
  CHKB200(b200_vec_aypx(y->map->n,yd,alpha,xd));
  if (alpha == 1.0 || alpha == -1.0) { ierr = PetscLogFlops(1.0*y->map->n);CHKERRQ(ierr); }
  else if (alpha != 0.0)             { ierr = PetscLogFlops(2.0*y->map->n);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecAXPBY_SeqB200"
static PetscErrorCode VecAXPBY_SeqB200(Vec y,PetscScalar alpha,PetscScalar beta,Vec x)   /* VecAXPBY_Seq bvec1.c:320 */
{
  const PetscScalar *xd;
  PetscScalar       *yd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVRW(y,yd);
  CHKB200(b200_vec_axpby(y->map->n,yd,alpha,beta,xd));
  ierr = PetscLogFlops(3.0*y->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecWAXPY_SeqB200"
static PetscErrorCode VecWAXPY_SeqB200(Vec w,PetscScalar alpha,Vec x,Vec y)      /* VecWAXPY_Seq dvec2.c:1082 */
{
  const PetscScalar *xd,*yd;
  PetscScalar       *wd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVR(y,yd); DEVW(w,wd);
  CHKB200(b200_vec_waxpy(w->map->n,wd,alpha,xd,yd));
  if (alpha == 1.0 || alpha == -1.0) { ierr = PetscLogFlops(1.0*w->map->n);CHKERRQ(ierr); }
  else if (alpha != 0.0)             { ierr = PetscLogFlops(2.0*w->map->n);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecAXPBYPCZ_SeqB200"
static PetscErrorCode VecAXPBYPCZ_SeqB200(Vec z,PetscScalar alpha,PetscScalar beta,PetscScalar gamma,Vec x,Vec y) /* bvec1.c:418 */
{
  const PetscScalar *xd,*yd;
  PetscScalar       *zd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVR(y,yd); DEVRW(z,zd);
  CHKB200(b200_vec_axpbypcz(z->map->n,zd,alpha,beta,gamma,xd,yd));
  ierr = PetscLogFlops(5.0*z->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecPointwiseMult_SeqB200"
static PetscErrorCode VecPointwiseMult_SeqB200(Vec w,Vec x,Vec y)                /* VecPointwiseMult_Seq bvec2.c:234 */
{
  const PetscScalar *xd,*yd;
  PetscScalar       *wd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVR(y,yd);
  if (w == x || w == y) DEVRW(w,wd); else DEVW(w,wd);
  CHKB200(b200_vec_pointwise_mult(w->map->n,wd,xd,yd));
  ierr = PetscLogFlops(w->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecPointwiseDivide_SeqB200"
static PetscErrorCode VecPointwiseDivide_SeqB200(Vec w,Vec x,Vec y)              /* VecPointwiseDivide_Seq bvec2.c:298 */
{
  const PetscScalar *xd,*yd;
  PetscScalar       *wd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVR(y,yd);
  if (w == x || w == y) DEVRW(w,wd); else DEVW(w,wd);
  CHKB200(b200_vec_pointwise_divide(w->map->n,wd,xd,yd));
  ierr = PetscLogFlops(w->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecReciprocal_SeqB200"
static PetscErrorCode VecReciprocal_SeqB200(Vec x)                               /* VecReciprocal_Default vecimpl.h:611 */
{
  PetscScalar    *xd;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  DEVRW(x,xd);
  CHKB200(b200_vec_reciprocal(x->map->n,xd));
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecMAXPY_SeqB200"
static PetscErrorCode VecMAXPY_SeqB200(Vec y,PetscInt nv,const PetscScalar *alpha,Vec *x)   /* VecMAXPY_Seq dvec2.c:836 */
{
  const PetscScalar *stackp[64],**xp = stackp;
  PetscScalar       *yd;
  PetscInt          k;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  if (nv > 64) { ierr = PetscMalloc(nv*sizeof(PetscScalar*),&xp);CHKERRQ(ierr); }
  for (k=0; k<nv; k++) DEVR(x[k],xp[k]);
  DEVRW(y,yd);
  CHKB200(b200_vec_maxpy(y->map->n,yd,nv,alpha,(const double *const*)xp));
  if (nv > 64) { ierr = PetscFree(xp);CHKERRQ(ierr); }
  ierr = PetscLogFlops(nv*2.0*y->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* ---- reductions: results land in host scalars and are valid on return (callers use them at once: cg.c:207-220) - */
#undef __FUNCT__
#define __FUNCT__ "VecDot_SeqB200"
static PetscErrorCode VecDot_SeqB200(Vec x,Vec y,PetscScalar *z)                 /* VecDot_Seq bvec1.c:57 */
{
  const PetscScalar *xd,*yd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd); DEVR(y,yd);
  CHKB200(b200_vec_dot_local(x->map->n,xd,yd,z));
  if (x->map->n > 0) { ierr = PetscLogFlops(2.0*x->map->n-1);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecNorm_SeqB200"
static PetscErrorCode VecNorm_SeqB200(Vec x,NormType type,PetscReal *z)          /* VecNorm_Seq bvec2.c:605 */
{
  const PetscScalar *xd;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(x,xd);
  CHKB200(b200_vec_norm_local(x->map->n,xd,(int)type,z));
  if (type == NORM_2 || type == NORM_FROBENIUS) { ierr = PetscLogFlops(PetscMax(2.0*x->map->n-1,0.0));CHKERRQ(ierr); }
  else if (type == NORM_1)                      { ierr = PetscLogFlops(PetscMax(x->map->n-1.0,0.0));CHKERRQ(ierr); }
  else if (type == NORM_1_AND_2)                { ierr = PetscLogFlops(PetscMax(3.0*x->map->n-2,0.0));CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecMDot_SeqB200"
static PetscErrorCode VecMDot_SeqB200(Vec x,PetscInt nv,const Vec y[],PetscScalar *z)   /* VecMDot_Seq dvec2.c:146 */
{
  const PetscScalar *stackp[64],**yp = stackp,*xd;
  PetscInt          k;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  if (nv > 64) { ierr = PetscMalloc(nv*sizeof(PetscScalar*),&yp);CHKERRQ(ierr); }
  DEVR(x,xd);
  for (k=0; k<nv; k++) DEVR(y[k],yp[k]);
  CHKB200(b200_vec_mdot_local(x->map->n,xd,nv,(const double *const*)yp,z));
  if (nv > 64) { ierr = PetscFree(yp);CHKERRQ(ierr); }
  ierr = PetscLogFlops(PetscMax(nv*(2.0*x->map->n-1),0.0));CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecDotNorm2_SeqB200"
static PetscErrorCode VecDotNorm2_SeqB200(Vec s,Vec t,PetscScalar *dp,PetscScalar *nm)  /* VecDotNorm2_SeqCUSP veccusp.cu:1816 */
{
  const PetscScalar *sd,*td;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  DEVR(s,sd); DEVR(t,td);
  CHKB200(b200_vec_dotnorm2(s->map->n,sd,td,dp,nm));
  ierr = PetscLogFlops(4.0*s->map->n);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* MPI forms: local device reduction, then the communicator's Allreduce (VecDot_MPI pbvec.c:9, VecNorm_MPI pvec2.c:49,
 * VecMDot_MPI pvec2.c:10) */
#undef __FUNCT__
#define __FUNCT__ "VecDot_MPIB200"
static PetscErrorCode VecDot_MPIB200(Vec x,Vec y,PetscScalar *z)
{
  PetscScalar    work;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecDot_SeqB200(x,y,&work);CHKERRQ(ierr);
  ierr = MPI_Allreduce(&work,z,1,MPIU_SCALAR,MPIU_SUM,((PetscObject)x)->comm);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecMDot_MPIB200"
static PetscErrorCode VecMDot_MPIB200(Vec x,PetscInt nv,const Vec y[],PetscScalar *z)
{
  PetscScalar    awork[128],*work = awork;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (nv > 128) { ierr = PetscMalloc(nv*sizeof(PetscScalar),&work);CHKERRQ(ierr); }
  ierr = VecMDot_SeqB200(x,nv,y,work);CHKERRQ(ierr);
  ierr = MPI_Allreduce(work,z,nv,MPIU_SCALAR,MPIU_SUM,((PetscObject)x)->comm);CHKERRQ(ierr);
  if (nv > 128) { ierr = PetscFree(work);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecNorm_MPIB200"
static PetscErrorCode VecNorm_MPIB200(Vec x,NormType type,PetscReal *z)
{
  PetscReal      work[2],sum[2];
  MPI_Comm       comm = ((PetscObject)x)->comm;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecNorm_SeqB200(x,type,work);CHKERRQ(ierr);
  if (type == NORM_2 || type == NORM_FROBENIUS) {
    work[0] *= work[0];                                                           /* pvec2.c:59-62 */
    ierr = MPI_Allreduce(work,sum,1,MPIU_REAL,MPIU_SUM,comm);CHKERRQ(ierr);
    *z = PetscSqrtReal(sum[0]);
  } else if (type == NORM_1) {
    ierr = MPI_Allreduce(work,z,1,MPIU_REAL,MPIU_SUM,comm);CHKERRQ(ierr);
  } else if (type == NORM_INFINITY) {
    ierr = MPI_Allreduce(work,z,1,MPIU_REAL,MPIU_MAX,comm);CHKERRQ(ierr);
  } else if (type == NORM_1_AND_2) {
    work[1] *= work[1];
    ierr = MPI_Allreduce(work,sum,2,MPIU_REAL,MPIU_SUM,comm);CHKERRQ(ierr);
    z[0] = sum[0]; z[1] = PetscSqrtReal(sum[1]);
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecDotNorm2_MPIB200"
static PetscErrorCode VecDotNorm2_MPIB200(Vec s,Vec t,PetscScalar *dp,PetscScalar *nm)   /* mpicusp.cu:168 */
{
  PetscScalar    work[2],sum[2];
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecDotNorm2_SeqB200(s,t,work,work+1);CHKERRQ(ierr);
  ierr = MPI_Allreduce(work,sum,2,MPIU_SCALAR,MPIU_SUM,((PetscObject)s)->comm);CHKERRQ(ierr);
  *dp = sum[0]; *nm = sum[1];
  PetscFunctionReturn(0);
}

/* ---- inherited host ops that touch ->array directly: make the host copy current around them -------------------- */
#define HOSTOP_RW(NAME,BASE,PROTO,ARGS) \
static PetscErrorCode NAME PROTO { PetscErrorCode ierr; PetscFunctionBegin; ierr = VecB200HostReadWrite(x);CHKERRQ(ierr); ierr = BASE ARGS;CHKERRQ(ierr); PetscFunctionReturn(0); }
/* read-only host ops: the base may itself go through VecGetArrayRead -> ops->getarray (which marks the host newer);
 * since we know it does not write, the state from before the call is put back */
#define HOSTOP_RO(NAME,BASE,PROTO,ARGS) \
static PetscErrorCode NAME PROTO { PetscErrorCode ierr; B200VecData *dev; B200VecWhere w; PetscFunctionBegin; ierr = VecB200HostRead(x);CHKERRQ(ierr); \
  ierr = VecB200Data(x,&dev);CHKERRQ(ierr); w = dev->where; ierr = BASE ARGS;CHKERRQ(ierr); dev->where = w; dev->hostpending = PETSC_FALSE; PetscFunctionReturn(0); }

#undef __FUNCT__
#define __FUNCT__ "VecHostOp_B200"
HOSTOP_RW(VecSetValues_SeqB200,VecSetValues_Seq,(Vec x,PetscInt ni,const PetscInt ix[],const PetscScalar y[],InsertMode m),(x,ni,ix,y,m))
HOSTOP_RW(VecSetValuesBlocked_SeqB200,VecSetValuesBlocked_Seq,(Vec x,PetscInt ni,const PetscInt ix[],const PetscScalar y[],InsertMode m),(x,ni,ix,y,m))
HOSTOP_RW(VecSetRandom_SeqB200,VecSetRandom_Seq,(Vec x,PetscRandom r),(x,r))      /* host generator: same stream as the reference */
HOSTOP_RW(VecPlaceArray_SeqB200,VecPlaceArray_Seq,(Vec x,const PetscScalar *a),(x,a))       /* veccusp.cu:1717 */
HOSTOP_RW(VecReplaceArray_SeqB200,VecReplaceArray_Seq,(Vec x,const PetscScalar *a),(x,a))   /* veccusp.cu:1751 */
HOSTOP_RW(VecResetArray_SeqB200,VecResetArray_Seq,(Vec x),(x))                              /* veccusp.cu:1734 */
HOSTOP_RO(VecGetValues_SeqB200,VecGetValues_Seq,(Vec x,PetscInt ni,const PetscInt ix[],PetscScalar y[]),(x,ni,ix,y))
HOSTOP_RO(VecView_SeqB200,VecView_Seq,(Vec x,PetscViewer v),(x,v))
HOSTOP_RO(VecMax_SeqB200,VecMax_Seq,(Vec x,PetscInt *i,PetscReal *r),(x,i,r))
HOSTOP_RO(VecMin_SeqB200,VecMin_Seq,(Vec x,PetscInt *i,PetscReal *r),(x,i,r))
/* MPI */
HOSTOP_RW(VecSetValues_MPIB200,VecSetValues_MPI,(Vec x,PetscInt ni,const PetscInt ix[],const PetscScalar y[],InsertMode m),(x,ni,ix,y,m))
HOSTOP_RW(VecSetValuesBlocked_MPIB200,VecSetValuesBlocked_MPI,(Vec x,PetscInt ni,const PetscInt ix[],const PetscScalar y[],InsertMode m),(x,ni,ix,y,m))
HOSTOP_RW(VecAssemblyBegin_MPIB200,VecAssemblyBegin_MPI,(Vec x),(x))
HOSTOP_RW(VecAssemblyEnd_MPIB200,VecAssemblyEnd_MPI,(Vec x),(x))
HOSTOP_RW(VecPlaceArray_MPIB200,VecPlaceArray_MPI,(Vec x,const PetscScalar *a),(x,a))
HOSTOP_RO(VecGetValues_MPIB200,VecGetValues_MPI,(Vec x,PetscInt ni,const PetscInt ix[],PetscScalar y[]),(x,ni,ix,y))
HOSTOP_RO(VecView_MPIB200,VecView_MPI,(Vec x,PetscViewer v),(x,v))
HOSTOP_RO(VecMax_MPIB200,VecMax_MPI,(Vec x,PetscInt *i,PetscReal *r),(x,i,r))
HOSTOP_RO(VecMin_MPIB200,VecMin_MPI,(Vec x,PetscInt *i,PetscReal *r),(x,i,r))

/* two- and three-vector host ops of the sequential type: all operands current on the host, the result host-newer */
#undef __FUNCT__
#define __FUNCT__ "VecPointwiseHost_B200"
static PetscErrorCode VecPointwiseHost_B200(PetscErrorCode (*base)(Vec,Vec,Vec),Vec w,Vec x,Vec y)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200HostRead(x);CHKERRQ(ierr);
  ierr = VecB200HostRead(y);CHKERRQ(ierr);
  ierr = VecB200HostReadWrite(w);CHKERRQ(ierr);
  ierr = (*base)(w,x,y);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
static PetscErrorCode VecPointwiseMax_SeqB200(Vec w,Vec x,Vec y)    { return VecPointwiseHost_B200(VecPointwiseMax_Seq,w,x,y); }
static PetscErrorCode VecPointwiseMin_SeqB200(Vec w,Vec x,Vec y)    { return VecPointwiseHost_B200(VecPointwiseMin_Seq,w,x,y); }
static PetscErrorCode VecPointwiseMaxAbs_SeqB200(Vec w,Vec x,Vec y) { return VecPointwiseHost_B200(VecPointwiseMaxAbs_Seq,w,x,y); }

#undef __FUNCT__
#define __FUNCT__ "VecMaxPointwiseDivide_SeqB200"
static PetscErrorCode VecMaxPointwiseDivide_SeqB200(Vec x,Vec y,PetscReal *max)
{
  PetscErrorCode ierr;
  B200VecData    *dx,*dy;
  B200VecWhere   wx,wy;

  PetscFunctionBegin;
  ierr = VecB200HostRead(x);CHKERRQ(ierr);
  ierr = VecB200HostRead(y);CHKERRQ(ierr);
  ierr = VecB200Data(x,&dx);CHKERRQ(ierr); wx = dx->where;
  ierr = VecB200Data(y,&dy);CHKERRQ(ierr); wy = dy->where;
  ierr = VecMaxPointwiseDivide_Seq(x,y,max);CHKERRQ(ierr);
  dx->where = wx; dy->where = wy; dx->hostpending = dy->hostpending = PETSC_FALSE;
  PetscFunctionReturn(0);
}

/* ---- life cycle -------------------------------------------------------------------------------------------------- */
#undef __FUNCT__
#define __FUNCT__ "VecB200FreeData"
static PetscErrorCode VecB200FreeData(Vec v,B200VecData *dev)
{
  Vec_Seq *s = (Vec_Seq*)v->data;

  PetscFunctionBegin;
  if (dev->d) { CHKB200(b200_free(dev->d)); dev->d = 0; }
  if (dev->pinned) {
    if (s->array == dev->pinned) s->array = 0;
    if (s->unplacedarray == dev->pinned) s->unplacedarray = 0;
    CHKB200(b200_host_free(dev->pinned)); dev->pinned = 0;
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecDestroy_SeqB200"
static PetscErrorCode VecDestroy_SeqB200(Vec v)                                   /* VecDestroy_SeqCUSP veccusp.cu:1877 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200FreeData(v,&((Vec_SeqB200*)v->data)->dev);CHKERRQ(ierr);
  ierr = VecDestroy_Seq(v);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "VecDestroy_MPIB200"
static PetscErrorCode VecDestroy_MPIB200(Vec v)                                   /* VecDestroy_MPICUSP mpicusp.cu:14 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecB200FreeData(v,&((Vec_MPIB200*)v->data)->dev);CHKERRQ(ierr);
  ierr = VecDestroy_MPI(v);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* the ops both types share */
#undef __FUNCT__
#define __FUNCT__ "VecSetOps_B200"
static PetscErrorCode VecSetOps_B200(Vec V)
{
  PetscFunctionBegin;
  V->petscnative          = PETSC_FALSE;            /* VecGetArray goes through ops->getarray (vecimpl.h:382) */
  V->ops->getarray        = VecGetArray_B200;
  V->ops->restorearray    = VecRestoreArray_B200;
  V->ops->scale           = VecScale_SeqB200;
  V->ops->copy            = VecCopy_SeqB200;
  V->ops->set             = VecSet_SeqB200;
  V->ops->swap            = VecSwap_SeqB200;
  V->ops->axpy            = VecAXPY_SeqB200;
  V->ops->axpby           = VecAXPBY_SeqB200;
  V->ops->axpbypcz        = VecAXPBYPCZ_SeqB200;
  V->ops->pointwisemult   = VecPointwiseMult_SeqB200;
  V->ops->pointwisedivide = VecPointwiseDivide_SeqB200;
  V->ops->reciprocal      = VecReciprocal_SeqB200;
  V->ops->maxpy           = VecMAXPY_SeqB200;
  V->ops->aypx            = VecAYPX_SeqB200;
  V->ops->waxpy           = VecWAXPY_SeqB200;
  V->ops->dot_local       = VecDot_SeqB200;
  V->ops->tdot_local      = VecDot_SeqB200;         /* real scalars: x'y == x^H y (bvec1.c:122) */
  V->ops->norm_local      = VecNorm_SeqB200;
  V->ops->mdot_local      = VecMDot_SeqB200;
  V->ops->mtdot_local     = VecMDot_SeqB200;
  V->ops->setrandom       = VecSetRandom_SeqB200;
  V->ops->pointwisemax    = VecPointwiseMax_SeqB200;
  V->ops->pointwisemin    = VecPointwiseMin_SeqB200;
  V->ops->pointwisemaxabs = VecPointwiseMaxAbs_SeqB200;
  V->ops->maxpointwisedivide = VecMaxPointwiseDivide_SeqB200;
  V->ops->replacearray    = VecReplaceArray_SeqB200;
  V->ops->resetarray      = VecResetArray_SeqB200;
  PetscFunctionReturn(0);
}

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "VecCreate_SeqB200"
PetscErrorCode VecCreate_SeqB200(Vec V)
{
  Vec_SeqB200    *s;
  PetscMPIInt    size;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MPI_Comm_size(((PetscObject)V)->comm,&size);CHKERRQ(ierr);
  if (size > 1) SETERRQ(PETSC_COMM_SELF,PETSC_ERR_ARG_WRONG,"Cannot create VECSEQB200 on more than one process");
  ierr = VecCreate_Seq_Private(V,0);CHKERRQ(ierr);
  /* grow the type data by the device part (Vec_Seq is only the VECHEADER) */
  ierr = PetscNewLog(V,Vec_SeqB200,&s);CHKERRQ(ierr);
  ierr = PetscMemcpy(&s->base,V->data,sizeof(Vec_Seq));CHKERRQ(ierr);
  ierr = PetscFree(V->data);CHKERRQ(ierr);
  V->data = (void*)s;
  ierr = PetscObjectChangeTypeName((PetscObject)V,VECSEQB200);CHKERRQ(ierr);
  ierr = VecSetOps_B200(V);CHKERRQ(ierr);
  V->ops->dot              = VecDot_SeqB200;
  V->ops->tdot             = VecDot_SeqB200;
  V->ops->norm             = VecNorm_SeqB200;
  V->ops->mdot             = VecMDot_SeqB200;
  V->ops->mtdot            = VecMDot_SeqB200;
  V->ops->dotnorm2         = VecDotNorm2_SeqB200;
  V->ops->setvalues        = VecSetValues_SeqB200;
  V->ops->setvaluesblocked = VecSetValuesBlocked_SeqB200;
  V->ops->getvalues        = VecGetValues_SeqB200;
  V->ops->view             = VecView_SeqB200;
  V->ops->max              = VecMax_SeqB200;
  V->ops->min              = VecMin_SeqB200;
  V->ops->placearray       = VecPlaceArray_SeqB200;
  V->ops->destroy          = VecDestroy_SeqB200;
  /* ops->duplicate stays VecDuplicate_Seq: it re-creates by type name (bvec2.c:1174), i.e. through this function */
  ierr = VecSet(V,0.0);CHKERRQ(ierr);                                             /* veccusp.cu:1944 */
  PetscFunctionReturn(0);
}
EXTERN_C_END

#undef __FUNCT__
#define __FUNCT__ "VecDuplicate_MPIB200"
static PetscErrorCode VecDuplicate_MPIB200(Vec win,Vec *v)                        /* VecDuplicate_MPICUSP mpicusp.cu:126 */
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (((Vec_MPI*)win->data)->localrep) SETERRQ(((PetscObject)win)->comm,PETSC_ERR_SUP,"Ghosted VECMPIB200 vectors are not supported");
  ierr = VecCreate(((PetscObject)win)->comm,v);CHKERRQ(ierr);
  ierr = PetscLayoutReference(win->map,&(*v)->map);CHKERRQ(ierr);
  ierr = VecCreate_MPIB200(*v);CHKERRQ(ierr);
  (*v)->stash.donotstash   = win->stash.donotstash;
  (*v)->stash.ignorenegidx = win->stash.ignorenegidx;
  ierr = PetscObjectListDuplicate(((PetscObject)win)->olist,&((PetscObject)(*v))->olist);CHKERRQ(ierr);
  ierr = PetscFunctionListDuplicate(((PetscObject)win)->qlist,&((PetscObject)(*v))->qlist);CHKERRQ(ierr);
  (*v)->map->bs   = win->map->bs;
  (*v)->bstash.bs = win->bstash.bs;
  PetscFunctionReturn(0);
}

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "VecCreate_MPIB200"
PetscErrorCode VecCreate_MPIB200(Vec V)
{
  Vec_MPIB200    *s;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecCreate_MPI_Private(V,PETSC_FALSE,0,0);CHKERRQ(ierr);
  ierr = PetscNewLog(V,Vec_MPIB200,&s);CHKERRQ(ierr);
  ierr = PetscMemcpy(&s->base,V->data,sizeof(Vec_MPI));CHKERRQ(ierr);
  ierr = PetscFree(V->data);CHKERRQ(ierr);
  V->data = (void*)s;
  ierr = PetscObjectChangeTypeName((PetscObject)V,VECMPIB200);CHKERRQ(ierr);
  ierr = VecSetOps_B200(V);CHKERRQ(ierr);
  V->ops->dot              = VecDot_MPIB200;
  V->ops->tdot             = VecDot_MPIB200;
  V->ops->norm             = VecNorm_MPIB200;
  V->ops->mdot             = VecMDot_MPIB200;
  V->ops->mtdot            = VecMDot_MPIB200;
  V->ops->dotnorm2         = VecDotNorm2_MPIB200;
  V->ops->setvalues        = VecSetValues_MPIB200;
  V->ops->setvaluesblocked = VecSetValuesBlocked_MPIB200;
  V->ops->assemblybegin    = VecAssemblyBegin_MPIB200;
  V->ops->assemblyend      = VecAssemblyEnd_MPIB200;
  V->ops->getvalues        = VecGetValues_MPIB200;
  V->ops->view             = VecView_MPIB200;
  V->ops->max              = VecMax_MPIB200;
  V->ops->min              = VecMin_MPIB200;
  V->ops->placearray       = VecPlaceArray_MPIB200;
  V->ops->duplicate        = VecDuplicate_MPIB200;
  V->ops->destroy          = VecDestroy_MPIB200;
  ierr = VecSet(V,0.0);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
EXTERN_C_END

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "VecCreate_B200"
PetscErrorCode VecCreate_B200(Vec v)                                              /* VecCreate_CUSP mpicusp.cu:232 */
{
  PetscMPIInt    size;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MPI_Comm_size(((PetscObject)v)->comm,&size);CHKERRQ(ierr);
  if (size == 1) { ierr = VecSetType(v,VECSEQB200);CHKERRQ(ierr); }
  else           { ierr = VecSetType(v,VECMPIB200);CHKERRQ(ierr); }
  PetscFunctionReturn(0);
}
EXTERN_C_END
