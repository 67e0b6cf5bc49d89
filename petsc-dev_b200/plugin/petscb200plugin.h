/* petscb200plugin.h — the PETSc-side binding of libpetscb200: new Vec and Mat types registered through
 * VecRegister / MatRegister / MatRegisterBaseName, the way VECSEQCUSP / MATSEQAIJCUSP plug in
 * (reference src/vec/vec/interface/vecregall.c:46-49, src/mat/interface/matregis.c:134-137).
 *
 *   -vec_type seqb200 | mpib200 | b200            -mat_type seqaijb200 | mpiaijb200 | aijb200
 *
 * Host side in C against the reference's own headers; every numerical op goes through the C ABI of
 * include/petscb200.h.  No CPU fallback: the ops fail with PETSC_ERR_LIB when the CUDA library reports an error.
 *
 * Loading: either call PetscB200Register() after PetscInitialize(), or start ANY unmodified PETSc program with
 *   -dll_prepend /path/to/libpetscb200plugin.so
 * and PetscDLLibraryRegister_petscb200plugin() (below) is run by PetscDLLibraryPrepend (src/sys/dll/dl.c:185-190).
 *
 * Coherence seam (option 2): the types set petscnative = PETSC_FALSE and provide
 * ops->getarray / ops->restorearray, so the inline VecGetArray/VecRestoreArray of a PETSc built WITHOUT
 * PETSC_HAVE_CUSP (include/petsc-private/vecimpl.h:382,434) pull the data back to the host and mark it host-newer. */
#ifndef PETSCB200PLUGIN_H
#define PETSCB200PLUGIN_H

#include <petscmat.h>
#include <petscvec.h>
#include <petscksp.h>

#define VECSEQB200    "seqb200"
#define VECMPIB200    "mpib200"
#define VECB200       "b200"
#define MATSEQAIJB200 "seqaijb200"
#define MATMPIAIJB200 "mpiaijb200"
#define MATAIJB200    "aijb200"
#define KSPCGB200     "cgb200"
#define KSPGMRESB200  "gmresb200"

EXTERN_C_BEGIN
extern PetscErrorCode PetscB200Register(void);
extern PetscErrorCode PetscDLLibraryRegister_petscb200plugin(const char path[]);
extern PetscErrorCode VecCreate_SeqB200(Vec);
extern PetscErrorCode VecCreate_MPIB200(Vec);
extern PetscErrorCode VecCreate_B200(Vec);
extern PetscErrorCode MatCreate_SeqAIJB200(Mat);
extern PetscErrorCode MatCreate_MPIAIJB200(Mat);
extern PetscErrorCode KSPCreate_CGB200(KSP);
extern PetscErrorCode KSPCreate_GMRESB200(KSP);
EXTERN_C_END

/* where the newest copy of a vector lives (the role of PetscCUSPFlag, include/petscsys.h) */
typedef enum {B200_VEC_UNALLOCATED = 0, B200_VEC_HOST = 1, B200_VEC_DEVICE = 2, B200_VEC_BOTH = 3} B200VecWhere;

/* device access for other plugins (cf. VecCUSPGetArrayRead/Write, src/vec/vec/impls/seq/seqcusp/cuspvecimpl.h:95-140) */
extern PetscErrorCode VecB200GetDeviceRead(Vec, const PetscScalar **);
extern PetscErrorCode VecB200GetDeviceWrite(Vec, PetscScalar **);      /* contents undefined, caller overwrites all of it */
extern PetscErrorCode VecB200GetDeviceReadWrite(Vec, PetscScalar **);
extern PetscErrorCode VecB200GetWhere(Vec, B200VecWhere *);
extern PetscErrorCode MatB200GetHandle(Mat, void **b200_mat_handle);   /* device operator of an aijb200 matrix, refreshed */
/* PETSc's VecScatter object on the device (b200scatter.c): after this call VecScatterBegin/End of `ctx` run GPU to GPU
 * whenever both vectors are B200 vectors and fall through to the reference's code otherwise.  The seam is the one
 * VecScatterInitializeForGPU uses (ctx->spptr, src/vec/vec/utils/veccusp/vscatcusp.c:29).  A PETSc with the plugin built
 * in would call it at the end of VecScatterCreate (vscat.c:920); for an unmodified PETSc,
 * LD_PRELOAD=libpetscb200scatterhook.so does exactly that (b200scatterhook.c). */
extern PetscErrorCode VecScatterB200Attach(VecScatter ctx);
extern PetscErrorCode PetscB200GetScatterCounters(PetscInt *ondevice, PetscInt *onhost);
/* counters for tests: host<->device copies the coherence protocol performed, device matrix uploads */
extern PetscErrorCode PetscB200GetCounters(PetscInt *h2d, PetscInt *d2h, PetscInt *matuploads, PetscInt *matvalueupdates);

#endif
