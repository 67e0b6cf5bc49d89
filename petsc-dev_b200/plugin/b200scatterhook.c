/* b200scatterhook.c — stand-in for the one line a PETSc maintainer would add at the end of VecScatterCreate
 * (src/vec/vec/utils/vscat.c:920-1560): `VecScatterB200Attach(*newctx)`.  For an UNMODIFIED PETSc shared library the
 * same effect comes from symbol interposition: start the program with
 *     LD_PRELOAD=/path/to/libpetscb200scatterhook.so  prog -dll_prepend /path/to/libpetscb200plugin.so -vec_type b200 ...
 * and every VecScatterCreate — the program's own and the library's internal ones (VecCreateGhost, DMDA global-to-local,
 * MatSetUpMultiply_MPIAIJ) — returns a scatter whose Begin/End run on the device when both vectors are B200 vectors.
 * Host vectors take the reference's code path untouched (b200scatter.c).  Holds no logic of its own. */
#define _GNU_SOURCE
#include <dlfcn.h>
#include "petscb200plugin.h"

#undef __FUNCT__
#define __FUNCT__ "VecScatterCreate"
PetscErrorCode VecScatterCreate(Vec xin,IS ix,Vec yin,IS iy,VecScatter *newctx)
{
  static PetscErrorCode (*real)(Vec,IS,Vec,IS,VecScatter*) = 0;
  PetscErrorCode        ierr;

  if (!real) {
    real = (PetscErrorCode (*)(Vec,IS,Vec,IS,VecScatter*))dlsym(RTLD_NEXT,"VecScatterCreate");
    if (!real) return PETSC_ERR_LIB;
  }
  ierr = (*real)(xin,ix,yin,iy,newctx);
  if (ierr) return ierr;
  return VecScatterB200Attach(*newctx);
}
