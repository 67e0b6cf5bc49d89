static char help[] = "Read access to a device vector must not make the host copy the newer one: VecGetArrayRead/VecRestoreArrayRead\n\
between device operations costs one download and no upload; VecGetArray/VecRestoreArray (a write) costs the upload.\n\
Run with -vec_type b200 -info and read the plugin's totals line; the printed numbers are the same for every vector type.\n";

#include <petscvec.h>

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc,char **argv)
{
  Vec               x;
  PetscInt          n = 10,i;
  const PetscScalar *r;
  PetscScalar       *w,sum = 0.0;
  PetscReal         nrm;
  PetscErrorCode    ierr;

  PetscInitialize(&argc,&argv,(char*)0,help);
  ierr = VecCreate(PETSC_COMM_SELF,&x);CHKERRQ(ierr);
  ierr = VecSetSizes(x,n,n);CHKERRQ(ierr);
  ierr = VecSetFromOptions(x);CHKERRQ(ierr);
  ierr = VecSet(x,1.0);CHKERRQ(ierr);                                   /* device */
  ierr = VecGetArrayRead(x,&r);CHKERRQ(ierr);                           /* download 1 */
  for (i=0; i<n; i++) sum += r[i];
  ierr = VecRestoreArrayRead(x,&r);CHKERRQ(ierr);
  ierr = VecScale(x,2.0);CHKERRQ(ierr);                                 /* device copy still current: no upload */
  ierr = VecGetArrayRead(x,&r);CHKERRQ(ierr);                           /* download 2 */
  for (i=0; i<n; i++) sum += r[i];
  ierr = VecRestoreArrayRead(x,&r);CHKERRQ(ierr);
  ierr = VecNorm(x,NORM_1,&nrm);CHKERRQ(ierr);                          /* device, no upload */
  ierr = PetscPrintf(PETSC_COMM_SELF,"sum %g norm1 %g\n",(double)PetscRealPart(sum),(double)nrm);CHKERRQ(ierr);
  ierr = VecGetArray(x,&w);CHKERRQ(ierr);                               /* both copies current: no download */
  w[0] = 5.0;
  ierr = VecRestoreArray(x,&w);CHKERRQ(ierr);                           /* state increases: the host copy is the newer one */
  ierr = VecNorm(x,NORM_1,&nrm);CHKERRQ(ierr);                          /* upload 1 */
  ierr = PetscPrintf(PETSC_COMM_SELF,"after the write: norm1 %g\n",(double)nrm);CHKERRQ(ierr);
  ierr = VecDestroy(&x);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
