static char help[] = "MatSetValuesBatch on a parallel AIJ matrix: the two-triangles-per-cell mesh of the reference's\n\
ksp/examples/tutorials/ex4.c on an X x Y vertex grid, every rank bringing a contiguous chunk of the element list\n\
(so some of its elements touch rows another rank owns).  Test driver of this repository: public PETSc API only, the\n\
matrix type comes from -mat_type (mpiaij = reference, mpiaijb200 = plugin).\n";
#include <petscmat.h>

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc,char **argv)
{
  Mat               A;
  PetscInt          X = 3,Y = 3,nl = 3,ne,e0,e1,e,i,j,k,*er;
  PetscScalar       *em;
  const PetscScalar t[9] = {1.0,-0.5,-0.5, -0.5,0.5,0.0, -0.5,0.0,0.5};
  PetscMPIInt       rank,size;
  PetscErrorCode    ierr;

  ierr = PetscInitialize(&argc,&argv,0,help);CHKERRQ(ierr);
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD,&rank);CHKERRQ(ierr);
  ierr = MPI_Comm_size(PETSC_COMM_WORLD,&size);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-X",&X,PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-Y",&Y,PETSC_NULL);CHKERRQ(ierr);
  ne = 2*(X-1)*(Y-1);
  e0 = ne*rank/size; e1 = ne*(rank+1)/size;
  ierr = PetscMalloc2((e1-e0)*nl+1,PetscInt,&er,(e1-e0)*nl*nl+1,PetscScalar,&em);CHKERRQ(ierr);
  for (e=e0; e<e1; e++) {
    PetscInt cell = e/2,a,b,c,d;
    j = cell/(X-1); i = cell%(X-1);
    a = j*X+i; b = a+1; c = a+X; d = c+1;
    k = (e-e0)*nl;
    if (e%2 == 0) { er[k] = a; er[k+1] = b; er[k+2] = c; }                        /* lower triangle A B C */
    else          { er[k] = d; er[k+1] = c; er[k+2] = b; }                        /* upper triangle D C B */
    for (i=0; i<9; i++) em[(e-e0)*9+i] = t[i];
  }
  ierr = MatCreate(PETSC_COMM_WORLD,&A);CHKERRQ(ierr);
  ierr = MatSetSizes(A,PETSC_DECIDE,PETSC_DECIDE,X*Y,X*Y);CHKERRQ(ierr);
  ierr = MatSetType(A,MATMPIAIJ);CHKERRQ(ierr);
  ierr = MatSetFromOptions(A);CHKERRQ(ierr);
  ierr = MatSetUp(A);CHKERRQ(ierr);
  ierr = MatSetOption(A,MAT_NEW_NONZERO_ALLOCATION_ERR,PETSC_FALSE);CHKERRQ(ierr);
  ierr = MatSetValuesBatch(A,e1-e0,nl,er,em);CHKERRQ(ierr);
  ierr = MatAssemblyBegin(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatAssemblyEnd(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatView(A,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);
  ierr = MatDestroy(&A);CHKERRQ(ierr);
  ierr = PetscFree2(er,em);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
