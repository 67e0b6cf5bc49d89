static char help[] = "Off-process MatSetValues: every rank contributes entries to rows the NEXT rank owns (and to its own), with\n\
ADD_VALUES, so MatAssemblyBegin/End ship them through the matrix stash (MatSetValues_MPIAIJ mpiaij.c:496-570,\n\
MatStashScatterBegin_Private matstash.c); a second assembly adds to the same pattern from the other side.  Prints y = A x and\n\
the matrix after each assembly.  Public PETSc API only: run with the reference types and with -mat_type aijb200 -vec_type b200.\n";

#include <petscmat.h>

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc,char **argv)
{
  Mat            A;
  Vec            x,y;
  PetscMPIInt    rank,size;
  PetscInt       nloc = 5,N,i,rstart,rend,row,col[3],pass;
  PetscScalar    v[3],one = 1.0;
  PetscErrorCode ierr;

  PetscInitialize(&argc,&argv,(char*)0,help);
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD,&rank);CHKERRQ(ierr);
  ierr = MPI_Comm_size(PETSC_COMM_WORLD,&size);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-nloc",&nloc,PETSC_NULL);CHKERRQ(ierr);
  N    = nloc*size;
  ierr = MatCreate(PETSC_COMM_WORLD,&A);CHKERRQ(ierr);
  ierr = MatSetSizes(A,nloc,nloc,N,N);CHKERRQ(ierr);
  ierr = MatSetFromOptions(A);CHKERRQ(ierr);
  ierr = MatSeqAIJSetPreallocation(A,5,PETSC_NULL);CHKERRQ(ierr);
  ierr = MatMPIAIJSetPreallocation(A,5,PETSC_NULL,5,PETSC_NULL);CHKERRQ(ierr);
  /* the second assembly brings NEW nonzeros on three or more ranks (previous rank != next rank): allow them, so the
     device copy has to notice a changed pattern and ghost list */
  ierr = MatSetOption(A,MAT_NEW_NONZERO_ALLOCATION_ERR,PETSC_FALSE);CHKERRQ(ierr);
  ierr = MatGetOwnershipRange(A,&rstart,&rend);CHKERRQ(ierr);
  ierr = MatGetVecs(A,&x,&y);CHKERRQ(ierr);
  for (i=rstart; i<rend; i++) { v[0] = 1.0 + 0.25*(PetscScalar)i; ierr = VecSetValues(x,1,&i,v,INSERT_VALUES);CHKERRQ(ierr); }
  ierr = VecAssemblyBegin(x);CHKERRQ(ierr);
  ierr = VecAssemblyEnd(x);CHKERRQ(ierr);

  for (pass=0; pass<2; pass++) {
    for (i=rstart; i<rend; i++) {
      /* own row: diagonal */
      v[0] = 4.0 + (PetscScalar)pass; ierr = MatSetValues(A,1,&i,1,&i,v,ADD_VALUES);CHKERRQ(ierr);
      /* a row of the next rank (pass 0) or of the previous one (pass 1): three entries, one of them in MY column range */
      row    = (pass ? i + N - nloc : i + nloc) % N;
      col[0] = i; col[1] = row; col[2] = (row + 1) % N;
      v[0] = -1.0 - 0.5*(PetscScalar)rank; v[1] = 0.5; v[2] = -0.25*(PetscScalar)(pass+1);
      ierr = MatSetValues(A,1,&row,3,col,v,ADD_VALUES);CHKERRQ(ierr);
    }
    ierr = MatAssemblyBegin(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
    ierr = MatAssemblyEnd(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
    ierr = MatMult(A,x,y);CHKERRQ(ierr);
    ierr = PetscPrintf(PETSC_COMM_WORLD,"after assembly %D\n",pass);CHKERRQ(ierr);
    ierr = VecView(y,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);
    ierr = MatMultTranspose(A,x,y);CHKERRQ(ierr);
    ierr = VecView(y,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);
    ierr = MatView(A,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);
    ierr = VecScale(x,one + 0.5);CHKERRQ(ierr);
  }
  ierr = VecDestroy(&x);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);
  ierr = MatDestroy(&A);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
