static char help[] = "Columns rewritten in place under unchanged row offsets: MatZeroRows with a diagonal value on rows whose\n\
single entry was off the diagonal (MatZeroRows_SeqAIJ, aij.c:1680-1683) keeps i and nz and changes j.  A device copy that\n\
only compared the row offsets would refresh the values and keep multiplying with the old columns.\n\
Prints y = A x before and after; run with the reference types and with -mat_type aijb200 -vec_type b200.\n";

#include <petscmat.h>

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc,char **argv)
{
  Mat            A;
  Vec            x,y;
  PetscInt       n = 12,i,rstart,rend,col,rows[2],nrows = 0;
  PetscScalar    v;
  PetscErrorCode ierr;

  PetscInitialize(&argc,&argv,(char*)0,help);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-n",&n,PETSC_NULL);CHKERRQ(ierr);
  ierr = MatCreate(PETSC_COMM_WORLD,&A);CHKERRQ(ierr);
  ierr = MatSetSizes(A,PETSC_DECIDE,PETSC_DECIDE,n,n);CHKERRQ(ierr);
  ierr = MatSetFromOptions(A);CHKERRQ(ierr);
  ierr = MatSeqAIJSetPreallocation(A,1,PETSC_NULL);CHKERRQ(ierr);
  ierr = MatMPIAIJSetPreallocation(A,1,PETSC_NULL,1,PETSC_NULL);CHKERRQ(ierr);
  ierr = MatGetOwnershipRange(A,&rstart,&rend);CHKERRQ(ierr);
  /* one entry per row, just right of the diagonal inside the owned block (so that it sits in the diagonal block of the
     parallel type too); the last owned row keeps a diagonal entry */
  for (i=rstart; i<rend; i++) {
    col = (i+1 < rend) ? i+1 : i; v = 2.0 + (PetscScalar)i;
    ierr = MatSetValues(A,1,&i,1,&col,&v,INSERT_VALUES);CHKERRQ(ierr);
  }
  ierr = MatAssemblyBegin(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatAssemblyEnd(A,MAT_FINAL_ASSEMBLY);CHKERRQ(ierr);
  ierr = MatGetVecs(A,&x,&y);CHKERRQ(ierr);
  for (i=rstart; i<rend; i++) { v = 1.0 + 0.5*(PetscScalar)i; ierr = VecSetValues(x,1,&i,&v,INSERT_VALUES);CHKERRQ(ierr); }
  ierr = VecAssemblyBegin(x);CHKERRQ(ierr);
  ierr = VecAssemblyEnd(x);CHKERRQ(ierr);

  ierr = MatMult(A,x,y);CHKERRQ(ierr);
  ierr = PetscPrintf(PETSC_COMM_WORLD,"before MatZeroRows\n");CHKERRQ(ierr);
  ierr = VecView(y,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);

  if (rend-rstart > 3) { rows[0] = rstart; rows[1] = rstart+2; nrows = 2; }
  ierr = MatZeroRows(A,nrows,rows,7.0,PETSC_NULL,PETSC_NULL);CHKERRQ(ierr);
  ierr = MatMult(A,x,y);CHKERRQ(ierr);
  ierr = PetscPrintf(PETSC_COMM_WORLD,"after MatZeroRows\n");CHKERRQ(ierr);
  ierr = VecView(y,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);

  ierr = VecDestroy(&x);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);
  ierr = MatDestroy(&A);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
