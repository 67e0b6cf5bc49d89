static char help[] = "VecScatter layouts between vectors whose type comes from the options database: parallel->sequential with\n\
general, stride and block index sets (forward/reverse, INSERT/ADD, targets named twice), parallel->parallel, sequential->parallel,\n\
the four sequential->sequential layouts, and a VecScatterCopy.  Public PETSc API only: run with the reference's vector types\n\
and with -vec_type b200 (LD_PRELOAD=libpetscb200scatterhook.so attaches the device scatter) and compare the output.\n";

#include <petscvec.h>

#undef __FUNCT__
#define __FUNCT__ "Fill"
static PetscErrorCode Fill(Vec v,PetscScalar scale,PetscScalar shift)
{
  PetscInt       i,lo,hi;
  PetscScalar    val;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecGetOwnershipRange(v,&lo,&hi);CHKERRQ(ierr);
  for (i=lo; i<hi; i++) { val = scale*(PetscScalar)i + shift + 0.125*(PetscScalar)(i%7); ierr = VecSetValues(v,1,&i,&val,INSERT_VALUES);CHKERRQ(ierr); }
  ierr = VecAssemblyBegin(v);CHKERRQ(ierr);
  ierr = VecAssemblyEnd(v);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "NewVec"
static PetscErrorCode NewVec(MPI_Comm comm,PetscInt nlocal,PetscInt N,Vec *v)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = VecCreate(comm,v);CHKERRQ(ierr);
  ierr = VecSetSizes(*v,nlocal,N);CHKERRQ(ierr);
  ierr = VecSetFromOptions(*v);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

/* the sequential vectors differ from rank to rank: print them in rank order */
#undef __FUNCT__
#define __FUNCT__ "ShowSeq"
static PetscErrorCode ShowSeq(const char *title,Vec y)
{
  PetscMPIInt    rank;
  PetscInt       i,n;
  PetscScalar    *a;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD,&rank);CHKERRQ(ierr);
  ierr = VecGetLocalSize(y,&n);CHKERRQ(ierr);
  ierr = PetscPrintf(PETSC_COMM_WORLD,"%s\n",title);CHKERRQ(ierr);
  ierr = VecGetArray(y,&a);CHKERRQ(ierr);
  for (i=0; i<n; i++) { ierr = PetscSynchronizedPrintf(PETSC_COMM_WORLD,"[%d] %D %.15g\n",rank,i,(double)PetscRealPart(a[i]));CHKERRQ(ierr); }
  ierr = VecRestoreArray(y,&a);CHKERRQ(ierr);
  ierr = PetscSynchronizedFlush(PETSC_COMM_WORLD);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "ShowPar"
static PetscErrorCode ShowPar(const char *title,Vec x)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscPrintf(PETSC_COMM_WORLD,"%s\n",title);CHKERRQ(ierr);
  ierr = VecView(x,PETSC_VIEWER_STDOUT_WORLD);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "main"
int main(int argc,char **argv)
{
  PetscMPIInt    rank,size;
  PetscInt       nloc = 6,N,i,m,bs = 2,*idx,*idy;
  Vec            x,x2,y,ys,s1,s2;
  IS             ix,iy;
  VecScatter     ctx,ctx2;
  PetscErrorCode ierr;

  PetscInitialize(&argc,&argv,(char*)0,help);
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD,&rank);CHKERRQ(ierr);
  ierr = MPI_Comm_size(PETSC_COMM_WORLD,&size);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-nloc",&nloc,PETSC_NULL);CHKERRQ(ierr);
  nloc = 2*(nloc/2); if (nloc < 4) nloc = 4;
  N    = nloc*size;
  ierr = NewVec(PETSC_COMM_WORLD,nloc,N,&x);CHKERRQ(ierr);
  ierr = VecDuplicate(x,&x2);CHKERRQ(ierr);
  ierr = PetscMalloc2(2*N+8,PetscInt,&idx,2*N+8,PetscInt,&idy);CHKERRQ(ierr);

  /* 1. parallel -> sequential, general index sets: every rank gathers a window that straddles its neighbours, plus entry 0 */
  m = nloc+3;
  for (i=0; i<nloc+2; i++) { idx[i] = (rank*nloc - 1 + i + N) % N; idy[i] = nloc+1-i; }
  idx[nloc+2] = 0; idy[nloc+2] = nloc+2;
  ierr = NewVec(PETSC_COMM_SELF,m,m,&y);CHKERRQ(ierr);
  ierr = ISCreateGeneral(PETSC_COMM_SELF,m,idx,PETSC_COPY_VALUES,&ix);CHKERRQ(ierr);
  ierr = ISCreateGeneral(PETSC_COMM_SELF,m,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(x,ix,y,iy,&ctx);CHKERRQ(ierr);
  ierr = Fill(x,1.0,0.5);CHKERRQ(ierr);
  ierr = VecSet(y,-1.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("1a general PtoS forward INSERT",y);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x,y,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x,y,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("1b general PtoS forward ADD",y);CHKERRQ(ierr);
  /* reverse ADD: ghost contributions summed into their owners (entries several ranks hold get several terms) */
  ierr = VecSet(x2,100.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,y,x2,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,y,x2,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowPar("1c general PtoS reverse ADD",x2);CHKERRQ(ierr);
  /* a copy of the scatter behaves the same */
  ierr = VecScatterCopy(ctx,&ctx2);CHKERRQ(ierr);
  ierr = VecSet(y,0.0);CHKERRQ(ierr);
  ierr = VecScale(x,2.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx2,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx2,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("1d VecScatterCopy forward INSERT",y);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx2);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);

  /* 2. parallel -> sequential with stride index sets (every second entry of the whole vector onto every rank) */
  m = N/2;
  ierr = NewVec(PETSC_COMM_SELF,m,m,&y);CHKERRQ(ierr);
  ierr = ISCreateStride(PETSC_COMM_SELF,m,rank%2,2,&ix);CHKERRQ(ierr);
  ierr = ISCreateStride(PETSC_COMM_SELF,m,0,1,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(x,ix,y,iy,&ctx);CHKERRQ(ierr);
  ierr = VecSet(y,0.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x,y,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("2a stride PtoS forward INSERT",y);CHKERRQ(ierr);
  ierr = VecSet(x2,0.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,y,x2,INSERT_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,y,x2,INSERT_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowPar("2b stride PtoS reverse INSERT",x2);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);

  /* 3. block index sets (bs = 2), the same target block named twice: ADD_VALUES sums both contributions */
  m = 3;
  idx[0] = (rank*(nloc/bs) + nloc/bs) % (N/bs);      /* first block of the next rank                                  */
  idx[1] = rank*(nloc/bs);                           /* own first block                                              */
  idx[2] = (rank*(nloc/bs) + nloc/bs + 1) % (N/bs);  /* second block of the next rank                                */
  idy[0] = 0; idy[1] = 1; idy[2] = 0;
  ierr = NewVec(PETSC_COMM_SELF,2*bs,2*bs,&y);CHKERRQ(ierr);
  ierr = ISCreateBlock(PETSC_COMM_SELF,bs,m,idx,PETSC_COPY_VALUES,&ix);CHKERRQ(ierr);
  ierr = ISCreateBlock(PETSC_COMM_SELF,bs,m,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(x,ix,y,iy,&ctx);CHKERRQ(ierr);
  ierr = VecSet(y,1.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x,y,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x,y,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("3a block PtoS forward ADD, repeated target",y);CHKERRQ(ierr);
  ierr = VecSet(x2,0.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,y,x2,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,y,x2,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowPar("3b block PtoS reverse ADD",x2);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = VecDestroy(&y);CHKERRQ(ierr);

  /* 4. parallel -> parallel: a cyclic shift by nloc/2 + 1 entries (vec ex22 style) */
  for (i=0; i<nloc; i++) { idx[i] = rank*nloc + i; idy[i] = (rank*nloc + i + nloc/2 + 1) % N; }
  ierr = ISCreateGeneral(PETSC_COMM_SELF,nloc,idx,PETSC_COPY_VALUES,&ix);CHKERRQ(ierr);
  ierr = ISCreateGeneral(PETSC_COMM_SELF,nloc,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(x,ix,x2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecSet(x2,0.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x,x2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x,x2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowPar("4a PtoP shift forward INSERT",x2);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x2,x,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x2,x,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowPar("4b PtoP shift reverse ADD",x);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);

  /* 5. sequential -> parallel (vec ex12 style): every rank drops its sequential entries into the NEXT rank's part */
  m = nloc/2;
  ierr = NewVec(PETSC_COMM_SELF,m,m,&ys);CHKERRQ(ierr);
  ierr = Fill(ys,3.0,(PetscScalar)(10*rank));CHKERRQ(ierr);
  for (i=0; i<m; i++) { idx[i] = i; idy[i] = (((rank+1)%size)*nloc + 2*i) % N; }
  ierr = ISCreateGeneral(PETSC_COMM_SELF,m,idx,PETSC_COPY_VALUES,&ix);CHKERRQ(ierr);
  ierr = ISCreateGeneral(PETSC_COMM_SELF,m,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(ys,ix,x2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecSet(x2,-2.0);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,ys,x2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,ys,x2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowPar("5a StoP forward INSERT",x2);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,x2,ys,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,x2,ys,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowSeq("5b StoP reverse ADD",ys);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = VecDestroy(&ys);CHKERRQ(ierr);

  /* 6. sequential -> sequential: general->general (repeated target, ADD), general->stride, stride->general, stride->stride */
  m = 8;
  ierr = NewVec(PETSC_COMM_SELF,m,m,&s1);CHKERRQ(ierr);
  ierr = NewVec(PETSC_COMM_SELF,2*m,2*m,&s2);CHKERRQ(ierr);
  ierr = Fill(s1,1.0,(PetscScalar)rank);CHKERRQ(ierr);
  for (i=0; i<5; i++) { idx[i] = (3*i+rank) % m; idy[i] = (i == 4) ? 1 : 2*i+1; }      /* target 1 named by entries 0 and 4 */
  ierr = ISCreateGeneral(PETSC_COMM_SELF,5,idx,PETSC_COPY_VALUES,&ix);CHKERRQ(ierr);
  ierr = ISCreateGeneral(PETSC_COMM_SELF,5,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(s1,ix,s2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecSet(s2,0.25);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s1,s2,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s1,s2,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("6a SGtoSG forward ADD, repeated target",s2);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("6b SGtoSG forward INSERT, last writer wins",s2);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s2,s1,INSERT_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s2,s1,INSERT_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowSeq("6c SGtoSG reverse INSERT",s1);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = ISCreateStride(PETSC_COMM_SELF,5,2,3,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(s1,ix,s2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s1,s2,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s1,s2,ADD_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("6d SGtoSS forward ADD",s2);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISCreateStride(PETSC_COMM_SELF,5,1,1,&ix);CHKERRQ(ierr);
  ierr = VecScatterCreate(s1,ix,s2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("6e SStoSS forward INSERT",s2);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s2,s1,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s2,s1,ADD_VALUES,SCATTER_REVERSE);CHKERRQ(ierr);
  ierr = ShowSeq("6f SStoSS reverse ADD",s1);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  for (i=0; i<5; i++) idy[i] = 15-2*i;
  ierr = ISCreateGeneral(PETSC_COMM_SELF,5,idy,PETSC_COPY_VALUES,&iy);CHKERRQ(ierr);
  ierr = VecScatterCreate(s1,ix,s2,iy,&ctx);CHKERRQ(ierr);
  ierr = VecScatterBegin(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = VecScatterEnd(ctx,s1,s2,INSERT_VALUES,SCATTER_FORWARD);CHKERRQ(ierr);
  ierr = ShowSeq("6g SStoSG forward INSERT",s2);CHKERRQ(ierr);
  ierr = VecScatterDestroy(&ctx);CHKERRQ(ierr);
  ierr = ISDestroy(&ix);CHKERRQ(ierr);
  ierr = ISDestroy(&iy);CHKERRQ(ierr);
  ierr = VecDestroy(&s1);CHKERRQ(ierr);
  ierr = VecDestroy(&s2);CHKERRQ(ierr);

  ierr = PetscFree2(idx,idy);CHKERRQ(ierr);
  ierr = VecDestroy(&x);CHKERRQ(ierr);
  ierr = VecDestroy(&x2);CHKERRQ(ierr);
  ierr = PetscFinalize();
  return 0;
}
