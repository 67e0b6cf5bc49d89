/* b200reg.c — registration of the B200 Vec/Mat types with a running PETSc.
 *
 * What MatRegisterAll / VecRegisterAll do for the CUSP types inside the library (reference
 * src/mat/interface/matregis.c:134-137, src/vec/vec/interface/vecregall.c:46-49) an out-of-tree plugin does after
 * PetscInitialize with the same public calls (include/petscmat.h:183-227, include/petscvec.h:319-364). */
#include "petscb200plugin.h"
#include "petscb200.h"

static PetscBool PetscB200Registered = PETSC_FALSE;

#undef __FUNCT__
#define __FUNCT__ "PetscB200Finalize"
static PetscErrorCode PetscB200Finalize(void)
{
  PetscInt       h2d,d2h,sdev,shost;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscB200GetCounters(&h2d,&d2h,PETSC_NULL,PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscB200GetScatterCounters(&sdev,&shost);CHKERRQ(ierr);
  ierr = PetscInfo4(0,"B200 totals: %D vector uploads, %D vector downloads, %D scatters on the device, %D through the host\n",h2d,d2h,sdev,shost);CHKERRQ(ierr);
  PetscB200Registered = PETSC_FALSE;
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "PetscB200Register"
PetscErrorCode PetscB200Register(void)
{
  PetscInt       device = -1;
  PetscMPIInt    rank;
  int            ndev = 0;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  if (PetscB200Registered) PetscFunctionReturn(0);
  /* one process per GPU: rank r drives device r mod (devices on the node) unless -b200_device says otherwise
     (the reference never picks a device, src/sys/objects/pinit.c:838-846) */
  ierr = MPI_Comm_rank(PETSC_COMM_WORLD,&rank);CHKERRQ(ierr);
  ierr = PetscOptionsGetInt(PETSC_NULL,"-b200_device",&device,PETSC_NULL);CHKERRQ(ierr);
  if (device < 0) {
    if (b200_device_count(&ndev) || ndev < 1) SETERRQ1(PETSC_COMM_SELF,PETSC_ERR_LIB,"No usable CUDA device for the B200 types: %s",b200_last_error());
    device = rank % ndev;
  }
  if (b200_init((int)device)) SETERRQ2(PETSC_COMM_SELF,PETSC_ERR_LIB,"b200_init(%D) failed: %s",device,b200_last_error());
  /* more than one rank: the library's own rank-to-rank transport (ghost exchange GPU to GPU inside MatMult_MPIAIJB200).
     The NCCL id is made on rank 0 and broadcast over PETSC_COMM_WORLD; -b200_p2p 0 keeps the NCCL transport instead of
     the NVLink peer-memory one.  Not exercised on this box (its only PETSc is the one-rank MPIUNI build). */
  {
    PetscMPIInt size;
    PetscBool   p2p = PETSC_TRUE;
    char        id[B200_COMM_ID_BYTES];
    ierr = MPI_Comm_size(PETSC_COMM_WORLD,&size);CHKERRQ(ierr);
    if (size > 1) {
      ierr = PetscMemzero(id,sizeof(id));CHKERRQ(ierr);
      if (!rank && b200_comm_get_unique_id(id)) SETERRQ1(PETSC_COMM_SELF,PETSC_ERR_LIB,"b200_comm_get_unique_id: %s",b200_last_error());
      ierr = MPI_Bcast(id,B200_COMM_ID_BYTES,MPI_CHAR,0,PETSC_COMM_WORLD);CHKERRQ(ierr);
      if (b200_comm_init((int)rank,(int)size,id)) SETERRQ1(PETSC_COMM_SELF,PETSC_ERR_LIB,"b200_comm_init: %s",b200_last_error());
      ierr = PetscOptionsGetBool(PETSC_NULL,"-b200_p2p",&p2p,PETSC_NULL);CHKERRQ(ierr);
      if (p2p && b200_comm_enable_p2p(1)) {     /* the library agrees on the outcome across ranks: all switched, or none */
        ierr = PetscInfo1(0,"B200 peer-memory transport unavailable (%s); using NCCL send/recv\n",b200_last_error());CHKERRQ(ierr);
      }
    }
  }
  ierr = VecRegister(VECSEQB200,0,"VecCreate_SeqB200",VecCreate_SeqB200);CHKERRQ(ierr);
  ierr = VecRegister(VECMPIB200,0,"VecCreate_MPIB200",VecCreate_MPIB200);CHKERRQ(ierr);
  ierr = VecRegister(VECB200,   0,"VecCreate_B200",   VecCreate_B200);CHKERRQ(ierr);
  ierr = MatRegisterBaseName(MATAIJB200,MATSEQAIJB200,MATMPIAIJB200);CHKERRQ(ierr);
  ierr = MatRegister(MATSEQAIJB200,0,"MatCreate_SeqAIJB200",MatCreate_SeqAIJB200);CHKERRQ(ierr);
  ierr = MatRegister(MATMPIAIJB200,0,"MatCreate_MPIAIJB200",MatCreate_MPIAIJB200);CHKERRQ(ierr);
  /* device-resident Krylov drivers as NEW KSP types (itregis.c:69 registers the reference's the same way) */
  ierr = KSPRegister(KSPCGB200,0,"KSPCreate_CGB200",KSPCreate_CGB200);CHKERRQ(ierr);
  ierr = KSPRegister(KSPGMRESB200,0,"KSPCreate_GMRESB200",KSPCreate_GMRESB200);CHKERRQ(ierr);
  ierr = PetscRegisterFinalize(PetscB200Finalize);CHKERRQ(ierr);
  PetscB200Registered = PETSC_TRUE;
  PetscFunctionReturn(0);
}

/* run by PetscDLLibraryPrepend/Append when the program is started with -dll_prepend <this library>
   (src/sys/dll/dl.c:185-190 looks the symbol up by the library's base name) */
#undef __FUNCT__
#define __FUNCT__ "PetscDLLibraryRegister_petscb200plugin"
PetscErrorCode PetscDLLibraryRegister_petscb200plugin(const char path[])
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscB200Register();CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
