/* b200ksp.c — KSPCGB200 ("cgb200") and KSPGMRESB200 ("gmresb200"): the device-resident Krylov drivers of libpetscb200
 * (b200_ksp_solve: KSPSolve_CG cg.c:92-286 / KSPSolve_GMRES gmres.c:118-238 restated with fused kernels, the scalar
 * recurrences and KSPDefaultConverged running on the device) as NEW KSP types registered with KSPRegister
 * (reference src/ksp/ksp/interface/itregis.c:69, include/petscksp.h:83) — SURVEY 8(f) row 4: they must be new types so
 * that the reference's KSPCG / KSPGMRES (which already run on the B200 Vec/Mat types op by op) stay untouched.
 *
 *   ./ex2 -dll_prepend libpetscb200plugin.so -mat_type aijb200 -vec_type b200 -ksp_type cgb200 -pc_type jacobi
 *
 * Supported: operators of type seqaijb200 / mpiaijb200, PCJACOBI (plain diagonal) or PCNONE, left preconditioning,
 * KSPDefaultConverged.  Anything else is refused with PETSC_ERR_SUP (no silent fallback).  Residual norms are logged
 * and monitored after the device loop returns, in iteration order, so -ksp_monitor output has the reference's shape. */
#include "petscb200plugin.h"
#include <petsc-private/kspimpl.h>
#include <petscpc.h>
#include "petscb200.h"

#define CHKB200(call) do { int b200rc_ = (call); if (b200rc_) SETERRQ2(PETSC_COMM_SELF,b200rc_ == B200_ERR_FP ? PETSC_ERR_FP : PETSC_ERR_LIB,"libpetscb200 error %d: %s",b200rc_,b200_last_error()); } while (0)

typedef struct {
  int       kind;              /* 0 cg, 1 gmres                                  */
  PetscInt  restart;           /* -ksp_gmres_restart                             */
  PetscInt  check_every;       /* -ksp_b200_check_every: host looks at the device-side convergence flag every k iterations */
  PetscBool fused;             /* -ksp_b200_fused 0: the reference's op-by-op call sequence on the device                  */
  PetscBool mgs;               /* -ksp_gmres_modifiedgramschmidt (gmres.c:534)                                            */
  KSPGMRESCGSRefinementType refine; /* -ksp_gmres_cgs_refinement_type (gmres.c:529)                                          */
} KSP_B200;

#undef __FUNCT__
#define __FUNCT__ "KSPSetUp_B200"
static PetscErrorCode KSPSetUp_B200(KSP ksp)
{
  PetscFunctionBegin;                                         /* work vectors live inside the library */
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPSolve_B200"
static PetscErrorCode KSPSolve_B200(KSP ksp)
{
  KSP_B200          *kb = (KSP_B200*)ksp->data;
  Mat               Amat,Pmat;
  MatStructure      flag;
  PetscBool         isjacobi,isnone,isdefault;
  b200_mat          A;
  b200_ksp_options  o;
  b200_ksp_result   res;
  const PetscScalar *b;
  PetscScalar       *x;
  double            *hist;
  PetscInt          k,histmax;
  Vec               ones = 0,dinv = 0;
  PetscErrorCode    ierr;

  PetscFunctionBegin;
  ierr = PCGetOperators(ksp->pc,&Amat,&Pmat,&flag);CHKERRQ(ierr);
  if (Amat != Pmat) SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types need Amat == Pmat");
  ierr = PetscObjectTypeCompare((PetscObject)ksp->pc,PCJACOBI,&isjacobi);CHKERRQ(ierr);
  ierr = PetscObjectTypeCompare((PetscObject)ksp->pc,PCNONE,&isnone);CHKERRQ(ierr);
  if (!isjacobi && !isnone) SETERRQ1(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types support -pc_type jacobi or none, not %s",((PetscObject)ksp->pc)->type_name);
  if (ksp->pc_side != PC_LEFT) SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types support left preconditioning only");
  if (ksp->transpose_solve) SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types do not support KSPSolveTranspose");
  if (ksp->dscale || ksp->nullsp) SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types do not support diagonal scaling or null spaces");
  /* KSPDefaultConverged only (the convergence test runs on the device); -ksp_converged_use_initial_residual_norm etc. refused */
  isdefault = (ksp->converged == KSPDefaultConverged) ? PETSC_TRUE : PETSC_FALSE;
  if (!isdefault) SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types support KSPDefaultConverged only");
  if (ksp->cnvP && (((KSPDefaultConvergedCtx*)ksp->cnvP)->initialrtol || ((KSPDefaultConvergedCtx*)ksp->cnvP)->mininitialrtol))
    SETERRQ(((PetscObject)ksp)->comm,PETSC_ERR_SUP,"KSP B200 types do not support the initial-residual variants of the default test");

  ierr = MatB200GetHandle(Amat,(void**)&A);CHKERRQ(ierr);      /* uploads / refreshes the device matrix */
  CHKB200(b200_ksp_default_options(&o));
  o.ksp_type      = kb->kind;
  o.pc_type       = isjacobi ? 1 : 0;
  o.rtol          = ksp->rtol; o.abstol = ksp->abstol; o.dtol = ksp->divtol;
  o.max_it        = (int)ksp->max_it;
  o.norm_type     = (int)ksp->normtype;                         /* KSP_NORM_NONE..NATURAL == B200_KSP_NORM_* */
  o.guess_nonzero = ksp->guess_zero ? 0 : 1;
  o.gmres_restart = (int)kb->restart;
  o.gmres_cgs_refine  = (int)kb->refine;                        /* KSP_GMRES_CGS_REFINE_NEVER, _IFNEEDED, _ALWAYS = 0, 1, 2 */
  o.gmres_modified_gs = kb->mgs ? 1 : 0;
  o.fused         = kb->fused ? 1 : 0;
  o.check_every   = (int)kb->check_every;
  histmax = ksp->max_it + 2;
  ierr = PetscMalloc(histmax*sizeof(double),&hist);CHKERRQ(ierr);
  ierr = VecB200GetDeviceRead(ksp->vec_rhs,&b);CHKERRQ(ierr);
  if (ksp->guess_zero) { ierr = VecB200GetDeviceWrite(ksp->vec_sol,&x);CHKERRQ(ierr); }
  else                 { ierr = VecB200GetDeviceReadWrite(ksp->vec_sol,&x);CHKERRQ(ierr); }
  if (isjacobi) {
    /* the inverse diagonal comes from the PC itself, which KSPSetUp has set up (itfunc.c:267): PCApply_Jacobi on a vector
       of ones (jacobi.c:266-278) is exactly jac->diag, whichever of -pc_jacobi_rowmax/_rowsum/_abs built it (jacobi.c:125-206) */
    ierr = VecDuplicate(ksp->vec_rhs,&ones);CHKERRQ(ierr);
    ierr = VecDuplicate(ksp->vec_rhs,&dinv);CHKERRQ(ierr);
    ierr = VecSet(ones,1.0);CHKERRQ(ierr);
    ierr = PCApply(ksp->pc,ones,dinv);CHKERRQ(ierr);
    ierr = VecB200GetDeviceRead(dinv,&o.pc_jacobi_inverse_diag);CHKERRQ(ierr);
  }
  CHKB200(b200_ksp_solve(A,&o,b,x,hist,(int)histmax,&res));
  ierr = VecDestroy(&ones);CHKERRQ(ierr);
  ierr = VecDestroy(&dinv);CHKERRQ(ierr);

  /* replay the residual history through the reference's bookkeeping (kspimpl.h:131 KSPLogResidualHistory, KSPMonitor) */
  ierr = PetscObjectTakeAccess(ksp);CHKERRQ(ierr);
  ksp->its = 0;
  ierr = PetscObjectGrantAccess(ksp);CHKERRQ(ierr);
  for (k=0; k<res.nhist; k++) {
    ksp->rnorm = hist[k];
    KSPLogResidualHistory(ksp,hist[k]);
    ierr = KSPMonitor(ksp,k,hist[k]);CHKERRQ(ierr);
  }
  ierr = PetscObjectTakeAccess(ksp);CHKERRQ(ierr);
  ksp->its   = res.its;
  ksp->rnorm = res.rnorm;
  ierr = PetscObjectGrantAccess(ksp);CHKERRQ(ierr);
  ksp->reason = (KSPConvergedReason)res.reason;
  ierr = PetscFree(hist);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPDestroy_B200"
static PetscErrorCode KSPDestroy_B200(KSP ksp)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscFree(ksp->data);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPSetFromOptions_B200"
static PetscErrorCode KSPSetFromOptions_B200(KSP ksp)
{
  KSP_B200       *kb = (KSP_B200*)ksp->data;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscOptionsHead("KSP B200 options");CHKERRQ(ierr);
  ierr = PetscOptionsInt("-ksp_gmres_restart","Number of Krylov search directions","KSPGMRESSetRestart",kb->restart,&kb->restart,PETSC_NULL);CHKERRQ(ierr);
  if (kb->kind == 1) {
    PetscBool flg;
    ierr = PetscOptionsEnum("-ksp_gmres_cgs_refinement_type","Type of iterative refinement for classical (unmodified) Gram-Schmidt","KSPGMRESSetCGSRefinementType",
                            KSPGMRESCGSRefinementTypes,(PetscEnum)kb->refine,(PetscEnum*)&kb->refine,PETSC_NULL);CHKERRQ(ierr);
    ierr = PetscOptionsBoolGroupBegin("-ksp_gmres_classicalgramschmidt","Classical (unmodified) Gram-Schmidt (fast)","KSPGMRESSetOrthogonalization",&flg);CHKERRQ(ierr);
    if (flg) kb->mgs = PETSC_FALSE;
    ierr = PetscOptionsBoolGroupEnd("-ksp_gmres_modifiedgramschmidt","Modified Gram-Schmidt (slow,more stable)","KSPGMRESSetOrthogonalization",&flg);CHKERRQ(ierr);
    if (flg) kb->mgs = PETSC_TRUE;
  }
  ierr = PetscOptionsInt("-ksp_b200_check_every","Iterations between host looks at the device-side convergence flag","",kb->check_every,&kb->check_every,PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsBool("-ksp_b200_fused","Fused kernels (0: the reference's op-by-op sequence)","",kb->fused,&kb->fused,PETSC_NULL);CHKERRQ(ierr);
  ierr = PetscOptionsTail();CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPView_B200"
static PetscErrorCode KSPView_B200(KSP ksp,PetscViewer viewer)
{
  KSP_B200       *kb = (KSP_B200*)ksp->data;
  PetscBool      iascii;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscObjectTypeCompare((PetscObject)viewer,PETSCVIEWERASCII,&iascii);CHKERRQ(ierr);
  if (iascii) {
    if (kb->kind && kb->mgs) { ierr = PetscViewerASCIIPrintf(viewer,"  GMRES on the B200: restart=%D, modified Gram-Schmidt (one VecDot + one VecAXPY per basis vector)\n",kb->restart);CHKERRQ(ierr); }
    else if (kb->kind) { ierr = PetscViewerASCIIPrintf(viewer,"  GMRES on the B200: restart=%D, classical Gram-Schmidt (one VecMDot + one VecMAXPY kernel pass per step), %s\n",kb->restart,KSPGMRESCGSRefinementTypes[kb->refine]);CHKERRQ(ierr); }
    else          { ierr = PetscViewerASCIIPrintf(viewer,"  CG on the B200: %s, host checks every %D iterations\n",kb->fused ? "fused kernels" : "op-by-op sequence",kb->check_every);CHKERRQ(ierr); }
  }
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPCreate_B200_Private"
static PetscErrorCode KSPCreate_B200_Private(KSP ksp,int kind)
{
  KSP_B200       *kb;
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = PetscNewLog(ksp,KSP_B200,&kb);CHKERRQ(ierr);
  kb->kind = kind; kb->restart = 30; kb->check_every = 16; kb->fused = PETSC_TRUE;
  ksp->data = (void*)kb;
  ierr = KSPSetSupportedNorm(ksp,KSP_NORM_PRECONDITIONED,PC_LEFT,2);CHKERRQ(ierr);
  if (!kind) {                                                  /* cg.c:434-437 */
    ierr = KSPSetSupportedNorm(ksp,KSP_NORM_UNPRECONDITIONED,PC_LEFT,1);CHKERRQ(ierr);
    ierr = KSPSetSupportedNorm(ksp,KSP_NORM_NATURAL,PC_LEFT,1);CHKERRQ(ierr);
    ierr = KSPSetSupportedNorm(ksp,KSP_NORM_NONE,PC_LEFT,1);CHKERRQ(ierr);
  }
  ksp->ops->setup          = KSPSetUp_B200;
  ksp->ops->solve          = KSPSolve_B200;
  ksp->ops->destroy        = KSPDestroy_B200;
  ksp->ops->view           = KSPView_B200;
  ksp->ops->setfromoptions = KSPSetFromOptions_B200;
  ksp->ops->buildsolution  = KSPDefaultBuildSolution;
  ksp->ops->buildresidual  = KSPDefaultBuildResidual;
  PetscFunctionReturn(0);
}

EXTERN_C_BEGIN
#undef __FUNCT__
#define __FUNCT__ "KSPCreate_CGB200"
PetscErrorCode KSPCreate_CGB200(KSP ksp)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = KSPCreate_B200_Private(ksp,0);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}

#undef __FUNCT__
#define __FUNCT__ "KSPCreate_GMRESB200"
PetscErrorCode KSPCreate_GMRESB200(KSP ksp)
{
  PetscErrorCode ierr;

  PetscFunctionBegin;
  ierr = KSPCreate_B200_Private(ksp,1);CHKERRQ(ierr);
  PetscFunctionReturn(0);
}
EXTERN_C_END
