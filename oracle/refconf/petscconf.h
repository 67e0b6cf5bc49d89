/* Hand-written configuration header for compiling the reference (erdc/petsc-dev, PETSc 3.3-dev) C sources
 * on this box as the parity oracle / CPU baseline.  The reference's configure engine (config/BuildSystem)
 * is an un-vendored submodule, so `./configure` cannot generate this file; the macro set follows
 * SURVEY.md appendix A.4.  Real double scalars, 32-bit PetscInt, MPIUNI (one rank), no external packages,
 * no X11, PETSC_USE_DEBUG off.  Test infrastructure only. */
#if !defined(INCLUDED_PETSCCONF_H)
#define INCLUDED_PETSCCONF_H

#define PETSC_ARCH "arch-b200-oracle"
#define PETSC_DIR "/root/reference"
#define PETSC_LIB_DIR "/root/repo/oracle/_ref"
#define PETSC_HAVE_MPIUNI 1
#define PETSC_USE_REAL_DOUBLE 1
#define PETSC_USE_SCALAR_REAL 1
#define PETSC_CLANGUAGE_C 1
#define PETSC_BLASLAPACK_UNDERSCORE 1
#define PETSC_USE_SINGLE_LIBRARY 1

#define PETSC_HAVE_STDLIB_H 1
#define PETSC_HAVE_STRING_H 1
#define PETSC_HAVE_STRINGS_H 1
#define PETSC_HAVE_MEMORY_H 1
#define PETSC_HAVE_MALLOC_H 1
#define PETSC_HAVE_MATH_H 1
#define PETSC_HAVE_STDINT_H 1
#define PETSC_HAVE_UNISTD_H 1
#define PETSC_HAVE_SYS_TIME_H 1
#define PETSC_HAVE_SYS_TYPES_H 1
#define PETSC_HAVE_SYS_PARAM_H 1
#define PETSC_HAVE_SYS_STAT_H 1
#define PETSC_HAVE_SYS_RESOURCE_H 1
#define PETSC_HAVE_SYS_UTSNAME_H 1
#define PETSC_HAVE_SYS_WAIT_H 1
#define PETSC_HAVE_LIMITS_H 1
#define PETSC_HAVE_FLOAT_H 1
#define PETSC_HAVE_FCNTL_H 1
#define PETSC_HAVE_PWD_H 1
#define PETSC_HAVE_TIME_H 1
#define PETSC_HAVE_DLFCN_H 1
#define PETSC_HAVE_NETDB_H 1
#define PETSC_HAVE_STDARG_H 1
#define PETSC_TIME_WITH_SYS_TIME 1
#define PETSC_HAVE_GETTIMEOFDAY 1
#define PETSC_HAVE_TIME 1
#define PETSC_HAVE_GETCWD 1
#define PETSC_HAVE_GETPWUID 1
#define PETSC_HAVE_GETHOSTNAME 1
#define PETSC_HAVE_GETDOMAINNAME 1
#define PETSC_HAVE_UNAME 1
#define PETSC_HAVE_REALPATH 1
#define PETSC_HAVE_READLINK 1
#define PETSC_HAVE_ACCESS 1
#define PETSC_HAVE_STAT 1
#define PETSC_HAVE_LSEEK 1
#define PETSC_HAVE_MEMMOVE 1
#define PETSC_HAVE_MEMALIGN 1
#define PETSC_HAVE_VSNPRINTF 1
#define PETSC_HAVE_VPRINTF 1
#define PETSC_HAVE_VFPRINTF 1
#define PETSC_HAVE_SNPRINTF 1
#define PETSC_HAVE_VA_COPY 1
#define PETSC_HAVE_SLEEP 1
#define PETSC_HAVE_USLEEP 1
#define PETSC_HAVE_NANOSLEEP 1
#define PETSC_HAVE_GETRUSAGE 1
#define PETSC_HAVE_CLOCK 1
#define PETSC_HAVE_DRAND48 1
#define PETSC_HAVE_RAND 1
#define PETSC_HAVE_STRCASECMP 1
#define PETSC_HAVE_GETPAGESIZE 1
#define PETSC_HAVE_POPEN 1
#define PETSC_HAVE_FORK 1
#define PETSC_HAVE_SIGNAL 1
#define PETSC_HAVE_SIGACTION 1
#define PETSC_HAVE_SIGSET 1
#define PETSC_HAVE_DLOPEN 1
#define PETSC_HAVE_DLSYM 1
#define PETSC_HAVE_DLERROR 1
/* -dll_prepend / -dll_append really dlopen() the named library (src/sys/dll/dlimpl.c:9,255): this is how the
   out-of-tree B200 plugin gets into an unmodified example program */
#define PETSC_HAVE_DYNAMIC_LIBRARIES 1
#define PETSC_HAVE_RTLD_LAZY 1
#define PETSC_HAVE_RTLD_NOW 1
#define PETSC_HAVE_RTLD_GLOBAL 1
#define PETSC_HAVE_RTLD_LOCAL 1
#define PETSC_HAVE_DLCLOSE 1
#define PETSC_HAVE_ISNAN 1
#define PETSC_HAVE_ISINF 1
#define PETSC_HAVE_ERF 1
#define PETSC_HAVE_FENV_H 1
#define PETSC_HAVE_BUILTIN_EXPECT 1
#define PETSC_HAVE_DOUBLE_ALIGN_MALLOC 1
#define PETSC_USE_PROC_FOR_SIZE 1
#define PETSC_HAVE_MPI_COMM_F2C 1
#define PETSC_HAVE_MPI_COMM_C2F 1
#define PETSC_HAVE_MPI_FINT 1

#define PETSC_SIZEOF_VOID_P 8
#define PETSC_SIZEOF_INT 4
#define PETSC_SIZEOF_LONG 8
#define PETSC_SIZEOF_LONG_LONG 8
#define PETSC_SIZEOF_SIZE_T 8
#define PETSC_SIZEOF_SHORT 2
#define PETSC_SIZEOF_CHAR 1
#define PETSC_SIZEOF_FLOAT 4
#define PETSC_SIZEOF_DOUBLE 8
#define PETSC_SIZEOF_MPI_COMM 4
#define PETSC_SIZEOF_MPI_FINT 4
#define PETSC_BITS_PER_BYTE 8
#define PETSC_MEMALIGN 16
#define PETSC_LEVEL1_DCACHE_LINESIZE 64
#define PETSC_LEVEL1_DCACHE_SIZE 32768
#define PETSC_LEVEL1_DCACHE_ASSOC 8
#define PETSC_RETSIGTYPE void
#define PETSC_FUNCTION_NAME_C __func__
#define PETSC_FUNCTION_NAME __func__
#define PETSC_STATIC_INLINE static inline
#define PETSC_C_STATIC_INLINE static inline
#define PETSC_RESTRICT __restrict
#define PETSC_C_RESTRICT __restrict
#define PETSC_UNUSED __attribute__((unused))
#define PETSC_Prefetch(a,b,c) __builtin_prefetch((a),(b),(c))
#define PETSC_PREFETCH_HINT_NTA 0
#define PETSC_PREFETCH_HINT_T2 1
#define PETSC_PREFETCH_HINT_T1 2
#define PETSC_PREFETCH_HINT_T0 3
#define PETSC_DIR_SEPARATOR '/'
#define PETSC_PATH_SEPARATOR ':'
#define PETSC_REPLACE_DIR_SEPARATOR '\\'
#define PETSC_USE_ERRORCHECKING 1
#define PETSC_USE_INFO 1
#define PETSC_USE_LOG 1
#define PETSC_USE_CTABLE 1
#define PETSC_USE_BACKWARD_LOOP 1
#define PETSC_IS_COLOR_VALUE_TYPE short
#define MPIU_COLORING_VALUE MPI_UNSIGNED_SHORT
#define IS_COLORING_MAX 65535
#define PETSC_UINTPTR_T uintptr_t
#define PETSC__GNU_SOURCE 1
#define PETSC__BSD_SOURCE 1
#define PETSC__DEFAULT_SOURCE 1
#define PETSC_USE_GDB_DEBUGGER 1
#define PETSC_VERSION_HG "unknown"
#define PETSC_VERSION_DATE_HG "unknown"
#define PETSC_SLSUFFIX "so"
#define PETSC_CXX_STATIC_INLINE static inline
#define PETSC_CXX_RESTRICT __restrict
#define PETSC_HAVE_SCHED_CPU_SET_T 1
#define PETSC_BLASLAPACK_STDCALL_DISABLED 1

#endif
