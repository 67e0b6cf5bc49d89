/* Empty on this box: the reference's configure would place prototypes for missing libc functions here. */
#if !defined(INCLUDED_PETSCFIX_H)
#define INCLUDED_PETSCFIX_H
#endif
