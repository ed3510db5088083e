/* TEST INFRASTRUCTURE: sparse multifrontal Cholesky of the 3x3-block normal equations (oracle/cpu_chol.c). */
#ifndef GS_CPU_CHOL_H
#define GS_CPU_CHOL_H
typedef struct chol chol;
/* nested-dissection ordering + symbolic factorisation of the block pattern (full, symmetric, diagonal included) */
chol* chol_analyze(int n, const int* rowptr, const int* colidx);
/* numeric factorisation of H + lambda*D; H = 9 doubles per pattern entry (row-major blocks); 0 ok, 1 not SPD */
int chol_factor(chol* C, const double* H, const int* diag, double lambda, const double* dscale);
/* x = (H + lambda D)^-1 rhs, 3 doubles per vertex in the caller's vertex order */
void chol_solve(chol* C, const double* rhs, double* x);
void chol_info(const chol* C, double* nnz_l, double* flops, int* max_front, int* nnodes);
void chol_free(chol* C);
#endif
