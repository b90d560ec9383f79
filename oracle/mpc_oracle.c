/*
 * mpc_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU oracle for the MPCGPU linear-system hot path (form_schur_system -> pcg),
 * fp32 and fp64 instantiations of oracle_impl.h.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product (mpcgpu_b200/csrc) never links or calls it.
 *
 * Exported: oracle_schur_{f32,f64}, oracle_bdmv_{f32,f64}, oracle_pcg_{f32,f64},
 *           oracle_csc_values_{f32,f64}, oracle_csc_pattern, oracle_dz_{f32,f64}.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>

#define ORACLE_MAX_DIM 32

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)

#define REAL float
#define F(name) CAT(name, _f32)
#define FABS(x) fabsf(x)
#include "oracle_impl.h"
#undef REAL
#undef F
#undef FABS

#define REAL double
#define F(name) CAT(name, _f64)
#define FABS(x) fabs(x)
#include "oracle_impl.h"
#undef REAL
#undef F
#undef FABS

/* CSC pattern of the upper triangle of the block-tridiagonal S; restates
 * prep_csr (include/utils/csr.cuh:40-74).  col_ptr: s*N+1, row_ind: nnz. */
void oracle_csc_pattern(int s, int N, int *col_ptr, int *row_ind)
{
    const int tri = (s + 1) * s / 2, brow_ct = s * s + tri;
    int b, row, col;
    col_ptr[0] = 0;
    for (b = 0; b < N; b++)
        for (row = 0; row < s; row++) {
            int off = (b > 0) * tri + (b > 0) * (b - 1) * brow_ct + (b > 0) * row * s
                    + ((row + 1) * row) / 2;
            int len = (b > 0) * s + row + 1;
            col_ptr[b * s + row + 1] = off + len;
            for (col = 0; col < len; col++)
                row_ind[off + col] = (b > 0) * (b - 1) * s + col;
        }
}
