/*
 * nls_oracle.h -- CPU restatement of the reference's host (Eigen) algorithm.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under nonlinear-solvers_b200/ or include/
 * may include, link or call this.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, as the checker or
 * the timed CPU baseline -- never as the product path.
 *
 * PARITY STATUS: pinned against the reference's own host sources; Eigen's
 * arithmetic substituted.  The reference's CPU headers are compiled UNMODIFIED,
 * where they lie, against a mini-Eigen stand-in (oracle/ref_shim, Eigen3 itself
 * is neither vendored nor installed) into
 * oracle/_ref/libnls_ref.so (oracle/ref_harness.cpp, oracle/Makefile), and
 * tests/test_reference_pin.py checks this restatement against it: matrices
 * identical, Lanczos / matrix functions / all step functions <= 1e-12.
 * Unpinned remainder: Eigen's own floating-point behaviour (dot summation order,
 * its eigensolver) -- the reference ships no golden vectors.  Independent
 * anchors: (i) the reference's own external check,
 * scipy.sparse.linalg.expm_multiply
 * (nlsolvers/scipy-test/check_krylov_compute.py:41-49), (ii) an independent
 * numpy/scipy twin (oracle/oracle_np.py) and (iii) dense eigh-based evaluation
 * of f(tL)u on small grids; see tests/test_oracle.py.
 *
 * The C API mirrors include/nls.h (same descriptor, same enums) so parity tests
 * drive both sides with identical calls.
 */
#ifndef NLS_ORACLE_H
#define NLS_ORACLE_H
#include "../include/nls.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_engine *orc_handle;

int orc_create(const nls_desc *desc, orc_handle *out);
int orc_destroy(orc_handle h);
int orc_set_field(orc_handle h, int field, const void *src, size_t bytes);
int orc_get_field(orc_handle h, int field, void *dst, size_t bytes);
int orc_step(orc_handle h, double dt, uint32_t nsteps);
int orc_stencil_apply(orc_handle h, const void *x, void *y);
int orc_matfunc(orc_handle h, int fn, double t_re, double t_im, const void *in, void *out);
int orc_lanczos(orc_handle h, const void *in, void *V, void *alpha, double *beta);
int orc_reduce(orc_handle h, int what, double *out);
int orc_get_field_resampled(orc_handle h, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *dst, size_t bytes);
/* CSR of the assembled operator (for the scipy anchor): nnz returned through *nnz;
 * pass NULL arrays to query sizes. values are real (the stencil weights). */
int orc_csr(orc_handle h, int64_t *nnz, int64_t *rowptr, int32_t *col, double *val);
/* number of OpenMP threads the batch loop will use */
int orc_threads(void);
void orc_set_threads(int n);
const char *orc_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
