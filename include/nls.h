/*
 * nls.h -- C ABI of the B200-native time-stepping engine (libnls_b200.so).
 *
 * This is the drop-in boundary "B3" of SURVEY.md section 8(b): the only layer
 * that touches CUDA.  Everything above it (the C++ solver classes with the
 * reference's names in include/nls/, the *_driver CLIs, bench.py, the tests)
 * talks to the GPU exclusively through these entry points.  Plain C, opaque
 * handle, int status (0 = ok, <0 = error; text via nls_last_error()), caller
 * owned host buffers, no exceptions across the boundary, no CPU fallback:
 * every entry point fails with NLS_ERR_CUDA when no sm_100 device is present.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   nls_create/nls_destroy   device/nlse_solver_dev.hpp:53-92 (ctor/dtor of
 *                            NLSESolverDevice), device/sg_solver_dev.hpp:92-150,
 *                            nlsolvers/device/include/nlse_dev.hpp:78-176
 *   nls_set_field            H2D copies in those ctors (nlse_solver_dev.hpp:64)
 *   nls_step                 NLSESolverDevice::step nlse_solver_dev.hpp:94-111,
 *                            SGESolverDevice::step sg_single_dev.hpp:81-90,
 *                            host steps nlsolvers/host/include/nlse_cubic_solver.hpp:55-74,
 *                            sg_single_solver.hpp:43-59, sg_single_sv_solver.hpp:9-19
 *                            + neumann_bc_no_velocity boundaries.hpp:40-57
 *   nls_get_field            transfer_snapshots nlse_solver_dev.hpp:113-118
 *   nls_matfunc              MatrixFunctionApplicator{Real,Complex}::apply
 *                            device/matfunc_complex.hpp:155-177, matfunc_real.hpp:171-234;
 *                            host expm_multiply eigen_krylov_complex.hpp:54-83 etc.
 *   nls_lanczos              lanczos_iteration(_complex) device/lanczos.hpp:126-194,
 *                            device/lanczos_complex.hpp:318-551
 *   nls_stencil_apply        DeviceSpMV<T>::multiply device/spmv.hpp:65-72
 *   nls_reduce               calculate_energy_kernel_2d device/sg_solver_dev.hpp:51-80,
 *                            mass: finalized_scripts/compare_utils_complex_2d.py:142-143
 */
#ifndef NLS_H
#define NLS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct nls_engine *nls_handle;

/* status codes */
enum {
  NLS_OK = 0,
  NLS_ERR_ARG = -1,      /* bad descriptor / argument */
  NLS_ERR_CUDA = -2,     /* CUDA runtime error (incl. "no device") */
  NLS_ERR_STATE = -3,    /* call not valid for this handle's equation/integrator */
  NLS_ERR_NONFINITE = -4 /* the on-device tridiagonal eigensolve did not converge (non-finite state) */
};
/* Device-side failures (eigensolver non-convergence -> NLS_ERR_NONFINITE; grid-barrier / z-slab peer /
 * TMA pipeline time-out -> NLS_ERR_CUDA) are detected by the next nls_sync / nls_get_field* / nls_reduce /
 * nls_step_timed / nls_matfunc and are STICKY: the state was only partially advanced, so the handle is
 * poisoned and every later call except nls_destroy returns the same error. */

/* equation = pointwise nonlinearity */
enum {
  NLS_EQ_NLSE_CUBIC = 0, /* g = m |u|^2                      nlse_cubic_solver.hpp:55-74 */
  NLS_EQ_NLSE_CQ = 1,    /* g = m (s1|u|^2 + s2|u|^4)        nlse_cubic_quintic_solver.hpp:10-29 */
  NLS_EQ_NLSE_SAT = 2,   /* g = m |u|^2/(1+kappa u) (sic)    nlse_saturating_solver.hpp:11-33 */
  NLS_EQ_SGE = 3,        /* f = sin u, filter mod-cosine     sg_single_solver.hpp:43-59 */
  NLS_EQ_KGE = 4,        /* f = u^3, filter id_sqrt          kg_solver.hpp:13-22 */
  NLS_EQ_PHI4 = 5,       /* f = u - u^3                      phi4_solver.hpp:13-23 */
  NLS_EQ_SGE_DOUBLE = 6, /* f = sin u + 0.6 sin 2u           sg_double_solver.hpp:13-23 */
  NLS_EQ_SGE_SINH = 7    /* f = sinh u                       sg_hyperbolic_solver.hpp:13-23 */
};

/* integrator */
enum {
  NLS_INT_SS2 = 0,      /* Strang splitting, NLSE only */
  NLS_INT_GAUTSCHI = 1, /* Gautschi-type, real equations only */
  NLS_INT_SV = 2,       /* Stoermer-Verlet, real equations only */
  NLS_INT_SEWI = 3      /* NLSE two-step symmetric exponential wave integrator ("Gautschi" NLSE):
                         * u+ = e^{2 tau L} u- - 2 tau e^{tau L} sinc(dt L) B(u), B = -m g(|u|^2) u, tau = i dt
                         * (nlse_cubic_gautschi_solver_3d.hpp:8-35, nlse_cubic_quintic_gautschi_solver.hpp:11-41,
                         * nlse_saturating_gautschi_solver.hpp:11-44, device nlse_dev.hpp:205-238); the first
                         * step after nls_set_field(U) is one SS2 step with u- = u0
                         * (nlse_cubic_gautschi_driver.cpp:122-125, nlse_dev.hpp:206-209) */
};

/* Laplacian flavour */
enum {
  NLS_LAP_UNIFORM = 0, /* build_laplacian_noflux[_3d]              laplacians.hpp:9-52, :105-156 */
  NLS_LAP_ANISO = 1    /* build_anisotropic_laplacian_noflux[_3d]  laplacians.hpp:54-103, :158-218 */
};

/* re-orthogonalisation mode of the Lanczos process */
enum {
  NLS_REORTH_FULL = 0, /* full re-orthogonalisation against all previous columns (reference: MGS sweep), in the form the
                        * engine prefers: NLS_REORTH_GRAM where it applies (multi-kernel path, m <= 13, no NLS_FLAG_TMA /
                        * NLS_FLAG_MARCH), else NLS_REORTH_BLOCK; the persistent kernel has its own one-pass block form */
  NLS_REORTH_NONE = 1, /* plain 3-term recurrence (not the reference algorithm; opt-in) */
  NLS_REORTH_BLOCK = 4, /* force the two-pass block form: pass A = stencil + recurrence term + ALL dots, pass B =
                         * Gram-corrected block update (2j + 5 vector passes per iteration); multi-kernel path */
  NLS_REORTH_GRAM = 3, /* force the Gram-delayed form (multi-kernel path, m <= 13; larger m fall back to
                        * NLS_REORTH_BLOCK): per iteration ONE narrow stencil pass that writes the raw vector L v_j and
                        * measures only its products with v_j and v_{j-1}, and ONE update pass over all columns whose
                        * projections on the older columns are derived from the Gram entries of the two newest columns
                        * through the three-term recurrence -- j + 6 vector passes per iteration instead of 2j + 5.
                        * Tracks the reference's MGS sweep at rounding level (tests/test_single_sweep_prototype.py,
                        * tests/test_gpu_parity.py::test_gram_delayed_*); C5 47.6 -> 36.4, C3 9.3 -> 7.1, C4 6.2 -> 5.2
                        * ms/step (profiles/r02v_bench_gram_delayed_ab.txt) */
  NLS_REORTH_MGS = 2   /* the reference's sequential modified Gram-Schmidt sweep, operation for operation
                        * (eigen_krylov_complex.hpp:28-38): one launch per dot product, multi-kernel path only.
                        * Slow by construction; a bisecting tool that separates re-orthogonalisation rounding of
                        * NLS_REORTH_FULL from a real defect when a parity run drifts */
};

/* matrix functions f(T) applied through the Krylov basis */
enum {
  NLS_FN_EXP = 0,         /* exp(t*lambda), t complex          eigen_krylov_complex.hpp:54-83 */
  NLS_FN_SINC = 1,        /* sinc(t*lambda)                    eigen_krylov_complex.hpp:85-103 */
  NLS_FN_COS_SQRT = 2,    /* cos(t sqrt|lambda|)               eigen_krylov_real.hpp:57-87 */
  NLS_FN_SINC2_SQRT = 3,  /* sinc^2(t sqrt|lambda|)            eigen_krylov_real.hpp:89-123 */
  NLS_FN_ID_SQRT = 4,     /* t sqrt|lambda|                    eigen_krylov_real.hpp:125-152 */
  NLS_FN_SINC2_HALF = 5,  /* sinc^2(t/2 sqrt|lambda|)          eigen_krylov_real.hpp:154-183 */
  NLS_FN_MOD_COSINE = 6   /* [cos^2(th/2) sin(th)/th]^2        eigen_krylov_real.hpp:185-209 */
};

/* fields addressable through nls_set_field / nls_get_field */
enum {
  NLS_FIELD_U = 0,      /* state u (complex128 for NLSE, float64 otherwise) */
  NLS_FIELD_U_PAST = 1, /* previous state (real equations; NLS_INT_SEWI: u-, setting it skips the SS2 start-up) */
  NLS_FIELD_V = 2,      /* get: (u-u_past)/dt  sg_single_driver.cpp:113; set: u_past = u - dt*v (:70) */
  NLS_FIELD_M = 3,      /* m(x): float64, n values, shared by the whole batch */
  NLS_FIELD_C = 4       /* c(x): float64, n values (NLS_LAP_ANISO only) */
};

/* reductions */
enum {
  NLS_RED_MASS = 0,     /* sum |u|^2 * dx*dy[*dz] (per trajectory) */
  NLS_RED_NONFINITE = 1, /* number of non-finite entries of u */
  /* the energy of the reference's drift scripts, per trajectory (not on z-slab handles):
   *  complex state: finalized_scripts/compare_utils_complex_2d.py:134-151 -- centred differences
   *    on the interior sites, sum(|u_x|^2+|u_y|^2[+|u_z|^2]) dV - 1/2 sum(|u|^4) dV (m_eff = 1);
   *  real state: energy_sge.py:6-14 -- numpy.gradient differences (one-sided on the faces; the
   *    script passes dx for array axis 0 and dy for axis 1, kept), v = (u-u_past)/dt,
   *    1/2 dV sum 1/2(|grad u|^2 + v^2 + 2(1-cos u)).  3D uses (dz,dy,dx) for (z,y,x). */
  NLS_RED_ENERGY = 2
};

typedef struct nls_desc {
  uint32_t struct_size; /* = sizeof(nls_desc) */
  uint32_t nx, ny, nz;  /* nz == 1 => 2D. Layout: idx = (k*ny + j)*nx + i */
  double dx, dy, dz;    /* scale: 1/(dx*dy) in 2D, 1/(dx*dx) in 3D (laplacians.hpp:49,:153) */
  int32_t equation;
  int32_t integrator;
  int32_t laplacian;
  int32_t reorth;
  uint32_t krylov_m;    /* 2..64 */
  uint32_t batch;       /* number of independent trajectories in this handle (>=1) */
  double s1, s2, kappa; /* cubic-quintic / saturable parameters */
  double op_sign;       /* +1: operator is L, -1: operator is -L (kg_driver.cpp:93). 0 => +1 */
  double dt;            /* time step of the real integrators (used by V get/set) */
  int32_t device;       /* CUDA device ordinal */
  /* z-slab decomposition (3D only): this handle owns planes [z0, z0+nz) of a
   * global grid with nz_global planes; 0 => not decomposed (nz_global = nz). */
  uint32_t nz_global;
  uint32_t z0;
  uint32_t flags;       /* NLS_FLAG_* */
} nls_desc;

enum {
  NLS_FLAG_NO_GRAPH = 1u, /* launch kernels directly instead of replaying a CUDA graph */
  NLS_FLAG_NO_PERSISTENT = 2u, /* never use the persistent cooperative step kernel */
  NLS_FLAG_MARCH = 8u,      /* 3D: plane-marching stencil pass (z window in registers); measured slower than
                             * the TMA-staged and flat passes on B200, kept as an option */
  NLS_FLAG_TMA = 32u,       /* TMA-staged Lanczos passes (nls_tma.cuh: cp.async.bulk + mbarrier ring, arithmetic from
                             * shared memory).  Correct and tested, but measured slower than the register-staged
                             * passes on B200 so far, hence opt-in */
  NLS_FLAG_TMAP = 128u,     /* 3D single-trajectory grids: pass A of the Lanczos iteration with the stencil input staged through
                             * shared memory by tensor-map TMA loads (nls_tmap.cuh: cp.async.bulk.tensor.3d boxes with halo).
                             * DEFAULT for complex states (C3 10.6 -> 9.5, C5 61.7 -> 50.5 ms/step); the flag opts real
                             * states in */
  NLS_FLAG_NO_TMAP = 256u,  /* never use the tensor-map pass (flat register-streaming pass A instead) */
  NLS_FLAG_PK_SV = 64u,     /* Stoermer-Verlet through the persistent kernel (k_pk_sv); off by default: at
                             * 1024^2 one grid barrier per step (15.7 us) costs more than a k_sv launch (11-14 us) */
  NLS_FLAG_DEBUG_PK_FAULT = 0x80000000u, /* test hook: one CTA of the persistent kernel never reaches its first
                             * grid barrier, so the barrier times out (~2 s) and the abort path is exercised */
  NLS_FLAG_SEWI_HOST2D = 4u /* NLS_INT_SEWI with the sign convention of the 2D cubic HOST solver:
                             * u+ = e^{-2 tau L} u- + 2 tau e^{-tau L} sinc(dt L) B(u)
                             * (nlse_cubic_gautschi_solver.hpp:21-39) */
};

const char *nls_last_error(void);
int nls_version(void);
/* number of CUDA devices visible, or a negative status */
int nls_device_count(void);

int nls_create(const nls_desc *desc, nls_handle *out);
int nls_destroy(nls_handle h);

/* host -> device. bytes must equal the field's size (batch*n*scalar for U/U_PAST/V, n*8 for M/C).
 * Pageable source: returns after the copy.  Page-locked source: asynchronous (see nls_set_field_async). */
int nls_set_field(nls_handle h, int field, const void *host_src, size_t bytes);
/* device -> host (synchronous on the handle's stream). */
int nls_get_field(nls_handle h, int field, void *host_dst, size_t bytes);
/* same, but source/destination are device pointers on the handle's device */
int nls_set_field_dev(nls_handle h, int field, const void *dev_src, size_t bytes);
int nls_get_field_dev(nls_handle h, int field, void *dev_dst, size_t bytes);

/* advance nsteps time steps (each followed by the no-flux boundary copy).
 * Asynchronous on the handle's stream; nls_sync waits. */
int nls_step(nls_handle h, double dt, uint32_t nsteps);
int nls_sync(nls_handle h);

/* y = op_sign * L x for the handle's stencil; x,y host arrays of batch*n scalars */
int nls_stencil_apply(nls_handle h, const void *host_x, void *host_y);
/* out = f(t L) in  (Krylov approximation with the handle's m); host arrays */
int nls_matfunc(nls_handle h, int fn, double t_re, double t_im, const void *host_in,
                void *host_out);
/* run the Lanczos process on host_in; returns V (m columns of n scalars, column
 * after column), alpha[m] (diag of T; scalar type), beta[m] (beta[0]=||in||,
 * beta[j+1]=T(j+1,j)); trajectory 0 of the batch only. Any output may be NULL. */
int nls_lanczos(nls_handle h, const void *host_in, void *host_V, void *host_alpha,
                double *host_beta);
/* the same two with DEVICE pointers on the handle's device (the reference's signatures: DeviceSpMV<T>::multiply
 * device/spmv.hpp:65-72, MatrixFunctionApplicator*::apply device/matfunc_complex.hpp:155-156, matfunc_real.hpp:171);
 * stream-ordered on the handle's stream (nls_stream), no host synchronisation */
int nls_stencil_apply_dev(nls_handle h, const void *dev_x, void *dev_y);
int nls_matfunc_dev(nls_handle h, int fn, double t_re, double t_im, const void *dev_in, void *dev_out);
/* CSR compatibility check for the B2 class layer: rel_err = max|CSR x - L x| / max|L x| for a deterministic
 * pseudo-random x, L = the handle's matrix-free stencil.  The reference ships an assembled CSR to its solver
 * classes (sg_single_dev.hpp:28-31, device/spmv.hpp:30-52); this engine accepts one only if it IS its stencil.
 * values: double (values_complex = 0) or interleaved complex128 (1); arrays on the host (on_device = 0) or device. */
int nls_csr_residual(nls_handle h, const int *row_ptr, const int *col_ind, const void *values, uint32_t n, uint64_t nnz,
                     int values_complex, int on_device, double *rel_err);
int nls_pointer_is_device(const void *p);
int nls_memcpy_to_host(void *dst, const void *dev_src, size_t bytes);
/* out[batch] */
int nls_reduce(nls_handle h, int what, double *out);

/* Allocation-free mass reduction.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *   dev_out[t] = cell * sum_i |dev_u[t * traj_stride + i]|^2,  i < n_sites, t < batch
 * -- the NLS_RED_MASS value of nls_reduce (finalized_scripts/compare_utils_complex_2d.py:142-143) for a state the
 * CALLER holds in device memory: no handle, no cudaMalloc / cudaFree / allocating library routine inside.  The per-CTA
 * partial sums and the completion tickets of the grid reduction live in ONE file-scope static __device__ array
 * (NLS_NOALLOC_SCRATCH_BYTES, held by every GPU that loads the library until exit), hence the limits:
 * batch <= NLS_NOALLOC_MAX_BATCH, n_sites * batch <= NLS_NOALLOC_MAX_SITES; anything larger returns NLS_ERR_ARG
 * with a message.  dev_u: complex128 (is_complex = 1) or double (0); traj_stride in elements (0 = n_sites);
 * dev_out: `batch` doubles in device memory; stream: a cudaStream_t (NULL = the default stream); the call is
 * stream-ordered (one memset node + one kernel node, capturable) and does not synchronise.
 * All calls SHARE the scratch array, which keeps the last call's contents: the tickets are cleared at the start of
 * every call; calls that overlap on the GPU (different streams, or several host threads on one device) RACE and must
 * be serialised by the caller.  The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress): make it outside a capture. */
#define NLS_NOALLOC_MAX_BATCH 256u
#define NLS_NOALLOC_MAX_CTAS 4096u
#define NLS_NOALLOC_MAX_SITES (1ull << 31)
#define NLS_NOALLOC_SCRATCH_BYTES (NLS_NOALLOC_MAX_CTAS * 8u + NLS_NOALLOC_MAX_BATCH * 4u)
int nls_reduce_mass_noalloc(int device, const void *dev_u, uint64_t n_sites, uint32_t batch, uint64_t traj_stride,
                            int is_complex, double cell, double *dev_out, void *stream);

/* Snapshot egress with on-device downsampling: the field (NLS_FIELD_U, or NLS_FIELD_V
 * for real equations) is resampled on the device to tnx x tny [x tnz] points per trajectory with the
 * launcher's `downsample_interpolation` (finalized_scripts/downsampling.py:68-104, 3D :215-258:
 * scipy RegularGridInterpolator(method="linear") from linspace(-L,L,n) onto linspace(-L,L,t) per axis,
 * i.e. multilinear interpolation at index position k*(n-1)/(t-1)), and only the small result crosses
 * PCIe.  host_dst: [batch][tnz][tny][tnx] scalars, bytes must match; tnz must be 1 for 2D handles;
 * every target extent must be >= 2 (or equal to 1 together with the source extent).  Not on z-slab
 * handles. */
int nls_get_field_resampled(nls_handle h, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *host_dst,
                            size_t bytes);

/* ---- on-device field / initial-condition generators ----
 * The closed-form m(x) / c(x) fields of finalized_scripts/m_fields_2d.py:14-147 and c_fields_2d.py:12-169 and the
 * analytic initial conditions, evaluated on the device at x_i = numpy.linspace(-L, L, n)[i] of the GLOBAL grid (a
 * z-slab handle fills its own planes: no 16 GiB host array for the 1024^3 case).  field = sum of terms, then one
 * post-operation.  The random PARAMETERS of the scripts' fields are drawn on the host by the caller (tools mirror
 * the scripts' numpy RNG call order: nonlinear-solvers_b200/fieldgen.py); generators that need n x n random arrays
 * or image filters (multiscale, turbulent, topological, waveguide) are not closed-form and stay host-side.
 * Axes are the ENGINE's: x = fastest index.  (The scripts build X, Y with indexing='ij', i.e. their X runs along
 * the slow array axis = the engine's y.)
 *   NLS_GEN_CONST          p0
 *   NLS_GEN_GAUSS          p0 amp, p1..p3 centre (x,y,z), p4 width:   amp exp(-r^2 / (2 w^2))
 *   NLS_GEN_COSWAVE        p0 amp, p1..p3 k, p4 phase, p5 f; iparam 1 = sin:   amp cos(f (k.x) + phase)
 *   NLS_GEN_TANH_BOUNDARY  p0 m1, p1 m2, p2 radius, p3 width; iparam 0 circle 1 square 2 x 3 y 4 diagonal
 *   NLS_GEN_SINSIN         p0 base, p1 amp, p2 k:   base (1 + amp sin(k x) sin(k y))
 *   NLS_GEN_LAYERS         p0 base, p1 contrast, p2 L, p3 axis (0 x, 1 y); iparam number of layers
 *   NLS_GEN_CHECKER        p0 base, p1 cell, p2 sharpness (<= 0: sign)
 *   NLS_GEN_HALFSPACE      p0 base, p1 sharpness, p2 L, p3 axis
 *   NLS_GEN_GAUSS_WAVE     p0 amp, p1..p3 centre, p4 sigma, p5..p7 k:   amp exp(-r^2/(2 s^2)) exp(i k.x)  (complex)
 *   NLS_GEN_KINK           p0 amp, p1 a, p2 b:   amp atan(exp(a - b sqrt((x-y)^2/3 + (x+y)^2/2)))
 * post: NONE; CLAMP_MIN max(v, p0); MINMAX p0 (v - min)/(max - min); MAXABS v / max|v|; GRADIENT_MASS (no terms)
 * p0 (1 + p1 g^2/(g^2 + p2^2)) with g = central differences / 2 of the handle's c field (m_fields_2d.py:38-49).
 * field: NLS_FIELD_U (trajectory `traj`, or all with -1), NLS_FIELD_M, NLS_FIELD_C.  MINMAX / MAXABS need the
 * whole field on one device: not on z-slab handles. */
typedef struct nls_gen_term {
  int32_t kind;   /* NLS_GEN_* */
  int32_t iparam;
  double p[10];
} nls_gen_term;
enum { NLS_GEN_CONST = 0, NLS_GEN_GAUSS = 1, NLS_GEN_COSWAVE = 2, NLS_GEN_TANH_BOUNDARY = 3, NLS_GEN_SINSIN = 4,
       NLS_GEN_LAYERS = 5, NLS_GEN_CHECKER = 6, NLS_GEN_HALFSPACE = 7, NLS_GEN_GAUSS_WAVE = 8, NLS_GEN_KINK = 9 };
enum { NLS_GENPOST_NONE = 0, NLS_GENPOST_CLAMP_MIN = 1, NLS_GENPOST_MINMAX = 2, NLS_GENPOST_MAXABS = 3,
       NLS_GENPOST_GRADIENT_MASS = 4 };
int nls_generate_field(nls_handle h, int field, int32_t traj, const nls_gen_term *terms, uint32_t nterms, int post,
                       const double *post_params /* [3] */, double Lx, double Ly, double Lz);

/* ---- asynchronous snapshot egress / state upload ----
 * Replaces the blocking cudaMemcpy per snapshot of nlsolvers/device/include/nlse_dev.hpp:323-334
 * (store_snapshot_online) and sg_single_dev.hpp:96-100: the snapshot is copied device-to-device into a 2-deep
 * ring at this point of the handle's stream order (i.e. after all steps issued so far) and leaves for the host
 * on a separate D2H stream while the following nls_step calls run.  dst should be page-locked
 * (nls_host_alloc / nls_host_register); it holds the data after the next nls_wait / nls_sync.
 * dst_traj_stride: bytes between consecutive trajectories of a batch in dst (0 = dense [batch][n]), so an
 * ensemble's snapshot lands directly in a (B, num_snapshots, ...) trajectory file image.
 * nls_set_field_async uploads through a 2-deep staging ring on an H2D stream (the copy overlaps the steps still
 * running); the source must stay unchanged until nls_wait.  Fields: U, U_PAST, M, C. */
int nls_get_field_async(nls_handle h, int field, void *host_dst, size_t bytes, size_t dst_traj_stride);
int nls_get_field_resampled_async(nls_handle h, int field, uint32_t tnx, uint32_t tny, uint32_t tnz, void *host_dst,
                                  size_t bytes, size_t dst_traj_stride);
int nls_set_field_async(nls_handle h, int field, const void *host_src, size_t bytes);
/* waits for the steps AND the asynchronous transfers issued so far; reports device-side failures (= nls_sync) */
int nls_wait(nls_handle h);
/* page-locked host memory for callers without a CUDA runtime of their own */
int nls_host_alloc(void **out, size_t bytes);
int nls_host_free(void *p);
int nls_host_register(void *p, size_t bytes);
int nls_host_unregister(void *p);

/* Spectral downsampling of a 2D field on the device = the launcher's `downsample_fft`
 * (finalized_scripts/downsampling.py:10-34: ortho fft2 -> fftshift -> centred tny x tnx crop -> ifftshift -> ortho
 * ifft2), evaluated as a direct cropped DFT (two small dense products, no FFT library).  host_dst:
 * [batch][tny][tnx] scalars (real states: the real part, like the script).  2D single-GPU handles. */
int nls_get_field_fft_resampled(nls_handle h, int field, uint32_t tnx, uint32_t tny, void *host_dst, size_t bytes);
int nls_get_field_fft_resampled_async(nls_handle h, int field, uint32_t tnx, uint32_t tny, void *host_dst, size_t bytes,
                                      size_t dst_traj_stride);

/* number of kernel launches issued (directly or through graph replays) since creation */
uint64_t nls_launch_count(nls_handle h);
/* how many of them were TMA-staged Lanczos passes (nls_tma.cuh) */
uint64_t nls_tma_launch_count(nls_handle h);
/* how many were tensor-map pass-A launches (nls_tmap.cuh; counted at issue, also inside graph capture) */
uint64_t nls_tmap_launch_count(nls_handle h);
/* the CUDA stream (cudaStream_t) the handle launches on */
void *nls_stream(nls_handle h);
/* time a run of nsteps steps with CUDA events on the handle's stream; ms returned */
int nls_step_timed(nls_handle h, double dt, uint32_t nsteps, float *ms);

/* ---- z-slab decomposition across GPUs (one handle per GPU) ----
 * A handle created with nz_global > nz owns planes [z0, z0+nz) of the global grid (batch must
 * be 1, nz >= 2) and keeps one ghost plane per side.  Halo planes and the scalar sums of every
 * Lanczos reduction travel INSIDE the kernels through peer-mapped memory (NVLink); the host
 * only wires the handles together once:
 *   1. every rank calls nls_slab_export and the 128-byte blobs are gathered in rank order
 *      (rank r owns the r-th slab counting from z = 0), e.g. with an all-gather;
 *   2. every rank calls nls_slab_connect with all blobs.  Ranks may be processes (CUDA IPC)
 *      or host threads of one process (direct peer access).
 * After that nls_set_field(U, C), nls_step, nls_matfunc, nls_lanczos, nls_stencil_apply and
 * nls_reduce are COLLECTIVE: every rank must issue the same calls in the same order (they
 * exchange ghost planes / sums; a missing peer makes the call fail after ~3 s instead of
 * hanging).  Host arrays passed to a slab handle hold the local planes only; reductions and
 * Lanczos scalars returned are global.  All ranks must be synchronised (nls_sync + a host
 * barrier) before any of them calls nls_destroy. */
int nls_slab_export(nls_handle h, void *buf128);
int nls_slab_connect(nls_handle h, const void *all_blobs /* nranks*128 bytes */, int rank, int nranks);

#ifdef __cplusplus
}
#endif
#endif /* NLS_H */
