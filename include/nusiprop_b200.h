/* nusiprop_b200.h -- C ABI of the B200-native nuSIprop redshift-evolution path.
 *
 * This is the drop-in boundary: plain C types, caller-owned buffers, no exceptions, no torch
 * types.  The reference has no FFI of its own -- its public surface for this path is the
 * header-only C++ class nuSIprop::calculate_flux (reference nuSIprop.hpp:22-540) and the Cython
 * wrapper pyprop (reference nuSIprop.pyx:12-144).  Each entry point below names the reference
 * interface it replaces; the drop-in class in nusiprop_b200/cxx/nuSIprop.hpp and the Python
 * mirror nusiprop_b200.pyprop (nusiprop_b200/api.py) are thin callers of this ABI.
 *
 * Threading: every call that uses a context takes that context's own lock, so concurrent calls on ONE
 * context are serialised by the library (objects of the drop-in class share contexts); distinct contexts
 * are independent.  All functions return NUSI_OK (0) or a negative NUSI_E_* code;
 * nusi_strerror() maps codes to text.  There is no CPU compute path: without a CUDA device
 * nusi_ctx_create fails with NUSI_E_NO_DEVICE.
 */
#ifndef NUSIPROP_B200_H
#define NUSIPROP_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define NUSI_ABI_VERSION 4

/* error codes (function return values) */
#define NUSI_OK 0
#define NUSI_E_INVALID (-1)   /* bad argument */
#define NUSI_E_NO_DEVICE (-2) /* no usable CUDA device / wrong architecture */
#define NUSI_E_CUDA (-3)      /* CUDA runtime error; see nusi_last_error() */
#define NUSI_E_TABLE (-4)     /* phi-phi table file missing or malformed (reference: interp.hpp:251-254 exit(1)) */
#define NUSI_E_NOMEM (-5)

/* per-point status bits written by the evolve calls (0 = clean) */
#define NUSI_ST_WARN_GAMMA 1      /* reference prints "Negative cross section when computing Gamma"      nuSIprop.hpp:909 */
#define NUSI_ST_WARN_ALPHATILDE 2 /* ... alphaTilde                                                     nuSIprop.hpp:1215 */
#define NUSI_ST_WARN_ALPHA 4      /* ... alpha                                                          nuSIprop.hpp:1505 */
#define NUSI_ST_PP_RANGE 8        /* phi-phi lookup outside the table (reference: exit(1))              interp.hpp:355-361 */
#define NUSI_ST_NO_MASS 256       /* no neutrino mass spectrum for this mntot (reference: exit(1))      aux.hpp:48-49 */
#define NUSI_ST_BAD_PARAM 512     /* mphi <= 0, or a non-finite mphi / g / mntot: the point is not evolved, its fluxes are NaN */

typedef struct nusi_ctx nusi_ctx;       /* one (device, option set): grids, source term, tables */
typedef struct nusi_points nusi_points; /* a batch of prepared parameter points resident in HBM */

/* Fixed options of a context = the constructor arguments of calculate_flux that are NOT the
 * mutable public members (reference nuSIprop.hpp:61-65): majorana, non_resonant,
 * normal_ordering, N_bins_E, lEmin, lEmax, zmax, flav, phiphi. */
typedef struct nusi_opts {
  int majorana;          /* default 1 */
  int non_resonant;      /* default 1 */
  int normal_ordering;   /* default 1 */
  int N_bins_E;          /* default 300 */
  double lEmin;          /* default 12.0 */
  double lEmax;          /* default 17.0 */
  double zmax;           /* default 5.0 */
  int flav;              /* default 2 */
  int phiphi;            /* default 0 (C++ default; the Python wrapper defaults to 1, nuSIprop.pyx:52) */
  const char *alphatilde_phiphi_path; /* NULL -> "xsec/alphatilde_phiphi.bin" relative to CWD (nuSIprop.hpp:168) */
  const char *alpha_phiphi_path;      /* NULL -> "xsec/alpha_phiphi.bin"      (nuSIprop.hpp:169) */
  int device;            /* CUDA device ordinal, default 0 */
  int max_points_per_launch; /* 0 = choose from free HBM */
  int elementary_mode;   /* alpha builder (Majorana, non_resonant):
                            0 = strict: every log/arg evaluated where the reference evaluates it;
                            1 = the logs/args of a table vertex are formed from one vertex log plus per-edge logs;
                            2 = 1 + per-vertex grouping of the s-t / t-u interference sums, 16x16-vertex
                                tiles, all three target masses per barrier (same closed forms);
                            3 (default) = 2, and runs of consecutive points with identical (mphi, mntot) -- the rows
                                of a (mphi, g) scan -- share everything that does not depend on g (evaluated once
                                per run, at most 32 points); bitwise identical to mode 2, 2.9x faster on a scan.
                                Put the points of a scan in mphi-major order to profit; any order is correct.
                            Parity with the reference is at the reference's own FMA-on/off noise floor in all
                            all modes. */
  int source_model;      /* 0 (default) = the reference's source: diffuse supernova background, Fermi-Dirac spectrum
                                (nuSIprop.hpp:638-662; si and norm are inert there);
                            1 = power law of the upstream code that the reference keeps commented out (:648-657):
                                L_i(z, E) = norm_total/3 SFR(z) (E (1+z)/E0)^-si, E0 = 1e14 eV,
                                norm_total = norm / flux_FS_E0()  (:205, :666-692) */
  double si;             /* spectral index, source_model 1 only (default 2; si == 1 is rejected: the bin integral of
                            :656 divides by 1 - si) */
  double norm;           /* free-streaming flux at Earth at E0, source_model 1 only (default 1) */
  /* ---- switches of the implementation (since ABI 3; host_copy: ABI 4; all default 0 = the product path).  They select cross-check
   *      kernels and tuning variants that produce the same numbers; tests and tools set them here, nothing in the
   *      library reads the environment. */
  int integrator;        /* 0 = default: role-split wavefront kernel, one CTA per point (32 <= N_bins_E <= 576) or one point over a
                                thread-block cluster of 2 / 4 CTAs (577 <= N_bins_E <= 993 / 2017); diagonal-owned wavefront kernel otherwise;
                            1 = bin-owned wavefront kernel (cross-check);
                            2 = serial, reference-shaped bin-by-bin kernel with true divisions (cross-check, N_bins_E <= 256);
                            3 = diagonal-owned wavefront kernel at every binning;
                            4 = the cluster kernel wherever it applies (N_bins_E > 224), else 3 */
  int fine_variant;      /* 0 = 512 threads per point for N_bins_E > 256; 1 = the 256-thread configurations (same bits) */
  int builder_per_entry; /* 1 = no tiling: one thread per table entry (cross-check of the tiled builders) */
  int group_max;         /* most points per scan-builder group; 0 = 32 */
  int group_always;      /* 1 = use the scan builder however short the groups are */
  int rows_split;        /* CTAs that share one group in the row kernels; 0 = chosen from the launch size */
  int rows_ungrouped;    /* 1 = per-point row kernels even on a scan */
  int keep_order;        /* 1 = do not reorder a batch: by default the points are sorted by (mphi, mntot) before the
                                groups are cut, so the scan builder's gain does not depend on the caller's order;
                                results always come back in the caller's order */
  int two_stream;        /* 1 = builder and rows + integrator on two streams: the integrator of a slice runs beside the builder
                                of the next one (r02: 20.31 instead of 20.54 ms per 8192 points, same bits; not the default:
                                the per-kernel event times then include each other -- the builder reads 16.7 instead of 14.4 ms --
                                and no longer add up to the step, which is what the bench's per-kernel roofline is built on) */
  int source_device;     /* source_model 0: 0 = Lum(z_i, bin j) computed on the host (same glibc calls as the reference,
                                bit-identical to it) and uploaded; 1 = computed by the CUDA source kernel (device exp / log
                                + Li2 / Li3; differs from the host array by up to 1e-6 in the four lowest bins, where the
                                difference of antiderivatives cancels) */
  int host_copy;         /* nusi_evolve_batch, device -> host copies of the results: 0 = default, slice by slice on a copy stream
                                under the kernels of the next slice (21.9 ms per 8192 points on B200, device-resident 21.1);
                            1 = after the last kernel (22.9 ms).  A page-locked caller buffer (cudaHostAlloc / cudaHostRegister,
                                detected per array) is written by the copy engine directly; a pageable one goes through the
                                context's page-locked staging area and a memcpy (78 MB per 8192 points: +2 ms) */
} nusi_opts;

/* Fills *o with the reference's C++ constructor defaults (nuSIprop.hpp:61-65). */
int nusi_opts_default(nusi_opts *o);

/* Replaces calculate_flux::calculate_flux(...) for everything that does not depend on
 * (mphi, g, mntot): energy and redshift grids (nuSIprop.hpp:113-128), PMNS matrix (:130-163),
 * phi-phi table load (:166-170), and the source term Lum(z_i, bin j) (:638-662), which is
 * evaluated once by a CUDA kernel and kept in HBM. */
int nusi_ctx_create(nusi_ctx **out, const nusi_opts *o);
void nusi_ctx_destroy(nusi_ctx *c);

/* get_N_bins_E() (nuSIprop.hpp:407); N_steps_z and the effective zmax = z[N_steps_z-1]
 * (nuSIprop.hpp:124-128); NT = N_bins_E + N_steps_z - 2 rows of the cross-section tables
 * (nuSIprop.hpp:218-221).  Any pointer may be NULL. */
int nusi_ctx_query(const nusi_ctx *c, int *N_bins_E, int *N_steps_z, int *NT, double *zmax_eff);
/* get_energy(i) for all i (nuSIprop.hpp:412-429) plus the bin edges; each [N_bins_E] or NULL. */
int nusi_ctx_energies(const nusi_ctx *c, double *Emin, double *Ecen, double *Emax);
/* z[0..N_steps_z-1] (nuSIprop.hpp:125-127). */
int nusi_ctx_redshifts(const nusi_ctx *c, double *z);
/* |U_ab|^2, [3][3] row-major, a = flavour, b = mass (nuSIprop.hpp:155-163, used as std::norm). */
int nusi_ctx_pmns2(const nusi_ctx *c, double *U2);
/* The source term resident in HBM: out[(i-1)*N + j] = Lum(z[i], Emin[j], Emax[j]), i = 1..N_steps_z-1
 * (nuSIprop.hpp:659-662; per mass eigenstate). */
int nusi_ctx_source(const nusi_ctx *c, double *out);
/* source_model 1 only: replace the power-law source of the context by the one for (si, norm) -- the public members
 * calculate_flux re-reads at each evolve() (nuSIprop.hpp:174, :205) -- without creating a new context. */
int nusi_ctx_set_powerlaw(nusi_ctx *c, double si, double norm);

/* The same grids WITHOUT a context or a GPU (host arithmetic only): what the constructor of
 * calculate_flux computes at nuSIprop.hpp:113-128.  Emin/Ecen/Emax are [N_bins_E] or NULL; the
 * drop-in class uses this so that get_energy() works before the first evolve(). */
int nusi_grid(int N_bins_E, double lEmin, double lEmax, double zmax, int *N_steps_z, double *zmax_eff,
              double *Emin, double *Ecen, double *Emax);

/* energy_FS() of the reference (nuSIprop.hpp:694-744 with flux_FS_E0 :666-692 and Lum_times_E
 * :731-744): total energy of the free-streaming POWER-LAW flux, the quantity
 * check_energy_conservation() (:339-357) compares the evolved flux with.  Host arithmetic
 * (300 pow() calls); zmax_eff is the context's effective zmax (z[N_steps_z-1]). */
int nusi_freestream_energy(double zmax_eff, double lEmin, double lEmax, double si, double norm, double *E_FS);

/* Free-streaming energy of the context's OWN source, the counterpart of energy_FS() (nuSIprop.hpp:694-729) for it:
 *     E_FS = \int_0^zmax dz/H(z) \int_Emin^Emax dE E sum_i L^i(z, E (1+z))
 * source_model 0 (this fork's diffuse supernova background, L^i = R_SN(z) dN/dE(E (1+z)), :607-662): the z integral by the
 * reference's 100 x 3 Gauss-Legendre panels, the E integral by Gauss-Legendre panels in ln E (the reference has no
 * closed form for it: it compares the supernova-sourced flux with the POWER-LAW energy, SURVEY 0.2, so its
 * check_energy_conservation() returns -1 in the fork's default mode).  source_model 1: nusi_freestream_energy() for the
 * context's (si, norm).  Host arithmetic.  With this E_FS, (E_int - E_FS)/E_FS is what check_energy_conservation() is
 * meant to be: the discretisation error in the free-streaming limit, the energy the interactions fail to conserve otherwise. */
int nusi_ctx_freestream_energy(nusi_ctx *c, double *E_FS);
/* E_int of check_energy_conservation() (nuSIprop.hpp:351-355) for n points whose mass-basis fluxes are in HBM
 * (d_flux_mass [n][3][N], as nusi_evolve_points leaves them): a device reduction, one double per point into the
 * HOST array E_int[n]. */
int nusi_energy_integral(nusi_ctx *c, int n, const double *d_flux_mass, double *E_int, void *stream);
/* nusi_energy_integral() that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same reduction, same bits; the per-point results are staged in a file-scope
 * `static __device__` array of NUSI_STATIC_MAX_POINTS doubles (512 KiB, held by every GPU that loads the library until
 * exit) instead of a buffer the context grows with cudaMalloc.  n > NUSI_STATIC_MAX_POINTS returns NUSI_E_INVALID
 * (nusi_last_error() names the limit).  The array is shared by every call of every context and keeps the last call's
 * contents (each call overwrites the n slots it reads back; nothing has to start at zero): calls that overlap on the
 * GPU -- different streams or threads -- race on it.  Calls no cudaMalloc / cudaFree / allocating library routine. */
#define NUSI_STATIC_MAX_POINTS 65536
int nusi_energy_integral_static(nusi_ctx *c, int n, const double *d_flux_mass, double *E_int, void *stream);

/* Bin-integrated power-law source of nusi_opts.source_model == 1 -- the upstream Lum() that the reference keeps as a
 * comment (nuSIprop.hpp:648-657) with norm_total = norm / flux_FS_E0() (:205, :666-692) -- for the grids of
 * nusi_grid(): out[(N_steps_z-1)][N_bins_E], row i at redshift z[i+1] (where the z-loop evaluates it, :283), per mass
 * eigenstate.  Host arithmetic, no GPU (what nusi_ctx_create uploads for source_model 1; test access). */
int nusi_powerlaw_source(int N_bins_E, double lEmin, double lEmax, double zmax, double si, double norm, double *out);

/* Neutrino masses for one mntot: nuSIaux::getmL + nuSIprop.hpp:184-203.  Host arithmetic.
 * Returns NUSI_OK and mn[3], or NUSI_E_INVALID when the reference would exit(1). */
int nusi_masses(double mntot, int normal_ordering, double mn[3]);

/* evolve() for a batch of points (reference nuSIprop.hpp:176-337, called once per object).
 * HOST buffers in, HOST buffers out; host<->device copies are inside the call.
 *   mphi, g, mntot   [n]         the public members calculate_flux re-reads at each evolve() (:174)
 *   flux_mass        [n][3][N]   get_flux(k, j)      (:359-381)   may be NULL
 *   flux_fla         [n][3][N]   get_flux_fla(f, j)  (:383-405)   may be NULL
 *   status           [n]         NUSI_ST_* bits, may be NULL
 * si and norm do not enter: in this fork they only feed check_energy_conservation (:205). */
int nusi_evolve_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                      double *flux_mass, double *flux_fla, int *status);

/* Device-resident variant: prepare once (host mass solve + upload), evolve many times with
 * outputs left in HBM.  d_flux_mass / d_flux_fla / d_status are DEVICE pointers with the shapes
 * above (either flux pointer may be NULL); stream is a cudaStream_t passed as void* (NULL =
 * default stream).  The call only enqueues work. */
int nusi_points_create(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                       nusi_points **out);
void nusi_points_destroy(nusi_points *p);
int nusi_points_count(const nusi_points *p);
int nusi_evolve_points(nusi_ctx *c, const nusi_points *p, double *d_flux_mass, double *d_flux_fla,
                       int *d_status, void *stream);

/* The three cross-section tables of one evolve() (nuSIprop.hpp:218-253), computed by the CUDA
 * builder, copied to HOST buffers.  Gamma, alphaTilde: [n][NT].  alpha: [n][NT][NT] dense
 * row-major, entries m > i filled, others 0 (resonant-only contexts fill only m = i+1, the
 * only entries that mode reads, nuSIprop.hpp:275).  Any pointer may be NULL. */
int nusi_tables_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                      double *Gamma, double *alphaTilde, double *alpha, int *status);

/* Generator of the two phi-phi cross-section tables in the reference's binary format (records of
 * float32 (x_0, .., x_{D-1}, f), last coordinate fastest: interp.hpp:259-290, xsec/text_to_binary.cpp).
 * Replaces xsec/funcs.c:12-39 (integrand) + xsec/tables_phiphi.py:21-59 (scipy dblquad per node, days
 * on a CPU) by fixed Gauss-Legendre quadrature in a CUDA kernel (seconds).  Node grids as the script:
 *   alphatilde: |tbar_plus| = geomspace(4, 1e4, n_t)  x  log10 delta = linspace(0.005, 0.05, n_d)
 *   alpha:      sbar_plus   = geomspace(4, 1e4, n_s)  x  n = 1..n_n  x  log10 delta (same)
 * Pass 0 for a count to get the reference's sizes (n_t 5000, n_s 1000, n_n 1000, n_d 100: the only
 * sizes the reference's own constructor can load, nuSIprop.hpp:168-169; alpha file = 1.6 GB).  Either
 * path may be NULL to skip that table.  seconds (may be NULL) receives the CUDA-event time of the kernels. */
int nusi_phiphi_tables_generate(int device, const char *alphatilde_path, const char *alpha_path, int n_t, int n_s,
                                int n_n, int n_d, double *seconds);

/* How the last batch prepared on this context (nusi_points_create / nusi_evolve_batch) was laid out for the scan
 * builder: number of groups, points in groups of more than one point, and whether the batch had to be reordered. */
int nusi_last_grouping(const nusi_ctx *c, int *n_groups, int *n_grouped_points, int *reordered);

/* CUDA-event time of the evolve calls on this context since the previous nusi_last_timing (at most the last 1024
 * pipeline slices), per kernel, summed over the launches [ms]: Gamma/alphaTilde rows, alpha table, z-stepping
 * integrator; and how many kernels were launched.  Blocks until that work has finished. */
int nusi_last_timing(const nusi_ctx *c, float *ms_rows, float *ms_alpha, float *ms_integrate, int *n_launches);

/* Device special functions evaluated on the GPU for n arguments (unit tests of the
 * dilogarithms that replace gsl_sf_dilog / gsl_sf_complex_dilog_xy_e / polylogarithm::Li2,Li3).
 * which: 0 Re Li2(x) ; 1 complex Li2(x + i y) -> (re, im) ; 2 Li2 on [-1,0] ; 3 Li3 on [-1,0]. */
int nusi_specfun_eval(int device, int which, int n, const double *x, const double *y, double *re, double *im);

/* Sustained FP64 FMA rate of the device [TFLOP/s] from a dependent-free DFMA loop (roofline
 * denominator; MEASURED_PEAKS.json has no FP64 figure). */
int nusi_measure_fp64_peak(int device, double *tflops);

const char *nusi_strerror(int code);
const char *nusi_last_error(void); /* text of the last CUDA/IO failure on this thread */
int nusi_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* NUSIPROP_B200_H */
