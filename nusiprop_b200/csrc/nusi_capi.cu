// nusi_capi.cu -- implementation of include/nusiprop_b200.h: contexts, batches of points, and the evolve pipeline
//     rows kernels + edges + alpha table builder  ->  z-stepping integrator (epilogue fused).
// The kernels live in k_*.cu and are reached through nusi_launch.h.  There is deliberately no host compute path
// in this library; nothing here reads the environment (every switch is a field of nusi_opts).
#include "../../include/nusiprop_b200.h"
#include "nusi_launch.h"
#include "nusi_host.hpp"

#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <array>
#include <mutex>
#include <numeric>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

using namespace nusi;

static thread_local std::string g_last_error;
static int fail(int code, const std::string &msg) {
  g_last_error = msg;
  return code;
}
#define CU(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(e_ == cudaErrorMemoryAllocation ? NUSI_E_NOMEM : NUSI_E_CUDA,                    \
                  std::string(#call) + ": " + cudaGetErrorString(e_));                             \
  } while (0)

template <typename T>
static cudaError_t upload(const std::vector<T> &h, T **d) {
  cudaError_t e = cudaMalloc((void **)d, h.size() * sizeof(T) + 8);
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
}

#ifndef NUSI_SLICE
#define NUSI_SLICE 2048  // points per slice of the evolve pipeline (run_pipeline)
#endif
#define NUSI_GROUP_CAP 256       // most points of one scan-builder group
#define NUSI_MASS_CACHE_MAX 4096  // distinct mass sums memoised per context

struct nusi_ctx {
  std::mutex mu;  // every entry point that uses the context holds it
  nusi_opts opt;
  std::string path_at, path_a;  // own copies of the table paths (opt's pointers belong to the caller)
  HostGrid H;
  DeviceGrid G;
  // device arrays owned by the context
  double *d_row_lo = nullptr, *d_row_hi = nullptr, *d_dE = nullptr, *d_cz = nullptr, *d_ndz = nullptr;
  double *d_source = nullptr, *d_Emin = nullptr, *d_Emax = nullptr, *d_zsrc = nullptr, *d_rsn = nullptr, *d_wE2 = nullptr;
  double *d_pp[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  double *d_edgeE = nullptr;  // distinct bin-edge energies of the table cells (tiled builders)
  int2 *d_tiles = nullptr;
  int n_tiles = 0;
  unsigned char *d_pool = nullptr;  // points block of nusi_evolve_batch, kept across calls
  size_t pool_cap = 0;
  int2 *d_unit_groups = nullptr;  // (0, L) for L = 0..NUSI_GROUP_CAP: one-group launches of the scan builder
  int4 *d_ctiles = nullptr;       // compact builder: per tile the cell ranges [a0, a1) x [b0, b1) of its row / column segment
  int n_ctiles = 0;
  int n_edges = 0;
  bool use_tiles = false;
  // per-launch work space (grown on demand): slot i of rows / tri / edges / work / stat belongs to point i of a chunk
  int cap_points = 0;
  size_t tstride = 0;
  double *d_rows = nullptr, *d_tri = nullptr, *d_work = nullptr, *d_edges = nullptr, *d_rowsh = nullptr;
  int rowsh_groups = 0;  // groups of one launch that d_rowsh holds (a quarter of the slots: launches with more, shorter groups use the one-kernel rows)
  int max_chunk = 0;
  // host-buffer path (nusi_evolve_batch): device staging of the outputs, page-locked host staging, streams
  double *d_out_fm = nullptr, *d_out_ff = nullptr, *d_eint = nullptr;
  int *d_out_st = nullptr;
  int out_cap = 0, eint_cap = 0;
  unsigned char *h_stage = nullptr;  // page-locked
  size_t h_stage_cap = 0;
  cudaStream_t s_main = nullptr, s_x = nullptr, s_i = nullptr, s_copy = nullptr;  // pipeline, builder, rows+integrator, D2H
  cudaEvent_t ev_fork = nullptr, ev_jx = nullptr, ev_ji = nullptr, ev_jc = nullptr;
  std::vector<cudaEvent_t> ev_alpha, ev_done, ev_copied;  // per slice: table ready / fluxes ready / fluxes on the host
  // timing of the last evolve call
  std::vector<cudaEvent_t> ev;  // per slice 6: rows begin/end, builder begin/end, integrator begin/end
  int ev_used = 0;
  int n_launches = 0;
  int last_groups = 0, last_grouped_points = 0, last_reordered = 0;
  std::unordered_map<uint64_t, std::pair<bool, std::array<double, 3>>> mass_cache;
};

struct nusi_points {
  int n = 0;
  int device = 0;
  // Points in EVOLVE order: sorted by (mphi, mntot) unless opts.keep_order, so that the points of a scan row are
  // neighbours whatever order the caller used.  out_index[i] = the caller's index of evolve-order point i.
  PointParams *d_params = nullptr;
  int *d_status0 = nullptr;   // host-side status (NO_MASS, BAD_PARAM), evolve order
  int *d_statw = nullptr;     // working copy the builder kernels OR their warning bits into
  int *d_out_index = nullptr; // NULL when the evolve order is the caller's order
  bool reordered = false;
  // scan builder: runs of consecutive points with identical (mphi, m_nu), cut at slice boundaries and at
  // group_max points; (first point relative to its slice, length)
  int2 *d_groups = nullptr;
  unsigned char *d_block = nullptr;  // the one device allocation behind the arrays above
  bool pooled = false;               // d_block belongs to the context (nusi_evolve_batch)
  std::vector<int2> h_groups;  // host copy with ABSOLUTE first point (evolve order)
  std::vector<int> h_out_index;
  std::vector<int> slice_g0;  // slice k owns groups [slice_g0[k], slice_g0[k+1])
  int SL = 0;                 // the slice length the groups were cut for
};

extern "C" {

int nusi_abi_version(void) { return NUSI_ABI_VERSION; }
const char *nusi_last_error(void) { return g_last_error.c_str(); }
const char *nusi_strerror(int code) {
  switch (code) {
    case NUSI_OK: return "ok";
    case NUSI_E_INVALID: return "invalid argument";
    case NUSI_E_NO_DEVICE: return "no usable CUDA device (this library has no CPU path)";
    case NUSI_E_CUDA: return "CUDA runtime error";
    case NUSI_E_TABLE: return "phi-phi table file missing or malformed";
    case NUSI_E_NOMEM: return "out of memory";
    default: return "unknown error";
  }
}

int nusi_opts_default(nusi_opts *o) {
  if (!o) return NUSI_E_INVALID;
  std::memset(o, 0, sizeof(*o));
  o->majorana = 1;
  o->non_resonant = 1;
  o->normal_ordering = 1;
  o->N_bins_E = 300;
  o->lEmin = 12.0;
  o->lEmax = 17.0;
  o->zmax = 5.0;
  o->flav = 2;
  o->phiphi = 0;
  o->device = 0;
  o->elementary_mode = 3;
  o->source_model = 0;
  o->si = 2.0;
  o->norm = 1.0;
  return NUSI_OK;
}

int nusi_masses(double mntot, int normal_ordering, double mn[3]) {
  if (!mn || !std::isfinite(mntot)) return NUSI_E_INVALID;
  return solve_masses(mntot, normal_ordering != 0, mn) ? NUSI_OK : NUSI_E_INVALID;
}

static void free_workspace(nusi_ctx *c) {
  for (double **p : {&c->d_rows, &c->d_tri, &c->d_work, &c->d_edges, &c->d_rowsh}) {
    if (*p) cudaFree(*p);
    *p = nullptr;
  }
  c->cap_points = 0;
  c->rowsh_groups = 0;
}

void nusi_ctx_destroy(nusi_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->opt.device);
  cudaDeviceSynchronize();
  free_workspace(c);
  double *ptrs[] = {c->d_row_lo, c->d_row_hi, c->d_dE, c->d_cz, c->d_ndz, c->d_source, c->d_Emin, c->d_Emax, c->d_zsrc,
                    c->d_rsn, c->d_wE2, c->d_pp[0], c->d_pp[1], c->d_pp[2], c->d_pp[3], c->d_pp[4], c->d_edgeE,
                    c->d_out_fm, c->d_out_ff, c->d_eint};
  for (double *p : ptrs)
    if (p) cudaFree(p);
  if (c->d_out_st) cudaFree(c->d_out_st);
  if (c->h_stage) cudaFreeHost(c->h_stage);
  for (cudaStream_t s : {c->s_main, c->s_x, c->s_i, c->s_copy})
    if (s) cudaStreamDestroy(s);
  for (auto *v : {&c->ev_done, &c->ev_alpha, &c->ev_copied, &c->ev})
    for (cudaEvent_t e : *v) cudaEventDestroy(e);
  for (cudaEvent_t e : {c->ev_fork, c->ev_jx, c->ev_ji, c->ev_jc})
    if (e) cudaEventDestroy(e);
  if (c->d_tiles) cudaFree(c->d_tiles);
  if (c->d_ctiles) cudaFree(c->d_ctiles);
  if (c->d_unit_groups) cudaFree(c->d_unit_groups);
  if (c->d_pool) cudaFree(c->d_pool);
  delete c;
}

// source term of the context (nuSIprop.hpp:638-662 or the power law of :648-657) -> c->d_source
static int upload_source(nusi_ctx *c) {
  const HostGrid &H = c->H;
  const int S = H.Nz - 1;
  if (!c->d_source) CU(cudaMalloc((void **)&c->d_source, (size_t)S * H.N * sizeof(double)));
  if (c->opt.source_model == 1 || !c->opt.source_device) {
    std::vector<double> src((size_t)S * H.N);
    if (c->opt.source_model == 1)
      powerlaw_source(H.N, S, H.zsrc.data(), H.Emin.data(), H.Emax.data(), H.zmax_eff, c->opt.si, c->opt.norm, src.data());
    else  // the reference's own arithmetic with the same glibc calls: bit-identical to its Lum()
      dsnb_source(H.N, S, H.zsrc.data(), H.rsn.data(), H.lum_pref, H.Emin.data(), H.Emax.data(), src.data());
    CU(cudaMemcpy(c->d_source, src.data(), src.size() * sizeof(double), cudaMemcpyHostToDevice));
  } else {
    CU(launch_source(H.N, S, c->d_zsrc, c->d_rsn, H.lum_pref, c->d_Emin, c->d_Emax, c->d_source, nullptr));
    CU(cudaDeviceSynchronize());
  }
  return NUSI_OK;
}

int nusi_ctx_create(nusi_ctx **out, const nusi_opts *o) {
  if (!out || !o) return NUSI_E_INVALID;
  *out = nullptr;
  if (o->N_bins_E < 2 || !(o->lEmax > o->lEmin) || !(o->zmax > 0) || o->flav < 0 || o->flav > 2)
    return fail(NUSI_E_INVALID, "bad binning / flavour options");
  if (o->source_model < 0 || o->source_model > 1 ||
      (o->source_model == 1 && (!(std::fabs(1 - o->si) > 1e-12) || !std::isfinite(o->si) || !std::isfinite(o->norm))))
    return fail(NUSI_E_INVALID, "bad source options (source_model 0 or 1; si != 1 and finite; norm finite)");
  if (o->integrator < 0 || o->integrator > 4 || o->group_max < 0 || o->rows_split < 0 || o->elementary_mode < 0)
    return fail(NUSI_E_INVALID, "bad implementation switches");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || o->device < 0 || o->device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device with the requested ordinal");
  CU(cudaSetDevice(o->device));

  nusi_ctx *c = new nusi_ctx;
  c->opt = *o;
  if (o->alphatilde_phiphi_path) c->path_at = o->alphatilde_phiphi_path;
  if (o->alpha_phiphi_path) c->path_a = o->alpha_phiphi_path;
  c->opt.alphatilde_phiphi_path = c->opt.alpha_phiphi_path = nullptr;
  build_grid(o->N_bins_E, o->lEmin, o->lEmax, o->zmax, o->normal_ordering != 0, c->H);
  const HostGrid &H = c->H;
  if (H.Nz < 2) {
    delete c;
    return fail(NUSI_E_INVALID, "zmax too small for this binning (fewer than 2 redshift steps)");
  }
  // the integrators' limits are known now: say so here, not at the first evolve()
  if (o->non_resonant && (!integrate_diag_supports(H.N, H.NT) || (o->integrator == 2 && H.N > 256))) {
    delete c;
    return fail(NUSI_E_INVALID, o->integrator == 2 ? "the serial cross-check integrator supports N_bins_E <= 256"
                                                   : "binning not supported by the integrator (N_bins_E <= 2041, N_bins_E + N_steps_z - 2 <= 3072)");
  }
#define UP(vec, ptr)                                                            \
  do {                                                                          \
    cudaError_t e_ = upload(vec, &ptr);                                         \
    if (e_ != cudaSuccess) {                                                    \
      nusi_ctx_destroy(c);                                                      \
      return fail(NUSI_E_CUDA, std::string("upload: ") + cudaGetErrorString(e_)); \
    }                                                                           \
  } while (0)
  UP(H.row_lo, c->d_row_lo);
  UP(H.row_hi, c->d_row_hi);
  UP(H.dE, c->d_dE);
  UP(H.cz, c->d_cz);
  UP(H.ndz, c->d_ndz);
  UP(H.Emin, c->d_Emin);
  UP(H.Emax, c->d_Emax);
  UP(H.zsrc, c->d_zsrc);
  UP(H.rsn, c->d_rsn);
  {
    std::vector<double> w(H.N);  // (ln Emax - ln Emin) E^2 of check_energy_conservation (nuSIprop.hpp:354)
    for (int i = 0; i < H.N; ++i) w[i] = (std::log(H.Emax[i]) - std::log(H.Emin[i])) * sqr(H.Ecen[i]);
    UP(w, c->d_wE2);
  }

  DeviceGrid &G = c->G;
  std::memset(&G, 0, sizeof(G));
  G.N = H.N;
  G.Nz = H.Nz;
  G.NT = H.NT;
  G.majorana = o->majorana != 0;
  G.non_resonant = o->non_resonant != 0;
  G.phiphi = (o->phiphi != 0) && (o->non_resonant != 0);  // tables are loaded only then (nuSIprop.hpp:166)
  G.flav = o->flav;
  G.row_lo = c->d_row_lo;
  G.row_hi = c->d_row_hi;
  G.dE = c->d_dE;
  G.cz = c->d_cz;
  G.ndz = c->d_ndz;
  for (int i = 0; i < 9; ++i) G.U2[i] = H.U2[i];

  if (G.phiphi) {
    HostPhiPhi T;
    const char *p_at = !c->path_at.empty() ? c->path_at.c_str() : "xsec/alphatilde_phiphi.bin";
    const char *p_a = !c->path_a.empty() ? c->path_a.c_str() : "xsec/alpha_phiphi.bin";
    const std::string err = load_phiphi(p_at, p_a, std::log10(H.Emax[0] / H.Emin[0]), T);
    if (!err.empty()) {
      nusi_ctx_destroy(c);
      return fail(NUSI_E_TABLE, err);
    }
    UP(T.at_x0, c->d_pp[0]);
    UP(T.at_f, c->d_pp[1]);
    UP(T.a_x0, c->d_pp[2]);
    UP(T.a_x1, c->d_pp[3]);
    UP(T.a_f, c->d_pp[4]);
    G.pp.at_n0 = (int)T.at_x0.size();
    G.pp.at_x0 = c->d_pp[0];
    G.pp.at_f = c->d_pp[1];
    G.pp.a_n0 = (int)T.a_x0.size();
    G.pp.a_n1 = (int)T.a_x1.size();
    G.pp.a_x0 = c->d_pp[2];
    G.pp.a_x1 = c->d_pp[3];
    G.pp.a_f = c->d_pp[4];
    G.pp.delta_in_range = T.delta_in_range;
  }
#undef UP

  // source term, once per context
  if (int rc = upload_source(c)) {
    nusi_ctx_destroy(c);
    return rc;
  }
  G.source = c->d_source;

  // tiled builders (Majorana + non-resonant): edge list with shared main-grid edges
  if (G.non_resonant && G.majorana && !o->builder_per_entry) {
    bool shared_ok = true;
    for (int a = 0; a + 1 < H.N; ++a) shared_ok = shared_ok && (H.row_hi[a] == H.row_lo[a + 1]);
    if (shared_ok) {
      std::vector<double> edge((size_t)H.N + 1 + 2 * (H.NT - H.N));
      for (int a = 0; a < H.N; ++a) edge[a] = H.row_lo[a];
      edge[H.N] = H.row_hi[H.N - 1];
      for (int a = H.N; a < H.NT; ++a) {
        edge[H.N + 1 + 2 * (a - H.N)] = H.row_lo[a];
        edge[H.N + 2 + 2 * (a - H.N)] = H.row_hi[a];
      }
      std::vector<int2> tiles;
      const int nt = (H.NT + NUSI_TS - 1) / NUSI_TS;
      for (int tb = 0; tb < nt; ++tb)
        for (int ta = 0; ta <= tb; ++ta) tiles.push_back(make_int2(ta, tb));
      // compact builder: segments of <= 15 cells and <= 16 edges
      std::vector<int> seg(1, 0);
      for (int s0 = 0; s0 < H.NT;) {
        int s1 = s0;
        while (s1 < H.NT && (s1 - s0) < NUSI_CE - 1 && edge_hi_id(s1, H.N) - edge_lo_id(s0, H.N) + 1 <= NUSI_CE) ++s1;
        seg.push_back(s1);
        s0 = s1;
      }
      std::vector<int4> ctiles;
      for (int sb = 0; sb + 1 < (int)seg.size(); ++sb)
        for (int sa = 0; sa <= sb; ++sa) ctiles.push_back(make_int4(seg[sa], seg[sa + 1], seg[sb], seg[sb + 1]));
      c->n_ctiles = (int)ctiles.size();
      cudaError_t e1 = upload(edge, &c->d_edgeE);
      if (e1 == cudaSuccess) e1 = upload(tiles, &c->d_tiles);
      if (e1 == cudaSuccess) e1 = upload(ctiles, &c->d_ctiles);
      std::vector<int2> unit;
      for (int L = 0; L <= NUSI_GROUP_CAP; ++L) unit.push_back(make_int2(0, L));
      if (e1 == cudaSuccess) e1 = upload(unit, &c->d_unit_groups);
      if (e1 == cudaSuccess) e1 = alpha_setup_attributes();
      if (e1 == cudaSuccess) e1 = alpha_check_setup_attributes();
      if (e1 != cudaSuccess) {
        nusi_ctx_destroy(c);
        return fail(NUSI_E_CUDA, std::string("tile setup: ") + cudaGetErrorString(e1));
      }
      c->n_tiles = (int)tiles.size();
      c->n_edges = (int)edge.size();
      c->use_tiles = true;
    }
  }
  c->tstride = G.non_resonant ? (size_t)H.NT * (H.NT - 1) / 2 : (size_t)H.NT;
  // chunk size: keep the packed tables of one launch within ~1/4 of free HBM, at most 8192 points
  size_t free_b = 0, total_b = 0;
  cudaMemGetInfo(&free_b, &total_b);
  const size_t per_point = (c->tstride + 2 * (size_t)H.NT + 3 * (size_t)H.N + edge_block_doubles(c->n_edges, H.NT)) * sizeof(double);
  size_t chunk = (free_b / 4) / per_point;
  if (chunk > 8192) chunk = 8192;
  if (chunk < 1) chunk = 1;
  if (o->max_points_per_launch > 0 && (size_t)o->max_points_per_launch < chunk) chunk = o->max_points_per_launch;
  c->max_chunk = (int)chunk;
  *out = c;
  return NUSI_OK;
}

int nusi_ctx_query(const nusi_ctx *c, int *N, int *Nz, int *NT, double *zmax_eff) {
  if (!c) return NUSI_E_INVALID;
  if (N) *N = c->H.N;
  if (Nz) *Nz = c->H.Nz;
  if (NT) *NT = c->H.NT;
  if (zmax_eff) *zmax_eff = c->H.zmax_eff;
  return NUSI_OK;
}
int nusi_ctx_energies(const nusi_ctx *c, double *Emin, double *Ecen, double *Emax) {
  if (!c) return NUSI_E_INVALID;
  const size_t b = c->H.N * sizeof(double);
  if (Emin) std::memcpy(Emin, c->H.Emin.data(), b);
  if (Ecen) std::memcpy(Ecen, c->H.Ecen.data(), b);
  if (Emax) std::memcpy(Emax, c->H.Emax.data(), b);
  return NUSI_OK;
}
int nusi_ctx_redshifts(const nusi_ctx *c, double *z) {
  if (!c || !z) return NUSI_E_INVALID;
  std::memcpy(z, c->H.z.data(), c->H.Nz * sizeof(double));
  return NUSI_OK;
}
int nusi_ctx_pmns2(const nusi_ctx *c, double *U2) {
  if (!c || !U2) return NUSI_E_INVALID;
  std::memcpy(U2, c->H.U2, 9 * sizeof(double));
  return NUSI_OK;
}
int nusi_ctx_source(const nusi_ctx *cc, double *out) {
  nusi_ctx *c = const_cast<nusi_ctx *>(cc);
  if (!c || !out) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  CU(cudaSetDevice(c->opt.device));
  CU(cudaDeviceSynchronize());
  CU(cudaMemcpy(out, c->d_source, (size_t)(c->H.Nz - 1) * c->H.N * sizeof(double), cudaMemcpyDeviceToHost));
  return NUSI_OK;
}
int nusi_ctx_set_powerlaw(nusi_ctx *c, double si, double norm) {
  if (!c) return NUSI_E_INVALID;
  if (c->opt.source_model != 1) return fail(NUSI_E_INVALID, "nusi_ctx_set_powerlaw: the context was created with source_model 0");
  if (!(std::fabs(1 - si) > 1e-12) || !std::isfinite(si) || !std::isfinite(norm)) return fail(NUSI_E_INVALID, "si != 1 and finite; norm finite");
  std::lock_guard<std::mutex> lk(c->mu);
  if (c->opt.si == si && c->opt.norm == norm) return NUSI_OK;
  CU(cudaSetDevice(c->opt.device));
  CU(cudaDeviceSynchronize());  // nothing in flight reads the old table
  c->opt.si = si;
  c->opt.norm = norm;
  return upload_source(c);
}

int nusi_grid(int N_bins_E, double lEmin, double lEmax, double zmax, int *N_steps_z, double *zmax_eff, double *Emin,
              double *Ecen, double *Emax) {
  if (N_bins_E < 2 || !(lEmax > lEmin) || !(zmax > 0)) return fail(NUSI_E_INVALID, "bad binning options");
  HostGrid H;
  build_grid(N_bins_E, lEmin, lEmax, zmax, true, H);
  if (N_steps_z) *N_steps_z = H.Nz;
  if (zmax_eff) *zmax_eff = H.zmax_eff;
  const size_t b = (size_t)H.N * sizeof(double);
  if (Emin) std::memcpy(Emin, H.Emin.data(), b);
  if (Ecen) std::memcpy(Ecen, H.Ecen.data(), b);
  if (Emax) std::memcpy(Emax, H.Emax.data(), b);
  return NUSI_OK;
}

int nusi_freestream_energy(double zmax_eff, double lEmin, double lEmax, double si, double norm, double *E_FS) {
  if (!E_FS || !(zmax_eff > 0) || !(lEmax > lEmin)) return NUSI_E_INVALID;
  *E_FS = freestream_energy(zmax_eff, lEmin, lEmax, si, norm);
  return NUSI_OK;
}

int nusi_ctx_freestream_energy(nusi_ctx *c, double *E_FS) {
  if (!c || !E_FS) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  if (c->opt.source_model == 1)
    *E_FS = freestream_energy(c->H.zmax_eff, c->opt.lEmin, c->opt.lEmax, c->opt.si, c->opt.norm);
  else
    *E_FS = freestream_energy_dsnb(c->H.zmax_eff, c->opt.lEmin, c->opt.lEmax);
  return NUSI_OK;
}

int nusi_powerlaw_source(int N_bins_E, double lEmin, double lEmax, double zmax, double si, double norm, double *out) {
  if (!out || N_bins_E < 2 || !(lEmax > lEmin) || !(zmax > 0) || !(std::fabs(1 - si) > 1e-12)) return NUSI_E_INVALID;
  HostGrid H;
  build_grid(N_bins_E, lEmin, lEmax, zmax, true, H);
  if (H.Nz < 2) return NUSI_E_INVALID;
  powerlaw_source(H.N, H.Nz - 1, H.zsrc.data(), H.Emin.data(), H.Emax.data(), H.zmax_eff, si, norm, out);
  return NUSI_OK;
}

// host preparation of a batch: masses (memoised per mntot), width, |U_fk|^2
static void prepare_points(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                           std::vector<PointParams> &pp, std::vector<int> &st) {
  pp.resize(n);
  st.assign(n, 0);
  const bool no = c->opt.normal_ordering != 0;
  if (c->mass_cache.size() > NUSI_MASS_CACHE_MAX) c->mass_cache.clear();  // a sampler with continuous mntot must not grow it for ever
  for (int i = 0; i < n; ++i) {
    PointParams &P = pp[i];
    std::memset(&P, 0, sizeof(P));
    P.majorana = c->G.majorana;
    P.non_resonant = c->G.non_resonant;
    P.phiphi = c->G.phiphi;
    P.mphi = -1.0;  // kernels skip the point and write NaN
    if (!std::isfinite(mphi[i]) || !std::isfinite(g[i]) || !(mphi[i] > 0)) {
      st[i] = NUSI_ST_BAD_PARAM;
      continue;
    }
    if (!std::isfinite(mntot[i])) {
      st[i] = NUSI_ST_NO_MASS | NUSI_ST_BAD_PARAM;
      continue;
    }
    uint64_t key;
    std::memcpy(&key, &mntot[i], 8);
    auto it = c->mass_cache.find(key);
    if (it == c->mass_cache.end()) {
      std::array<double, 3> m = {0, 0, 0};
      const bool ok = solve_masses(mntot[i], no, m.data());
      it = c->mass_cache.emplace(key, std::make_pair(ok, m)).first;
    }
    if (!it->second.first) {
      st[i] = NUSI_ST_NO_MASS;
      continue;
    }
    P.mphi = mphi[i];
    P.g = g[i];
    P.Ga = scalar_width(mphi[i], g[i], c->G.majorana != 0);
    for (int k = 0; k < 3; ++k) {
      P.mn[k] = it->second.second[k];
      P.U2f[k] = c->H.U2[c->opt.flav * 3 + k];
    }
  }
}

// Grow the work space to `npts` slots.  The new buffers are allocated first and the old ones released only when
// every allocation has succeeded, so a failure leaves the context usable at its previous capacity.
static int ensure_workspace(nusi_ctx *c, int npts) {
  if (npts <= c->cap_points) return NUSI_OK;
  CU(cudaDeviceSynchronize());  // nothing in flight uses the old buffers
  double *rows = nullptr, *tri = nullptr, *work = nullptr, *edges = nullptr, *rowsh = nullptr;
  const int rowsh_groups = npts / 4 + 64;
  cudaError_t e = cudaMalloc((void **)&rows, (size_t)npts * 2 * c->H.NT * sizeof(double));
  // + 16 bytes: the integrator's 16-byte-granular bulk copies may read one double past the last column
  if (e == cudaSuccess) e = cudaMalloc((void **)&tri, (size_t)npts * c->tstride * sizeof(double) + 16);
  if (e == cudaSuccess && !c->G.non_resonant) e = cudaMalloc((void **)&work, (size_t)npts * 3 * c->H.N * sizeof(double));
  if (e == cudaSuccess && c->use_tiles)
    e = cudaMalloc((void **)&edges, (size_t)npts * edge_block_doubles(c->n_edges, c->H.NT) * sizeof(double));
  if (e == cudaSuccess && c->G.majorana && c->G.non_resonant && !c->G.phiphi)
    e = cudaMalloc((void **)&rowsh, (size_t)rowsh_groups * c->H.NT * 3 * NUSI_ROWSH * sizeof(double));
  if (e != cudaSuccess) {
    for (double *p : {rows, tri, work, edges, rowsh})
      if (p) cudaFree(p);
    cudaGetLastError();
    return fail(NUSI_E_NOMEM, std::string("work space for ") + std::to_string(npts) + " points: " + cudaGetErrorString(e));
  }
  free_workspace(c);
  c->d_rows = rows;
  c->d_tri = tri;
  c->d_work = work;
  c->d_edges = edges;
  c->d_rowsh = rowsh;
  c->rowsh_groups = rowsh ? rowsh_groups : 0;
  c->cap_points = npts;
  return NUSI_OK;
}

static int next_events(nusi_ctx *c, cudaEvent_t **e3) {
  if ((int)c->ev.size() < c->ev_used + 6) {
    for (int i = 0; i < 6; ++i) {
      cudaEvent_t e;
      CU(cudaEventCreate(&e));
      c->ev.push_back(e);
    }
  }
  *e3 = &c->ev[c->ev_used];
  c->ev_used += 6;
  return NUSI_OK;
}

// Builder arguments of one launch: point i of the launch uses slot wo + i of rows / tri / edges.
static BuildArgs build_args(const nusi_ctx *c, int wo, const PointParams *d_params, int cnt, int *d_status, const int2 *d_groups,
                            int ngroups) {
  BuildArgs A;
  A.G = c->G;
  A.nE = c->n_edges;
  A.params = d_params;
  A.cnt = cnt;
  A.status = d_status;
  A.rows = c->d_rows + (size_t)wo * 2 * c->H.NT;
  A.tri = c->d_tri + (size_t)wo * c->tstride;
  A.tstride = c->tstride;
  A.edges = c->d_edges ? c->d_edges + (size_t)wo * edge_block_doubles(c->n_edges, c->H.NT) : nullptr;
  A.edgeE = c->d_edgeE;
  A.tiles = c->d_tiles;
  A.n_tiles = c->n_tiles;
  A.ctiles = c->d_ctiles;
  A.n_ctiles = c->n_ctiles;
  A.groups = d_groups;
  A.ngroups = ngroups;
  A.rowsh = c->d_rowsh;  // one launch at a time per stream uses it (the row kernels of consecutive slices are stream-ordered)
  A.rowsh_groups = c->rowsh_groups;
  return A;
}

// Gamma / alphaTilde rows of one launch
static int launch_rows_stage(nusi_ctx *c, const BuildArgs &A, cudaStream_t s) {
  const DeviceGrid &G = c->G;
  const bool grouped = c->opt.elementary_mode >= 3 && A.groups && A.ngroups > 0 && G.majorana && G.non_resonant && !G.phiphi &&
                       !c->opt.rows_ungrouped;
  int split = 1;
  if (grouped) {
    // scan: the g-independent part of a row once per (part of a) group.  A group is shared by `split` CTAs: 4 when the
    // launch is large (the per-point part dominates, so repeating the shared part four times is cheap and the grid is
    // 4x finer), more for small launches so that the grid still fills the 148 SMs x 4 CTAs twice (each CTA keeps >= 1 point).
    split = 4;
    const long long nblk = (G.NT + NUSI_ROWS_PER_CTA - 1) / NUSI_ROWS_PER_CTA;
    const long long want = 1184, have = nblk * A.ngroups * split;
    if (have < want) split = (int)std::min<long long>(32, (want + nblk * A.ngroups - 1) / (nblk * A.ngroups));
    // two-kernel rows (shared part once per group in its own kernel): nothing is repeated, so the per-point kernel takes two
    // points of a 32-point group per CTA (B200, 8192 points: 4 / 8 / 16 / 32 CTAs per group 2.97 / 2.70 / 2.61 / 2.63 ms)
    if (A.rowsh && A.ngroups <= A.rowsh_groups) split = std::max(split, 16);
    if (c->opt.rows_split > 0) split = c->opt.rows_split;
  }
  CU(launch_rows(A, grouped, split, s));
  c->n_launches += (grouped && A.rowsh && A.ngroups <= A.rowsh_groups) ? 4 : 2;  // shared part + per-point part per function
  return NUSI_OK;
}

// alpha table of one launch
static int launch_alpha_stage(nusi_ctx *c, const BuildArgs &A, cudaStream_t s) {
  const DeviceGrid &G = c->G;
  const int mode = c->opt.elementary_mode;
  if (c->use_tiles) {
    // the scan builder pays off from ~1.3 points per group on (measured on B200, 8192 points: groups of 1: 45.3 ms,
    // of 2: 31.7, of 8: 18.3, of 32: 14.5; compact builder 42.0 ms)
    const bool scan = mode >= 3 && A.groups && A.ngroups > 0 && !G.phiphi && (4 * (long long)A.ngroups <= 3 * (long long)A.cnt || c->opt.group_always);
    if (scan) {
      CU(launch_edges_group(A, s));  // full edge data for the first point of a group, the g-dependent fields for the others
      CU(launch_alpha_group(A, s));  // one CTA per (tile, group of points with the same mphi and neutrino masses)
    } else {
      CU(launch_edges(A, s));
      if (mode >= 2)
        CU(launch_alpha_compact(A, s));
      else
        CU(launch_alpha_tile(A, mode, s));
    }
    c->n_launches += 2;
  } else {
    CU(launch_alpha_entry(A, mode >= 2 ? 1 : 0, s));
    c->n_launches += 1;
  }
  return NUSI_OK;
}

static int launch_integrate_stage(nusi_ctx *c, const IntegArgs &A, cudaStream_t s) {
  cudaError_t e;
  if (!c->G.non_resonant)
    e = launch_integrate_resonant(A, s);
  else if (c->opt.integrator == 2)
    e = launch_integrate_serial(A, s);
  else if (c->opt.integrator == 1)
    e = launch_integrate_wave(A, s);
  else if (c->opt.integrator == 0 && integrate_split_supports(c->G.N, c->G.NT))
    e = launch_integrate_split(A, c->opt.fine_variant, s);
  // beyond the one-CTA role-split kernel's reach (N_bins_E > 576: 20 warps x 96 registers fill the SM): one point over a 4-CTA
  // cluster.  The choice depends on the binning alone, so that a point's bits do not depend on the size of its batch.
  else if ((c->opt.integrator == 0 || c->opt.integrator == 4) && integrate_cluster_supports(c->G.N, c->G.NT)) {
    e = launch_integrate_cluster(A, c->opt.fine_variant, s);
    // a device (or partition) that cannot co-schedule a 4-CTA cluster with this much shared memory: the one-CTA kernel
    if ((e == cudaErrorLaunchOutOfResources || e == cudaErrorInvalidConfiguration || e == cudaErrorInvalidClusterSize || e == cudaErrorInvalidValue) &&
        integrate_diag_supports(c->G.N, c->G.NT)) {
      cudaGetLastError();
      e = launch_integrate_diag(A, c->opt.fine_variant, s);
    }
  }
  else
    e = launch_integrate_diag(A, c->opt.fine_variant, s);
  if (e == cudaErrorInvalidValue) return fail(NUSI_E_INVALID, "binning not supported by the selected integrator");
  CU(e);
  c->n_launches += 1;
  return NUSI_OK;
}

// One device block per points object: [params | groups | status0 | statw | out_index], filled by ONE host->device
// copy.  pooled: the block is the context's own (grown on demand, kept across calls) -- nusi_evolve_batch, which would
// otherwise pay a cudaMalloc + cudaFree (each a device synchronisation) per call.
static int points_create_impl(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                              nusi_points **out, bool pooled) {
  if (!c || !out || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  std::vector<PointParams> pp0, pp;
  std::vector<int> st0;
  nusi_points *p = new nusi_points;
  p->n = n;
  p->device = c->opt.device;
  prepare_points(c, n, mphi, g, mntot, pp0, st0);
  // evolve order: stable sort by (mphi, m_nu) -- the points of a scan row become neighbours whatever order the caller
  // used (g-major grids, shuffled samples); rejected points (mphi = -1) go first
  std::vector<int> order(n);
  std::iota(order.begin(), order.end(), 0);
  auto key_less = [&](int a, int b) {
    const PointParams &A = pp0[a], &B = pp0[b];
    if (A.mphi != B.mphi) return A.mphi < B.mphi;
    if (A.mn[0] != B.mn[0]) return A.mn[0] < B.mn[0];
    return A.mn[2] < B.mn[2];
  };
  if (!c->opt.keep_order && c->opt.elementary_mode >= 3 && !std::is_sorted(order.begin(), order.end(), key_less))
    std::stable_sort(order.begin(), order.end(), key_less);
  p->reordered = false;
  for (int i = 0; i < n; ++i) p->reordered = p->reordered || order[i] != i;
  if (p->reordered) {
    pp.resize(n);
    std::vector<int> st(n);
    for (int i = 0; i < n; ++i) {
      pp[i] = pp0[order[i]];
      st[i] = st0[order[i]];
    }
    st0.swap(st);
    p->h_out_index = order;
  } else {
    pp.swap(pp0);
  }
  std::vector<int2> groups;
  {
    const int gmax = c->opt.group_max > 0 ? std::min(c->opt.group_max, NUSI_GROUP_CAP) : 32;
    p->SL = c->max_chunk < NUSI_SLICE ? c->max_chunk : NUSI_SLICE;
    for (int off = 0; off < n; off += p->SL) {
      p->slice_g0.push_back((int)groups.size());
      const int end = (n - off < p->SL) ? n : off + p->SL;
      for (int i = off; i < end;) {
        int j = i + 1;
        while (j < end && j - i < gmax && std::memcmp(&pp[j].mphi, &pp[i].mphi, sizeof(double)) == 0 &&
               std::memcmp(pp[j].mn, pp[i].mn, sizeof(pp[i].mn)) == 0)
          ++j;
        groups.push_back(make_int2(i - off, j - i));
        p->h_groups.push_back(make_int2(i, j - i));
        i = j;
      }
    }
    p->slice_g0.push_back((int)groups.size());
  }
  c->last_groups = (int)groups.size();
  c->last_grouped_points = 0;
  for (const int2 &gp : groups)
    if (gp.y > 1) c->last_grouped_points += gp.y;
  c->last_reordered = p->reordered ? 1 : 0;

  auto al16 = [](size_t b) { return (b + 15) & ~(size_t)15; };
  const size_t o_groups = al16((size_t)n * sizeof(PointParams));
  const size_t o_status = o_groups + al16(groups.size() * sizeof(int2));
  const size_t o_statw = o_status + al16((size_t)n * sizeof(int));
  const size_t o_index = o_statw + al16((size_t)n * sizeof(int));
  const size_t bytes = o_index + (p->reordered ? al16((size_t)n * sizeof(int)) : 0);
  std::vector<unsigned char> stage(bytes);
  std::memcpy(stage.data(), pp.data(), (size_t)n * sizeof(PointParams));
  std::memcpy(stage.data() + o_groups, groups.data(), groups.size() * sizeof(int2));
  std::memcpy(stage.data() + o_status, st0.data(), (size_t)n * sizeof(int));
  std::memcpy(stage.data() + o_statw, st0.data(), (size_t)n * sizeof(int));
  if (p->reordered) std::memcpy(stage.data() + o_index, order.data(), (size_t)n * sizeof(int));
  cudaError_t e = cudaSuccess;
  unsigned char *base = nullptr;
  if (pooled) {
    if (bytes > c->pool_cap) {
      cudaDeviceSynchronize();
      if (c->d_pool) cudaFree(c->d_pool);
      c->d_pool = nullptr;
      c->pool_cap = 0;
      e = cudaMalloc((void **)&c->d_pool, bytes + bytes / 4);
      if (e == cudaSuccess) c->pool_cap = bytes + bytes / 4;
    }
    base = c->d_pool;
    p->pooled = true;
  } else {
    e = cudaMalloc((void **)&base, bytes);
  }
  if (e == cudaSuccess) {
    p->d_block = base;
    p->d_params = reinterpret_cast<PointParams *>(base);
    p->d_groups = reinterpret_cast<int2 *>(base + o_groups);
    p->d_status0 = reinterpret_cast<int *>(base + o_status);
    p->d_statw = reinterpret_cast<int *>(base + o_statw);
    p->d_out_index = p->reordered ? reinterpret_cast<int *>(base + o_index) : nullptr;
    e = cudaMemcpy(base, stage.data(), bytes, cudaMemcpyHostToDevice);
  }
  if (e != cudaSuccess) {
    nusi_points_destroy(p);
    return fail(NUSI_E_CUDA, std::string("points upload: ") + cudaGetErrorString(e));
  }
  *out = p;
  return NUSI_OK;
}
int nusi_points_create(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                       nusi_points **out) {
  if (!c) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  return points_create_impl(c, n, mphi, g, mntot, out, false);
}
void nusi_points_destroy(nusi_points *p) {
  if (!p) return;
  if (!p->pooled && p->d_block) {
    cudaSetDevice(p->device);
    cudaFree(p->d_block);
  }
  delete p;
}
int nusi_points_count(const nusi_points *p) { return p ? p->n : 0; }
int nusi_last_grouping(const nusi_ctx *c, int *n_groups, int *n_grouped_points, int *reordered) {
  if (!c) return NUSI_E_INVALID;
  if (n_groups) *n_groups = c->last_groups;
  if (n_grouped_points) *n_grouped_points = c->last_grouped_points;
  if (reordered) *reordered = c->last_reordered;
  return NUSI_OK;
}

// ------------------------------------------------------------------------------------------------------
// The evolve pipeline.  A batch is cut into slices of NUSI_SLICE points (evolve order).  Per slice k, in stream order:
//     rows(k)   edges(k) + alpha table(k)   integrate(k)  [-> D2H(k) on the copy stream, under slice k+1's kernels]
// Outputs land in slot out_index[i] of the output arrays (the caller's order).  With host buffers and an unchanged
// order, each slice's fluxes go home on the copy stream as soon as its integrator has finished; with a reordered
// batch a slice's outputs are scattered over the arrays, so they go home in caller-order pieces after the last slice.
// opts.two_stream puts the builder and rows + integrator on two streams (measured neutral on B200: the builder
// already owns every register of an SM, so the integrator's CTAs displace builder CTAs one for one).
// ------------------------------------------------------------------------------------------------------
static int ensure_streams(nusi_ctx *c) {
  if (c->s_main) return NUSI_OK;
  int lo = 0, hi = 0;
  CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CU(cudaStreamCreateWithPriority(&c->s_main, cudaStreamNonBlocking, lo));
  CU(cudaStreamCreateWithPriority(&c->s_x, cudaStreamNonBlocking, lo));
  CU(cudaStreamCreateWithPriority(&c->s_i, cudaStreamNonBlocking, hi));
  CU(cudaStreamCreateWithPriority(&c->s_copy, cudaStreamNonBlocking, hi));
  for (cudaEvent_t *e : {&c->ev_fork, &c->ev_jx, &c->ev_ji, &c->ev_jc}) CU(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  return NUSI_OK;
}
static int slice_event(std::vector<cudaEvent_t> &v, size_t k, cudaEvent_t *out) {
  while (v.size() <= k) {
    cudaEvent_t ev;
    CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    v.push_back(ev);
  }
  *out = v[k];
  return NUSI_OK;
}

// Host destinations of nusi_evolve_batch.  direct: the caller's buffers are page-locked, copies go straight there;
// otherwise through the context's page-locked staging area and a host memcpy per piece (a cudaMemcpyAsync into
// pageable memory would block the host until the slice has finished and serialise the whole pipeline).
struct HostOut {
  double *fm = nullptr, *ff = nullptr;
  int *st = nullptr;
  // per array: the caller's buffer is page-locked (the copy engine writes it directly), or it is pageable and the copy goes
  // through the context's page-locked staging area followed by a memcpy
  bool fm_direct = true, ff_direct = true, st_direct = true;
  double *s_fm = nullptr, *s_ff = nullptr;  // staging (whole batch, same indexing as the destinations)
  int *s_st = nullptr;
  bool any() const { return fm || ff || st; }
  bool staged() const { return (fm && !fm_direct) || (ff && !ff_direct) || (st && !st_direct); }
};

// d_fm / d_ff / d_st: device outputs indexed by the caller's point index.
static int run_pipeline(nusi_ctx *c, const nusi_points *p, double *d_fm, double *d_ff, int *d_st, cudaStream_t caller,
                        const HostOut *ho) {
  const int N = c->H.N;
  const bool to_host = ho && ho->any();
  const bool overlap = c->opt.two_stream != 0;
  int rc = ensure_streams(c);
  if (rc) return rc;
  // slice size SL; the workspace holds a whole number of slices (all of them when the batch fits), or just the batch
  const int SL = c->max_chunk < NUSI_SLICE ? c->max_chunk : NUSI_SLICE;
  const int chunk = p->n <= SL ? p->n : (p->n <= c->max_chunk ? ((p->n + SL - 1) / SL) * SL : std::max(1, c->max_chunk / SL) * SL);
  if ((rc = ensure_workspace(c, chunk))) return rc;
  if (c->ev_used >= 6 * 1024) {  // nobody reads the timings: start over rather than grow for ever
    c->ev_used = 0;
    c->n_launches = 0;
  }
  cudaStream_t sx = overlap ? c->s_x : caller, si = overlap ? c->s_i : caller, sc = to_host ? c->s_copy : caller;
  if (overlap || to_host) {
    CU(cudaEventRecord(c->ev_fork, caller));
    if (overlap) {
      CU(cudaStreamWaitEvent(sx, c->ev_fork, 0));
      CU(cudaStreamWaitEvent(si, c->ev_fork, 0));
    }
  }
  // the builder kernels OR their warning bits into the working copy of the host-side status
  CU(cudaMemcpyAsync(p->d_statw, p->d_status0, (size_t)p->n * sizeof(int), cudaMemcpyDeviceToDevice, si));
  if (overlap) {
    CU(cudaEventRecord(c->ev_ji, si));
    CU(cudaStreamWaitEvent(sx, c->ev_ji, 0));
  }
  const bool per_slice_copy = to_host && !p->reordered && c->opt.host_copy == 0;
  auto copy_piece = [&](int off, int cnt, size_t k) -> int {  // outputs [off, off + cnt) (caller order) -> host
    const size_t o = (size_t)off * 3 * N, b = (size_t)cnt * 3 * N * sizeof(double);
    if (ho->fm) CU(cudaMemcpyAsync((ho->fm_direct ? ho->fm : ho->s_fm) + o, d_fm + o, b, cudaMemcpyDeviceToHost, sc));
    if (ho->ff) CU(cudaMemcpyAsync((ho->ff_direct ? ho->ff : ho->s_ff) + o, d_ff + o, b, cudaMemcpyDeviceToHost, sc));
    if (ho->st) CU(cudaMemcpyAsync((ho->st_direct ? ho->st : ho->s_st) + off, d_st + off, (size_t)cnt * sizeof(int), cudaMemcpyDeviceToHost, sc));
    cudaEvent_t evc;
    int r = slice_event(c->ev_copied, k, &evc);
    if (r) return r;
    CU(cudaEventRecord(evc, sc));
    return NUSI_OK;
  };
  size_t k = 0;
  for (int off = 0; off < p->n; off += SL, ++k) {
    const int cnt = (p->n - off < SL) ? p->n - off : SL;
    const int wo = off % chunk;  // the slices of one chunk use distinct workspace slots
    cudaEvent_t *t6, ev_tab, ev_out;
    if ((rc = next_events(c, &t6))) return rc;
    if ((rc = slice_event(c->ev_alpha, k, &ev_tab))) return rc;
    if (wo == 0 && off > 0 && overlap) {  // a new chunk reuses the workspace: everything queued so far must be done with it
      CU(cudaEventRecord(c->ev_ji, si));
      CU(cudaStreamWaitEvent(sx, c->ev_ji, 0));
    }
    const bool grouped = (p->SL == SL) && p->d_groups && k + 1 < p->slice_g0.size();
    const BuildArgs B = build_args(c, wo, p->d_params + off, cnt, p->d_statw + off, grouped ? p->d_groups + p->slice_g0[k] : nullptr,
                                   grouped ? p->slice_g0[k + 1] - p->slice_g0[k] : 0);
    CU(cudaEventRecord(t6[0], si));
    if ((rc = launch_rows_stage(c, B, si))) return rc;
    CU(cudaEventRecord(t6[1], si));
    CU(cudaEventRecord(t6[2], sx));
    if ((rc = launch_alpha_stage(c, B, sx))) return rc;
    CU(cudaEventRecord(t6[3], sx));
    if (overlap) {
      CU(cudaEventRecord(ev_tab, sx));
      CU(cudaStreamWaitEvent(si, ev_tab, 0));
    }
    CU(cudaEventRecord(t6[4], si));
    IntegArgs I;
    I.G = c->G;
    I.params = p->d_params + off;
    I.cnt = cnt;
    I.rows = B.rows;
    I.tri = B.tri;
    I.tstride = c->tstride;
    I.work = c->d_work ? c->d_work + (size_t)wo * 3 * N : nullptr;
    // identity order: offset the output arrays instead of indexing them
    I.flux_mass = d_fm ? (p->reordered ? d_fm : d_fm + (size_t)off * 3 * N) : nullptr;
    I.flux_fla = d_ff ? (p->reordered ? d_ff : d_ff + (size_t)off * 3 * N) : nullptr;
    I.out_index = p->reordered ? p->d_out_index + off : nullptr;
    I.status_in = p->d_statw + off;
    I.status_out = d_st ? (p->reordered ? d_st : d_st + off) : nullptr;
    if ((rc = launch_integrate_stage(c, I, si))) return rc;
    CU(cudaEventRecord(t6[5], si));
    if (per_slice_copy) {
      if ((rc = slice_event(c->ev_done, k, &ev_out))) return rc;
      CU(cudaEventRecord(ev_out, si));
      CU(cudaStreamWaitEvent(sc, ev_out, 0));
      if ((rc = copy_piece(off, cnt, k))) return rc;
    }
  }
  // join
  if (overlap) {
    CU(cudaEventRecord(c->ev_jx, sx));
    CU(cudaEventRecord(c->ev_ji, si));
    CU(cudaStreamWaitEvent(caller, c->ev_jx, 0));
    CU(cudaStreamWaitEvent(caller, c->ev_ji, 0));
  }
  size_t npieces = per_slice_copy ? k : 0;
  if (to_host && !per_slice_copy) {  // reordered batch: everything is computed; home in caller-order pieces
    CU(cudaEventRecord(c->ev_jx, caller));
    CU(cudaStreamWaitEvent(sc, c->ev_jx, 0));
    for (int off = 0; off < p->n; off += SL, ++npieces)
      if ((rc = copy_piece(off, std::min(SL, p->n - off), npieces))) return rc;
  }
  if (to_host && ho->staged()) {  // staging -> the caller's pageable buffers, piece by piece as the copies complete
    size_t q = 0;
    for (int off = 0; off < p->n; off += SL, ++q) {
      const int cnt = std::min(SL, p->n - off);
      CU(cudaEventSynchronize(c->ev_copied[q]));
      const size_t o = (size_t)off * 3 * N, b = (size_t)cnt * 3 * N * sizeof(double);
      if (ho->fm && !ho->fm_direct) std::memcpy(ho->fm + o, ho->s_fm + o, b);
      if (ho->ff && !ho->ff_direct) std::memcpy(ho->ff + o, ho->s_ff + o, b);
      if (ho->st && !ho->st_direct) std::memcpy(ho->st + off, ho->s_st + off, (size_t)cnt * sizeof(int));
    }
  }
  if (to_host) {
    CU(cudaEventRecord(c->ev_jc, sc));
    CU(cudaStreamWaitEvent(caller, c->ev_jc, 0));
  }
  return NUSI_OK;
}

int nusi_evolve_points(nusi_ctx *c, const nusi_points *p, double *d_fm, double *d_ff, int *d_status, void *stream) {
  if (!c || !p) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  CU(cudaSetDevice(c->opt.device));
  return run_pipeline(c, p, d_fm, d_ff, d_status, (cudaStream_t)stream, nullptr);
}

static bool is_pinned(const void *ptr) {
  if (!ptr) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeHost;
}

// Host buffers in, host buffers out.
int nusi_evolve_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                      double *flux_mass, double *flux_fla, int *status) {
  if (!c || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  CU(cudaSetDevice(c->opt.device));
  const int N = c->H.N;
  if (n > c->out_cap) {
    CU(cudaDeviceSynchronize());
    for (double **q : {&c->d_out_fm, &c->d_out_ff}) {
      if (*q) cudaFree(*q);
      *q = nullptr;
    }
    if (c->d_out_st) cudaFree(c->d_out_st);
    c->d_out_st = nullptr;
    c->out_cap = 0;
    CU(cudaMalloc((void **)&c->d_out_fm, (size_t)n * 3 * N * sizeof(double)));
    CU(cudaMalloc((void **)&c->d_out_ff, (size_t)n * 3 * N * sizeof(double)));
    CU(cudaMalloc((void **)&c->d_out_st, (size_t)n * sizeof(int)));
    c->out_cap = n;
  }
  int rc = ensure_streams(c);
  if (rc) return rc;
  HostOut ho;
  ho.fm = flux_mass;
  ho.ff = flux_fla;
  ho.st = status;
  ho.fm_direct = is_pinned(flux_mass);
  ho.ff_direct = is_pinned(flux_fla);
  ho.st_direct = is_pinned(status);
  if (ho.staged()) {
    const size_t bf = (size_t)n * 3 * N * sizeof(double), need = 2 * bf + (size_t)n * sizeof(int);
    if (need > c->h_stage_cap) {
      if (c->h_stage) cudaFreeHost(c->h_stage);
      c->h_stage = nullptr;
      c->h_stage_cap = 0;
      CU(cudaMallocHost((void **)&c->h_stage, need));
      c->h_stage_cap = need;
    }
    ho.s_fm = reinterpret_cast<double *>(c->h_stage);
    ho.s_ff = reinterpret_cast<double *>(c->h_stage + bf);
    ho.s_st = reinterpret_cast<int *>(c->h_stage + 2 * bf);
  }
  nusi_points *p = nullptr;
  if ((rc = points_create_impl(c, n, mphi, g, mntot, &p, true))) return rc;
  rc = run_pipeline(c, p, flux_mass ? c->d_out_fm : nullptr, flux_fla ? c->d_out_ff : nullptr, c->d_out_st, c->s_main, &ho);
  cudaError_t e2 = cudaDeviceSynchronize();
  nusi_points_destroy(p);
  if (rc) return rc;
  if (e2 != cudaSuccess) return fail(NUSI_E_CUDA, std::string("evolve_batch: ") + cudaGetErrorString(e2));
  return NUSI_OK;
}

int nusi_energy_integral(nusi_ctx *c, int n, const double *d_flux_mass, double *E_int, void *stream) {
  if (!c || n <= 0 || !d_flux_mass || !E_int) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  CU(cudaSetDevice(c->opt.device));
  if (n > c->eint_cap) {
    if (c->d_eint) cudaFree(c->d_eint);
    c->d_eint = nullptr;
    c->eint_cap = 0;
    CU(cudaMalloc((void **)&c->d_eint, (size_t)n * sizeof(double)));
    c->eint_cap = n;
  }
  cudaStream_t s = (cudaStream_t)stream;
  CU(launch_energy_integral(n, c->H.N, d_flux_mass, c->d_wE2, c->d_eint, s));
  CU(cudaMemcpyAsync(E_int, c->d_eint, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  return NUSI_OK;
}

int nusi_tables_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot, double *Gamma,
                      double *alphaTilde, double *alpha, int *status) {
  if (!c || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  CU(cudaSetDevice(c->opt.device));
  const int NT = c->H.NT;
  nusi_points *p = nullptr;
  int rc = points_create_impl(c, n, mphi, g, mntot, &p, false);
  if (rc) return rc;
  c->ev_used = 0;
  c->n_launches = 0;
  std::vector<double> hrows((size_t)2 * NT), htri(c->tstride);
  std::vector<int> hst(n);
  // group by group, through the same builder a batch of these points gets in nusi_evolve_batch
  for (size_t gi = 0; gi < p->h_groups.size() && rc == NUSI_OK; ++gi) {
    const int i0 = p->h_groups[gi].x, len = p->h_groups[gi].y;
    rc = ensure_workspace(c, len);
    if (rc) break;
    const BuildArgs B = build_args(c, 0, p->d_params + i0, len, p->d_statw + i0, c->d_unit_groups ? c->d_unit_groups + len : nullptr, 1);
    rc = launch_rows_stage(c, B, nullptr);
    if (rc == NUSI_OK) rc = launch_alpha_stage(c, B, nullptr);
    if (rc) break;
    for (int i = i0; i < i0 + len; ++i) {
      const int dst = p->reordered ? p->h_out_index[i] : i;  // the caller's index of this point
      cudaError_t e = cudaMemcpy(hrows.data(), c->d_rows + (size_t)(i - i0) * 2 * NT, hrows.size() * sizeof(double), cudaMemcpyDeviceToHost);
      if (e == cudaSuccess)
        e = cudaMemcpy(htri.data(), c->d_tri + (size_t)(i - i0) * c->tstride, htri.size() * sizeof(double), cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) {
        rc = fail(NUSI_E_CUDA, std::string("tables_batch: ") + cudaGetErrorString(e));
        break;
      }
      if (Gamma) std::memcpy(Gamma + (size_t)dst * NT, hrows.data(), NT * sizeof(double));
      if (alphaTilde) std::memcpy(alphaTilde + (size_t)dst * NT, hrows.data() + NT, NT * sizeof(double));
      if (alpha) {
        double *A = alpha + (size_t)dst * NT * NT;
        std::memset(A, 0, (size_t)NT * NT * sizeof(double));
        if (c->G.non_resonant) {
          for (int b = 1; b < NT; ++b)
            for (int a = 0; a < b; ++a) A[(size_t)a * NT + b] = htri[(size_t)b * (b - 1) / 2 + a];
        } else {
          for (int a = 0; a + 1 < NT; ++a) A[(size_t)a * NT + a + 1] = htri[a];
        }
      }
    }
  }
  if (rc == NUSI_OK && status) {
    cudaError_t e = cudaMemcpy(hst.data(), p->d_statw, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = fail(NUSI_E_CUDA, cudaGetErrorString(e));
    for (int i = 0; i < n && rc == NUSI_OK; ++i) status[p->reordered ? p->h_out_index[i] : i] = hst[i];
  }
  nusi_points_destroy(p);
  return rc;
}

int nusi_last_timing(const nusi_ctx *cc, float *ms_rows, float *ms_alpha, float *ms_integrate, int *n_launches) {
  nusi_ctx *c = const_cast<nusi_ctx *>(cc);
  if (!c) return NUSI_E_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  float tr = 0, ta = 0, ti = 0;
  for (int i = 0; i + 5 < c->ev_used && i + 5 < (int)c->ev.size(); i += 6) {
    CU(cudaEventSynchronize(c->ev[i + 5]));
    CU(cudaEventSynchronize(c->ev[i + 3]));
    float a = 0, b = 0, d = 0;
    CU(cudaEventElapsedTime(&a, c->ev[i], c->ev[i + 1]));
    CU(cudaEventElapsedTime(&b, c->ev[i + 2], c->ev[i + 3]));
    CU(cudaEventElapsedTime(&d, c->ev[i + 4], c->ev[i + 5]));
    tr += a;
    ta += b;
    ti += d;
  }
  if (ms_rows) *ms_rows = tr;
  if (ms_alpha) *ms_alpha = ta;
  if (ms_integrate) *ms_integrate = ti;
  if (n_launches) *n_launches = c->n_launches;
  c->ev_used = 0;  // the next call reports what was launched after this one
  c->n_launches = 0;
  return NUSI_OK;
}

// coordinates go through the reference pipeline's text step ("%.7e", tables_phiphi.py:36,59) before
// they become float32 (text_to_binary.cpp:32); the node VALUE is computed at the unrounded double
static float text_rounded(double v) {
  char buf[64];
  std::snprintf(buf, sizeof buf, "%.7e", v);
  return std::strtof(buf, nullptr);
}
static std::vector<double> geomspace(double a, double b, int n) {
  std::vector<double> v(n);
  const double la = std::log10(a), lb = std::log10(b);
  for (int i = 0; i < n; ++i) v[i] = std::pow(10.0, la + (lb - la) * i / (n - 1));
  v[0] = a;
  v[n - 1] = b;
  return v;
}
static std::vector<double> linspace(double a, double b, int n) {
  std::vector<double> v(n);
  for (int i = 0; i < n; ++i) v[i] = a + (b - a) * i / (n - 1);
  v[n - 1] = b;
  return v;
}

// one table: kernel in slabs of <= 2^24 nodes, records interleaved on the host and appended to the file
static int generate_table(int which, const char *path, const std::vector<double> &x0, const std::vector<double> &x1,
                          const std::vector<double> &x2, float *ms_total) {
  const int D = which == 0 ? 2 : 3;
  const long long n0 = (long long)x0.size(), n1 = (long long)x1.size(), n2 = which == 0 ? 1 : (long long)x2.size();
  const long long total = n0 * n1 * n2;
  FILE *fp = std::fopen(path, "wb");
  if (!fp) return fail(NUSI_E_TABLE, std::string("cannot open ") + path + " for writing");
  double *d_x[3] = {nullptr, nullptr, nullptr};
  float *d_out = nullptr;
  const long long slab = 1LL << 24;
  cudaError_t e = upload(x0, &d_x[0]);
  if (e == cudaSuccess) e = upload(x1, &d_x[1]);
  if (e == cudaSuccess && which == 1) e = upload(x2, &d_x[2]);
  if (e == cudaSuccess) e = cudaMalloc((void **)&d_out, (size_t)slab * sizeof(float));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (e == cudaSuccess) e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  std::vector<float> vals((size_t)slab), rec((size_t)slab * (D + 1));
  std::vector<float> c0(n0), c1(n1), c2(n2);
  for (long long i = 0; i < n0; ++i) c0[i] = text_rounded(x0[i]);
  for (long long i = 0; i < n1; ++i) c1[i] = which == 0 ? text_rounded(x1[i]) : (float)x1[i];  // "%4u" for n
  for (long long i = 0; i < n2 && which == 1; ++i) c2[i] = text_rounded(x2[i]);
  bool io_ok = true;
  for (long long first = 0; first < total && e == cudaSuccess && io_ok; first += slab) {
    const long long cnt = std::min(slab, total - first);
    cudaEventRecord(e0);
    e = launch_phiphi_table(which, first, cnt, (int)n1, (int)n2, d_x[0], d_x[1], d_x[2], d_out);
    cudaEventRecord(e1);
    if (e == cudaSuccess) e = cudaMemcpy(vals.data(), d_out, (size_t)cnt * sizeof(float), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) break;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    *ms_total += ms;
    for (long long i = 0; i < cnt; ++i) {
      const long long node = first + i;
      float *r = &rec[(size_t)i * (D + 1)];
      if (which == 0) {
        r[0] = c0[node / n1];
        r[1] = c1[node % n1];
        r[2] = vals[i];
      } else {
        const long long i01 = node / n2;
        r[0] = c0[i01 / n1];
        r[1] = c1[i01 % n1];
        r[2] = c2[node % n2];
        r[3] = vals[i];
      }
    }
    io_ok = std::fwrite(rec.data(), sizeof(float) * (D + 1), (size_t)cnt, fp) == (size_t)cnt;
  }
  io_ok = (std::fclose(fp) == 0) && io_ok;
  for (double *p : d_x)
    if (p) cudaFree(p);
  if (d_out) cudaFree(d_out);
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  if (e != cudaSuccess) return fail(NUSI_E_CUDA, std::string("phiphi_table_kernel: ") + cudaGetErrorString(e));
  if (!io_ok) return fail(NUSI_E_TABLE, std::string("short write on ") + path);
  return NUSI_OK;
}

int nusi_phiphi_tables_generate(int device, const char *alphatilde_path, const char *alpha_path, int n_t, int n_s,
                                int n_n, int n_d, double *seconds) {
  if (n_t == 0) n_t = 5000;
  if (n_s == 0) n_s = 1000;
  if (n_n == 0) n_n = 1000;
  if (n_d == 0) n_d = 100;
  if (n_t < 4 || n_s < 4 || n_n < 4 || n_d < 4) return fail(NUSI_E_INVALID, "phi-phi tables need at least 4 nodes per axis");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(device));
  float ms = 0;
  int rc = NUSI_OK;
  const std::vector<double> l10d = linspace(0.005, 0.05, n_d);
  if (alphatilde_path) rc = generate_table(0, alphatilde_path, geomspace(4.0, 1e4, n_t), l10d, std::vector<double>(), &ms);
  if (rc == NUSI_OK && alpha_path) {
    std::vector<double> nn(n_n);
    for (int i = 0; i < n_n; ++i) nn[i] = i + 1;
    rc = generate_table(1, alpha_path, geomspace(4.0, 1e4, n_s), nn, l10d, &ms);
  }
  if (seconds) *seconds = ms * 1e-3;
  return rc;
}

int nusi_specfun_eval(int device, int which, int n, const double *x, const double *y, double *re, double *im) {
  if (n <= 0 || !x || !re || which < 0 || which > 7 || (which == 1 && (!y || !im)) || (which == 6 && !y)) return NUSI_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(device));
  double *d[4] = {nullptr, nullptr, nullptr, nullptr};
  const size_t b = (size_t)n * sizeof(double);
  for (int i = 0; i < 4; ++i) CU(cudaMalloc((void **)&d[i], b));
  CU(cudaMemcpy(d[0], x, b, cudaMemcpyHostToDevice));
  if (y) CU(cudaMemcpy(d[1], y, b, cudaMemcpyHostToDevice));
  CU(launch_specfun(which, n, d[0], d[1], d[2], d[3]));
  CU(cudaMemcpy(re, d[2], b, cudaMemcpyDeviceToHost));
  if (im && which == 1) CU(cudaMemcpy(im, d[3], b, cudaMemcpyDeviceToHost));
  for (int i = 0; i < 4; ++i) cudaFree(d[i]);
  return NUSI_OK;
}

int nusi_measure_fp64_peak(int device, double *tflops) {
  if (!tflops) return NUSI_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  const int blocks = prop.multiProcessorCount * 8, thr = 256, iters = 1 << 15;
  double *d = nullptr;
  CU(cudaMalloc((void **)&d, (size_t)blocks * thr * sizeof(double)));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  double best = 0;
  for (int rep = 0; rep < 4; ++rep) {
    CU(cudaEventRecord(e0));
    CU(launch_fp64_peak(blocks, thr, d, iters));
    CU(cudaEventRecord(e1));
    CU(cudaEventSynchronize(e1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    const double fl = 2.0 * 8.0 * (double)iters * blocks * thr;
    const double tf = fl / (ms * 1e-3) / 1e12;
    if (rep > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *tflops = best;
  return NUSI_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// nusi_energy_integral_static: nusi_energy_integral() for callers that must not allocate (code under CUDA graph capture,
// serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same reduction as energy_integral_kernel (k_misc.cu), same bits; the per-point results are staged in a
// file-scope device array instead of the context's cudaMalloc'ed d_eint, so the call allocates nothing.
//   * the array is NUSI_STATIC_MAX_POINTS doubles (512 KiB), held by every GPU that loads the library until exit;
//   * every call (of any context, any thread) shares it: calls that overlap on the GPU race; it keeps the last call's
//     contents, and nothing needs to start at zero -- each call writes slots [0, n) before it reads them back;
//   * in host code `g_eint_static` is only a stand-in for the symbol: the device address comes from
//     cudaGetSymbolAddress, once per device with that device current.
static __device__ __align__(256) double g_eint_static[NUSI_STATIC_MAX_POINTS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
double *g_eint_static_ptr[kMaxDevices] = {};  // device address of g_eint_static on device d, or NULL

// one warp per point; the arithmetic (and its order) of energy_integral_kernel, k_misc.cu
__global__ void __launch_bounds__(128) energy_integral_static_kernel(int n, int N, const double *__restrict__ flux_mass,
                                                                     const double *__restrict__ wE2, double *__restrict__ scratch) {
  const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (p >= n) return;
  const double *f = flux_mass + (size_t)p * 3 * N;
  double acc = 0.0;
  for (int i = lane; i < N; i += 32) acc += wE2[i] * ((f[i] + f[N + i]) + f[2 * N + i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) scratch[p] = acc;
}
}  // namespace

extern "C" int nusi_energy_integral_static(nusi_ctx *c, int n, const double *d_flux_mass, double *E_int, void *stream) {
  if (!c || n <= 0 || !d_flux_mass || !E_int) return NUSI_E_INVALID;
  if (n > NUSI_STATIC_MAX_POINTS)
    return fail(NUSI_E_INVALID, "nusi_energy_integral_static: n = " + std::to_string(n) + " exceeds NUSI_STATIC_MAX_POINTS = " +
                                    std::to_string(NUSI_STATIC_MAX_POINTS) + " (use nusi_energy_integral, or split the batch)");
  std::lock_guard<std::mutex> lk(c->mu);
  const int dev = c->opt.device;
  if (dev < 0 || dev >= kMaxDevices) return fail(NUSI_E_INVALID, "nusi_energy_integral_static: device index out of range");
  CU(cudaSetDevice(dev));
  double *scratch = nullptr;
  {
    std::lock_guard<std::mutex> ls(g_static_mu);
    if (!g_eint_static_ptr[dev]) {
      void *p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_eint_static);
      if (e != cudaSuccess || !p)
        return fail(NUSI_E_CUDA, std::string("nusi_energy_integral_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
      g_eint_static_ptr[dev] = static_cast<double *>(p);
    }
    scratch = g_eint_static_ptr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  energy_integral_static_kernel<<<(n + 3) / 4, 128, 0, s>>>(n, c->H.N, d_flux_mass, c->d_wE2, scratch);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(E_int, scratch, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  return NUSI_OK;
}
