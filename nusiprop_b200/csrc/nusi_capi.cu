// nusi_capi.cu -- implementation of include/nusiprop_b200.h: contexts, batches of points, and the
// launch sequence  rows_kernel + alpha_kernel (cross-section builder)  ->  integrate_kernel.
// There is deliberately no host compute path in this library.
#include "../../include/nusiprop_b200.h"
#include "nusi_kernels.cuh"
#include "nusi_host.hpp"

#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <array>
#include <map>
#include <string>
#include <vector>
#include <utility>

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
      return fail(NUSI_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                \
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
#ifndef NUSI_DIAG_SB_SMALL
#define NUSI_DIAG_SB_SMALL 32  // z-steps per band of the integrator for N_bins_E <= 225
#endif
#define NUSI_GROUP_CAP 256  // most points of one scan-builder group

struct nusi_ctx {
  nusi_opts opt;
  HostGrid H;
  DeviceGrid G;
  // device arrays owned by the context
  double *d_row_lo = nullptr, *d_row_hi = nullptr, *d_dE = nullptr, *d_cz = nullptr, *d_ndz = nullptr;
  double *d_source = nullptr, *d_Emin = nullptr, *d_Emax = nullptr, *d_zsrc = nullptr, *d_rsn = nullptr;
  double *d_pp[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  double *d_edgeE = nullptr;  // distinct bin-edge energies of the table cells (tiled builder)
  int2 *d_tiles = nullptr;
  int n_tiles = 0;
  unsigned char *d_pool = nullptr;  // points block of nusi_evolve_batch, kept across calls
  size_t pool_cap = 0;
  int2 *d_unit_groups = nullptr;  // (0, L) for L = 0..NUSI_GROUP_CAP: one-group launches of the scan builder
  int4 *d_ctiles = nullptr;  // compact builder: per tile the cell ranges [a0, a1) x [b0, b1) of its row / column segment
  int n_ctiles = 0;
  int n_edges = 0;
  bool use_tiles = false;
  double *d_edges = nullptr;  // per-launch edge blocks [cap_points][edge_block_doubles]
  // per-launch work space (grown on demand)
  int cap_points = 0;
  size_t tstride = 0;
  double *d_rows = nullptr, *d_tri = nullptr, *d_work = nullptr;
  int max_chunk = 0;
  // host-buffer path (nusi_evolve_batch): device staging of the outputs, a compute and a copy stream
  double *d_out_fm = nullptr, *d_out_ff = nullptr;
  int *d_out_st = nullptr;
  int out_cap = 0;
  cudaStream_t s_x = nullptr, s_i = nullptr, s_copy = nullptr;  // builder, rows+integrator (high priority), D2H
  cudaEvent_t ev_fork = nullptr, ev_jx = nullptr, ev_ji = nullptr, ev_jc = nullptr;
  std::vector<cudaEvent_t> ev_alpha, ev_done;  // per slice: table ready / fluxes ready
  // timing of the last evolve call
  std::vector<cudaEvent_t> ev;  // per slice 6: rows begin/end, builder begin/end, integrator begin/end
  int ev_used = 0;
  int n_launches = 0;
  std::map<std::pair<double, int>, std::pair<bool, std::array<double, 3>>> mass_cache;
};

struct nusi_points {
  int n = 0;
  int device = 0;
  PointParams *d_params = nullptr;
  int *d_status0 = nullptr;  // host-side status (NO_MASS) pre-uploaded, OR'ed into the outputs
  std::vector<int> h_status0;
  // scan builder (elementary_mode 3): runs of consecutive points with identical (mphi, m_nu), cut at slice boundaries
  // and at NUSI_GROUP_MAX points; (first point relative to its slice, length)
  int2 *d_groups = nullptr;
  unsigned char *d_block = nullptr;  // the one device allocation behind d_params / d_groups / d_status0
  bool pooled = false;               // d_block belongs to the context (nusi_evolve_batch)
  std::vector<int2> h_groups;  // host copy with ABSOLUTE first point (nusi_tables_batch)
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
  if (!mn) return NUSI_E_INVALID;
  return solve_masses(mntot, normal_ordering != 0, mn) ? NUSI_OK : NUSI_E_INVALID;
}

void nusi_ctx_destroy(nusi_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->opt.device);
  double *ptrs[] = {c->d_row_lo, c->d_row_hi, c->d_dE, c->d_cz, c->d_ndz, c->d_source, c->d_Emin, c->d_Emax,
                    c->d_zsrc, c->d_rsn, c->d_pp[0], c->d_pp[1], c->d_pp[2], c->d_pp[3], c->d_pp[4],
                    c->d_rows, c->d_tri, c->d_work, c->d_edgeE, c->d_edges};
  if (c->d_out_fm) cudaFree(c->d_out_fm);
  if (c->d_out_ff) cudaFree(c->d_out_ff);
  if (c->d_out_st) cudaFree(c->d_out_st);
  if (c->s_x) cudaStreamDestroy(c->s_x);
  if (c->s_i) cudaStreamDestroy(c->s_i);
  if (c->s_copy) cudaStreamDestroy(c->s_copy);
  for (cudaEvent_t e : c->ev_done) cudaEventDestroy(e);
  for (cudaEvent_t e : c->ev_alpha) cudaEventDestroy(e);
  for (cudaEvent_t e : {c->ev_fork, c->ev_jx, c->ev_ji, c->ev_jc})
    if (e) cudaEventDestroy(e);
  if (c->d_tiles) cudaFree(c->d_tiles);
  if (c->d_ctiles) cudaFree(c->d_ctiles);
  if (c->d_unit_groups) cudaFree(c->d_unit_groups);
  if (c->d_pool) cudaFree(c->d_pool);
  for (double *p : ptrs)
    if (p) cudaFree(p);
  for (cudaEvent_t e : c->ev) cudaEventDestroy(e);
  delete c;
}

int nusi_ctx_create(nusi_ctx **out, const nusi_opts *o) {
  if (!out || !o) return NUSI_E_INVALID;
  *out = nullptr;
  if (o->N_bins_E < 2 || !(o->lEmax > o->lEmin) || !(o->zmax > 0) || o->flav < 0 || o->flav > 2)
    return fail(NUSI_E_INVALID, "bad binning / flavour options");
  if (o->source_model < 0 || o->source_model > 1 ||
      (o->source_model == 1 && (!(std::fabs(1 - o->si) > 1e-12) || !std::isfinite(o->si) || !std::isfinite(o->norm))))
    return fail(NUSI_E_INVALID, "bad source options (source_model 0 or 1; si != 1 and finite; norm finite)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || o->device < 0 || o->device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device with the requested ordinal");
  CU(cudaSetDevice(o->device));

  nusi_ctx *c = new nusi_ctx;
  c->opt = *o;
  build_grid(o->N_bins_E, o->lEmin, o->lEmax, o->zmax, o->normal_ordering != 0, c->H);
  const HostGrid &H = c->H;
  if (H.Nz < 2) {
    delete c;
    return fail(NUSI_E_INVALID, "zmax too small for this binning (fewer than 2 redshift steps)");
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
    const char *p_at = o->alphatilde_phiphi_path ? o->alphatilde_phiphi_path : "xsec/alphatilde_phiphi.bin";
    const char *p_a = o->alpha_phiphi_path ? o->alpha_phiphi_path : "xsec/alpha_phiphi.bin";
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

  // source term, once per context, on the device
  const int S = H.Nz - 1;
  cudaError_t e = cudaMalloc((void **)&c->d_source, (size_t)S * H.N * sizeof(double));
  if (e == cudaSuccess && o->source_model == 1) {
    // upstream power law (nuSIprop.hpp:648-657): host arithmetic (glibc pow), uploaded once
    std::vector<double> src((size_t)S * H.N);
    powerlaw_source(H.N, S, H.zsrc.data(), H.Emin.data(), H.Emax.data(), H.zmax_eff, o->si, o->norm, src.data());
    e = cudaMemcpy(c->d_source, src.data(), src.size() * sizeof(double), cudaMemcpyHostToDevice);
  } else if (e == cudaSuccess) {
    source_kernel<<<(S * H.N + 127) / 128, 128>>>(H.N, S, c->d_zsrc, c->d_rsn, H.lum_pref, c->d_Emin, c->d_Emax,
                                                  c->d_source);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
  }
  if (e != cudaSuccess) {
    nusi_ctx_destroy(c);
    return fail(NUSI_E_CUDA, std::string("source_kernel: ") + cudaGetErrorString(e));
  }
  G.source = c->d_source;

  // tiled builder (Majorana + non-resonant): edge list with shared main-grid edges
  if (G.non_resonant && G.majorana && !getenv("NUSI_NO_TILES")) {
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
      if (e1 == cudaSuccess)
        e1 = cudaFuncSetAttribute(alpha_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(GroupSmem));
      if (e1 == cudaSuccess)
        e1 = cudaFuncSetAttribute(alpha_tile_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem));
      if (e1 == cudaSuccess)
        e1 = cudaFuncSetAttribute(alpha_tile_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TileSmem));

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
int nusi_ctx_source(const nusi_ctx *c, double *out) {
  if (!c || !out) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  CU(cudaMemcpy(out, c->d_source, (size_t)(c->H.Nz - 1) * c->H.N * sizeof(double), cudaMemcpyDeviceToHost));
  return NUSI_OK;
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

int nusi_powerlaw_source(int N_bins_E, double lEmin, double lEmax, double zmax, double si, double norm, double *out) {
  if (!out || N_bins_E < 2 || !(lEmax > lEmin) || !(zmax > 0) || !(std::fabs(1 - si) > 1e-12)) return NUSI_E_INVALID;
  HostGrid H;
  build_grid(N_bins_E, lEmin, lEmax, zmax, true, H);
  if (H.Nz < 2) return NUSI_E_INVALID;
  powerlaw_source(H.N, H.Nz - 1, H.zsrc.data(), H.Emin.data(), H.Emax.data(), H.zmax_eff, si, norm, out);
  return NUSI_OK;
}

// host preparation of a batch: masses (cached per mntot), width, |U_fk|^2
static void prepare_points(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                           std::vector<PointParams> &pp, std::vector<int> &st) {
  pp.resize(n);
  st.assign(n, 0);
  const bool no = c->opt.normal_ordering != 0;
  for (int i = 0; i < n; ++i) {
    PointParams &P = pp[i];
    std::memset(&P, 0, sizeof(P));
    auto key = std::make_pair(mntot[i], (int)no);
    auto it = c->mass_cache.find(key);
    if (it == c->mass_cache.end()) {
      std::array<double, 3> m = {0, 0, 0};
      const bool ok = solve_masses(mntot[i], no, m.data());
      it = c->mass_cache.emplace(key, std::make_pair(ok, m)).first;
    }
    P.majorana = c->G.majorana;
    P.non_resonant = c->G.non_resonant;
    P.phiphi = c->G.phiphi;
    if (!it->second.first || !(mphi[i] > 0)) {
      st[i] = NUSI_ST_NO_MASS;
      P.mphi = -1.0;  // kernels skip the point and write NaN
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

static int ensure_workspace(nusi_ctx *c, int npts) {
  if (npts <= c->cap_points) return NUSI_OK;
  if (c->d_rows) cudaFree(c->d_rows);
  if (c->d_tri) cudaFree(c->d_tri);
  if (c->d_work) cudaFree(c->d_work);
  if (c->d_edges) cudaFree(c->d_edges);
  c->d_rows = c->d_tri = c->d_work = c->d_edges = nullptr;
  c->cap_points = 0;
  CU(cudaMalloc((void **)&c->d_rows, (size_t)npts * 2 * c->H.NT * sizeof(double)));
  CU(cudaMalloc((void **)&c->d_tri, (size_t)npts * c->tstride * sizeof(double)));
  if (!c->G.non_resonant) CU(cudaMalloc((void **)&c->d_work, (size_t)npts * 3 * c->H.N * sizeof(double)));
  if (c->use_tiles) CU(cudaMalloc((void **)&c->d_edges, (size_t)npts * edge_block_doubles(c->n_edges, c->H.NT) * sizeof(double)));
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

// Workspace of one launch: point i of the launch uses slot wo + i of rows / tri / edges.
struct WorkSlot {
  double *rows, *tri, *edges;
};
static WorkSlot work_slot(const nusi_ctx *c, int wo) {
  WorkSlot w;
  w.rows = c->d_rows + (size_t)wo * 2 * c->H.NT;
  w.tri = c->d_tri + (size_t)wo * c->tstride;
  w.edges = c->d_edges ? c->d_edges + (size_t)wo * edge_block_doubles(c->n_edges, c->H.NT) : nullptr;
  return w;
}

// Gamma / alphaTilde rows of `cnt` points
static int launch_rows(nusi_ctx *c, const PointParams *d_params, int cnt, int *d_status, const WorkSlot &W, cudaStream_t s,
                       const int2 *d_groups = nullptr, int ngroups = 0) {
  const DeviceGrid &G = c->G;
  dim3 gr((G.NT + NUSI_ROWS_PER_CTA - 1) / NUSI_ROWS_PER_CTA, cnt);
  const char *mode_env = getenv("NUSI_ALPHA_MODE");
  const int mode = mode_env ? atoi(mode_env) : c->opt.elementary_mode;
  if (mode >= 3 && d_groups && ngroups > 0 && G.majorana && G.non_resonant && !G.phiphi && !getenv("NUSI_ROWS_UNGROUPED")) {
    const char *rs = getenv("NUSI_ROWS_SPLIT");
    // scan: the g-independent part of a row once per (part of a) group.  A group is shared by `split` CTAs: 4 when the
    // launch is large (the per-point part dominates, so repeating the shared part four times is cheap and the grid is
    // 4x finer), more for small launches so that the grid still fills the 148 SMs x 4 CTAs twice (each CTA keeps >= 1 point).
    int split = 4;
    const long long want = 1184, have = (long long)gr.x * ngroups * split;
    if (have < want) split = (int)std::min<long long>(32, (want + (long long)gr.x * ngroups - 1) / ((long long)gr.x * ngroups));
    if (rs && atoi(rs) > 0) split = atoi(rs);
    dim3 gg(gr.x, ngroups, split);
    rows_group_kernel<0><<<gg, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, d_params, d_groups, W.rows, d_status);
    rows_group_kernel<1><<<gg, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, d_params, d_groups, W.rows, d_status);
    c->n_launches += 2;
    CU(cudaGetLastError());
    return NUSI_OK;
  }
  rows_kernel<0><<<gr, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, d_params, W.rows, d_status);
  rows_kernel<1><<<gr, NUSI_ROWS_PER_CTA * 3, 0, s>>>(G, d_params, W.rows, d_status);
  c->n_launches += 2;
  CU(cudaGetLastError());
  return NUSI_OK;
}

// alpha table of `cnt` points
static int launch_alpha(nusi_ctx *c, const PointParams *d_params, int cnt, int *d_status, const WorkSlot &W, cudaStream_t s,
                        const int2 *d_groups = nullptr, int ngroups = 0) {
  const DeviceGrid &G = c->G;
  const char *mode_env = getenv("NUSI_ALPHA_MODE");
  const int mode = mode_env ? atoi(mode_env) : c->opt.elementary_mode;
  if (c->use_tiles) {
    const int ntask = 3 * (2 * c->n_edges + G.NT);
    dim3 ge((ntask + 127) / 128, cnt);
    edges_kernel<<<ge, 128, 0, s>>>(G, c->n_edges, d_params, c->d_edgeE, W.edges);
    dim3 gt(c->n_tiles, cnt);
    // the scan builder pays off from ~1.3 points per group on (measured on B200, 8192 points: groups of 1: 45.3 ms,
    // of 2: 31.7, of 8: 18.3, of 32: 14.5; compact builder 42.0 ms)
    if (mode >= 3 && d_groups && ngroups > 0 && !G.phiphi && (4 * (long long)ngroups <= 3 * (long long)cnt || getenv("NUSI_GROUP_ALWAYS"))) {
      // scan builder: one CTA per (tile, group of points with the same mphi and neutrino masses)
      dim3 gg(c->n_ctiles, ngroups);
      alpha_group_kernel<<<gg, 256, sizeof(GroupSmem), s>>>(G, c->n_edges, d_params, W.edges, c->d_ctiles, d_groups, W.tri, c->tstride, d_status);
    } else if (mode >= 2) {
      dim3 gc(c->n_ctiles, cnt);
      alpha_compact_kernel<<<gc, 256, 0, s>>>(G, c->n_edges, d_params, W.edges, c->d_ctiles, W.tri, c->tstride, d_status);
    } else if (mode == 0)
      alpha_tile_kernel<0><<<gt, 256, sizeof(TileSmem), s>>>(G, c->n_edges, d_params, c->d_edgeE, W.edges, c->d_tiles, W.tri,
                                                            c->tstride, d_status);
    else
      alpha_tile_kernel<1><<<gt, 256, sizeof(TileSmem), s>>>(G, c->n_edges, d_params, c->d_edgeE, W.edges, c->d_tiles, W.tri,
                                                            c->tstride, d_status);
    c->n_launches += 2;
  } else {
    const long long nent = G.non_resonant ? (long long)G.NT * (G.NT - 1) / 2 : (long long)(G.NT - 1);
    dim3 ga((unsigned)((nent + 127) / 128), cnt);
    alpha_kernel<<<ga, 128, 0, s>>>(G, d_params, W.tri, c->tstride, d_status, mode >= 2 ? 1 : 0);
    c->n_launches += 1;
  }
  CU(cudaGetLastError());
  return NUSI_OK;
}

static int launch_integrate(nusi_ctx *c, const PointParams *d_params, int cnt, double *d_fm, double *d_ff,
                            const WorkSlot &W, int wo, cudaStream_t s) {
  const DeviceGrid &G = c->G;
  double *const d_work = c->d_work ? c->d_work + (size_t)wo * 3 * c->H.N : nullptr;
  if (G.non_resonant) {
    const int N = G.N, NT = G.NT, S = G.Nz - 1;
    if (getenv("NUSI_SERIAL_INTEGRATOR")) {  // reference-shaped bin-by-bin kernel, kept for cross-checks
      if (N > 256) return fail(NUSI_E_INVALID, "serial integrator supports N_bins_E <= 256");
      integrate_kernel<1><<<cnt, ((N + 31) / 32) * 32, 0, s>>>(G, d_params, W.rows, W.tri, c->tstride, d_fm, d_ff);
    } else if (getenv("NUSI_WAVE_INTEGRATOR")) {  // bin-owned accumulators, kept for cross-checks
#define NUSI_WAVE(BPT, SB, NST)                                                                             \
  do {                                                                                                      \
    const size_t sh = (size_t)((NST + 2) * NT + 2 * 32 + 2 * S) * sizeof(double);                           \
    CU(cudaFuncSetAttribute(integrate_wave_kernel<BPT, SB, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
    integrate_wave_kernel<BPT, SB, NST><<<cnt, 256, sh, s>>>(G, d_params, W.rows, W.tri, c->tstride, d_fm, d_ff); \
  } while (0)
      if (N <= 256 && NT <= 256 * 3)
        NUSI_WAVE(1, 32, 8);
      else if (N <= 512 && NT <= 256 * 4)
        NUSI_WAVE(2, 16, 8);
      else if (N <= 1024 && NT <= 256 * 6)
        NUSI_WAVE(4, 8, 8);
      else if (N <= 2048 && NT <= 256 * 10)
        NUSI_WAVE(8, 4, 6);
      else
        return fail(NUSI_E_INVALID, "N_bins_E > 2048 is not supported by the integrator");
#undef NUSI_WAVE
    } else {
      // diagonal-owned accumulators: N <= 256 BPT bins, N + SB - 1 <= 256 DPT diagonals, NT <= 256 (BPT + 2) table rows
#define NUSI_DIAG(BPT, SB, DPT, NST)                                                                        \
  do {                                                                                                      \
    const size_t sh = (size_t)((NST + 2) * NT + 2 * SB + 2 * SB * NUSI_PRE_DOUBLES + 2 * S + 2 * N + 2) * sizeof(double); \
    CU(cudaFuncSetAttribute(integrate_diag_kernel<BPT, SB, DPT, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
    integrate_diag_kernel<BPT, SB, DPT, NST><<<cnt, 256, sh, s>>>(G, d_params, W.rows, W.tri, c->tstride, d_fm, d_ff); \
  } while (0)
      // More threads per point for finer binning = fewer bins and diagonals per thread, so wide bands still fit the
      // registers and the table is streamed fewer times (bands x waves is what a latency-bound point pays for).
      // NUSI_FINE_VARIANT=0 selects the 256-thread configurations throughout (kept for comparison).
#define NUSI_DIAG_T(BPT, SB, DPT, NST, NTHR)                                                                \
  do {                                                                                                      \
    const size_t sh = (size_t)((NST + 2) * NT + 2 * SB + 2 * SB * NUSI_PRE_DOUBLES + 2 * S + 2 * N + 2) * sizeof(double); \
    CU(cudaFuncSetAttribute(integrate_diag_kernel<BPT, SB, DPT, NST, NTHR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh)); \
    integrate_diag_kernel<BPT, SB, DPT, NST, NTHR><<<cnt, NTHR, sh, s>>>(G, d_params, W.rows, W.tri, c->tstride, d_fm, d_ff); \
  } while (0)
      const char *fv = getenv("NUSI_FINE_VARIANT");
      const bool wide = !(fv && atoi(fv) == 0);
      if (N <= 225 && NT <= 256 && !getenv("NUSI_DIAG_PRE3")) {  // a table column fits one row per thread (N_E = 200: NT = 231)
        const size_t sh = (size_t)((8 + 2) * NT + 2 * NUSI_DIAG_SB_SMALL + 2 * NUSI_DIAG_SB_SMALL * NUSI_PRE_DOUBLES + 2 * S + 2 * N + 2) * sizeof(double);
        CU(cudaFuncSetAttribute(integrate_diag_kernel<1, NUSI_DIAG_SB_SMALL, 1, 8, 256, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
        integrate_diag_kernel<1, NUSI_DIAG_SB_SMALL, 1, 8, 256, 1><<<cnt, 256, sh, s>>>(G, d_params, W.rows, W.tri, c->tstride, d_fm, d_ff);
      } else if (N <= 225 && NT <= 256 * 3)
        NUSI_DIAG(1, NUSI_DIAG_SB_SMALL, 1, 8);
      else if (N <= 256 && NT <= 256 * 3)
        NUSI_DIAG(1, 32, 2, 8);
      else if (wide && N + 32 - 1 <= 512 && NT <= 512 * 3)
        NUSI_DIAG_T(1, 32, 1, 8, 512);
      else if (N <= 512 && NT <= 256 * 4)
        NUSI_DIAG(2, 16, 3, 8);
      else if (wide && N <= 1024 && NT <= 512 * 4)
        NUSI_DIAG_T(2, 16, 3, 8, 512);
      else if (N <= 1024 && NT <= 256 * 6)
        NUSI_DIAG(4, 8, 5, 8);
      else if (wide && N + 8 - 1 <= 512 * 4 && NT <= 512 * 6)
        NUSI_DIAG_T(4, 8, 4, 6, 512);
      else if (N <= 2048 && NT <= 256 * 10)
        NUSI_DIAG(8, 4, 9, 6);
      else
        return fail(NUSI_E_INVALID, "N_bins_E > 2048 is not supported by the integrator");
#undef NUSI_DIAG_T
#undef NUSI_DIAG
    }
  } else {
    integrate_resonant_kernel<<<(cnt + 63) / 64, 64, 0, s>>>(G, cnt, d_params, W.rows, W.tri, c->tstride,
                                                             d_work, d_fm, d_ff);
  }
  c->n_launches += 1;
  CU(cudaGetLastError());
  return NUSI_OK;
}

// One device block per points object: [params | groups | status0], filled by ONE host->device copy.  pooled: the block
// is the context's own (grown on demand, kept across calls) -- nusi_evolve_batch, which would otherwise pay three
// cudaMalloc + three cudaFree (each a device synchronisation) per call.
static int points_create_impl(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                              nusi_points **out, bool pooled) {
  if (!c || !out || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  std::vector<PointParams> pp;
  nusi_points *p = new nusi_points;
  p->n = n;
  p->device = c->opt.device;
  prepare_points(c, n, mphi, g, mntot, pp, p->h_status0);
  std::vector<int2> groups;
  {
    const char *gm = getenv("NUSI_GROUP_MAX");
    const int gmax = gm ? (atoi(gm) > 0 ? (atoi(gm) < NUSI_GROUP_CAP ? atoi(gm) : NUSI_GROUP_CAP) : 1) : 32;
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
  const size_t o_groups = ((size_t)n * sizeof(PointParams) + 15) & ~(size_t)15;
  const size_t o_status = o_groups + ((groups.size() * sizeof(int2) + 15) & ~(size_t)15);
  const size_t bytes = o_status + (size_t)n * sizeof(int);
  std::vector<unsigned char> stage(bytes);
  std::memcpy(stage.data(), pp.data(), (size_t)n * sizeof(PointParams));
  std::memcpy(stage.data() + o_groups, groups.data(), groups.size() * sizeof(int2));
  std::memcpy(stage.data() + o_status, p->h_status0.data(), (size_t)n * sizeof(int));
  cudaError_t e = cudaSuccess;
  unsigned char *base = nullptr;
  if (pooled) {
    if (bytes > c->pool_cap) {
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

// ------------------------------------------------------------------------------------------------------
// The evolve pipeline.  A batch is cut into slices of NUSI_SLICE points.  Per slice k:
//     stream s_i (high priority):  rows(k)                      wait table(k)   integrate(k)
//     stream s_x                :  edges(k) + alpha table(k) ------^       edges(k+1) + alpha table(k+1) ...
// With host buffers each slice's fluxes go home on a copy stream under the next slice's kernels (always on).
// The two-stream compute overlap drawn above is OPT-IN (NUSI_OVERLAP=1): measured on B200 it is neutral
// (29.8 vs 29.9 ms per 4096 points) -- the builder already owns every register of an SM at 4 CTAs/SM, so the
// integrator's CTAs displace builder CTAs one for one -- and it blurs the per-kernel event times.  By default
// all kernels run on the caller's stream in order.
// ------------------------------------------------------------------------------------------------------
static int ensure_streams(nusi_ctx *c) {
  if (c->s_x) return NUSI_OK;
  int lo = 0, hi = 0;
  CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
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

// d_fm / d_ff / d_st: device outputs indexed from the batch start.  h_*: optional host destinations (same shapes).
static int run_pipeline(nusi_ctx *c, const nusi_points *p, double *d_fm, double *d_ff, int *d_st, cudaStream_t caller,
                        double *h_fm, double *h_ff, int *h_st) {
  const int N = c->H.N;
  const bool to_host = h_fm || h_ff || h_st;
  const bool overlap = getenv("NUSI_OVERLAP") != nullptr;
  int rc = ensure_streams(c);
  if (rc) return rc;
  // slice size SL; the workspace holds a whole number of slices (all of them when the batch fits)
  const int SL = c->max_chunk < NUSI_SLICE ? c->max_chunk : NUSI_SLICE;
  const int chunk = p->n <= c->max_chunk ? ((p->n + SL - 1) / SL) * SL : (c->max_chunk / SL) * SL;
  if ((rc = ensure_workspace(c, chunk))) return rc;
  c->ev_used = 0;
  c->n_launches = 0;
  cudaStream_t sx = overlap ? c->s_x : caller, si = overlap ? c->s_i : caller, sc = (overlap || to_host) ? c->s_copy : caller;
  if (overlap || to_host) {
    CU(cudaEventRecord(c->ev_fork, caller));
    if (overlap) {
      CU(cudaStreamWaitEvent(sx, c->ev_fork, 0));
      CU(cudaStreamWaitEvent(si, c->ev_fork, 0));
    }
  }
  size_t k = 0;
  for (int off = 0; off < p->n; off += SL, ++k) {
    const int cnt = (p->n - off < SL) ? p->n - off : SL;
    const int wo = off % chunk;  // the slices of one chunk use distinct workspace slots
    const WorkSlot W = work_slot(c, wo);
    cudaEvent_t *t6, ev_tab, ev_out;
    if ((rc = next_events(c, &t6))) return rc;
    if ((rc = slice_event(c->ev_alpha, k, &ev_tab))) return rc;
    if (wo == 0 && off > 0 && overlap) {  // a new chunk reuses the workspace: everything queued so far must be done with it
      CU(cudaEventRecord(c->ev_ji, si));
      CU(cudaStreamWaitEvent(sx, c->ev_ji, 0));
    }
    CU(cudaEventRecord(t6[0], si));
    const bool grouped = (p->SL == SL) && p->d_groups && k + 1 < p->slice_g0.size();
    if ((rc = launch_rows(c, p->d_params + off, cnt, d_st + off, W, si, grouped ? p->d_groups + p->slice_g0[k] : nullptr,
                          grouped ? p->slice_g0[k + 1] - p->slice_g0[k] : 0)))
      return rc;
    CU(cudaEventRecord(t6[1], si));
    CU(cudaEventRecord(t6[2], sx));
    if ((rc = launch_alpha(c, p->d_params + off, cnt, d_st + off, W, sx, grouped ? p->d_groups + p->slice_g0[k] : nullptr,
                           grouped ? p->slice_g0[k + 1] - p->slice_g0[k] : 0)))
      return rc;
    CU(cudaEventRecord(t6[3], sx));
    if (overlap) {
      CU(cudaEventRecord(ev_tab, sx));
      CU(cudaStreamWaitEvent(si, ev_tab, 0));
    }
    CU(cudaEventRecord(t6[4], si));
    double *fm = d_fm ? d_fm + (size_t)off * 3 * N : nullptr;
    double *ff = d_ff ? d_ff + (size_t)off * 3 * N : nullptr;
    if ((rc = launch_integrate(c, p->d_params + off, cnt, fm, ff, W, wo, si))) return rc;
    CU(cudaEventRecord(t6[5], si));
    if (to_host) {
      if ((rc = slice_event(c->ev_done, k, &ev_out))) return rc;
      if (overlap) {  // the status words are complete when rows, builder and integrator of the slice are
        CU(cudaEventRecord(ev_out, si));
        CU(cudaStreamWaitEvent(sc, ev_out, 0));
      } else {
        CU(cudaEventRecord(ev_out, caller));
        CU(cudaStreamWaitEvent(sc, ev_out, 0));
      }
      const size_t o = (size_t)off * 3 * N, b = (size_t)cnt * 3 * N * sizeof(double);
      if (h_fm) CU(cudaMemcpyAsync(h_fm + o, d_fm + o, b, cudaMemcpyDeviceToHost, sc));
      if (h_ff) CU(cudaMemcpyAsync(h_ff + o, d_ff + o, b, cudaMemcpyDeviceToHost, sc));
      if (h_st) CU(cudaMemcpyAsync(h_st + off, d_st + off, (size_t)cnt * sizeof(int), cudaMemcpyDeviceToHost, sc));
    }
  }
  // join
  if (overlap) {
    CU(cudaEventRecord(c->ev_jx, sx));
    CU(cudaEventRecord(c->ev_ji, si));
    CU(cudaStreamWaitEvent(caller, c->ev_jx, 0));
    CU(cudaStreamWaitEvent(caller, c->ev_ji, 0));
  }
  if (to_host) {
    CU(cudaEventRecord(c->ev_jc, sc));
    CU(cudaStreamWaitEvent(caller, c->ev_jc, 0));
  }
  return NUSI_OK;
}

int nusi_evolve_points(nusi_ctx *c, const nusi_points *p, double *d_fm, double *d_ff, int *d_status, void *stream) {
  if (!c || !p) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  cudaStream_t s = (cudaStream_t)stream;
  int *d_st = d_status ? d_status : p->d_status0;
  if (d_status) CU(cudaMemcpyAsync(d_status, p->d_status0, (size_t)p->n * sizeof(int), cudaMemcpyDeviceToDevice, s));
  return run_pipeline(c, p, d_fm, d_ff, d_st, s, nullptr, nullptr, nullptr);
}

// Host buffers in, host buffers out: the pipeline above with the device->host copy of every finished slice on
// a copy stream (effective when the caller's buffers are page-locked; with pageable memory the driver stages
// the copies, still correct).
int nusi_evolve_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot,
                      double *flux_mass, double *flux_fla, int *status) {
  if (!c || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  const int N = c->H.N;
  if (n > c->out_cap) {
    if (c->d_out_fm) cudaFree(c->d_out_fm);
    if (c->d_out_ff) cudaFree(c->d_out_ff);
    if (c->d_out_st) cudaFree(c->d_out_st);
    c->d_out_fm = c->d_out_ff = nullptr;
    c->d_out_st = nullptr;
    c->out_cap = 0;
    CU(cudaMalloc((void **)&c->d_out_fm, (size_t)n * 3 * N * sizeof(double)));
    CU(cudaMalloc((void **)&c->d_out_ff, (size_t)n * 3 * N * sizeof(double)));
    CU(cudaMalloc((void **)&c->d_out_st, (size_t)n * sizeof(int)));
    c->out_cap = n;
  }
  int rc = ensure_streams(c);
  if (rc) return rc;
  nusi_points *p = nullptr;
  if ((rc = points_create_impl(c, n, mphi, g, mntot, &p, true))) return rc;
  cudaStream_t s = c->s_copy;  // any stream of ours serves as the "caller" of the pipeline
  cudaError_t e = cudaMemcpyAsync(c->d_out_st, p->d_status0, (size_t)n * sizeof(int), cudaMemcpyDeviceToDevice, s);
  if (e == cudaSuccess)
    rc = run_pipeline(c, p, flux_mass ? c->d_out_fm : nullptr, flux_fla ? c->d_out_ff : nullptr, c->d_out_st, s, flux_mass,
                      flux_fla, status);
  cudaError_t e2 = cudaDeviceSynchronize();
  nusi_points_destroy(p);
  if (rc) return rc;
  if (e == cudaSuccess) e = e2;
  if (e != cudaSuccess) return fail(NUSI_E_CUDA, std::string("evolve_batch: ") + cudaGetErrorString(e));
  return NUSI_OK;
}

int nusi_tables_batch(nusi_ctx *c, int n, const double *mphi, const double *g, const double *mntot, double *Gamma,
                      double *alphaTilde, double *alpha, int *status) {
  if (!c || n <= 0 || !mphi || !g || !mntot) return NUSI_E_INVALID;
  CU(cudaSetDevice(c->opt.device));
  const int NT = c->H.NT;
  nusi_points *p = nullptr;
  int rc = nusi_points_create(c, n, mphi, g, mntot, &p);
  if (rc) return rc;
  c->ev_used = 0;
  c->n_launches = 0;
  std::vector<double> hrows((size_t)2 * NT), htri(c->tstride);
  // group by group, through the same builder a batch of these points gets in nusi_evolve_batch
  for (size_t gi = 0; gi < p->h_groups.size() && rc == NUSI_OK; ++gi) {
    const int i0 = p->h_groups[gi].x, len = p->h_groups[gi].y;
    rc = ensure_workspace(c, len);
    if (rc) break;
    const WorkSlot W = work_slot(c, 0);
    rc = launch_rows(c, p->d_params + i0, len, p->d_status0 + i0, W, nullptr, c->d_unit_groups ? c->d_unit_groups + len : nullptr, 1);
    if (rc == NUSI_OK)
      rc = launch_alpha(c, p->d_params + i0, len, p->d_status0 + i0, W, nullptr, c->d_unit_groups ? c->d_unit_groups + len : nullptr, 1);
    if (rc) break;
    for (int i = i0; i < i0 + len; ++i) {
      cudaError_t e = cudaMemcpy(hrows.data(), c->d_rows + (size_t)(i - i0) * 2 * NT, hrows.size() * sizeof(double), cudaMemcpyDeviceToHost);
      if (e == cudaSuccess)
        e = cudaMemcpy(htri.data(), c->d_tri + (size_t)(i - i0) * c->tstride, htri.size() * sizeof(double), cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) {
        rc = fail(NUSI_E_CUDA, std::string("tables_batch: ") + cudaGetErrorString(e));
        break;
      }
      if (Gamma) std::memcpy(Gamma + (size_t)i * NT, hrows.data(), NT * sizeof(double));
      if (alphaTilde) std::memcpy(alphaTilde + (size_t)i * NT, hrows.data() + NT, NT * sizeof(double));
      if (alpha) {
        double *A = alpha + (size_t)i * NT * NT;
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
    cudaError_t e = cudaMemcpy(status, p->d_status0, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = fail(NUSI_E_CUDA, cudaGetErrorString(e));
  }
  nusi_points_destroy(p);
  return rc;
}

int nusi_last_timing(const nusi_ctx *c, float *ms_rows, float *ms_alpha, float *ms_integrate, int *n_launches) {
  if (!c) return NUSI_E_INVALID;
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
    phiphi_table_kernel<<<(unsigned)((cnt + 127) / 128), 128>>>(which, first, cnt, (int)n1, (int)n2, d_x[0], d_x[1], d_x[2], d_out);
    cudaEventRecord(e1);
    e = cudaGetLastError();
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
  if (n <= 0 || !x || !re || which < 0 || which > 3 || (which == 1 && (!y || !im))) return NUSI_E_INVALID;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(NUSI_E_NO_DEVICE, "no CUDA device");
  CU(cudaSetDevice(device));
  double *d[4] = {nullptr, nullptr, nullptr, nullptr};
  const size_t b = (size_t)n * sizeof(double);
  for (int i = 0; i < 4; ++i) CU(cudaMalloc((void **)&d[i], b));
  CU(cudaMemcpy(d[0], x, b, cudaMemcpyHostToDevice));
  if (y) CU(cudaMemcpy(d[1], y, b, cudaMemcpyHostToDevice));
  specfun_kernel<<<(n + 127) / 128, 128>>>(which, n, d[0], d[1], d[2], d[3]);
  CU(cudaGetLastError());
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
    fp64_peak_kernel<<<blocks, thr>>>(d, iters, 0.999999, 1e-9);
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
