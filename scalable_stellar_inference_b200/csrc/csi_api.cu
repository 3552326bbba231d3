// csi_api.cu -- the C ABI of libcsi (include/csi.h): handle management, constant preparation, data upload
// and the stream-ordered evaluation entry points.  No CPU fallback: creation fails without an sm_100 device.
#include <math.h>
#include <cmath>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <limits>
#include <mutex>

#include "csi_internal.h"

using namespace csi;

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CK(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (call);                                                                      \
    if (e_ != cudaSuccess)                                                                        \
      return fail(CSI_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                 \
  } while (0)

extern "C" const char *csi_last_error(void) { return g_err.c_str(); }
extern "C" int csi_version(void) { return CSI_VERSION; }

// -------------------------------------------------------------------------------------------------
// host-side constants (double precision)
// -------------------------------------------------------------------------------------------------
static double h_log_ndtr(double x) {
  if (x > 0) return log1p(-0.5 * erfc(x * M_SQRT1_2));
  double v = 0.5 * erfc(-x * M_SQRT1_2);
  if (v > 0) return log(v);
  // far tail: asymptotic expansion of the Mills ratio
  double x2 = x * x;
  return -0.5 * x2 - log(-x) - 0.5 * log(2 * M_PI) + log1p(-1.0 / x2 + 3.0 / (x2 * x2));
}

static double h_log_ndtr_diff(double alpha, double beta) {
  bool flip = alpha > 0;
  double a = flip ? -beta : alpha, b = flip ? -alpha : beta;
  if (b <= 0) {
    double lb = h_log_ndtr(b);
    double la = std::isinf(a) ? -INFINITY : h_log_ndtr(a);
    return lb + log1p(-exp(la - lb));
  }
  double ta = std::isinf(a) ? 0.0 : 0.5 * erfc(-a * M_SQRT1_2);
  double tb = std::isinf(b) ? 0.0 : 0.5 * erfc(b * M_SQRT1_2);
  return log1p(-(ta + tb));
}

static double tn_logz(double loc, double scale, double low, double high) {
  double alpha = std::isfinite(low) ? (low - loc) / scale : -INFINITY;
  double beta = std::isfinite(high) ? (high - loc) / scale : INFINITY;
  return h_log_ndtr_diff(alpha, beta);
}

static bool make_tn(const csi_tn &t, TnC &o) {
  if (!(t.scale > 0) || !(t.low < t.high)) return false;
  o.loc = (float)t.loc; o.scale = (float)t.scale; o.low = (float)t.low; o.high = (float)t.high;
  bool lo = std::isfinite(t.low), hi = std::isfinite(t.high);
  o.kind = lo && hi ? TN_INTERVAL : hi ? TN_UPPER : lo ? TN_LOWER : TN_REAL;
  const double lz = tn_logz(t.loc, t.scale, t.low, t.high);
  o.logz = (float)lz;
  o.inv_scale = (float)(1.0 / t.scale);
  o.lognorm = (float)(lz + log(t.scale));
  o.logw = lo && hi ? (float)log(t.high - t.low) : 0.0f;
  return true;
}

static const double kGrav = 6.6743e-8;

static int build_model(const csi_config &c, csi_handle *h) {
  DevModel &m = h->model;
  memset(&m, 0, sizeof(m));
  if (c.level < CSI_LEVEL_SINGLE || c.level > CSI_LEVEL_HIERARCHICAL) return fail(CSI_EINVAL, "level must be 0..2");
  if (c.kind != CSI_KIND_DIAG && c.kind != CSI_KIND_FULL)
    return fail(CSI_EINVAL, "Kind must be one of 'diag' or 'full' ('none' is unusable in the reference).");
  if (c.prior != CSI_PRIOR_FLAT && c.prior != CSI_PRIOR_INFORMATIVE)
    return fail(CSI_EINVAL, "Variable 'prior' must be one of ['flat', 'informative'].");
  if (c.bands < 0 || c.bands > CSI_MAX_BANDS) return fail(CSI_EINVAL, "bands must be 0..4");
  if (c.bands > 0 && c.bc_mode != CSI_BC_GRID && c.bc_mode != CSI_BC_NET)
    return fail(CSI_EINVAL, "photometry needs bc_mode grid or net");
  m.prior_mode = c.level == CSI_LEVEL_HIERARCHICAL ? P_HIER : (c.prior == CSI_PRIOR_FLAT ? P_FLAT : P_INFORMATIVE);
  m.kind = c.kind;
  m.bands = c.bands;
  m.bc_mode = c.bands > 0 ? c.bc_mode : 0;
  m.n_mean = 7 + (c.bands > 0 ? c.bands + 1 : 0);
  if (c.n_obs < 1 || c.n_obs > CSI_MAX_OBS) return fail(CSI_EINVAL, "n_obs must be 1..12");
  m.n_obs = c.n_obs;
  for (int k = 0; k < kMaxMean; ++k) m.col_of_mean[k] = -1;
  for (int j = 0; j < c.n_obs; ++j) {
    int idx = c.obs_index[j];
    if (idx < 0 || idx >= m.n_mean) return fail(CSI_EINVAL, "obs_index out of range for this model");
    if (c.kind == CSI_KIND_DIAG && m.col_of_mean[idx] >= 0)
      return fail(CSI_EINVAL, "duplicate observable");
    m.col_of_mean[idx] = j;
    m.obs_idx[j] = idx;
  }
  if (!make_tn(c.log_evol, m.log_evol) || !make_tn(c.log_mass, m.log_mass) || !make_tn(c.m_h, m.mh) ||
      !make_tn(c.dydz, m.dydz) || !make_tn(c.mu_a, m.mu_a))
    return fail(CSI_EINVAL, "truncated-normal prior needs scale > 0 and low < high");
  if (!(c.y_low < c.y_high) || !(c.a_low < c.a_high)) return fail(CSI_EINVAL, "Y / a_MLT bounds");
  m.y_lo = (float)c.y_low; m.y_hi = (float)c.y_high; m.y_w = (float)(c.y_high - c.y_low);
  m.log_y_w = (float)log(c.y_high - c.y_low);
  m.a_lo = (float)c.a_low; m.a_hi = (float)c.a_high; m.a_w = (float)(c.a_high - c.a_low);
  m.log_a_w = (float)log(c.a_high - c.a_low);
  m.y0_loc = (float)c.y0_loc; m.y0_scale = (float)c.y0_scale;
  m.py_conc = (float)c.py_conc; m.py_rate = (float)c.py_rate;
  m.py_const = (float)(c.py_conc * log(c.py_rate) - lgamma(c.py_conc));
  m.pa_conc = (float)c.pa_conc; m.pa_rate = (float)c.pa_rate;
  m.pa_const = (float)(c.pa_conc * log(c.pa_rate) - lgamma(c.pa_conc));
  const double hd = 0.5 * c.delta_df;
  m.tau_conc = (float)hd; m.tau_rate = (float)hd; m.tau_const = (float)(hd * log(hd) - lgamma(hd));
  // constants of SingleStarModel (numpyro_models.py:42-48, 61-84)
  const double log_m_sun = log10(1.988410) + 33, log_r_sun = log10(6.957) + 10;
  const double log_g_sun = log10(kGrav) + log_m_sun - 2 * log_r_sun;
  const double log_teff_sun = log10(5772.0), log_numax_sun = log10(3090.0);
  const double A[7][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1},
                          {0, 4, 2, 0}, {0, 0, -2, 0}, {0, -0.5, -2, 0}};
  const double cv[7] = {-9.0, 0, 0, 0, -4.0 * log_teff_sun, log_g_sun, log_numax_sun + 0.5 * log_teff_sun};
  double Aext[kMaxMean][4] = {{0}};
  for (int k = 0; k < 7; ++k) {
    double s = cv[k];
    for (int j = 0; j < 4; ++j) { s += A[k][j] * c.delta_loc[j]; Aext[k][j] = A[k][j]; }
    m.cc[k] = (float)s;
  }
  for (int b = 0; b < c.bands; ++b)
    for (int j = 0; j < 4; ++j) Aext[7 + b][j] = -2.5 * A[4][j];
  for (int a = 0; a < kMaxMean; ++a)
    for (int b = 0; b < kMaxMean; ++b) {
      double s = 0;
      for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) s += Aext[a][i] * c.delta_cov[i * 4 + j] * Aext[b][j];
      m.cov_ext[a * kMaxMean + b] = (float)s;
      if (a == b) m.var_ext[a] = (float)s;
    }
  m.sd_c1 = (float)c.sd_c1; m.sd_c0 = (float)c.sd_c0;
  m.sd_const = c.bands > 0 ? (float)(-(lgamma(c.sd_c1) + lgamma(c.sd_c0) - lgamma(c.sd_c1 + c.sd_c0))) : 0.0f;
  m.max_distance = (float)c.max_distance; m.av = (float)c.av;
  m.log_zx_sun = (float)log10(0.0181);
  m.bol_mag_sun = 4.75f;
  m.inf_dydz = (float)c.dydz.loc; m.inf_y0 = (float)c.y0_loc;
  m.inf_s1 = (float)c.dydz.scale; m.inf_s0 = (float)c.y0_scale;
  m.inf_mua_loc = (float)c.mu_a.loc; m.inf_mua_scale = (float)c.mu_a.scale;
  m.hyper_on = 1;

  // site layout in the reference's trace order
  static const char *hyper_names[5] = {"Y_0", "dY_dZ", "precision_Y", "mu_a", "precision_a"};
  int ns = 0;
  auto add = [&](int slot, const char *name) { h->slot_of_site[ns] = slot; h->names[h->n_hyper + ns] = name; ++ns; };
  h->n_hyper = m.prior_mode == P_HIER ? 5 : 0;
  for (int j = 0; j < h->n_hyper; ++j) h->names[j] = hyper_names[j];
  if (m.prior_mode == P_HIER) {
    add(S_TAU, "scaled_precision"); add(S_EVOL, "log_evol"); add(S_MASS, "log_mass"); add(S_MH, "M_H");
    add(S_Y, "Y_decentered"); add(S_A, "a_decentered");
  } else if (m.prior_mode == P_FLAT) {
    add(S_EVOL, "log_evol"); add(S_MASS, "log_mass"); add(S_MH, "M_H"); add(S_Y, "Y"); add(S_A, "a_MLT");
    add(S_TAU, "scaled_precision");
  } else {
    add(S_EVOL, "log_evol"); add(S_MASS, "log_mass"); add(S_MH, "M_H"); add(S_PY, "precision_Y"); add(S_Y, "Y");
    add(S_PA, "precision_a"); add(S_A, "a_MLT"); add(S_TAU, "scaled_precision");
  }
  if (c.bands > 0) add(S_SD, "scaled_distance");
  h->n_sites = ns;
  h->level = c.level; h->prior = c.prior;
  return CSI_OK;
}

// -------------------------------------------------------------------------------------------------
// uploads
// -------------------------------------------------------------------------------------------------
template <class T>
static int dev_upload(csi_handle *h, const std::vector<T> &v, T **out) {
  void *p = nullptr;
  CK(cudaMalloc(&p, std::max<size_t>(v.size() * sizeof(T), 16)));
  h->owned.push_back(p);
  if (!v.empty()) CK(cudaMemcpy(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *out = (T *)p;
  return CSI_OK;
}

static int check_net(const csi_net &n, const char *what) {
  if (n.n_layers < 2 || n.n_layers > CSI_MAX_LAYERS) return fail(CSI_EINVAL, std::string(what) + ": n_layers");
  if (n.n_in < 1 || n.n_in > 8 || n.n_out < 1 || n.n_out > 4 || n.n_hidden < 4 || n.n_hidden > 128 || n.n_hidden % 4)
    return fail(CSI_EINVAL, std::string(what) + ": unsupported widths");
  for (int l = 0; l < n.n_layers; ++l)
    if (!n.weights[l] || !n.bias[l]) return fail(CSI_EINVAL, std::string(what) + ": null weights");
  if (!n.in_loc || !n.in_scale || !n.out_loc || !n.out_scale) return fail(CSI_EINVAL, std::string(what) + ": null scaling");
  return CSI_OK;
}

static int upload_net(csi_handle *h, const csi_net &n, NetDev &d) {
  memset(&d, 0, sizeof(d));
  d.n_layers = n.n_layers; d.n_in = n.n_in; d.n_hidden = n.n_hidden; d.n_out = n.n_out;
  for (int l = 0; l < n.n_layers; ++l) {
    const int K = l == 0 ? n.n_in : n.n_hidden, N = l == n.n_layers - 1 ? n.n_out : n.n_hidden;
    const int np = (N + 3) & ~3, kp = (K + 3) & ~3;
    d.kdim[l] = K; d.ndim[l] = N; d.np[l] = np; d.kp[l] = kp;
    std::vector<float> W((size_t)K * np, 0.f), Wt((size_t)N * kp, 0.f), b(np, 0.f);
    for (int k = 0; k < K; ++k)
      for (int j = 0; j < N; ++j) {
        float w = n.weights[l][(size_t)k * N + j];
        W[(size_t)k * np + j] = w;
        Wt[(size_t)j * kp + k] = w;
      }
    for (int j = 0; j < N; ++j) b[j] = n.bias[l][j];
    float *p;
    int rc;
    if ((rc = dev_upload(h, W, &p))) return rc; d.W[l] = p;
    if ((rc = dev_upload(h, Wt, &p))) return rc; d.Wt[l] = p;
    if ((rc = dev_upload(h, b, &p))) return rc; d.b[l] = p;
  }
  for (int k = 0; k < n.n_in; ++k) { d.in_loc[k] = (float)n.in_loc[k]; d.in_scale[k] = (float)n.in_scale[k]; }
  for (int j = 0; j < n.n_out; ++j) { d.out_loc[j] = (float)n.out_loc[j]; d.out_scale[j] = (float)n.out_scale[j]; }
  return CSI_OK;
}

static int upload_grid(csi_handle *h, const csi_grid &g, int bands, float av, GridDev &d) {
  memset(&d, 0, sizeof(d));
  size_t nn = 1;
  for (int a = 0; a < 4; ++a) {
    if (g.shape[a] < 2 || !g.points[a]) return fail(CSI_EINVAL, "bc_grid: every axis needs >= 2 points");
    for (int i = 1; i < g.shape[a]; ++i)
      if (!(g.points[a][i] > g.points[a][i - 1])) return fail(CSI_EINVAL, "bc_grid: points must be strictly increasing");
    d.shape[a] = g.shape[a];
    nn *= (size_t)g.shape[a];
    std::vector<float> p(g.points[a], g.points[a] + g.shape[a]);
    float *dp; int rc;
    if ((rc = dev_upload(h, p, &dp))) return rc;
    d.pts[a] = dp;
    // uniform-bin lookup table over [p0, p_last]: lut[b] = (number of points < left edge of bin b) - 1, in [0, n-2]
    if (g.shape[a] > 65535) return fail(CSI_EINVAL, "bc_grid: axis too long");
    const float x0 = p.front(), x1 = p.back();
    const float inv = (float)kLutBins / (x1 - x0);
    std::vector<unsigned short> lut(kLutBins);
    for (int b = 0; b < kLutBins; ++b) {
      const float edge = x0 + (float)b / inv;
      int cnt = 0;
      while (cnt < g.shape[a] && p[cnt] < edge) ++cnt;
      lut[b] = (unsigned short)std::min(std::max(cnt - 1, 0), g.shape[a] - 2);
    }
    unsigned short *dl;
    if ((rc = dev_upload(h, lut, &dl))) return rc;
    d.lut[a] = dl; d.lut_x0[a] = x0; d.lut_inv[a] = inv;
  }
  if (!g.values) return fail(CSI_EINVAL, "bc_grid: null values");
  std::vector<float4> v(nn);
  for (size_t i = 0; i < nn; ++i) {
    float t[4] = {0.f, 0.f, 0.f, 0.f};
    for (int b = 0; b < bands; ++b) t[b] = (float)g.values[i * bands + b];
    v[i] = make_float4(t[0], t[1], t[2], t[3]);
  }
  float4 *dv; int rc;
  if ((rc = dev_upload(h, v, &dv))) return rc;
  d.values = dv;
  // Av-collapsed copy for the potential (GridDev::values3): locate `av` on axis 3 and interpolate every (T, g, F)
  // node along it with the SAME fp32 operations as the first level of grid_eval (fmaf(t, c1 - c0, c0)), so the
  // 3-D path returns bit-identical values to the 4-D one.
  {
    const int n3 = g.shape[3];
    std::vector<float> p3(g.points[3], g.points[3] + n3);
    const float x = av;
    int i3 = 0;
    while (i3 < n3 - 2 && p3[i3 + 1] < x) ++i3;
    const float g0 = p3[i3], g1 = p3[i3 + 1];
    const float iw = 1.0f / (g1 - g0);
    const float t3 = (x - g0) * iw;
    d.av_oob = (x < p3.front() || x > p3.back() || !(x == x)) ? 1 : 0;
    const size_t n012 = nn / (size_t)n3;
    std::vector<float4> v3(n012);
    for (size_t j = 0; j < n012; ++j) {
      const float4 c0 = v[j * n3 + i3], c1 = v[j * n3 + i3 + 1];
      v3[j] = make_float4(std::fmaf(t3, c1.x - c0.x, c0.x), std::fmaf(t3, c1.y - c0.y, c0.y),
                          std::fmaf(t3, c1.z - c0.z, c0.z), std::fmaf(t3, c1.w - c0.w, c0.w));
    }
    float4 *dv3;
    if ((rc = dev_upload(h, v3, &dv3))) return rc;
    d.values3 = dv3;
  }
  return CSI_OK;
}

// -------------------------------------------------------------------------------------------------
// create / destroy
// -------------------------------------------------------------------------------------------------
extern "C" int csi_create(const csi_config *cfg, csi_handle **out) {
  if (!cfg || !out) return fail(CSI_EINVAL, "null argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(CSI_ENODEVICE, "no CUDA device: libcsi has no CPU fallback");
  if (cfg->device < 0 || cfg->device >= ndev) return fail(CSI_EINVAL, "device ordinal out of range");
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10)
    return fail(CSI_ENODEVICE, std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
                                   "; libcsi is built for sm_100a only");
  CK(cudaSetDevice(cfg->device));
  csi_handle *h = new csi_handle();
  h->device = cfg->device;
  h->sm_count = prop.multiProcessorCount;
  int rc = build_model(*cfg, h);
  if (rc == CSI_OK) rc = check_net(cfg->emulator, "emulator");
  if (rc == CSI_OK && (cfg->emulator.n_in != 5 || cfg->emulator.n_out != 4))
    rc = fail(CSI_EINVAL, "emulator must map 5 inputs to 4 outputs");
  if (rc == CSI_OK) rc = upload_net(h, cfg->emulator, h->emu);
  if (rc == CSI_OK && cfg->bc_net.n_layers > 0) {
    rc = check_net(cfg->bc_net, "bc_net");
    if (rc == CSI_OK && cfg->bc_net.n_in != 4) rc = fail(CSI_EINVAL, "bc_net must take 4 inputs");
    if (rc == CSI_OK && cfg->bands > 0 && cfg->bc_mode == CSI_BC_NET && cfg->bc_net.n_out < cfg->bands)
      rc = fail(CSI_EINVAL, "bc_net has fewer outputs than bands");
    if (rc == CSI_OK) rc = upload_net(h, cfg->bc_net, h->bcnet);
  } else if (rc == CSI_OK && cfg->bands > 0 && cfg->bc_mode == CSI_BC_NET) {
    rc = fail(CSI_EINVAL, "bc_mode net without bc_net weights");
  }
  if (rc == CSI_OK && cfg->bc_grid.shape[0] > 0) {
    rc = upload_grid(h, cfg->bc_grid, std::max(cfg->bands, 1), (float)cfg->av, h->grid);
  } else if (rc == CSI_OK && cfg->bands > 0 && cfg->bc_mode == CSI_BC_GRID) {
    rc = fail(CSI_EINVAL, "bc_mode grid without a table");
  }
  h->engine = cfg->engine;
  if (rc == CSI_OK && cfg->engine != CSI_ENGINE_TCGEN05 && cfg->engine != CSI_ENGINE_FFMA)
    rc = fail(CSI_EINVAL, "unknown engine");
  if (rc == CSI_OK && cfg->engine == CSI_ENGINE_TCGEN05) {
    if (!(cfg->emulator.n_hidden == 128 && tc_supported(cfg->emulator)))
      rc = fail(CSI_EINVAL, "tcgen05 engine needs the 5->128x6->4 emulator with biases within 64x the largest weight");
    else if (cfg->bands > 0 && cfg->bc_mode == CSI_BC_NET)
      rc = fail(CSI_EINVAL, "tcgen05 engine: BC-emulator net inside the potential is not supported yet; use "
                            "CSI_ENGINE_FFMA or the BC table");
  }
  // The tensor-core weight images are built whenever the emulator has the supported shape: the potential uses them
  // with CSI_ENGINE_TCGEN05, the emulator operator (csi_net_forward / csi_net_vjp, which = 0) with either engine.
  if (rc == CSI_OK && cfg->emulator.n_hidden == 128 && tc_supported(cfg->emulator)) {
    {
      std::vector<uint8_t> img(tc_image_bytes(cfg->emulator)), resident(tc_resident_bytes(cfg->emulator));
      std::vector<float> small(tc_small_floats(cfg->emulator));
      tc_build_images(cfg->emulator, img.data(), resident.data(), small.data(), &h->tc.act_scale, &h->tc.w_scale);
      uint8_t *di, *dr; float *ds;
      rc = dev_upload(h, img, &di);
      if (rc == CSI_OK) rc = dev_upload(h, resident, &dr);
      if (rc == CSI_OK) rc = dev_upload(h, small, &ds);
      if (rc == CSI_OK) {
        h->tc.debug = getenv("CSI_TC_DEBUG") ? atoi(getenv("CSI_TC_DEBUG")) : 0;
        if (h->tc.debug & 8) {
          std::vector<long long> z(8 * 1024, 0);
          long long *dt;
          rc = dev_upload(h, z, &dt);
          h->tc.trace = dt;
        }
        h->tc.img = di; h->tc.resident = dr; h->tc.small = ds; h->tc.n_steps = 2 * (cfg->emulator.n_layers - 2);
      }
    }
  }
  // The BC emulator as an OPERATOR (csi_net_forward / csi_net_vjp, which = 1) runs on the same tensor-core kernel with
  // zero-padded images (4 -> 64 x 6 -> 3 inside the 5 -> 128 x 6 -> 4 layout, N = 64 MMAs over the first K-half only).
  // Inside the potential it stays on the CUDA-core engine: a second fused MLP would need its own ReLU-mask and
  // operand space in shared memory, which the stellar emulator's weight ring fills.
  if (rc == CSI_OK && cfg->bc_net.n_layers > 0 && tc_supported(cfg->bc_net)) {
    std::vector<uint8_t> img(tc_image_bytes(cfg->bc_net)), resident(tc_resident_bytes(cfg->bc_net));
    std::vector<float> small(tc_small_floats(cfg->bc_net));
    tc_build_images(cfg->bc_net, img.data(), resident.data(), small.data(), &h->tc_bc.act_scale, &h->tc_bc.w_scale);
    uint8_t *di, *dr; float *ds;
    rc = dev_upload(h, img, &di);
    if (rc == CSI_OK) rc = dev_upload(h, resident, &dr);
    if (rc == CSI_OK) rc = dev_upload(h, small, &ds);
    if (rc == CSI_OK) {
      h->tc_bc.img = di; h->tc_bc.resident = dr; h->tc_bc.small = ds; h->tc_bc.n_steps = 10;
      h->tc_bc.half_k = tc_half_k(cfg->bc_net) ? 1 : 0;
    }
  }
  if (rc == CSI_OK) {
    void *p;
    if (cudaMalloc(&p, sizeof(unsigned int)) != cudaSuccess) rc = fail(CSI_ECUDA, "cudaMalloc counter");
    else { h->owned.push_back(p); h->d_counter = (unsigned int *)p; cudaMemset(p, 0, sizeof(unsigned int)); }
  }
  if (rc == CSI_OK && cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess)
    rc = fail(CSI_ECUDA, "cudaStreamCreate");
  if (rc == CSI_OK) {
    void *p;
    if (cudaMalloc(&p, sizeof(unsigned long long)) != cudaSuccess) rc = fail(CSI_ECUDA, "cudaMalloc exchange epoch");
    else { h->owned.push_back(p); h->d_epoch = (unsigned long long *)p; cudaMemset(p, 0, sizeof(unsigned long long)); }
  }
  if (rc != CSI_OK) { std::string keep = g_err; csi_destroy(h); g_err = keep; return rc; }
  *out = h;
  return CSI_OK;
}

static void free_data(csi_handle *h) {
  for (void *p : {(void *)h->d_obs, (void *)h->d_diag, (void *)h->d_mhloc, (void *)h->d_mhscale, (void *)h->d_mhlogz,
                  (void *)h->d_partials, (void *)h->d_z, (void *)h->d_grad, (void *)h->d_out})
    if (p) cudaFree(p);
  if (h->h_z) cudaFreeHost(h->h_z);
  if (h->h_grad) cudaFreeHost(h->h_grad);
  if (h->h_out) cudaFreeHost(h->h_out);
  h->d_obs = h->d_diag = h->d_mhloc = h->d_mhscale = h->d_mhlogz = nullptr;
  h->d_partials = nullptr; h->d_z = h->d_grad = nullptr; h->d_out = nullptr;
  h->h_z = h->h_grad = nullptr; h->h_out = nullptr;
  h->n_local = 0;
}

extern "C" void csi_destroy(csi_handle *h) {
  if (!h) return;
  cudaSetDevice(h->device);
  free_data(h);
  for (void *p : h->owned) cudaFree(p);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

extern "C" int csi_set_shard(csi_handle *h, int32_t rank, int32_t world) {
  if (!h || world < 1 || rank < 0 || rank >= world) return fail(CSI_EINVAL, "bad shard");
  h->rank = rank; h->world = world;
  h->model.hyper_on = rank == 0 ? 1 : 0;
  return CSI_OK;
}

extern "C" int csi_set_peer_exchange(csi_handle *h, int32_t world, int32_t rank, void *const *peer_bufs) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  if (world <= 1 || !peer_bufs) { h->px = csi::PeerX{}; return CSI_OK; }  // off: out6 holds the local sums
  if (world > csi::kMaxPeers || rank < 0 || rank >= world) return fail(CSI_EINVAL, "peer exchange: world <= 8, 0 <= rank < world");
  csi::PeerX px{};
  for (int r = 0; r < world; ++r) {
    if (!peer_bufs[r]) return fail(CSI_EINVAL, "peer exchange: null buffer");
    px.bufs[r] = static_cast<double *>(peer_bufs[r]);
  }
  px.world = world; px.rank = rank; px.epoch = h->d_epoch;
  // a missing peer turns into NaN after this long (default 5 s; CSI_PEER_TIMEOUT_MS for tests)
  const char *tmo = getenv("CSI_PEER_TIMEOUT_MS");
  px.timeout_ns = (unsigned long long)(tmo && atof(tmo) > 0 ? atof(tmo) : 5000.0) * 1000000ull;
  CK(cudaSetDevice(h->device));
  CK(cudaMemset(h->d_epoch, 0, sizeof(unsigned long long)));  // every rank restarts the count with its (zeroed) buffer
  h->px = px;
  return CSI_OK;
}

extern "C" int csi_set_grid_limit(csi_handle *h, int32_t max_ctas) {
  if (!h || max_ctas < 0) return fail(CSI_EINVAL, "bad grid limit");
  h->grid_limit = max_ctas;
  return CSI_OK;
}

extern "C" int csi_layout(const csi_handle *h, int32_t *n_hyper, int32_t *n_star_sites, const char **names) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  if (n_hyper) *n_hyper = h->n_hyper;
  if (n_star_sites) *n_star_sites = h->n_sites;
  if (names)
    for (int i = 0; i < h->n_hyper + h->n_sites; ++i) names[i] = h->names[i];
  return CSI_OK;
}

extern "C" int csi_set_data(csi_handle *h, int64_t n, const double *obs, const double *diag, const double *mh_loc,
                            const double *mh_scale) {
  if (!h || !obs) return fail(CSI_EINVAL, "null argument");
  if (n < 1) return fail(CSI_EINVAL, "n_local must be >= 1");
  if (h->level != CSI_LEVEL_SINGLE && (long long)n * h->world < 2)
    return fail(CSI_EINVAL, "Variable num_stars must be greater than 1.");
  if (h->level == CSI_LEVEL_SINGLE && n != 1) return fail(CSI_EINVAL, "SingleStarModel takes exactly one star");
  if ((mh_loc == nullptr) != (mh_scale == nullptr)) return fail(CSI_EINVAL, "mh_loc and mh_scale go together");
  CK(cudaSetDevice(h->device));
  free_data(h);
  const int k = h->model.n_obs;
  std::vector<float> t((size_t)k * n);
  for (int64_t i = 0; i < n; ++i)
    for (int j = 0; j < k; ++j) t[(size_t)j * n + i] = (float)obs[i * k + j];
  CK(cudaMalloc(&h->d_obs, t.size() * sizeof(float)));
  CK(cudaMemcpy(h->d_obs, t.data(), t.size() * sizeof(float), cudaMemcpyHostToDevice));
  h->has_diag = diag != nullptr;
  if (diag) {
    for (int64_t i = 0; i < n; ++i)
      for (int j = 0; j < k; ++j) {
        double d = diag[i * k + j];
        if (!(d >= 0)) return fail(CSI_EINVAL, "diag must be >= 0");
        t[(size_t)j * n + i] = (float)d;
      }
    CK(cudaMalloc(&h->d_diag, t.size() * sizeof(float)));
    CK(cudaMemcpy(h->d_diag, t.data(), t.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  h->has_mh = mh_loc != nullptr;
  if (mh_loc) {
    std::vector<float> a(n), b(n), c(n);
    const TnC &tn = h->model.mh;
    for (int64_t i = 0; i < n; ++i) {
      if (!(mh_scale[i] > 0)) return fail(CSI_EINVAL, "mh_scale must be > 0");
      // device arrays: loc, 1 / scale, logz + log(scale)  (see StarIO::mh_*)
      a[i] = (float)mh_loc[i]; b[i] = (float)(1.0 / mh_scale[i]);
      c[i] = (float)(tn_logz(mh_loc[i], mh_scale[i], tn.kind & 1 ? (double)tn.low : -INFINITY,
                             tn.kind & 2 ? (double)tn.high : INFINITY) + log(mh_scale[i]));
    }
    CK(cudaMalloc(&h->d_mhloc, n * sizeof(float)));
    CK(cudaMalloc(&h->d_mhscale, n * sizeof(float)));
    CK(cudaMalloc(&h->d_mhlogz, n * sizeof(float)));
    CK(cudaMemcpy(h->d_mhloc, a.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->d_mhscale, b.data(), n * sizeof(float), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->d_mhlogz, c.data(), n * sizeof(float), cudaMemcpyHostToDevice));
  }
  h->partial_cap = max_partial_blocks(h->sm_count, n);
  CK(cudaMalloc(&h->d_partials, (size_t)h->partial_cap * 8 * sizeof(double)));
  CK(cudaMemset(h->d_partials, 0, (size_t)h->partial_cap * 8 * sizeof(double)));
  h->stage_len = h->n_hyper + (long long)h->n_sites * n;
  CK(cudaMallocHost(&h->h_z, h->stage_len * sizeof(float)));
  CK(cudaMallocHost(&h->h_grad, h->stage_len * sizeof(float)));
  CK(cudaMallocHost(&h->h_out, 8 * sizeof(double)));
  CK(cudaMalloc(&h->d_z, h->stage_len * sizeof(float)));
  CK(cudaMalloc(&h->d_grad, h->stage_len * sizeof(float)));
  CK(cudaMalloc(&h->d_out, 8 * sizeof(double)));
  h->n_local = n;
  return CSI_OK;
}

// -------------------------------------------------------------------------------------------------
// evaluation
// -------------------------------------------------------------------------------------------------
static void fill_launch(const csi_handle *h, const float *z, double *out6, float *grad, Launch &L) {
  memset(&L, 0, sizeof(L));
  L.model = h->model;
  L.emu = h->emu; L.bcnet = h->bcnet; L.grid = h->grid; L.tc = h->tc;
  const long long n = h->n_local;
  StarIO &io = L.io;
  io.n = n;
  io.z_hyper = z;
  for (int s = 0; s < h->n_sites; ++s) {
    int slot = h->slot_of_site[s];
    io.z[slot] = z + h->n_hyper + (long long)s * n;
    io.g[slot] = grad ? grad + h->n_hyper + (long long)s * n : nullptr;
  }
  for (int j = 0; j < h->model.n_obs; ++j) {
    io.obs[j] = h->d_obs + (long long)j * n;
    io.diag[j] = h->has_diag ? h->d_diag + (long long)j * n : nullptr;
    io.obsm[h->model.obs_idx[j]] = io.obs[j];
    io.diagm[h->model.obs_idx[j]] = io.diag[j];
  }
  if (h->has_mh) { io.mh_loc = h->d_mhloc; io.mh_scale = h->d_mhscale; io.mh_logz = h->d_mhlogz; }
  L.partials = h->d_partials;
  L.counter = h->d_counter;
  L.out6 = out6;
  L.grad_hyper = (grad && h->n_hyper) ? grad : nullptr;
  L.px = h->px;
  L.sm_count = h->sm_count;
  L.stop = h->stop_flag;
  L.grid_limit = h->grid_limit;
}

// The stream-ordered entry points never change the caller's current device; they refuse to launch on another one
// (the handle's pointers belong to h->device, and the null stream is the current device's).
static int check_device(const csi_handle *h) {
  int cur = -1;
  if (cudaGetDevice(&cur) != cudaSuccess || cur != h->device)
    return fail(CSI_ESTATE, "current CUDA device " + std::to_string(cur) + " is not the handle's device " +
                                std::to_string(h->device) + ": call cudaSetDevice first");
  return CSI_OK;
}

extern "C" int csi_potential(csi_handle *h, const float *z_dev, double *out_dev, float *grad_dev, void *stream) {
  if (!h || !z_dev || !out_dev || !grad_dev) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  Launch L;
  fill_launch(h, z_dev, out_dev, grad_dev, L);
  cudaStream_t st = (cudaStream_t)stream;
  int rc = h->engine == CSI_ENGINE_TCGEN05 ? launch_potential_tc(L, st, &h->last_grid, &h->last_smem)
                                           : launch_potential_ffma(L, st, &h->last_grid, &h->last_smem);
  if (rc < 0) return fail(CSI_ECUDA, std::string("potential launch: ") + cudaGetErrorString((cudaError_t)(-rc)));
  h->last_launches = rc;
  return CSI_OK;
}

// device-visible alias of a pinned / registered host pointer (unified addressing), or null for pageable memory
static void *device_alias_of_host(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

extern "C" int csi_potential_host(csi_handle *h, const float *z_host, double *out_host, float *grad_host) {
  if (!h || !z_host || !out_host || !grad_host) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  CK(cudaSetDevice(h->device));
  const size_t bytes = (size_t)h->stage_len * sizeof(float);
  // Page-locked caller buffers: one kernel launch, no whole-buffer copy before or after it.  The gradient and the
  // 6 sums are written by the kernel straight into the caller's memory (posted PCIe writes) and it reads the sites
  // over PCIe itself (its input staging runs half a tile ahead).  Pageable buffers go through the handle's pinned
  // staging area.  (Copy-engine staging of the sites behind a progress word was measured slower and removed:
  // profiles/r1_host_gate_ab.txt)
  const float *z_alias = static_cast<const float *>(device_alias_of_host(z_host));
  float *grad_alias = static_cast<float *>(device_alias_of_host(grad_host));
  const float *z_in = z_alias;
  if (!z_alias) {
    memcpy(h->h_z, z_host, bytes);
    CK(cudaMemcpyAsync(h->d_z, h->h_z, bytes, cudaMemcpyHostToDevice, h->stream));
    z_in = h->d_z;
  }
  double *out_alias = static_cast<double *>(device_alias_of_host(h->h_out));
  int rc = csi_potential(h, z_in, out_alias ? out_alias : h->d_out, grad_alias ? grad_alias : h->d_grad, h->stream);
  if (rc != CSI_OK) { cudaStreamSynchronize(h->stream); return rc; }
  if (!grad_alias) CK(cudaMemcpyAsync(h->h_grad, h->d_grad, bytes, cudaMemcpyDeviceToHost, h->stream));
  if (!out_alias) CK(cudaMemcpyAsync(h->h_out, h->d_out, 6 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  CK(cudaStreamSynchronize(h->stream));
  if (!grad_alias) memcpy(grad_host, h->h_grad, bytes);
  memcpy(out_host, h->h_out, 6 * sizeof(double));
  return CSI_OK;
}

extern "C" int csi_predict(csi_handle *h, const float *z_dev, float *mean_dev, float *params_dev, void *stream) {
  if (!h || !z_dev) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  Launch L;
  fill_launch(h, z_dev, nullptr, nullptr, L);
  L.stop = nullptr;
  L.mean_out = mean_dev; L.params_out = params_dev;
  int rc = h->engine == CSI_ENGINE_TCGEN05 ? launch_predict_tc(L, (cudaStream_t)stream)
                                           : launch_predict_ffma(L, (cudaStream_t)stream);
  if (rc < 0) return fail(CSI_ECUDA, std::string("predict launch: ") + cudaGetErrorString((cudaError_t)(-rc)));
  return CSI_OK;
}

static const NetDev *pick_net(csi_handle *h, int which) {
  if (which == 0) return &h->emu;
  if (which == 1 && h->bcnet.n_layers > 0) return &h->bcnet;
  return nullptr;
}

static const int64_t kNetTcMinRows = 2048;

extern "C" int csi_net_forward(csi_handle *h, int32_t which, const float *x, float *y, int64_t n, void *stream) {
  if (!h || !x || !y || n < 0) return fail(CSI_EINVAL, "bad argument");
  const NetDev *net = pick_net(h, which);
  if (!net) return fail(CSI_EINVAL, "network not loaded");
  int rc;
  const TcWeights &tcw = which == 0 ? h->tc : h->tc_bc;
  if (tcw.img && n >= kNetTcMinRows) {
    // large batches: the tcgen05 kernel in operator mode (small ones keep the CUDA-core kernel and its plain fp32
    // k-order accumulation, which the single-point known-answer tests pin to 5e-7)
    Launch L;
    memset(&L, 0, sizeof(L));
    L.tc = tcw; L.io.n = n; L.net_x = x; L.net_y = y; L.sm_count = h->sm_count;
    L.net_nin = net->n_in; L.net_nout = net->n_out;
    rc = launch_net_tc(L, false, (cudaStream_t)stream);
  } else {
    rc = launch_net_forward(*net, x, y, n, (cudaStream_t)stream);
  }
  if (rc < 0) return fail(CSI_ECUDA, std::string("net_forward: ") + cudaGetErrorString((cudaError_t)(-rc)));
  return CSI_OK;
}

extern "C" int csi_net_vjp(csi_handle *h, int32_t which, const float *x, const float *gy, float *gx, int64_t n,
                           void *stream) {
  if (!h || !x || !gy || !gx || n < 0) return fail(CSI_EINVAL, "bad argument");
  const NetDev *net = pick_net(h, which);
  if (!net) return fail(CSI_EINVAL, "network not loaded");
  int rc;
  const TcWeights &tcw = which == 0 ? h->tc : h->tc_bc;
  if (tcw.img && n >= kNetTcMinRows) {
    Launch L;
    memset(&L, 0, sizeof(L));
    L.tc = tcw; L.io.n = n; L.net_x = x; L.net_gy = gy; L.net_gx = gx; L.sm_count = h->sm_count;
    L.net_nin = net->n_in; L.net_nout = net->n_out;
    rc = launch_net_tc(L, true, (cudaStream_t)stream);
  } else {
    rc = launch_net_vjp(*net, x, gy, gx, n, (cudaStream_t)stream);
  }
  if (rc < 0) return fail(CSI_ECUDA, std::string("net_vjp: ") + cudaGetErrorString((cudaError_t)(-rc)));
  return CSI_OK;
}

extern "C" int csi_grid_forward(csi_handle *h, const float *x, float *y, int64_t n, void *stream) {
  if (!h || !x || !y || n < 0) return fail(CSI_EINVAL, "bad argument");
  if (!h->grid.values) return fail(CSI_EINVAL, "no BC table loaded");
  int rc = launch_grid_forward(h->grid, std::max(h->model.bands, 1), x, y, n, (cudaStream_t)stream);
  if (rc < 0) return fail(CSI_ECUDA, std::string("grid_forward: ") + cudaGetErrorString((cudaError_t)(-rc)));
  return CSI_OK;
}

extern "C" int csi_grid_vjp(csi_handle *h, const float *x, const float *gy, float *gx, int64_t n, void *stream) {
  if (!h || !x || !gy || !gx || n < 0) return fail(CSI_EINVAL, "bad argument");
  if (!h->grid.values) return fail(CSI_EINVAL, "no BC table loaded");
  int rc = launch_grid_vjp(h->grid, std::max(h->model.bands, 1), x, gy, gx, n, (cudaStream_t)stream);
  if (rc < 0) return fail(CSI_ECUDA, std::string("grid_vjp: ") + cudaGetErrorString((cudaError_t)(-rc)));
  return CSI_OK;
}

// -------------------------------------------------------------------------------------------------
// helpers
// -------------------------------------------------------------------------------------------------
extern "C" int csi_dev_alloc(csi_handle *h, int64_t bytes, void **out) {
  if (!h || !out || bytes < 0) return fail(CSI_EINVAL, "bad argument");
  CK(cudaSetDevice(h->device));
  CK(cudaMalloc(out, std::max<int64_t>(bytes, 16)));
  return CSI_OK;
}
extern "C" int csi_dev_free(csi_handle *h, void *p) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  CK(cudaFree(p));
  return CSI_OK;
}
extern "C" int csi_memcpy_h2d(csi_handle *h, void *dst, const void *src, int64_t bytes) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  CK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return CSI_OK;
}
extern "C" int csi_memcpy_d2h(csi_handle *h, void *dst, const void *src, int64_t bytes) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  CK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return CSI_OK;
}
extern "C" int csi_synchronize(csi_handle *h) {
  if (!h) return fail(CSI_EINVAL, "null handle");
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  return CSI_OK;
}

// debugging aid (not part of the public header): copies the clock64 trace of CTA 0 recorded under CSI_TC_DEBUG=8
extern "C" int csi_tc_trace(csi_handle *h, long long *out, int n) {
  if (!h || !h->tc.trace) return fail(CSI_ESTATE, "no trace");
  CK(cudaMemcpy(out, h->tc.trace, sizeof(long long) * n, cudaMemcpyDeviceToHost));
  return CSI_OK;
}

extern "C" int csi_info(const csi_handle *h, int64_t info[8]) {
  if (!h || !info) return fail(CSI_EINVAL, "null argument");
  info[0] = h->sm_count; info[1] = h->engine; info[2] = h->last_launches; info[3] = h->n_local;
  info[4] = h->last_smem; info[5] = h->last_grid; info[6] = h->rank; info[7] = h->world;
  return CSI_OK;
}

extern "C" int csi_time_potential(csi_handle *h, const float *z_dev, double *out_dev, float *grad_dev, int32_t iters,
                                  float *ms_per_call) {
  if (!h || !ms_per_call || iters < 1) return fail(CSI_EINVAL, "bad argument");
  CK(cudaSetDevice(h->device));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  CK(cudaStreamSynchronize(h->stream));
  CK(cudaEventRecord(e0, h->stream));
  for (int i = 0; i < iters; ++i) {
    int rc = csi_potential(h, z_dev, out_dev, grad_dev, h->stream);
    if (rc != CSI_OK) return rc;
  }
  CK(cudaEventRecord(e1, h->stream));
  CK(cudaEventSynchronize(e1));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ms_per_call = ms / iters;
  return CSI_OK;
}

#include "csi_nuts.cuh"  // NUTS leapfrog / U-turn vector kernels (csi_leapfrog_pre, csi_leapfrog_post, csi_uturn)

// ---------------------------------------------------------------------------------------------------------------
// csi_uturn_static: csi_uturn for callers that may not allocate (no counterpart in the reference).
// The block partials and the ticket of the fixed-order grid reduction live in a
// file-scope device array instead of the handle's cudaMalloc'ed scratch, so the call does no cudaMalloc / cudaFree.
// Every GPU that loads the library holds the array (CSI_STATIC_SCRATCH_BYTES) until the process exits, and every
// call on a device shares it: calls that overlap on the GPU race.
// ---------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kStaticBlocks = 1024;                               // >= 4 CTAs per SM on a B200 (592)
constexpr int kStaticTicketDoubles = 32;                          // the ticket gets its own 256 bytes at the end
constexpr size_t kStaticDoubles = (size_t)kStaticBlocks * csi::nuts::kNV + kStaticTicketDoubles;
static_assert(kStaticDoubles * sizeof(double) == CSI_STATIC_SCRATCH_BYTES, "csi.h states the size of the array");
constexpr int kStaticMaxDevices = 64;
}  // namespace

static __device__ __align__(256) double g_uturn_static_scratch[kStaticDoubles];

// In host code `g_uturn_static_scratch` is only a stand-in for the symbol: the device address has to be asked for,
// once per device, with that device current.
static double *g_uturn_static_ptr[kStaticMaxDevices];
static std::mutex g_uturn_static_mu;

static int uturn_static_scratch(int device, double **out) {
  if (device < 0 || device >= kStaticMaxDevices) return fail(CSI_EINVAL, "csi_uturn_static: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_uturn_static_mu);
  if (!g_uturn_static_ptr[device]) {
    double *p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, g_uturn_static_scratch));   // the caller checked that `device` is current
    if (!p) return fail(CSI_ECUDA, "csi_uturn_static: cudaGetSymbolAddress returned a null address");
    g_uturn_static_ptr[device] = p;
  }
  *out = g_uturn_static_ptr[device];
  return CSI_OK;
}

extern "C" int csi_uturn_static(csi_handle *h, const float *r_left, const float *r_right, const float *r_sum,
                                const float *inv_mass, double *dots4, void *stream) {
  if (!h || !r_left || !r_right || !r_sum || !inv_mass || !dots4) return fail(CSI_EINVAL, "null argument");
  if (h->n_local < 1) return fail(CSI_ESTATE, "csi_set_data has not been called");
  if (int rcd = check_device(h)) return rcd;
  const long long n = h->stage_len;
  if (n > (long long)CSI_UTURN_STATIC_MAX_LEN)
    return fail(CSI_EINVAL, "csi_uturn_static: vector length " + std::to_string(n) + " exceeds CSI_UTURN_STATIC_MAX_LEN (" +
                                std::to_string((long long)CSI_UTURN_STATIC_MAX_LEN) + "); use csi_uturn");
  double *scratch = nullptr;
  if (int rcs = uturn_static_scratch(h->device, &scratch)) return rcs;
  unsigned int *ticket = reinterpret_cast<unsigned int *>(scratch + (size_t)kStaticBlocks * csi::nuts::kNV);
  // the array keeps the last call's contents: the partials are overwritten before they are read, the ticket must
  // start at zero (the finalising block resets it, but a launch that was aborted would not have)
  CK(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), (cudaStream_t)stream));
  const int grid = std::min(csi::nuts::grid_for(h, n), kStaticBlocks);   // same grid as csi_uturn -> same sums
  csi::nuts::uturn_kernel<<<grid, csi::nuts::kBlock, 0, (cudaStream_t)stream>>>(
      n, h->n_hyper, r_left, r_right, r_sum, inv_mass, dots4, scratch, ticket);
  CK(cudaGetLastError());
  return CSI_OK;
}
