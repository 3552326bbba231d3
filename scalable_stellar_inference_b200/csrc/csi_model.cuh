// csi_model.cuh -- per-star device arithmetic of the NUTS potential (shared by both MLP engines).
//
// One CUDA thread owns one star.  The evaluation of a star is split in three phases around the emulator
// MLP so that the MLP engine (tcgen05 tiles or CUDA-core FFMA) can sit in between:
//
//   star_pre  : unconstrained sites -> constrained parameters, prior log-density + log|J| and its
//               gradient, emulator inputs x[5]               (numpyro_models.py:148-201 / :299-339)
//   [MLP fwd] : x[5] -> o[4]                                 (_jax/emulator.py:44-50)
//   star_mid  : predict + bolometric corrections + likelihood, cotangent of o
//                                                            (numpyro_models.py:203-240, _jax/star.py:39-65)
//   [MLP vjp] : g_o[4] -> g_x[5]
//   star_post : chain rule back to the unconstrained sites, gradient stores
//
// numpyro's transforms / log-densities (third-party, absent from the reference tree) are restated from
// their closed forms; see oracle/dists.py for the same list with the formulas spelled out.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define CSI_DEV __device__ __forceinline__

namespace csi {

constexpr int kMaxMean = 12;
constexpr int kMaxObs = 12;
constexpr int kMaxBands = 4;
constexpr int kSlots = 9;  // internal per-star site slots
enum Slot { S_TAU = 0, S_EVOL = 1, S_MASS = 2, S_MH = 3, S_Y = 4, S_A = 5, S_PY = 6, S_PA = 7, S_SD = 8 };
enum PriorMode { P_FLAT = 0, P_INFORMATIVE = 1, P_HIER = 2 };

constexpr float kLn10 = 2.302585092994046f;
constexpr float kLog2_10 = 3.321928094887362f;  // |argument| < 2 where it is used: 10^v = 2^(v log2 10) to ~1 ulp
constexpr float kHalfLog2Pi = 0.9189385332046727f;
constexpr float kInvSqrt2 = 0.7071067811865476f;

// support kinds of numpyro's biject_to
enum TnKind { TN_REAL = 0, TN_LOWER = 1, TN_UPPER = 2, TN_INTERVAL = 3 };

struct TnC {            // TruncatedNormal with constant bounds (loc/scale may be overridden per star)
  float loc, scale, low, high;
  float logz;           // log(Phi(beta) - Phi(alpha)) for the constant loc/scale (host, fp64)
  float inv_scale;      // 1 / scale and
  float lognorm;        // logz + log(scale), both formed in fp64 on the host: no per-star divide / logf for them
  float logw;           // log(high - low) for the interval transform
  int kind;
};

struct DevModel {
  int prior_mode, kind, n_obs, bands, bc_mode, n_mean, hyper_on, pad0;
  int col_of_mean[kMaxMean];  // observation column that observes mean index m, or -1
  int obs_idx[kMaxObs];
  TnC log_evol, log_mass, mh, dydz, mu_a;
  float y_lo, y_hi, y_w, log_y_w, a_lo, a_hi, a_w, log_a_w;
  float y0_loc, y0_scale;
  float py_conc, py_rate, py_const, pa_conc, pa_rate, pa_const;  // const = c*log(rate) - lgamma(c)
  float tau_conc, tau_rate, tau_const;
  float cc[7];                 // c + A @ delta_loc   (numpyro_models.py:205-209)
  float var_ext[kMaxMean];     // diag(A cov A^T), extended with the photometric rows
  float cov_ext[kMaxMean * kMaxMean];
  float sd_c1, sd_c0, sd_const, max_distance, av;
  float log_zx_sun, bol_mag_sun;
  float inf_dydz, inf_y0, inf_s1, inf_s0, inf_mua_loc, inf_mua_scale;
};

struct StarIO {
  const float *z[kSlots];
  float *g[kSlots];
  const float *obs[kMaxObs];
  const float *diag[kMaxObs];   // entries may be null (= 0)
  const float *obsm[kMaxMean];  // the same columns indexed by the MEAN they observe (null = unobserved); kind diag
  const float *diagm[kMaxMean];
  const float *mh_loc, *mh_scale, *mh_logz;  // per-star M_H prior: loc, 1 / scale, logz + log(scale); null = DevModel::mh
  const float *z_hyper;         // 5 values (hierarchical only)
  long long n;
};

// 4-D BC table, float4-padded bands
constexpr int kLutBins = 256;
struct GridDev {
  int shape[4];
  const float *pts[4];
  const unsigned short *lut[4];  // [kLutBins] per axis: an interval index <= the answer for every x of the bin
  float lut_x0[4], lut_inv[4];   // bin = (x - lut_x0) * lut_inv
  const float4 *values;  // [n0*n1*n2*n3]
  // The potential always evaluates the table at the model constant Av (numpyro_models `const`), so the table is
  // interpolated along axis 3 ONCE at creation (same fp32 arithmetic as grid_eval's first level): the per-star work
  // is then a 3-D interpolation, 8 gathers instead of 16.  null when not built; av_oob = Av outside the axis.
  const float4 *values3;  // [n0*n1*n2]
  int av_oob;
};

// ---------------------------------------------------------------------------------------------
// scalar helpers
// ---------------------------------------------------------------------------------------------
// Fast forms for operands that are normal numbers of moderate size (MUFU-based, 1-2 ulp, no slow-path branch: an
// IEEE divide or a libm call splits the basic block and serialises the otherwise independent per-site chains).
CSI_DEV float rcp_(float x) { return __fdividef(1.0f, x); }
CSI_DEV float exp_(float x) { return exp2f(x * 1.4426950408889634f); }
CSI_DEV float log_(float x) { return __logf(x); }
CSI_DEV float sigmoidf_(float u) { return rcp_(1.0f + exp_(-u)); }

// log(s(1-s)) = -softplus(u) - softplus(-u) = -|u| - 2 log1p(exp(-|u|))
CSI_DEV float log_sigmoid_prod(float u) {
  float a = fabsf(u);
  return -a - 2.0f * log1pf(expf(-a));
}

// sigmoid(u) together with l1 = log(1 + exp(-|u|)) from ONE exponential:
//   log(s (1 - s)) = -|u| - 2 l1,   log s = min(u, 0) - l1,   log(1 - s) = -max(u, 0) - l1
CSI_DEV void sig_pair(float u, float &s, float &l1) {
  const float a = fabsf(u);
  const float e = exp_(-a);
  const float w = 1.0f + e;
  const float inv = rcp_(w);
  s = (u >= 0.0f) ? inv : e * inv;
  l1 = log_(w);  // e in (0, 1]: absolute error ~2e-7 on a term added to O(1) log-densities
}

// log Phi(x), x <= 0 branch uses the scaled complementary error function (no underflow)
CSI_DEV float log_ndtr(float x) {
  if (x <= 0.0f) {
    float t = -x * kInvSqrt2;
    return logf(0.5f * erfcxf(t)) - t * t;
  }
  return log1pf(-0.5f * erfcf(x * kInvSqrt2));
}

// log(Phi(beta) - Phi(alpha)), alpha < beta, either may be infinite.  Also phi(alpha)/Z and phi(beta)/Z.
// (oracle/dists.py log_ndtr_diff; numpyro evaluates the same difference on the tail nearer to loc.)
CSI_DEV void log_ndtr_diff(float alpha, float beta, float &logz, float &ra, float &rb) {
  bool flip = alpha > 0.0f;
  float a = flip ? -beta : alpha;
  float b = flip ? -alpha : beta;
  if (b <= 0.0f) {
    float lb = log_ndtr(b);
    float la = isinf(a) ? -INFINITY : log_ndtr(a);
    logz = lb + log1pf(-expf(la - lb));
  } else {
    float ta = isinf(a) ? 0.0f : 0.5f * erfcf(-a * kInvSqrt2);
    float tb = isinf(b) ? 0.0f : 0.5f * erfcf(b * kInvSqrt2);
    logz = log1pf(-(ta + tb));
  }
  ra = isinf(alpha) ? 0.0f : expf(-0.5f * alpha * alpha - kHalfLog2Pi - logz);
  rb = isinf(beta) ? 0.0f : expf(-0.5f * beta * beta - kHalfLog2Pi - logz);
}

// numpyro biject_to: value, dvalue/du, log|J|, dlog|J|/du
CSI_DEV void bounded_transform(const TnC &c, float u, float &v, float &dv, float &ld, float &dld) {
  if (c.kind == TN_INTERVAL) {
    float s, l1;
    sig_pair(u, s, l1);
    float w = c.high - c.low;
    v = c.low + w * s;
    dv = w * s * (1.0f - s);
    ld = c.logw - fabsf(u) - 2.0f * l1;
    dld = 1.0f - 2.0f * s;
  } else if (c.kind == TN_UPPER) {
    float e = exp_(u);
    v = c.high - e; dv = -e; ld = u; dld = 1.0f;
  } else if (c.kind == TN_LOWER) {
    float e = exp_(u);
    v = c.low + e; dv = e; ld = u; dld = 1.0f;
  } else {
    v = u; dv = 1.0f; ld = 0.0f; dld = 0.0f;
  }
}

// constant-bound TruncatedNormal site in unconstrained space: value, dv/du, log-term, dlog-term/du.
// loc / scale / logz may be per-star overrides (the M_H prior of scripts/model_star.py:126-128).
// inv_scale = 1 / scale, lognorm = logz + log(scale) (TnC::inv_scale / lognorm or the per-star arrays)
CSI_DEV void tn_site(const TnC &c, float u, float loc, float inv_scale, float lognorm, float &v, float &dv,
                     float &lt, float &dlt) {
  float ld, dld;
  bounded_transform(c, u, v, dv, ld, dld);
  float r = (v - loc) * inv_scale;
  lt = -0.5f * r * r - kHalfLog2Pi - lognorm + ld;
  dlt = -r * inv_scale * dv + dld;
}

// log TN(x; mu, sigma, low, high) and partials wrt x, mu, sigma (finite bounds)
// (inv = 1 / sigma and logsig = log sigma are passed in: they are per-evaluation constants in the hierarchical model)
CSI_DEV void tn_logpdf_grads_pre(float x, float mu, float inv, float logsig, float low, float high, float &lp,
                                 float &dx, float &dmu, float &dsig) {
  float alpha = (low - mu) * inv, beta = (high - mu) * inv;
  float logz, ra, rb;
  log_ndtr_diff(alpha, beta, logz, ra, rb);
  float r = (x - mu) * inv;
  lp = -0.5f * r * r - logsig - kHalfLog2Pi - logz;
  float dlogz_dmu = -(rb - ra) * inv;
  float dlogz_dsig = -(beta * rb - alpha * ra) * inv;
  dx = -r * inv;
  dmu = r * inv - dlogz_dmu;
  dsig = (r * r - 1.0f) * inv - dlogz_dsig;
}
CSI_DEV void tn_logpdf_grads(float x, float mu, float sigma, float low, float high, float &lp, float &dx,
                             float &dmu, float &dsig) {
  float inv = 1.0f / sigma;
  float alpha = (low - mu) * inv, beta = (high - mu) * inv;
  float logz, ra, rb;
  log_ndtr_diff(alpha, beta, logz, ra, rb);
  float r = (x - mu) * inv;
  lp = -0.5f * r * r - logf(sigma) - kHalfLog2Pi - logz;
  float dlogz_dmu = -(rb - ra) * inv;
  float dlogz_dsig = -(beta * rb - alpha * ra) * inv;
  dx = -r * inv;
  dmu = r * inv - dlogz_dmu;
  dsig = (r * r - 1.0f) * inv - dlogz_dsig;
}

// Gamma(conc, rate) site with exp transform: log-term (log_prob + u) and d/du
CSI_DEV void gamma_site(float u, float conc, float rate, float cst, float &x, float &lt, float &dlt) {
  x = exp_(u);
  lt = cst + (conc - 1.0f) * u - rate * x + u;  // log x = u
  dlt = (conc - 1.0f) - rate * x + 1.0f;
}

// ---------------------------------------------------------------------------------------------
// hyper-parameters (numpyro_models.py:277-297), evaluated once per block
// ---------------------------------------------------------------------------------------------
struct HyperVals {
  float y0, dydz, sigma_y, mu_a, sigma_a;
  float ddydz, dmu_a;   // dvalue/du of the two interval sites
  float prior_lp;       // sum of the 5 hyper log-terms
  float gprior[5];      // d prior_lp / du
  // truncation terms of a_MLT ~ TN(mu_a, sigma_a, a_lo, a_hi): they depend on the hyper-parameters only
  float a_inv, a_logsig, a_alpha, a_beta, a_logz, a_ra, a_rb;
  float y_inv, y_logsig;  // 1 / sigma_Y, log sigma_Y
};

CSI_DEV void hyper_eval(const DevModel &m, const float *zh, HyperVals &h) {
  float lp = 0.0f;
  h.y0 = zh[0];
  float r = (h.y0 - m.y0_loc) / m.y0_scale;
  lp += -0.5f * r * r - logf(m.y0_scale) - kHalfLog2Pi;
  h.gprior[0] = -r / m.y0_scale;
  float lt, dlt;
  tn_site(m.dydz, zh[1], m.dydz.loc, m.dydz.inv_scale, m.dydz.lognorm, h.dydz, h.ddydz, lt, dlt);
  lp += lt; h.gprior[1] = dlt;
  float py;
  gamma_site(zh[2], m.py_conc, m.py_rate, m.py_const, py, lt, dlt);
  lp += lt; h.gprior[2] = dlt;
  h.sigma_y = rsqrtf(py);
  tn_site(m.mu_a, zh[3], m.mu_a.loc, m.mu_a.inv_scale, m.mu_a.lognorm, h.mu_a, h.dmu_a, lt, dlt);
  lp += lt; h.gprior[3] = dlt;
  float pa;
  gamma_site(zh[4], m.pa_conc, m.pa_rate, m.pa_const, pa, lt, dlt);
  lp += lt; h.gprior[4] = dlt;
  h.sigma_a = rsqrtf(pa);
  h.prior_lp = lp;
  h.a_inv = 1.0f / h.sigma_a;
  h.a_logsig = logf(h.sigma_a);
  h.a_alpha = (m.a_lo - h.mu_a) * h.a_inv;
  h.a_beta = (m.a_hi - h.mu_a) * h.a_inv;
  log_ndtr_diff(h.a_alpha, h.a_beta, h.a_logz, h.a_ra, h.a_rb);
  h.y_inv = 1.0f / h.sigma_y;
  h.y_logsig = logf(h.sigma_y);
}

// ---------------------------------------------------------------------------------------------
// per-star state carried between the phases
// ---------------------------------------------------------------------------------------------
struct StarState {
  float x[5];        // evol, mass, M_H, Y, a_MLT  (emulator inputs)
  float log_mass, tau;
  float gp[kSlots];  // d(log prior + log|J|)/du per slot (prior part only)
  float c_evol, c_mass, dlm, dmh, dy, da;  // chain factors d(param)/du
  float dist, dsd;   // photometry: distance, d distance/du
  float lp;          // per-star log-density (prior, then + likelihood)
  // likelihood cotangents that bypass the emulator
  float g_logmass, g_mh_dir, g_tau_lik, g_dist;
};

struct HyperAcc {  // per-thread partial sums of dlogp/d(y0, dydz, sigma_y, mu_a, sigma_a)
  double lp;
  double g[5];
};

// the inputs of star_pre, loaded ahead of use so that their latency hides under the preceding MMA stage
struct StarIn {
  float zv[kSlots];
  float mloc, mscale, mlogz;
};
CSI_DEV void star_load_z(const DevModel &m, const StarIO &io, long long i, StarIn &in) {
#pragma unroll
  for (int s = 0; s < kSlots; ++s) in.zv[s] = io.z[s] ? io.z[s][i] : 0.0f;
  in.mloc = io.mh_loc ? io.mh_loc[i] : m.mh.loc;
  in.mscale = io.mh_scale ? io.mh_scale[i] : m.mh.inv_scale;
  in.mlogz = io.mh_logz ? io.mh_logz[i] : m.mh.lognorm;
}

// The tcgen05 kernel stages a star's inputs through a per-thread column of shared memory with asynchronous copies
// issued most of a tile ahead (csi_tc.cu): rows 0..8 the sites and 9..11 the [Fe/H] prior while the priors are
// pending, then rows k / kMaxMean + k the observation / extra variance of mean k until star_mid has read them.
constexpr int kStageRows = 2 * kMaxMean;
constexpr int kStageStride = 256;  // one column per epilogue thread
// (all rows are read unconditionally -- independent LDS in one basic block -- and absent inputs are replaced by
// a select afterwards: a never-written row may hold any bit pattern, a select does not propagate it)
CSI_DEV void star_take_z(const DevModel &m, const StarIO &io, const float *col, StarIn &in) {
  float v[kSlots + 3];
#pragma unroll
  for (int s = 0; s < kSlots + 3; ++s) v[s] = col[s * kStageStride];
#pragma unroll
  for (int s = 0; s < kSlots; ++s) in.zv[s] = io.z[s] ? v[s] : 0.0f;
  in.mloc = io.mh_loc ? v[9] : m.mh.loc;
  in.mscale = io.mh_scale ? v[10] : m.mh.inv_scale;
  in.mlogz = io.mh_logz ? v[11] : m.mh.lognorm;
}

// the arrays a tile copies into its staging column: global base pointer (star 0) and row, built once per CTA
struct StageList {
  const float *src[kStageRows];
  int row[kStageRows];
  int n;
};
CSI_DEV void stage_list_sites(const StarIO &io, StageList &l) {
  int c = 0;
  for (int s = 0; s < kSlots; ++s)
    if (io.z[s]) { l.src[c] = io.z[s]; l.row[c++] = s; }
  if (io.mh_loc) { l.src[c] = io.mh_loc; l.row[c++] = 9; }
  if (io.mh_scale) { l.src[c] = io.mh_scale; l.row[c++] = 10; }
  if (io.mh_logz) { l.src[c] = io.mh_logz; l.row[c++] = 11; }
  l.n = c;
}
CSI_DEV void stage_list_obs(const DevModel &m, const StarIO &io, StageList &l) {
  int c = 0;
  if (m.kind == 0)
    for (int k = 0; k < kMaxMean; ++k)
      if (io.obsm[k]) {
        l.src[c] = io.obsm[k]; l.row[c++] = k;
        if (io.diagm[k]) { l.src[c] = io.diagm[k]; l.row[c++] = kMaxMean + k; }
      }
  l.n = c;
}

CSI_DEV void star_pre(const DevModel &m, const HyperVals &h, const StarIn &in, StarState &st, HyperAcc &acc) {
  float lp = 0.0f;
  const float *zv = in.zv;
  const float mloc = in.mloc, mscale = in.mscale, mlogz = in.mlogz;
#pragma unroll
  for (int s = 0; s < kSlots; ++s) st.gp[s] = 0.0f;
  // scaled_precision ~ Gamma(df/2, df/2)   (:195-196 / :301-302)
  float lt, dlt;
  gamma_site(zv[S_TAU], m.tau_conc, m.tau_rate, m.tau_const, st.tau, lt, dlt);
  lp += lt; st.gp[S_TAU] = dlt;
  // log_evol, log_mass, M_H: constant-bound truncated normals (:149-161 / :304-316)
  float v, dv;
  tn_site(m.log_evol, zv[S_EVOL], m.log_evol.loc, m.log_evol.inv_scale, m.log_evol.lognorm, v, dv, lt, dlt);
  lp += lt; st.gp[S_EVOL] = dlt;
  float evol = exp2f(v * kLog2_10);
  st.x[0] = evol; st.c_evol = kLn10 * evol * dv;
  tn_site(m.log_mass, zv[S_MASS], m.log_mass.loc, m.log_mass.inv_scale, m.log_mass.lognorm, v, dv, lt, dlt);
  lp += lt; st.gp[S_MASS] = dlt;
  float mass = exp2f(v * kLog2_10);
  st.log_mass = v; st.x[1] = mass; st.c_mass = kLn10 * mass * dv; st.dlm = dv;
  float mh;
  tn_site(m.mh, zv[S_MH], mloc, mscale, mlogz, mh, dv, lt, dlt);
  lp += lt; st.gp[S_MH] = dlt;
  st.x[2] = mh; st.dmh = dv;

  float t = exp2f(-(mh + m.log_zx_sun) * kLog2_10);
  float q = rcp_(1.0f + t);
  float dq_dmh = kLn10 * q * (1.0f - q);
  float g_mh_prior = 0.0f;

  if (m.prior_mode == P_HIER) {
    // decentred Y and a_MLT with star-/hyper-dependent bounds (:318-333); Y = ylo + w*sigmoid(u)
    float f = h.dydz * q;
    float inv1f = rcp_(1.0f + f);
    float mu_y = (h.y0 + f) * inv1f;
    float dmu_dy0 = inv1f;
    float dmu_df = (1.0f - h.y0) * inv1f * inv1f;
    float uy = zv[S_Y];
    float sy, l1y;
    sig_pair(uy, sy, l1y);
    float y = m.y_lo + m.y_w * sy;
    float dy = m.y_w * sy * (1.0f - sy);
    float lpy, dl_dy, dl_dmu, dl_dsig;
    tn_logpdf_grads_pre(y, mu_y, h.y_inv, h.y_logsig, m.y_lo, m.y_hi, lpy, dl_dy, dl_dmu, dl_dsig);
    lp += lpy + m.log_y_w - fabsf(uy) - 2.0f * l1y;
    st.gp[S_Y] = dl_dy * dy + (1.0f - 2.0f * sy);
    g_mh_prior += dl_dmu * dmu_df * h.dydz * dq_dmh;
    acc.g[0] += (double)(dl_dmu * dmu_dy0);
    acc.g[1] += (double)(dl_dmu * dmu_df * q);
    acc.g[2] += (double)dl_dsig;
    st.x[3] = y; st.dy = dy;
    float ua = zv[S_A];
    float sa, l1a;
    sig_pair(ua, sa, l1a);
    float a = m.a_lo + m.a_w * sa;
    float da = m.a_w * sa * (1.0f - sa);
    // TN(mu_a, sigma_a) has the same normaliser for every star: evaluated once in hyper_eval
    const float ra_ = (a - h.mu_a) * h.a_inv;
    const float lpa = -0.5f * ra_ * ra_ - h.a_logsig - kHalfLog2Pi - h.a_logz;
    const float dl_da = -ra_ * h.a_inv;
    const float dl_dmua = ra_ * h.a_inv + (h.a_rb - h.a_ra) * h.a_inv;
    const float dl_dsiga = (ra_ * ra_ - 1.0f) * h.a_inv + (h.a_beta * h.a_rb - h.a_alpha * h.a_ra) * h.a_inv;
    lp += lpa + m.log_a_w - fabsf(ua) - 2.0f * l1a;
    st.gp[S_A] = dl_da * da + (1.0f - 2.0f * sa);
    acc.g[3] += (double)dl_dmua;
    acc.g[4] += (double)dl_dsiga;
    st.x[4] = a; st.da = da;
  } else {
    float uy = zv[S_Y], ua = zv[S_A];
    float sy = sigmoidf_(uy), sa = sigmoidf_(ua);
    float y = m.y_lo + m.y_w * sy, a = m.a_lo + m.a_w * sa;
    float dy = m.y_w * sy * (1.0f - sy), da = m.a_w * sa * (1.0f - sa);
    float ldy = m.log_y_w + log_sigmoid_prod(uy), lda = m.log_a_w + log_sigmoid_prod(ua);
    float dldy = 1.0f - 2.0f * sy, dlda = 1.0f - 2.0f * sa;
    st.x[3] = y; st.dy = dy; st.x[4] = a; st.da = da;
    if (m.prior_mode == P_FLAT) {
      // Uniform priors (:163-165)
      lp += -m.log_y_w + ldy - m.log_a_w + lda;
      st.gp[S_Y] = dldy; st.gp[S_A] = dlda;
    } else {
      // informative (:167-193): per-star precision sites; enrichment-law uncertainty folded in
      float py, pa;
      gamma_site(zv[S_PY], m.py_conc, m.py_rate, m.py_const, py, lt, dlt);
      lp += lt; float g_upy = dlt;
      float sig_y0sq = 1.0f / py;
      float zxp1 = t + 1.0f;
      float fp1 = m.inf_dydz / zxp1 + 1.0f;
      float mu_y = (m.inf_y0 + m.inf_dydz / zxp1) / fp1;
      float f_dydz = (1.0f - m.inf_y0) / (fp1 * fp1) / zxp1;
      float f_y0 = 1.0f / fp1;
      float s1 = m.inf_s1, s0 = m.inf_s0;
      float var_y = sig_y0sq + (f_dydz * s1) * (f_dydz * s1) + (f_y0 * s0) * (f_y0 * s0);
      float sig_yt = sqrtf(var_y);
      float lpy, dl_dy, dl_dmu, dl_dsig;
      tn_logpdf_grads(y, mu_y, sig_yt, m.y_lo, m.y_hi, lpy, dl_dy, dl_dmu, dl_dsig);
      lp += lpy + ldy;
      st.gp[S_Y] = dl_dy * dy + dldy;
      float dz = -kLn10 * t;
      float dfp1 = -m.inf_dydz / (zxp1 * zxp1) * dz;
      float dmu = (1.0f - m.inf_y0) / (fp1 * fp1) * dfp1;
      float dfd = (1.0f - m.inf_y0) * (-2.0f * dfp1 / (fp1 * fp1 * fp1) / zxp1 - dz / (fp1 * fp1) / (zxp1 * zxp1));
      float dfy = -dfp1 / (fp1 * fp1);
      float dsig = (f_dydz * s1 * s1 * dfd + f_y0 * s0 * s0 * dfy) / sig_yt;
      g_mh_prior += dl_dmu * dmu + dl_dsig * dsig;
      st.gp[S_PY] = g_upy + dl_dsig * (-0.5f * sig_y0sq / sig_yt);
      gamma_site(zv[S_PA], m.pa_conc, m.pa_rate, m.pa_const, pa, lt, dlt);
      lp += lt; float g_upa = dlt;
      float sig_a0sq = 1.0f / pa;
      float sig_at = sqrtf(sig_a0sq + m.inf_mua_scale * m.inf_mua_scale);
      float lpa, dl_da, dl_dmua, dl_dsiga;
      tn_logpdf_grads(a, m.inf_mua_loc, sig_at, m.a_lo, m.a_hi, lpa, dl_da, dl_dmua, dl_dsiga);
      lp += lpa + lda;
      st.gp[S_A] = dl_da * da + dlda;
      st.gp[S_PA] = g_upa + dl_dsiga * (-0.5f * sig_a0sq / sig_at);
    }
  }
  st.gp[S_MH] += g_mh_prior * st.dmh;

  st.dist = 1.0f; st.dsd = 0.0f;
  if (m.bands > 0) {
    // scaled_distance ~ Beta(c1, c0) on (0,1), sigmoid transform (notebook/single-star-phot.ipynb cell [7])
    float u = zv[S_SD];
    float sd, l1;
    sig_pair(u, sd, l1);
    const float log_sd = fminf(u, 0.0f) - l1, log_1msd = -fmaxf(u, 0.0f) - l1;  // log s, log(1 - s)
    lp += m.sd_const + (m.sd_c1 - 1.0f) * log_sd + (m.sd_c0 - 1.0f) * log_1msd + (log_sd + log_1msd);
    st.gp[S_SD] = m.sd_c1 * (1.0f - sd) - m.sd_c0 * sd;
    st.dist = m.max_distance * sd;
    st.dsd = m.max_distance * sd * (1.0f - sd);
  }
  st.lp = lp;
  st.g_logmass = 0.0f; st.g_mh_dir = 0.0f; st.g_tau_lik = 0.0f; st.g_dist = 0.0f;
}

CSI_DEV void star_pre(const DevModel &m, const HyperVals &h, const StarIO &io, long long i, StarState &st,
                      HyperAcc &acc) {
  StarIn in;
  star_load_z(m, io, i, in);
  star_pre(m, h, in, st, acc);
}

// ---------------------------------------------------------------------------------------------
// BC table: multilinear interpolation with analytic Jacobian (jax.scipy RegularGridInterpolator, linear)
// ---------------------------------------------------------------------------------------------
// searchsorted(g, x, 'left') - 1, clipped to [0, n-2]: the largest i with g[i] < x.  A uniform-bin lookup table
// gives a start index that is never past the answer (one bin of slack absorbs the rounding of the bin number);
// a short linear walk finishes.  Pointers may be global or shared (the tcgen05 kernel keeps the axes in smem).
CSI_DEV int grid_locate(const float *g, const unsigned short *lut, float x0, float inv_bin, int n, float x, float &t,
                        float &inv_w, bool &oob) {
  int b = (int)((x - x0) * inv_bin) - 1;
  b = min(max(b, 0), kLutBins - 1);
  int i = (x == x) ? (int)lut[b] : 0;
  float g1 = g[i + 1];
  while (i < n - 2 && g1 < x) { ++i; g1 = g[i + 1]; }
  const float g0 = g[i];
  inv_w = rcp_(g1 - g0);
  t = (x - g0) * inv_w;
  oob = oob || (x < g[0]) || (x > g[n - 1]) || !(x == x);
  return i;
}

// value[b] and d value[b] / d x[d]; out of bounds: value NaN, Jacobian 0 (jnp.where semantics).
// The 16 corners are reduced by successive 1-D interpolations (axis 3 first, axis 0 last); the difference
// vectors of each level are kept, and the derivative along axis d is the same interpolation of the level-d
// differences over the axes 0..d-1 -- 15 + 11 float4 lerps instead of 16 x (4 x 5) weighted accumulations.
// D3 = false skips the derivative along axis 3 (Av is a constant of the model: its cotangent is never used).
struct F4 { float v[4]; };
CSI_DEV F4 f4_sub(const F4 &a, const F4 &b) { return F4{{a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2], a.v[3] - b.v[3]}}; }
CSI_DEV F4 f4_fma(float t, const F4 &d, const F4 &a) {
  return F4{{fmaf(t, d.v[0], a.v[0]), fmaf(t, d.v[1], a.v[1]), fmaf(t, d.v[2], a.v[2]), fmaf(t, d.v[3], a.v[3])}};
}
CSI_DEV F4 f4_lerp(float t, const F4 &a, const F4 &b) { return f4_fma(t, f4_sub(b, a), a); }

template <bool D3 = true>
CSI_DEV void grid_eval(const GridDev &gd, const float x[4], float val[4], float jac[4][4]) {
  int idx[4]; float t[4], iw[4]; bool oob = false;
#pragma unroll
  for (int d = 0; d < 4; ++d)
    idx[d] = grid_locate(gd.pts[d], gd.lut[d], gd.lut_x0[d], gd.lut_inv[d], gd.shape[d], x[d], t[d], iw[d], oob);
  const long long s3 = 1, s2 = gd.shape[3], s1 = s2 * gd.shape[2], s0 = s1 * gd.shape[1];
  const float4 *base = gd.values + idx[0] * s0 + idx[1] * s1 + idx[2] * s2 + idx[3] * s3;
  F4 c[16];  // corner k: bit 3 = axis 0, bit 2 = axis 1, bit 1 = axis 2, bit 0 = axis 3
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    long long off = ((k >> 3) & 1) * s0 + ((k >> 2) & 1) * s1 + ((k >> 1) & 1) * s2 + (k & 1) * s3;
    const float4 q = __ldg(base + off);
    c[k] = F4{{q.x, q.y, q.z, q.w}};
  }
  F4 e3[8], d3[8], e2[4], d2[4], e1[2], d1[2];
#pragma unroll
  for (int j = 0; j < 8; ++j) { d3[j] = f4_sub(c[2 * j + 1], c[2 * j]); e3[j] = f4_fma(t[3], d3[j], c[2 * j]); }
#pragma unroll
  for (int j = 0; j < 4; ++j) { d2[j] = f4_sub(e3[2 * j + 1], e3[2 * j]); e2[j] = f4_fma(t[2], d2[j], e3[2 * j]); }
#pragma unroll
  for (int j = 0; j < 2; ++j) { d1[j] = f4_sub(e2[2 * j + 1], e2[2 * j]); e1[j] = f4_fma(t[1], d1[j], e2[2 * j]); }
  const F4 d0 = f4_sub(e1[1], e1[0]);
  const F4 v = f4_fma(t[0], d0, e1[0]);
  const F4 g1 = f4_lerp(t[0], d1[0], d1[1]);
  const F4 g2 = f4_lerp(t[0], f4_lerp(t[1], d2[0], d2[1]), f4_lerp(t[1], d2[2], d2[3]));
  F4 g3 = F4{{0.f, 0.f, 0.f, 0.f}};
  if (D3) {
    const F4 a0 = f4_lerp(t[2], d3[0], d3[1]), a1 = f4_lerp(t[2], d3[2], d3[3]);
    const F4 a2 = f4_lerp(t[2], d3[4], d3[5]), a3 = f4_lerp(t[2], d3[6], d3[7]);
    g3 = f4_lerp(t[0], f4_lerp(t[1], a0, a1), f4_lerp(t[1], a2, a3));
  }
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    val[b] = oob ? nanf("") : v.v[b];
    jac[b][0] = oob ? 0.0f : d0.v[b] * iw[0];
    jac[b][1] = oob ? 0.0f : g1.v[b] * iw[1];
    jac[b][2] = oob ? 0.0f : g2.v[b] * iw[2];
    jac[b][3] = oob ? 0.0f : g3.v[b] * iw[3];
  }
}

// The same interpolation on the Av-collapsed table (GridDev::values3): 3 axes, 8 corners.  The three axis searches
// run interleaved (their lookup-table and axis loads are independent), one combined walk loop.
CSI_DEV void grid_eval3(const GridDev &gd, const float x[4], float val[4], float jac[4][4]) {
  int idx[3]; float g1[3], t[3], iw[3];
  bool oob = gd.av_oob != 0;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    int b = (int)((x[d] - gd.lut_x0[d]) * gd.lut_inv[d]) - 1;
    b = min(max(b, 0), kLutBins - 1);
    idx[d] = (x[d] == x[d]) ? (int)gd.lut[d][b] : 0;
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) g1[d] = gd.pts[d][idx[d] + 1];
  bool more = true;
  while (more) {
    more = false;
#pragma unroll
    for (int d = 0; d < 3; ++d)
      if (idx[d] < gd.shape[d] - 2 && g1[d] < x[d]) { ++idx[d]; g1[d] = gd.pts[d][idx[d] + 1]; more = true; }
  }
  const long long s2 = 1, s1 = gd.shape[2], s0 = s1 * gd.shape[1];
  const float4 *base = gd.values3 + idx[0] * s0 + idx[1] * s1 + idx[2] * s2;
  F4 e3[8];  // corner j: bit 2 = axis 0, bit 1 = axis 1, bit 0 = axis 2
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float4 q = __ldg(base + ((j >> 2) & 1) * s0 + ((j >> 1) & 1) * s1 + (j & 1) * s2);
    e3[j] = F4{{q.x, q.y, q.z, q.w}};
  }
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const float g0 = gd.pts[d][idx[d]];
    iw[d] = rcp_(g1[d] - g0);
    t[d] = (x[d] - g0) * iw[d];
    oob = oob || (x[d] < gd.pts[d][0]) || (x[d] > gd.pts[d][gd.shape[d] - 1]) || !(x[d] == x[d]);
  }
  F4 e2[4], d2[4], e1[2], d1[2];
#pragma unroll
  for (int j = 0; j < 4; ++j) { d2[j] = f4_sub(e3[2 * j + 1], e3[2 * j]); e2[j] = f4_fma(t[2], d2[j], e3[2 * j]); }
#pragma unroll
  for (int j = 0; j < 2; ++j) { d1[j] = f4_sub(e2[2 * j + 1], e2[2 * j]); e1[j] = f4_fma(t[1], d1[j], e2[2 * j]); }
  const F4 d0 = f4_sub(e1[1], e1[0]);
  const F4 v = f4_fma(t[0], d0, e1[0]);
  const F4 gg1 = f4_lerp(t[0], d1[0], d1[1]);
  const F4 gg2 = f4_lerp(t[0], f4_lerp(t[1], d2[0], d2[1]), f4_lerp(t[1], d2[2], d2[3]));
#pragma unroll
  for (int b = 0; b < 4; ++b) {
    val[b] = oob ? nanf("") : v.v[b];
    jac[b][0] = oob ? 0.0f : d0.v[b] * iw[0];
    jac[b][1] = oob ? 0.0f : gg1.v[b] * iw[1];
    jac[b][2] = oob ? 0.0f : gg2.v[b] * iw[2];
    jac[b][3] = 0.0f;
  }
}

// ---------------------------------------------------------------------------------------------
// predict + likelihood (numpyro_models.py:203-240), photometry (_jax/star.py:39-65)
// BcFn: object with forward(const float in[4], float val[4]) and vjp(const float gb[4], float gin[4])
// ---------------------------------------------------------------------------------------------
// observations of star i for the diagonal likelihood, indexed by the mean they observe (no column search);
// loaded ahead of star_mid so that their latency hides under the last forward MMA stage
struct StarObs { float ob[kMaxMean], dg[kMaxMean]; };
CSI_DEV void star_load_obs(const DevModel &m, const StarIO &io, long long i, StarObs &so) {
#pragma unroll
  for (int k = 0; k < kMaxMean; ++k) {
    const bool on = m.kind == 0 && io.obsm[k] != nullptr;
    so.ob[k] = on ? io.obsm[k][i] : 0.0f;
    so.dg[k] = (on && io.diagm[k]) ? io.diagm[k][i] : 0.0f;
  }
}

CSI_DEV void star_take_obs(const DevModel &m, const StarIO &io, const float *col, StarObs &so) {
  float v[kStageRows];
#pragma unroll
  for (int k = 0; k < kStageRows; ++k) v[k] = col[k * kStageStride];
#pragma unroll
  for (int k = 0; k < kMaxMean; ++k) {
    const bool on = m.kind == 0 && io.obsm[k] != nullptr;
    so.ob[k] = on ? v[k] : 0.0f;
    so.dg[k] = (on && io.diagm[k]) ? v[kMaxMean + k] : 0.0f;
  }
}

template <class BcFn>
CSI_DEV void star_mid(const DevModel &m, const StarIO &io, long long i, StarState &st, const float o[4],
                      float g_o[4], BcFn &&bc, const StarObs &so, float *mean_out = nullptr) {
  const float *ob_ = so.ob, *dg_ = so.dg;
  float mean[kMaxMean];
  // loc = A @ (o + delta_loc) + B @ params + c
  mean[0] = o[0] + m.cc[0];
  mean[1] = o[1] + m.cc[1];
  mean[2] = o[2] + m.cc[2];
  mean[3] = o[3] + m.cc[3];
  mean[4] = 4.0f * o[1] + 2.0f * o[2] + m.cc[4];
  mean[5] = -2.0f * o[2] + m.cc[5] + st.log_mass;
  mean[6] = -0.5f * o[1] - 2.0f * o[2] + m.cc[6] + st.log_mass;
#pragma unroll
  for (int k = 7; k < kMaxMean; ++k) mean[k] = 0.0f;
  float teff = 0.0f;
  if (m.bands > 0) {
    float bcin[4], bcv[4];
    teff = exp2f(mean[1] * kLog2_10);
    bcin[0] = (m.bc_mode == 1) ? teff : mean[1];
    bcin[1] = mean[5]; bcin[2] = st.x[2]; bcin[3] = m.av;
    bc.forward(bcin, bcv);
    float bol = m.bol_mag_sun - 2.5f * mean[4];
    const float inv_dist = rcp_(st.dist);
    float dm = 5.0f * (1.0f - __log2f(st.dist) * 0.30102999566398120f);
#pragma unroll
    for (int b = 0; b < kMaxBands; ++b)
      if (b < m.bands) mean[7 + b] = (bol - bcv[b]) - dm;
#pragma unroll
    for (int b = 0; b <= kMaxBands; ++b)
      if (b == m.bands) mean[7 + b] = 1e3f * inv_dist;
  }
  if (mean_out) {
#pragma unroll
    for (int k = 0; k < kMaxMean; ++k) mean_out[k] = mean[k];
  }

  float gm[kMaxMean];
#pragma unroll
  for (int k = 0; k < kMaxMean; ++k) gm[k] = 0.0f;
  float ll = 0.0f, g_tau = 0.0f;
  const float inv_tau = rcp_(st.tau);
  if (m.kind == 0) {
    // Normal(mean[idx], sqrt(var[idx]/tau + diag)).log_prob(obs)   (:231-235)
#pragma unroll
    for (int k = 0; k < kMaxMean; ++k) {
      if (io.obsm[k] != nullptr) {
        float vk = m.var_ext[k] * inv_tau;
        float s2 = vk + dg_[k];
        float inv_s2 = rcp_(s2);
        float r = ob_[k] - mean[k];
        float q = r * inv_s2;
        ll += -0.5f * r * q - 0.5f * log_(s2) - kHalfLog2Pi;
        gm[k] = q;
        float dl_ds2 = 0.5f * (q * q - inv_s2);
        g_tau += dl_ds2 * (-vk * inv_tau);
      }
    }
  } else {
    // MultivariateNormal(mean[idx], cov_sub/tau + diag(diag_i)) -- per-star k x k (the batched form at
    // numpyro_models.py:239 is broken at HEAD)
    const int k = m.n_obs;
    float S[kMaxObs * kMaxObs], r[kMaxObs], yv[kMaxObs];
    for (int a = 0; a < k; ++a) {
      int ia = m.obs_idx[a];
      const float oba = io.obs[a][i];
      const float dg = io.diag[a] ? io.diag[a][i] : 0.0f;
      r[a] = oba - mean[ia];
      for (int b = 0; b < k; ++b)
        S[a * kMaxObs + b] = m.cov_ext[ia * kMaxMean + m.obs_idx[b]] * inv_tau + (a == b ? dg : 0.0f);
    }
    // Cholesky S = L L^T (in place, lower)
    float logdet = 0.0f;
    for (int a = 0; a < k; ++a) {
      for (int b = 0; b <= a; ++b) {
        float s = S[a * kMaxObs + b];
        for (int c = 0; c < b; ++c) s -= S[a * kMaxObs + c] * S[b * kMaxObs + c];
        if (a == b) { s = sqrtf(s); logdet += 2.0f * logf(s); S[a * kMaxObs + a] = s; }
        else S[a * kMaxObs + b] = s / S[b * kMaxObs + b];
      }
    }
    // inverse of L (lower) in the strict upper part + separate diagonal array
    float Li[kMaxObs * kMaxObs];
    for (int c = 0; c < k; ++c) {
      for (int a = 0; a < k; ++a) {
        if (a < c) { Li[a * kMaxObs + c] = 0.0f; continue; }
        float s = (a == c) ? 1.0f : 0.0f;
        for (int b = c; b < a; ++b) s -= S[a * kMaxObs + b] * Li[b * kMaxObs + c];
        Li[a * kMaxObs + c] = s / S[a * kMaxObs + a];
      }
    }
    // yv = S^-1 r = Li^T (Li r);  tr(S^-1 C) and yv^T C yv with C = cov_sub
    float w[kMaxObs];
    for (int a = 0; a < k; ++a) { float s = 0.0f; for (int b = 0; b <= a; ++b) s += Li[a * kMaxObs + b] * r[b]; w[a] = s; }
    float quad_r = 0.0f;
    for (int a = 0; a < k; ++a) quad_r += w[a] * w[a];
    for (int a = 0; a < k; ++a) { float s = 0.0f; for (int b = a; b < k; ++b) s += Li[b * kMaxObs + a] * w[b]; yv[a] = s; }
    float tr = 0.0f, quad = 0.0f;
    for (int a = 0; a < k; ++a)
      for (int b = 0; b < k; ++b) {
        float sinv = 0.0f;  // (S^-1)[a][b] = sum_c Li[c][a] Li[c][b]
        for (int c = max(a, b); c < k; ++c) sinv += Li[c * kMaxObs + a] * Li[c * kMaxObs + b];
        float cab = m.cov_ext[m.obs_idx[a] * kMaxMean + m.obs_idx[b]];
        tr += sinv * cab;
        quad += yv[a] * cab * yv[b];
      }
    ll = -0.5f * quad_r - 0.5f * logdet - (float)k * kHalfLog2Pi;
    g_tau = 0.5f * inv_tau * inv_tau * (tr - quad);
    for (int a = 0; a < k; ++a) {
      int ia = m.obs_idx[a];
#pragma unroll
      for (int kk = 0; kk < kMaxMean; ++kk)
        if (kk == ia) gm[kk] += yv[a];
    }
  }
  st.lp += ll;
  st.g_tau_lik = g_tau;

  // reverse mode through predict / photometry
  float g_logmass = gm[5] + gm[6];
  float g_mh = 0.0f, g_dist = 0.0f;
  float gm1 = gm[1], gm4 = gm[4], gm5 = gm[5];
  if (m.bands > 0) {
    float gsum = 0.0f, gb[4] = {0.f, 0.f, 0.f, 0.f};  // cotangent of bcv is -g_mag
#pragma unroll
    for (int b = 0; b < kMaxBands; ++b)
      if (b < m.bands) { gsum += gm[7 + b]; gb[b] = -gm[7 + b]; }
    float g_plx = 0.0f;
#pragma unroll
    for (int b = 0; b <= kMaxBands; ++b)
      if (b == m.bands) g_plx = gm[7 + b];
    float gin[4];
    bc.vjp(gb, gin);
    gm4 += -2.5f * gsum;
    gm1 += gin[0] * ((m.bc_mode == 1) ? kLn10 * teff : 1.0f);
    gm5 += gin[1];
    g_logmass += gin[1];
    g_mh += gin[2];
    const float inv_d = rcp_(st.dist);
    g_dist = gsum * (5.0f / kLn10) * inv_d - g_plx * 1e3f * inv_d * inv_d;
  }
  st.g_logmass = g_logmass; st.g_mh_dir = g_mh; st.g_dist = g_dist;
  // g_out = g_mean[:7] @ A
  g_o[0] = gm[0];
  g_o[1] = gm1 + 4.0f * gm4 - 0.5f * gm[6];
  g_o[2] = gm[2] + 2.0f * gm4 - 2.0f * gm5 - 2.0f * gm[6];
  g_o[3] = gm[3];
}

// chain rule back to the unconstrained sites; stores dU/du = -dlogp/du
CSI_DEV void star_post(const DevModel &m, const StarIO &io, long long i, const StarState &st, const float g_x[5]) {
  io.g[S_TAU][i] = -(st.gp[S_TAU] + st.g_tau_lik * st.tau);
  io.g[S_EVOL][i] = -(st.gp[S_EVOL] + g_x[0] * st.c_evol);
  io.g[S_MASS][i] = -(st.gp[S_MASS] + g_x[1] * st.c_mass + st.g_logmass * st.dlm);
  io.g[S_MH][i] = -(st.gp[S_MH] + (g_x[2] + st.g_mh_dir) * st.dmh);
  io.g[S_Y][i] = -(st.gp[S_Y] + g_x[3] * st.dy);
  io.g[S_A][i] = -(st.gp[S_A] + g_x[4] * st.da);
  if (m.prior_mode == P_INFORMATIVE) {
    io.g[S_PY][i] = -st.gp[S_PY];
    io.g[S_PA][i] = -st.gp[S_PA];
  }
  if (m.bands > 0) io.g[S_SD][i] = -(st.gp[S_SD] + st.g_dist * st.dsd);
}

// BC sources for star_mid
struct BcNone {
  CSI_DEV void forward(const float *, float val[4]) {
#pragma unroll
    for (int b = 0; b < 4; ++b) val[b] = 0.0f;
  }
  CSI_DEV void vjp(const float *, float gin[4]) {
#pragma unroll
    for (int d = 0; d < 4; ++d) gin[d] = 0.0f;
  }
};
struct BcGrid {  // table interpolation; the Jacobian is formed with the value so corners are gathered once
  const GridDev &gd;
  float jac[4][4];
  CSI_DEV void forward(const float in[4], float val[4]) {
    if (gd.values3 != nullptr) grid_eval3(gd, in, val, jac);  // in[3] is the model constant the table was collapsed at
    else grid_eval<false>(gd, in, val, jac);
  }
  CSI_DEV void vjp(const float gb[4], float gin[4]) {
#pragma unroll
    for (int d = 0; d < 3; ++d) gin[d] = gb[0] * jac[0][d] + gb[1] * jac[1][d] + gb[2] * jac[2][d] + gb[3] * jac[3][d];
    gin[3] = 0.0f;  // Av is a model constant (numpyro_models `const`): its cotangent is never consumed
  }
};

// ---------------------------------------------------------------------------------------------
// block-level reduction of (U, hyper-gradient) partials and the single-pass grid reduce
// out6 = [U, dU/d(u_Y0, u_dYdZ, u_precY, u_mua, u_preca)] in double
// ---------------------------------------------------------------------------------------------
CSI_DEV double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace csi
