/* csi.h -- C ABI of libcsi: the B200-native NUTS potential (log-density + gradient) of celestify's
 * single-, multi- and hierarchical-star models (alexlyttle/scalable-stellar-inference).
 *
 * The reference has no native code and therefore no FFI for this path: everything below REPLACES what
 * XLA compiles from the reference's Python.  Each entry point cites the reference interface it stands in
 * for (paths relative to the reference root).  Plain pointers and sizes only -- no torch / jax types.
 *
 * Conventions
 *   - "dev" pointers are CUDA device pointers on the handle's device; "host" pointers are ordinary memory.
 *   - All functions return 0 (CSI_OK) on success, a negative CSI_E* code otherwise; csi_last_error()
 *     returns a thread-local message.  Nothing throws across the ABI.  A non-finite potential is a VALUE
 *     (NUTS treats it as a divergence), not an error.
 *   - A handle is bound to one device; calls on one handle must not race, different handles are independent.
 *     Evaluation calls are stream-ordered and never synchronise the host (except the *_host variants).
 *   - There is NO CPU fallback: creation fails with CSI_ENODEVICE when no sm_100 device is present.
 */
#ifndef CSI_H_
#define CSI_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CSI_VERSION 200

#define CSI_OK 0
#define CSI_EINVAL (-1)    /* bad argument / unsupported configuration */
#define CSI_ENODEVICE (-2) /* no CUDA device of compute capability 10.x */
#define CSI_ECUDA (-3)     /* CUDA runtime error (message has the detail) */
#define CSI_ESTATE (-4)    /* call out of order (e.g. potential before set_data) */

#define CSI_MAX_LAYERS 8
#define CSI_MAX_BANDS 4
#define CSI_MAX_MEAN 12 /* 7 emulator-derived outputs + CSI_MAX_BANDS magnitudes + parallax */
#define CSI_MAX_OBS 12
#define CSI_MAX_STAR_SITES 9
#define CSI_MAX_HYPER 5

/* model level: which numpyro model callable is reproduced (src/celestify/numpyro_models.py:54,249,263) */
#define CSI_LEVEL_SINGLE 0
#define CSI_LEVEL_MULTI 1
#define CSI_LEVEL_HIERARCHICAL 2
/* prior of the non-hierarchical models (numpyro_models.py:54 `prior=`, :163-193) */
#define CSI_PRIOR_FLAT 0
#define CSI_PRIOR_INFORMATIVE 1
/* likelihood kind (numpyro_models.py:54 `kind=`, :231-240); "none" is unusable in the reference and rejected */
#define CSI_KIND_DIAG 0
#define CSI_KIND_FULL 1
/* bolometric-correction source (src/celestify/_jax/star.py:39-52) */
#define CSI_BC_NONE 0
#define CSI_BC_GRID 1 /* 4-D table, BolometricCorrections.interpolate, input [Teff, log_g, M_H, Av] */
#define CSI_BC_NET 2  /* BC emulator, BolometricCorrections.model, input [log_Teff, log_g, M_H, Av] */
/* MLP engine */
#define CSI_ENGINE_TCGEN05 0 /* tcgen05 fp16 hi/lo split (3 passes), fp32 accumulate in TMEM -- default */
#define CSI_ENGINE_FFMA 1    /* CUDA-core fp32 FFMA engine (validation / small problems) */

/* Dense ReLU network  y = out_loc + out_scale * MLP((x - in_loc) / in_scale).
 * Replaces EmulatorBase.__init__ (src/celestify/emulator.py:8-21) and BolometricCorrectionsBase.__init__
 * (src/celestify/bolometric_corrections.py:11-26): same arrays, row-major kernel[in, out] per layer. */
typedef struct csi_net {
  int32_t n_layers;                      /* number of dense layers (7 for both shipped nets), 0 = absent */
  int32_t n_in, n_hidden, n_out;         /* 5,128,4 (stellar emulator) / 4,64,3 (BC emulator) */
  const float *weights[CSI_MAX_LAYERS];  /* host */
  const float *bias[CSI_MAX_LAYERS];     /* host */
  const double *in_loc, *in_scale;       /* host, n_in each  (in_scale = variance**0.5 or maxval-minval) */
  const double *out_loc, *out_scale;     /* host, n_out each */
} csi_net;

/* 4-D bolometric-correction table (src/celestify/bolometric_corrections.py:28-45): values[nT,ng,nF,nA,bands]. */
typedef struct csi_grid {
  int32_t shape[4];        /* all 0 = absent */
  const double *points[4]; /* host, strictly increasing, shape[d] each */
  const double *values;    /* host, row-major [shape0, shape1, shape2, shape3, bands] */
} csi_grid;

/* Truncated-normal prior parameters; low = -INFINITY / high = +INFINITY for an open side. */
typedef struct csi_tn { double loc, scale, low, high; } csi_tn;

/* Everything the reference bakes into the traced potential as constants: the `const` dict
 * (numpyro_models.py:117-134), the emulator-error model (`:107-115`), the observable selection (`:86-88`)
 * and the model class / kind / prior.  Photometric fields reconstruct the model of
 * notebook/single-star-phot.ipynb cell [7] (not in the tree at HEAD). */
typedef struct csi_config {
  int32_t level, prior, kind;
  int32_t n_obs;                  /* number of observed COLUMNS k */
  int32_t obs_index[CSI_MAX_OBS]; /* index into [log_age, log_Teff, log_R, log_Dnu, log_L, log_g, log_numax,
                                     mag_0..mag_{bands-1}, plx] */
  int32_t bands, bc_mode;
  csi_tn log_evol, log_mass, m_h; /* per-star priors (numpyro_models.py:120-123) */
  double y_low, y_high, a_low, a_high;
  double y0_loc, y0_scale;        /* Y_0 ~ Normal */
  csi_tn dydz, mu_a;              /* hyper-priors; loc/scale also feed the informative prior (:171-189) */
  double py_conc, py_rate, pa_conc, pa_rate; /* Gamma(concentration, rate) of precision_Y / precision_a */
  double delta_df, delta_loc[4], delta_cov[16]; /* emulator error: Student-t df, offset, covariance */
  double sd_c1, sd_c0, max_distance, av;        /* scaled_distance ~ Beta(c1, c0); distance = max * sd */
  csi_net emulator;
  csi_net bc_net;
  csi_grid bc_grid;
  int32_t device; /* CUDA ordinal */
  int32_t engine; /* CSI_ENGINE_* */
} csi_config;

typedef struct csi_handle csi_handle;

const char *csi_last_error(void);
int csi_version(void);

/* Replaces model construction: SingleStarModel/MultiStarModel/HierarchicalStarModel.__init__
 * (numpyro_models.py:54-105,249-253,263-264) + Emulator("jax") (emulator.py:39-51).  Copies weights,
 * tables and constants to the device; pre-splits the hidden-layer weights for the tensor-core engine. */
int csi_create(const csi_config *cfg, csi_handle **out);
void csi_destroy(csi_handle *h);

/* Shard bookkeeping for star-axis data parallelism: rank r of w owns a contiguous block of stars; the
 * (replicated) hyper-prior terms are added on rank 0 only so that a SUM-allreduce of `out` is exact. */
int csi_set_shard(csi_handle *h, int32_t rank, int32_t world);

/* Fused allreduce of `out` (no reference equivalent: the reference never shards the star axis).  peer_bufs[r] is
 * a device pointer, valid in THIS process, to rank r's exchange buffer of CSI_PEER_BUF_DOUBLES(world) doubles,
 * zero-initialised and writable by all ranks over NVLink peer memory (CUDA IPC / symmetric memory; peer_bufs[rank]
 * is the caller's own).  Afterwards the finalising block of every csi_potential stores its 6 sums into all ranks'
 * buffers and adds the world's contributions in rank order, so `out` (and the hyper slots of grad) hold the GLOBAL
 * sums and no separate collective call is needed; the device-side sampler (csi_nuts_*) sends its inner products
 * through the same buffers.  The exchange epoch lives in device memory and is advanced by the kernels themselves,
 * so launches may be captured into CUDA graphs and replayed.  All ranks must issue the same sequence of calls.
 * world <= 1 or peer_bufs == NULL switches it off. */
#define CSI_PEER_SLOT_DOUBLES 32
#define CSI_PEER_BUF_DOUBLES(world) (2 * (world) * CSI_PEER_SLOT_DOUBLES)
int csi_set_peer_exchange(csi_handle *h, int32_t world, int32_t rank, void *const *peer_bufs);

/* Observations of the local stars.  Replaces the `obs=` / `diag=` arguments of model.__call__
 * (numpyro_models.py:242,255,341) and the per-star M_H prior arrays of scripts/model_star.py:126-128.
 * obs, diag: host, row-major [n_local, n_obs] (diag may be NULL = 0).  mh_loc / mh_scale: host [n_local]
 * or NULL (= the constants of cfg.m_h). */
int csi_set_data(csi_handle *h, int64_t n_local, const double *obs, const double *diag,
                 const double *mh_loc, const double *mh_scale);

/* At most max_ctas thread blocks per csi_potential launch (0 = one per SM, the default): lets several chains, each
 * with its own handle and stream, evaluate side by side on one GPU. */
int csi_set_grid_limit(csi_handle *h, int32_t max_ctas);

/* Layout of the unconstrained vector: [hyper sites (n_hyper) | star site 0 (n_local) | star site 1 ...],
 * sites in the reference's trace order (numpyro_models.py:279-288 then :148-196 / :301-331).
 * names (optional) receives pointers to static strings. */
int csi_layout(const csi_handle *h, int32_t *n_hyper, int32_t *n_star_sites, const char **names);

/* THE hot path: U(z) = -log joint (unconstrained space) and dU/dz for the local shard.  Replaces
 * jax.value_and_grad(potential_fn) that numpyro's NUTS calls per leapfrog (scripts/model_star.py:134-152).
 *   z_dev    : float[n_hyper + n_sites*n_local]
 *   out_dev  : double[1 + CSI_MAX_HYPER] = [U_local, dU/dhyper (local partial sums)]  -- allreduce this
 *   grad_dev : float[n_hyper + n_sites*n_local]; the hyper slots receive the LOCAL partials in fp32
 * Stream-ordered, no host synchronisation, no allocation. `stream` is a cudaStream_t (NULL = legacy default).
 * The handle's device must be the caller's current device (CSI_ESTATE otherwise; the call never switches it).
 * Domain: the tcgen05 engine carries activations as fp16 pairs scaled by 2^9, so hidden activations must stay
 * below ~127 (the shipped emulator reaches ~6 inside the prior box); beyond that the result is NaN -- a value. */
int csi_potential(csi_handle *h, const float *z_dev, double *out_dev, float *grad_dev, void *stream);

/* Same through HOST buffers, synchronous.  Page-locked (cudaHostAlloc / cudaHostRegister) z_host / grad_host are
 * accessed by the kernel itself over PCIe -- no copy calls, one launch; pageable ones go through pinned staging
 * inside the handle (H2D of z, evaluate, D2H of grad).  This is the end-to-end call timed as `e2e` in bench.py. */
int csi_potential_host(csi_handle *h, const float *z_host, double *out_host, float *grad_host);

/* Emulator plug point: replaces Emulator.model (src/celestify/_jax/emulator.py:35-50) and the VJP JAX
 * derives from it.  x: [n,5] row-major, y / gy: [n,4], gx: [n,5]; all dev.  which = 0 stellar emulator,
 * 1 BC emulator (x [n,4], y [n,3]; src/celestify/_jax/bolometric_corrections.py:17-25). */
int csi_net_forward(csi_handle *h, int32_t which, const float *x_dev, float *y_dev, int64_t n, void *stream);
int csi_net_vjp(csi_handle *h, int32_t which, const float *x_dev, const float *gy_dev, float *gx_dev,
                int64_t n, void *stream);

/* BC table interpolation: replaces BolometricCorrections.interpolate
 * (src/celestify/_jax/bolometric_corrections.py:13-15; jax.scipy RegularGridInterpolator, linear, NaN
 * outside the grid).  x: [n,4] = [Teff, log_g, M_H, Av]; y / gy: [n,bands]; gx: [n,4] (0 where out of bounds). */
int csi_grid_forward(csi_handle *h, const float *x_dev, float *y_dev, int64_t n, void *stream);
int csi_grid_vjp(csi_handle *h, const float *x_dev, const float *gy_dev, float *gx_dev, int64_t n, void *stream);

/* Deterministic per-star outputs at z (numpyro.deterministic sites + predict's `loc`): replaces the
 * forward-only use in Predictive (scripts/model_star.py:157-173).  mean_dev: float[n_local, CSI_MAX_MEAN]
 * row-major (unused columns 0); params_dev: float[n_local, 8] = [evol, mass, M_H, Y, a_MLT, log_mass, tau,
 * distance]. Either may be NULL. */
int csi_predict(csi_handle *h, const float *z_dev, float *mean_dev, float *params_dev, void *stream);

/* Device memory helpers so that a host language without a CUDA binding can drive the dev-pointer calls. */
int csi_dev_alloc(csi_handle *h, int64_t bytes, void **out);
int csi_dev_free(csi_handle *h, void *p);
int csi_memcpy_h2d(csi_handle *h, void *dst_dev, const void *src_host, int64_t bytes);
int csi_memcpy_d2h(csi_handle *h, void *dst_host, const void *src_dev, int64_t bytes);
int csi_synchronize(csi_handle *h);

/* Introspection: info[0]=SM count, [1]=engine in use, [2]=kernel launches issued by the last
 * csi_potential, [3]=n_local, [4]=dynamic shared memory of the potential kernel, [5]=grid size. */
int csi_info(const csi_handle *h, int64_t info[8]);

/* ---- NUTS vector arithmetic around csi_potential (the caller of the path) ------------------------
 * One leapfrog of numpyro's velocity-Verlet (numpyro/infer/hmc_util.py velocity_verlet, driven from
 * scripts/model_star.py:134-152) is  csi_leapfrog_pre -> csi_potential(z1) [-> allreduce of out6] ->
 * csi_leapfrog_post; all vectors are float[n_hyper + n_sites*n_local] in csi_layout order, device memory,
 * stream-ordered, no synchronisation.
 *   pre : r_half = r - eps/2 g ;  z1 = z + eps inv_mass r_half ;  zeroes kin2
 *   post: g1 = grad with the hyper slots taken from out6[1..] (the REDUCED sums) ;  r1 = r_half - eps/2 g1 ;
 *         kin2[0] = sum over local stars of inv_mass r1^2 (allreduce this when sharded), kin2[1] = the hyper part;
 *         kinetic energy = (kin2[0] + kin2[1]) / 2. */
int csi_leapfrog_pre(csi_handle *h, const float *z, const float *r, const float *g, const float *inv_mass,
                     float eps, float *z1, float *r_half, double *kin2, void *stream);
int csi_leapfrog_post(csi_handle *h, const float *r_half, const float *grad, const double *out6,
                      const float *inv_mass, float eps, float *r1, float *g1, double *kin2, void *stream);

/* Generalised U-turn test (hmc_util._is_turning): rc = r_sum - (r_left + r_right)/2;
 * dots4 = [(inv_mass r_left).rc, (inv_mass r_right).rc] over the local stars (allreduce when sharded), then the
 * same two over the hyper entries; turning iff dots4[0]+dots4[2] <= 0 or dots4[1]+dots4[3] <= 0. */
int csi_uturn(csi_handle *h, const float *r_left, const float *r_right, const float *r_sum, const float *inv_mass,
              double *dots4, void *stream);

/* csi_uturn for callers that may not allocate (code under CUDA-graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments,
 * same launch and same sums as csi_uturn, but the block partials and the ticket of the grid reduction live in a
 * file-scope device array of CSI_STATIC_SCRATCH_BYTES (1024 blocks x 25 doubles + a 256-byte ticket line) that
 * every GPU loading the library holds until the process exits, so the call performs no cudaMalloc / cudaFree.
 * Vectors longer than CSI_UTURN_STATIC_MAX_LEN elements are refused with CSI_EINVAL.  All calls on a device share
 * the array: calls that overlap on the GPU (different streams, or a second handle on the same device) race --
 * serialise them, or use csi_uturn, whose scratch is per handle. */
#define CSI_STATIC_SCRATCH_BYTES 205056
#define CSI_UTURN_STATIC_MAX_LEN (1 << 24)
int csi_uturn_static(csi_handle *h, const float *r_left, const float *r_right, const float *r_sum,
                     const float *inv_mass, double *dots4, void *stream);

/* ---- the sampler itself: No-U-Turn transitions with the tree logic on the device ----------------------------------
 * Replaces numpyro.infer.NUTS / MCMC as driven by scripts/model_star.py:132-155 (multinomial NUTS, iterative
 * sub-tree building with checkpointed U-turn tests, biased progressive sampling; numpyro/infer/hmc_util.py).
 * z_state / g_state / inv_mass are device vectors float[n_hyper + n_sites*n_local] owned by the CALLER:
 * z_state / g_state hold the current position and dU/dz and are updated in place by every transition; inv_mass
 * (diagonal inverse mass matrix) may be rewritten between transitions.  Everything else lives in the object.
 * One leapfrog = 3 launches (half step, csi_potential, half step + tree bookkeeping), a doubling = one CUDA graph;
 * the host synchronises once per doubling.  Random decisions come from a counter-based hash of (seed, draw, ...):
 * runs are bit-reproducible, and for sharded runs (csi_set_shard + csi_set_peer_exchange; star_offset = global
 * index of the first local star, n_total = all stars) every rank takes the same decisions. */
typedef struct csi_nuts csi_nuts;
typedef struct csi_nuts_stats {
  double potential;    /* U at the new state */
  double accept_prob;  /* mean acceptance probability over the leaves built (dual-averaging statistic) */
  double energy0;      /* H at the start of the transition */
  int64_t num_leapfrog;
  int64_t launches;    /* kernel launches queued (graph nodes counted) */
  int32_t depth, diverging, turning, host_syncs;
} csi_nuts_stats;
int csi_nuts_create(csi_handle *h, float *z_state_dev, float *g_state_dev, const float *inv_mass_dev,
                    int32_t max_tree_depth, double max_delta_energy, uint64_t seed, int64_t star_offset,
                    int64_t n_total, csi_nuts **out);
void csi_nuts_destroy(csi_nuts *s);
/* Evaluates U and dU/dz at z_state (into g_state); call once before the first transition and whenever z_state has
 * been overwritten by the caller. */
int csi_nuts_init_state(csi_nuts *s, double *u_out);
/* One NUTS transition with step size `step_size`; `draw` numbers the transition (random stream). */
int csi_nuts_transition(csi_nuts *s, double step_size, uint64_t draw, csi_nuts_stats *stats);
/* Energy change of ONE leapfrog with a fresh momentum from the current state (left untouched): the probe of
 * find_heuristic_step_size (scripts/model_star.py:20). */
int csi_nuts_probe_step(csi_nuts *s, double step_size, uint64_t draw, double *delta_energy);
/* Test access: run only the momentum draw of a transition; copy an internal vector to the host
 * (0 z_left, 1 r_left, 2 g_left, 3 z_right, 4 r_right, 5 g_right, 6 r_sum). */
int csi_nuts_begin_only(csi_nuts *s, double step_size, uint64_t draw);
int csi_nuts_peek(csi_nuts *s, int32_t which, float *out_host);
/* The counter-based hash behind every random decision (kind 1 momentum, 2 direction, 3 leaf accept, 4 merge);
 * needs no device.  uniform = (hash >> 11) * 2^-53. */
uint64_t csi_nuts_hash(uint64_t seed, uint64_t draw, uint64_t kind, uint64_t idx);

/* Times `iters` back-to-back csi_potential calls on the handle's own stream with CUDA events recorded on
 * that stream; returns the mean milliseconds per call.  Used by bench.py for the roofline numbers. */
int csi_time_potential(csi_handle *h, const float *z_dev, double *out_dev, float *grad_dev, int32_t iters,
                       float *ms_per_call);

#ifdef __cplusplus
}
#endif
#endif /* CSI_H_ */
