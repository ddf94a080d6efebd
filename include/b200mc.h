/* b200mc -- C ABI of the B200-native HMC/NUTS sampling hot path.
 *
 * The reference (NumPyro 0.18.0) has no FFI of its own: its hot path sits behind the
 * Python MCMCKernel interface (numpyro/infer/mcmc.py:32-158) implemented by HMC/NUTS
 * (numpyro/infer/hmc.py:541-948).  Each entry point below names the reference
 * function it replaces; INTEGRATION.md shows the ctypes / jax.ffi binding.
 *
 * Conventions: plain pointers and sizes only.  Pointers named *_dev are DEVICE
 * pointers on the handle's device, owned by the caller; *_host are host pointers.
 * Work is enqueued on the caller's stream (a cudaStream_t passed as void*; NULL = default
 * stream).  Asynchronous -- return once the work is enqueued: b200mc_init / b200mc_sample /
 * b200mc_potential_and_grad on the single-launch engines (tall-data logistic with fewer
 * chains than the lock-step threshold, eight schools, stochastic volatility, horseshoe
 * below the threshold), the b200mc_random_* / b200mc_threefry2x32 hooks, b200mc_get_vector /
 * b200mc_set_vector, b200mc_memcpy_* / b200mc_memset.  SYNCHRONISING -- they wait for the
 * stream (or the device) before they return: b200mc_create (data re-tiling, operand images)
 * and b200mc_destroy (device synchronise); b200mc_init / b200mc_sample on the lock-step
 * many-chain engine (>= 64 chains of the logistic / horseshoe families, >= 16 with a dense
 * mass matrix: the host reads the count of unfinished chains every 16 global steps);
 * b200mc_get_scalars / b200mc_set_scalars / b200mc_get_profile (device synchronise + host
 * copy); b200mc_gemm3xtf32; b200mc_stream_synchronize; b200mc_event_elapsed_ms;
 * b200mc_pool_trim; b200mc_diag_* (host results).  Functions return 0 on success, non-zero on failure;
 * b200mc_last_error() describes the last failure of the calling thread.  A handle is
 * not thread-safe; distinct handles are independent.  float32 / int32 / uint32 only,
 * flat vectors in sorted-site-name order (jax ravel_pytree order).
 */
#ifndef B200MC_H_
#define B200MC_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200mc_handle b200mc_handle;

enum { B200MC_LOGISTIC = 0, B200MC_EIGHT_SCHOOLS = 1, B200MC_HORSESHOE = 2, B200MC_SV = 3,
       /* LOGISTIC with a subsampled plate, for HMCECS (hmc_gibbs.py:502-665): see subsample_size */
       B200MC_LOGISTIC_ECS = 4,
       /* EIGHT_SCHOOLS data; the centred model under LocScaleReparam(centered), reparam.py:64-145 */
       B200MC_EIGHT_SCHOOLS_LOCSCALE = 5,
       /* test/infer/test_mcmc.py:239-244: p ~ Beta(alpha, beta)[N], obs ~ Bernoulli(p);
        * data0 = column counts of ones s[N] float32, data1 = [N][4] float32: number of
        * observations, alpha, beta, lgamma(alpha) + lgamma(beta) - lgamma(alpha + beta) */
       B200MC_BETA_BERNOULLI = 6,
       /* test/infer/test_mcmc.py:283-288: p ~ Dirichlet(c[N]), obs ~ Categorical(p); flat
        * dimension N - 1 (StickBreakingTransform); data0 = class counts[N] float32, data1 =
        * c[N] followed by sum lgamma(c) - lgamma(sum c), float32 */
       B200MC_DIRICHLET_CATEGORICAL = 7,
       /* GLMs on a dense design matrix: coefs ~ Normal(0, 1)[ncols]; likelihood 0: obs ~
        * Poisson(exp(X coefs)); likelihood 1: prec_obs ~ Gamma(3, 1), obs ~ Normal(X coefs,
        * prec_obs^-1/2).  data0 = X[N][ncols] float32 row-major, data1 = y[N] float32 followed
        * by one float32: sum_i lgamma(y_i + 1) (Poisson; anything for Normal) */
       B200MC_GLM = 8 };
enum { B200MC_NUTS = 0, B200MC_HMC = 1 };

/* Arguments of NUTS.__init__/HMC.__init__ (hmc.py:625-642, 913-930) + MCMC's
 * num_warmup/num_chains (mcmc.py:323-335) + the registered model family and its data
 * (the `model_args` of MCMC.run). */
typedef struct {
  /* sizeof(b200mc_config) of the CALLER's header: b200mc_create refuses a struct of another
   * size (an older or newer binding) instead of reading fields at the wrong offsets.  Fields
   * are only ever appended; B200MC_ABI_VERSION / b200mc_version() change with the layout. */
  uint32_t struct_size;
  int family;
  int num_chains;
  /* data, device pointers:
   *   LOGISTIC      data0 = X[N][ncols] float32 row-major, data1 = labels[N] uint8
   *   EIGHT_SCHOOLS data0 = y[N] float32, data1 = sigma[N] float32
   *   HORSESHOE     data0 = X[N][ncols] float32 row-major, data1 = Y[N] float32
   *   SV            data0 = returns[N] float32, data1 = NULL                         */
  int N;
  int ncols;
  const void* data0_dev;
  const void* data1_dev;
  int algo;
  float step_size;
  int adapt_step_size;
  int adapt_mass_matrix;
  int dense_mass;               /* 1: full mass matrix; inverse_mass_matrix_dev is then [D][D] */
  int regularize_mass_matrix;
  float target_accept_prob;
  int max_tree_depth_warmup;    /* max_tree_depth=(d1, d2): hmc.py:299-303 */
  int max_tree_depth_sample;
  int num_warmup;
  int has_trajectory_length;    /* HMC */
  float trajectory_length;
  int num_steps;                /* HMC: <= 0 means None */
  int threefry_partitionable;   /* jax_threefry_partitionable: 1 = JAX >= 0.5 default */
  float init_radius;            /* init_to_uniform(radius=2) (initialization.py:95) */
  const float* inverse_mass_matrix_dev; /* [D] diagonal ([D][D] if dense_mass) or NULL (identity) */
  int find_heuristic_step_size; /* hmc_util.py:314-384 at init and at every window end */
  int likelihood;               /* HORSESHOE: 0 Bernoulli(logits), 1 Normal(mean, prec_obs^-1/2) */
  /* LOGISTIC row sharding (the reference's GSPMD row sharding, mcmc.py:239-265): this handle
   * holds rows [rank*N_total/nranks, ...) of X only (N, data0, data1 describe the LOCAL slice);
   * the sampler state is replicated on every rank and the D+1 partial sums are exchanged per
   * leapfrog over NVLink peer memory.  nranks <= 1: single GPU. */
  int nranks;
  int rank;
  /* Structured mass matrix (dense_mass=[("a","b"), ...] / dict-valued inverse_mass_matrix:
   * hmc_util.py:439-486, hmc.py:92-102, 760-769).  num_blocks <= 1: one block over the whole
   * flat vector.  Otherwise HOST tables: block_perm[D] = flat indices in concatenated block
   * order (each block in its own site order, blocks in the reference's dict order = the order
   * of the per-block momentum keys), block_sizes[num_blocks], block_dense[num_blocks] (1 =
   * full sub-matrix, 0 = diagonal).  dense_mass must be 1 iff any block is dense;
   * inverse_mass_matrix_dev is then the full [D][D] matrix in FLAT order with zeros outside
   * the blocks. */
  int num_blocks;
  const int* block_perm_host;
  const int* block_sizes_host;
  const int* block_dense_host;
  /* LOGISTIC_ECS (examples/covtype.py:66-71 with subsample_size, under HMCECS): the potential is
   * evaluated on each chain's current subsample of `subsample_size` of the N rows.
   * ecs_degree 0: no proxy, the subsample log-likelihood scaled by N/m (plate scale);
   * 1 / 2: difference estimator with a first / second order Taylor control variate around
   * ecs_ref_dev[D] (hmc_gibbs.py:584-588, contrib/ecs_proxies.py:23-52,94-278); the sums over
   * ALL rows at the reference point are computed once by b200mc_create. */
  int subsample_size;
  int ecs_degree;
  const float* ecs_ref_dev;
  /* EIGHT_SCHOOLS_LOCSCALE: 0 <= centered <= 1 (1 = the centred model, 0 = fully decentred) */
  float centered;
  /* Conditioned latent sites (HMCGibbs: the wrapped model `condition`s + `substitute`s the Gibbs
   * sites, hmc_gibbs.py:30-34,139-152): cond_count coordinates of the family's flat vector,
   * cond_index_host[cond_count] ascending, whole sites only, are taken out of the sampler's
   * vector (D shrinks by cond_count) and held at the values of b200mc_set_conditioned.  Families
   * that run through the chain-batched kernels only (not LOGISTIC / the lock-step engine). */
  int cond_count;
  const int* cond_index_host;
} b200mc_config;

/* Scalar leaves of HMCState (hmc.py:31-48) and HMCAdaptState (hmc_util.py:18-30). */
typedef struct {
  int32_t i;
  float potential_energy;
  float energy;
  int32_t num_steps;
  float accept_prob;
  float mean_accept_prob;
  int32_t diverging;
  uint32_t rng_key[2];
  float step_size;
  float ss_x_t, ss_x_avg, ss_g_avg;
  int32_t ss_t;
  float ss_prox_center;
  int32_t mm_n;
  int32_t window_idx;
  uint32_t adapt_key[2];
  int32_t status;       /* 1: "Cannot find valid initial parameters" (infer/util.py:795) */
  int32_t init_tries;
  int64_t grad_evals;   /* cumulative potential+gradient evaluations of this chain */
} b200mc_chain_scalars;

/* Vector leaves, each [num_chains][D] float32. */
enum {
  B200MC_VEC_Z = 0,
  B200MC_VEC_Z_GRAD = 1,
  B200MC_VEC_INVERSE_MASS_MATRIX = 2,
  B200MC_VEC_MASS_MATRIX_SQRT = 3,
  B200MC_VEC_WELFORD_MEAN = 4,
  B200MC_VEC_WELFORD_M2 = 5,
  /* dense_mass only, each [num_chains][D][D] */
  B200MC_MAT_INVERSE_MASS_MATRIX = 6,
  B200MC_MAT_MASS_MATRIX_SQRT = 7,
  B200MC_MAT_MASS_MATRIX_SQRT_INV = 8,
  B200MC_MAT_WELFORD_M2 = 9
};

/* Collection buffers of fori_collect (numpyro/util.py:277-407): [n_keep][num_chains][...]
 * device arrays, any of which may be NULL. */
typedef struct {
  int n_keep;
  float* z_dev;                 /* [n_keep][C][D] unconstrained draws */
  float* potential_energy_dev;  /* [n_keep][C] */
  float* energy_dev;
  int32_t* num_steps_dev;
  float* accept_prob_dev;
  float* mean_accept_prob_dev;
  int32_t* diverging_dev;
  float* step_size_dev;         /* adapt_state.step_size */
} b200mc_collect;

const char* b200mc_last_error(void);
/* ABI version of the loaded library; a binding compares it with the B200MC_ABI_VERSION it was
 * written against before calling anything else (numpyro_b200/_lib.py does). */
#define B200MC_ABI_VERSION 4
int b200mc_version(void);

/* HMC.__init__ + the data binding of MCMC.run.  Raises (returns non-zero) for an
 * unregistered family or unsupported option: there is no fallback path. */
int b200mc_create(const b200mc_config* cfg, b200mc_handle** out);
void b200mc_destroy(b200mc_handle* h);
int b200mc_dim(const b200mc_handle* h);           /* flat unconstrained dimension D */

/* HMC.init (hmc.py:740-799): per-chain keys uint32[C][2] (already split by the caller
 * as MCMC.run does, mcmc.py:661-662); init_z_dev = init_params [C][D] or NULL for
 * init_to_uniform via find_valid_initial_params (infer/util.py:335-477). */
int b200mc_init(b200mc_handle* h, const uint32_t* keys_dev, const float* init_z_dev, void* stream);

/* HMC.sample applied num_iters times inside one launch = the fori_collect loop
 * (util.py:359-369).  Local iteration k is collected into slot
 * (k - collect_start) / thinning when k >= collect_start (last write wins). */
int b200mc_sample(b200mc_handle* h, int num_iters, int collect_start, int thinning,
                  const b200mc_collect* out, void* stream);

/* mcmc.last_state / post_warmup_state round trip (mcmc.py:549-578). */
int b200mc_get_scalars(b200mc_handle* h, b200mc_chain_scalars* out_host);
int b200mc_set_scalars(b200mc_handle* h, const b200mc_chain_scalars* in_host);
int b200mc_get_vector(b200mc_handle* h, int which, float* out_dev, void* stream);
int b200mc_set_vector(b200mc_handle* h, int which, const float* in_dev, void* stream);

/* Row sharding: every rank exports the IPC handles of its exchange buffers (128 bytes), the
 * host all-gathers them (any transport), and every rank imports the table [nranks][128].
 * Must be called on all ranks between b200mc_create and b200mc_init. */
#define B200MC_COMM_HANDLE_BYTES 128
int b200mc_comm_export(b200mc_handle* h, void* out_host /*[128]*/);
int b200mc_comm_import(b200mc_handle* h, const void* all_host /*[nranks][128]*/);

/* Cycle counters of the streaming (logistic) kernel, accumulated over launches by CTA 0:
 * out[0] stream+partial, [1] grid-arrival wait, [2] grid reduction, [3] sampler step,
 * [4] passes (= gradient evaluations); SM clock cycles.  [8..15]: sampler sub-phases when
 * the library is built with -DB200MC_SAMPLER_PROF, else 0.  No reference counterpart
 * (diagnostic). */
int b200mc_get_profile(b200mc_handle* h, uint64_t* out_host /*[16]*/, int reset);

/* Launch geometry of the streaming kernel of a handle (diagnostic): out[0] CTAs, [1] TMA ring
 * stages, [2] tiles pinned in shared memory per CTA, [3] bytes per 32-row tile, [4] tiles,
 * [5] dynamic shared memory bytes, [6] chains per launch, [7] reserved. */
int b200mc_glm_layout(b200mc_handle* h, int* out_host /*[8]*/);

/* Memory-roofline microbenchmark with the streaming kernel's access pattern (diagnostic; no
 * reference counterpart): one persistent CTA per SM streams `bytes` of device memory in tiles
 * of `tile_bytes` through an `n_stages` shared-memory ring with 1-D TMA bulk copies, `passes`
 * times (odd passes backwards if reverse_odd).  Writes the achieved GB/s: the L2 roof when the
 * buffer is L2-resident, the HBM read roof when it is several times the L2 size.  Synchronises. */
int b200mc_membench(uint64_t bytes, int tile_bytes, int n_stages, int passes, int reverse_odd,
                    float* gbs_host, void* stream);

/* potential_energy + value_and_grad (infer/util.py:302-327, hmc_util.py:242-252) at
 * n points z[n][D]: tier-2 parity hook. */
int b200mc_potential_and_grad(b200mc_handle* h, const float* z_dev, int n, float* U_dev,
                              float* grad_dev, void* stream);

/* ---- HMCECS (numpyro/infer/hmc_gibbs.py:502-665) on a LOGISTIC_ECS handle; all on `stream`.
 * b200mc_ecs_init_subsample: every chain's initial subsample = the plate's value in the
 *   prototype trace `trace(substitute(seed(model, key_u), init_to_sample))` (hmc_gibbs.py:547-552):
 *   key_u_dev[C][2]; the seed handler's splits for the sites before the plate, then the
 *   Fisher-Yates of primitives.py:443-462 (one jax.random.randint per index).
 * b200mc_ecs_get_subsample / _set_subsample: idx[C][subsample_size] int32, device pointers.
 * b200mc_ecs_update: one Gibbs step per chain (hmc_gibbs.py:619-652): rng_key = split(key)[0];
 *   _update_block(rng_key, num_blocks, ...) (contrib/ecs_proxies.py:58-73) proposes one block of
 *   new indices; pe_new at the chain's current z; accept_prob = min(exp(pe - pe_new), 1);
 *   bernoulli(rng_key, accept_prob); on acceptance the chain's subsample, potential energy and
 *   z_grad are replaced.  keys_dev[C][2] is overwritten with rng_key; accept_prob_dev[C]. */
int b200mc_ecs_init_subsample(b200mc_handle* h, const uint32_t* key_u_dev, void* stream);
int b200mc_ecs_get_subsample(b200mc_handle* h, int32_t* idx_dev, void* stream);
int b200mc_ecs_set_subsample(b200mc_handle* h, const int32_t* idx_dev, void* stream);
int b200mc_ecs_update(b200mc_handle* h, uint32_t* keys_dev, int num_blocks, float* accept_prob_dev,
                      void* stream);
/* Values of the conditioned coordinates of every chain (b200mc_config.cond_count):
 * values_dev[num_chains][cond_count] in UNCONSTRAINED space, in cond_index order, and
 * offset_dev[num_chains] (or null: zeros) added to the potential -- the Jacobian terms of the
 * conditioned sites, which the reference's conditioned density does not carry.  The chains'
 * cached potential_energy / z_grad are NOT refreshed: the caller re-evaluates them
 * (b200mc_potential_and_grad, point i uses chain i's values) as hmc_gibbs.py:170-185 does. */
int b200mc_set_conditioned(b200mc_handle* h, const float* values_dev, const float* offset_dev,
                           void* stream);

/* The whole HMCECS loop on the device: num_iters times { the Gibbs step of b200mc_ecs_update; one
 * transition of the inner kernel }, in ONE launch, collecting like b200mc_sample.
 * accept_prob_dev[C]: the last step's accept_prob; accept_all_dev[num_iters][C] (or NULL): every
 * step's (HMCECSState.accept_prob of each iteration, hmc_gibbs.py:665). */
int b200mc_ecs_sample(b200mc_handle* h, uint32_t* keys_dev, int num_blocks, int num_iters,
                      int collect_start, int thinning, const b200mc_collect* out,
                      float* accept_prob_dev, float* accept_all_dev, void* stream);
/* tier-1 hook: jax.random.randint(key, (n,), minval, maxval) int32 (n = 0: shape ()) */
int b200mc_random_randint(const uint32_t* key_dev, int n, int minval, int maxval,
                            int partitionable, int32_t* out_dev, void* stream);

/* log_likelihood / Predictive (numpyro/infer/util.py:803-1148) for the registered families'
 * observation site over S posterior draws: m[s][i] = X[i] . B[s] (X_dev [N][K] row-major and
 * B [S][K], the GLM families) or m = loc [S][N] when X_dev is NULL.  likelihood 0
 * Bernoulli(logits = m), 1 Normal(m, scale), 2 StudentT(df[s], m, scale); scale_kind 0 none,
 * 1 per draw [S], 2 per observation [N], 3 full [S][N].  With y_dev (and keys_dev NULL):
 * out[s][i] = log_prob(y_i).  With keys_dev [S][2] (the observation site's key of every draw, as
 * the seed handler hands it out): out[s][i] = one draw of the site -- Bernoulli as int32
 * (uniform(key_s, (N,))[i] < expit(m)), Normal as float32 (m + scale * normal(key_s, (N,))[i]). */
int b200mc_obs_forward(const float* X_dev, int N, int K, const float* B_or_loc_dev, int S,
                       const float* scale_dev, int scale_kind, const float* df_dev,
                       const float* y_dev, const uint32_t* keys_dev, int likelihood,
                       int partitionable, float* out_dev, void* stream);

/* On-device diagnostics over collected draws x[n_total][C][D] (the collection layout of
 * b200mc_collect.z_dev), draws first .. first+n-1 of every chain: effective_sample_size (Geyer
 * initial monotone sequence on biased autocovariances, numpyro/diagnostics.py:158-203),
 * gelman_rubin (:45-61; written only when C > 1) and split_gelman_rubin (:64-80).  Outputs are
 * device arrays of D float64, any may be NULL.  Synchronises the stream. */
int b200mc_diag_ess_rhat(const float* x_dev, int n_total, int first, int n, int chains, int D,
                         double* n_eff_dev, double* r_hat_dev, double* split_r_hat_dev, void* stream);

/* Tier-2 hook of the streaming (tall-data logistic) kernel in the regime it is benchmarked in:
 * the SAME point z[D] is evaluated on `passes` consecutive passes of ONE launch -- forward
 * sweep, reversed sweep, forward, ... with the TMA ring wrapping and the L2-resident ping-pong
 * in between -- and every pass's result is written: U_dev[passes], grad_dev[passes][D].  All
 * passes must agree with the oracle (and with each other up to summation order). */
int b200mc_potential_and_grad_passes(b200mc_handle* h, const float* z_dev, int passes,
                                     float* U_dev, float* grad_dev, void* stream);

/* Allocation-free variant of the horseshoe potential + gradient (the kernel behind
 * b200mc_potential_and_grad for the HORSESHOE family below the lock-step threshold).  NO
 * COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  It needs no handle and allocates nothing (no cudaMalloc / cudaFree / allocating
 * library call): the kernel's scratch -- one residual per (point, data row), n * N floats --
 * lives in a file-scope device array of B200MC_NOALLOC_MAX_FLOATS floats (4 MiB) that every GPU
 * that loads the library holds until exit; n * N beyond that is refused with an error.
 * Inputs and outputs are the caller's device buffers: X[N][ncols] f32 row-major, Y[N] f32,
 * likelihood 0 Bernoulli(logits) / 1 Normal, z[n][D] with D = 2 * ncols + 1 (+ 1 if Normal) in
 * flat sorted-site order, U[n], grad[n][D].  Asynchronous on `stream`.  All calls share the one
 * scratch array (it keeps the last call's contents; nothing in it has to start at zero): calls
 * that overlap on the GPU -- different streams, different host threads -- RACE; serialise them. */
#define B200MC_NOALLOC_MAX_FLOATS (1u << 20)
int b200mc_horseshoe_potential_noalloc(const float* X_dev, const float* Y_dev, int N, int ncols,
                                       int likelihood, const float* z_dev, int n, float* U_dev,
                                       float* grad_dev, void* stream);

/* jax.random on Threefry-2x32 (tier-1 parity hooks). */
int b200mc_threefry2x32(const uint32_t* key_dev /*[n][2]*/, const uint32_t* ctr_dev /*[n][2]*/,
                        uint32_t* out_dev /*[n][2]*/, int n, void* stream);
int b200mc_random_split(const uint32_t* key_dev /*[n][2]*/, int n, int num,
                        uint32_t* out_dev /*[n][num][2]*/, int partitionable, void* stream);
int b200mc_random_bits(const uint32_t* key_dev /*[2]*/, int n, uint32_t* out_dev,
                       int partitionable, void* stream);
int b200mc_random_uniform(const uint32_t* key_dev, int n, float minval, float maxval,
                          float* out_dev, int partitionable, void* stream);
int b200mc_random_normal(const uint32_t* key_dev, int n, float* out_dev, int partitionable,
                         void* stream);

/* Split-precision tensor-core contraction of the many-chain potentials (tier-2 hook):
 * out[c][m] = sum_k A[m][k] * B[c][k] for row-major f32 A[M][K], B[Nc][K], Nc <= 256, computed
 * as hi*hi + hi*lo + lo*hi in TF32 on tcgen05 with f32 accumulation over `splits` k-ranges.
 * Stands where the reference has jnp.dot / matmul inside the traced potential
 * (examples/horseshoe_regression.py:61,78; examples/covtype.py:68). */
int b200mc_gemm3xtf32(const float* A_dev, const float* B_dev, int M, int Nc, int K, int splits,
                      float* out_dev /*[Nc][M]*/, void* stream);

/* build_adaptation_schedule (hmc_util.py:387-436), host only: writes up to cap windows,
 * returns the number of windows. */
int b200mc_adaptation_schedule(int num_steps, int* starts_host, int* ends_host, int cap);

/* Device-memory plumbing so a host binding needs nothing but this library. */
int b200mc_device_count(void);
int b200mc_set_device(int device);
/* b200mc_malloc / b200mc_free (and the handle's own buffers) go through a caching allocator:
 * freed blocks are kept and reused on an exact-size match (B200MC_POOL_BYTES caps the cache,
 * 0 disables it); free a buffer only after the work that uses it has been waited for.
 * b200mc_pool_trim returns every cached block to the driver. */
void* b200mc_malloc(uint64_t bytes);
int b200mc_free(void* p_dev);
int b200mc_pool_trim(void);
int b200mc_memcpy_h2d(void* dst_dev, const void* src_host, uint64_t bytes, void* stream);
int b200mc_memcpy_d2h(void* dst_host, const void* src_dev, uint64_t bytes, void* stream);
int b200mc_memset(void* dst_dev, int value, uint64_t bytes, void* stream);
int b200mc_stream_synchronize(void* stream);
void* b200mc_host_alloc(uint64_t bytes);          /* pinned host memory */
int b200mc_host_free(void* p_host);
/* CUDA-event timing on the launching stream. */
void* b200mc_event_create(void);
int b200mc_event_record(void* event, void* stream);
int b200mc_event_elapsed_ms(void* start, void* stop, float* ms_host);
int b200mc_event_destroy(void* event);
/* number of this library's kernels launched so far by the calling process */
int64_t b200mc_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* B200MC_H_ */
