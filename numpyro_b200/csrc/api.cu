// C ABI of libb200mc (declared in include/b200mc.h): handle management, launch logic.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <unordered_map>
#include <string>
#include <vector>

#include <cusolverDn.h>  // types only: the library is loaded on demand (dlopen)
#include <dlfcn.h>

#include "../../include/b200mc.h"
#include "launchers.h"  // launch structs + the per-family translation units' launch wrappers

using namespace b200mc;

static_assert(sizeof(b200mc_chain_scalars) == sizeof(ChainScalars), "ABI mirror out of sync");

static thread_local std::string g_err;
static std::atomic<int64_t> g_launches{0};

static int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}
#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return fail("%s failed: %s", #call, cudaGetErrorString(e_)); \
  } while (0)

struct b200mc_handle {
  b200mc_config cfg;
  int device = 0, num_sms = 0, max_smem = 0;
  int C = 0, D = 0;
  SamplerParams P;
  ModelData M;
  ChainScalars* cs = nullptr;
  ChainVectors cv{};
  float* work = nullptr;
  int mode = 0, c_pad = 0;
  // conditioned sites (b200mc_config.cond_count)
  int cond_count = 0, cond_dfull = 0;
  int* cond_map = nullptr;     // [D] free -> full
  int* cond_index = nullptr;   // [cond_count] conditioned -> full
  float* cond_full = nullptr;  // [C][2 * cond_dfull]
  float* cond_offset = nullptr;
  // glm
  int glm_ctas = 0, glm_nt_total = 0, glm_stages = 0, glm_pinned = 0, glm_smem = 0;
  unsigned char* glm_xt = nullptr;  // re-tiled X + labels
  float* partials = nullptr;     // [2][ctas][D+1]
  unsigned* counters = nullptr;  // [0] arrive, [3] error
  unsigned long long* prof = nullptr;  // [kGlmProfSlots]
  int glm_reverse = 1;
  // several chains on one stream (glm_mc_kernel.cuh): own ring / pinned split and partial buffer
  bool glm_mc = false;
  int mc_stages = 0, mc_pinned = 0, mc_smem = 0, mc_nt_total = 0, mc_ctas = 0;
  unsigned char* mc_xt = nullptr;  // X + labels re-tiled in 16-row tiles (glm_mc_kernel.cuh)
  float* mc_partials = nullptr;    // [2][ctas][kMcChains][D+1]
  float* dense_replicas = nullptr;
  int* blk_tables = nullptr;  // structured mass matrix: blk_of, blk_pos, blk_perm [D] + size, dense [nb]
  // row sharding
  int nranks = 1, rank = 0;
  float* xchg_local = nullptr;
  unsigned* xarrive_local = nullptr;
  float* xchg[kGlmMaxRanks] = {};
  unsigned* xarrive[kGlmMaxRanks] = {};
  bool comm_ready = false;
  bool initialized = false;
  // HMCECS (LOGISTIC_ECS): subsamples, proposals, proxy sums, scratch
  int *ecs_idx = nullptr, *ecs_prop = nullptr, *ecs_hash = nullptr;
  float *ecs_sums = nullptr, *ecs_ref = nullptr, *ecs_gtmp = nullptr;
  int ecs_hcap = 0;
  // many-chain lock-step engine (mode 3): operand images and GEMM plumbing
  struct Batch {
    unsigned char *x_hi = nullptr, *x_lo = nullptr;    // X   [RA1][K1]
    unsigned char *xt_hi = nullptr, *xt_lo = nullptr;  // X^T [RA2][RA1]
    unsigned char *b_hi = nullptr, *b_lo = nullptr;    // B   [256][K1]
    unsigned char *r_hi = nullptr, *r_lo = nullptr;    // R   [256][RA1]
    float *ll_part = nullptr, *gpart = nullptr, *col_scale = nullptr;
    unsigned char* snap = nullptr;
    int* status = nullptr;
    unsigned* active = nullptr;
    MatvecRequest* reqs = nullptr;
    float* y_f32 = nullptr;  // logistic: labels as floats
    int RA1 = 0, K1 = 0, RA2 = 0, splits = 1, snap_stride = 0;
  } bt;
};
constexpr int kBatchStepsPerSync = 16;

extern "C" const char* b200mc_last_error(void) { return g_err.c_str(); }
extern "C" int b200mc_version(void) { return B200MC_ABI_VERSION; }
extern "C" int64_t b200mc_kernel_launches(void) { return g_launches.load(); }

// build_adaptation_schedule: hmc_util.py:387-436
static std::vector<std::pair<int, int>> adaptation_schedule(int num_steps) {
  std::vector<std::pair<int, int>> sch;
  if (num_steps < 20) {
    sch.push_back({0, num_steps - 1});
    return sch;
  }
  int start_buffer = 75, end_buffer = 50, init_window = 25;
  if (start_buffer + end_buffer + init_window > num_steps) {
    start_buffer = (int)(0.15 * num_steps);
    end_buffer = (int)(0.1 * num_steps);
    init_window = num_steps - start_buffer - end_buffer;
  }
  sch.push_back({0, start_buffer - 1});
  const int end_window_start = num_steps - end_buffer;
  int next_size = init_window, next_start = start_buffer;
  while (next_start < end_window_start) {
    const int cur_start = next_start;
    int cur_size = next_size;
    if (3 * cur_size <= end_window_start - cur_start) next_size = 2 * cur_size;
    else cur_size = end_window_start - cur_start;
    next_start = cur_start + cur_size;
    sch.push_back({cur_start, next_start - 1});
  }
  sch.push_back({end_window_start, num_steps - 1});
  return sch;
}

extern "C" int b200mc_adaptation_schedule(int num_steps, int* starts, int* ends, int cap) {
  auto sch = adaptation_schedule(num_steps);
  for (size_t i = 0; i < sch.size() && (int)i < cap; ++i) {
    starts[i] = sch[i].first;
    ends[i] = sch[i].second;
  }
  return (int)sch.size();
}

static int family_sites(const b200mc_config& c, SamplerParams& P, int& D) {
  // sites in TRACE order with their offsets in the flat (sorted-name) vector
  switch (c.family) {
    case B200MC_LOGISTIC:  // coefs
    case B200MC_LOGISTIC_ECS:
      D = c.ncols;
      P.nsites = 1; P.site_off[0] = 0; P.site_size[0] = D;
      return 0;
    case B200MC_EIGHT_SCHOOLS_LOCSCALE:  // trace: mu, tau, theta(_decentered) == sorted
      if (!(c.centered >= 0.f && c.centered <= 1.f))
        return fail("`centered` argument does not satisfy `0 <= centered <= 1`.");
      [[fallthrough]];
    case B200MC_EIGHT_SCHOOLS:  // trace: mu, tau, theta_base == sorted
      D = 2 + c.N;
      P.nsites = 3;
      P.site_off[0] = 0; P.site_size[0] = 1;
      P.site_off[1] = 1; P.site_size[1] = 1;
      P.site_off[2] = 2; P.site_size[2] = c.N;
      return 0;
    case B200MC_BETA_BERNOULLI:  // p_latent
      D = c.N;
      P.nsites = 1; P.site_off[0] = 0; P.site_size[0] = D;
      return 0;
    case B200MC_DIRICHLET_CATEGORICAL:  // p_latent on the simplex: N - 1 unconstrained coordinates
      if (c.N < 2) return fail("dirichlet_categorical: need at least two classes");
      D = c.N - 1;
      P.nsites = 1; P.site_off[0] = 0; P.site_size[0] = D;
      return 0;
    case B200MC_SV:  // trace: sigma, s, nu; sorted: nu, s, sigma
      D = 2 + c.N;
      P.nsites = 3;
      P.site_off[0] = 1 + c.N; P.site_size[0] = 1;
      P.site_off[1] = 1;       P.site_size[1] = c.N;
      P.site_off[2] = 0;       P.site_size[2] = 1;
      return 0;
    case B200MC_HORSESHOE: {
      // trace: lambdas, tau, unscaled_betas(, prec_obs); sorted: lambdas(, prec_obs), tau,
      // unscaled_betas
      const int Dh = c.ncols, nrm = c.likelihood == 1 ? 1 : 0;
      if (Dh < 1) return fail("horseshoe: ncols must be >= 1");
      D = 2 * Dh + 1 + nrm;
      P.nsites = 3 + nrm;
      P.site_off[0] = 0;            P.site_size[0] = Dh;
      P.site_off[1] = Dh + nrm;     P.site_size[1] = 1;
      P.site_off[2] = Dh + nrm + 1; P.site_size[2] = Dh;
      if (nrm) { P.site_off[3] = Dh; P.site_size[3] = 1; }
      return 0;
    }
    case B200MC_GLM: {  // trace: coefs(, prec_obs) == sorted
      const int nrm = c.likelihood == 1 ? 1 : 0;
      if (c.ncols < 1) return fail("glm: ncols must be >= 1");
      D = c.ncols + nrm;
      P.nsites = 1 + nrm;
      P.site_off[0] = 0; P.site_size[0] = c.ncols;
      if (nrm) { P.site_off[1] = c.ncols; P.site_size[1] = 1; }
      return 0;
    }
    default:
      return fail("model family %d is not a registered potential family (logistic=0, "
                  "eight_schools=1, horseshoe=2, stochastic_volatility=3, logistic_ecs=4, eight_schools_locscale=5, beta_bernoulli=6, dirichlet_categorical=7, glm=8); no fallback", c.family);
  }
}

// ---------------------------------------------------------------- 3xTF32 GEMM plumbing
__global__ void labels_to_float_kernel(const uint8_t* y, int n, float* out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = (float)y[i];
}
// Caching device allocator.  cudaMalloc / cudaFree of a 128 MB buffer cost ~5-10 ms each (the
// free also synchronises the device) -- more than the H2D copy of the design matrix itself --
// and the end-to-end path creates a handle per MCMC.run.  Freed blocks are therefore kept in
// per-device free lists keyed by their (512-byte rounded) size and handed out again on an
// exact-size match; the same shapes come back run after run.  Blocks go back to the lists only
// from b200mc_destroy (after a device synchronise) and b200mc_free (the Python mirror frees a
// buffer only after the copy or kernel that used it has been waited for), so a cached block
// has no work in flight.  B200MC_POOL_BYTES caps the cached bytes (default 16 GiB, 0 = off).
namespace {
struct DevicePool {
  std::mutex mu;
  std::unordered_map<void*, std::pair<int, size_t>> live;   // ptr -> (device, rounded bytes)
  std::multimap<std::pair<int, size_t>, void*> cached;
  size_t cached_bytes = 0;
  size_t cap = (size_t)16 << 30;
  DevicePool() {
    if (const char* e = getenv("B200MC_POOL_BYTES")) cap = (size_t)strtoull(e, nullptr, 10);
  }
};
DevicePool& pool() {
  static DevicePool* p = new DevicePool();  // never destroyed: no frees after the driver is gone
  return *p;
}
}  // namespace
static cudaError_t dmalloc(void** p, size_t bytes) {
  bytes = (bytes + 511) & ~(size_t)511;
  if (bytes == 0) bytes = 512;
  int dev = 0;
  cudaGetDevice(&dev);
  DevicePool& P = pool();
  {
    std::lock_guard<std::mutex> lk(P.mu);
    auto it = P.cached.find({dev, bytes});
    if (it != P.cached.end()) {
      *p = it->second;
      P.cached.erase(it);
      P.cached_bytes -= bytes;
      P.live[*p] = {dev, bytes};
      return cudaSuccess;
    }
  }
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) {  // out of memory: give the cached blocks back and retry once
    cudaGetLastError();
    std::vector<void*> drop;
    {
      std::lock_guard<std::mutex> lk(P.mu);
      for (auto& kv : P.cached) drop.push_back(kv.second);
      P.cached.clear();
      P.cached_bytes = 0;
    }
    for (void* q : drop) cudaFree(q);
    e = cudaMalloc(p, bytes);
  }
  if (e == cudaSuccess) {
    std::lock_guard<std::mutex> lk(P.mu);
    P.live[*p] = {dev, bytes};
  }
  return e;
}
static void dfree(void* p) {
  if (!p) return;
  DevicePool& P = pool();
  {
    std::lock_guard<std::mutex> lk(P.mu);
    auto it = P.live.find(p);
    if (it != P.live.end()) {
      const auto key = it->second;
      P.live.erase(it);
      if (P.cached_bytes + key.second <= P.cap) {
        P.cached.insert({key, p});
        P.cached_bytes += key.second;
        return;
      }
    }
  }
  cudaFree(p);
}
extern "C" int b200mc_pool_trim(void) {
  DevicePool& P = pool();
  std::vector<void*> drop;
  {
    std::lock_guard<std::mutex> lk(P.mu);
    for (auto& kv : P.cached) drop.push_back(kv.second);
    P.cached.clear();
    P.cached_bytes = 0;
  }
  cudaDeviceSynchronize();
  for (void* q : drop) cudaFree(q);
  return 0;
}

namespace {
__global__ void sum_splits_kernel(const float* part, int splits, int Nc, int RA, int M, float* out) {
  const long total = (long)Nc * M;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i / M), m = (int)(i - (long)c * M);
    float a = 0.f;
    for (int s = 0; s < splits; ++s) a += part[((long)s * kGemmN + c) * RA + m];
    out[i] = a;
  }
}
int build_image(const float* src, int R_src, int K_src, int transpose, int R, int K,
                unsigned char** hi, unsigned char** lo, cudaStream_t st) {
  const size_t bytes = (size_t)R * K * 4;
  CK(dmalloc((void**)hi, bytes));
  CK(dmalloc((void**)lo, bytes));
  CK(kl::image_build_launch(src, R_src, K_src, transpose, R, K, *hi, *lo, st));
  g_launches++;
  return 0;
}
int launch_gemm3x(const Gemm3xLaunch& L, int epi, cudaStream_t st) {
  static bool attr_set = false;
  static int persist = -1, sms = 0;
  if (!attr_set) {
    CK(kl::gemm_prepare());
    const char* e = getenv("B200MC_GEMM_PERSIST");
    persist = e ? atoi(e) : 1;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    attr_set = true;
  }
  const int grid = (L.RA / kGemmM) * L.splits;
  if (epi == EPI_RESID && persist && L.KB <= kPersistMaxKB && L.splits == 1) {
    // short inner dimension (logistic: K = 64): persistent CTAs, B resident, TMEM double-buffered
    CK(kl::gemm3x_persistent_launch(L, grid < sms ? grid : sms, st));
    g_launches++;
    return 0;
  }
  CK(kl::gemm3x_launch(L, epi, grid, st));
  g_launches++;
  return 0;
}
}  // namespace

extern "C" void b200mc_destroy(b200mc_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();  // nothing of this handle is in flight when its blocks are recycled
  dfree(h->cs);
  dfree(h->cv.z); dfree(h->cv.z_grad); dfree(h->cv.inverse_mass_matrix);
  dfree(h->cv.mass_matrix_sqrt); dfree(h->cv.welford_mean); dfree(h->cv.welford_m2);
  dfree(h->work); dfree(h->partials); dfree(h->mc_partials); dfree(h->counters); dfree(h->prof); dfree(h->glm_xt); dfree(h->mc_xt);
  dfree(h->cv.dense); dfree(h->dense_replicas); dfree(h->M.scratch);
  dfree(h->blk_tables);
  dfree(h->ecs_idx); dfree(h->ecs_prop); dfree(h->ecs_hash); dfree(h->ecs_sums); dfree(h->ecs_ref);
  dfree(h->ecs_gtmp);
  for (int r = 0; r < h->nranks; ++r) {
    if (r == h->rank) continue;
    if (h->xchg[r]) cudaIpcCloseMemHandle(h->xchg[r]);
    if (h->xarrive[r]) cudaIpcCloseMemHandle(h->xarrive[r]);
  }
  cudaFree(h->xchg_local); cudaFree(h->xarrive_local);
  dfree(h->cond_map); dfree(h->cond_index); dfree(h->cond_full); dfree(h->cond_offset);
  dfree(h->bt.x_hi); dfree(h->bt.x_lo); dfree(h->bt.xt_hi); dfree(h->bt.xt_lo);
  dfree(h->bt.b_hi); dfree(h->bt.b_lo); dfree(h->bt.r_hi); dfree(h->bt.r_lo);
  dfree(h->bt.ll_part); dfree(h->bt.gpart); dfree(h->bt.col_scale);
  dfree(h->bt.snap); dfree(h->bt.status); dfree(h->bt.active); dfree(h->bt.reqs); dfree(h->bt.y_f32);
  delete h;
}

extern "C" int b200mc_dim(const b200mc_handle* h) { return h ? h->D : -1; }

extern "C" int b200mc_create(const b200mc_config* cfg, b200mc_handle** out) {
  if (!cfg || !out) return fail("null argument");
  *out = nullptr;
  if (cfg->struct_size != (uint32_t)sizeof(b200mc_config))
    return fail("b200mc_config.struct_size = %u but this library's b200mc_config has %u bytes "
                "(ABI version %d): the binding and the library are out of sync",
                cfg->struct_size, (unsigned)sizeof(b200mc_config), B200MC_ABI_VERSION);
  const b200mc_config& c = *cfg;
  if (c.num_chains < 1) return fail("num_chains must be >= 1");
  if (c.algo != B200MC_NUTS && c.algo != B200MC_HMC)
    return fail("`algo` must be one of `HMC` or `NUTS`.");
  if (c.algo == B200MC_HMC && c.num_steps <= 0 && !c.has_trajectory_length)
    return fail("At least one of `num_steps` or `trajectory_length` must be specified.");
  if (c.max_tree_depth_warmup < 1 || c.max_tree_depth_sample < 1 ||
      c.max_tree_depth_warmup > kMaxTreeDepth || c.max_tree_depth_sample > kMaxTreeDepth)
    return fail("max_tree_depth must be in [1, %d]", kMaxTreeDepth);
  if (c.N < 1 || !c.data0_dev) return fail("model data missing");

  auto* h = new b200mc_handle();
  h->cfg = c;
  SamplerParams& P = h->P;
  memset(&P, 0, sizeof(P));
  if (family_sites(c, P, h->D)) { delete h; return 1; }
  std::vector<int> cond_map_h;
  if (c.cond_count > 0) {
    // Conditioned sites: whole sites leave the sampler's vector; the site table (trace order,
    // offsets in the flat vector) and D describe the free coordinates from here on.
    if (c.family == B200MC_LOGISTIC || c.family == B200MC_LOGISTIC_ECS || c.num_blocks > 1 || c.nranks > 1) {
      delete h;
      return fail("conditioned sites: not available for this family / configuration");
    }
    if (!c.cond_index_host || c.cond_count >= h->D) { delete h; return fail("conditioned sites: bad index list"); }
    const int Dfull = h->D;
    std::vector<int> is_cond(Dfull, 0), to_free(Dfull, -1);
    for (int k = 0; k < c.cond_count; ++k) {
      const int i = c.cond_index_host[k];
      if (i < 0 || i >= Dfull || (k > 0 && i <= c.cond_index_host[k - 1])) {
        delete h;
        return fail("conditioned sites: indices must be ascending and inside the vector");
      }
      is_cond[i] = 1;
    }
    for (int i = 0; i < Dfull; ++i)
      if (!is_cond[i]) { to_free[i] = (int)cond_map_h.size(); cond_map_h.push_back(i); }
    int ns = 0;
    for (int k = 0; k < P.nsites; ++k) {
      int n_in = 0;
      for (int j = 0; j < P.site_size[k]; ++j) n_in += is_cond[P.site_off[k] + j];
      if (n_in == P.site_size[k]) continue;  // conditioned: observed for the init search's key tree
      if (n_in != 0) { delete h; return fail("conditioned sites: a site is conditioned as a whole or not at all"); }
      P.site_off[ns] = to_free[P.site_off[k]];
      P.site_size[ns] = P.site_size[k];
      ++ns;
    }
    P.nsites = ns;
    h->cond_count = c.cond_count;
    h->cond_dfull = Dfull;
    h->D = (int)cond_map_h.size();
  }
  cudaGetDevice(&h->device);
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, h->device);
  cudaDeviceGetAttribute(&h->max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device);
  {
    // the step machine's call chains (transition end -> adaptation -> next transition) are
    // deeper than ptxas' static estimate accounts for: give the device stack head-room
    size_t stack = 2048;
    if (const char* e = getenv("B200MC_STACK_BYTES")) stack = (size_t)atoi(e);
    size_t cur = 0;
    cudaDeviceGetLimit(&cur, cudaLimitStackSize);
    if (cur < stack) cudaDeviceSetLimit(cudaLimitStackSize, stack);
  }
  const int C = h->C = c.num_chains, D = h->D;
  P.D = D;
  P.algo = c.algo;
  P.max_depth_warmup = c.max_tree_depth_warmup;
  P.max_depth_sample = c.max_tree_depth_sample;
  P.max_depth = c.max_tree_depth_warmup > c.max_tree_depth_sample ? c.max_tree_depth_warmup
                                                                  : c.max_tree_depth_sample;
  P.num_warmup = c.num_warmup;
  P.adapt_step_size = c.adapt_step_size;
  P.adapt_mass_matrix = c.adapt_mass_matrix;
  P.regularize_mass_matrix = c.regularize_mass_matrix;
  P.partitionable = c.threefry_partitionable;
  P.target_accept_prob = c.target_accept_prob;
  P.init_step_size = c.step_size;
  P.has_trajectory_length = c.has_trajectory_length;
  P.trajectory_length = c.trajectory_length;
  P.fixed_num_steps = c.num_steps;
  P.max_delta_energy = 1000.0f;  // hmc.py:188
  P.dense_mass = c.dense_mass ? 1 : 0;
  P.find_heuristic = c.find_heuristic_step_size ? 1 : 0;
  P.factor_at_init = (c.dense_mass && c.inverse_mass_matrix_dev) ? 1 : 0;
  P.init_radius = c.init_radius > 0.f ? c.init_radius : 2.0f;
  auto sch = adaptation_schedule(c.num_warmup);
  if ((int)sch.size() > kMaxWindows) { delete h; return fail("too many adaptation windows"); }
  P.num_windows = (int)sch.size();
  for (size_t i = 0; i < sch.size(); ++i) P.window_end[i] = sch[i].second;

  h->M.family = c.family;
  h->M.N = c.N;
  h->M.D = D;
  h->M.a = (const float*)c.data0_dev;
  h->M.b = (const float*)c.data1_dev;
  h->M.ncols = c.ncols;
  h->M.likelihood = c.likelihood;
  h->M.scratch = nullptr;
  h->M.y8 = nullptr; h->M.ecs_idx = nullptr; h->M.ecs_ref = nullptr; h->M.ecs_sums = nullptr;
  h->M.ecs_m = 0; h->M.ecs_degree = 0; h->M.scratch_stride = 0;
  h->M.centered = c.centered;
  h->M.cond_map = nullptr; h->M.cond_full = nullptr; h->M.cond_offset = nullptr;
  h->M.cond_dfull = 0; h->M.cond_nfree = 0;

  const size_t vec_bytes = (size_t)C * D * sizeof(float);
  auto alloc = [&](void** p, size_t bytes) {
    cudaError_t e = dmalloc(p, bytes);
    if (e == cudaSuccess) e = cudaMemset(*p, 0, bytes);
    return e;
  };
  auto alloc_ipc = [&](void** p, size_t bytes) {  // cudaIpcGetMemHandle needs cudaMalloc memory
    cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaSuccess) e = cudaMemset(*p, 0, bytes);
    return e;
  };
#define ACK(call)                                             \
  do {                                                        \
    cudaError_t e_ = (call);                                  \
    if (e_ != cudaSuccess) {                                  \
      fail("%s failed: %s", #call, cudaGetErrorString(e_));   \
      b200mc_destroy(h);                                      \
      return 1;                                               \
    }                                                         \
  } while (0)
  ACK(alloc((void**)&h->cs, sizeof(ChainScalars) * C));
  ACK(alloc((void**)&h->cv.z, vec_bytes));
  ACK(alloc((void**)&h->cv.z_grad, vec_bytes));
  ACK(alloc((void**)&h->cv.inverse_mass_matrix, vec_bytes));
  ACK(alloc((void**)&h->cv.mass_matrix_sqrt, vec_bytes));
  ACK(alloc((void**)&h->cv.welford_mean, vec_bytes));
  ACK(alloc((void**)&h->cv.welford_m2, vec_bytes));
  if (c.dense_mass) ACK(alloc((void**)&h->cv.dense, sizeof(float) * 4 * (size_t)C * D * D));
  if (h->cond_count > 0) {
    ACK(alloc((void**)&h->cond_map, sizeof(int) * D));
    ACK(alloc((void**)&h->cond_index, sizeof(int) * h->cond_count));
    ACK(alloc((void**)&h->cond_full, sizeof(float) * 2 * (size_t)C * h->cond_dfull));
    ACK(alloc((void**)&h->cond_offset, sizeof(float) * C));
    ACK(cudaMemcpy(h->cond_map, cond_map_h.data(), sizeof(int) * D, cudaMemcpyHostToDevice));
    ACK(cudaMemcpy(h->cond_index, c.cond_index_host, sizeof(int) * h->cond_count, cudaMemcpyHostToDevice));
    h->M.cond_map = h->cond_map; h->M.cond_full = h->cond_full; h->M.cond_offset = h->cond_offset;
    h->M.cond_dfull = h->cond_dfull; h->M.cond_nfree = D;
    h->cfg.cond_index_host = nullptr;  // the caller's array is not referenced after create
  }

  P.num_blocks = 1;
  if (c.num_blocks > 1) {  // structured mass matrix: validate and upload the block tables
    if (!c.block_perm_host || !c.block_sizes_host || !c.block_dense_host) {
      b200mc_destroy(h); return fail("num_blocks > 1 needs block_perm/sizes/dense tables");
    }
    const int nb = c.num_blocks;
    std::vector<int> t(3 * (size_t)D + 2 * (size_t)nb, -1);
    int *of = t.data(), *pos = of + D, *perm = pos + D, *size = perm + D, *dn = size + nb;
    int a = 0, any_dense = 0;
    for (int b = 0; b < nb; ++b) {
      size[b] = c.block_sizes_host[b];
      dn[b] = c.block_dense_host[b] ? 1 : 0;
      any_dense |= dn[b];
      if (size[b] < 1 || a + size[b] > D) { b200mc_destroy(h); return fail("block sizes do not add up to D=%d", D); }
      for (int j = 0; j < size[b]; ++j, ++a) {
        const int f = c.block_perm_host[a];
        if (f < 0 || f >= D || of[f] != -1) { b200mc_destroy(h); return fail("block_perm is not a permutation of 0..D-1"); }
        perm[a] = f; of[f] = b; pos[f] = j;
      }
    }
    if (a != D) { b200mc_destroy(h); return fail("block sizes do not add up to D=%d", D); }
    if ((any_dense != 0) != (c.dense_mass != 0)) {
      b200mc_destroy(h); return fail("dense_mass must be set iff a block is dense");
    }
    ACK(dmalloc((void**)&h->blk_tables, t.size() * sizeof(int)));
    ACK(cudaMemcpy(h->blk_tables, t.data(), t.size() * sizeof(int), cudaMemcpyHostToDevice));
    P.num_blocks = nb;
    P.blk_of = h->blk_tables; P.blk_pos = P.blk_of + D; P.blk_perm = P.blk_pos + D;
    P.blk_size = P.blk_perm + D; P.blk_dense = P.blk_size + nb;
  }

  const int nslots = num_slots(P.max_depth, P.dense_mass);
  int batch_min = 64;  // north_star: from 64 chains on, X.Theta is a real dense contraction
  // with a dense mass matrix the batched whole-machine sweeps pay off much earlier (a single
  // CTA cannot stream its chain's 16 MB matrix at speed): e.g. 256 chains sharded over 8 GPUs
  if (c.family == B200MC_HORSESHOE && c.dense_mass) batch_min = 16;
  if (const char* e = getenv("B200MC_BATCH_MIN_CHAINS")) batch_min = atoi(e);
  const bool batch = (c.family == B200MC_HORSESHOE || c.family == B200MC_LOGISTIC) &&
                     h->cond_count == 0 && C >= batch_min && !(c.nranks > 1) && P.num_blocks <= 1;  // (> kBatchCols: column groups)
  if (c.family == B200MC_LOGISTIC && !batch) {
    if (D > kGlmMaxD) { b200mc_destroy(h); return fail("logistic: D=%d > %d not supported", D, kGlmMaxD); }
    if (!c.data1_dev) { b200mc_destroy(h); return fail("logistic: labels missing"); }
    // Re-tile X and the labels once (layout: glm_kernel.cuh); the caller's buffers are not
    // referenced after b200mc_create returns.
    const int nt_total = (c.N + kGlmTileRows - 1) / kGlmTileRows;
    h->glm_nt_total = nt_total;
    ACK(dmalloc((void**)&h->glm_xt, (size_t)nt_total * glm_tile_bytes(D)));
    ACK(kl::glm_retile(h->num_sms * 8, (const float*)c.data0_dev, (const uint8_t*)c.data1_dev, c.N, D,
                       nt_total, h->glm_xt, 0));
    g_launches++;
    ACK(cudaDeviceSynchronize());
    int ctas = nt_total < h->num_sms ? nt_total : h->num_sms;
    h->glm_ctas = ctas;
    // Shared memory holds a TMA ring (stages: a multiple of the consumer-warp count -- every
    // stage is owned by one warp -- two per warp when they fit) and, in what is left, tiles
    // PINNED for the whole launch (at most one per producer lane): they never go through
    // L2 again, which also shrinks the L2 working set of the ping-pong stream.
    // Budget: the 196 KB carve-out step (1 KB of it is reserved by the system), not the full
    // 227 KB -- the next step leaves 28 KB of L1, too little for the control warp's stack:
    // measured +4.4k cycles per sampler step (profiles/r01_glm_smem_sweep.txt).
    int budget = 196 * 1024 - 1024;
    if (const char* e = getenv("B200MC_GLM_SMEM_MAX")) budget = atoi(e);
    if (budget > h->max_smem ||
        glm_smem_layout(D, kGlmConsumerWarps, 0, P.max_depth, P.dense_mass).total > budget)
      budget = h->max_smem;
    auto fits = [&](int ns_, int np_) {
      return glm_smem_layout(D, ns_, np_, P.max_depth, P.dense_mass).total <= budget;
    };
    int ns = 2 * kGlmConsumerWarps;
    {
      // A slice that (nearly) fits the shared memory of the grid -- e.g. 581,012 rows sharded
      // over 8 GPUs: 16 tiles of 7.2 KB per CTA -- is PINNED instead of streamed: one ring stage
      // per consumer warp, every other tile resident for the whole launch, so a pass reads
      // shared memory only and the leapfrog is left with its serial exchange chain.
      const int per_cta = (nt_total + ctas - 1) / ctas;
      int np1 = 0;
      while (np1 < kGlmMaxStages && fits(kGlmConsumerWarps, np1 + 1)) ++np1;
      if (per_cta <= np1 + kGlmConsumerWarps / 2) ns = kGlmConsumerWarps;
    }
    if (const char* e = getenv("B200MC_GLM_STAGES")) {  // tuning knobs
      const int v = atoi(e) / kGlmConsumerWarps * kGlmConsumerWarps;
      if (v >= kGlmConsumerWarps && v <= kGlmMaxStages) ns = v;
    }
    while (ns > kGlmConsumerWarps && !fits(ns, 0)) ns -= kGlmConsumerWarps;
    if (!fits(ns, 0)) {
      b200mc_destroy(h);
      return fail("logistic: shared memory layout does not fit");
    }
    int np = 0;
    while (np < kGlmMaxStages && fits(ns, np + 1)) ++np;
    if (const char* e = getenv("B200MC_GLM_PINNED")) {
      const int v = atoi(e);
      if (v >= 0 && v < np) np = v;
    }
    h->glm_stages = ns;
    h->glm_pinned = np;
    h->glm_smem = glm_smem_layout(D, ns, np, P.max_depth, P.dense_mass).total;
    int occ = 0;
    ACK(kl::glm_prepare(h->max_smem, h->glm_smem, &occ));
    if (occ < 1) { b200mc_destroy(h); return fail("logistic kernel does not fit on an SM"); }
    ACK(alloc((void**)&h->partials, sizeof(float) * 2 * (size_t)ctas * glm_partial_stride(D)));
    ACK(alloc((void**)&h->counters, sizeof(unsigned) * 8));
    ACK(alloc((void**)&h->prof, sizeof(unsigned long long) * kGlmProfSlots));
    if (c.dense_mass && ctas > 1)
      ACK(alloc((void**)&h->dense_replicas, sizeof(float) * 4 * (size_t)(ctas - 1) * D * D));
    if (const char* e = getenv("B200MC_GLM_REVERSE")) h->glm_reverse = atoi(e);
    {
      // 2+ chains (below the lock-step threshold): up to kMcChains chains share one stream of X
      // per launch.  B200MC_GLM_MC=0 turns it off (one launch per chain, X streamed per chain);
      // =1 forces it for a single chain too (parity aid: a chain alone through the same kernel).
      int want = (C >= 2) ? 1 : 0;
      if (const char* e = getenv("B200MC_GLM_MC")) want = atoi(e);
      const bool can = D <= kMcMaxD && !c.dense_mass && !(c.nranks > 1) && P.num_blocks <= 1;
      if (want && can) {
        auto mfits = [&](int ns_, int np_) {
          return glm_mc_smem_layout(D, ns_, np_, P.max_depth).total <= budget;
        };
        // ring: every consumer warp owns its stages; three per warp when they fit, else two
        int mns = 3 * kMcConsumerWarps;
        if (const char* e = getenv("B200MC_GLM_MC_STAGES")) {
          const int v = atoi(e) / kMcConsumerWarps * kMcConsumerWarps;
          if (v >= kMcConsumerWarps && v <= kMcMaxStages) mns = v;
        }
        while (mns > kMcConsumerWarps && !mfits(mns, 0)) mns -= kMcConsumerWarps;
        if (mfits(mns, 0)) {
          int mnp = 0;
          while (mnp < kMcMaxStages && mfits(mns, mnp + 1)) ++mnp;
          if (const char* e = getenv("B200MC_GLM_MC_PINNED")) {
            const int v = atoi(e);
            if (v >= 0 && v < mnp) mnp = v;
          }
          h->mc_stages = mns;
          h->mc_pinned = mnp;
          h->mc_smem = glm_mc_smem_layout(D, mns, mnp, P.max_depth).total;
          int mocc = 0;
          ACK(kl::glm_mc_prepare(h->max_smem, h->mc_smem, &mocc));
          if (mocc >= 1) {
            // this handle runs the multi-chain kernel only: its own 16-row image replaces the
            // 32-row one
            h->mc_nt_total = (c.N + kMcTileRows - 1) / kMcTileRows;
            h->mc_ctas = h->mc_nt_total < h->num_sms ? h->mc_nt_total : h->num_sms;
            ACK(dmalloc((void**)&h->mc_xt, (size_t)h->mc_nt_total * glm_mc_tile_bytes(D)));
            ACK(kl::glm_mc_retile(h->num_sms * 8, (const float*)c.data0_dev, (const uint8_t*)c.data1_dev,
                                  c.N, D, h->mc_nt_total, h->mc_xt, 0));
            g_launches++;
            ACK(cudaDeviceSynchronize());
            dfree(h->glm_xt);
            h->glm_xt = nullptr;
            ACK(alloc((void**)&h->mc_partials,
                      sizeof(float) * 2 * (size_t)h->mc_ctas * kMcChains * glm_partial_stride(D)));
            h->glm_mc = true;
          }
        }
      }
    }
    h->nranks = c.nranks > 1 ? c.nranks : 1;
    h->rank = c.nranks > 1 ? c.rank : 0;
    if (h->nranks > kGlmMaxRanks || h->rank < 0 || h->rank >= h->nranks) {
      b200mc_destroy(h);
      return fail("row sharding: need 0 <= rank < nranks <= %d", kGlmMaxRanks);
    }
    if (h->nranks > 1) {
      if (C != 1) { b200mc_destroy(h); return fail("row sharding runs one chain per handle"); }
      ACK(alloc_ipc((void**)&h->xchg_local,
                sizeof(float) * 2 * 2 * (size_t)h->nranks * glm_partial_stride(D)));
      ACK(alloc_ipc((void**)&h->xarrive_local, sizeof(unsigned) * 4));
    }
  } else {
    if (c.family == B200MC_HORSESHOE) {
      if (!c.data1_dev) { b200mc_destroy(h); return fail("horseshoe: Y missing"); }
      if (!batch) ACK(alloc((void**)&h->M.scratch, sizeof(float) * (size_t)C * c.N));
    }
    if (c.family == B200MC_GLM) {
      if (!c.data1_dev) { b200mc_destroy(h); return fail("glm: y missing"); }
      ACK(alloc((void**)&h->M.scratch, sizeof(float) * (size_t)C * c.N));
    }
    // thread per chain (tiny D, diagonal mass), warp per chain, CTA per chain; or the
    // lock-step tensor-core engine (batch_kernel.cuh)
    h->mode = batch ? 3 : ((D <= 32 && !c.dense_mass) ? 0 : (D <= 512 ? 1 : 2));
    // a GLM's potential is N x ncols work: a warp per chain from a few thousand products on,
    // a CTA per chain beyond that, whatever D is
    if (c.family == B200MC_GLM && (long)c.N * c.ncols >= 2048)
      h->mode = ((long)c.N * c.ncols >= 65536 || D > 512) ? 2 : 1;
    if (c.family == B200MC_LOGISTIC_ECS) {
      const int m = c.subsample_size, deg = c.ecs_degree;
      if (!c.data0_dev || !c.data1_dev) { b200mc_destroy(h); return fail("logistic_ecs: data missing"); }
      if (m < 1 || m >= c.N) { b200mc_destroy(h); return fail("logistic_ecs: need 0 < subsample_size < N"); }
      if (deg < 0 || deg > 2 || (deg > 0 && !c.ecs_ref_dev)) {
        b200mc_destroy(h);
        return fail("logistic_ecs: ecs_degree must be 0, 1 or 2 (with ecs_ref_dev for 1, 2)");
      }
      if (deg == 2 && D > 62) { b200mc_destroy(h); return fail("logistic_ecs: second-order proxy needs D <= 62"); }
      h->mode = 2;  // a CTA per chain: the subsample's rows are spread over its threads
      h->M.y8 = (const uint8_t*)c.data1_dev;
      h->M.ecs_m = m;
      h->M.ecs_degree = deg;
      h->M.scratch_stride = 2L * m + (long)kEcsSlices * D;
      ACK(alloc((void**)&h->M.scratch, sizeof(float) * (size_t)C * h->M.scratch_stride));
      ACK(alloc((void**)&h->ecs_idx, sizeof(int) * (size_t)C * m));
      ACK(alloc((void**)&h->ecs_prop, sizeof(int) * (size_t)C * m));
      ACK(alloc((void**)&h->ecs_gtmp, sizeof(float) * (size_t)C * D));
      h->ecs_hcap = 16;
      while (h->ecs_hcap < 4 * m) h->ecs_hcap *= 2;
      ACK(alloc((void**)&h->ecs_hash, sizeof(int) * 2 * (size_t)C * h->ecs_hcap));
      ACK(alloc((void**)&h->ecs_sums, sizeof(float) * (size_t)(1 + D + D * D)));
      h->M.ecs_idx = h->ecs_idx;
      h->M.ecs_sums = h->ecs_sums;
      if (deg > 0) {
        ACK(alloc((void**)&h->ecs_ref, sizeof(float) * D));
        ACK(cudaMemcpy(h->ecs_ref, c.ecs_ref_dev, sizeof(float) * D, cudaMemcpyDeviceToDevice));
        h->M.ecs_ref = h->ecs_ref;
        const int nparts = 4 * h->num_sms;
        float* parts = nullptr;
        ACK(alloc((void**)&parts, sizeof(float) * (size_t)nparts * (1 + D + D * D)));
        cudaError_t e = kl::ecs_taylor_sums_launch(h->M.a, h->M.y8, c.N, D, h->ecs_ref, deg, parts, nparts,
                                                   h->ecs_sums, 0);
        g_launches += 2;
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        dfree(parts);
        ACK(e);
      }
      // until b200mc_ecs_init_subsample / _set_subsample: rows 0 .. m-1
      std::vector<int> id0((size_t)C * m);
      for (size_t i = 0; i < id0.size(); ++i) id0[i] = (int)(i % (size_t)m);
      ACK(cudaMemcpy(h->ecs_idx, id0.data(), sizeof(int) * id0.size(), cudaMemcpyHostToDevice));
    }
    if (const char* e = (c.family == B200MC_LOGISTIC_ECS) ? nullptr : getenv("B200MC_CHAIN_MODE")) {  // tuning aid: force the chain grouping
      const int m = atoi(e);
      if (!batch && ((m == 0 && D <= 32 && !c.dense_mass) || (m == 1 && D <= 512) || m == 2)) h->mode = m;
    }
    h->c_pad = (C + 31) / 32 * 32;
    const size_t chains = (h->mode == 0) ? (size_t)h->c_pad : (size_t)C;
    ACK(alloc((void**)&h->work, sizeof(float) * chains * nslots * D));
    if (batch) {
      auto& b = h->bt;
      b.RA1 = (c.N + kGemmM - 1) / kGemmM * kGemmM;
      b.K1 = (c.ncols + 31) / 32 * 32;
      b.RA2 = (c.ncols + kGemmM - 1) / kGemmM * kGemmM;
      const int kb2 = b.RA1 / 32, tiles2 = b.RA2 / kGemmM;
      b.splits = h->num_sms / tiles2;
      if (b.splits < 1) b.splits = 1;
      // The tensor core's f32 accumulator truncates instead of rounding to nearest: every MMA
      // pulls the running sum towards zero by ~half an ulp, so a long accumulation chain drifts
      // (covtype shape, 148 split-K ranges = 123 k-blocks = 1,476 MMAs per TMEM accumulator:
      // every gradient component 2.5e-5 low against the float64 oracle, same sign --
      // tests/test_gpu_baseline_shapes.py; at 5,000 rows the chain is 2 k-blocks and the bias
      // is invisible).  Chains are therefore capped at kMaxChainKB k-blocks; the split partials
      // are summed in f32 on the CUDA cores (round to nearest, fixed order).
      constexpr int kMaxChainKB = 64;
      const int need = (kb2 + kMaxChainKB - 1) / kMaxChainKB;
      if (b.splits < need) b.splits = (need + b.splits - 1) / b.splits * b.splits;  // whole waves
      if (b.splits > kb2) b.splits = kb2;
      if (build_image((const float*)c.data0_dev, c.N, c.ncols, 0, b.RA1, b.K1, &b.x_hi, &b.x_lo, 0) ||
          build_image((const float*)c.data0_dev, c.N, c.ncols, 1, b.RA2, b.RA1, &b.xt_hi, &b.xt_lo, 0)) {
        b200mc_destroy(h);
        return 1;
      }
      ACK(alloc((void**)&b.b_hi, (size_t)kBatchCols * b.K1 * 4));
      ACK(alloc((void**)&b.b_lo, (size_t)kBatchCols * b.K1 * 4));
      ACK(alloc((void**)&b.r_hi, (size_t)kBatchCols * b.RA1 * 4));
      ACK(alloc((void**)&b.r_lo, (size_t)kBatchCols * b.RA1 * 4));
      ACK(alloc((void**)&b.ll_part, sizeof(float) * (size_t)(b.RA1 / 32) * kBatchCols));
      ACK(alloc((void**)&b.gpart, sizeof(float) * (size_t)b.splits * kBatchCols * b.RA2));
      ACK(alloc((void**)&b.col_scale, sizeof(float) * kBatchCols));
      b.snap_stride = (int)((sizeof(Sampler<BlockGroup>) + 255) / 256 * 256);
      ACK(alloc((void**)&b.snap, (size_t)C * b.snap_stride));
      ACK(alloc((void**)&b.status, sizeof(int) * C));
      ACK(alloc((void**)&b.reqs, sizeof(MatvecRequest) * C));
      ACK(alloc((void**)&b.active, sizeof(unsigned) * (kBatchStepsPerSync + 2)));
      if (c.family == B200MC_LOGISTIC) {  // labels arrive as bytes; the epilogue reads floats
        if (!c.data1_dev) { b200mc_destroy(h); return fail("logistic: labels missing"); }
        ACK(alloc((void**)&b.y_f32, sizeof(float) * (size_t)c.N));
        labels_to_float_kernel<<<592, 256>>>((const uint8_t*)c.data1_dev, c.N, b.y_f32);
        g_launches++;
        h->M.b = b.y_f32;
        h->M.likelihood = 0;
      }
      ACK(cudaDeviceSynchronize());
    }
  }
#undef ACK
  *out = h;
  return 0;
}

// ---------------------------------------------------------------- small utility kernels
__global__ void init_vectors_kernel(ChainVectors cv, const float* imm_in, const float* z_in, int C,
                                    int D) {
  const long n = (long)C * D;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const int d = (int)(i % D);
    // _initialize_mass_matrix (hmc_util.py:488-515), diagonal branch
    const float im = imm_in ? imm_in[d] : 1.0f;
    const float sq_inv = sqrtf(im);
    cv.inverse_mass_matrix[i] = im;
    cv.mass_matrix_sqrt[i] = imm_in ? 1.0f / sq_inv : 1.0f;
    cv.welford_mean[i] = 0.f;
    cv.welford_m2[i] = 0.f;
    cv.z[i] = z_in ? z_in[i] : 0.f;
    cv.z_grad[i] = 0.f;
  }
}

// _initialize_mass_matrix, dense branch (hmc_util.py:488-507): identity, or the user's M^-1
// in slot 0 (its factors are computed on device by Sampler::factor_dense in init_kernel)
__global__ void init_dense_kernel(float* dense, const float* imm_in, int C, int D) {
  const long dd = (long)D * D, n = (long)C * 4 * dd;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const long e = i % dd;
    const int which = (int)((i / dd) % 4), r = (int)(e / D), col = (int)(e % D);
    float v = 0.f;
    if (which == 0) v = imm_in ? imm_in[e] : (r == col ? 1.f : 0.f);
    else if (which < 3) v = (r == col) ? 1.f : 0.f;
    dense[i] = v;
  }
}

static CollectSpec make_collect(int num_iters, int collect_start, int thinning,
                                const b200mc_collect* o) {
  CollectSpec c;
  memset(&c, 0, sizeof(c));
  c.num_iters = num_iters;
  c.collect_start = collect_start;
  c.thinning = thinning < 1 ? 1 : thinning;
  if (o) {
    c.n_keep = o->n_keep;
    c.z = o->z_dev;
    c.potential_energy = o->potential_energy_dev;
    c.energy = o->energy_dev;
    c.num_steps = o->num_steps_dev;
    c.accept_prob = o->accept_prob_dev;
    c.mean_accept_prob = o->mean_accept_prob_dev;
    c.diverging = o->diverging_dev;
    c.step_size = o->step_size_dev;
  }
  return c;
}

static int launch_chain(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys,
                        int user_init, cudaStream_t st, const EcsStep* ecs = nullptr) {
  ChainLaunch L;
  if (ecs) L.ecs = *ecs;
  else memset(&L.ecs, 0, sizeof(L.ecs));
  L.P = h->P; L.M = h->M; L.col = col; L.cs = h->cs; L.cv = h->cv; L.work = h->work;
  L.init_keys = keys; L.user_init = user_init; L.C = h->C; L.c_pad = h->c_pad;
  {
    static int thresh = -1;
    if (thresh < 0) {
      const char* e = getenv("B200MC_END_THRESH");
      thresh = e ? atoi(e) : 28;  // sweep in profiles/r01_chain_kernel_tuning.txt
      if (thresh < 1) thresh = 1;
      if (thresh > 32) thresh = 32;
    }
    L.end_thresh = thresh;
  }
  L.hot_mask = 0u;
  L.cpw = 32;
  if (L.ecs.enabled && !keys) {
    // HMCECS loop: step machine in one warp, state and subsample rows in shared memory, when
    // they fit (B200MC_ECS_FAST=0: the CTA-per-chain kernel below)
    static int fast_on = -1;
    if (fast_on < 0) {
      const char* e = getenv("B200MC_ECS_FAST");
      fast_on = e ? atoi(e) : 1;
    }
    const size_t smem = ecs_chain_smem(h->D, h->P.max_depth, h->P.dense_mass, h->M.ecs_m, h->M.ncols);
    if (fast_on && h->D <= kEcsCtaCols && h->M.ecs_m <= kEcsCtaRows * 256 && !h->P.dense_mass &&
        smem + kl::ecs_chain_static_smem() <= (size_t)h->max_smem) {
      L.hot_mask = 0u;
      L.ecs.rows_in_smem = 1;
      CK(kl::ecs_chain_launch(L, h->C, smem, st));
      g_launches++;
      return 0;
    }
  }
  if (h->mode == 0) {
    const int threads = 64;  // small blocks spread the chains over all SMs
    // chains per warp (B200MC_CHAINS_PER_WARP overrides)
    static int cpw_env = -1;
    if (cpw_env < 0) {
      const char* e = getenv("B200MC_CHAINS_PER_WARP");
      cpw_env = e ? atoi(e) : 0;
    }
    // Measured on eight schools (profiles/r02_chains_per_warp.txt): a launch takes
    // max(latency of one warp, warps x issue cost of a warp); both grow with the chains per warp
    // (divergence), so few chains want a warp each and many chains want full warps.  Relative
    // figures of that sweep (ms, 148 SMs):
    static const float kLat[6] = {24.5f, 28.8f, 28.5f, 28.3f, 30.0f, 32.2f};      // cpw 1, 2, .., 32
    static const float kIssue[6] = {0.020f, 0.025f, 0.031f, 0.040f, 0.053f, 0.070f};
    int cpw = 32;
    float best = 1e30f;
    for (int i = 0; i < 6; ++i) {
      const int c = 1 << i;
      const float warps = (float)((h->C + c - 1) / c) * (148.0f / (float)h->num_sms);
      const float t = fmaxf(kLat[i], warps * kIssue[i]);
      if (t < best) { best = t; cpw = c; }
    }
    if (cpw_env >= 1 && cpw_env <= 32) cpw = cpw_env;
    L.cpw = cpw;
    const int warps = (h->C + cpw - 1) / cpw;
    CK(kl::chain_launch(0, L, (warps * 32 + threads - 1) / threads, threads, 0, st));
  } else if (h->mode == 1) {
    const int wpb = 4;
    CK(kl::chain_launch(1, L, (h->C + wpb - 1) / wpb, wpb * 32, 0, st));
  } else {
    // hot slots in shared memory when four resident CTAs per SM still fit (B200MC_CHAIN_HOT=0: off)
    static int hot_on = -1;
    if (hot_on < 0) {
      const char* e = getenv("B200MC_CHAIN_HOT");
      hot_on = e ? atoi(e) : 1;
    }
    size_t smem = 0;
    if (hot_on && !h->P.dense_mass && 4 * (chain_hot_bytes(h->D, kChainHotMask) + 1024) <= (size_t)h->max_smem) {
      L.hot_mask = kChainHotMask;
      smem = chain_hot_bytes(h->D, kChainHotMask);
    }
    if (L.ecs.enabled) {
      const size_t rows = sizeof(float) * (size_t)h->M.ecs_m * h->M.ncols;
      static int rows_on = -1;  // B200MC_ECS_ROWS_SMEM=0: gather the rows from X every time (A/B aid)
      if (rows_on < 0) {
        const char* e = getenv("B200MC_ECS_ROWS_SMEM");
        rows_on = e ? atoi(e) : 1;
      }
      L.ecs.rows_in_smem = (rows_on && smem + rows + 2048 <= (size_t)h->max_smem) ? 1 : 0;
      if (L.ecs.rows_in_smem) smem += rows;
    }
    CK(kl::chain_launch(2, L, h->C, kChainCtaThreads, smem, st));
  }
  g_launches++;
  return 0;
}

// ---------------------------------------------------------------- window-end factorisation
// Dense mass, lock-step engine: the Cholesky factor and the triangular inverse of a parked
// chain's covariance (hmc_util.py:229-233: jnp.linalg.cholesky, solve_triangular) come from
// cuSOLVER, a plain library call five times per warm-up and chain, not a hot path.  The
// library is dlopen'ed on first use so that libb200mc.so itself keeps no dependency on it.
namespace {
struct Solver {
  void* lib = nullptr;
  cusolverDnHandle_t handle = nullptr;
  decltype(&cusolverDnCreate) create = nullptr;
  decltype(&cusolverDnSetStream) set_stream = nullptr;
  decltype(&cusolverDnSpotrf_bufferSize) potrf_size = nullptr;
  decltype(&cusolverDnSpotrf) potrf = nullptr;
  decltype(&cusolverDnXtrtri_bufferSize) trtri_size = nullptr;
  decltype(&cusolverDnXtrtri) trtri = nullptr;
  void* work = nullptr;
  size_t work_bytes = 0;
  void* host_work = nullptr;
  size_t host_bytes = 0;
  int* info = nullptr;
};
Solver* get_solver() {
  static Solver S;
  if (S.handle) return &S;
  const char* names[] = {"libcusolver.so.11", "libcusolver.so", "/usr/local/cuda/lib64/libcusolver.so.11"};
  for (const char* n : names)
    if (!S.lib) S.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
  if (!S.lib) { fail("dense mass: cannot load libcusolver (%s)", dlerror()); return nullptr; }
#define SYM(field, name)                                                   \
  S.field = (decltype(S.field))dlsym(S.lib, name);                         \
  if (!S.field) { fail("dense mass: libcusolver lacks %s", name); return nullptr; }
  SYM(create, "cusolverDnCreate") SYM(set_stream, "cusolverDnSetStream")
  SYM(potrf_size, "cusolverDnSpotrf_bufferSize") SYM(potrf, "cusolverDnSpotrf")
  SYM(trtri_size, "cusolverDnXtrtri_bufferSize") SYM(trtri, "cusolverDnXtrtri")
#undef SYM
  if (S.create(&S.handle) != CUSOLVER_STATUS_SUCCESS) { fail("cusolverDnCreate failed"); S.handle = nullptr; return nullptr; }
  if (cudaMalloc((void**)&S.info, sizeof(int) * 2) != cudaSuccess) { fail("out of memory"); return nullptr; }
  return &S;
}
}  // namespace

// mats: one chain's [4][D][D]; n = Welford count.  Everything on stream st.
static int factor_chain_dense(float* mats, int D, int n, int regularize, cudaStream_t st) {
  Solver* S = get_solver();
  if (!S) return 1;
  if (S->set_stream(S->handle, st) != CUSOLVER_STATUS_SUCCESS) return fail("cusolverDnSetStream failed");
  const long DD = (long)D * D;
  float *msq = mats + DD, *W = mats + 2 * DD, *m2 = mats + 3 * DD;
  int lwork = 0;
  size_t dev_b = 0, host_b = 0;
  if (S->potrf_size(S->handle, CUBLAS_FILL_MODE_LOWER, D, W, D, &lwork) != CUSOLVER_STATUS_SUCCESS)
    return fail("cusolverDnSpotrf_bufferSize failed");
  if (S->trtri_size(S->handle, CUBLAS_FILL_MODE_UPPER, CUBLAS_DIAG_NON_UNIT, D, CUDA_R_32F, msq, D, &dev_b,
                    &host_b) != CUSOLVER_STATUS_SUCCESS)
    return fail("cusolverDnXtrtri_bufferSize failed");
  size_t need = sizeof(float) * (size_t)lwork;
  if (dev_b > need) need = dev_b;
  if (need > S->work_bytes) {
    cudaFree(S->work);
    CK(cudaMalloc(&S->work, need));
    S->work_bytes = need;
  }
  if (host_b > S->host_bytes) {
    free(S->host_work);
    S->host_work = malloc(host_b);
    S->host_bytes = host_b;
  }
  CK(kl::factor_cov_launch(mats, D, (float)n, regularize, st));
  // row-major symmetric W == column-major W; LOWER (column-major) leaves Lc[a][b] at W[b*D + a]
  if (S->potrf(S->handle, CUBLAS_FILL_MODE_LOWER, D, W, D, (float*)S->work, lwork, S->info) != CUSOLVER_STATUS_SUCCESS)
    return fail("cusolverDnSpotrf failed");
  CK(kl::factor_tinv_launch(mats, D, st));
  CK(cudaMemcpyAsync(W, m2, sizeof(float) * DD, cudaMemcpyDeviceToDevice, st));  // mass_matrix_sqrt_inv
  CK(cudaMemsetAsync(m2, 0, sizeof(float) * DD, st));                             // Welford restarts
  // row-major lower tril_inv == column-major UPPER of its transpose: inverting that in place
  // leaves the row-major lower inverse, mass_matrix_sqrt (:231-233)
  if (S->trtri(S->handle, CUBLAS_FILL_MODE_UPPER, CUBLAS_DIAG_NON_UNIT, D, CUDA_R_32F, msq, D, S->work, dev_b,
               S->host_work, host_b, S->info + 1) != CUSOLVER_STATUS_SUCCESS)
    return fail("cusolverDnXtrtri failed");
  g_launches += 2;
  CK(cudaGetLastError());
  return 0;
}

// The two contractions of one batched potential evaluation (batch_kernel.cuh).
static int launch_batch_gemms(b200mc_handle* h, cudaStream_t st) {
  auto& b = h->bt;
  Gemm3xLaunch g1 = {};
  g1.a_hi = b.x_hi; g1.a_lo = b.x_lo; g1.b_hi = b.b_hi; g1.b_lo = b.b_lo;
  g1.RA = b.RA1; g1.KB = b.K1 / 32; g1.splits = 1;
  g1.y = h->M.b; g1.n_valid = h->M.N; g1.likelihood = h->M.likelihood; g1.col_scale = b.col_scale;
  g1.r_hi = b.r_hi; g1.r_lo = b.r_lo; g1.ll_part = b.ll_part;
  if (launch_gemm3x(g1, EPI_RESID, st)) return 1;
  Gemm3xLaunch g2 = {};
  g2.a_hi = b.xt_hi; g2.a_lo = b.xt_lo; g2.b_hi = b.r_hi; g2.b_lo = b.r_lo;
  g2.RA = b.RA2; g2.KB = b.RA1 / 32; g2.splits = b.splits; g2.out = b.gpart;
  return launch_gemm3x(g2, EPI_STORE, st);
}

static void fill_batch_launch(b200mc_handle* h, BatchLaunch& L) {
  auto& b = h->bt;
  memset(&L, 0, sizeof(L));
  L.P = h->P; L.M = h->M; L.cs = h->cs; L.cv = h->cv; L.work = h->work;
  L.C = h->C < kBatchCols ? h->C : kBatchCols; L.c0 = 0; L.C_total = h->C;
  L.snap = b.snap; L.snap_stride = b.snap_stride; L.status = b.status; L.active = b.active;
  L.reqs = b.reqs;
  L.b_hi = b.b_hi; L.b_lo = b.b_lo; L.K1 = b.K1;
  L.gpart = b.gpart; L.splits = b.splits; L.RA2 = b.RA2;
  L.ll_part = b.ll_part; L.n_rowgroups = b.RA1 / 32; L.col_scale = b.col_scale;
}

// Lock-step engine: every global step evaluates the potential of ALL chains with two tensor-core
// contractions and advances every chain by one gradient.  The host only looks at the device
// every kBatchStepsPerSync steps (how many chains still wait for a gradient), so this call
// returns after the work is DONE, unlike the single-launch paths.
static int launch_batch_group(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys,
                              int user_init, int c0, int Cg, cudaStream_t st);

// More than kBatchCols chains: one column group of the operand images after the other, each
// run to the end of the call's quota (chains are independent; a group's chains keep their
// global indices in every per-chain array and in the collected draws).
static int launch_batch(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys,
                        int user_init, cudaStream_t st) {
  for (int c0 = 0; c0 < h->C; c0 += kBatchCols) {
    const int Cg = (h->C - c0 < kBatchCols) ? h->C - c0 : kBatchCols;
    if (launch_batch_group(h, col, keys, user_init, c0, Cg, st)) return 1;
  }
  return 0;
}

static int launch_batch_group(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys,
                              int user_init, int c0, int Cg, cudaStream_t st) {
  BatchLaunch L;
  fill_batch_launch(h, L);
  L.col = col; L.init_keys = keys; L.user_init = user_init;
  L.c0 = c0; L.C = Cg;
  auto& b = h->bt;
  unsigned* const slot_begin = b.active + kBatchStepsPerSync;     // count after the first launch
  unsigned* const slot_scrap = b.active + kBatchStepsPerSync + 1;  // launches nobody reads
  const bool dense = h->P.dense_mass != 0;
  std::vector<int> status_h(Cg);
  std::vector<MatvecRequest> reqs_h(Cg);
  const dim3 mv_grid((h->D + kMatvecRows - 1) / kMatvecRows, Cg);
  const size_t mv_smem = sizeof(float) * 2 * (size_t)h->D;
  // a chain needs at most three sweeps between two gradients (leaf, then the two of a
  // transition start); chains that still wait afterwards simply continue in the next step
  auto sweeps_and_resumes = [&](unsigned* count_slot) -> int {
    for (int k = 0; k < 3; ++k) {
      CK(kl::batch_matvec_launch(L, mv_grid, mv_smem, st));
      L.kind = BATCH_RESUME;
      L.active = (k == 2) ? count_slot : slot_scrap;
      CK(kl::batch_step_launch(L, Cg, st));
      g_launches += 2;
    }
    return 0;
  };
  CK(cudaMemsetAsync(b.active, 0, sizeof(unsigned) * (kBatchStepsPerSync + 2), st));
  L.kind = BATCH_BEGIN;
  L.active = dense ? slot_scrap : slot_begin;
  CK(kl::batch_step_launch(L, Cg, st));
  g_launches++;
  L.init_keys = nullptr;
  if (dense && sweeps_and_resumes(slot_begin)) return 1;
  CK(cudaGetLastError());
  unsigned active = 0;
  CK(cudaMemcpyAsync(&active, slot_begin, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  while (active > 0) {
    CK(cudaMemsetAsync(b.active, 0, sizeof(unsigned) * kBatchStepsPerSync, st));
    for (int i = 0; i < kBatchStepsPerSync; ++i) {
      if (launch_batch_gemms(h, st)) return 1;
      L.kind = BATCH_GRAD;
      L.active = dense ? slot_scrap : b.active + i;
      CK(kl::batch_step_launch(L, Cg, st));
      g_launches++;
      if (dense && sweeps_and_resumes(b.active + i)) return 1;
    }
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(&active, b.active + kBatchStepsPerSync - 1, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    if (dense && h->P.adapt_mass_matrix) {
      // chains parked at a window end: covariance, Cholesky, triangular inverse, then resume
      CK(cudaMemcpyAsync(status_h.data(), b.status + c0, sizeof(int) * Cg, cudaMemcpyDeviceToHost, st));
      CK(cudaMemcpyAsync(reqs_h.data(), b.reqs + c0, sizeof(MatvecRequest) * Cg, cudaMemcpyDeviceToHost, st));
      CK(cudaStreamSynchronize(st));
      bool any = false;
      for (int c = 0; c < Cg; ++c) {
        if (status_h[c] != 3) continue;
        if (factor_chain_dense(h->cv.dense + (long)(c0 + c) * 4 * h->D * h->D, h->D, reqs_h[c].src1,
                               h->P.regularize_mass_matrix, st))
          return 1;
        status_h[c] = 2;  // its request has mat = -1: the next resume launch just continues
        any = true;
      }
      if (any) CK(cudaMemcpyAsync(b.status + c0, status_h.data(), sizeof(int) * Cg, cudaMemcpyHostToDevice, st));
    }
    CK(cudaStreamSynchronize(st));
  }
  return 0;
}

static int launch_glm(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys, int user_init,
                      int chain, const float* eval_z, float* eval_U, float* eval_grad,
                      cudaStream_t st, int eval_passes = 1) {
  GlmLaunch L;
  memset(&L, 0, sizeof(L));
  L.P = h->P; L.col = col; L.cs = h->cs; L.cv = h->cv;
  L.Xt = h->glm_xt;
  L.N = h->cfg.N; L.D = h->D;
  L.nt_total = h->glm_nt_total;
  L.n_stages = h->glm_stages;
  L.n_pinned = h->glm_pinned;
  L.reverse_odd = h->glm_reverse;
  L.partials = h->partials;
  L.arrive = h->counters + 0;
  L.error = (int*)(h->counters + 3);
  L.prof = h->prof;
  L.dense_replicas = h->dense_replicas;
  L.nranks = h->nranks;
  L.rank = h->rank;
  if (h->nranks > 1) {
    if (!h->comm_ready) return fail("row sharding: b200mc_comm_import has not been called");
    for (int r = 0; r < h->nranks; ++r) {
      L.xchg[r] = h->xchg[r];
      L.xarrive[r] = h->xarrive[r];
    }
  }
  L.init_keys = keys; L.user_init = user_init; L.chain = chain; L.C = h->C;
  L.eval_z = eval_z; L.eval_U = eval_U; L.eval_grad = eval_grad; L.eval_passes = eval_passes;
  CK(cudaMemsetAsync(h->counters, 0, sizeof(unsigned) * 3, st));  // arrive
  CK(kl::glm_launch(L, h->glm_ctas, (size_t)h->glm_smem, st));
  g_launches++;
  return 0;
}

static int launch_glm_mc(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys, int user_init,
                         int chain0, int n_live, const float* eval_z, float* eval_U, float* eval_grad,
                         cudaStream_t st) {
  GlmMcLaunch L;
  memset(&L, 0, sizeof(L));
  L.P = h->P; L.col = col; L.cs = h->cs; L.cv = h->cv;
  L.Xt = h->mc_xt;
  L.N = h->cfg.N; L.D = h->D;
  L.nt_total = h->mc_nt_total;
  L.n_stages = h->mc_stages;
  L.n_pinned = h->mc_pinned;
  L.reverse_odd = h->glm_reverse;
  L.partials = h->mc_partials;
  L.arrive = h->counters + 0;
  L.error = (int*)(h->counters + 3);
  L.prof = h->prof;
  L.init_keys = keys; L.user_init = user_init;
  L.chain0 = chain0; L.n_live = n_live; L.C = h->C;
  L.eval_z = eval_z; L.eval_U = eval_U; L.eval_grad = eval_grad;
  CK(cudaMemsetAsync(h->counters, 0, sizeof(unsigned) * 3, st));
  CK(kl::glm_mc_launch(L, h->mc_ctas, (size_t)h->mc_smem, st));
  g_launches++;
  return 0;
}

// all chains of a tall-data logistic handle: groups of kMcChains on one stream, or one by one
static int launch_glm_all(b200mc_handle* h, const CollectSpec& col, const uint32_t* keys, int user_init,
                          cudaStream_t st) {
  if (h->glm_mc) {
    for (int c0 = 0; c0 < h->C; c0 += kMcChains) {
      const int n = h->C - c0 < kMcChains ? h->C - c0 : kMcChains;
      if (launch_glm_mc(h, col, keys, user_init, c0, n, nullptr, nullptr, nullptr, st)) return 1;
    }
    return 0;
  }
  for (int c = 0; c < h->C; ++c)
    if (launch_glm(h, col, keys, user_init, c, nullptr, nullptr, nullptr, st)) return 1;
  return 0;
}

extern "C" int b200mc_init(b200mc_handle* h, const uint32_t* keys_dev, const float* init_z_dev,
                           void* stream) {
  if (!h || !keys_dev) return fail("null argument");
  cudaStream_t st = (cudaStream_t)stream;
  init_vectors_kernel<<<148, 256, 0, st>>>(
      h->cv, h->cfg.dense_mass ? nullptr : h->cfg.inverse_mass_matrix_dev, init_z_dev, h->C, h->D);
  g_launches++;
  if (h->cfg.dense_mass) {
    init_dense_kernel<<<148, 256, 0, st>>>(h->cv.dense, h->cfg.inverse_mass_matrix_dev, h->C, h->D);
    g_launches++;
  }
  CK(cudaGetLastError());
  CK(cudaMemsetAsync(h->cs, 0, sizeof(ChainScalars) * h->C, st));
  const CollectSpec col = make_collect(0, 0, 1, nullptr);
  const int user_init = init_z_dev ? 1 : 0;
  if (h->mode == 3) {
    if (launch_batch(h, col, keys_dev, user_init, st)) return 1;
  } else if (h->cfg.family == B200MC_LOGISTIC) {
    if (launch_glm_all(h, col, keys_dev, user_init, st)) return 1;
  } else {
    if (launch_chain(h, col, keys_dev, user_init, st)) return 1;
  }
  h->initialized = true;
  return 0;
}

extern "C" int b200mc_sample(b200mc_handle* h, int num_iters, int collect_start, int thinning,
                             const b200mc_collect* out, void* stream) {
  if (!h) return fail("null handle");
  if (!h->initialized) return fail("b200mc_sample called before b200mc_init / b200mc_set_*");
  if (num_iters <= 0) return 0;
  if (thinning < 1) return fail("thinning must be a positive integer");
  cudaStream_t st = (cudaStream_t)stream;
  const CollectSpec col = make_collect(num_iters, collect_start, thinning, out);
  if (h->mode == 3) return launch_batch(h, col, nullptr, 0, st);
  if (h->cfg.family == B200MC_LOGISTIC) return launch_glm_all(h, col, nullptr, 0, st);
  return launch_chain(h, col, nullptr, 0, st);
}

__global__ void cond_scatter_kernel(float* full, const int* index, const float* values, int C,
                                    int count, int dfull) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= C * count) return;
  const int c = i / count, k = i - c * count;
  full[(long)c * 2 * dfull + index[k]] = values[i];
}

extern "C" int b200mc_set_conditioned(b200mc_handle* h, const float* values_dev, const float* offset_dev,
                                      void* stream) {
  if (!h || !values_dev) return fail("null argument");
  if (h->cond_count <= 0) return fail("the handle has no conditioned sites (b200mc_config.cond_count)");
  cudaStream_t st = (cudaStream_t)stream;
  const int n = h->C * h->cond_count;
  cond_scatter_kernel<<<(n + 255) / 256, 256, 0, st>>>(h->cond_full, h->cond_index, values_dev, h->C,
                                                      h->cond_count, h->cond_dfull);
  g_launches++;
  CK(cudaGetLastError());
  if (offset_dev)
    CK(cudaMemcpyAsync(h->cond_offset, offset_dev, sizeof(float) * h->C, cudaMemcpyDeviceToDevice, st));
  else
    CK(cudaMemsetAsync(h->cond_offset, 0, sizeof(float) * h->C, st));
  return 0;
}

extern "C" int b200mc_potential_and_grad(b200mc_handle* h, const float* z_dev, int n, float* U_dev,
                                         float* grad_dev, void* stream) {
  if (!h || !z_dev || !U_dev || !grad_dev) return fail("null argument");
  cudaStream_t st = (cudaStream_t)stream;
  if (h->mode == 3) {  // tensor-core path: up to 256 points per pair of contractions
    for (int i0 = 0; i0 < n; i0 += kBatchCols) {
      const int m = (n - i0 < kBatchCols) ? n - i0 : kBatchCols;
      BatchLaunch L;
      fill_batch_launch(h, L);
      L.C = m;
      L.eval_z = z_dev + (long)i0 * h->D;
      L.eval_U = U_dev + i0;
      L.eval_grad = grad_dev + (long)i0 * h->D;
      CK(kl::batch_potential_launch(L, m, 0, st));
      g_launches++;
      if (launch_batch_gemms(h, st)) return 1;
      CK(kl::batch_potential_launch(L, m, 1, st));
      g_launches++;
      CK(cudaGetLastError());
    }
    return 0;
  }
  if (h->cfg.family == B200MC_LOGISTIC && h->glm_mc) {  // kMcChains points per stream of X
    const CollectSpec col = make_collect(0, 0, 1, nullptr);
    for (int i0 = 0; i0 < n; i0 += kMcChains) {
      const int m = n - i0 < kMcChains ? n - i0 : kMcChains;
      if (launch_glm_mc(h, col, nullptr, 0, 0, m, z_dev + (long)i0 * h->D, U_dev + i0,
                        grad_dev + (long)i0 * h->D, st))
        return 1;
    }
    return 0;
  }
  if (h->cfg.family == B200MC_LOGISTIC) {
    const CollectSpec col = make_collect(0, 0, 1, nullptr);
    for (int i = 0; i < n; ++i)
      if (launch_glm(h, col, nullptr, 0, 0, z_dev + (long)i * h->D, U_dev + i,
                     grad_dev + (long)i * h->D, st))
        return 1;
    return 0;
  }
  if (h->cond_count > 0 && n > h->C)
    return fail("conditioned sites: point i is evaluated with chain i's values, so n <= num_chains");
  // per-chain scratch (horseshoe residuals) is sized for num_chains: evaluate in batches
  for (int i0 = 0; i0 < n; i0 += h->C) {
    const int m = (n - i0 < h->C) ? n - i0 : h->C;
    const float* zp = z_dev + (long)i0 * h->D;
    float* gp = grad_dev + (long)i0 * h->D;
    CK(kl::chain_potential_launch(h->mode, h->M, zp, U_dev + i0, gp, m, st));
    g_launches++;
  }
  return 0;
}

extern "C" int b200mc_potential_and_grad_passes(b200mc_handle* h, const float* z_dev, int passes,
                                                float* U_dev, float* grad_dev, void* stream) {
  if (!h || !z_dev || !U_dev || !grad_dev || passes < 1) return fail("bad argument");
  if (h->cfg.family != B200MC_LOGISTIC || h->mode == 3 || h->glm_mc)
    return fail("multi-pass evaluation exists for the one-chain streaming (tall-data logistic) kernel only");
  const CollectSpec col = make_collect(0, 0, 1, nullptr);
  return launch_glm(h, col, nullptr, 0, 0, z_dev, U_dev, grad_dev, (cudaStream_t)stream, passes);
}

// ---------------------------------------------------------------- HMCECS (hmc_gibbs.py:502-665)
static int ecs_check(b200mc_handle* h) {
  if (!h) return fail("null handle");
  if (h->cfg.family != B200MC_LOGISTIC_ECS) return fail("not a LOGISTIC_ECS handle");
  return 0;
}
static void fill_ecs_launch(b200mc_handle* h, EcsLaunch& L) {
  L.M = h->M; L.cs = h->cs; L.cv = h->cv; L.C = h->C; L.partitionable = h->P.partitionable;
  L.idx = h->ecs_idx; L.idx_prop = h->ecs_prop; L.grad_tmp = h->ecs_gtmp;
  L.keys = nullptr; L.accept_prob = nullptr; L.num_blocks = 1; L.key_u = nullptr;
  L.sites_before_plate = 1;  // `coefs` precedes plate `N` in examples/covtype.py:66-71
  L.hkeys = h->ecs_hash; L.hvals = h->ecs_hash + (size_t)h->C * h->ecs_hcap; L.hcap = h->ecs_hcap;
}
extern "C" int b200mc_ecs_init_subsample(b200mc_handle* h, const uint32_t* key_u_dev, void* stream) {
  if (ecs_check(h)) return 1;
  if (!key_u_dev) return fail("null argument");
  EcsLaunch L;
  fill_ecs_launch(h, L);
  L.key_u = key_u_dev;
  CK(kl::ecs_subsample_init_launch(L, (cudaStream_t)stream));
  g_launches++;
  return 0;
}
extern "C" int b200mc_ecs_get_subsample(b200mc_handle* h, int32_t* idx_dev, void* stream) {
  if (ecs_check(h)) return 1;
  CK(cudaMemcpyAsync(idx_dev, h->ecs_idx, sizeof(int) * (size_t)h->C * h->M.ecs_m,
                     cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_ecs_set_subsample(b200mc_handle* h, const int32_t* idx_dev, void* stream) {
  if (ecs_check(h)) return 1;
  CK(cudaMemcpyAsync(h->ecs_idx, idx_dev, sizeof(int) * (size_t)h->C * h->M.ecs_m,
                     cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_ecs_update(b200mc_handle* h, uint32_t* keys_dev, int num_blocks,
                                 float* accept_prob_dev, void* stream) {
  if (ecs_check(h)) return 1;
  if (!h->initialized) return fail("b200mc_ecs_update called before b200mc_init / b200mc_set_*");
  if (!keys_dev || !accept_prob_dev || num_blocks < 1) return fail("bad argument");
  EcsLaunch L;
  fill_ecs_launch(h, L);
  L.keys = keys_dev;
  L.accept_prob = accept_prob_dev;
  L.num_blocks = num_blocks;
  CK(kl::ecs_gibbs_launch(L, (cudaStream_t)stream));
  g_launches++;
  return 0;
}
extern "C" int b200mc_ecs_sample(b200mc_handle* h, uint32_t* keys_dev, int num_blocks, int num_iters,
                                 int collect_start, int thinning, const b200mc_collect* out,
                                 float* accept_prob_dev, float* accept_all_dev, void* stream) {
  if (ecs_check(h)) return 1;
  if (!h->initialized) return fail("b200mc_ecs_sample called before b200mc_init / b200mc_set_*");
  if (!keys_dev || !accept_prob_dev || num_blocks < 1) return fail("bad argument");
  if (num_iters <= 0) return 0;
  if (thinning < 1) return fail("thinning must be a positive integer");
  EcsStep E;
  E.enabled = 1; E.idx = h->ecs_idx; E.prop = h->ecs_prop; E.gtmp = h->ecs_gtmp; E.keys = keys_dev;
  E.accept_prob = accept_prob_dev; E.accept_all = accept_all_dev; E.num_blocks = num_blocks;
  E.partitionable = h->P.partitionable; E.rows_in_smem = 0;
  const CollectSpec col = make_collect(num_iters, collect_start, thinning, out);
  return launch_chain(h, col, nullptr, 0, (cudaStream_t)stream, &E);
}
extern "C" int b200mc_random_randint(const uint32_t* key_dev, int n, int minval, int maxval,
                                     int partitionable, int32_t* out_dev, void* stream) {
  if (!key_dev || !out_dev || n < 0) return fail("bad argument");
  CK(kl::ecs_randint_launch(key_dev, n, minval, maxval, partitionable, out_dev, (cudaStream_t)stream));
  g_launches++;
  return 0;
}

// ---------------------------------------------------------------- roofline microbenchmark
extern "C" int b200mc_membench(uint64_t bytes, int tile_bytes, int n_stages, int passes,
                               int reverse_odd, float* gbs_host, void* stream) {
  if (!gbs_host || tile_bytes < 512 || (tile_bytes & 15) || passes < 1) return fail("membench: bad argument");
  if (n_stages < kGlmConsumerWarps || n_stages > kGlmMaxStages || n_stages % kGlmConsumerWarps)
    return fail("membench: n_stages must be a multiple of %d, at most %d", kGlmConsumerWarps, kGlmMaxStages);
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 0, max_smem = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  CK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const int smem = n_stages * tile_bytes + 2 * kGlmMaxStages * 8;
  if (smem > max_smem) return fail("membench: ring of %d x %d bytes does not fit", n_stages, tile_bytes);
  MemBenchLaunch L;
  L.nt_total = (int)(bytes / (uint64_t)tile_bytes);
  if (L.nt_total < sms) return fail("membench: buffer too small");
  L.tile_bytes = tile_bytes; L.n_stages = n_stages; L.reverse_odd = reverse_odd;
  unsigned char* buf = nullptr;
  unsigned* sink = nullptr;
  CK(dmalloc((void**)&buf, (size_t)L.nt_total * tile_bytes));
  CK(dmalloc((void**)&sink, 512));
  CK(cudaMemsetAsync(buf, 1, (size_t)L.nt_total * tile_bytes, st));
  CK(cudaMemsetAsync(sink, 0, 512, st));
  L.buf = buf; L.sink = sink;
  CK(kl::membench_prepare(max_smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  L.passes = 2;  // warm-up: page in, fill L2
  CK(kl::membench_launch(L, sms, smem, st));
  L.passes = passes;
  CK(cudaEventRecord(e0, st));
  CK(kl::membench_launch(L, sms, smem, st));
  CK(cudaEventRecord(e1, st));
  g_launches += 2;
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  dfree(buf);
  dfree(sink);
  if (e != cudaSuccess) return fail("membench: %s", cudaGetErrorString(e));
  *gbs_host = (float)((double)passes * L.nt_total * tile_bytes / (ms * 1e-3) / 1e9);
  return 0;
}

// Launch geometry of the streaming (tall-data logistic) kernel of a handle: out[0] CTAs,
// [1] ring stages, [2] pinned tiles per CTA, [3] tile bytes, [4] tiles, [5] dynamic smem bytes,
// [6] chains per launch, [7] reserved.
extern "C" int b200mc_glm_layout(b200mc_handle* h, int* out_host) {
  if (!h || !out_host) return fail("null argument");
  if (!h->glm_xt && !h->mc_xt) return fail("handle does not run the streaming kernel");
  out_host[0] = h->glm_ctas; out_host[1] = h->glm_stages; out_host[2] = h->glm_pinned;
  out_host[3] = glm_tile_bytes(h->D); out_host[4] = h->glm_nt_total; out_host[5] = h->glm_smem;
  out_host[6] = h->glm_mc ? kMcChains : 1; out_host[7] = 0;
  if (h->glm_mc) {
    out_host[0] = h->mc_ctas; out_host[1] = h->mc_stages; out_host[2] = h->mc_pinned;
    out_host[3] = glm_mc_tile_bytes(h->D); out_host[4] = h->mc_nt_total; out_host[5] = h->mc_smem;
  }
  return 0;
}

// ---------------------------------------------------------------- state round trip
extern "C" int b200mc_get_scalars(b200mc_handle* h, b200mc_chain_scalars* out_host) {
  if (!h || !out_host) return fail("null argument");
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(out_host, h->cs, sizeof(ChainScalars) * h->C, cudaMemcpyDeviceToHost));
  if (h->counters) {
    int err = 0;
    CK(cudaMemcpy(&err, h->counters + 3, sizeof(int), cudaMemcpyDeviceToHost));
    if (err) CK(cudaMemset(h->counters + 3, 0, sizeof(int)));  // reported once, not sticky
    if (err == 2) return fail("cross-rank exchange timed out inside the streaming kernel");
    if (err) return fail("grid barrier timed out inside the streaming kernel");
  }
  return 0;
}

extern "C" int b200mc_set_scalars(b200mc_handle* h, const b200mc_chain_scalars* in_host) {
  if (!h || !in_host) return fail("null argument");
  CK(cudaMemcpy(h->cs, in_host, sizeof(ChainScalars) * h->C, cudaMemcpyHostToDevice));
  h->initialized = true;
  return 0;
}

extern "C" int b200mc_comm_export(b200mc_handle* h, void* out_host) {
  if (!h || !out_host) return fail("null argument");
  if (h->nranks <= 1) return fail("handle is not row sharded");
  static_assert(2 * sizeof(cudaIpcMemHandle_t) <= B200MC_COMM_HANDLE_BYTES, "handle size");
  cudaIpcMemHandle_t hs[2];
  CK(cudaIpcGetMemHandle(&hs[0], h->xchg_local));
  CK(cudaIpcGetMemHandle(&hs[1], h->xarrive_local));
  memcpy(out_host, hs, sizeof(hs));
  return 0;
}

extern "C" int b200mc_comm_import(b200mc_handle* h, const void* all_host) {
  if (!h || !all_host) return fail("null argument");
  if (h->nranks <= 1) return fail("handle is not row sharded");
  const char* base = (const char*)all_host;
  for (int r = 0; r < h->nranks; ++r) {
    if (r == h->rank) {
      h->xchg[r] = h->xchg_local;
      h->xarrive[r] = h->xarrive_local;
      continue;
    }
    cudaIpcMemHandle_t hs[2];
    memcpy(hs, base + (size_t)r * B200MC_COMM_HANDLE_BYTES, sizeof(hs));
    CK(cudaIpcOpenMemHandle((void**)&h->xchg[r], hs[0], cudaIpcMemLazyEnablePeerAccess));
    CK(cudaIpcOpenMemHandle((void**)&h->xarrive[r], hs[1], cudaIpcMemLazyEnablePeerAccess));
  }
  h->comm_ready = true;
  return 0;
}

extern "C" int b200mc_get_profile(b200mc_handle* h, uint64_t* out_host, int reset) {
  if (!h || !out_host) return fail("null argument");
  for (int i = 0; i < kGlmProfSlots; ++i) out_host[i] = 0;
  if (!h->prof) return 0;
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(out_host, h->prof, sizeof(unsigned long long) * kGlmProfSlots,
                cudaMemcpyDeviceToHost));
  if (reset) CK(cudaMemset(h->prof, 0, sizeof(unsigned long long) * kGlmProfSlots));
  return 0;
}

static float* vec_ptr(b200mc_handle* h, int which) {
  switch (which) {
    case B200MC_VEC_Z: return h->cv.z;
    case B200MC_VEC_Z_GRAD: return h->cv.z_grad;
    case B200MC_VEC_INVERSE_MASS_MATRIX: return h->cv.inverse_mass_matrix;
    case B200MC_VEC_MASS_MATRIX_SQRT: return h->cv.mass_matrix_sqrt;
    case B200MC_VEC_WELFORD_MEAN: return h->cv.welford_mean;
    case B200MC_VEC_WELFORD_M2: return h->cv.welford_m2;
    default: return nullptr;
  }
}

// dense matrices are stored [C][4][D][D]; the ABI exposes each kind as [C][D][D]
static int copy_matrix(b200mc_handle* h, int which, float* user, bool to_user, cudaStream_t st) {
  if (!h->cv.dense) return fail("matrix state exists only with dense_mass");
  const size_t dd = (size_t)h->D * h->D;
  float* base = h->cv.dense + (size_t)(which - B200MC_MAT_INVERSE_MASS_MATRIX) * dd;
  if (to_user)
    CK(cudaMemcpy2DAsync(user, dd * 4, base, 4 * dd * 4, dd * 4, h->C, cudaMemcpyDeviceToDevice, st));
  else
    CK(cudaMemcpy2DAsync(base, 4 * dd * 4, user, dd * 4, dd * 4, h->C, cudaMemcpyDeviceToDevice, st));
  return 0;
}

extern "C" int b200mc_get_vector(b200mc_handle* h, int which, float* out_dev, void* stream) {
  if (h && out_dev && which >= B200MC_MAT_INVERSE_MASS_MATRIX && which <= B200MC_MAT_WELFORD_M2)
    return copy_matrix(h, which, out_dev, true, (cudaStream_t)stream);
  float* p = h ? vec_ptr(h, which) : nullptr;
  if (!p || !out_dev) return fail("bad argument");
  CK(cudaMemcpyAsync(out_dev, p, sizeof(float) * (size_t)h->C * h->D, cudaMemcpyDeviceToDevice,
                     (cudaStream_t)stream));
  return 0;
}

extern "C" int b200mc_set_vector(b200mc_handle* h, int which, const float* in_dev, void* stream) {
  if (h && in_dev && which >= B200MC_MAT_INVERSE_MASS_MATRIX && which <= B200MC_MAT_WELFORD_M2)
    return copy_matrix(h, which, const_cast<float*>(in_dev), false, (cudaStream_t)stream);
  float* p = h ? vec_ptr(h, which) : nullptr;
  if (!p || !in_dev) return fail("bad argument");
  CK(cudaMemcpyAsync(p, in_dev, sizeof(float) * (size_t)h->C * h->D, cudaMemcpyDeviceToDevice,
                     (cudaStream_t)stream));
  return 0;
}

// ---------------------------------------------------------------- jax.random test hooks
__global__ void threefry_kernel(const uint32_t* key, const uint32_t* ctr, uint32_t* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t a = ctr[2 * i], b = ctr[2 * i + 1];
  threefry2x32(key[2 * i], key[2 * i + 1], a, b);
  out[2 * i] = a;
  out[2 * i + 1] = b;
}
__global__ void split_kernel(const uint32_t* key, int n, int num, uint32_t* out, int part) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * num) return;
  const int k = i / num, j = i % num;
  Key o;
  if (num == 2) {  // exercise the specialised paths the sampler uses
    Key a, b;
    split2(Key{key[2 * k], key[2 * k + 1]}, part != 0, a, b);
    o = j ? b : a;
  } else if (num == 3) {
    Key a, b, c;
    split3(Key{key[2 * k], key[2 * k + 1]}, part != 0, a, b, c);
    o = j == 0 ? a : (j == 1 ? b : c);
  } else {
    o = split_at(Key{key[2 * k], key[2 * k + 1]}, (uint32_t)num, (uint32_t)j, part != 0);
  }
  out[2 * i] = o.k0;
  out[2 * i + 1] = o.k1;
}
// what: 0 bits, 1 uniform, 2 normal
__global__ void draw_kernel(const uint32_t* key, int n, int what, float lo, float hi, void* out,
                            int part) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Key k{key[0], key[1]};
  if (what == 0) ((uint32_t*)out)[i] = random_bits_at(k, (uint32_t)n, (uint32_t)i, part != 0);
  else if (what == 1) ((float*)out)[i] = uniform_at(k, (uint32_t)n, (uint32_t)i, lo, hi, part != 0);
  else ((float*)out)[i] = normal_at(k, (uint32_t)n, (uint32_t)i, part != 0);
}

// ---------------------------------------------------------------- 3xTF32 GEMM (test hook)
extern "C" int b200mc_gemm3xtf32(const float* A_dev, const float* B_dev, int M, int Nc, int K,
                                 int splits, float* out_dev, void* stream) {
  if (M <= 0 || K <= 0 || Nc <= 0 || Nc > kGemmN) return fail("gemm3xtf32: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int RA = (M + kGemmM - 1) / kGemmM * kGemmM, Kp = (K + 31) / 32 * 32, KB = Kp / 32;
  if (splits < 1) splits = 1;
  if (splits > KB) splits = KB;
  unsigned char *a_hi = nullptr, *a_lo = nullptr, *b_hi = nullptr, *b_lo = nullptr;
  float* part = nullptr;
  int rc = build_image(A_dev, M, K, 0, RA, Kp, &a_hi, &a_lo, st);
  if (!rc) rc = build_image(B_dev, Nc, K, 0, kGemmN, Kp, &b_hi, &b_lo, st);
  if (!rc && cudaMalloc((void**)&part, sizeof(float) * (size_t)splits * kGemmN * RA) != cudaSuccess)
    rc = fail("gemm3xtf32: out of memory");
  if (!rc) {
    Gemm3xLaunch L = {};
    L.a_hi = a_hi; L.a_lo = a_lo; L.b_hi = b_hi; L.b_lo = b_lo;
    L.RA = RA; L.KB = KB; L.splits = splits; L.out = part;
    rc = launch_gemm3x(L, EPI_STORE, st);
  }
  if (!rc) {
    sum_splits_kernel<<<592, 256, 0, st>>>(part, splits, Nc, RA, M, out_dev);
    g_launches++;
    if (cudaGetLastError() != cudaSuccess) rc = fail("gemm3xtf32: reduce launch failed");
  }
  if (cudaStreamSynchronize(st) != cudaSuccess && !rc) rc = fail("gemm3xtf32: %s", cudaGetErrorString(cudaGetLastError()));
  dfree(a_hi); dfree(a_lo); dfree(b_hi); dfree(b_lo);  // build_image allocates from the pool
  cudaFree(part);
  return rc;
}

extern "C" int b200mc_threefry2x32(const uint32_t* key_dev, const uint32_t* ctr_dev,
                                   uint32_t* out_dev, int n, void* stream) {
  if (n <= 0) return 0;
  threefry_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(key_dev, ctr_dev, out_dev, n);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}
extern "C" int b200mc_random_split(const uint32_t* key_dev, int n, int num, uint32_t* out_dev,
                                   int partitionable, void* stream) {
  if (n <= 0 || num <= 0) return 0;
  split_kernel<<<(n * num + 255) / 256, 256, 0, (cudaStream_t)stream>>>(key_dev, n, num, out_dev,
                                                                       partitionable);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}
static int draw(const uint32_t* key_dev, int n, int what, float lo, float hi, void* out, int part,
                void* stream) {
  if (n <= 0) return 0;
  draw_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(key_dev, n, what, lo, hi, out, part);
  g_launches++;
  CK(cudaGetLastError());
  return 0;
}
extern "C" int b200mc_random_bits(const uint32_t* key_dev, int n, uint32_t* out_dev,
                                  int partitionable, void* stream) {
  return draw(key_dev, n, 0, 0.f, 0.f, out_dev, partitionable, stream);
}
extern "C" int b200mc_random_uniform(const uint32_t* key_dev, int n, float minval, float maxval,
                                     float* out_dev, int partitionable, void* stream) {
  return draw(key_dev, n, 1, minval, maxval, out_dev, partitionable, stream);
}
extern "C" int b200mc_random_normal(const uint32_t* key_dev, int n, float* out_dev,
                                    int partitionable, void* stream) {
  return draw(key_dev, n, 2, 0.f, 0.f, out_dev, partitionable, stream);
}

// ---------------------------------------------------------------- device plumbing
extern "C" int b200mc_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
extern "C" int b200mc_set_device(int device) {
  CK(cudaSetDevice(device));
  return 0;
}
extern "C" void* b200mc_malloc(uint64_t bytes) {
  void* p = nullptr;
  if (dmalloc(&p, bytes) != cudaSuccess) {
    fail("device allocation of %llu bytes failed: %s", (unsigned long long)bytes,
         cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}
extern "C" int b200mc_free(void* p) {
  dfree(p);
  return 0;
}
extern "C" int b200mc_memcpy_h2d(void* dst, const void* src, uint64_t bytes, void* stream) {
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_memcpy_d2h(void* dst, const void* src, uint64_t bytes, void* stream) {
  CK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_memset(void* dst, int value, uint64_t bytes, void* stream) {
  CK(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_stream_synchronize(void* stream) {
  CK(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}
extern "C" void* b200mc_host_alloc(uint64_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 16) != cudaSuccess) {
    fail("cudaMallocHost failed: %s", cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return p;
}
extern "C" int b200mc_host_free(void* p) {
  CK(cudaFreeHost(p));
  return 0;
}
extern "C" void* b200mc_event_create(void) {
  cudaEvent_t e = nullptr;
  if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
  return (void*)e;
}
extern "C" int b200mc_event_record(void* event, void* stream) {
  CK(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return 0;
}
extern "C" int b200mc_event_elapsed_ms(void* start, void* stop, float* ms) {
  CK(cudaEventSynchronize((cudaEvent_t)stop));
  CK(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return 0;
}
extern "C" int b200mc_event_destroy(void* event) {
  CK(cudaEventDestroy((cudaEvent_t)event));
  return 0;
}

// ---------------------------------------------------------------- on-device diagnostics
// effective_sample_size + split_gelman_rubin + gelman_rubin (numpyro/diagnostics.py:45-80,
// 158-203) of draws x[n_total][C][D] as the samplers collect them, draws first .. first+n-1.
// Outputs are DEVICE arrays of D doubles (any may be null).  Scratch comes from the caching
// allocator; synchronises the stream before it returns.
extern "C" int b200mc_diag_ess_rhat(const float* x_dev, int n_total, int first, int n, int chains, int D,
                                    double* n_eff_dev, double* r_hat_dev, double* split_r_hat_dev,
                                    void* stream) {
  if (!x_dev || n_total < 1 || chains < 1 || D < 1 || first < 0 || n < 2 || first + n > n_total)
    return fail("diag: bad argument");
  if (split_r_hat_dev && n < 4) return fail("diag: split R-hat needs at least 4 draws");
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, max_smem = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  CK(kl::diag_prepare(max_smem - 4096));  // the kernel also holds 2 KB of static shared memory
  int pb = 32;
  while (pb > 1 && (size_t)n * pb * sizeof(float) > (size_t)max_smem - 4096) pb >>= 1;
  if ((size_t)n * pb * sizeof(float) > (size_t)max_smem - 4096)
    return fail("diag: %d draws per chain do not fit the shared-memory series buffer (max %d)", n,
                (max_smem - 4096) / 4);
  double *mean = nullptr, *ssd = nullptr, *gamma = nullptr;
  const size_t stat_bytes = sizeof(double) * 2 * (size_t)chains * D;
  CK(dmalloc((void**)&mean, stat_bytes));
  CK(dmalloc((void**)&ssd, stat_bytes));
  int rc = 0;
  auto run = [&](int split, int max_lag, double* rhat_out, double* neff_out) -> int {
    DiagLaunch L;
    L.x = x_dev; L.n_total = n_total; L.C = chains; L.D = D; L.first = first; L.n = n;
    L.split = split; L.max_lag = max_lag; L.pb = pb; L.mean = mean; L.ssd = ssd; L.gamma_sum = gamma;
    const int nn = split ? n / 2 : n;
    CK(kl::diag_stats_launch(L, (size_t)nn * pb * sizeof(float), st));
    DiagFinish F;
    F.mean = mean; F.ssd = ssd; F.M = chains * (split ? 2 : 1); F.n = nn; F.D = D;
    F.gamma_sum = max_lag > 0 ? gamma : nullptr; F.max_lag = max_lag;
    F.r_hat = rhat_out; F.n_eff = neff_out;
    CK(kl::diag_finish_launch(F, st));
    g_launches += 2;
    return 0;
  };
  if (n_eff_dev || r_hat_dev) {
    if (n_eff_dev) {
      if (cudaSuccess != dmalloc((void**)&gamma, sizeof(double) * (size_t)n * D)) rc = fail("diag: out of memory");
      if (!rc && cudaMemsetAsync(gamma, 0, sizeof(double) * (size_t)n * D, st) != cudaSuccess) rc = fail("diag: memset failed");
    }
    if (!rc) rc = run(0, n_eff_dev ? n : 0, chains > 1 ? r_hat_dev : nullptr, n_eff_dev);
  }
  if (!rc && split_r_hat_dev) rc = run(1, 0, split_r_hat_dev, nullptr);
  if (cudaStreamSynchronize(st) != cudaSuccess && !rc) rc = fail("diag: %s", cudaGetErrorString(cudaGetLastError()));
  dfree(mean); dfree(ssd); dfree(gamma);
  return rc;
}

// ---------------------------------------------------------------- log_likelihood / Predictive
extern "C" int b200mc_obs_forward(const float* X_dev, int N, int K, const float* B_or_loc_dev, int S,
                                  const float* scale_dev, int scale_kind, const float* df_dev,
                                  const float* y_dev, const uint32_t* keys_dev, int likelihood,
                                  int partitionable, float* out_dev, void* stream) {
  if (!B_or_loc_dev || !out_dev || N < 1 || S < 1) return fail("obs_forward: bad argument");
  if (X_dev && (K < 1 || K > 4096)) return fail("obs_forward: K must be in [1, 4096]");
  if (likelihood < 0 || likelihood > 2) return fail("obs_forward: likelihood must be 0 (Bernoulli logits), 1 (Normal) or 2 (StudentT)");
  if (scale_kind < 0 || scale_kind > 3 || (scale_kind && !scale_dev)) return fail("obs_forward: bad scale");
  if (likelihood == 2 && !df_dev) return fail("obs_forward: StudentT needs df");
  if (!keys_dev && !y_dev) return fail("obs_forward: give observations (log-likelihood) or keys (sampling)");
  if (keys_dev && likelihood == 2)
    return fail("obs_forward: sampling a StudentT site is not implemented (jax.random.t draws through a "
                "rejection-sampled gamma); log_likelihood is");
  ObsLaunch L;
  L.X = X_dev; L.N = N; L.K = X_dev ? K : 0; L.S = S; L.B = B_or_loc_dev;
  L.scale = scale_dev; L.scale_kind = scale_kind; L.df = df_dev; L.y = keys_dev ? nullptr : y_dev;
  L.keys = keys_dev; L.likelihood = likelihood; L.partitionable = partitionable; L.out = out_dev;
  int dev = 0, max_smem = 0;
  CK(cudaGetDevice(&dev));
  CK(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if (X_dev && obs_smem(K) > (size_t)max_smem)
    return fail("obs_forward: K = %d columns do not fit the shared-memory tile", K);
  CK(kl::obs_forward_launch(L, (cudaStream_t)stream));
  g_launches++;
  return 0;
}

// ---------------------------------------------------------------- allocation-free entry point
// b200mc_horseshoe_potential_noalloc: the horseshoe potential + gradient kernel
// (chain_potential_kernel, models.cuh) with its per-point residual scratch -- [n][N] floats, which
// the handle-based path takes from the caching allocator at b200mc_create -- in a file-scope
// device array instead, for callers that must not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the 4 MiB
// until exit.  No counterpart in the reference.
// Calls no cudaMalloc / cudaFree / allocating library routine.  The array is shared by every
// call and keeps the last call's contents; nothing in it has to start at zero (the kernel writes
// every residual before it reads it), but two calls that overlap on the GPU (different streams)
// race on it.
namespace {
constexpr size_t kNoAllocScratchFloats = B200MC_NOALLOC_MAX_FLOATS;
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];
constexpr int kNoAllocMaxDevices = 64;
float* g_noalloc_ptr[kNoAllocMaxDevices] = {};  // device address of the array, per device
std::mutex g_noalloc_mu;
}  // namespace

extern "C" int b200mc_horseshoe_potential_noalloc(const float* X_dev, const float* Y_dev, int N,
                                                  int ncols, int likelihood, const float* z_dev, int n,
                                                  float* U_dev, float* grad_dev, void* stream) {
  if (!X_dev || !Y_dev || !z_dev || !U_dev || !grad_dev) return fail("null argument");
  if (N < 1 || ncols < 1 || n < 1) return fail("horseshoe_potential_noalloc: N, ncols and n must be >= 1");
  if (likelihood != 0 && likelihood != 1) return fail("horseshoe_potential_noalloc: likelihood must be 0 or 1");
  if ((size_t)n * (size_t)N > kNoAllocScratchFloats)
    return fail("horseshoe_potential_noalloc: n * N = %zu residuals exceed the static scratch of %zu floats "
                "(B200MC_NOALLOC_MAX_FLOATS); use the handle-based b200mc_potential_and_grad",
                (size_t)n * (size_t)N, kNoAllocScratchFloats);
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail("horseshoe_potential_noalloc: device index %d", dev);
  float* scratch = nullptr;
  {
    // In host code the array's name is only a stand-in: the device address comes from
    // cudaGetSymbolAddress, once per device, with that device current.
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
      float* p = nullptr;
      CK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
      g_noalloc_ptr[dev] = p;
    }
    scratch = g_noalloc_ptr[dev];
  }
  ModelData M;
  memset(&M, 0, sizeof(M));
  M.family = B200MC_HORSESHOE;
  M.N = N;
  M.ncols = ncols;
  M.likelihood = likelihood;
  M.D = 2 * ncols + 1 + (likelihood == 1 ? 1 : 0);
  M.a = X_dev;
  M.b = Y_dev;
  M.scratch = scratch;
  cudaStream_t st = (cudaStream_t)stream;
  // same chain grouping as a handle of this shape (b200mc_create): thread / warp / CTA per point
  const int mode = (M.D <= 32) ? 0 : (M.D <= 512 ? 1 : 2);
  CK(kl::chain_potential_launch(mode, M, z_dev, U_dev, grad_dev, n, st));
  g_launches++;
  return 0;
}
