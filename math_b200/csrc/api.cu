// api.cu -- host-level entry points: transparent device cache + evaluation.
//
// stan_cuda_glm_eval is what the STAN_CUDA branch of bernoulli_logit_glm_lpmf /
// poisson_log_glm_lpmf / normal_id_glm_lpdf calls between value_of_rec and ops_partials.build
// (SURVEY 3.1).  Per evaluation it moves K+2 doubles to the device and K+4 back; X and y are
// resident from the first call on ("caches the data matrix X and outcome vector y on the device
// the first time it sees them").  The cache is keyed by host identity plus a content
// fingerprint, or -- inside a map_rect shard scope -- by (call_id, job), which is the identity
// the reference's MPI backend already uses for immutable job data
// (stan/math/prim/mat/functor/mpi_parallel_call.hpp:40-89).
#include <math.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <map>
#include <memory>
#include <tuple>

#include "glm_device.cuh"

namespace stancuda {

struct Shard {
  stan_cuda_data* d = nullptr;
  int64_t row0 = 0;
};

struct Entry {
  std::vector<Shard> shards;   // GLM data, one block per device
  double* vec = nullptr;       // normal_lpdf: cached y
  int vec_device = -1;
  uint64_t fingerprint = 0;
  size_t bytes = 0;
  double lgamma_sum = 0.0;     // cached integer outcomes of a density: sum lgamma(v + 1)
  uint64_t last_use = 0;       // tick of the most recent hit (LRU eviction under a byte budget)
  ~Entry() {
    for (Shard& s : shards) free_data_block(s.d);
    if (vec) {
      cudaSetDevice(vec_device);
      cudaFree(vec);
    }
  }
};

// key: (kind, a, b, n, k, ldx, misc)   kind 0 = host pointers, 1 = shard scope, 2 = normal_lpdf y
typedef std::tuple<int, uintptr_t, uintptr_t, int64_t, int64_t, int64_t, int64_t> Key;

static std::mutex g_cache_mu;
static std::map<Key, std::shared_ptr<Entry>> g_cache;
static uint64_t g_cache_tick = 0;       // guarded by g_cache_mu
static int64_t g_cache_limit = 0;       // bytes; 0 = automatic (a fraction of the device memory)

// Byte budget of the cache.  Data sets are immutable for the life of a Stan program, so the common case never
// evicts; the budget keeps a long-lived process that keeps presenting NEW data (or a test-suite) from filling
// the 180 GB of a B200.  Automatic budget: 80 % of the smallest device's memory.
static int64_t cache_limit_locked() {
  if (g_cache_limit > 0) return g_cache_limit;
  static int64_t auto_limit = 0;
  if (auto_limit == 0) {
    const char* v = getenv("STAN_CUDA_CACHE_BYTES");
    if (v && atoll(v) > 0) {
      auto_limit = atoll(v);
    } else {
      int64_t smallest = 0;
      for (int dev : context().devices) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, dev) == cudaSuccess && (smallest == 0 || int64_t(prop.totalGlobalMem) < smallest))
          smallest = int64_t(prop.totalGlobalMem);
      }
      auto_limit = smallest > 0 ? smallest / 5 * 4 : (int64_t(1) << 62);
    }
  }
  return auto_limit;
}

// Drops least-recently-used entries until `incoming` more bytes fit the budget.  Entries are shared_ptr:
// an evaluation that still holds one keeps its device memory until it returns.  Caller holds g_cache_mu.
static void make_room_locked(int64_t incoming) {
  const int64_t limit = cache_limit_locked();
  int64_t total = incoming;
  for (auto& kv : g_cache) total += int64_t(kv.second->bytes);
  while (total > limit && !g_cache.empty()) {
    auto lru = g_cache.begin();
    for (auto it = g_cache.begin(); it != g_cache.end(); ++it)
      if (it->second->last_use < lru->second->last_use) lru = it;
    total -= int64_t(lru->second->bytes);
    g_cache.erase(lru);
    counters().cache_evictions.fetch_add(1, std::memory_order_relaxed);
  }
}

struct Scope {
  bool active = false;
  int call_id = 0, job = 0, device = 0;
  int ordinal = 0;  // GLM calls made inside this scope so far: the n-th call of a job is a different data set than the first
};
static thread_local Scope tl_scope;

// One multiply per sampled word (a byte-wise FNV has a dependent chain of 8 multiplies per word, which made the
// hashing -- not the 128 scattered loads -- the cost of a fingerprint: 2.0 us warm / 18 us cold per evaluation at
// C2 on the dev container's CPU, 0.4 / 7.5 us with this mix: the short chain also lets the loads overlap).
static inline uint64_t fnv(uint64_t h, uint64_t v) {
  h ^= v;
  h *= 0x9E3779B97F4A7C15ull;
  return h ^ (h >> 29);
}

// Content fingerprint from <= 64 sampled elements of each array: catches an address that is
// reused for different data (the reference once reverted a transparent GPU cache for
// exactly this reason, RELEASE-NOTES.txt:36).
// `samples` elements of each array.  Host-pointer keys use 64: an address can be reused for other data.  Inside a
// map_rect shard scope the identity is (call_id, job) and immutability is map_rect's own contract
// (stan/math/prim/mat/functor/map_rect.hpp:97-100), so 4 samples per array (first, last, two in between) remain as
// a tripwire only: the 128 scattered host reads per evaluation cost more than the launch of a 26 MB shard.
static std::atomic<int> g_cache_verify{-1};  // -1: read STAN_CUDA_CACHE_VERIFY on first use
static bool cache_verify() {
  int v = g_cache_verify.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("STAN_CUDA_CACHE_VERIFY");
    v = (e && *e && *e != '0') ? 1 : 0;
    g_cache_verify.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}

// Hash of EVERY element (verify mode): four independent chains so the multiplies overlap.
static uint64_t hash_words(const void* base, int64_t words8, uint64_t h) {
  const unsigned char* p = static_cast<const unsigned char*>(base);
  uint64_t h0 = h, h1 = h ^ 0x1111111111111111ull, h2 = h ^ 0x2222222222222222ull, h3 = h ^ 0x3333333333333333ull;
  int64_t i = 0;
  for (; i + 4 <= words8; i += 4) {
    uint64_t w[4];
    memcpy(w, p + i * 8, 32);
    h0 = fnv(h0, w[0]); h1 = fnv(h1, w[1]); h2 = fnv(h2, w[2]); h3 = fnv(h3, w[3]);
  }
  for (; i < words8; ++i) {
    uint64_t w;
    memcpy(&w, p + i * 8, 8);
    h0 = fnv(h0, w);
  }
  return fnv(fnv(fnv(h0, h1), h2), h3);
}

static uint64_t fingerprint_glm_full(const stan_cuda_glm_args* a, bool with_y) {
  uint64_t h = 1469598103934665603ull;
  const int64_t outer = a->x_row_major ? a->n : a->k, inner = a->x_row_major ? a->k : a->n;
  for (int64_t o = 0; o < outer; ++o) h = hash_words(a->x + o * a->ldx, inner, h);
  if (with_y) {
    if (a->y_is_double) {
      h = hash_words(a->y, a->n, h);
    } else {
      const int32_t* y = static_cast<const int32_t*>(a->y);
      h = hash_words(y, a->n / 2, h);
      if (a->n & 1) h = fnv(h, uint64_t(uint32_t(y[a->n - 1])));
    }
  }
  return h;
}

static uint64_t fingerprint_glm(const stan_cuda_glm_args* a, int64_t samples, bool with_y) {
  if (cache_verify()) return fingerprint_glm_full(a, with_y);
  uint64_t h = 1469598103934665603ull;
  const int64_t cnt = a->n * a->k;
  const int64_t step = cnt > samples ? (cnt - 1) / (samples - 1) : 1;
  for (int64_t q = 0; q < cnt; q += step) {
    int64_t i, j;
    if (a->x_row_major) { i = q / a->k; j = q % a->k; } else { j = q / a->n; i = q % a->n; }
    const double v = a->x_row_major ? a->x[i * a->ldx + j] : a->x[j * a->ldx + i];
    uint64_t bits;
    memcpy(&bits, &v, 8);
    h = fnv(h, bits);
  }
  const int64_t ystep = a->n > samples ? (a->n - 1) / (samples - 1) : 1;
  for (int64_t i = 0; with_y && i < a->n; i += ystep) {
    uint64_t bits = 0;
    if (a->y_is_double) memcpy(&bits, static_cast<const double*>(a->y) + i, 8);
    else bits = uint64_t(uint32_t(static_cast<const int32_t*>(a->y)[i]));
    h = fnv(h, bits);
  }
  return h;
}

static uint64_t fingerprint_vec(const double* v, int64_t n) {
  uint64_t h = 1469598103934665603ull;
  const int64_t step = n > 128 ? n / 128 : 1;
  for (int64_t i = 0; i < n; i += step) {
    uint64_t bits;
    memcpy(&bits, v + i, 8);
    h = fnv(h, bits);
  }
  return h;
}

// Which devices serve this call: the scope's device, all devices (row-sharded) or the first.
static std::vector<int> pick_devices(int64_t n, int64_t k) {
  Context& c = context();
  std::vector<int> out;
  if (tl_scope.active) {
    out.push_back(c.devices[size_t(tl_scope.device) % c.devices.size()]);
  } else if (c.devices.size() > 1 && n * (k > 0 ? k : 1) >= c.shard_min_elems && n >= int64_t(c.devices.size()) * 1024) {
    out = c.devices;
  } else {
    out.push_back(c.devices[0]);
  }
  return out;
}

static int get_glm_entry(const stan_cuda_glm_args* a, std::shared_ptr<Entry>* out) {
  set_categorical_classes_hint(a->family == STAN_CUDA_CATEGORICAL_LOGIT ? a->n_classes : 0);  // picks the data layout
  if (a->x_per_call) {
    // var x: nothing to recognise -- the calling thread keeps one set of device buffers per (shape, family) and
    // refills it on every call (upload + re-tiling; no cudaMalloc after the first call of that shape)
    static thread_local char tl_thread_tag;
    const int64_t bits = int64_t(a->family) * 8 + a->x_row_major * 2 + a->y_is_double + 64 * tuning().layout + 256 * tuning().cat_mma;
    const Key key(4, uintptr_t(&tl_thread_tag), uintptr_t(tl_scope.active ? tl_scope.device + 1 : 0), a->n, a->k, a->ldx, bits);
    std::shared_ptr<Entry> e;
    {
      std::lock_guard<std::mutex> lk(g_cache_mu);
      auto it = g_cache.find(key);
      if (it != g_cache.end()) {
        e = it->second;
        e->last_use = ++g_cache_tick;
      }
    }
    if (e) {
      for (Shard& sh : e->shards) {
        ThreadDev* td;
        SC_TRY(thread_dev(sh.d->device, &td));
        SC_TRY(refill_data_block(sh.d, a->x, a->ldx, a->x_row_major, a->y, sh.row0, td->stream));
      }
      *out = e;
      return STAN_CUDA_OK;
    }
    std::lock_guard<std::mutex> lk(g_cache_mu);
    make_room_locked(a->n * ((a->k + 15) / 16 * 16 + 1) * int64_t(sizeof(double)));
    e = std::make_shared<Entry>();
    const std::vector<int> devs = pick_devices(a->n, a->k);
    const int64_t G = int64_t(devs.size()), per = (a->n + G - 1) / G;
    for (int64_t g = 0; g < G; ++g) {
      const int64_t r0 = g * per, rows = (a->n - r0 < per) ? (a->n - r0) : per;
      if (rows <= 0) break;
      ThreadDev* td;
      SC_TRY(thread_dev(devs[g], &td));
      Shard sh;
      sh.row0 = r0;
      SC_TRY(make_data_block(devs[g], a->family, rows, a->k, a->x, a->ldx, a->x_row_major, a->y, a->y_is_double, r0,
                             td->stream, &sh.d));
      e->bytes += sh.d->bytes;
      e->shards.push_back(sh);
    }
    e->last_use = ++g_cache_tick;
    g_cache[key] = e;
    *out = e;
    return STAN_CUDA_OK;
  }
  Key key;
  const bool with_y = !a->y_per_call;  // a per-call y is no part of the identity: keyed on x alone
  const bool cat_wide = a->family == STAN_CUDA_CATEGORICAL_LOGIT && categorical_is_wide(a->k, a->n_classes);
  const int64_t kind_bits = int64_t(a->family) * 8 + (a->y_per_call ? 4 : 0) + a->x_row_major * 2 + a->y_is_double +
                            64 * tuning().layout + 256 * tuning().cat_mma + 512 * (cat_wide ? 1 : 0);
  if (tl_scope.active) {
    // (call_id, job) is the identity of a job's immutable data (mpi_parallel_call.hpp:40-89); a functor may make
    // SEVERAL GLM calls per job (one design matrix, two outcome vectors ...): the ordinal of the call inside the
    // job tells them apart (a functor runs the same sequence of calls on every evaluation of the job)
    const int ordinal = tl_scope.ordinal++;
    key = Key(1, uintptr_t(uint32_t(tl_scope.call_id)) | (uintptr_t(uint32_t(ordinal)) << 32), uintptr_t(tl_scope.job), a->n,
              a->k, kind_bits, tl_scope.device);
  } else {
    key = Key(0, uintptr_t(a->x), with_y ? uintptr_t(a->y) : uintptr_t(0), a->n, a->k, a->ldx, kind_bits);
  }
  const uint64_t fp = fingerprint_glm(a, tl_scope.active ? 16 : 64, with_y);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_cache.find(key);
  if (it != g_cache.end() && it->second->fingerprint == fp) {
    counters().cache_hits.fetch_add(1, std::memory_order_relaxed);
    it->second->last_use = ++g_cache_tick;
    *out = it->second;
    return STAN_CUDA_OK;
  }
  counters().cache_misses.fetch_add(1, std::memory_order_relaxed);
  NvtxRange fill_range("stan_cuda cache fill (upload + re-tiling)");
  if (it != g_cache.end()) g_cache.erase(it);  // stale: same identity, different content
  make_room_locked(a->n * ((a->k + 15) / 16 * 16 + 1) * int64_t(sizeof(double)));  // tiled x (padded) + y
  auto t0 = std::chrono::steady_clock::now();
  std::shared_ptr<Entry> e = std::make_shared<Entry>();
  e->fingerprint = fp;
  const std::vector<int> devs = pick_devices(a->n, a->k);
  const int64_t G = int64_t(devs.size());
  const int64_t per = (a->n + G - 1) / G;
  for (int64_t g = 0; g < G; ++g) {
    const int64_t r0 = g * per;
    const int64_t rows = (a->n - r0 < per) ? (a->n - r0) : per;
    if (rows <= 0) break;
    ThreadDev* td;
    SC_TRY(thread_dev(devs[g], &td));
    Shard s;
    s.row0 = r0;
    SC_TRY(make_data_block(devs[g], a->family, rows, a->k, a->x, a->ldx, a->x_row_major, a->y, a->y_is_double,
                           r0, td->stream, &s.d));
    e->bytes += s.d->bytes;
    e->shards.push_back(s);
  }
  auto t1 = std::chrono::steady_clock::now();
  counters().last_fill_ms.store(std::chrono::duration<double, std::milli>(t1 - t0).count());
  e->last_use = ++g_cache_tick;
  g_cache[key] = e;
  *out = e;
  return STAN_CUDA_OK;
}

// families with a fifth operand in the sigma slot (normal_id: sigma, neg_binomial_2_log: phi)
static inline bool has_scale(int family) {
  return family == STAN_CUDA_NORMAL_ID || family == STAN_CUDA_NEG_BINOMIAL_2_LOG;
}

static int check_glm_args(const stan_cuda_glm_args* a) {
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null args");
  if (a->family < 0 || a->family > STAN_CUDA_CATEGORICAL_LOGIT) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: unknown family");
  if (a->n < 0 || a->k < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: negative size");
  if (a->n > 0 && (!a->y || (!a->x && a->k > 0))) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null x or y");
  if (a->k > 0 && !a->beta) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null beta");
  if (a->family == STAN_CUDA_CATEGORICAL_LOGIT) {
    if (a->n_classes < 1 || !a->alpha || a->alpha_len != a->n_classes)
      return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: categorical_logit needs n_classes >= 1 and alpha_len == n_classes");
  } else if (a->family == STAN_CUDA_ORDERED_LOGISTIC) {
    if (a->cuts_len < 0 || (a->cuts_len > 0 && !a->cuts)) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: null cuts");
  } else if (!a->alpha || (a->alpha_len != 1 && a->alpha_len != a->n)) {
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: alpha_len must be 1 or n");
  }
  if (has_scale(a->family)) {
    if (!a->sigma || (a->sigma_len != 1 && a->sigma_len != a->n))
      return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: sigma_len must be 1 or n");
  }
  if (a->family == STAN_CUDA_NORMAL_ID) {
    if (!a->y_is_double) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: normal_id needs double y");
  } else if (a->y_is_double) {
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: bernoulli/poisson/neg_binomial need int32 y");
  }
  if (a->k > 0 && a->ldx < (a->x_row_major ? a->k : a->n))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: ldx too small");
  return STAN_CUDA_OK;
}

// How long the host spins on the completion flag before it hands the wait to the driver.  Round 1 gave up after 250 us,
// so every evaluation longer than that (C3: 1.4 ms) paid a blocking cudaStreamSynchronize and its wake-up latency on top
// of the kernel.  A sampler's thread has nothing else to do until the gradient is there: spin for up to 50 ms
// (STAN_CUDA_SPIN_US), then fall back (a faulted kernel never raises the flag; the driver call then reports the error).
static int64_t spin_limit_us() {
  static const int64_t us = [] {
    const char* v = getenv("STAN_CUDA_SPIN_US");
    const long long x = v ? atoll(v) : 50000;
    return int64_t(x < 0 ? 0 : x);
  }();
  return us;
}

static bool low_latency_enabled() {
  static const bool on = [] {
    const char* v = getenv("STAN_CUDA_LOW_LATENCY");
    return !(v && *v == '0');
  }();
  return on;
}

struct ShardRun {
  ThreadDev* td = nullptr;
  double *d_params = nullptr, *d_result = nullptr, *d_alpha = nullptr, *d_sigma = nullptr, *d_rowa = nullptr,
         *d_rows = nullptr;
  double* d_dx = nullptr;     // var x: d * beta^T, column-major rows x k
  void* d_y = nullptr;        // per-call outcomes (y_per_call)
  double y_lgamma_sum = 0.0;
};

int glm_eval(const stan_cuda_glm_args* a) {
  NvtxRange range("stan_cuda_glm_eval");
  SC_TRY(check_glm_args(a));
  if (a->n == 0) {  // size_zero(y): the reference returns 0 before touching anything (:72-74)
    if (a->lp) *a->lp = 0.0;
    if (a->dbeta) for (int64_t j = 0; j < a->k; ++j) a->dbeta[j] = 0.0;
    if (a->family == STAN_CUDA_ORDERED_LOGISTIC && a->dcuts) for (int64_t j = 0; j < a->cuts_len; ++j) a->dcuts[j] = 0.0;
    if (a->family == STAN_CUDA_CATEGORICAL_LOGIT) {
      if (a->dbeta) for (int64_t j = 0; j < a->k * a->n_classes; ++j) a->dbeta[j] = 0.0;
      if (a->dalpha) for (int64_t j = 0; j < a->n_classes; ++j) a->dalpha[j] = 0.0;
    }
    return STAN_CUDA_OK;
  }
  const bool categorical = a->family == STAN_CUDA_CATEGORICAL_LOGIT;
  const size_t ncls = categorical ? size_t(a->n_classes) : 0;
  const bool ordered = a->family == STAN_CUDA_ORDERED_LOGISTIC;
  const size_t ncuts = ordered ? size_t(a->cuts_len) : 0;
  SC_TRY(ensure_init());
  std::shared_ptr<Entry> e;
  SC_TRY(get_glm_entry(a, &e));
  const size_t S = e->shards.size();
  std::vector<ShardRun> runs(S);
  // Per-call outcomes: y goes to the device now (padded to whole tiles like the cached copy) into a buffer of
  // its own; integer outcomes also get their range and sum lgamma(y + 1) recomputed.
  int64_t call_lo = INT64_MAX, call_hi = INT64_MIN;
  const bool y_call = a->y_per_call && !a->x_per_call;  // (a per-call x refills y along with it)
  if (y_call) {
    const size_t ysz = a->y_is_double ? sizeof(double) : sizeof(int32_t);
    for (size_t s = 0; s < S; ++s) {
      stan_cuda_data* d = e->shards[s].d;
      ShardRun& run = runs[s];
      SC_TRY(thread_dev(d->device, &run.td));
      SC_CUDA(cudaSetDevice(d->device));
      const size_t padded = size_t(d->t.tiles) * d->t.R * ysz, used = size_t(d->n) * ysz;
      SC_TRY(run.td->reserve_y((padded + 15) / 16 * 2));
      run.d_y = run.td->d_ybuf;
      SC_CUDA(cudaMemcpyAsync(run.d_y, static_cast<const char*>(a->y) + size_t(e->shards[s].row0) * ysz, used,
                              cudaMemcpyHostToDevice, run.td->stream));
      if (padded > used)
        SC_CUDA(cudaMemsetAsync(static_cast<char*>(run.d_y) + used, 0, padded - used, run.td->stream));
      if (!a->y_is_double) {
        const int want_lg = a->family == STAN_CUDA_POISSON_LOG || a->family == STAN_CUDA_NEG_BINOMIAL_2_LOG;
        int64_t lo = 0, hi = 0;
        SC_TRY(int_vector_stats(static_cast<const int*>(run.d_y), d->n, want_lg, run.td->stream, &run.y_lgamma_sum, &lo, &hi));
        call_lo = std::min(call_lo, lo);
        call_hi = std::max(call_hi, hi);
      }
    }
  }
  // eager domain check of the outcomes (bernoulli_logit_glm_lpmf.hpp:70, poisson_log_glm_lpmf.hpp:70)
  // from the range computed at fill: no O(N) host loop per call
  if (a->family != STAN_CUDA_NORMAL_ID) {
    int64_t lo = e->shards[0].d->y_min, hi = e->shards[0].d->y_max;
    for (size_t s = 1; s < S; ++s) {
      lo = std::min(lo, e->shards[s].d->y_min);
      hi = std::max(hi, e->shards[s].d->y_max);
    }
    if (y_call) {
      lo = call_lo;
      hi = call_hi;
    }
    const auto bad_outcome = [&]() {
      if (tl_scope.active && tl_scope.ordinal > 0) --tl_scope.ordinal;  // the wrapper's retry is the same call of the job
      return STAN_CUDA_BAD_OUTCOME;
    };
    if (ordered) {  // ordered_logistic_glm_lpmf.hpp:78 check_bounded(y, 1, N_classes)
      if (lo < 1 || hi > int64_t(ncuts) + 1) return bad_outcome();
    } else if (categorical) {  // categorical_logit_glm_lpmf.hpp:63-64
      if (lo < 1 || hi > int64_t(ncls)) return bad_outcome();
    } else if (lo < 0 || (a->family == STAN_CUDA_BERNOULLI_LOGIT && hi > 1)) {
      return bad_outcome();
    }
  }
  const bool alpha_vec = !ordered && !categorical && a->alpha_len != 1;
  const bool sigma_vec = has_scale(a->family) && a->sigma_len != 1;
  bool want_rowa = ((a->flags & STAN_CUDA_NEED_DALPHA) && alpha_vec && a->dalpha) ||
                   ((a->flags & STAN_CUDA_NEED_DY) && a->dy && a->family == STAN_CUDA_NORMAL_ID);
  const bool want_rows = (a->flags & STAN_CUDA_NEED_DSIGMA) && sigma_vec && a->dsigma;
  // var x: d/dx = d * beta^T needs d per row on the device (the ROWIO instantiation), then one outer-product kernel
  const bool want_dx = (a->flags & STAN_CUDA_NEED_DX) && a->dx && a->k > 0;
  if (want_dx && (categorical || ordered))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: d/dx is not available for ordered_logistic / categorical_logit");
  if (want_dx && a->lddx < a->n) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval: lddx < n");
  const bool want_rowa_user = want_rowa;
  want_rowa = want_rowa || want_dx;
  const int world = comm_world_size();
  if (world > 1 && S > 1) return fail(STAN_CUDA_ERR_STATE, "multi-process communicator with several devices per process");

  // Sharded evaluations of one process are serialised: the devices then see the collectives of
  // different host threads in the same order (required by the in-kernel exchange and by NCCL).
  static std::mutex sharded_mu;
  std::unique_lock<std::mutex> sharded_lock(sharded_mu, std::defer_lock);
  PeerXchg* xchg = nullptr;   // [S] views (in-process) or this rank's view (multi-process)
  unsigned long long xchg_seq = 0;
  if (S > 1) {
    sharded_lock.lock();
    std::vector<int> devs(S);
    for (size_t s = 0; s < S; ++s) devs[s] = e->shards[s].d->device;
    xchg = comm_local_xchg(devs);
    if (xchg) xchg_seq = ++xchg[0].seq;
  } else if (world > 1) {
    xchg = comm_rank_xchg();
    if (xchg) xchg_seq = ++xchg->seq;
  }

  const int kpad = e->shards[0].d->t.kpad;
  const size_t np = categorical ? ncls * (size_t(kpad) + 1) : size_t(kpad) + 2 + ncuts;
  const size_t nr = categorical ? 4 + ncls * (size_t(kpad) + 1) : size_t(kpad) + 4 + ncuts;
  if (xchg && nr > size_t(kXchgCols)) xchg = nullptr;  // result block wider than an exchange slot: NCCL behind the kernel
  // Low-latency path (no per-row vectors in or out, and no separate collective behind the kernel):
  // the parameter block travels in the kernel arguments, the kernel writes the result block and a
  // completion flag straight into mapped pinned host memory, the host spins on the flag.  Saves the
  // H2D copy, the D2H copy and the stream synchronisation of the general path (~15 us per call).
  const bool per_row = alpha_vec || sigma_vec || want_rowa || want_rows;
  const bool low_latency = !per_row && !categorical && low_latency_enabled() && ((S == 1 && world == 1) || xchg != nullptr) &&
                           np <= size_t(kInlineParams);
  for (size_t s = 0; s < S; ++s) {
    stan_cuda_data* d = e->shards[s].d;
    const int64_t rows = d->n, r0 = e->shards[s].row0;
    ShardRun& run = runs[s];
    SC_TRY(thread_dev(d->device, &run.td));
    ThreadDev* td = run.td;
    SC_CUDA(cudaSetDevice(d->device));
    SC_TRY(td->reserve_host(np + nr + 2));
    size_t need = np + nr + 8;
    const size_t rows_al = (size_t(rows) + 1) & ~size_t(1);
    if (alpha_vec) need += rows_al;
    if (sigma_vec) need += rows_al;
    if (want_rowa) need += rows_al;
    if (want_rows) need += rows_al;
    if (want_dx) need += size_t(rows) * size_t(a->k);
    SC_TRY(td->reserve_dev(need));
    double* p = td->d_buf;
    run.d_params = p; p += (np + 1) & ~size_t(1);
    run.d_result = p; p += (nr + 1) & ~size_t(1);
    if (alpha_vec) { run.d_alpha = p; p += rows_al; }
    if (sigma_vec) { run.d_sigma = p; p += rows_al; }
    if (want_rowa) { run.d_rowa = p; p += rows_al; }
    if (want_rows) { run.d_rows = p; p += rows_al; }
    if (want_dx) { run.d_dx = p; p += size_t(rows) * size_t(a->k); }
    double* hp = td->h_pinned;
    if (categorical) {  // weights class-major, every class padded to kpad; intercepts behind
      for (size_t c = 0; c < ncls; ++c) {
        for (int j = 0; j < kpad; ++j) hp[c * kpad + j] = j < a->k ? a->beta[c * a->k + j] : 0.0;
        hp[ncls * kpad + c] = a->alpha[c];
      }
    } else {
      for (int j = 0; j < kpad; ++j) hp[j] = j < a->k ? a->beta[j] : 0.0;
      hp[kpad] = (alpha_vec || ordered) ? 0.0 : a->alpha[0];
      hp[kpad + 1] = (has_scale(a->family) && !sigma_vec) ? a->sigma[0] : 1.0;
      for (size_t j = 0; j < ncuts; ++j) hp[kpad + 2 + j] = a->cuts[j];
    }
    if (!low_latency)
      SC_CUDA(cudaMemcpyAsync(run.d_params, hp, np * sizeof(double), cudaMemcpyHostToDevice, td->stream));
    if (alpha_vec)
      SC_CUDA(cudaMemcpyAsync(run.d_alpha, a->alpha + r0, size_t(rows) * sizeof(double), cudaMemcpyHostToDevice, td->stream));
    if (sigma_vec)
      SC_CUDA(cudaMemcpyAsync(run.d_sigma, a->sigma + r0, size_t(rows) * sizeof(double), cudaMemcpyHostToDevice, td->stream));
    LaunchIO io{a->flags, run.d_params, run.d_alpha, run.d_sigma, run.d_result, run.d_rowa, run.d_rows};
    io.ncuts = int(ncuts);
    io.ncls = int(ncls);
    io.shared_gpu = tl_scope.active;
    io.d_y = run.d_y;
    io.y_lgamma_sum = run.y_lgamma_sum;
    if (xchg) {
      io.xchg = S > 1 ? &xchg[s] : xchg;
      io.xchg_seq = xchg_seq;
    }
    if (low_latency) {
      io.h_params_inline = hp;
      io.d_result = hp + np;  // pinned + mapped: same address on the device (unified addressing)
      io.host_flag = reinterpret_cast<unsigned long long*>(hp + np + nr);
      io.host_seq = ++td->host_seq;
    }
    SC_TRY(launch_glm(d, io, &td->ws, td->stream));
    if (want_dx) SC_TRY(launch_dx_outer(run.d_rowa, run.d_params, rows, a->k, run.d_dx, td->stream));
  }
  // combine the additive result blocks of the row shards (unless the kernel already did)
  if (xchg) {
    counters().fused_allreduces.fetch_add(1, std::memory_order_relaxed);
  } else if (S > 1) {
    std::vector<int> devs(S);
    std::vector<double*> bufs(S);
    std::vector<cudaStream_t> streams(S);
    for (size_t s = 0; s < S; ++s) {
      devs[s] = e->shards[s].d->device;
      bufs[s] = runs[s].d_result;
      streams[s] = runs[s].td->stream;
    }
    SC_TRY(comm_allreduce_local(devs, bufs, int64_t(nr), streams));
  } else if (world > 1) {
    SC_CUDA(cudaSetDevice(e->shards[0].d->device));
    SC_TRY(comm_allreduce_rank(runs[0].d_result, int64_t(nr), runs[0].td->stream));
  }
  // results back
  NvtxRange wait_range("stan_cuda wait for the result block");
  if (low_latency) {
    for (size_t s = 0; s < S; ++s) {
      ThreadDev* td = runs[s].td;
      volatile unsigned long long* flag = reinterpret_cast<volatile unsigned long long*>(td->h_pinned + np + nr);
      const auto t0 = std::chrono::steady_clock::now();
      int spins = 0;
      while (*flag != td->host_seq) {
        if ((++spins & 1023) == 0 &&
            std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(spin_limit_us())) {
          // long kernel (or a fault): let the driver block / report
          SC_CUDA(cudaSetDevice(e->shards[s].d->device));
          SC_CUDA(cudaStreamSynchronize(td->stream));
          break;
        }
#if defined(__x86_64__)
        __builtin_ia32_pause();
#endif
      }
      std::atomic_thread_fence(std::memory_order_acquire);
    }
  }
  for (size_t s = 0; s < S && !low_latency; ++s) {
    stan_cuda_data* d = e->shards[s].d;
    const int64_t rows = d->n, r0 = e->shards[s].row0;
    ThreadDev* td = runs[s].td;
    SC_CUDA(cudaSetDevice(d->device));
    if (s == 0)
      SC_CUDA(cudaMemcpyAsync(td->h_pinned + np, runs[s].d_result, nr * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
    if (want_dx)
      SC_CUDA(cudaMemcpy2DAsync(a->dx + r0, size_t(a->lddx) * sizeof(double), runs[s].d_dx, size_t(rows) * sizeof(double),
                                size_t(rows) * sizeof(double), size_t(a->k), cudaMemcpyDeviceToHost, td->stream));
    if (want_rowa_user) {
      double* dst = (alpha_vec && (a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha) ? a->dalpha : a->dy;
      SC_CUDA(cudaMemcpyAsync(dst + r0, runs[s].d_rowa, size_t(rows) * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
    }
    if (want_rows)
      SC_CUDA(cudaMemcpyAsync(a->dsigma + r0, runs[s].d_rows, size_t(rows) * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
  }
  for (size_t s = 0; s < S && !low_latency; ++s) {
    SC_CUDA(cudaSetDevice(e->shards[s].d->device));
    SC_CUDA(cudaStreamSynchronize(runs[s].td->stream));
  }
  const double* res = runs[0].td->h_pinned + np;
  if (xchg) {
    // the in-kernel exchange gives up after 30 s without a peer's block and marks every result with this NaN payload
    uint64_t bits;
    memcpy(&bits, &res[0], 8);
    if (bits == 0x7ff8dead00000000ull)
      return fail(STAN_CUDA_ERR_NCCL, "fused all-reduce: a peer device did not publish its result block within 30 s "
                                      "(a rank died, or the ranks did not launch the same sequence of sharded evaluations)");
  }
  if (a->lp) *a->lp = res[0];
  if (categorical) {
    if ((a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha) memcpy(a->dalpha, res + 4, ncls * sizeof(double));
    if ((a->flags & STAN_CUDA_NEED_DBETA) && a->dbeta)
      for (size_t c = 0; c < ncls; ++c) memcpy(a->dbeta + c * a->k, res + 4 + ncls + c * kpad, size_t(a->k) * sizeof(double));
  } else if ((a->flags & STAN_CUDA_NEED_DBETA) && a->dbeta) memcpy(a->dbeta, res + 4, size_t(a->k) * sizeof(double));
  if ((a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha && !alpha_vec && !ordered && !categorical) a->dalpha[0] = res[1];
  if (ordered && (a->flags & STAN_CUDA_NEED_DCUTS) && a->dcuts) memcpy(a->dcuts, res + 4 + kpad, ncuts * sizeof(double));
  if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma && has_scale(a->family) && !sigma_vec)
    a->dsigma[0] = res[2];
  if ((a->flags & STAN_CUDA_NEED_DY) && a->dy && a->family == STAN_CUDA_NORMAL_ID) {
    // dy = -mu_derivative (normal_id_glm_lpdf.hpp:118-120); the kernel wrote mu_derivative
    const bool via_dalpha = alpha_vec && (a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha;
    const double* src = via_dalpha ? a->dalpha : a->dy;
    for (int64_t i = 0; i < a->n; ++i) a->dy[i] = -src[i];
  }
  if (a->aux) {
    a->aux[0] = res[1];
    a->aux[1] = res[3];
    a->aux[2] = 0.0;
    a->aux[3] = y_call ? runs[0].y_lgamma_sum : e->shards[0].d->lgamma_sum;
  }
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  return isfinite(res[3]) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

}  // namespace stancuda

// =============================================================================================
// Batched evaluation: `count` independent GLMs of one family and one K (map_rect shards), item j on device
// j % G (the dispatch rule of stan/math/cuda/map_rect_devices.hpp) and keyed (call_id, j) like a shard scope,
// ONE kernel launch per device (glm_warp_batch_kernel, grid.y = items of that device).  SURVEY 8(e): "shards
// resident on the same device should be evaluated by one batched launch when the functor is recognisably a
// single GLM call".  A functor is opaque to map_rect, so the unmodified C++ path cannot know that; this entry
// point is for callers that do (tests, bench tools, a code generator), and for measuring both ways.
// =============================================================================================
namespace stancuda {

int glm_eval_batch(const stan_cuda_glm_args* items, int count, int call_id, int* item_rc) {
  NvtxRange range("stan_cuda_glm_eval_batch");
  if (!items || count < 1) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: no items");
  const int family = items[0].family;
  if (family < STAN_CUDA_BERNOULLI_LOGIT || family > STAN_CUDA_NORMAL_ID)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: bernoulli_logit, poisson_log and normal_id only");
  for (int i = 0; i < count; ++i) {
    const stan_cuda_glm_args* a = &items[i];
    SC_TRY(check_glm_args(a));
    if (a->family != family || a->k != items[0].k || a->k < 1)
      return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: items must share family and k >= 1");
    if (a->n < 1 || a->alpha_len != 1 || (family == STAN_CUDA_NORMAL_ID && a->sigma_len != 1) || (a->flags & STAN_CUDA_NEED_DY))
      return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: non-empty items with scalar alpha / sigma and data y only");
    if (item_rc) item_rc[i] = STAN_CUDA_OK;
  }
  SC_TRY(ensure_init());
  Context& c = context();
  const int G = int(c.devices.size());
  std::vector<std::shared_ptr<Entry>> ents;
  ents.resize(size_t(count));
  {
    const Scope saved = tl_scope;
    int rc = STAN_CUDA_OK;
    for (int i = 0; i < count && rc >= 0; ++i) {
      tl_scope.active = true;
      tl_scope.call_id = call_id;
      tl_scope.job = i;
      tl_scope.device = i % G;
      tl_scope.ordinal = 0;
      rc = get_glm_entry(&items[i], &ents[size_t(i)]);
    }
    tl_scope = saved;
    SC_TRY(rc);
  }
  // eager domain check of the outcomes from the ranges computed at fill (bernoulli_logit_glm_lpmf.hpp:70,
  // poisson_log_glm_lpmf.hpp:70): nothing is evaluated if any item fails
  bool bad = false;
  if (family != STAN_CUDA_NORMAL_ID) {
    for (int i = 0; i < count; ++i) {
      const stan_cuda_data* d = ents[size_t(i)]->shards[0].d;
      if (d->y_min < 0 || (family == STAN_CUDA_BERNOULLI_LOGIT && d->y_max > 1)) {
        bad = true;
        if (item_rc) item_rc[i] = STAN_CUDA_BAD_OUTCOME;
      }
    }
  }
  if (bad) return STAN_CUDA_BAD_OUTCOME;

  if (!ents[0]->shards[0].d->colblocks.empty())
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: k <= 512 only (wider matrices run as several launches per item: "
                                   "use stan_cuda_glm_eval)");
  const int kpad = ents[0]->shards[0].d->t.kpad;
  const size_t np = size_t(kpad) + 2, nr = size_t(kpad) + 4;
  const size_t npa = (np + 1) & ~size_t(1), nra = (nr + 1) & ~size_t(1);
  const size_t bs = (sizeof(GlmKernelParams) + 7) / 8;  // doubles per parameter block
  struct DevRun {
    ThreadDev* td = nullptr;
    std::vector<int> idx;
    double* h_res = nullptr;
  };
  std::vector<DevRun> runs;
  runs.resize(size_t(G));
  for (int i = 0; i < count; ++i) runs[size_t(i % G)].idx.push_back(i);
  for (int g = 0; g < G; ++g) {
    DevRun& run = runs[size_t(g)];
    const int cnt = int(run.idx.size());
    if (cnt == 0) continue;
    if (cnt > 65535) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_glm_eval_batch: more than 65535 items per device (grid.y)");
    const int dev = c.devices[size_t(g)];
    SC_TRY(thread_dev(dev, &run.td));
    ThreadDev* td = run.td;
    SC_CUDA(cudaSetDevice(dev));
    std::vector<const stan_cuda_data*> ds;
    ds.resize(size_t(cnt));
    int64_t min_tiles = INT64_MAX;
    for (int j = 0; j < cnt; ++j) {
      ds[size_t(j)] = ents[size_t(run.idx[size_t(j)])]->shards[0].d;
      min_tiles = std::min(min_tiles, ds[size_t(j)]->t.tiles);
    }
    int gx = 1, gs = 1;
    size_t scr = 0;
    SC_TRY(glm_warp_batch_plan(ds[0], cnt, min_tiles, &gx, &gs, &scr));
    const size_t scr_al = (scr + 1) & ~size_t(1);
    const size_t tick_d = 32;  // 64 ticket words per problem
    SC_TRY(td->reserve_dev(size_t(cnt) * (tick_d + scr_al + npa + nra + bs) + 8));
    SC_TRY(td->reserve_host(size_t(cnt) * (npa + nra + bs) + 8));
    double* dp = td->d_buf;
    double* d_tick = dp; dp += size_t(cnt) * tick_d;
    double* d_scr = dp; dp += size_t(cnt) * scr_al;
    double* d_par = dp; dp += size_t(cnt) * npa;
    double* d_res = dp; dp += size_t(cnt) * nra;
    GlmKernelParams* d_batch = reinterpret_cast<GlmKernelParams*>(dp);
    double* h_par = td->h_pinned;
    run.h_res = h_par + size_t(cnt) * npa;
    GlmKernelParams* h_batch = reinterpret_cast<GlmKernelParams*>(run.h_res + size_t(cnt) * nra);
    SC_CUDA(cudaMemsetAsync(d_tick, 0, size_t(cnt) * tick_d * sizeof(double), td->stream));
    std::vector<Workspace> wss;
    wss.resize(size_t(cnt));  // views into d_buf: not owned, never freed through workspace_free
    std::vector<LaunchIO> ios;
    ios.reserve(size_t(cnt));
    for (int j = 0; j < cnt; ++j) {
      const stan_cuda_glm_args* a = &items[run.idx[size_t(j)]];
      double* hp = h_par + size_t(j) * npa;
      for (int q = 0; q < kpad; ++q) hp[q] = q < a->k ? a->beta[q] : 0.0;
      hp[kpad] = a->alpha[0];
      hp[kpad + 1] = family == STAN_CUDA_NORMAL_ID ? a->sigma[0] : 1.0;
      Workspace& w = wss[size_t(j)];
      w.device = dev;
      w.scratch = d_scr + size_t(j) * scr_al;
      w.ticket = reinterpret_cast<unsigned int*>(d_tick + size_t(j) * tick_d);
      w.scratch_doubles = scr_al;
      ios.push_back(LaunchIO{a->flags, d_par + size_t(j) * npa, nullptr, nullptr, d_res + size_t(j) * nra, nullptr, nullptr});
    }
    SC_CUDA(cudaMemcpyAsync(d_par, h_par, size_t(cnt) * npa * sizeof(double), cudaMemcpyHostToDevice, td->stream));
    SC_TRY(launch_glm_warp_batch(ds.data(), ios.data(), wss.data(), cnt, gx, gs, h_batch, d_batch, td->stream));
    SC_CUDA(cudaMemcpyAsync(run.h_res, d_res, size_t(cnt) * nra * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
  }
  int worst = STAN_CUDA_OK;
  for (int g = 0; g < G; ++g) {
    DevRun& run = runs[size_t(g)];
    if (run.idx.empty()) continue;
    SC_CUDA(cudaSetDevice(c.devices[size_t(g)]));
    SC_CUDA(cudaStreamSynchronize(run.td->stream));
    for (size_t j = 0; j < run.idx.size(); ++j) {
      const stan_cuda_glm_args* a = &items[run.idx[j]];
      const double* res = run.h_res + j * nra;
      if (a->lp) *a->lp = res[0];
      if ((a->flags & STAN_CUDA_NEED_DBETA) && a->dbeta) memcpy(a->dbeta, res + 4, size_t(a->k) * sizeof(double));
      if ((a->flags & STAN_CUDA_NEED_DALPHA) && a->dalpha) a->dalpha[0] = res[1];
      if ((a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma && family == STAN_CUDA_NORMAL_ID) a->dsigma[0] = res[2];
      if (a->aux) {
        a->aux[0] = res[1];
        a->aux[1] = res[3];
        a->aux[2] = 0.0;
        a->aux[3] = ents[size_t(run.idx[j])]->shards[0].d->lgamma_sum;
      }
      if (!isfinite(res[3])) {
        worst = STAN_CUDA_NONFINITE;
        if (item_rc) item_rc[run.idx[j]] = STAN_CUDA_NONFINITE;
      }
    }
  }
  counters().evals.fetch_add(uint64_t(count), std::memory_order_relaxed);
  return worst;
}

}  // namespace stancuda

// =============================================================================================
// normal_lpdf
// =============================================================================================
namespace stancuda {

static int get_vec_entry(const double* v, int64_t n, int device, cudaStream_t stream, std::shared_ptr<Entry>* out) {
  Key key(2, uintptr_t(v), 0, n, 0, 0, device);
  const uint64_t fp = fingerprint_vec(v, n);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_cache.find(key);
  if (it != g_cache.end() && it->second->fingerprint == fp) {
    counters().cache_hits.fetch_add(1, std::memory_order_relaxed);
    it->second->last_use = ++g_cache_tick;
    *out = it->second;
    return STAN_CUDA_OK;
  }
  counters().cache_misses.fetch_add(1, std::memory_order_relaxed);
  if (it != g_cache.end()) g_cache.erase(it);
  make_room_locked(n * int64_t(sizeof(double)));
  std::shared_ptr<Entry> e = std::make_shared<Entry>();
  e->fingerprint = fp;
  e->last_use = ++g_cache_tick;
  e->vec_device = device;
  SC_CUDA(cudaSetDevice(device));
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->vec), size_t(n) * sizeof(double)));
  e->bytes = size_t(n) * sizeof(double);
  SC_CUDA(cudaMemcpyAsync(e->vec, v, size_t(n) * sizeof(double), cudaMemcpyHostToDevice, stream));
  SC_CUDA(cudaStreamSynchronize(stream));
  counters().bytes_uploaded.fetch_add(uint64_t(n) * sizeof(double));
  g_cache[key] = e;
  *out = e;
  return STAN_CUDA_OK;
}

// Vectors of the elementwise densities between host and device.  A copy from or to PAGEABLE host memory crawls on
// this platform once it is large (80 MB: 6.8 GB/s direct, 18 GB/s staged; 8 MB: no difference), so large pageable vectors
// size go through the page-locked staging buffers and copy threads of gemm.cu; page-locked ones (the C++ wrappers
// gather operands into stan_cuda_host_staging) are copied directly.
int dense_copy_to_device(double* dev, int64_t ld_dev, const double* host, int64_t ld_host, int64_t rows, int64_t cols, cudaStream_t stream);
int dense_copy_to_host(double* host, int64_t ld_host, const double* dev, int64_t ld_dev, int64_t rows, int64_t cols, cudaStream_t stream);

static bool is_pageable(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return at.type == cudaMemoryTypeUnregistered;
}
constexpr int64_t kStagedVecMin = int64_t(1) << 21;  // doubles (16 MiB): below that a direct copy is as fast (measured at 8 MB)

static int vec_to_device(double* dev, const double* host, int64_t n, cudaStream_t stream) {
  if (n >= kStagedVecMin && is_pageable(host)) return dense_copy_to_device(dev, n, host, n, n, 1, stream);
  SC_CUDA(cudaMemcpyAsync(dev, host, size_t(n) * sizeof(double), cudaMemcpyHostToDevice, stream));
  return STAN_CUDA_OK;
}
// asynchronous for small or page-locked destinations; returns with the data in place otherwise (the callers
// synchronise the stream right after their last copy either way)
static int vec_to_host(double* host, const double* dev, int64_t n, cudaStream_t stream) {
  if (n >= kStagedVecMin && is_pageable(host)) return dense_copy_to_host(host, n, dev, n, n, 1, stream);
  SC_CUDA(cudaMemcpyAsync(host, dev, size_t(n) * sizeof(double), cudaMemcpyDeviceToHost, stream));
  return STAN_CUDA_OK;
}

int locscale_lpdf_eval(int family, const stan_cuda_normal_args* a);
int normal_lpdf_eval(const stan_cuda_normal_args* a) { return locscale_lpdf_eval(STAN_CUDA_LOCSCALE_NORMAL, a); }

// normal_lpdf / cauchy_lpdf / lognormal_lpdf: three real operands, each a scalar or a vector of n
int locscale_lpdf_eval(int family, const stan_cuda_normal_args* a) {
  NvtxRange range("stan_cuda_locscale_lpdf_eval");
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_normal_lpdf_eval: null args");
  if (family < STAN_CUDA_LOCSCALE_NORMAL || family > STAN_CUDA_LOCSCALE_INV_GAMMA)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_locscale_lpdf_eval: unknown density");
  if (a->n < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_normal_lpdf_eval: negative n");
  if (a->n == 0) {  // size_zero(y, mu, sigma) -> 0 (normal_lpdf.hpp:45-47)
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  if (!a->y || !a->mu || !a->sigma) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_normal_lpdf_eval: null operand");
  auto ok_len = [&](int64_t len) { return len == 1 || len == a->n; };
  if (!ok_len(a->y_len) || !ok_len(a->mu_len) || !ok_len(a->sigma_len))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_normal_lpdf_eval: operand length must be 1 or n");
  SC_TRY(ensure_init());
  Context& c = context();
  const int dev = tl_scope.active ? c.devices[size_t(tl_scope.device) % c.devices.size()] : c.devices[0];
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  const int64_t n = a->n;
  const bool yv = a->y_len != 1, mv = a->mu_len != 1, sv = a->sigma_len != 1;
  const bool need_mu = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dmu;
  const bool need_sigma = (a->flags & STAN_CUDA_NEED_DSIGMA) && a->dsigma;
  const bool need_y = (a->flags & STAN_CUDA_NEED_DY) && a->dy;
  std::shared_ptr<Entry> ycache;
  if (yv && a->y_is_data) SC_TRY(get_vec_entry(a->y, n, dev, td->stream, &ycache));
  const size_t nal = (size_t(n) + 1) & ~size_t(1);
  size_t need = 8;
  if (yv && !ycache) need += nal;
  if (mv) need += nal;
  if (sv) need += nal;
  if (need_mu && mv) need += nal;
  if (need_sigma && sv) need += nal;
  if (need_y && yv) need += nal;
  SC_TRY(td->reserve_dev(need));
  SC_TRY(td->reserve_host(8));
  double* p = td->d_buf;
  NormalIO io{};
  io.flags = a->flags;
  io.family = family;
  io.n = n;
  io.result = p; p += 8;
  io.y_vec = yv; io.mu_vec = mv; io.sigma_vec = sv;
  io.y0 = a->y[0]; io.mu0 = a->mu[0]; io.sigma0 = a->sigma[0];
  if (yv) {
    if (ycache) {
      io.y = ycache->vec;
    } else {
      SC_TRY(vec_to_device(p, a->y, n, td->stream));
      io.y = p; p += nal;
    }
  }
  if (mv) {
    SC_TRY(vec_to_device(p, a->mu, n, td->stream));
    io.mu = p; p += nal;
  }
  if (sv) {
    SC_TRY(vec_to_device(p, a->sigma, n, td->stream));
    io.sigma = p; p += nal;
  }
  if (need_mu && mv) { io.dmu_row = p; p += nal; }
  if (need_sigma && sv) { io.dsigma_row = p; p += nal; }
  if (need_y && yv) { io.dy_row = p; p += nal; }
  SC_TRY(launch_normal_lpdf(io, &td->ws, dev, td->stream));
  SC_CUDA(cudaMemcpyAsync(td->h_pinned, io.result, 6 * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
  if (io.dmu_row) SC_TRY(vec_to_host(a->dmu, io.dmu_row, n, td->stream));
  if (io.dsigma_row) SC_TRY(vec_to_host(a->dsigma, io.dsigma_row, n, td->stream));
  if (io.dy_row) SC_TRY(vec_to_host(a->dy, io.dy_row, n, td->stream));
  SC_CUDA(cudaStreamSynchronize(td->stream));
  const double* res = td->h_pinned;
  if (a->lp) *a->lp = res[0];
  if (need_mu && !mv) a->dmu[0] = res[1];
  if (need_y && !yv) a->dy[0] = family >= STAN_CUDA_LOCSCALE_LOGNORMAL ? res[5] : -res[1];
  if (need_sigma && !sv) a->dsigma[0] = res[2];
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  const unsigned bits = unsigned(res[4]);
  if (bits & 7u) return STAN_CUDA_BAD_OPERAND;
  if (bits & 8u) {  // lognormal_lpdf with an outcome equal to zero (lognormal_lpdf.hpp:41-45), gamma_lpdf with a negative one
                    // (gamma_lpdf.hpp:69-74): LOG_ZERO, no partials
    if (a->lp) *a->lp = -INFINITY;
    return STAN_CUDA_LOG_ZERO;
  }
  return isfinite(res[3]) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

// device of the calling thread's shard scope, else the first device (gemm.cu)
int scope_device_ordinal() {
  Context& c = context();
  return tl_scope.active ? c.devices[size_t(tl_scope.device) % c.devices.size()] : c.devices[0];
}

// student_t_lpdf (student_t.cu): the four operands travel like those of normal_lpdf -- a length-n data y is cached
// on the device, every other vector is uploaded per call, scalars go by value
int student_t_lpdf_eval(const stan_cuda_student_t_args* a) {
  NvtxRange range("stan_cuda_student_t_lpdf_eval");
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_student_t_lpdf_eval: null args");
  if (a->n < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_student_t_lpdf_eval: negative n");
  if (a->n == 0) {  // size_zero(y, nu, mu, sigma) -> 0 (student_t_lpdf.hpp:57-59)
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  if (!a->y || !a->nu || !a->mu || !a->sigma) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_student_t_lpdf_eval: null operand");
  const int64_t n = a->n;
  const double* src[4] = {a->y, a->nu, a->mu, a->sigma};
  const int64_t len[4] = {a->y_len, a->nu_len, a->mu_len, a->sigma_len};
  double* grad[4] = {a->dy, a->dnu, a->dmu, a->dsigma};
  const uint32_t need_bit[4] = {STAN_CUDA_NEED_DY, STAN_CUDA_NEED_DNU, STAN_CUDA_NEED_DALPHA, STAN_CUDA_NEED_DSIGMA};
  for (int q = 0; q < 4; ++q)
    if (len[q] != 1 && len[q] != n) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_student_t_lpdf_eval: operand length must be 1 or n");
  SC_TRY(ensure_init());
  Context& c = context();
  const int dev = tl_scope.active ? c.devices[size_t(tl_scope.device) % c.devices.size()] : c.devices[0];
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  bool vec[4], need[4];
  uint32_t flags = a->flags & ~uint32_t(STAN_CUDA_NEED_DY | STAN_CUDA_NEED_DNU | STAN_CUDA_NEED_DALPHA | STAN_CUDA_NEED_DSIGMA);
  for (int q = 0; q < 4; ++q) {
    vec[q] = len[q] != 1;
    need[q] = (a->flags & need_bit[q]) && grad[q];
    if (need[q]) flags |= need_bit[q];
  }
  std::shared_ptr<Entry> ycache;
  if (vec[0] && a->y_is_data) SC_TRY(get_vec_entry(a->y, n, dev, td->stream, &ycache));
  const size_t nal = (size_t(n) + 1) & ~size_t(1);
  size_t want = 8;
  for (int q = 0; q < 4; ++q) {
    if (vec[q] && !(q == 0 && ycache)) want += nal;
    if (vec[q] && need[q]) want += nal;
  }
  SC_TRY(td->reserve_dev(want));
  SC_TRY(td->reserve_host(8));
  double* p = td->d_buf;
  StudentTIO io{};
  io.flags = flags;
  io.n = n;
  io.result = p; p += 8;
  io.y0 = a->y[0]; io.nu0 = a->nu[0]; io.mu0 = a->mu[0]; io.sigma0 = a->sigma[0];
  if (!vec[1]) student_t_scalar_dof_terms(&io);  // io.nu stays null for a scalar nu
  const double** in_slot[4] = {&io.y, &io.nu, &io.mu, &io.sigma};
  double** out_slot[4] = {&io.dy_row, &io.dnu_row, &io.dmu_row, &io.dsigma_row};
  for (int q = 0; q < 4; ++q) {
    if (!vec[q]) continue;
    if (q == 0 && ycache) {
      io.y = ycache->vec;
      continue;
    }
    SC_TRY(vec_to_device(p, src[q], n, td->stream));
    *in_slot[q] = p; p += nal;
  }
  for (int q = 0; q < 4; ++q)
    if (vec[q] && need[q]) { *out_slot[q] = p; p += nal; }
  SC_TRY(launch_student_t_lpdf(io, &td->ws, dev, td->stream));
  SC_CUDA(cudaMemcpyAsync(td->h_pinned, io.result, 8 * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
  for (int q = 0; q < 4; ++q)
    if (*out_slot[q]) SC_TRY(vec_to_host(grad[q], *out_slot[q], n, td->stream));
  SC_CUDA(cudaStreamSynchronize(td->stream));
  const double* res = td->h_pinned;
  if (a->lp) *a->lp = res[0];
  for (int q = 0; q < 4; ++q)
    if (need[q] && !vec[q]) grad[q][0] = res[1 + q];
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  if (res[6] != 0.0) return STAN_CUDA_BAD_OPERAND;
  return isfinite(res[5]) ? STAN_CUDA_OK : STAN_CUDA_NONFINITE;
}

// =============================================================================================
// vectorised elementwise densities (density.cu)
// =============================================================================================
static uint64_t fingerprint_ivec(const int32_t* v, int64_t n) {
  uint64_t h = 1469598103934665603ull;
  const int64_t step = n > 128 ? n / 128 : 1;
  for (int64_t i = 0; i < n; i += step) h = fnv(h, uint64_t(uint32_t(v[i])));
  return h;
}

// integer outcomes are immutable data: cached on the device like the y of a GLM (key kind 3)
static int get_ivec_entry(const int32_t* v, int64_t n, int device, int want_lgamma, cudaStream_t stream,
                          std::shared_ptr<Entry>* out) {
  Key key(3, uintptr_t(v), 0, n, 0, want_lgamma, device);
  const uint64_t fp = fingerprint_ivec(v, n);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_cache.find(key);
  if (it != g_cache.end() && it->second->fingerprint == fp) {
    counters().cache_hits.fetch_add(1, std::memory_order_relaxed);
    it->second->last_use = ++g_cache_tick;
    *out = it->second;
    return STAN_CUDA_OK;
  }
  counters().cache_misses.fetch_add(1, std::memory_order_relaxed);
  if (it != g_cache.end()) g_cache.erase(it);
  make_room_locked(n * int64_t(sizeof(double)));
  std::shared_ptr<Entry> e = std::make_shared<Entry>();
  e->fingerprint = fp;
  e->last_use = ++g_cache_tick;
  e->vec_device = device;
  const size_t bytes = ((size_t(n) + 1) & ~size_t(1)) * sizeof(int32_t);
  SC_CUDA(cudaSetDevice(device));
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&e->vec), bytes));
  e->bytes = bytes;
  SC_CUDA(cudaMemcpyAsync(e->vec, v, size_t(n) * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
  if (want_lgamma) {  // data-only term of poisson_log_lpmf (:79-81), once per data set instead of N lgamma per call
    SC_TRY(int_vector_stats(reinterpret_cast<const int*>(e->vec), n, 1, stream, &e->lgamma_sum, nullptr, nullptr));
  } else {
    SC_CUDA(cudaStreamSynchronize(stream));
  }
  counters().bytes_uploaded.fetch_add(uint64_t(n) * sizeof(int32_t));
  g_cache[key] = e;
  *out = e;
  return STAN_CUDA_OK;
}

int density_eval(const stan_cuda_density_args* a) {
  NvtxRange range("stan_cuda_density_eval");
  if (!a) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: null args");
  if (a->density < STAN_CUDA_DENSITY_BERNOULLI_LOGIT || a->density > STAN_CUDA_DENSITY_BERNOULLI)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: unknown density");
  if (a->n < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: negative n");
  if (a->n == 0) {  // size_zero(n, theta) -> 0 (bernoulli_logit_lpmf.hpp:36-38, poisson_log_lpmf.hpp:32-34)
    if (a->lp) *a->lp = 0.0;
    return STAN_CUDA_OK;
  }
  if (!a->y || !a->theta) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: null operand");
  if ((a->y_len != 1 && a->y_len != a->n) || (a->theta_len != 1 && a->theta_len != a->n))
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_density_eval: operand length must be 1 or n");
  SC_TRY(ensure_init());
  Context& c = context();
  const int dev = tl_scope.active ? c.devices[size_t(tl_scope.device) % c.devices.size()] : c.devices[0];
  ThreadDev* td;
  SC_TRY(thread_dev(dev, &td));
  SC_CUDA(cudaSetDevice(dev));
  const int64_t n = a->n;
  const bool yv = a->y_len != 1, tv = a->theta_len != 1;
  const bool need_d = (a->flags & STAN_CUDA_NEED_DALPHA) && a->dtheta;
  std::shared_ptr<Entry> ycache;
  const bool poisson = a->density == STAN_CUDA_DENSITY_POISSON_LOG || a->density == STAN_CUDA_DENSITY_POISSON;
  if (yv) SC_TRY(get_ivec_entry(a->y, n, dev, poisson ? 1 : 0, td->stream, &ycache));
  const size_t nal = (size_t(n) + 1) & ~size_t(1);
  size_t need = 8;
  if (tv) need += nal;
  if (need_d && tv) need += nal;
  SC_TRY(td->reserve_dev(need));
  SC_TRY(td->reserve_host(8));
  double* p = td->d_buf;
  DensityIO io{};
  io.density = a->density;
  io.flags = a->flags;
  io.n = n;
  io.result = p; p += 8;
  io.y_vec = yv; io.theta_vec = tv;
  io.y0 = a->y[0]; io.theta0 = a->theta[0];
  if (yv) io.y = reinterpret_cast<const int*>(ycache->vec);
  if (poisson) {
    io.lgamma_cached = 1;
    io.lgamma_total = yv ? ycache->lgamma_sum : (a->y[0] >= 0 ? double(n) * lgamma(double(a->y[0]) + 1.0) : 0.0);
  }
  if (tv) {
    SC_TRY(vec_to_device(p, a->theta, n, td->stream));
    io.theta = p; p += nal;
  }
  if (need_d && tv) { io.dtheta_row = p; p += nal; }
  SC_TRY(launch_density(io, &td->ws, dev, td->stream));
  SC_CUDA(cudaMemcpyAsync(td->h_pinned, io.result, 5 * sizeof(double), cudaMemcpyDeviceToHost, td->stream));
  // the per-element partials travel in the same batch (one synchronisation); on the two codes below the caller
  // discards them, as the reference has not built any partial at that point either
  if (io.dtheta_row)
    SC_TRY(vec_to_host(a->dtheta, io.dtheta_row, n, td->stream));
  SC_CUDA(cudaStreamSynchronize(td->stream));
  const double* res = td->h_pinned;
  const unsigned bits = unsigned(res[4]);
  counters().evals.fetch_add(1, std::memory_order_relaxed);
  if (bits & 3u) return STAN_CUDA_BAD_OPERAND;
  if (bits & 4u) {
    if (a->lp) *a->lp = -INFINITY;
    return STAN_CUDA_LOG_ZERO;
  }
  if (a->density == STAN_CUDA_DENSITY_BERNOULLI && !tv) {
    // scalar probability: the reference's closed form from the number of successes (bernoulli_lpmf.hpp:63-91; the
    // two one-sided branches avoid 0 * log(0)); the kernel counted them (exact in a double)
    const double N = double(n), ones = res[1], th = a->theta[0];
    double lp = 0.0, d = 0.0;
    if (ones == N) {
      lp = N * log(th);
      d = N / th;
    } else if (ones == 0.0) {
      lp = N * log1p(-th);
      d = N / (th - 1.0);
    } else {
      lp = ones * log(th);
      lp += (N - ones) * log1p(-th);
      d = ones / th;
      d += (N - ones) / (th - 1.0);
    }
    if (a->lp) *a->lp = (a->flags & STAN_CUDA_LP_MAIN) ? lp : 0.0;
    if (need_d) a->dtheta[0] = d;
    return STAN_CUDA_OK;
  }
  if (a->lp) *a->lp = res[0];
  if (need_d && !tv) a->dtheta[0] = res[1];
  return STAN_CUDA_OK;
}

// Pinned (page-locked, portable) staging memory of the calling thread: wrappers gather the values of
// var operands straight into it, so the upload is a full-speed DMA without a bounce buffer.
void* host_staging(size_t bytes) {
  struct Staging {
    void* p = nullptr;
    size_t cap = 0;
  };
  static thread_local Staging st;
  if (bytes <= st.cap) return st.p;
  if (ensure_init() < 0) return nullptr;
  if (st.p) cudaFreeHost(st.p);
  st.p = nullptr;
  st.cap = 0;
  const size_t cap = bytes + bytes / 4 + 4096;
  if (cudaHostAlloc(&st.p, cap, cudaHostAllocPortable) != cudaSuccess) {
    cudaGetLastError();
    st.p = nullptr;
    fail(STAN_CUDA_ERR_CUDA, "stan_cuda_host_staging: cudaHostAlloc failed");
    return nullptr;
  }
  st.cap = cap;
  return st.p;
}

}  // namespace stancuda

using namespace stancuda;

extern "C" {

int stan_cuda_glm_eval(const stan_cuda_glm_args* a) { return glm_eval(a); }
int stan_cuda_glm_eval_batch(const stan_cuda_glm_args* items, int count, int call_id, int* item_rc) {
  return glm_eval_batch(items, count, call_id, item_rc);
}
int stan_cuda_normal_lpdf_eval(const stan_cuda_normal_args* a) { return normal_lpdf_eval(a); }
int stan_cuda_student_t_lpdf_eval(const stan_cuda_student_t_args* a) { return student_t_lpdf_eval(a); }
int stan_cuda_density_eval(const stan_cuda_density_args* a) { return density_eval(a); }
int stan_cuda_locscale_lpdf_eval(int32_t density, const stan_cuda_normal_args* a) { return locscale_lpdf_eval(density, a); }
void* stan_cuda_host_staging(size_t bytes) { return host_staging(bytes); }

int stan_cuda_cache_invalidate(const void* host_ptr) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  int dropped = 0;
  for (auto it = g_cache.begin(); it != g_cache.end();) {
    const Key& k = it->first;
    const bool by_ptr = std::get<0>(k) != 1;
    // inside a shard scope the entries of the CURRENT job (every call ordinal) are dropped: their identity is
    // (call_id, job), not a pointer
    const bool scope_hit = !by_ptr && tl_scope.active && uint32_t(std::get<1>(k)) == uint32_t(tl_scope.call_id) &&
                           std::get<2>(k) == uintptr_t(tl_scope.job);
    if (scope_hit || (by_ptr && (std::get<1>(k) == uintptr_t(host_ptr) || std::get<2>(k) == uintptr_t(host_ptr)))) {
      it = g_cache.erase(it);
      ++dropped;
    } else {
      ++it;
    }
  }
  return dropped;
}

void stan_cuda_cache_clear(void) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  g_cache.clear();
}

void stan_cuda_cache_set_verify(int verify) { g_cache_verify.store(verify ? 1 : 0, std::memory_order_relaxed); }

void stan_cuda_cache_set_limit(int64_t bytes) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  g_cache_limit = bytes > 0 ? bytes : 0;
  if (g_cache_limit > 0) make_room_locked(0);
}

int64_t stan_cuda_cache_bytes(void) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  int64_t b = 0;
  for (auto& kv : g_cache) b += int64_t(kv.second->bytes);
  return b;
}

// Scopes nest: exit restores the scope that was current at the matching enter (a job whose functor runs nested
// TBB parallelism of its own can be handed another job of the same map_rect call while it waits).
static thread_local Scope tl_scope_saved[8];
static thread_local int tl_scope_depth = 0;

void stan_cuda_shard_scope_enter(int call_id, int job, int device) {
  if (tl_scope_depth < 8) tl_scope_saved[tl_scope_depth] = tl_scope;
  ++tl_scope_depth;
  tl_scope.active = true;
  tl_scope.call_id = call_id;
  tl_scope.job = job;
  tl_scope.device = device < 0 ? 0 : device;
  tl_scope.ordinal = 0;
}

void stan_cuda_shard_scope_exit(void) {
  if (tl_scope_depth > 0) --tl_scope_depth;
  if (tl_scope_depth < 8) tl_scope = tl_scope_saved[tl_scope_depth];
  if (tl_scope_depth == 0) tl_scope.active = false;
}

}  // extern "C"
