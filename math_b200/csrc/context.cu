// context.cu -- device enumeration, per-thread streams/staging, error text, counters.
// Plays the role of opencl_context (stan/math/opencl/opencl_context.hpp:54-383) for the CUDA
// subsystem, with two differences the north star asks for: several devices per process, and one
// stream + staging area per (host thread, device) so TBB workers never share a queue.
#include "context.cuh"

#include <cstdlib>
#include <cstring>
#include <sstream>

namespace stancuda {

static thread_local std::string tl_error;

void set_last_error(const std::string& s) { tl_error = s; }

int fail(int code, const std::string& what) {
  tl_error = what;
  return code;
}

int fail_cuda(cudaError_t e, const char* expr, const char* file, int line) {
  std::ostringstream os;
  os << "CUDA error " << int(e) << " (" << cudaGetErrorName(e) << ": " << cudaGetErrorString(e) << ") at "
     << file << ":" << line << " in " << expr;
  tl_error = os.str();
  cudaGetLastError();  // clear the sticky-less error state
  return STAN_CUDA_ERR_CUDA;
}

Counters& counters() {
  static Counters c;
  return c;
}

Context& context() {
  static Context c;
  return c;
}

static int describe_locked(Context& c) {
  std::ostringstream os;
  os << "stan::math::cuda  devices=" << c.devices.size();
  for (size_t i = 0; i < c.devices.size(); ++i) {
    cudaDeviceProp prop;
    SC_CUDA(cudaGetDeviceProperties(&prop, c.devices[i]));
    os << "\n  [" << i << "] cuda:" << c.devices[i] << " " << prop.name << " sm_" << prop.major << prop.minor
       << " SMs=" << prop.multiProcessorCount << " mem=" << (prop.totalGlobalMem >> 20) << " MiB"
       << " smem/block(optin)=" << (prop.sharedMemPerBlockOptin >> 10) << " KiB";
  }
  c.description = os.str();
  return STAN_CUDA_OK;
}

int init_devices(int n_devices, int single_ordinal) {
  Context& c = context();
  std::lock_guard<std::mutex> lk(c.mu);
  if (c.inited) {
    if (single_ordinal >= 0 && (c.devices.size() != 1 || c.devices[0] != single_ordinal))
      return fail(STAN_CUDA_ERR_STATE, "stan_cuda_init_device: already initialised with other devices");
    return STAN_CUDA_OK;
  }
  int visible = 0;
  cudaError_t e = cudaGetDeviceCount(&visible);
  if (e != cudaSuccess || visible <= 0) {
    cudaGetLastError();
    return fail(STAN_CUDA_ERR_CUDA,
                std::string("no CUDA device available: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "count is 0") +
                    " (this library has no CPU fallback)");
  }
  std::vector<int> devs;
  if (single_ordinal >= 0) {
    if (single_ordinal >= visible) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_init_device: ordinal out of range");
    devs.push_back(single_ordinal);
  } else {
    if (n_devices <= 0) {
      const char* env = getenv("STAN_CUDA_DEVICES");
      n_devices = (env && *env) ? atoi(env) : visible;
      if (n_devices <= 0) n_devices = visible;
    }
    if (n_devices > visible) n_devices = visible;
    for (int i = 0; i < n_devices; ++i) devs.push_back(i);
  }
  for (int d : devs) {
    cudaDeviceProp prop;
    SC_CUDA(cudaGetDeviceProperties(&prop, d));
    if (prop.major < 10)
      return fail(STAN_CUDA_ERR_STATE, std::string("device ") + prop.name +
                                           " is not sm_100-class; this library is built for sm_100a only");
  }
  c.devices = devs;
  // peer access between the devices of the process (needed by nothing on the data path today:
  // shards never read each other's memory; kept so NCCL can pick P2P transports)
  c.inited = true;
  return describe_locked(c);
}

int ensure_init() {
  Context& c = context();
  if (c.inited) return STAN_CUDA_OK;
  return init_devices(0, -1);
}

// ---- per-thread, per-device state ------------------------------------------------------------
struct ThreadState {
  std::vector<ThreadDev*> devs;  // indexed by device ordinal; leaked at thread exit on purpose
                                 // (CUDA may already be torn down when TLS destructors run)
};
static thread_local ThreadState tl_state;

int thread_dev(int device, ThreadDev** out) {
  if (device < 0) return fail(STAN_CUDA_ERR_ARG, "thread_dev: negative device");
  if (size_t(device) >= tl_state.devs.size()) tl_state.devs.resize(device + 1, nullptr);
  ThreadDev* td = tl_state.devs[device];
  if (!td) {
    SC_CUDA(cudaSetDevice(device));
    td = new ThreadDev();
    td->device = device;
    SC_CUDA(cudaStreamCreateWithFlags(&td->stream, cudaStreamNonBlocking));
    tl_state.devs[device] = td;
  }
  *out = td;
  return STAN_CUDA_OK;
}

int ThreadDev::reserve_host(size_t doubles) {
  if (doubles <= h_cap) return STAN_CUDA_OK;
  SC_CUDA(cudaSetDevice(device));
  if (h_pinned) SC_CUDA(cudaFreeHost(h_pinned));
  h_pinned = nullptr;
  h_cap = 0;
  size_t cap = doubles < 1024 ? 1024 : doubles + doubles / 4;
  // mapped + portable: kernels of any device may write results into it (low-latency path, api.cu)
  SC_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&h_pinned), cap * sizeof(double), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(h_pinned, 0, cap * sizeof(double));
  h_cap = cap;
  return STAN_CUDA_OK;
}

int ThreadDev::reserve_dev(size_t doubles) {
  if (doubles <= d_cap) return STAN_CUDA_OK;
  SC_CUDA(cudaSetDevice(device));
  if (d_buf) {
    SC_CUDA(cudaStreamSynchronize(stream));
    SC_CUDA(cudaFree(d_buf));
  }
  d_buf = nullptr;
  d_cap = 0;
  size_t cap = doubles < 4096 ? 4096 : doubles + doubles / 4;
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_buf), cap * sizeof(double)));
  d_cap = cap;
  return STAN_CUDA_OK;
}

int ThreadDev::reserve_y(size_t doubles) {
  if (doubles <= d_ycap) return STAN_CUDA_OK;
  SC_CUDA(cudaSetDevice(device));
  if (d_ybuf) {
    SC_CUDA(cudaStreamSynchronize(stream));
    SC_CUDA(cudaFree(d_ybuf));
  }
  d_ybuf = nullptr;
  d_ycap = 0;
  const size_t cap = doubles < 4096 ? 4096 : doubles + doubles / 4;
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_ybuf), cap * sizeof(double)));
  d_ycap = cap;
  return STAN_CUDA_OK;
}

int workspace_reserve(Workspace* w, int device, size_t doubles) {
  if (w->device == device && w->scratch && doubles <= w->scratch_doubles) return STAN_CUDA_OK;
  if (w->fixed)
    return fail(STAN_CUDA_ERR_ARG, "this problem needs " + std::to_string(doubles) + " doubles of reduction scratch, the fixed "
                                   "scratch of the no-allocation entry point holds " + std::to_string(w->scratch_doubles));
  SC_CUDA(cudaSetDevice(device));
  if (w->scratch && w->device == device) {
    SC_CUDA(cudaDeviceSynchronize());
    SC_CUDA(cudaFree(w->scratch));
    w->scratch = nullptr;
  }
  if (!w->ticket || w->device != device) {
    SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&w->ticket), 256));
    SC_CUDA(cudaMemset(w->ticket, 0, 256));
  }
  size_t cap = doubles < 65536 ? 65536 : doubles + doubles / 2;
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&w->scratch), cap * sizeof(double)));
  w->scratch_doubles = cap;
  w->device = device;
  return STAN_CUDA_OK;
}

void workspace_free(Workspace* w) {
  if (w->device < 0) return;
  cudaSetDevice(w->device);
  if (w->scratch) cudaFree(w->scratch);
  if (w->ticket) cudaFree(w->ticket);
  w->scratch = nullptr;
  w->ticket = nullptr;
  w->scratch_doubles = 0;
  w->device = -1;
}

}  // namespace stancuda

using namespace stancuda;

extern "C" {

int stan_cuda_init(int n_devices) { return init_devices(n_devices, -1); }

int stan_cuda_init_device(int device_ordinal) {
  if (device_ordinal < 0) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_init_device: negative ordinal");
  return init_devices(1, device_ordinal);
}

int stan_cuda_device_count(void) {
  if (ensure_init() < 0) return 0;
  return int(context().devices.size());
}

const char* stan_cuda_describe(void) {
  if (ensure_init() < 0) return stan_cuda_last_error();
  return context().description.c_str();
}

const char* stan_cuda_last_error(void) {
  static thread_local std::string copy;
  copy = tl_error;
  return copy.c_str();
}

int stan_cuda_stats(stan_cuda_stats_t* out) {
  if (!out) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_stats: null");
  Counters& c = counters();
  out->cache_hits = c.cache_hits.load();
  out->cache_misses = c.cache_misses.load();
  out->bytes_uploaded = c.bytes_uploaded.load();
  out->evals = c.evals.load();
  out->kernel_launches = c.launches.load();
  out->allreduces = c.allreduces.load();
  out->last_fill_ms = c.last_fill_ms.load();
  out->fused_allreduces = c.fused_allreduces.load();
  out->cache_evictions = c.cache_evictions.load();
  return STAN_CUDA_OK;
}

void stan_cuda_stats_reset(void) {
  Counters& c = counters();
  c.cache_hits = 0;
  c.cache_misses = 0;
  c.bytes_uploaded = 0;
  c.evals = 0;
  c.launches = 0;
  c.allreduces = 0;
  c.fused_allreduces = 0;
  c.cache_evictions = 0;
  c.last_fill_ms = 0.0;
}

void stan_cuda_set_shard_min_elems(int64_t elems) { context().shard_min_elems = elems; }

}  // extern "C"
