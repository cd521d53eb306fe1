// comm.cu -- NCCL all-reduce of the per-shard result block [lp, sum d, dsigma, witness, dbeta...].
//
// The reference has no device collective (its OpenCL backend is single-device,
// stan/math/opencl/opencl_context.hpp:48-53); its multi-process map_rect gathers with Boost.MPI
// (stan/math/prim/mat/functor/mpi_parallel_call.hpp:387-392).  Here every row shard produces the
// same K+4 additive doubles and NCCL sums them over NVLink/NVSwitch, on the evaluation stream,
// right behind the fused kernel.  Two shapes:
//   * one process per GPU (torchrun): the host plumbing hands every rank the 128-byte unique id,
//     stan_cuda_comm_init_rank builds the communicator;
//   * one process, several GPUs (a Stan program): ncclCommInitAll over the context's devices,
//     group call from the evaluating thread.
// libnccl.so.2 is loaded with dlopen on first use so the library itself has no link-time NCCL
// dependency (torch's bundled NCCL is reused when the process already has it).
#include <dlfcn.h>

#include "context.cuh"

namespace stancuda {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclSum = 0 };
enum { ncclFloat64 = 8 };

struct NcclApi {
  void* handle = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
};

static std::mutex g_mu;
static NcclApi g_api;
static ncclComm_t g_rank_comm = nullptr;
static int g_world = 1, g_rank = 0, g_rank_device = -1;
static std::vector<ncclComm_t> g_local_comms;
static std::vector<int> g_local_devices;

static int fail_nccl(int rc, const char* what) {
  std::string s = std::string("NCCL error in ") + what + ": ";
  s += g_api.GetErrorString ? g_api.GetErrorString(rc) : "unknown";
  return fail(STAN_CUDA_ERR_NCCL, s);
}

static int load_nccl_locked() {
  if (g_api.handle) return STAN_CUDA_OK;
  const char* names[] = {getenv("STAN_CUDA_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* nm : names) {
    if (!nm || !*nm) continue;
    h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) return fail(STAN_CUDA_ERR_NCCL, std::string("cannot load libnccl.so.2: ") + dlerror());
#define SC_SYM(field, name)                                                        \
  g_api.field = reinterpret_cast<decltype(g_api.field)>(dlsym(h, name));            \
  if (!g_api.field) return fail(STAN_CUDA_ERR_NCCL, std::string("libnccl lacks ") + name);
  SC_SYM(GetUniqueId, "ncclGetUniqueId")
  SC_SYM(CommInitRank, "ncclCommInitRank")
  SC_SYM(CommInitAll, "ncclCommInitAll")
  SC_SYM(CommDestroy, "ncclCommDestroy")
  SC_SYM(AllReduce, "ncclAllReduce")
  SC_SYM(GroupStart, "ncclGroupStart")
  SC_SYM(GroupEnd, "ncclGroupEnd")
  SC_SYM(GetErrorString, "ncclGetErrorString")
  SC_SYM(GetVersion, "ncclGetVersion")
#undef SC_SYM
  g_api.handle = h;
  return STAN_CUDA_OK;
}

int comm_world_size() { return g_world; }

int comm_allreduce_rank(double* d_buf, int64_t count, cudaStream_t stream) {
  if (!g_rank_comm) return fail(STAN_CUDA_ERR_STATE, "no multi-process communicator (stan_cuda_comm_init_rank)");
  int rc = g_api.AllReduce(d_buf, d_buf, size_t(count), ncclFloat64, ncclSum, g_rank_comm, stream);
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclAllReduce");
  counters().allreduces.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

int comm_allreduce_local(const std::vector<int>& devices, const std::vector<double*>& bufs, int64_t count,
                         const std::vector<cudaStream_t>& streams) {
  std::lock_guard<std::mutex> lk(g_mu);
  SC_TRY(load_nccl_locked());
  if (g_local_comms.empty() || g_local_devices != devices) {
    for (ncclComm_t c : g_local_comms) g_api.CommDestroy(c);
    g_local_comms.assign(devices.size(), nullptr);
    int rc = g_api.CommInitAll(g_local_comms.data(), int(devices.size()), devices.data());
    if (rc != ncclSuccess) {
      g_local_comms.clear();
      return fail_nccl(rc, "ncclCommInitAll");
    }
    g_local_devices = devices;
  }
  int rc = g_api.GroupStart();
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclGroupStart");
  for (size_t i = 0; i < devices.size(); ++i) {
    rc = g_api.AllReduce(bufs[i], bufs[i], size_t(count), ncclFloat64, ncclSum, g_local_comms[i], streams[i]);
    if (rc != ncclSuccess) {
      g_api.GroupEnd();
      return fail_nccl(rc, "ncclAllReduce (group)");
    }
  }
  rc = g_api.GroupEnd();
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclGroupEnd");
  counters().allreduces.fetch_add(1, std::memory_order_relaxed);
  return STAN_CUDA_OK;
}

}  // namespace stancuda

using namespace stancuda;

extern "C" {

int stan_cuda_comm_unique_id(void* id128) {
  if (!id128) return fail(STAN_CUDA_ERR_ARG, "stan_cuda_comm_unique_id: null");
  std::lock_guard<std::mutex> lk(g_mu);
  SC_TRY(load_nccl_locked());
  ncclUniqueId id;
  int rc = g_api.GetUniqueId(&id);
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclGetUniqueId");
  memcpy(id128, &id, sizeof(id));
  return STAN_CUDA_OK;
}

int stan_cuda_comm_init_rank(int world_size, int rank, const void* id128) {
  if (!id128 || world_size < 1 || rank < 0 || rank >= world_size)
    return fail(STAN_CUDA_ERR_ARG, "stan_cuda_comm_init_rank: bad arguments");
  SC_TRY(ensure_init());
  if (context().devices.size() != 1)
    return fail(STAN_CUDA_ERR_STATE,
                "stan_cuda_comm_init_rank: a multi-process communicator needs one device per process "
                "(call stan_cuda_init_device first)");
  std::lock_guard<std::mutex> lk(g_mu);
  SC_TRY(load_nccl_locked());
  if (g_rank_comm) return fail(STAN_CUDA_ERR_STATE, "stan_cuda_comm_init_rank: communicator already exists");
  const int dev = context().devices[0];
  SC_CUDA(cudaSetDevice(dev));
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclComm_t comm = nullptr;
  int rc = g_api.CommInitRank(&comm, world_size, id, rank);
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclCommInitRank");
  g_rank_comm = comm;
  g_world = world_size;
  g_rank = rank;
  g_rank_device = dev;
  return STAN_CUDA_OK;
}

int stan_cuda_comm_world_size(void) { return g_world; }

void stan_cuda_comm_destroy(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_rank_comm && g_api.CommDestroy) g_api.CommDestroy(g_rank_comm);
  g_rank_comm = nullptr;
  g_world = 1;
  g_rank = 0;
  for (ncclComm_t c : g_local_comms)
    if (c && g_api.CommDestroy) g_api.CommDestroy(c);
  g_local_comms.clear();
  g_local_devices.clear();
}

int stan_cuda_allreduce(double* d_buf, int64_t count, void* stream) {
  if (g_world <= 1) return STAN_CUDA_OK;
  return comm_allreduce_rank(d_buf, count, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
