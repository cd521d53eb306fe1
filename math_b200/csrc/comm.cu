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

#include <cstring>

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
  int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
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
// peer exchange state (fused all-reduce)
static PeerXchg g_rank_xchg;                    // multi-process: this rank's view
static std::vector<void*> g_rank_ipc_opened;    // peer blocks opened with cudaIpcOpenMemHandle
static std::vector<PeerXchg> g_local_xchg;      // in-process: one view per device
static std::vector<int> g_local_xchg_devices;
static std::string g_xchg_note = "not initialised";

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
  SC_SYM(AllGather, "ncclAllGather")
  SC_SYM(GroupStart, "ncclGroupStart")
  SC_SYM(GroupEnd, "ncclGroupEnd")
  SC_SYM(GetErrorString, "ncclGetErrorString")
  SC_SYM(GetVersion, "ncclGetVersion")
#undef SC_SYM
  g_api.handle = h;
  return STAN_CUDA_OK;
}

bool fused_allreduce_enabled() {
  static const bool on = [] {
    const char* v = getenv("STAN_CUDA_FUSED_ALLREDUCE");
    return !(v && *v == '0');
  }();
  return on;
}

// ---- peer exchange set-up ----------------------------------------------------------------------
// Multi-process: every rank allocates its block, the CUDA IPC handles travel through one
// ncclAllGather, every rank maps the others' blocks; a final ncclAllReduce of an ok-flag makes the
// decision unanimous (and doubles as the barrier after which peers may write).
static int setup_rank_xchg_locked(int world, int rank, int dev) {
  g_rank_xchg = PeerXchg();
  if (!fused_allreduce_enabled()) { g_xchg_note = "disabled by STAN_CUDA_FUSED_ALLREDUCE=0"; return STAN_CUDA_OK; }
  if (world > kXchgMaxWorld) { g_xchg_note = "world larger than 16"; return STAN_CUDA_OK; }
  SC_CUDA(cudaSetDevice(dev));
  double ok = 1.0;
  double* local = nullptr;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (cudaMalloc(reinterpret_cast<void**>(&local), xchg_block_bytes(world)) != cudaSuccess ||
      cudaMemset(local, 0, xchg_block_bytes(world)) != cudaSuccess ||
      cudaIpcGetMemHandle(&mine, local) != cudaSuccess) {
    cudaGetLastError();
    ok = 0.0;
  }
  // gather the handles
  const size_t hb = sizeof(cudaIpcMemHandle_t);
  unsigned char* d_handles = nullptr;
  double* d_ok = nullptr;
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_handles), hb * (world + 1)));
  SC_CUDA(cudaMalloc(reinterpret_cast<void**>(&d_ok), sizeof(double)));
  SC_CUDA(cudaMemcpy(d_handles + hb * world, &mine, hb, cudaMemcpyHostToDevice));
  int rc = g_api.AllGather(d_handles + hb * world, d_handles, hb, /*ncclChar*/ 0, g_rank_comm, nullptr);
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclAllGather (ipc handles)");
  SC_CUDA(cudaStreamSynchronize(nullptr));
  std::vector<cudaIpcMemHandle_t> all(world);
  SC_CUDA(cudaMemcpy(all.data(), d_handles, hb * world, cudaMemcpyDeviceToHost));
  PeerXchg x;
  x.world = world;
  x.rank = rank;
  x.local = local;
  std::vector<void*> opened;
  for (int q = 0; q < world && ok > 0; ++q) {
    if (q == rank) { x.peers[q] = local; continue; }
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      ok = 0.0;
      break;
    }
    opened.push_back(ptr);
    x.peers[q] = static_cast<double*>(ptr);
  }
  SC_CUDA(cudaMemcpy(d_ok, &ok, sizeof(double), cudaMemcpyHostToDevice));
  rc = g_api.AllReduce(d_ok, d_ok, 1, ncclFloat64, ncclSum, g_rank_comm, nullptr);
  if (rc != ncclSuccess) return fail_nccl(rc, "ncclAllReduce (peer exchange agreement)");
  SC_CUDA(cudaStreamSynchronize(nullptr));
  double total = 0.0;
  SC_CUDA(cudaMemcpy(&total, d_ok, sizeof(double), cudaMemcpyDeviceToHost));
  cudaFree(d_handles);
  cudaFree(d_ok);
  if (total == double(world)) {
    g_rank_xchg = x;
    g_rank_ipc_opened = opened;
    g_xchg_note = "fused peer exchange over CUDA IPC";
  } else {
    for (void* ptr : opened) cudaIpcCloseMemHandle(ptr);
    if (local) cudaFree(local);
    g_xchg_note = "CUDA IPC peer mapping unavailable on some rank: ncclAllReduce is used";
  }
  return STAN_CUDA_OK;
}

PeerXchg* comm_rank_xchg() { return g_rank_xchg.world > 1 ? &g_rank_xchg : nullptr; }

// In-process: peer access between every pair of devices, one block per device.
PeerXchg* comm_local_xchg(const std::vector<int>& devices) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!fused_allreduce_enabled()) return nullptr;
  const int world = int(devices.size());
  if (world < 2 || world > kXchgMaxWorld) return nullptr;
  if (g_local_xchg_devices == devices) return g_local_xchg.empty() ? nullptr : g_local_xchg.data();
  g_local_xchg_devices = devices;
  g_local_xchg.clear();
  std::vector<double*> blocks(world, nullptr);
  bool ok = true;
  for (int i = 0; i < world && ok; ++i) {
    ok = cudaSetDevice(devices[i]) == cudaSuccess;
    for (int j = 0; j < world && ok; ++j) {
      if (i == j) continue;
      int can = 0;
      ok = cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) == cudaSuccess && can;
      if (ok) {
        cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        ok = e == cudaSuccess;
      }
    }
    if (ok) ok = cudaMalloc(reinterpret_cast<void**>(&blocks[i]), xchg_block_bytes(world)) == cudaSuccess &&
                 cudaMemset(blocks[i], 0, xchg_block_bytes(world)) == cudaSuccess && cudaDeviceSynchronize() == cudaSuccess;
  }
  if (!ok) {
    cudaGetLastError();
    for (int i = 0; i < world; ++i)
      if (blocks[i]) { cudaSetDevice(devices[i]); cudaFree(blocks[i]); }
    g_xchg_note = "peer access between the devices of the process unavailable: ncclAllReduce is used";
    return nullptr;
  }
  g_local_xchg.resize(world);
  for (int i = 0; i < world; ++i) {
    g_local_xchg[i].world = world;
    g_local_xchg[i].rank = i;
    g_local_xchg[i].local = blocks[i];
    for (int q = 0; q < world; ++q) g_local_xchg[i].peers[q] = blocks[q];
  }
  g_xchg_note = "fused peer exchange (in-process peer access)";
  return g_local_xchg.data();
}

int comm_world_size() { return g_world; }

int comm_allreduce_rank(double* d_buf, int64_t count, cudaStream_t stream) {
  NvtxRange range("stan_cuda ncclAllReduce (result block)");
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
  if (world_size > 1) SC_TRY(setup_rank_xchg_locked(world_size, rank, dev));
  return STAN_CUDA_OK;
}

int stan_cuda_comm_world_size(void) { return g_world; }

int stan_cuda_comm_fused_available(void) { return g_rank_xchg.world > 1 ? 1 : 0; }

const char* stan_cuda_comm_describe(void) { return g_xchg_note.c_str(); }

void stan_cuda_comm_destroy(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_rank_xchg.world > 1) {
    cudaDeviceSynchronize();
    for (void* ptr : g_rank_ipc_opened) cudaIpcCloseMemHandle(ptr);
    g_rank_ipc_opened.clear();
    if (g_rank_xchg.local) cudaFree(g_rank_xchg.local);
    g_rank_xchg = PeerXchg();
  }
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
