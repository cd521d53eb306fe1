// context.cuh -- process-wide context and per-(thread, device) state.
#pragma once
#include "common.cuh"

namespace stancuda {

struct Context {
  std::mutex mu;
  bool inited = false;
  std::vector<int> devices;       // CUDA ordinals used by this process
  std::string description;
  int64_t shard_min_elems = int64_t(1) << 24;
};
Context& context();
int init_devices(int n_devices, int single_ordinal);
int ensure_init();

// One per (host thread, device): stream, pinned staging, device scratch, reduction workspace.
struct ThreadDev {
  int device = -1;
  cudaStream_t stream = nullptr;
  double* h_pinned = nullptr;
  size_t h_cap = 0;  // doubles
  double* d_buf = nullptr;
  size_t d_cap = 0;  // doubles
  Workspace ws;
  int reserve_host(size_t doubles);
  int reserve_dev(size_t doubles);
};
int thread_dev(int device, ThreadDev** out);

// data blocks (data.cu)
int make_data_block(int device, int family, int64_t n_rows, int64_t k, const double* x, int64_t ldx,
                    int x_row_major, const void* y, int y_is_double, int64_t row0, cudaStream_t stream,
                    stan_cuda_data** out);
void free_data_block(stan_cuda_data* d);

// NCCL (comm.cu) -------------------------------------------------------------------------------
int comm_world_size();                       // ranks of the multi-process communicator (1 if none)
int comm_allreduce_rank(double* d_buf, int64_t count, cudaStream_t stream);
// in-process communicator over the context's devices; bufs[i]/streams[i] belong to devices[i]
int comm_allreduce_local(const std::vector<int>& devices, const std::vector<double*>& bufs, int64_t count,
                         const std::vector<cudaStream_t>& streams);

}  // namespace stancuda
