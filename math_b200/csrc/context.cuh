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
  unsigned long long host_seq = 0;  // completion-flag sequence of the low-latency path (api.cu)
  double* d_buf = nullptr;
  size_t d_cap = 0;  // doubles
  double* d_ybuf = nullptr;  // per-call outcome vector (stan_cuda_glm_args.y_per_call), separate from d_buf
  size_t d_ycap = 0;         // doubles
  Workspace ws;
  int reserve_host(size_t doubles);
  int reserve_dev(size_t doubles);
  int reserve_y(size_t doubles);
};
int thread_dev(int device, ThreadDev** out);

// data blocks (data.cu)
int make_data_block(int device, int family, int64_t n_rows, int64_t k, const double* x, int64_t ldx,
                    int x_row_major, const void* y, int y_is_double, int64_t row0, cudaStream_t stream,
                    stan_cuda_data** out);
void free_data_block(stan_cuda_data* d);
int refill_data_block(stan_cuda_data* d, const double* x, int64_t ldx, int x_row_major, const void* y, int64_t row0,
                      cudaStream_t stream);

// Peer exchange (comm.cu): all-reduce fused into the GLM kernel over NVLink peer memory --------
// Every device owns one exchange block; the last CTA of the fused kernel pushes its result block
// into slot [seq & 1][rank] of EVERY device's block (plain stores to peer-mapped memory), raises
// a flag per peer, waits for the world's flags in its own block and adds the blocks in rank order.
constexpr int kXchgMaxWorld = 16;
constexpr int kXchgCols = 520;  // >= kpad_max (512) + 4, per-rank stride in doubles
struct PeerXchg {
  int world = 0;                       // 0: not available
  int rank = 0;
  double* local = nullptr;             // this device's block: [2][world][kXchgCols] doubles + flags
  double* peers[kXchgMaxWorld] = {};   // peers[q]: device q's block as mapped into this device
  unsigned long long seq = 0;          // sequence number of the collective (same on every rank)
};
inline size_t xchg_flag_offset_doubles(int world) { return size_t(2) * world * kXchgCols; }
inline size_t xchg_block_bytes(int world) { return (xchg_flag_offset_doubles(world) + size_t(2) * world + 8) * sizeof(double); }
// multi-process communicator: exchange of this rank, or world == 0 (fall back to ncclAllReduce)
PeerXchg* comm_rank_xchg();
// in-process communicator over `devices`: exchanges per device (index = position), or nullptr
PeerXchg* comm_local_xchg(const std::vector<int>& devices);
bool fused_allreduce_enabled();

// NCCL (comm.cu) -------------------------------------------------------------------------------
int comm_world_size();                       // ranks of the multi-process communicator (1 if none)
int comm_allreduce_rank(double* d_buf, int64_t count, cudaStream_t stream);
// in-process communicator over the context's devices; bufs[i]/streams[i] belong to devices[i]
int comm_allreduce_local(const std::vector<int>& devices, const std::vector<double*>& bufs, int64_t count,
                         const std::vector<cudaStream_t>& streams);

}  // namespace stancuda
