// Collectives for the data-parallel learner.  The reference's LGBM_NetworkInit(machines, port, timeout, n)
// (lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/NetworkManager.scala:195-218) builds a TCP mesh
// inside lightgbmlib and LGBM_NetworkFree (BasePartitionTask.scala:282) tears it down.  Here the same two calls open a TCP
// star through rank 0 (the first entry of the machine list the Scala NetworkManager already assembles) that stays up for
// the life of the network: it carries the few host-side values of a fit (row counts, label sums, CUDA IPC handles).  Bins
// and split records move over NVLink peer memory (batch_tree.cuh); NCCL is loaded and a communicator is built only if peer
// memory cannot be mapped (EnsureNccl), so a fit on one NVSwitch node pays no communicator set-up.
// One process per GPU; rank = position of this process's "host:port" in the machine list.
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <mutex>
#include <string>
#include <vector>

namespace lgb {

class Network {
 public:
  static Network* Get();
  void Init(const char* machines, int local_listen_port, int listen_time_out_s, int num_machines);
  void Free();
  int rank() const { return rank_; }
  int world() const { return world_; }
  bool active() const { return world_ > 1; }
  int timeout_s() const { return timeout_s_; }
  // NCCL fallback: loads libnccl and builds the communicator (collective: every rank must call it).  No-op when it exists.
  void EnsureNccl(int device);
  bool has_nccl() const { return comm_ != nullptr; }

  // device-buffer collectives of the NCCL fallback, enqueued on `stream`
  void AllReduceSumInt64(long long* buf, size_t n, cudaStream_t stream);
  void AllReduceMaxU32(unsigned int* buf, size_t n, cudaStream_t stream);
  void ReduceScatterSumInt64(const long long* send, long long* recv, size_t recv_count, cudaStream_t stream);
  void AllGatherBytes(const void* send, void* recv, size_t bytes_per_rank, cudaStream_t stream);
  // small host-value collectives over the TCP star (blocking; rank order, so every rank computes the same sums)
  void HostAllGatherBytes(const void* mine, size_t bytes, std::vector<unsigned char>* all);
  void HostAllReduceSum(double* vals, int n);
  void HostAllGatherInt64(long long mine, std::vector<long long>* all);
  // Peer-memory exchange (all ranks on one NVLink / NVSwitch node): every rank contributes one cudaMalloc'ed buffer and gets
  // the device addresses of all ranks' buffers in its own address space (own rank: the local pointer).  Blocking, set-up time.
  // Returns false (and leaves `peers` empty) when a peer's memory cannot be mapped; the caller then falls back to NCCL.
  bool ExchangePeerPointers(int device, void* local, std::vector<void*>* peers);
  void ClosePeerPointers(const std::vector<void*>& peers);

  // parse "h:p,h:p,..." and find this process's rank; exposed for CPU tests
  static std::vector<std::pair<std::string, int>> ParseMachines(const std::string& machines);
  static int FindRank(const std::vector<std::pair<std::string, int>>& m, int local_listen_port);
  // one-shot TCP hand-off of an opaque blob from rank 0 to everyone; exposed for CPU tests
  static void BroadcastBlob(const std::vector<std::pair<std::string, int>>& m, int rank, int timeout_s, void* blob, size_t bytes);

 private:
  int rank_ = 0, world_ = 1, device_ = -1, timeout_s_ = 120;
  std::mutex mu_;
  std::vector<int> fds_;       // rank 0: socket to every other rank ([0] unused); other ranks: [0] = socket to rank 0
  void CloseSockets();
  void* comm_ = nullptr;
  void* nccl_lib_ = nullptr;
  void LoadNccl();
};

}  // namespace lgb
