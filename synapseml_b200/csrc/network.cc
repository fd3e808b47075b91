#include "network.h"

#include <arpa/inet.h>
#include <dlfcn.h>
#include <ifaddrs.h>
#include <netdb.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>

#include <chrono>
#include <cstring>
#include <thread>

#include "util.h"

namespace lgb {

// ---- minimal NCCL ABI (stable across 2.x); resolved with dlsym so the library loads on a box without NCCL and a
// single-GPU fit never touches it.
namespace {
typedef struct { char internal[128]; } NcclUniqueId;
typedef int (*FnGetUniqueId)(NcclUniqueId*);
typedef int (*FnCommInitRank)(void**, int, NcclUniqueId, int);
typedef int (*FnCommDestroy)(void*);
typedef const char* (*FnGetErrorString)(int);
typedef int (*FnAllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*FnReduceScatter)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*FnAllGather)(const void*, void*, size_t, int, void*, cudaStream_t);
FnGetUniqueId p_GetUniqueId;
FnCommInitRank p_CommInitRank;
FnCommDestroy p_CommDestroy;
FnGetErrorString p_GetErrorString;
FnAllReduce p_AllReduce;
FnReduceScatter p_ReduceScatter;
FnAllGather p_AllGather;
// ncclDataType_t / ncclRedOp_t values (nccl.h)
constexpr int kNcclInt8 = 0, kNcclUint32 = 3, kNcclInt64 = 4, kNcclFloat64 = 8;
constexpr int kNcclSum = 0, kNcclMax = 2;

#define LGB_NCCL(expr)                                                                                            \
  do {                                                                                                            \
    int _r = (expr);                                                                                              \
    if (_r != 0) throw std::runtime_error(StrFormat("NCCL error '%s' at %s:%d", p_GetErrorString(_r), __FILE__, __LINE__)); \
  } while (0)

void SendAll(int fd, const void* buf, size_t n) {
  const char* p = static_cast<const char*>(buf);
  while (n > 0) {
    ssize_t k = send(fd, p, n, MSG_NOSIGNAL);
    LGB_CHECK(k > 0, "Network bootstrap: send failed (%s)", strerror(errno));
    p += k; n -= k;
  }
}
void RecvAll(int fd, void* buf, size_t n) {
  char* p = static_cast<char*>(buf);
  while (n > 0) {
    ssize_t k = recv(fd, p, n, 0);
    LGB_CHECK(k > 0, "Network bootstrap: recv failed (%s)", k == 0 ? "peer closed" : strerror(errno));
    p += k; n -= k;
  }
}
}  // namespace

Network* Network::Get() {
  static Network inst;
  return &inst;
}

std::vector<std::pair<std::string, int>> Network::ParseMachines(const std::string& machines) {
  std::vector<std::pair<std::string, int>> out;
  size_t pos = 0;
  while (pos < machines.size()) {
    size_t c = machines.find(',', pos);
    if (c == std::string::npos) c = machines.size();
    std::string tok = machines.substr(pos, c - pos);
    pos = c + 1;
    if (tok.empty()) continue;
    size_t colon = tok.rfind(':');
    LGB_CHECK(colon != std::string::npos, "Malformed machine entry \"%s\" (expected host:port)", tok.c_str());
    out.emplace_back(tok.substr(0, colon), atoi(tok.substr(colon + 1).c_str()));
  }
  return out;
}

int Network::FindRank(const std::vector<std::pair<std::string, int>>& m, int local_listen_port) {
  std::vector<int> cand;
  for (size_t i = 0; i < m.size(); ++i)
    if (m[i].second == local_listen_port) cand.push_back(static_cast<int>(i));
  LGB_CHECK(!cand.empty(), "local_listen_port %d is not in the machine list", local_listen_port);
  if (cand.size() == 1) return cand[0];
  // several hosts use the same port: match the host part against this machine's addresses
  std::vector<std::string> mine = {"127.0.0.1", "localhost"};
  char hn[256];
  if (gethostname(hn, sizeof(hn)) == 0) mine.push_back(hn);
  struct ifaddrs* ifa = nullptr;
  if (getifaddrs(&ifa) == 0) {
    for (auto* p = ifa; p; p = p->ifa_next) {
      if (!p->ifa_addr || p->ifa_addr->sa_family != AF_INET) continue;
      char b[64];
      inet_ntop(AF_INET, &reinterpret_cast<sockaddr_in*>(p->ifa_addr)->sin_addr, b, sizeof(b));
      mine.push_back(b);
    }
    freeifaddrs(ifa);
  }
  for (int i : cand)
    for (auto& a : mine)
      if (m[i].first == a) return i;
  throw std::runtime_error("Cannot determine this process's rank from the machine list");
}

void Network::BroadcastBlob(const std::vector<std::pair<std::string, int>>& m, int rank, int timeout_s, void* blob, size_t bytes) {
  const int world = static_cast<int>(m.size());
  if (world <= 1) return;
  if (rank == 0) {
    int ls = socket(AF_INET, SOCK_STREAM, 0);
    LGB_CHECK(ls >= 0, "Network bootstrap: socket() failed");
    int one = 1;
    setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_addr.s_addr = htonl(INADDR_ANY);
    addr.sin_port = htons(static_cast<uint16_t>(m[0].second));
    if (bind(ls, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
      close(ls);
      throw std::runtime_error(StrFormat("Network bootstrap: cannot bind port %d (%s)", m[0].second, strerror(errno)));
    }
    listen(ls, world);
    timeval tv{timeout_s, 0};
    setsockopt(ls, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    for (int i = 1; i < world; ++i) {
      int fd = accept(ls, nullptr, nullptr);
      if (fd < 0) { close(ls); throw std::runtime_error("Network bootstrap: timed out waiting for the other ranks"); }
      try { SendAll(fd, blob, bytes); } catch (...) { close(fd); close(ls); throw; }
      close(fd);
    }
    close(ls);
  } else {
    addrinfo hints{}, *res = nullptr;
    hints.ai_family = AF_INET;
    hints.ai_socktype = SOCK_STREAM;
    LGB_CHECK(getaddrinfo(m[0].first.c_str(), std::to_string(m[0].second).c_str(), &hints, &res) == 0 && res,
              "Network bootstrap: cannot resolve %s", m[0].first.c_str());
    auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout_s);
    int fd = -1;
    while (true) {
      fd = socket(AF_INET, SOCK_STREAM, 0);
      if (fd >= 0 && connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
      if (fd >= 0) close(fd);
      fd = -1;
      if (std::chrono::steady_clock::now() > deadline) break;
      std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
    freeaddrinfo(res);
    LGB_CHECK(fd >= 0, "Network bootstrap: cannot connect to rank 0 at %s:%d", m[0].first.c_str(), m[0].second);
    timeval tv{timeout_s, 0};
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
    try { RecvAll(fd, blob, bytes); } catch (...) { close(fd); throw; }
    close(fd);
  }
}

void Network::LoadNccl() {
  if (nccl_lib_) return;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (auto n : names) {
    nccl_lib_ = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (nccl_lib_) break;
  }
  LGB_CHECK(nccl_lib_ != nullptr, "num_machines > 1 needs NCCL, but libnccl.so.2 could not be loaded (%s)", dlerror());
#define LOAD(sym, var) var = reinterpret_cast<decltype(var)>(dlsym(nccl_lib_, sym)); LGB_CHECK(var != nullptr, "NCCL symbol %s missing", sym)
  LOAD("ncclGetUniqueId", p_GetUniqueId);
  LOAD("ncclCommInitRank", p_CommInitRank);
  LOAD("ncclCommDestroy", p_CommDestroy);
  LOAD("ncclGetErrorString", p_GetErrorString);
  LOAD("ncclAllReduce", p_AllReduce);
  LOAD("ncclReduceScatter", p_ReduceScatter);
  LOAD("ncclAllGather", p_AllGather);
#undef LOAD
}

namespace {
void SetSockOpts(int fd, int timeout_s) {
  int one = 1;
  setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
  timeval tv{timeout_s, 0};
  setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
  setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof(tv));
}
}  // namespace

// Opens the star: rank 0 listens on its own port of the machine list and accepts one connection per other rank (which
// introduces itself with its rank); the sockets stay open until Free().
void Network::Init(const char* machines, int local_listen_port, int listen_time_out_s, int num_machines) {
  std::lock_guard<std::mutex> lock(mu_);
  if (num_machines <= 1) { rank_ = 0; world_ = 1; return; }
  LGB_CHECK(fds_.empty() && world_ == 1, "LGBM_NetworkInit called twice without LGBM_NetworkFree");
  auto m = ParseMachines(machines ? machines : "");
  LGB_CHECK(static_cast<int>(m.size()) == num_machines, "Machine list has %d entries but num_machines=%d",
            static_cast<int>(m.size()), num_machines);
  const int r = FindRank(m, local_listen_port);
  const int timeout = listen_time_out_s > 0 ? listen_time_out_s : 120;
  std::vector<int> fds;
  try {
    if (r == 0) {
      fds.assign(num_machines, -1);
      int ls = socket(AF_INET, SOCK_STREAM, 0);
      LGB_CHECK(ls >= 0, "Network: socket() failed");
      int one = 1;
      setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
      sockaddr_in addr{};
      addr.sin_family = AF_INET;
      addr.sin_addr.s_addr = htonl(INADDR_ANY);
      addr.sin_port = htons(static_cast<uint16_t>(m[0].second));
      if (bind(ls, reinterpret_cast<sockaddr*>(&addr), sizeof(addr)) != 0) {
        close(ls);
        throw std::runtime_error(StrFormat("Network: cannot bind port %d (%s)", m[0].second, strerror(errno)));
      }
      listen(ls, num_machines);
      timeval tv{timeout, 0};
      setsockopt(ls, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof(tv));
      for (int i = 1; i < num_machines; ++i) {
        int fd = accept(ls, nullptr, nullptr);
        if (fd < 0) { close(ls); throw std::runtime_error("Network: timed out waiting for the other ranks to connect"); }
        SetSockOpts(fd, timeout);
        int32_t who = -1;
        try { RecvAll(fd, &who, sizeof(who)); } catch (...) { close(fd); close(ls); throw; }
        if (who < 1 || who >= num_machines || fds[who] >= 0) { close(fd); close(ls); throw std::runtime_error("Network: unexpected rank in the handshake"); }
        fds[who] = fd;
      }
      close(ls);
    } else {
      addrinfo hints{}, *res = nullptr;
      hints.ai_family = AF_INET;
      hints.ai_socktype = SOCK_STREAM;
      LGB_CHECK(getaddrinfo(m[0].first.c_str(), std::to_string(m[0].second).c_str(), &hints, &res) == 0 && res,
                "Network: cannot resolve %s", m[0].first.c_str());
      auto deadline = std::chrono::steady_clock::now() + std::chrono::seconds(timeout);
      int fd = -1;
      while (true) {
        fd = socket(AF_INET, SOCK_STREAM, 0);
        if (fd >= 0 && connect(fd, res->ai_addr, res->ai_addrlen) == 0) break;
        if (fd >= 0) close(fd);
        fd = -1;
        if (std::chrono::steady_clock::now() > deadline) break;
        std::this_thread::sleep_for(std::chrono::milliseconds(5));
      }
      freeaddrinfo(res);
      LGB_CHECK(fd >= 0, "Network: cannot connect to rank 0 at %s:%d", m[0].first.c_str(), m[0].second);
      SetSockOpts(fd, timeout);
      fds.assign(1, fd);
      const int32_t who = r;
      SendAll(fd, &who, sizeof(who));
    }
  } catch (...) {
    for (int fd : fds) if (fd >= 0) close(fd);
    throw;
  }
  fds_ = fds;
  rank_ = r;
  world_ = num_machines;
  timeout_s_ = timeout;
}

void Network::CloseSockets() {
  for (int fd : fds_) if (fd >= 0) close(fd);
  fds_.clear();
}

void Network::Free() {
  std::lock_guard<std::mutex> lock(mu_);
  if (comm_) {
    cudaSetDevice(device_);
    p_CommDestroy(comm_);
    comm_ = nullptr;
  }
  CloseSockets();
  rank_ = 0;
  world_ = 1;
}

// Every rank sends `bytes` to rank 0, which answers with the concatenation in rank order.
void Network::HostAllGatherBytes(const void* mine, size_t bytes, std::vector<unsigned char>* all) {
  std::lock_guard<std::mutex> lock(mu_);
  all->assign(bytes * world_, 0);
  if (world_ <= 1) { memcpy(all->data(), mine, bytes); return; }
  LGB_CHECK(!fds_.empty(), "Network is not initialised (LGBM_NetworkInit)");
  if (rank_ == 0) {
    memcpy(all->data(), mine, bytes);
    for (int r = 1; r < world_; ++r) RecvAll(fds_[r], all->data() + bytes * r, bytes);
    for (int r = 1; r < world_; ++r) SendAll(fds_[r], all->data(), all->size());
  } else {
    SendAll(fds_[0], mine, bytes);
    RecvAll(fds_[0], all->data(), all->size());
  }
}

void Network::HostAllReduceSum(double* vals, int n) {
  if (world_ <= 1) return;
  std::vector<unsigned char> all;
  HostAllGatherBytes(vals, sizeof(double) * n, &all);
  const double* a = reinterpret_cast<const double*>(all.data());
  for (int i = 0; i < n; ++i) {
    double s = 0.0;
    for (int r = 0; r < world_; ++r) s += a[static_cast<size_t>(r) * n + i];      // rank order on every rank: identical sums
    vals[i] = s;
  }
}

void Network::HostAllGatherInt64(long long mine, std::vector<long long>* all) {
  std::vector<unsigned char> raw;
  HostAllGatherBytes(&mine, sizeof(mine), &raw);
  all->resize(world_);
  memcpy(all->data(), raw.data(), sizeof(long long) * world_);
}

void Network::EnsureNccl(int device) {
  if (world_ <= 1 || comm_ != nullptr) return;
  LoadNccl();
  NcclUniqueId id;
  memset(&id, 0, sizeof(id));
  if (rank_ == 0) LGB_NCCL(p_GetUniqueId(&id));
  std::vector<unsigned char> all;
  HostAllGatherBytes(&id, sizeof(id), &all);
  memcpy(&id, all.data(), sizeof(id));      // rank 0's
  device_ = device;
  LGB_CUDA(cudaSetDevice(device));
  LGB_NCCL(p_CommInitRank(&comm_, world_, id, rank_));
}

void Network::AllReduceSumInt64(long long* buf, size_t n, cudaStream_t stream) {
  LGB_NCCL(p_AllReduce(buf, buf, n, kNcclInt64, kNcclSum, comm_, stream));
}
void Network::AllReduceMaxU32(unsigned int* buf, size_t n, cudaStream_t stream) {
  LGB_NCCL(p_AllReduce(buf, buf, n, kNcclUint32, kNcclMax, comm_, stream));
}
void Network::ReduceScatterSumInt64(const long long* send, long long* recv, size_t recv_count, cudaStream_t stream) {
  LGB_NCCL(p_ReduceScatter(send, recv, recv_count, kNcclInt64, kNcclSum, comm_, stream));
}
void Network::AllGatherBytes(const void* send, void* recv, size_t bytes_per_rank, cudaStream_t stream) {
  LGB_NCCL(p_AllGather(send, recv, bytes_per_rank, kNcclInt8, comm_, stream));
}

bool Network::ExchangePeerPointers(int device, void* local, std::vector<void*>* peers) {
  peers->clear();
  if (world_ <= 1) { peers->push_back(local); return true; }
  LGB_CUDA(cudaSetDevice(device));
  struct Msg { cudaIpcMemHandle_t h; int ok; int device; } mine;
  memset(&mine, 0, sizeof(mine));
  mine.ok = cudaIpcGetMemHandle(&mine.h, local) == cudaSuccess;
  if (!mine.ok) cudaGetLastError();
  mine.device = device;
  std::vector<unsigned char> all;
  HostAllGatherBytes(&mine, sizeof(mine), &all);
  const Msg* msgs = reinterpret_cast<const Msg*>(all.data());
  bool ok = true;
  for (int r = 0; r < world_; ++r) ok = ok && msgs[r].ok != 0;
  peers->assign(world_, nullptr);
  if (ok) {
    for (int r = 0; r < world_; ++r) {
      if (r == rank_) { (*peers)[r] = local; continue; }
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, msgs[r].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
      (*peers)[r] = p;
    }
  }
  // everybody must agree, or some ranks would wait for flags nobody writes
  std::vector<long long> agree;
  HostAllGatherInt64(ok ? 1 : 0, &agree);
  bool all_ok = true;
  for (long long v : agree) all_ok = all_ok && v != 0;
  if (!all_ok) { ClosePeerPointers(*peers); peers->clear(); }
  return all_ok;
}

void Network::ClosePeerPointers(const std::vector<void*>& peers) {
  for (int r = 0; r < static_cast<int>(peers.size()); ++r)
    if (r != rank_ && peers[r] != nullptr) cudaIpcCloseMemHandle(peers[r]);
}

}  // namespace lgb
