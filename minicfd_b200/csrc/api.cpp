// api.cpp -- the C-ABI of include/minicfd_b200.h: solver object, device buffers, stage
// orchestration (halo -> face fluxes -> fused cell update) and the RK loop.
//
// Orchestration follows spatial::FiniteElement::GetResidualColumn (spatial/fem.hpp:173-183):
//   ShareGhostCellCoeffs -> volume + local faces + boundaries -> UpdateGhostCellCoeffs -> ghost faces
// with the MPI Isend/Irecv (mesh/part.hpp:1258-1302) replaced by NCCL send/recv on a second
// stream, overlapped with the local-face and boundary-face kernels.
#include <dlfcn.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <memory>
#include <string>
#include <vector>

#include "solver.hpp"

using namespace mcfd;

namespace {

// ---- NCCL through dlopen: reuse the copy torch already loaded when there is one ----
struct UniqueId { char internal[128]; };  // ncclUniqueId
struct NcclApi {
  void *handle = nullptr;
  int (*GetUniqueId)(void *) = nullptr;
  int (*CommInitRank)(void **, int, UniqueId /* by value */, int) = nullptr;
  int (*CommDestroy)(void *) = nullptr;
  int (*Send)(const void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*Recv)(void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(int) = nullptr;
};
constexpr int kNcclFloat64 = 8;  // ncclDouble / ncclFloat64 in nccl.h
constexpr int kNcclInt8 = 0, kNcclInt32 = 2, kNcclMin = 3;  // ncclChar, ncclInt, ncclMin in nccl.h

NcclApi *LoadNccl(std::string *error) {
  static NcclApi api;
  if (api.handle) return &api;
  const char *names[] = { "libnccl.so.2", "libnccl.so" };
  void *h = nullptr;
  for (const char *nm : names) { h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (h) break; }
  if (!h) for (const char *nm : names) { h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
  if (!h) { *error = std::string("cannot load NCCL: ") + dlerror(); return nullptr; }
  api.handle = h;
#define MCFD_SYM(field, name) \
  *reinterpret_cast<void **>(&api.field) = dlsym(h, name); \
  if (!api.field) { *error = std::string("NCCL symbol missing: ") + name; api.handle = nullptr; return nullptr; }
  MCFD_SYM(GetUniqueId, "ncclGetUniqueId")
  MCFD_SYM(CommInitRank, "ncclCommInitRank")
  MCFD_SYM(CommDestroy, "ncclCommDestroy")
  MCFD_SYM(Send, "ncclSend")
  MCFD_SYM(Recv, "ncclRecv")
  MCFD_SYM(AllReduce, "ncclAllReduce")
  MCFD_SYM(GroupStart, "ncclGroupStart")
  MCFD_SYM(GroupEnd, "ncclGroupEnd")
  MCFD_SYM(GetErrorString, "ncclGetErrorString")
#undef MCFD_SYM
  return &api;
}

std::string g_create_error;

template <class T>
cudaError_t Upload(const std::vector<T> &host, T **dev) {
  *dev = nullptr;
  if (host.empty()) return cudaSuccess;
  cudaError_t e = cudaMalloc(reinterpret_cast<void **>(dev), host.size() * sizeof(T));
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*dev, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice);
}

}  // namespace

struct mcfd_solver {
  mcfd_config cfg{};
  Basis basis{};
  HostTables host;
  DeviceTables dev{};
  int fields = 0;            // K * N
  int64_t column = 0;        // fields * n_local
  int64_t column_all = 0;    // fields * (n_local + n_ghost)
  // device memory
  double *d_metrics = nullptr, *d_face_frame = nullptr, *d_bface_xyz = nullptr, *d_corner = nullptr;
  int32_t *d_cell_face = nullptr, *d_face_holder = nullptr, *d_face_sharer = nullptr, *d_bface_group = nullptr,
          *d_bface_local = nullptr, *d_send_cells = nullptr, *d_pt_holder = nullptr, *d_pt_sharer = nullptr;
  uint8_t *d_cell_code = nullptr;
  uint16_t *d_face_code = nullptr;
  BcGroup *d_groups = nullptr;
  double *d_u[3] = { nullptr, nullptr, nullptr };
  double *d_fstar = nullptr, *d_send = nullptr, *h_pinned = nullptr;
  int *d_ticket = nullptr;   // 8 ints of device counters; [4] is pack_to_peers_kernel's blocks-done counter
  double *d_star_u = nullptr;
  int32_t *d_class_cells = nullptr, *d_class_ptr = nullptr;   // shape-class tables of a repetitive modal mesh
  double *d_mm_A1 = nullptr, *d_mm_A2 = nullptr, *d_mm_A3 = nullptr, *d_cls_S = nullptr, *d_cls_P = nullptr, *d_cls_W = nullptr;
  int32_t *d_cell_class = nullptr;  // Exact::star_u per interior / ghost face point (DeviceTables::star_u)
  // modal DG + limiter
  double *d_m_center = nullptr, *d_m_S = nullptr, *d_m_volume = nullptr, *d_m_qxyzw = nullptr, *d_f_xyzw = nullptr;
  double *d_fresh = nullptr;
  int64_t lazy_rebuilds = 0;  // weno::Lazy rebuilds every cell of every pass: counted on the host
  const double *tay_of = nullptr;  // the column whose LOCAL cells' Taylor form d_tay currently holds (nullptr: stale)
  double *d_tay = nullptr;   // modal DG: Taylor form S^t coeff of every local and ghost cell (face kernel input)
  int32_t *d_adj_ptr = nullptr, *d_adj_cell = nullptr, *d_adj_face = nullptr, *d_flags = nullptr;
  int cur = 0;
  std::vector<BcGroup> groups;
  std::vector<double *> group_values;
  double t_curr = 0.0;
  cudaStream_t stream = nullptr, comm_stream = nullptr;
  cudaEvent_t ev_packed = nullptr, ev_halo = nullptr, ev_begin = nullptr, ev_end = nullptr;
  NcclApi *nccl = nullptr;
  void *comm = nullptr;
  // Halo exchange over peer-mapped memory (SetupPeerLinks): the neighbours' columns and arrival counters as seen
  // from this GPU.  Used by mcfd_rk_update when no limiter is configured; everything else goes through NCCL.
  struct PeerHalo {
    bool ready = false;
    std::vector<void *> opened;                  // cudaIpcOpenMemHandle results (closed at destroy)
    double **d_peer_u = nullptr;                 // [3][n_neighbours] neighbour's d_u[b] + its ghost offset for this rank
    unsigned long long **d_peer_flag = nullptr;  // [n_neighbours]    neighbour's arrival counter for this rank
    unsigned long long *d_flags = nullptr;       // [n_neighbours]    this rank's arrival counters (written by neighbours)
    int32_t *d_send_nbr = nullptr;               // [n_send_cells]    neighbour index of each send cell
    int64_t *d_send_dst = nullptr;               // [n_send_cells]    offset (doubles) inside that neighbour's ghost block
    int *d_error = nullptr;
    int *d_sync = nullptr;                       // operand of the all-reduce that opens every mcfd_rk_update
    unsigned long long seq = 0;                  // exchanges so far; identical on every rank (same call sequence)
  } peer;
  bool profiling = false;
  mcfd_timing timing{};
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> face_events, cell_events;
  std::string error;
};

namespace {

int Fail(mcfd_solver *s, int status, const std::string &msg) {
  if (s) s->error = msg; else g_create_error = msg;
  return status;
}

#define MCFD_CUDA(s, call) \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return Fail((s), MCFD_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

int SyncGroups(mcfd_solver *s) {
  if (s->groups.empty()) return MCFD_OK;
  MCFD_CUDA(s, cudaMemcpyAsync(s->d_groups, s->groups.data(), s->groups.size() * sizeof(BcGroup), cudaMemcpyHostToDevice, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  return MCFD_OK;
}

struct ScopedTimer {
  mcfd_solver *s; bool face;
  cudaEvent_t a = nullptr, b = nullptr;
  ScopedTimer(mcfd_solver *s_, bool face_) : s(s_), face(face_) {
    if (!s->profiling) return;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, s->stream);
  }
  ~ScopedTimer() {
    if (!s->profiling) return;
    cudaEventRecord(b, s->stream);
    (face ? s->face_events : s->cell_events).emplace_back(a, b);
  }
};

// ---- halo exchange over peer memory -------------------------------------------------------------------
// What each rank tells every neighbour at attach time (sent through NCCL, once): where its three column buffers
// and its arrival counters live (CUDA IPC handles + the offset of the pointer inside the IPC allocation; raw
// pointers when both parts live in one process), and where that neighbour's cells go in its ghost tail.
struct PeerHello {
  cudaIpcMemHandle_t u[3], flags;
  uint64_t u_offset[3], flags_offset;  // pointer - base of the allocation the handle names
  uint64_t raw_u[3], raw_flags;        // same-process case
  int64_t ghost_offset;                // doubles: (n_local + recv_ptr[n]) * fields
  int32_t slot, pid, device, pad;
  uint64_t process_nonce;              // drawn once per process: pids repeat across hosts and PID namespaces
};

// Identifies THIS process among the ranks of a communicator: a pid alone is not enough (rank processes on different
// nodes, or in containers with their own PID namespace, routinely share one).
uint64_t ProcessNonce() {
  static const uint64_t nonce = [] {
    uint64_t v = 0;
    if (FILE *f = std::fopen("/dev/urandom", "rb")) { if (std::fread(&v, sizeof(v), 1, f) != 1) v = 0; std::fclose(f); }
    if (!v) v = (uint64_t(getpid()) << 32) ^ uint64_t(reinterpret_cast<uintptr_t>(&v)) ^ uint64_t(std::clock());
    return v | 1;
  }();
  return nonce;
}

cudaError_t AllocationOffset(const void *ptr, uint64_t *offset) {
  // cudaIpcGetMemHandle names the whole allocation a pointer lies in; small cudaMalloc's may share one
  typedef int (*GetRange)(unsigned long long *, size_t *, unsigned long long);
  static GetRange get_range = nullptr;
  if (!get_range) {
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &q);
    if (e != cudaSuccess) return e;
    if (!fn) return cudaErrorNotSupported;
    get_range = reinterpret_cast<GetRange>(fn);
  }
  unsigned long long base = 0;
  size_t size = 0;
  if (get_range(&base, &size, reinterpret_cast<unsigned long long>(ptr)) != 0) return cudaErrorInvalidValue;
  *offset = reinterpret_cast<unsigned long long>(ptr) - base;
  return cudaSuccess;
}

// Collective over the communicator (called from mcfd_attach_nccl on every rank).  Leaves s->peer.ready set only
// when EVERY rank could map all of its neighbours; otherwise all ranks keep the NCCL transport.
int SetupPeerLinks(mcfd_solver *s) {
  const HostTables &h = s->host;
  mcfd_solver::PeerHalo &p = s->peer;
  const char *env = std::getenv("MCFD_HALO_P2P");
  const bool wanted = !(env && std::atoi(env) == 0) && s->cfg.limiter == MCFD_LIMITER_NONE;
  const int nn = int(h.nbr_rank.size());
  int ok = wanted ? 1 : 0;
  std::vector<PeerHello> mine(nn), theirs(nn);
  PeerHello *d_mine = nullptr, *d_theirs = nullptr;
  int *d_ok = nullptr;
  cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&d_ok), sizeof(int));
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  if (ok && nn > 0) {
    e = cudaMalloc(reinterpret_cast<void **>(&p.d_flags), nn * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(p.d_flags, 0, nn * sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d_mine), nn * sizeof(PeerHello));
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d_theirs), nn * sizeof(PeerHello));
    for (int n = 0; n < nn && e == cudaSuccess; ++n) {
      PeerHello &m = mine[n];
      std::memset(&m, 0, sizeof(m));
      for (int b = 0; b < 3 && e == cudaSuccess; ++b) {
        e = cudaIpcGetMemHandle(&m.u[b], s->d_u[b]);
        if (e == cudaSuccess) e = AllocationOffset(s->d_u[b], &m.u_offset[b]);
        m.raw_u[b] = reinterpret_cast<uint64_t>(s->d_u[b]);
      }
      if (e == cudaSuccess) e = cudaIpcGetMemHandle(&m.flags, p.d_flags);
      if (e == cudaSuccess) e = AllocationOffset(p.d_flags, &m.flags_offset);
      m.raw_flags = reinterpret_cast<uint64_t>(p.d_flags);
      m.ghost_offset = (h.n_local + h.recv_ptr[n]) * int64_t(s->fields);
      m.slot = n; m.pid = int32_t(getpid()); m.device = s->cfg.device; m.process_nonce = ProcessNonce();
    }
    if (e == cudaSuccess) e = cudaMemcpy(d_mine, mine.data(), nn * sizeof(PeerHello), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { ok = 0; cudaGetLastError(); }
  }
  // the hello round happens only if every rank got this far (an allreduce everyone takes part in)
  auto all_agree = [&](int *flag) -> int {
    cudaMemcpy(d_ok, flag, sizeof(int), cudaMemcpyHostToDevice);
    int rc = s->nccl->AllReduce(d_ok, d_ok, 1, kNcclInt32, kNcclMin, s->comm, s->stream);
    if (rc != 0) return rc;
    if (cudaStreamSynchronize(s->stream) != cudaSuccess) return -1;
    cudaMemcpy(flag, d_ok, sizeof(int), cudaMemcpyDeviceToHost);
    return 0;
  };
  int rc = all_agree(&ok);
  if (rc == 0 && ok && nn > 0) {
    rc = s->nccl->GroupStart();
    for (int n = 0; n < nn && rc == 0; ++n) {
      rc = s->nccl->Send(d_mine + n, sizeof(PeerHello), kNcclInt8, h.nbr_rank[n], s->comm, s->stream);
      if (rc == 0) rc = s->nccl->Recv(d_theirs + n, sizeof(PeerHello), kNcclInt8, h.nbr_rank[n], s->comm, s->stream);
    }
    int rc_end = s->nccl->GroupEnd();
    if (rc == 0) rc = rc_end;
    if (rc == 0 && cudaStreamSynchronize(s->stream) != cudaSuccess) rc = -1;
    if (rc == 0) cudaMemcpy(theirs.data(), d_theirs, nn * sizeof(PeerHello), cudaMemcpyDeviceToHost);
  }
  int mapped = (rc == 0 && ok) ? 1 : 0;
  std::vector<double *> peer_u(size_t(3) * nn, nullptr);
  std::vector<unsigned long long *> peer_flag(nn, nullptr);
  for (int n = 0; n < nn && mapped; ++n) {
    const PeerHello &t = theirs[n];
    char *base_u[3] = { nullptr, nullptr, nullptr }, *base_f = nullptr;
    if (t.process_nonce == ProcessNonce() && t.pid == int32_t(getpid())) {  // both parts in this process: plain peer access
      if (t.device != s->cfg.device) {
        cudaError_t pe = cudaDeviceEnablePeerAccess(t.device, 0);
        if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) mapped = 0;
        cudaGetLastError();
      }
      for (int b = 0; b < 3; ++b) base_u[b] = reinterpret_cast<char *>(t.raw_u[b]) - t.u_offset[b];
      base_f = reinterpret_cast<char *>(t.raw_flags) - t.flags_offset;
    } else {
      // the three columns (or the counters) may share one IPC allocation; a handle can be opened once per process
      auto open = [&](const cudaIpcMemHandle_t &hd, char **out) {
        void *ptr = nullptr;
        if (cudaIpcOpenMemHandle(&ptr, hd, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return false; }
        p.opened.push_back(ptr);
        *out = static_cast<char *>(ptr);
        return true;
      };
      for (int b = 0; b < 3 && mapped; ++b) {
        for (int a = 0; a < b; ++a)
          if (std::memcmp(&t.u[a], &t.u[b], sizeof(cudaIpcMemHandle_t)) == 0) base_u[b] = base_u[a];
        if (!base_u[b] && !open(t.u[b], &base_u[b])) mapped = 0;
      }
      for (int a = 0; a < 3 && mapped; ++a)
        if (std::memcmp(&t.u[a], &t.flags, sizeof(cudaIpcMemHandle_t)) == 0) base_f = base_u[a];
      if (mapped && !base_f && !open(t.flags, &base_f)) mapped = 0;
    }
    if (!mapped) break;
    for (int b = 0; b < 3; ++b)
      peer_u[size_t(b) * nn + n] = reinterpret_cast<double *>(base_u[b] + t.u_offset[b]) + t.ghost_offset;
    peer_flag[n] = reinterpret_cast<unsigned long long *>(base_f + t.flags_offset) + t.slot;
  }
  if (rc == 0) rc = all_agree(&mapped);
  if (rc == 0 && mapped && nn > 0) {
    std::vector<int32_t> send_nbr(h.send_cells.size());
    std::vector<int64_t> send_dst(h.send_cells.size());
    for (int n = 0; n < nn; ++n)
      for (int64_t i = h.send_ptr[n]; i < h.send_ptr[n + 1]; ++i) { send_nbr[i] = n; send_dst[i] = (i - h.send_ptr[n]) * int64_t(s->fields); }
    e = Upload(peer_u, &p.d_peer_u);
    if (e == cudaSuccess) e = Upload(peer_flag, &p.d_peer_flag);
    if (e == cudaSuccess) e = Upload(send_nbr, &p.d_send_nbr);
    if (e == cudaSuccess) e = Upload(send_dst, &p.d_send_dst);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&p.d_error), sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(p.d_error, 0, sizeof(int));
    if (e != cudaSuccess) { cudaFree(d_ok); cudaFree(d_mine); cudaFree(d_theirs); return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e)); }
  }
  if (rc == 0 && mapped) {
    e = cudaMalloc(reinterpret_cast<void **>(&p.d_sync), sizeof(int));
    if (e == cudaSuccess) e = cudaMemset(p.d_sync, 0, sizeof(int));
    if (e != cudaSuccess) { cudaFree(d_ok); cudaFree(d_mine); cudaFree(d_theirs); return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e)); }
  }
  p.ready = rc == 0 && mapped;  // ranks without neighbours carry the flag too: it selects the code path, nothing more
  cudaFree(d_ok); cudaFree(d_mine); cudaFree(d_theirs);
  if (rc != 0) return Fail(s, MCFD_ERR_NCCL, "peer-memory handshake failed");
  return MCFD_OK;
}

// Part::ShareGhostCellCoeffs over peer memory: gather + remote store + arrival counter, on the comm stream
int PostHaloToPeers(mcfd_solver *s, double *u) {
  const HostTables &h = s->host;
  mcfd_solver::PeerHalo &p = s->peer;
  ++p.seq;
  if (h.nbr_rank.empty()) return MCFD_OK;
  int b = 0;
  while (b < 3 && s->d_u[b] != u) ++b;
  if (b == 3) return Fail(s, MCFD_ERR_ARG, "halo exchange of a foreign column");
  const int nn = int(h.nbr_rank.size());
  // MCFD_PACK_INLINE=1 (tuning): the pack kernel goes on the MAIN stream, ahead of the local faces
  static const bool inline_pack = std::getenv("MCFD_PACK_INLINE") && std::atoi(std::getenv("MCFD_PACK_INLINE")) == 1;
  cudaStream_t where = inline_pack ? s->stream : s->comm_stream;
  if (!inline_pack) {
    MCFD_CUDA(s, cudaEventRecord(s->ev_packed, s->stream));
    MCFD_CUDA(s, cudaStreamWaitEvent(s->comm_stream, s->ev_packed, 0));
  }
  MCFD_CUDA(s, LaunchPackToPeers(u, s->d_send_cells, p.d_send_nbr, p.d_send_dst, p.d_peer_u + size_t(b) * nn, h.send_ptr.back(),
                                 s->fields, p.d_peer_flag, nn, p.seq, reinterpret_cast<unsigned int *>(s->d_ticket + 4), where));
  s->timing.n_other_launches += 1;
  MCFD_CUDA(s, cudaEventRecord(s->ev_halo, where));
  return MCFD_OK;
}

// Part::ShareGhostCellCoeffs (part.hpp:1258-1284) on the comm stream
int PostHalo(mcfd_solver *s, double *u) {
  const HostTables &h = s->host;
  if (!s->comm || h.nbr_rank.empty()) return MCFD_OK;
  cudaError_t e = LaunchPackCells(u, s->d_send_cells, h.send_ptr.back(), s->fields, s->d_send, s->stream);
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  ++s->timing.n_other_launches;
  MCFD_CUDA(s, cudaEventRecord(s->ev_packed, s->stream));
  MCFD_CUDA(s, cudaStreamWaitEvent(s->comm_stream, s->ev_packed, 0));
  int rc = s->nccl->GroupStart();
  for (size_t n = 0; n < h.nbr_rank.size() && rc == 0; ++n) {
    size_t n_send = size_t(h.send_ptr[n + 1] - h.send_ptr[n]) * s->fields;
    size_t n_recv = size_t(h.recv_ptr[n + 1] - h.recv_ptr[n]) * s->fields;
    rc = s->nccl->Send(s->d_send + h.send_ptr[n] * s->fields, n_send, kNcclFloat64, h.nbr_rank[n], s->comm, s->comm_stream);
    if (rc == 0)
      rc = s->nccl->Recv(u + (h.n_local + h.recv_ptr[n]) * s->fields, n_recv, kNcclFloat64, h.nbr_rank[n], s->comm, s->comm_stream);
  }
  int rc_end = s->nccl->GroupEnd();
  if (rc != 0 || rc_end != 0) return Fail(s, MCFD_ERR_NCCL, s->nccl->GetErrorString(rc ? rc : rc_end));
  MCFD_CUDA(s, cudaEventRecord(s->ev_halo, s->comm_stream));
  return MCFD_OK;
}

// the scheme's face-flux and fused cell-update kernels
cudaError_t FaceFlux(mcfd_solver *s, double t, const double *u, int64_t f0, int64_t f1) {
  if (s->cfg.scheme == MCFD_SCHEME_DG_MODAL)
    return LaunchModalFaceFlux(s->dev, s->cfg.degree, s->cfg.riemann, s->cfg.gamma, t, s->d_tay, s->d_fstar, f0, f1, s->stream);
  return LaunchFaceFlux(s->dev, s->basis, s->cfg.scheme, s->cfg.riemann, s->cfg.gamma, t, u, s->d_fstar, f0, f1, s->stream);
}
cudaError_t CellUpdate(mcfd_solver *s, const double *u, const double *u0, double *out, StageOp op, int64_t cell_begin = 0,
                       int64_t cell_end = -1) {
  if (cell_end < 0) cell_end = s->host.n_local;
  if (s->cfg.scheme == MCFD_SCHEME_DG_MODAL) {
    // the cell kernel for p <= 2 leaves the Taylor form of the new column behind (not of a bare residual, kind 4)
    double *tay_out = (s->cfg.degree <= 2 && op.kind != 4) ? s->d_tay : nullptr;
    s->tay_of = tay_out ? out : (s->tay_of == out ? nullptr : s->tay_of);
    return LaunchModalCellUpdate(s->dev, s->cfg.degree, s->cfg.gamma, u, s->d_fstar, u0, out, tay_out, op, s->stream);
  }
  return LaunchCellUpdate(s->dev, s->basis, s->cfg.scheme, s->cfg.gamma, u, s->d_fstar, u0, out, op, cell_begin, cell_end, s->stream);
}

// One residual evaluation fused with a stage update: out = op(u, R(u, t), u0)
// `over_peers`: the halo goes through peer memory (mcfd_rk_update only: consecutive stages use different columns,
// which is what makes writing into a neighbour's ghost tail safe without an acknowledgement)
int Stage(mcfd_solver *s, double *u, const double *u0, double *out, StageOp op, double t, bool over_peers = false) {
  const HostTables &h = s->host;
  over_peers = over_peers && s->peer.ready && s->comm;
  int rc = over_peers ? PostHaloToPeers(s, u) : PostHalo(s, u);
  if (rc) return rc;
  cudaError_t e;
  const int64_t f_bnd = h.n_local_faces + h.n_ghost_faces;
  const bool modal = s->cfg.scheme == MCFD_SCHEME_DG_MODAL;
  {
    ScopedTimer timer(s, true);
    e = cudaSuccess;
    if (modal && s->tay_of != u) {
      e = LaunchModalTaylor(s->dev, s->cfg.degree, u, s->d_tay, 0, h.n_local, s->stream);
      s->tay_of = u;
      ++s->timing.n_other_launches;
    }
    if (!modal) {  // local and boundary faces in one launch: neither waits for the halo
      if (e == cudaSuccess)
        e = LaunchFaceFluxLocalAndBoundary(s->dev, s->basis, s->cfg.scheme, s->cfg.riemann, s->cfg.gamma, t, u, s->d_fstar, s->stream);
      s->timing.n_face_launches += (h.n_local_faces + h.n_boundary_faces) > 0;
    } else {
      if (e == cudaSuccess) e = FaceFlux(s, t, u, 0, h.n_local_faces);
      if (e == cudaSuccess) e = FaceFlux(s, t, u, f_bnd, h.n_faces);
      s->timing.n_face_launches += (h.n_local_faces > 0) + (h.n_boundary_faces > 0);
    }
  }
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  // Inner cells touch no ghost face (part.hpp:818-835): their update does not wait for the halo, which then has the
  // whole of it to arrive in (fem.hpp:173-183 overlaps the volume term and the local faces; here the local faces and
  // the inner cells).  The split sits on an even cell so that both ranges start on a 16-byte boundary (TMA tiles).
  int64_t split = 0;
  // MCFD_CELL_SPLIT=1 (tuning) turns it on.  Measured on 8 B200s, 64^3 FR p = 2 (32^3 cells per GPU, profiles/r02_multi/r02r):
  // 0.376 ms per step with the split against 0.348-0.364 without -- the second launch of the cell kernel costs a tail of
  // its own (0.066 against 0.055 ms per stage) and the halo was already hidden behind the local faces; off by default.
  static const bool split_on = std::getenv("MCFD_CELL_SPLIT") && std::atoi(std::getenv("MCFD_CELL_SPLIT")) == 1;
  if (!modal && split_on && h.n_ghost_faces > 0 && h.n_inner >= 2) {
    split = h.n_inner & ~int64_t(1);
    ScopedTimer timer(s, false);
    e = CellUpdate(s, u, u0, out, op, 0, split);
    ++s->timing.n_cell_launches;
    if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  }
  if (h.n_ghost_faces > 0) {
    if (s->comm) MCFD_CUDA(s, cudaStreamWaitEvent(s->stream, s->ev_halo, 0));  // UpdateGhostCellCoeffs: my own stores are out
    // ... and the neighbours' must have arrived: the ghost-face kernel polls the arrival counters itself (nodal schemes);
    // the modal path keeps the separate wait in front of its Taylor kernel
    PeerWait wait{ nullptr, 0, 0, 0, nullptr };
    if (over_peers) {
      wait = PeerWait{ s->peer.d_flags, int(h.nbr_rank.size()), s->peer.seq, 20LL * 1000 * 1000 * 1000, s->peer.d_error };
      if (modal) {
        MCFD_CUDA(s, LaunchWaitPeers(s->peer.d_flags, int(h.nbr_rank.size()), s->peer.seq, s->peer.d_error, s->stream));
        ++s->timing.n_other_launches;
      }
    }
    ScopedTimer timer(s, true);
    e = cudaSuccess;
    if (modal) {
      e = LaunchModalTaylor(s->dev, s->cfg.degree, u, s->d_tay, h.n_local, h.n_local + h.n_ghost, s->stream);
      ++s->timing.n_other_launches;
      if (e == cudaSuccess) e = FaceFlux(s, t, u, h.n_local_faces, f_bnd);
    } else {
      e = LaunchFaceFlux(s->dev, s->basis, s->cfg.scheme, s->cfg.riemann, s->cfg.gamma, t, u, s->d_fstar, h.n_local_faces, f_bnd,
                         s->stream, &wait);
    }
    if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
    ++s->timing.n_face_launches;
  }
  {
    ScopedTimer timer(s, false);
    e = CellUpdate(s, u, u0, out, op, split);
    ++s->timing.n_cell_launches;
  }
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  return MCFD_OK;
}

// Part::Reconstruct(limiter) (mesh/part.hpp:1304-1315), what WithLimiterAndSource::SetSolutionColumn runs after
// every column it is given (spatial/dg/general.hpp:247-250): post the halo, limit the inner cells, wait for the
// halo, limit the inter cells.  Each set is rebuilt from the data as it was before the set was touched.
int Limit(mcfd_solver *s, double *u) {
  if (s->cfg.limiter == MCFD_LIMITER_NONE) return MCFD_OK;
  const HostTables &h = s->host;
  int rc = PostHalo(s, u);
  if (rc) return rc;
  if (s->cfg.limiter == MCFD_LIMITER_DUMMY) {  // weno.hpp:266-281: every cell "troubled", rebuilt as a copy of itself
    if (s->comm && !h.nbr_rank.empty()) MCFD_CUDA(s, cudaStreamWaitEvent(s->stream, s->ev_halo, 0));
    s->lazy_rebuilds += h.n_local;
    return MCFD_OK;
  }
  const LimiterOp lim{ s->cfg.limiter, s->cfg.weno_w0, s->cfg.weno_eps };
  cudaError_t e = cudaSuccess;
  if (s->tay_of != u) {  // the detector reads the Taylor form of the column being limited
    e = LaunchModalTaylor(s->dev, s->cfg.degree, u, s->d_tay, 0, h.n_local, s->stream);
    if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
    s->tay_of = u;
    ++s->timing.n_other_launches;
  }
  e = LaunchModalLimit(s->dev, s->cfg.degree, lim, s->cfg.gamma, u, s->d_tay, s->d_fresh, s->d_flags, 0, h.n_inner, s->stream);
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  if (s->comm && !h.nbr_rank.empty()) MCFD_CUDA(s, cudaStreamWaitEvent(s->stream, s->ev_halo, 0));
  if (h.n_ghost > 0 && h.n_local > h.n_inner) {  // inter cells look at ghost neighbours
    e = LaunchModalTaylor(s->dev, s->cfg.degree, u, s->d_tay, h.n_local, h.n_local + h.n_ghost, s->stream);
    if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
    ++s->timing.n_other_launches;
  }
  e = LaunchModalLimit(s->dev, s->cfg.degree, lim, s->cfg.gamma, u, s->d_tay, s->d_fresh, s->d_flags, h.n_inner, h.n_local, s->stream);
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  s->timing.n_other_launches += (2 + (s->cfg.limiter == MCFD_LIMITER_EIGEN)) * ((h.n_inner > 0) + (h.n_local > h.n_inner));
  if (s->cfg.limiter == MCFD_LIMITER_LAZY) s->lazy_rebuilds += h.n_local;
  return MCFD_OK;
}

void CollectKernelTimes(mcfd_solver *s) {
  auto sum = [](std::vector<std::pair<cudaEvent_t, cudaEvent_t>> &v) {
    double ms = 0;
    for (auto &p : v) { float x = 0; cudaEventElapsedTime(&x, p.first, p.second); ms += x; cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    v.clear();
    return ms;
  };
  s->timing.ms_face = sum(s->face_events);
  s->timing.ms_cell = sum(s->cell_events);
}

}  // namespace

extern "C" {

const char *mcfd_last_error(const mcfd_solver *s) { return s ? s->error.c_str() : g_create_error.c_str(); }

int mcfd_create(const mcfd_config *cfg, const mcfd_mesh *mesh, mcfd_solver **out) {
  if (!cfg || !mesh || !out) return Fail(nullptr, MCFD_ERR_ARG, "null argument");
  *out = nullptr;
  if (cfg->scheme != MCFD_SCHEME_FR && cfg->scheme != MCFD_SCHEME_DGSEM && cfg->scheme != MCFD_SCHEME_DG_MODAL)
    return Fail(nullptr, MCFD_ERR_ARG, "scheme must be MCFD_SCHEME_FR, MCFD_SCHEME_DGSEM or MCFD_SCHEME_DG_MODAL");
  if (cfg->limiter < MCFD_LIMITER_NONE || cfg->limiter > MCFD_LIMITER_DUMMY) return Fail(nullptr, MCFD_ERR_ARG, "limiter kind");
  // limiter/weno.hpp:22-24,40,133 : the limiters need Projection::Wrapper / Taylor / average(), i.e. the modal basis
  if (cfg->limiter != MCFD_LIMITER_NONE && cfg->scheme != MCFD_SCHEME_DG_MODAL)
    return Fail(nullptr, MCFD_ERR_UNSUPPORTED, "limiters exist for the modal DG scheme only (as in the reference)");
  if (cfg->riemann < 0 || cfg->riemann > 3) return Fail(nullptr, MCFD_ERR_ARG, "riemann kind");
  if (cfg->rk_order < 1 || cfg->rk_order > 3) return Fail(nullptr, MCFD_ERR_ARG, "rk_order must be 1, 2 or 3");
  if (!(cfg->gamma > 1.0)) return Fail(nullptr, MCFD_ERR_ARG, "gamma must exceed 1");
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return Fail(nullptr, MCFD_ERR_CUDA, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= n_dev) return Fail(nullptr, MCFD_ERR_ARG, "device ordinal out of range");
  auto s = std::make_unique<mcfd_solver>();
  s->cfg = *cfg;
  int rc = BuildBasis(cfg->degree, cfg->scheme, &s->basis);
  if (rc) return Fail(nullptr, rc, "degree must be 1, 2 or 3");
  std::string err;
  rc = BuildTables(*cfg, *mesh, s->basis, &s->host, &err);
  if (rc) return Fail(nullptr, rc, err);
  HostTables &h = s->host;
  s->fields = MCFD_K * h.n;
  s->column = int64_t(s->fields) * h.n_local;
  s->column_all = int64_t(s->fields) * (h.n_local + h.n_ghost);

#define MCFD_CREATE_CUDA(call) \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { std::string m_ = std::string(#call) + ": " + cudaGetErrorString(e_); mcfd_destroy(s.release()); return Fail(nullptr, MCFD_ERR_CUDA, m_); } } while (0)
  MCFD_CREATE_CUDA(cudaSetDevice(cfg->device));
  MCFD_CREATE_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  {  // the halo kernels are small and on the critical path of the NEIGHBOURS: first in line whenever an SM slot frees up
    int least = 0, greatest = 0;
    MCFD_CREATE_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    MCFD_CREATE_CUDA(cudaStreamCreateWithPriority(&s->comm_stream, cudaStreamNonBlocking, greatest));
  }
  MCFD_CREATE_CUDA(cudaEventCreateWithFlags(&s->ev_packed, cudaEventDisableTiming));
  MCFD_CREATE_CUDA(cudaEventCreateWithFlags(&s->ev_halo, cudaEventDisableTiming));
  MCFD_CREATE_CUDA(cudaEventCreate(&s->ev_begin));
  MCFD_CREATE_CUDA(cudaEventCreate(&s->ev_end));
  MCFD_CREATE_CUDA(Upload(h.metrics, &s->d_metrics));
  MCFD_CREATE_CUDA(Upload(h.cell_face, &s->d_cell_face));
  MCFD_CREATE_CUDA(Upload(h.cell_code, &s->d_cell_code));
  MCFD_CREATE_CUDA(Upload(h.face_holder, &s->d_face_holder));
  MCFD_CREATE_CUDA(Upload(h.face_sharer, &s->d_face_sharer));
  MCFD_CREATE_CUDA(Upload(h.face_code, &s->d_face_code));
  MCFD_CREATE_CUDA(Upload(h.face_frame, &s->d_face_frame));
  MCFD_CREATE_CUDA(Upload(h.pt_holder, &s->d_pt_holder));
  MCFD_CREATE_CUDA(Upload(h.pt_sharer, &s->d_pt_sharer));
  MCFD_CREATE_CUDA(Upload(h.bface_group, &s->d_bface_group));
  MCFD_CREATE_CUDA(Upload(h.bface_local, &s->d_bface_local));
  MCFD_CREATE_CUDA(Upload(h.bface_xyz, &s->d_bface_xyz));
  MCFD_CREATE_CUDA(Upload(h.send_cells, &s->d_send_cells));
  MCFD_CREATE_CUDA(Upload(h.m_center, &s->d_m_center));
  MCFD_CREATE_CUDA(Upload(h.m_S, &s->d_m_S));
  MCFD_CREATE_CUDA(Upload(h.m_volume, &s->d_m_volume));
  MCFD_CREATE_CUDA(Upload(h.m_qxyzw, &s->d_m_qxyzw));
  MCFD_CREATE_CUDA(Upload(h.f_xyzw, &s->d_f_xyzw));
  MCFD_CREATE_CUDA(Upload(h.adj_ptr, &s->d_adj_ptr));
  MCFD_CREATE_CUDA(Upload(h.adj_cell, &s->d_adj_cell));
  MCFD_CREATE_CUDA(Upload(h.adj_face, &s->d_adj_face));
  if (h.n_classes > 0) {
    MCFD_CREATE_CUDA(Upload(h.class_cells, &s->d_class_cells));
    MCFD_CREATE_CUDA(Upload(h.class_ptr, &s->d_class_ptr));
    MCFD_CREATE_CUDA(Upload(h.mm_A1, &s->d_mm_A1));
    MCFD_CREATE_CUDA(Upload(h.mm_A2, &s->d_mm_A2));
    MCFD_CREATE_CUDA(Upload(h.mm_A3, &s->d_mm_A3));
    MCFD_CREATE_CUDA(Upload(h.cls_S, &s->d_cls_S));
    MCFD_CREATE_CUDA(Upload(h.cls_P, &s->d_cls_P));
    MCFD_CREATE_CUDA(Upload(h.cls_W, &s->d_cls_W));
    MCFD_CREATE_CUDA(Upload(h.cell_class, &s->d_cell_class));
  }
  std::vector<double>().swap(h.m_S); std::vector<double>().swap(h.m_qxyzw); std::vector<double>().swap(h.f_xyzw);
  // the big host tables are no longer needed once uploaded (corner_xyz is kept for Approximate)
  std::vector<double>().swap(h.metrics); std::vector<double>().swap(h.face_frame);
  std::vector<int32_t>().swap(h.cell_face); std::vector<uint8_t>().swap(h.cell_code);
  std::vector<int32_t>().swap(h.face_holder); std::vector<int32_t>().swap(h.face_sharer);
  std::vector<uint16_t>().swap(h.face_code);
  std::vector<int32_t>().swap(h.pt_holder); std::vector<int32_t>().swap(h.pt_sharer);
  s->groups.resize(std::max(h.n_groups, 1));
  for (auto &g : s->groups) { std::memset(&g, 0, sizeof(g)); g.kind = -1; }
  s->group_values.assign(s->groups.size(), nullptr);
  MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_groups), s->groups.size() * sizeof(BcGroup)));
  MCFD_CREATE_CUDA(cudaMemcpy(s->d_groups, s->groups.data(), s->groups.size() * sizeof(BcGroup), cudaMemcpyHostToDevice));
  for (int i = 0; i < 3; ++i) {
    MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_u[i]), s->column_all * sizeof(double)));
    MCFD_CREATE_CUDA(cudaMemset(s->d_u[i], 0, s->column_all * sizeof(double)));
  }
  const size_t fstar_bytes = size_t(FstarSize(std::max<int64_t>(h.n_faces, 1) * h.fq)) * sizeof(double);
  MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_fstar), fstar_bytes));
  MCFD_CREATE_CUDA(cudaMemset(s->d_fstar, 0, fstar_bytes));
  const size_t star_bytes = size_t(std::max<int64_t>(h.n_faces, 1)) * h.fq * sizeof(double);
  MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_star_u), star_bytes));
  MCFD_CREATE_CUDA(cudaMemset(s->d_star_u, 0, star_bytes));
  MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_ticket), 8 * sizeof(int)));
  MCFD_CREATE_CUDA(cudaMemset(s->d_ticket, 0, 8 * sizeof(int)));
  if (!h.send_cells.empty())
    MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_send), h.send_cells.size() * s->fields * sizeof(double)));
  if (cfg->scheme == MCFD_SCHEME_DG_MODAL)
    MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_tay), s->column_all * sizeof(double)));
  if (cfg->limiter != MCFD_LIMITER_NONE) {
    if (cfg->limiter != MCFD_LIMITER_DUMMY)   // the rebuilt projections; weno::Dummy rebuilds nothing
      MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_fresh), s->column * sizeof(double)));
    // troubled-cell flags [n_local], the compact list of troubled cells [n_local], its length [1], pad [1], and the
    // running 64-bit number of rebuilds [2 ints]
    MCFD_CREATE_CUDA(cudaMalloc(reinterpret_cast<void **>(&s->d_flags), (2 * h.n_local + 4) * sizeof(int32_t)));
    MCFD_CREATE_CUDA(cudaMemset(s->d_flags, 0, (2 * h.n_local + 4) * sizeof(int32_t)));
  }
#undef MCFD_CREATE_CUDA

  DeviceTables &d = s->dev;
  d.q = h.q; d.n = h.n; d.fq = h.fq;
  d.n_local = h.n_local; d.n_ghost = h.n_ghost; d.n_faces = h.n_faces;
  d.n_local_faces = h.n_local_faces; d.n_ghost_faces = h.n_ghost_faces; d.n_boundary_faces = h.n_boundary_faces;
  d.per_node_metrics = h.per_node_metrics; d.per_point_frames = h.per_point_frames;
  d.metrics = s->d_metrics; d.cell_face = s->d_cell_face; d.cell_code = s->d_cell_code;
  d.face_holder = s->d_face_holder; d.face_sharer = s->d_face_sharer; d.face_code = s->d_face_code;
  d.face_frame = s->d_face_frame; d.pt_holder = s->d_pt_holder; d.pt_sharer = s->d_pt_sharer; d.bface_group = s->d_bface_group; d.bface_local = s->d_bface_local;
  d.bface_xyz = s->d_bface_xyz; d.groups = s->d_groups; d.n_groups = int(s->groups.size());
  d.nq = h.nq; d.elem = h.elem; d.n_inner = h.n_inner;
  d.m_center = s->d_m_center; d.m_S = s->d_m_S; d.m_volume = s->d_m_volume; d.m_qxyzw = s->d_m_qxyzw; d.f_xyzw = s->d_f_xyzw;
  d.adj_ptr = s->d_adj_ptr; d.adj_cell = s->d_adj_cell; d.adj_face = s->d_adj_face;
  d.star_u = s->d_star_u;
  d.n_classes = h.n_classes; d.class_cells = s->d_class_cells; d.class_ptr = s->d_class_ptr;
  d.n_batches = 0;
  for (int k = 0; k < h.n_classes; ++k) d.n_batches += (h.class_ptr[k + 1] - h.class_ptr[k] + 7) / 8;
  d.mm_A1 = s->d_mm_A1; d.mm_A2 = s->d_mm_A2; d.mm_A3 = s->d_mm_A3; d.cls_S = s->d_cls_S;
  d.cls_P = s->d_cls_P; d.cls_W = s->d_cls_W; d.cell_class = s->d_cell_class;
  *out = s.release();
  return MCFD_OK;
}

int mcfd_destroy(mcfd_solver *s) {
  if (!s) return MCFD_OK;
  cudaSetDevice(s->cfg.device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  if (s->comm_stream) cudaStreamSynchronize(s->comm_stream);
  for (void *ptr : s->peer.opened) cudaIpcCloseMemHandle(ptr);
  // Neighbours may still have this rank's columns and counters open through cudaIpcOpenMemHandle, and freeing an
  // exported allocation before every importer has closed it is undefined.  mcfd_detach_nccl (collective) makes all
  // ranks close first; a solver destroyed WITHOUT it keeps its exported buffers until the process exits.
  const bool exported_in_use = s->peer.ready && s->comm && !s->host.nbr_rank.empty();
  if (exported_in_use)   // not an error of this call, but the caller should hear of it: mcfd_last_error(NULL)
    g_create_error = "mcfd_destroy: the solver still had peer links (mcfd_detach_nccl was not called on every rank first): "
                     "its three state columns and arrival counters stay allocated until the process exits";
  {
    void *mine[] = { s->peer.d_peer_u, s->peer.d_peer_flag, exported_in_use ? nullptr : s->peer.d_flags, s->peer.d_send_nbr, s->peer.d_send_dst, s->peer.d_error,
                     s->peer.d_sync };
    for (void *ptr : mine) if (ptr) cudaFree(ptr);
  }
  if (s->comm && s->nccl) s->nccl->CommDestroy(s->comm);
  void *ptrs[] = { s->d_metrics, s->d_face_frame, s->d_bface_xyz, s->d_corner, s->d_cell_face, s->d_face_holder,
                   s->d_face_sharer, s->d_bface_group, s->d_bface_local, s->d_send_cells, s->d_cell_code, s->d_face_code,
                   s->d_groups, s->d_pt_holder, s->d_pt_sharer, exported_in_use ? nullptr : s->d_u[0],
                   exported_in_use ? nullptr : s->d_u[1], exported_in_use ? nullptr : s->d_u[2], s->d_fstar, s->d_send,
                   s->d_m_center, s->d_m_S, s->d_m_volume, s->d_m_qxyzw, s->d_f_xyzw, s->d_fresh, s->d_adj_ptr, s->d_adj_cell,
                   s->d_adj_face, s->d_flags, s->d_tay, s->d_ticket, s->d_star_u, s->d_class_cells, s->d_class_ptr, s->d_mm_A1,
                   s->d_mm_A2, s->d_mm_A3, s->d_cls_S, s->d_cls_P, s->d_cls_W, s->d_cell_class };
  for (void *p : ptrs) if (p) cudaFree(p);
  for (double *p : s->group_values) if (p) cudaFree(p);
  if (s->h_pinned) cudaFreeHost(s->h_pinned);
  if (s->ev_packed) cudaEventDestroy(s->ev_packed);
  if (s->ev_halo) cudaEventDestroy(s->ev_halo);
  if (s->ev_begin) cudaEventDestroy(s->ev_begin);
  if (s->ev_end) cudaEventDestroy(s->ev_end);
  if (s->stream) cudaStreamDestroy(s->stream);
  if (s->comm_stream) cudaStreamDestroy(s->comm_stream);
  delete s;
  return MCFD_OK;
}

int64_t mcfd_column_size(const mcfd_solver *s) { return s ? s->column : -1; }
int32_t mcfd_fields_per_cell(const mcfd_solver *s) { return s ? s->fields : -1; }

int mcfd_set_time(mcfd_solver *s, double t) {
  if (!s) return MCFD_ERR_ARG;
  s->t_curr = t;
  return MCFD_OK;
}

int mcfd_set_solution(mcfd_solver *s, const double *column, int64_t n) {
  if (!s || !column) return MCFD_ERR_ARG;
  if (n != s->column) return Fail(s, MCFD_ERR_ARG, "column size mismatch");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  MCFD_CUDA(s, cudaMemcpyAsync(s->d_u[s->cur], column, n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  s->tay_of = nullptr;
  int rc = Limit(s, s->d_u[s->cur]);  // WithLimiterAndSource::SetSolutionColumn (dg/general.hpp:247-250)
  if (rc) return rc;
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  if (s->comm_stream) MCFD_CUDA(s, cudaStreamSynchronize(s->comm_stream));
  return MCFD_OK;
}

int mcfd_get_solution(mcfd_solver *s, double *column, int64_t n) {
  if (!s || !column) return MCFD_ERR_ARG;
  if (n != s->column) return Fail(s, MCFD_ERR_ARG, "column size mismatch");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  MCFD_CUDA(s, cudaMemcpyAsync(column, s->d_u[s->cur], n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  return MCFD_OK;
}

int mcfd_set_ghost_solution(mcfd_solver *s, const double *ghost_column, int64_t n) {
  if (!s || (!ghost_column && n)) return MCFD_ERR_ARG;
  if (n != s->column_all - s->column) return Fail(s, MCFD_ERR_ARG, "ghost column size mismatch");
  if (n == 0) return MCFD_OK;
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  MCFD_CUDA(s, cudaMemcpyAsync(s->d_u[s->cur] + s->column, ghost_column, n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  return MCFD_OK;
}

int mcfd_get_residual(mcfd_solver *s, double *column, int64_t n) {
  if (!s || !column) return MCFD_ERR_ARG;
  if (n != s->column) return Fail(s, MCFD_ERR_ARG, "column size mismatch");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  double *scratch = s->d_u[(s->cur + 1) % 3];
  int rc = Stage(s, s->d_u[s->cur], nullptr, scratch, StageOp{ 4, 0.0 }, s->t_curr);
  if (rc) return rc;
  MCFD_CUDA(s, cudaMemcpyAsync(column, scratch, n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  return MCFD_OK;
}

int mcfd_approximate(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t) {
  if (!s || n_params < 0 || n_params > MCFD_MAX_FUNC_PARAMS || (n_params && !params)) return MCFD_ERR_ARG;
  if (func <= MCFD_FUNC_HOST_VALUES || func > MCFD_FUNC_PLANE_SPLIT) return Fail(s, MCFD_ERR_ARG, "unknown function id");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  if (!s->d_corner) MCFD_CUDA(s, Upload(s->host.corner_xyz, &s->d_corner));
  double p[MCFD_MAX_FUNC_PARAMS] = { 0 };
  for (int i = 0; i < n_params; ++i) p[i] = params[i];
  s->tay_of = nullptr;
  if (s->cfg.scheme == MCFD_SCHEME_DG_MODAL)  // polynomial/projection.hpp:159-162 (no limiter: Cell::Approximate does not limit)
    MCFD_CUDA(s, LaunchModalApproximate(s->dev, s->cfg.degree, func, p, t, s->d_u[s->cur], s->stream));
  else
    MCFD_CUDA(s, LaunchApproximate(s->dev, s->basis, s->cfg.scheme, s->d_corner, func, p, t, s->d_u[s->cur], s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  return MCFD_OK;
}

int mcfd_measure_l1_error(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t, double *l1_error) {
  if (!s || !l1_error || n_params < 0 || n_params > MCFD_MAX_FUNC_PARAMS || (n_params && !params)) return MCFD_ERR_ARG;
  if (func <= MCFD_FUNC_HOST_VALUES || func > MCFD_FUNC_PLANE_SPLIT) return Fail(s, MCFD_ERR_ARG, "unknown function id");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  if (!s->d_corner) MCFD_CUDA(s, Upload(s->host.corner_xyz, &s->d_corner));
  double p[MCFD_MAX_FUNC_PARAMS] = { 0 };
  for (int i = 0; i < n_params; ++i) p[i] = params[i];
  int64_t n_blocks = 0;
  LaunchL1Error(s->dev, s->basis, s->cfg.scheme, s->cfg.degree, s->d_corner, func, p, t, s->d_u[s->cur], nullptr, &n_blocks, s->stream);
  // one of the two rotating stage buffers is free between steps and holds 5 doubles per block with room to spare
  double *partial = s->d_u[(s->cur + 1) % 3];
  if (n_blocks * 5 > s->column_all) return Fail(s, MCFD_ERR_UNSUPPORTED, "scratch column too small for the partial sums");
  MCFD_CUDA(s, LaunchL1Error(s->dev, s->basis, s->cfg.scheme, s->cfg.degree, s->d_corner, func, p, t, s->d_u[s->cur], partial,
                             &n_blocks, s->stream));
  std::vector<double> host(size_t(n_blocks) * 5);
  MCFD_CUDA(s, cudaMemcpyAsync(host.data(), partial, host.size() * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  for (int c = 0; c < 5; ++c) l1_error[c] = 0.0;
  for (int64_t b = 0; b < n_blocks; ++b)
    for (int c = 0; c < 5; ++c) l1_error[c] += host[size_t(b) * 5 + c];
  return MCFD_OK;
}

// No counterpart in the reference: mcfd_measure_l1_error for callers that cannot allocate (include/minicfd_b200.h).
// No cudaMalloc / cudaFree / host container on this path; the partial sums go to the static array of output_kernels.cu.
int mcfd_measure_l1_error_static(mcfd_solver *s, int32_t func, const double *params, int32_t n_params, double t,
                                 double *l1_error) {
  if (!s || !l1_error || n_params < 0 || n_params > MCFD_MAX_FUNC_PARAMS || (n_params && !params)) return MCFD_ERR_ARG;
  if (func <= MCFD_FUNC_HOST_VALUES || func > MCFD_FUNC_PLANE_SPLIT) return Fail(s, MCFD_ERR_ARG, "unknown function id");
  const bool modal = s->cfg.scheme == MCFD_SCHEME_DG_MODAL;
  const int64_t points = s->dev.n_local * (modal ? s->dev.nq : s->dev.n);
  if (points > kL1StaticMaxPoints)
    return Fail(s, MCFD_ERR_UNSUPPORTED, "mcfd_measure_l1_error_static: this part has " + std::to_string(points) +
                " (cell, quadrature point) pairs, the static scratch array holds the partial sums of at most " +
                std::to_string(kL1StaticMaxPoints) + "; use mcfd_measure_l1_error (which borrows a stage column)");
  if (!modal && !s->d_corner)
    return Fail(s, MCFD_ERR_UNSUPPORTED, "mcfd_measure_l1_error_static: the corner table of a nodal scheme is not on the "
                "device yet and this call may not allocate it; call mcfd_approximate, mcfd_measure_l1_error or "
                "mcfd_vtk_sample once during set-up");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  double p[MCFD_MAX_FUNC_PARAMS] = { 0 };
  for (int i = 0; i < n_params; ++i) p[i] = params[i];
  MCFD_CUDA(s, MeasureL1ErrorStatic(s->dev, s->basis, s->cfg.scheme, s->cfg.degree, s->d_corner, func, p, t, s->d_u[s->cur],
                                    l1_error, s->stream));
  return MCFD_OK;
}

int mcfd_set_boundary(mcfd_solver *s, int32_t group, int32_t kind, int32_t func, const double *params, int32_t n_params) {
  if (!s) return MCFD_ERR_ARG;
  if (group < 0 || group >= int(s->groups.size())) return Fail(s, MCFD_ERR_ARG, "boundary group out of range");
  if (kind < MCFD_BC_SOLID_WALL || kind > MCFD_BC_SMART) return Fail(s, MCFD_ERR_ARG, "unknown boundary kind");
  if (n_params < 0 || n_params > MCFD_MAX_FUNC_PARAMS || (n_params && !params)) return Fail(s, MCFD_ERR_ARG, "bad params");
  if (func < MCFD_FUNC_HOST_VALUES || func > MCFD_FUNC_PLANE_SPLIT) return Fail(s, MCFD_ERR_ARG, "unknown function id");
  bool needs_state = kind != MCFD_BC_SOLID_WALL && kind != MCFD_BC_SUPERSONIC_OUTLET;
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  BcGroup &g = s->groups[group];
  g.kind = kind; g.func = func;
  std::memset(g.params, 0, sizeof(g.params));
  for (int i = 0; i < n_params; ++i) g.params[i] = params[i];
  if (needs_state && func == MCFD_FUNC_HOST_VALUES && !s->group_values[group]) {
    // values must be supplied by mcfd_set_bc_values before the next residual; start from zeros
    int64_t n = s->host.group_faces[group] * s->host.fq * MCFD_K;
    if (n > 0) {
      MCFD_CUDA(s, cudaMalloc(reinterpret_cast<void **>(&s->group_values[group]), n * sizeof(double)));
      MCFD_CUDA(s, cudaMemset(s->group_values[group], 0, n * sizeof(double)));
    }
  }
  g.host_values = s->group_values[group];
  return SyncGroups(s);
}

int64_t mcfd_bc_point_count(const mcfd_solver *s, int32_t group) {
  if (!s || group < 0 || group >= int(s->groups.size())) return -1;
  return s->host.group_faces[group] * s->host.fq;
}

int mcfd_bc_points(const mcfd_solver *s, int32_t group, double *xyz, int64_t n) {
  if (!s || !xyz) return MCFD_ERR_ARG;
  if (n != mcfd_bc_point_count(s, group)) return MCFD_ERR_ARG;
  const HostTables &h = s->host;
  for (int64_t b = 0; b < h.n_boundary_faces; ++b) {
    if (h.bface_group[b] != group) continue;
    std::memcpy(xyz + int64_t(h.bface_local[b]) * h.fq * 3, &h.bface_xyz[size_t(b) * h.fq * 3], sizeof(double) * h.fq * 3);
  }
  return MCFD_OK;
}

int mcfd_set_bc_values(mcfd_solver *s, int32_t group, const double *values, int64_t n) {
  if (!s || !values) return MCFD_ERR_ARG;
  if (n != mcfd_bc_point_count(s, group)) return Fail(s, MCFD_ERR_ARG, "wrong number of boundary points");
  if (n == 0) return MCFD_OK;
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  if (!s->group_values[group])
    MCFD_CUDA(s, cudaMalloc(reinterpret_cast<void **>(&s->group_values[group]), n * MCFD_K * sizeof(double)));
  MCFD_CUDA(s, cudaMemcpyAsync(s->group_values[group], values, n * MCFD_K * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  s->groups[group].host_values = s->group_values[group];
  return SyncGroups(s);
}

/* temporal::RungeKutta<rk_order>::Update, n_steps times (rk.hpp:22-24,36-43,55-72).
 * Buffers: A = u(t) (kept as u_curr), B, C rotate; every stage is Stage() above. */
int mcfd_rk_update(mcfd_solver *s, double t, double dt, int32_t n_steps) {
  if (!s || n_steps < 0) return MCFD_ERR_ARG;
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  s->timing = mcfd_timing{};
  MCFD_CUDA(s, cudaEventRecord(s->ev_begin, s->stream));
  if (s->peer.ready && s->comm && n_steps > 0) {
    // The peer-memory exchange relies on neighbours being at most one stage apart, which holds from stage to stage
    // but not across API calls: a neighbour may still be inside an earlier call that reads its ghost cells (a bare
    // mcfd_get_residual, say) when this rank's first stage would store into them.  One small all-reduce per call
    // puts every rank at the door of its mcfd_rk_update (each rank's earlier calls have completed: they synchronise).
    if (s->nccl->AllReduce(s->peer.d_sync, s->peer.d_sync, 1, kNcclInt32, kNcclMin, s->comm, s->stream) != 0)
      return Fail(s, MCFD_ERR_NCCL, "all-reduce at the start of mcfd_rk_update failed");
  }
  for (int i = 0; i < n_steps; ++i) {
    const double t_curr = t + dt * i;
    const int a = s->cur, b = (s->cur + 1) % 3, c = (s->cur + 2) % 3;
    int rc = Stage(s, s->d_u[a], nullptr, s->d_u[b], StageOp{ 0, dt }, t_curr, true);
    if (!rc) rc = Limit(s, s->d_u[b]);
    if (rc) return rc;
    if (s->cfg.rk_order == 1) {
      s->cur = b;
    } else if (s->cfg.rk_order == 2) {
      rc = Stage(s, s->d_u[b], s->d_u[a], s->d_u[a], StageOp{ 3, dt }, t_curr + dt, true);
      if (!rc) rc = Limit(s, s->d_u[a]);
      if (rc) return rc;
    } else {
      rc = Stage(s, s->d_u[b], s->d_u[a], s->d_u[c], StageOp{ 1, dt }, t_curr + dt, true);
      if (!rc) rc = Limit(s, s->d_u[c]);
      if (rc) return rc;
      rc = Stage(s, s->d_u[c], s->d_u[a], s->d_u[a], StageOp{ 2, dt }, t_curr + dt / 2, true);
      if (!rc) rc = Limit(s, s->d_u[a]);
      if (rc) return rc;
    }
  }
  MCFD_CUDA(s, cudaEventRecord(s->ev_end, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->comm_stream));
  if (s->peer.ready && s->peer.d_error) {
    int who = 0;
    MCFD_CUDA(s, cudaMemcpy(&who, s->peer.d_error, sizeof(int), cudaMemcpyDeviceToHost));
    if (who) return Fail(s, MCFD_ERR_NCCL, "halo exchange over peer memory timed out waiting for rank " + std::to_string(s->host.nbr_rank[who - 1]));
  }
  float ms = 0;
  MCFD_CUDA(s, cudaEventElapsedTime(&ms, s->ev_begin, s->ev_end));
  s->timing.ms_total = ms;
  if (s->profiling) CollectKernelTimes(s);
  s->t_curr = t + dt * n_steps;
  return MCFD_OK;
}

int32_t mcfd_shape_classes(const mcfd_solver *s, int32_t *n_classes) {
  if (!s) return 0;
  if (n_classes) *n_classes = s->host.n_classes;
  return s->host.n_classes > 0 ? s->host.n_shapes : 0;
}

int mcfd_nccl_unique_id(void *unique_id_128) {
  if (!unique_id_128) return MCFD_ERR_ARG;
  std::string err;
  NcclApi *api = LoadNccl(&err);
  if (!api) { g_create_error = err; return MCFD_ERR_NCCL; }
  return api->GetUniqueId(unique_id_128) == 0 ? MCFD_OK : MCFD_ERR_NCCL;
}

int mcfd_attach_nccl(mcfd_solver *s, const void *unique_id_128, int32_t rank, int32_t world) {
  if (!s || !unique_id_128) return MCFD_ERR_ARG;
  std::string err;
  s->nccl = LoadNccl(&err);
  if (!s->nccl) return Fail(s, MCFD_ERR_NCCL, err);
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  UniqueId id;
  std::memcpy(&id, unique_id_128, sizeof(id));
  int rc = s->nccl->CommInitRank(&s->comm, world, id, rank);
  if (rc != 0) { s->comm = nullptr; return Fail(s, MCFD_ERR_NCCL, s->nccl->GetErrorString(rc)); }
  return SetupPeerLinks(s);
}

int mcfd_detach_nccl(mcfd_solver *s) {
  if (!s) return MCFD_ERR_ARG;
  if (!s->comm) return MCFD_OK;
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->comm_stream));
  for (void *ptr : s->peer.opened) cudaIpcCloseMemHandle(ptr);
  s->peer.opened.clear();
  int rc = 0;
  if (s->peer.ready && s->peer.d_sync) {  // every rank has closed its imports once this all-reduce completes
    rc = s->nccl->AllReduce(s->peer.d_sync, s->peer.d_sync, 1, kNcclInt32, kNcclMin, s->comm, s->stream);
    if (rc == 0 && cudaStreamSynchronize(s->stream) != cudaSuccess) rc = -1;
  }
  s->peer.ready = false;
  s->nccl->CommDestroy(s->comm);
  s->comm = nullptr;
  if (rc != 0) return Fail(s, MCFD_ERR_NCCL, "all-reduce in mcfd_detach_nccl failed");
  return MCFD_OK;
}

int mcfd_halo_transport(const mcfd_solver *s) {
  if (!s || !s->comm) return 0;
  return s->peer.ready ? 2 : 1;
}

int mcfd_get_timing(const mcfd_solver *s, mcfd_timing *out) {
  if (!s || !out) return MCFD_ERR_ARG;
  *out = s->timing;
  return MCFD_OK;
}

int mcfd_troubled_cells(mcfd_solver *s, int32_t *flags, int64_t *count) {
  if (!s || (!flags && !count)) return MCFD_ERR_ARG;
  if (!s->d_flags) return Fail(s, MCFD_ERR_UNSUPPORTED, "no limiter configured");
  cudaSetDevice(s->cfg.device);
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  if (flags && s->cfg.limiter == MCFD_LIMITER_DUMMY) {  // Dummy::IsNotSmooth is true for every cell
    for (int64_t i = 0; i < s->host.n_local; ++i) flags[i] = 1;
  } else if (flags) {
    std::vector<int32_t> host(size_t(s->host.n_local));
    MCFD_CUDA(s, cudaMemcpy(host.data(), s->d_flags, host.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    for (size_t i = 0; i < host.size(); ++i) flags[i] = host[i] != 0;
  }
  if (count) {
    unsigned long long total = 0;
    MCFD_CUDA(s, cudaMemcpy(&total, s->d_flags + 2 * s->host.n_local + 2, sizeof(total), cudaMemcpyDeviceToHost));
    *count = int64_t(total) + s->lazy_rebuilds;
  }
  return MCFD_OK;
}

int mcfd_set_profiling(mcfd_solver *s, int32_t per_kernel_events) {
  if (!s) return MCFD_ERR_ARG;
  s->profiling = per_kernel_events != 0;
  return MCFD_OK;
}

// host only (no device is touched): the tables mcfd_create would build, reduced to what it found
int mcfd_mesh_shape_classes(const mcfd_config *cfg, const mcfd_mesh *mesh, int32_t *n_shapes, int32_t *n_classes,
                            int32_t *per_point_frames) {
  if (!cfg || !mesh) return MCFD_ERR_ARG;
  Basis basis;
  int rc = BuildBasis(cfg->degree, cfg->scheme, &basis);
  if (rc) return rc;
  HostTables host;
  std::string error;
  rc = BuildTables(*cfg, *mesh, basis, &host, &error);
  if (rc) { g_create_error = error; return rc; }
  if (n_shapes) *n_shapes = host.n_classes > 0 ? host.n_shapes : 0;
  if (n_classes) *n_classes = host.n_classes;
  if (per_point_frames) *per_point_frames = host.per_point_frames ? 1 : 0;
  return MCFD_OK;
}

int64_t mcfd_struct_size(int32_t which) {
  switch (which) {
    case 0: return sizeof(mcfd_mesh);
    case 1: return sizeof(mcfd_config);
    case 2: return sizeof(mcfd_timing);
    default: return -1;
  }
}

/* ---- row f2: Section::GatherFields / ScatterFields, vtk::Writer ---- */
int mcfd_gather_fields(mcfd_solver *s, double *fields, int64_t n) {
  if (!s || !fields) return MCFD_ERR_ARG;
  if (n != s->column) return Fail(s, MCFD_ERR_ARG, "column size mismatch");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  double *scratch = s->d_u[(s->cur + 1) % 3];
  MCFD_CUDA(s, cudaEventRecord(s->ev_begin, s->stream));
  MCFD_CUDA(s, LaunchTransposeFields(true, s->host.n_local, s->fields, s->d_u[s->cur], scratch, s->stream));
  MCFD_CUDA(s, cudaEventRecord(s->ev_end, s->stream));
  MCFD_CUDA(s, cudaMemcpyAsync(fields, scratch, n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  s->timing = mcfd_timing{};  // ms_total = the kernel alone (bench.py --what output)
  float ms = 0;
  MCFD_CUDA(s, cudaEventElapsedTime(&ms, s->ev_begin, s->ev_end));
  s->timing.ms_total = ms;
  return MCFD_OK;
}

int mcfd_scatter_fields(mcfd_solver *s, const double *fields, int64_t n) {
  if (!s || !fields) return MCFD_ERR_ARG;
  if (n != s->column) return Fail(s, MCFD_ERR_ARG, "column size mismatch");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  double *scratch = s->d_u[(s->cur + 1) % 3];
  MCFD_CUDA(s, cudaMemcpyAsync(scratch, fields, n * sizeof(double), cudaMemcpyHostToDevice, s->stream));
  MCFD_CUDA(s, LaunchTransposeFields(false, s->host.n_local, s->fields, scratch, s->d_u[s->cur], s->stream));
  s->tay_of = nullptr;
  int rc = Limit(s, s->d_u[s->cur]);
  if (rc) return rc;
  MCFD_CUDA(s, cudaStreamSynchronize(s->stream));
  if (s->comm_stream) MCFD_CUDA(s, cudaStreamSynchronize(s->comm_stream));
  return MCFD_OK;
}

int mcfd_vtk_sample(mcfd_solver *s, double *coords, double *values, int64_t n_points) {
  if (!s || !coords || !values) return MCFD_ERR_ARG;
  if (s->host.elem != MCFD_CELL_HEXA_8) return Fail(s, MCFD_ERR_UNSUPPORTED, "vtk sampling is built for hexahedra only (vtk::Hexahedron27)");
  if (n_points != 27 * s->host.n_local) return Fail(s, MCFD_ERR_ARG, "n_points must be 27 * n_local_cells");
  MCFD_CUDA(s, cudaSetDevice(s->cfg.device));
  if (!s->d_corner) MCFD_CUDA(s, Upload(s->host.corner_xyz, &s->d_corner));
  // the two rotating stage buffers are free between steps: 8 doubles per sample fit when 27 * 8 <= 2 * fields,
  // otherwise (p = 1) a temporary is allocated
  double *d_coords = nullptr, *d_values = nullptr, *tmp = nullptr;
  if (int64_t(27) * 3 <= s->fields && int64_t(27) * 5 <= s->fields) {
    d_coords = s->d_u[(s->cur + 1) % 3];
    d_values = s->d_u[(s->cur + 2) % 3];
  } else {
    MCFD_CUDA(s, cudaMalloc(reinterpret_cast<void **>(&tmp), size_t(n_points) * 8 * sizeof(double)));
    d_coords = tmp;
    d_values = tmp + n_points * 3;
  }
  cudaError_t e = cudaEventRecord(s->ev_begin, s->stream);
  if (e == cudaSuccess) e = LaunchVtkSample(s->dev, s->basis, s->cfg.scheme, s->cfg.degree, s->d_corner, s->d_u[s->cur], d_coords,
                                            d_values, s->stream);
  if (e == cudaSuccess) e = cudaEventRecord(s->ev_end, s->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(coords, d_coords, size_t(n_points) * 3 * sizeof(double), cudaMemcpyDeviceToHost, s->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(values, d_values, size_t(n_points) * 5 * sizeof(double), cudaMemcpyDeviceToHost, s->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
  if (tmp) cudaFree(tmp);
  if (e != cudaSuccess) return Fail(s, MCFD_ERR_CUDA, cudaGetErrorString(e));
  s->timing = mcfd_timing{};  // ms_total = the sampling kernel alone (bench.py --what output)
  float ms = 0;
  MCFD_CUDA(s, cudaEventElapsedTime(&ms, s->ev_begin, s->ev_end));
  s->timing.ms_total = ms;
  return MCFD_OK;
}

int mcfd_write_vtu(mcfd_solver *s, const char *directory, const char *soln_name, const char *const *field_names) {
  if (!s || !directory || !soln_name) return MCFD_ERR_ARG;
  const int64_t n_points = 27 * s->host.n_local;
  std::vector<double> coords(size_t(n_points) * 3), values(size_t(n_points) * 5);
  int rc = mcfd_vtk_sample(s, coords.data(), values.data(), n_points);
  if (rc) return rc;
  std::string err;
  rc = WriteVtuFiles(directory, soln_name, field_names, s->cfg.rank, std::max(s->cfg.world, 1), s->host.n_local,
                     coords.data(), values.data(), &err);
  if (rc) return Fail(s, rc, err);
  return MCFD_OK;
}

int mcfd_write_vtu_points(const char *directory, const char *soln_name, const char *const *field_names, int32_t rank,
                          int32_t world, int64_t n_cells, const double *coords, const double *values) {
  if (!directory || !soln_name || n_cells < 0 || rank < 0 || world <= rank || (n_cells && (!coords || !values))) return MCFD_ERR_ARG;
  std::string err;
  int rc = WriteVtuFiles(directory, soln_name, field_names, rank, world, n_cells, coords, values, &err);
  if (rc) g_create_error = err;
  return rc;
}

void *mcfd_device_solution(mcfd_solver *s) {
  if (!s) return nullptr;
  s->tay_of = nullptr;  // the caller may write through the pointer
  return s->d_u[s->cur];
}
void *mcfd_stream(mcfd_solver *s) { return s ? s->stream : nullptr; }

/* Standalone device entry points for the known-answer tests. Host pointers in, host pointers out. */
static int RunPointwise(int which, int32_t kind, double gamma, int64_t n, const double *frames, const double *a,
                        const double *b, double *flux) {
  if (n < 0 || (n && (!frames || !a || !b || !flux))) return MCFD_ERR_ARG;
  if (n == 0) return MCFD_OK;
  int n_dev = 0;
  if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { g_create_error = "no CUDA device (there is no CPU fallback)"; return MCFD_ERR_CUDA; }
  double *d_fr = nullptr, *d_a = nullptr, *d_b = nullptr, *d_f = nullptr;
  cudaError_t e = cudaMalloc(reinterpret_cast<void **>(&d_fr), n * 9 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d_a), n * 5 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d_b), n * 5 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void **>(&d_f), n * 5 * sizeof(double));
  if (e == cudaSuccess) e = cudaMemcpy(d_fr, frames, n * 9 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_a, a, n * 5 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(d_b, b, n * 5 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = which == 0 ? LaunchRiemannFlux(kind, gamma, n, d_fr, d_a, d_b, d_f, nullptr)
                   : LaunchBoundaryFlux(kind, gamma, n, d_fr, d_a, d_b, d_f, nullptr);
  if (e == cudaSuccess) e = cudaMemcpy(flux, d_f, n * 5 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d_fr); cudaFree(d_a); cudaFree(d_b); cudaFree(d_f);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); return MCFD_ERR_CUDA; }
  return MCFD_OK;
}

int mcfd_riemann_flux(int32_t riemann, double gamma, int64_t n, const double *frames, const double *u_left,
                      const double *u_right, double *flux) {
  return RunPointwise(0, riemann, gamma, n, frames, u_left, u_right, flux);
}

int mcfd_boundary_flux(int32_t kind, double gamma, int64_t n, const double *frames, const double *u_inner,
                       const double *u_given, double *flux) {
  return RunPointwise(1, kind, gamma, n, frames, u_inner, u_given, flux);
}

}  // extern "C"
