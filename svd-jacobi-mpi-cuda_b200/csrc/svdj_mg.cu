// svdj_mg.cu -- the multi-GPU driver in C++ behind the C ABI: one collective call solves on N GPUs.
//
// The reference's distributed entry point is one collective call with the matrix on rank 0 only
// (main.cu:1455, 1587; rank and size taken inside the solver, lib/JacobiMethods.cu:209-211); every round-robin step
// rank 0 packs and MPI_Sends each rank its columns, receives them back and barriers (main.cu:583-942).  Here:
//
//   * n columns -> 2P column blocks of w columns; rank r keeps two resident (A-block m x w and the matching V-block
//     packed in one buffer).  A sweep is a tournament of the 2P blocks (circle method): round 0 sweeps all pairs of
//     the 2w local columns, rounds 1..2P-2 only the pairs across the two local blocks; between rounds the blocks move
//     one place round the ring.
//   * The move is fused into the compute: the update kernel of a round's last step stores the travelling blocks
//     straight into the buffers their next owners will read (peer stores over NVLink; four buffers per rank with
//     deterministic roles, so a sender knows which buffer of its neighbour is free).
//   * Rounds are ordered by FLAGS in peer memory, not by collectives: after its last kernel of a round a rank stores
//     the round's stamp into its two neighbours' flag words (st.release.sys), and before the next round it waits for
//     the stamps of the ranks it receives from (ld.acquire.sys in a one-thread kernel on the compute stream).  The
//     per-sweep convergence data (rotations, largest cosine) travels the same way: every rank posts its counters
//     into every rank's mailbox, waits for all P, and all ranks take the same decision.  No NCCL call, no host
//     synchronisation inside a sweep.
//   * Scatter and gather are rank 0's job, as in the reference: it copies the blocks out of its device-resident
//     matrix into the ranks' buffers and, at the end, gathers the columns in sigma order from wherever they are.
//
// One rank = one GPU = one host thread that owns a CUDA stream.  The ranks of one process are threads
// (svdj_dgesvd_mg: peer access between the devices); ranks in different processes (one per GPU, e.g. under
// torchrun) map each other's buffers through CUDA IPC (svdj_mg_rank_*).  Same code either way: the only thing a
// communicator has to provide is an all-gather of a few hundred host bytes at set-up.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <string>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <unistd.h>
#include <vector>

#include "svdj_common.cuh"
#include "svdj_host.cuh"
#include "svdj_qr.cuh"

namespace svdj {

typedef unsigned long long u64;

// ---------------------------------------------------------------------------------------------------------
// device side: flags, mailboxes, block copies
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(u64 *p, u64 v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ u64 ld_acquire_sys(const u64 *p) {
  u64 v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 global_timer_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Everything this stream did before (the update kernel's peer stores included) becomes visible before the flags.
__global__ void mg_signal_kernel(u64 *f0, u64 *f1, u64 stamp) {
  __threadfence_system();
  if (f0) st_release_sys(f0, stamp);
  if (f1) st_release_sys(f1, stamp);
}

// Root -> rank: the numerically-zero floor of this solve, then the start stamp
__global__ void mg_start_kernel(u64 *ctl, u64 floor_bits, u64 stamp) {
  ctl[64] = floor_bits;
  __threadfence_system();
  st_release_sys(ctl + 32, stamp);
}

// Spins until both flags have reached `stamp`.  A peer that never arrives (it failed) must not hang this GPU for
// ever: after timeout_ns the kernel records the failure in *err and returns; the host checks *err once per sweep.
// *err = (kind << 56) | (which flag << 48) | stamp: what was waited for when the time ran out
__global__ void mg_wait_kernel(const u64 *f0, const u64 *f1, u64 stamp, u64 *err, u64 timeout_ns, int kind = 1) {
  if (*reinterpret_cast<volatile u64 *>(err) != 0) return;  // an earlier wait of this solve already gave up
  const u64 t0 = global_timer_ns();
  for (int i = 0; i < 2; ++i) {
    const u64 *f = i == 0 ? f0 : f1;
    if (f == nullptr) continue;
    while (ld_acquire_sys(f) < stamp) {
      if (global_timer_ns() - t0 > timeout_ns) {
        *err = ((u64)kind << 56) | ((u64)i << 48) | stamp;
        return;
      }
      __nanosleep(200);
    }
  }
}

// mail layout per rank and parity: P slots of 4 words {applied, offdiag bits, stamp, -}
__global__ void mg_post_counters_kernel(const SweepCounters *c, u64 *const *mail_of_rank, int rank, int P, u64 stamp) {
  const int q = threadIdx.x;
  if (q >= P) return;
  u64 *slot = mail_of_rank[q] + 4 * rank;
  slot[0] = c->applied;
  slot[1] = c->offdiag_bits;
  __threadfence_system();
  st_release_sys(slot + 2, stamp);
}

__global__ void mg_collect_kernel(const u64 *mail, int P, u64 stamp, u64 *out, u64 *err, u64 timeout_ns) {
  if (*reinterpret_cast<volatile u64 *>(err) != 0) {
    out[0] = ~0ull;
    return;
  }
  const u64 t0 = global_timer_ns();
  u64 applied = 0, bits = 0;
  for (int q = 0; q < P; ++q) {
    const u64 *slot = mail + 4 * q;
    while (ld_acquire_sys(slot + 2) < stamp) {
      if (global_timer_ns() - t0 > timeout_ns) {
        *err = (2ull << 56) | ((u64)q << 48) | stamp;
        out[0] = ~0ull;
        return;
      }
      __nanosleep(200);
    }
    applied += slot[0];
    bits = max(bits, slot[1]);
  }
  out[0] = applied;
  out[1] = bits;
}

// dst (rows x cols_total, ldd) <- src (rows x cols_valid, lds), zero columns after cols_valid; dst may be peer memory
__global__ void __launch_bounds__(256) mg_copy_block_kernel(const double *__restrict__ src, size_t lds,
                                                            double *__restrict__ dst, size_t ldd, unsigned rows,
                                                            unsigned cols_valid) {
  const unsigned c = blockIdx.x;
  double *d = dst + (size_t)c * ldd;
  if (c < cols_valid) {
    const double *s = src + (size_t)c * lds;
    for (unsigned i = threadIdx.x; i < rows; i += 256) d[i] = s[i];
  } else {
    for (unsigned i = threadIdx.x; i < rows; i += 256) d[i] = 0.0;
  }
}

// V-block of w columns: column j is the unit vector e_(col0 + j), rows 0 .. rows-1
__global__ void __launch_bounds__(256) mg_init_v_kernel(double *__restrict__ V, size_t ldv, unsigned rows,
                                                        unsigned col0) {
  double *d = V + (size_t)blockIdx.x * ldv;
  const unsigned one = col0 + blockIdx.x;
  for (unsigned i = threadIdx.x; i < rows; i += 256) d[i] = (i == one) ? 1.0 : 0.0;
}

// out[:, k] = column at src[k] (anywhere: this GPU or a peer); null = zero column
__global__ void __launch_bounds__(256) mg_gather_kernel(const double *const *__restrict__ src, double *__restrict__ out,
                                                        size_t ldo, unsigned rows) {
  const double *s = src[blockIdx.x];
  double *d = out + (size_t)blockIdx.x * ldo;
  if (s != nullptr) {
    for (unsigned i = threadIdx.x; i < rows; i += 256) d[i] = s[i];
  } else {
    for (unsigned i = threadIdx.x; i < rows; i += 256) d[i] = 0.0;
  }
}

// sigma of the local columns goes to the root's mailbox: [2w sigmas][stamp as a double-sized word]
__global__ void __launch_bounds__(256) mg_post_sigma_kernel(const double *__restrict__ sig, unsigned count,
                                                            double *__restrict__ dst, u64 stamp) {
  for (unsigned i = threadIdx.x; i < count; i += 256) dst[i] = sig[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) st_release_sys(reinterpret_cast<u64 *>(dst + count), stamp);
}

// ---------------------------------------------------------------------------------------------------------
// communicators
// ---------------------------------------------------------------------------------------------------------
struct Comm {
  int rank = 0, size = 1;
  virtual ~Comm() {}
  virtual int allgather(const void *send, void *recv, size_t bytes) = 0;  // host memory; recv = size * bytes
  // Ranks that are threads of one process may share a GPU (the tests do that on a one-GPU box).  cudaMalloc
  // synchronises the whole device, so on a shared device it would wait for a peer's spinning wait kernel, which in
  // turn waits for this rank: every rank therefore finishes its allocations before any rank launches a kernel.
  // Ranks in different processes own their device and need no such barrier.
  // Returns the first non-zero status among the ranks (so that one rank's failed allocation stops all of them).
  virtual int before_first_launch(int my_status) { return my_status; }
  // Ranks that SHARE a GPU (threads of one process on one device: the one-GPU test mode).  A kernel that spins on a
  // flag may only be launched once the work that will set the flag has been submitted: streams of one device can
  // share a hardware queue, and a spinning kernel at its head then blocks the very kernel it is waiting for.  For
  // such a pair the signalling rank records an event after its signal and publishes the stamp on the host; the
  // waiting rank waits (on the host) for the publication and makes its stream wait for the event, so the spinning
  // kernel finds its flag already set.  Ranks on different GPUs have their own queues: no-ops there.
  virtual void set_devices(const std::vector<int> &) {}
  virtual void publish(int kind, int dst, u64 stamp, cudaStream_t s) { (void)kind, (void)dst, (void)stamp, (void)s; }
  virtual void await(int kind, int src, u64 stamp, cudaStream_t s) { (void)kind, (void)src, (void)stamp, (void)s; }
};
enum { kChanStart = 0, kChanRound = 1, kChanMail = 2, kChanSigma = 3, kChanDone = 4, kChans = 5 };

struct CallbackComm : Comm {
  svdj_allgather_fn fn;
  void *ctx;
  int allgather(const void *send, void *recv, size_t bytes) override { return fn(ctx, send, recv, bytes); }
};

// ranks = threads of this process
struct LocalHub {
  std::mutex mu;
  std::condition_variable cv;
  int size = 1, arrived = 0;
  unsigned long generation = 0;
  std::vector<unsigned char> buf;
  // same-device ordering (Comm::publish / await): one event and one published stamp per (channel, src, dst)
  std::vector<cudaEvent_t> ev;
  std::unique_ptr<std::atomic<u64>[]> pub;
  size_t slot(int kind, int src, int dst) const { return ((size_t)kind * size + src) * size + dst; }
  void init_board(int P) {
    size = P;
    ev.assign((size_t)kChans * P * P, nullptr);
    pub.reset(new std::atomic<u64>[(size_t)kChans * P * P]);
    for (size_t i = 0; i < (size_t)kChans * P * P; ++i) pub[i].store(0);
  }
};
struct LocalComm : Comm {
  LocalHub *hub = nullptr;
  std::vector<int> devices;
  void set_devices(const std::vector<int> &d) override {
    devices = d;
    for (int q = 0; q < size; ++q) {
      if (q == rank || devices[q] != devices[rank]) continue;
      for (int k = 0; k < kChans; ++k) {
        cudaEvent_t e = nullptr;
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) == cudaSuccess) hub->ev[hub->slot(k, rank, q)] = e;
      }
    }
  }
  bool shares_device(int q) const { return q != rank && !devices.empty() && devices[q] == devices[rank]; }
  void publish(int kind, int dst, u64 stamp, cudaStream_t s) override {
    if (!shares_device(dst)) return;
    const size_t i = hub->slot(kind, rank, dst);
    if (hub->ev[i]) cudaEventRecord(hub->ev[i], s);
    hub->pub[i].store(stamp, std::memory_order_release);
  }
  void await(int kind, int src, u64 stamp, cudaStream_t s) override {
    if (!shares_device(src)) return;
    const size_t i = hub->slot(kind, src, rank);
    const auto t0 = std::chrono::steady_clock::now();
    while (hub->pub[i].load(std::memory_order_acquire) < stamp) {
      if (std::chrono::steady_clock::now() - t0 > std::chrono::seconds(60)) return;  // the device-side wait reports it
      std::this_thread::yield();
    }
    if (hub->ev[i]) cudaStreamWaitEvent(s, hub->ev[i], 0);
  }
  int before_first_launch(int my_status) override {
    std::vector<int> all((size_t)size);
    allgather(&my_status, all.data(), sizeof(int));
    for (int v : all)
      if (v != 0) return v;
    return 0;
  }
  int allgather(const void *send, void *recv, size_t bytes) override {
    std::unique_lock<std::mutex> lk(hub->mu);
    if (hub->buf.size() < bytes * (size_t)size) hub->buf.resize(bytes * (size_t)size);
    std::memcpy(hub->buf.data() + bytes * (size_t)rank, send, bytes);
    const unsigned long gen = hub->generation;
    if (++hub->arrived == size) {
      hub->arrived = 0;
      ++hub->generation;
      hub->cv.notify_all();
    } else {
      hub->cv.wait(lk, [&] { return hub->generation != gen; });
    }
    std::memcpy(recv, hub->buf.data(), bytes * (size_t)size);
    // second phase: nobody may overwrite buf for the next all-gather before everyone has copied out
    const unsigned long gen2 = hub->generation;
    if (++hub->arrived == size) {
      hub->arrived = 0;
      ++hub->generation;
      hub->cv.notify_all();
    } else {
      hub->cv.wait(lk, [&] { return hub->generation != gen2; });
    }
    return 0;
  }
};

// ---------------------------------------------------------------------------------------------------------
// tournament of the 2P column blocks (host logic; tests/test_mg_cpu.py checks it through svdj_mg_schedule)
// ---------------------------------------------------------------------------------------------------------
struct Role {
  int top, bottom, spare[2];
};
struct Transfer {
  int src_rank, src_buf, dst_rank, dst_buf;
};

static std::vector<Role> initial_roles(int P) { return std::vector<Role>((size_t)P, Role{0, 1, {2, 3}}); }

// Buffer roles after one move of the circle method (top[0] stays; the other 2P-1 places shift one step round
// top[1] -> ... -> top[P-1] -> bottom[P-1] -> ... -> bottom[0] -> top[1]) and the transfers that realise it.
static void advance_roles(int P, const std::vector<Role> &roles, std::vector<Role> &next, std::vector<Transfer> &tr) {
  next.clear();
  tr.clear();
  if (P == 1) {
    next = roles;
    return;
  }
  for (int r = 0; r < P; ++r) {
    int spare[2] = {roles[r].spare[0], roles[r].spare[1]};
    int ns = 2;
    int nt, nb;
    if (r == 0) {
      nb = spare[--ns];
      tr.push_back({1, roles[1].bottom, 0, nb});  // new bottom[0] = old bottom[1]
      nt = roles[0].top;                          // block 0 never moves
    } else {
      nt = spare[--ns];
      if (r >= 2) tr.push_back({r - 1, roles[r - 1].top, r, nt});  // new top[r] = old top[r-1]
      else tr.push_back({0, roles[0].bottom, r, nt});              // new top[1] = old bottom[0]
      if (r < P - 1) {
        nb = spare[--ns];
        tr.push_back({r + 1, roles[r + 1].bottom, r, nb});  // new bottom[r] = old bottom[r+1]
      } else {
        nb = roles[r].top;  // the last rank's top turns the corner
      }
    }
    Role nr;
    nr.top = nt;
    nr.bottom = nb;
    int k = 0;
    for (int i = 0; i < 4; ++i)
      if (i != nt && i != nb) nr.spare[k++] = i;
    next.push_back(nr);
  }
}

// one move of the circle method on the block ids (the same move advance_roles makes on the buffers)
static void move_ids(int P, std::vector<int> &top, std::vector<int> &bottom) {
  if (P == 1) return;
  std::vector<int> nt(P), nb(P);
  nt[0] = top[0];
  nt[1] = bottom[0];
  for (int r = 2; r < P; ++r) nt[r] = top[r - 1];
  for (int r = 0; r < P - 1; ++r) nb[r] = bottom[r + 1];
  nb[P - 1] = top[P - 1];
  top.swap(nt);
  bottom.swap(nb);
}

// block ids (top, bottom) per rank after `round` moves
static void arrangement(int P, int round, std::vector<int> &top, std::vector<int> &bottom) {
  top.resize(P);
  bottom.resize(P);
  for (int r = 0; r < P; ++r) {
    top[r] = r;
    bottom[r] = P + r;
  }
  for (int it = 0; it < round; ++it) move_ids(P, top, bottom);
}

// ---------------------------------------------------------------------------------------------------------
// one rank
// ---------------------------------------------------------------------------------------------------------
static constexpr int kBufs = 4;
static constexpr size_t kCtlFlagWords = 128;  // flags: [0], [16] arrivals; [32] start; [48] done; [80] error
// how long a rank waits for a peer before it gives up (SVDJ_MG_TIMEOUT_S overrides the 60 s)
static u64 timeout_ns() {
  static const u64 t = [] {
    const char *e = std::getenv("SVDJ_MG_TIMEOUT_S");
    const long v = e ? std::atol(e) : 60;
    return (u64)(v > 0 ? v : 60) * 1000000000ull;
  }();
  return t;
}
#define kTimeoutNs (timeout_ns())

struct PeerBlob {
  int status, pid, device;
  u64 ptr[kBufs + 1];
  cudaIpcMemHandle_t ipc[kBufs + 1];
};

struct MgRank {
  std::unique_ptr<Comm> comm;
  int rank = 0, P = 1, device = 0;
  size_t max_m = 0, max_n = 0;  // as given at creation
  size_t w_max = 0, lda = 0, ldv = 0;  // block width capacity and strides of the buffers
  cudaStream_t stream = nullptr;
  double *buf[kBufs] = {};       // own buffers: A-part w_max x lda, then V-part w_max x ldv
  u64 *ctl = nullptr;            // own control block
  size_t ctl_words = 0;
  std::vector<std::vector<double *>> peer_buf;  // [rank][buffer]
  std::vector<u64 *> peer_ctl;                  // [rank]
  std::vector<void *> ipc_opened;
  SweepCounters *d_counters = nullptr;
  u64 *d_out = nullptr;   // collected counters
  u64 *h_out = nullptr;   // pinned
  u64 **d_mail_ptrs[2] = {nullptr, nullptr};  // device arrays of P pointers (every rank's mail of a parity)
  double *d_sigma = nullptr;                  // 2 w_max
  const double **d_src = nullptr;             // root: gather tables (2 x max_n pointers)
  std::map<std::pair<size_t, size_t>, void *> workspaces;  // (m, 2w) -> RingWorkspace
  // probes of the update kernel (every 8th step), read back once per sweep
  static constexpr int kProbes = 192;
  std::vector<BlockProbe> probe_store;
  u64 *d_probe_pairs = nullptr;
  u64 epoch = 0;        // solves done (start / done / sigma stamps)
  u64 round_stamp = 0;  // rounds done
  u64 sweep_stamp = 0;  // sweeps done
  std::string error;

  size_t mail_off(int parity) const { return kCtlFlagWords + (size_t)parity * 4 * (size_t)P; }
  size_t sigma_off() const { return kCtlFlagWords + 8 * (size_t)P; }  // doubles: P slots of (2 w_max + 2)
  double *A_of(double *b) const { return b; }
  double *V_of(double *b) const { return b + w_max * lda; }
};

#define MG_TRY(expr)                          \
  do {                                        \
    int rc__ = (expr);                        \
    if (rc__ != SVDJ_OK) return rc__;         \
  } while (0)

static void mg_destroy(MgRank *R) {
  if (!R) return;
  cudaSetDevice(R->device);
  if (R->stream) cudaStreamSynchronize(R->stream);
  for (void *p : R->ipc_opened) cudaIpcCloseMemHandle(p);
  for (auto &kv : R->workspaces) svdj_ring_ws_destroy(kv.second);
  for (int i = 0; i < kBufs; ++i)
    if (R->buf[i]) cudaFree(R->buf[i]);
  if (R->ctl) cudaFree(R->ctl);
  if (R->d_counters) cudaFree(R->d_counters);
  if (R->d_out) cudaFree(R->d_out);
  if (R->h_out) cudaFreeHost(R->h_out);
  if (R->d_mail_ptrs[0]) cudaFree(R->d_mail_ptrs[0]);
  if (R->d_mail_ptrs[1]) cudaFree(R->d_mail_ptrs[1]);
  if (R->d_sigma) cudaFree(R->d_sigma);
  if (R->d_src) cudaFree(R->d_src);
  if (R->d_probe_pairs) cudaFree(R->d_probe_pairs);
  for (BlockProbe &pb : R->probe_store) {
    if (pb.before) cudaEventDestroy(pb.before);
    if (pb.after) cudaEventDestroy(pb.after);
  }
  if (R->stream) cudaStreamDestroy(R->stream);
  cudaGetLastError();
  delete R;
}

static size_t round_up(size_t x, size_t k) { return (x + k - 1) / k * k; }

// columns are padded with zero columns to a whole number of 32-column blocks per travelling block
static size_t padded_cols(size_t n, int P) { return round_up(n, 64 * (size_t)P); }

// local allocations of a rank (no communication)
static int mg_alloc_local(MgRank *R) {
  const int P = R->P;
  if (P > 32) {
    set_error("svdj_mg: at most 32 ranks");
    return SVDJ_ERR_UNSUPPORTED;
  }
  if (cudaGetDevice(&R->device) != cudaSuccess) {
    set_error("svdj_mg: no CUDA device");
    cudaGetLastError();
    return SVDJ_ERR_CUDA;
  }
  const size_t npad = padded_cols(R->max_n, P);
  R->w_max = npad / (2 * (size_t)P);
  R->lda = round_up(std::max(R->max_m, (size_t)2), 16);
  R->ldv = round_up(npad, 16);
  const size_t buf_doubles = R->w_max * (R->lda + R->ldv);
  R->ctl_words = R->sigma_off() + (size_t)P * (2 * R->w_max + 2);
  cudaError_t e = cudaStreamCreateWithFlags(&R->stream, cudaStreamNonBlocking);
  for (int i = 0; i < kBufs && e == cudaSuccess; ++i) {
    e = cudaMalloc(&R->buf[i], sizeof(double) * buf_doubles);
    if (e == cudaSuccess) e = cudaMemset(R->buf[i], 0, sizeof(double) * buf_doubles);
  }
  if (e == cudaSuccess) e = cudaMalloc(&R->ctl, sizeof(u64) * R->ctl_words);
  if (e == cudaSuccess) e = cudaMemset(R->ctl, 0, sizeof(u64) * R->ctl_words);
  if (e == cudaSuccess) e = cudaMalloc(&R->d_counters, sizeof(SweepCounters));
  if (e == cudaSuccess) e = cudaMalloc(&R->d_out, 2 * sizeof(u64));
  if (e == cudaSuccess) e = cudaMallocHost(&R->h_out, 2 * sizeof(u64));
  if (e == cudaSuccess) e = cudaMalloc(&R->d_mail_ptrs[0], sizeof(u64 *) * P);
  if (e == cudaSuccess) e = cudaMalloc(&R->d_mail_ptrs[1], sizeof(u64 *) * P);
  if (e == cudaSuccess) e = cudaMalloc(&R->d_sigma, sizeof(double) * 2 * R->w_max);
  if (e == cudaSuccess && R->rank == 0) e = cudaMalloc(&R->d_src, sizeof(double *) * 2 * npad);
  if (e == cudaSuccess) e = cudaMalloc(&R->d_probe_pairs, sizeof(u64));
  R->probe_store.assign(MgRank::kProbes, BlockProbe());
  for (int i = 0; i < MgRank::kProbes && e == cudaSuccess; ++i) {
    e = cudaEventCreate(&R->probe_store[i].before);
    if (e == cudaSuccess) e = cudaEventCreate(&R->probe_store[i].after);
    R->probe_store[i].active_pairs = R->d_probe_pairs;
  }
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("svdj_mg: set-up allocation failed on rank %d: %s", R->rank, cudaGetErrorString(e));
    cudaGetLastError();
    return SVDJ_ERR_ALLOC;
  }
  return SVDJ_OK;
}

// maps every rank's four block buffers and control block: raw pointers inside one process (peer access between the
// devices), CUDA IPC handles across processes
static int mg_map_peers(MgRank *R, const std::vector<PeerBlob> &all) {
  const int P = R->P;
  const int my_pid = (int)getpid();
  R->peer_buf.assign((size_t)P, std::vector<double *>(kBufs, nullptr));
  R->peer_ctl.assign((size_t)P, nullptr);
  cudaError_t e = cudaSuccess;
  for (int q = 0; q < P; ++q) {
    if (q == R->rank) {
      for (int i = 0; i < kBufs; ++i) R->peer_buf[q][i] = R->buf[i];
      R->peer_ctl[q] = R->ctl;
      continue;
    }
    if (all[q].pid == my_pid) {
      if (all[q].device != R->device) {
        e = cudaDeviceEnablePeerAccess(all[q].device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) {
          cudaGetLastError();
          e = cudaSuccess;
        }
        if (e != cudaSuccess) {
          set_error("svdj_mg: cudaDeviceEnablePeerAccess(%d) from device %d -> %s", all[q].device, R->device,
                    cudaGetErrorString(e));
          cudaGetLastError();
          return SVDJ_ERR_CUDA;
        }
      }
      for (int i = 0; i < kBufs; ++i) R->peer_buf[q][i] = (double *)(uintptr_t)all[q].ptr[i];
      R->peer_ctl[q] = (u64 *)(uintptr_t)all[q].ptr[kBufs];
    } else {
      for (int i = 0; i <= kBufs; ++i) {
        void *p = nullptr;
        e = cudaIpcOpenMemHandle(&p, all[q].ipc[i], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
          set_error("svdj_mg: cudaIpcOpenMemHandle(rank %d, buffer %d) -> %s", q, i, cudaGetErrorString(e));
          cudaGetLastError();
          return SVDJ_ERR_CUDA;
        }
        R->ipc_opened.push_back(p);
        if (i < kBufs) R->peer_buf[q][i] = (double *)p;
        else R->peer_ctl[q] = (u64 *)p;
      }
    }
  }
  for (int par = 0; par < 2; ++par) {
    std::vector<u64 *> ptrs((size_t)P);
    for (int q = 0; q < P; ++q) ptrs[q] = R->peer_ctl[q] + R->mail_off(par);
    e = cudaMemcpy(R->d_mail_ptrs[par], ptrs.data(), sizeof(u64 *) * P, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      set_error("svdj_mg: cudaMemcpy -> %s", cudaGetErrorString(e));
      cudaGetLastError();
      return SVDJ_ERR_CUDA;
    }
  }
  return SVDJ_OK;
}

// Collective.  A rank whose local part fails still takes part in both all-gathers, so that the others learn of it
// and return an error instead of waiting for ever.
static int mg_create(std::unique_ptr<Comm> comm, size_t max_m, size_t max_n, MgRank **out) {
  *out = nullptr;
  MgRank *R = new MgRank();
  R->rank = comm->rank;
  R->P = comm->size;
  R->comm = std::move(comm);
  R->max_m = max_m;
  R->max_n = max_n;
  const int P = R->P;
  int rc = (max_m == 0 || max_n == 0 || max_m > 0x7fffffffu || max_n > 0x7fffffffu) ? -2 : mg_alloc_local(R);
  PeerBlob mine;
  std::memset(&mine, 0, sizeof(mine));
  mine.status = rc;
  mine.pid = (int)getpid();
  mine.device = R->device;
  if (rc == SVDJ_OK) {
    for (int i = 0; i <= kBufs; ++i) {
      void *p = (i < kBufs) ? (void *)R->buf[i] : (void *)R->ctl;
      mine.ptr[i] = (u64)(uintptr_t)p;
      if (cudaIpcGetMemHandle(&mine.ipc[i], p) != cudaSuccess) {
        cudaGetLastError();  // only the multi-process mode needs the handles; opening a zero handle fails there
        std::memset(&mine.ipc[i], 0, sizeof(mine.ipc[i]));
      }
    }
  }
  std::vector<PeerBlob> all((size_t)P);
  if (R->comm->allgather(&mine, all.data(), sizeof(PeerBlob)) != 0) {
    set_error("svdj_mg: the communicator's all-gather failed");
    mg_destroy(R);
    return SVDJ_ERR_CUDA;
  }
  for (int q = 0; q < P; ++q) {
    if (all[q].status != SVDJ_OK) {
      if (q != R->rank) set_error("svdj_mg: rank %d failed during set-up (status %d)", q, all[q].status);
      const int bad = all[q].status;
      mg_destroy(R);
      return rc != SVDJ_OK ? rc : bad;
    }
  }
  {
    std::vector<int> devs((size_t)P);
    for (int q = 0; q < P; ++q) devs[q] = (all[q].pid == mine.pid) ? all[q].device : -1 - q;  // other processes: never shared
    R->comm->set_devices(devs);
  }
  int map_rc = mg_map_peers(R, all);
  std::vector<int> map_all((size_t)P);
  if (R->comm->allgather(&map_rc, map_all.data(), sizeof(int)) != 0) map_rc = map_rc ? map_rc : SVDJ_ERR_CUDA;
  for (int q = 0; q < P && map_rc == SVDJ_OK; ++q) {
    if (map_all[q] != SVDJ_OK) {
      set_error("svdj_mg: rank %d could not map its peers (status %d)", q, map_all[q]);
      map_rc = map_all[q];
    }
  }
  if (map_rc != SVDJ_OK) {
    mg_destroy(R);
    return map_rc;
  }
  *out = R;
  return SVDJ_OK;
}

static int mg_workspace(MgRank *R, size_t m, size_t ncols, void **ws) {
  auto key = std::make_pair(m, ncols);
  auto it = R->workspaces.find(key);
  if (it != R->workspaces.end()) {
    *ws = it->second;
    return SVDJ_OK;
  }
  void *w = nullptr;
  if (int rc = svdj_ring_ws_create(m, ncols, &w)) return rc;
  R->workspaces[key] = w;
  *ws = w;
  return SVDJ_OK;
}

// The collective solve on device-resident data: rank 0 passes dA (m x n, lda_in), ds, dV (n x n, ldv_in) on ITS
// device; the other ranks pass null pointers and the same sizes and options.
static int mg_solve_device(MgRank *R, int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda_in, double *ds,
                           double *dV, size_t ldv_in, const svdj_opts &o, svdj_stats &st, int pre_status = SVDJ_OK) {
  const int P = R->P, me = R->rank;
  const bool root = (me == 0);
  const bool want_u = jobu != SVDJ_NOVEC, want_v = jobv != SVDJ_NOVEC;
  const size_t npad = padded_cols(n, P);
  const size_t w = npad / (2 * (size_t)P);
  if (m > R->lda || w > R->w_max || npad > R->ldv) {
    set_error("svdj_mg: %zu x %zu exceeds the sizes this handle was created for (%zu x %zu)", m, n, R->max_m, R->max_n);
    return SVDJ_ERR_UNSUPPORTED;
  }
  cudaStream_t s = R->stream;
  const auto t_begin = std::chrono::steady_clock::now();
  // every allocation of this solve happens here, before anybody launches (Comm::before_first_launch)
  void *ws = nullptr;
  Event ev0, ev1;
  DevBuf d_norms;
  int pre = pre_status;
  if (pre == SVDJ_OK) pre = mg_workspace(R, m, 2 * w, &ws);
  if (pre == SVDJ_OK) pre = ev0.create();
  if (pre == SVDJ_OK) pre = ev1.create();
  if (pre == SVDJ_OK && root) pre = d_norms.alloc(sizeof(double) * n);
  const int agreed = R->comm->before_first_launch(pre);
  if (agreed != SVDJ_OK) {
    if (pre == SVDJ_OK) set_error("svdj_mg: another rank failed before the solve started (status %d)", agreed);
    return pre != SVDJ_OK ? pre : agreed;
  }
  const u64 epoch = ++R->epoch;
  u64 *err = R->ctl + 80;
  unsigned long long launches = 0;

  // ---- scatter (root): floor, blocks, start flag ----
  std::vector<Role> roles = initial_roles(P);
  if (root) {
    // numerically-zero floor: tol * largest column norm of the input (as in the single-GPU driver)
    MG_TRY(launch_column_norms(dA, lda_in, m, n, d_norms.as<double>(), s));
    std::vector<double> hn(n);
    SVDJ_CUDA_TRY(cudaMemcpyAsync(hn.data(), d_norms.p, sizeof(double) * n, cudaMemcpyDeviceToHost, s));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
    double mx = 0.0;
    for (double v : hn) mx = std::max(mx, v);
    const double fl = o.tol * mx;
    u64 fl_bits;
    std::memcpy(&fl_bits, &fl, sizeof(fl_bits));
    for (int q = 0; q < P; ++q) {
      for (int slot = 0; slot < 2; ++slot) {
        const size_t bid = slot == 0 ? (size_t)q : (size_t)(P + q);
        const size_t c0 = bid * w;
        const unsigned valid = c0 < n ? (unsigned)std::min(w, n - c0) : 0u;
        double *dst = R->peer_buf[q][slot == 0 ? roles[q].top : roles[q].bottom];
        mg_copy_block_kernel<<<(unsigned)w, 256, 0, s>>>(dA + c0 * lda_in, lda_in, dst, R->lda, (unsigned)m, valid);
        ++launches;
      }
    }
    for (int q = 0; q < P; ++q) {
      mg_start_kernel<<<1, 1, 0, s>>>(R->peer_ctl[q], fl_bits, epoch);
      R->comm->publish(kChanStart, q, epoch, s);
      ++launches;
    }
    SVDJ_CUDA_TRY(cudaGetLastError());
  }
  // own V blocks: the identity's columns of the two blocks held at round 0
  if (want_v || want_u) {  // (the V part travels with the block in any case; it is only rotated when wanted)
    mg_init_v_kernel<<<(unsigned)w, 256, 0, s>>>(R->V_of(R->buf[roles[me].top]), R->ldv, (unsigned)npad,
                                                 (unsigned)((size_t)me * w));
    mg_init_v_kernel<<<(unsigned)w, 256, 0, s>>>(R->V_of(R->buf[roles[me].bottom]), R->ldv, (unsigned)npad,
                                                 (unsigned)((size_t)(P + me) * w));
    launches += 2;
  }
  R->comm->await(kChanStart, 0, epoch, s);
  mg_wait_kernel<<<1, 1, 0, s>>>(R->ctl + 32, nullptr, epoch, err, kTimeoutNs, 3);
  ++launches;
  SVDJ_CUDA_TRY(cudaGetLastError());
  // the floor arrives with the start flag
  u64 floor_bits = 0;
  SVDJ_CUDA_TRY(cudaMemcpyAsync(&floor_bits, R->ctl + 64, sizeof(u64), cudaMemcpyDeviceToHost, s));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
  double floor_norm = 0.0;
  std::memcpy(&floor_norm, &floor_bits, sizeof(double));
  MG_TRY(svdj_ring_ws_set_floor(ws, floor_norm));

  // ---- sweeps ----
  ProbeSet ps;
  ps.probes = R->probe_store.data();
  ps.count = (int)R->probe_store.size();
  double probe_ms = 0.0;
  unsigned long long probe_launches = 0;
  SVDJ_CUDA_TRY(cudaMemsetAsync(R->d_probe_pairs, 0, sizeof(u64), s));
  SVDJ_CUDA_TRY(cudaEventRecord(ev0.e, s));
  const int rounds = 2 * P - 1;
  const bool rotate_v = want_v;
  const bool trace = std::getenv("SVDJ_MG_TRACE") != nullptr;
  bool converged = false;
  std::vector<Role> next;
  std::vector<Transfer> tr;
  // Inner sweeps of the small eigen-solves.  On the ring a second inner sweep does not pay up to n = 8192 (2 GPUs:
  // 4096^2 0.249 s with one, 0.265 s with two; 8192^2 1.46 s against 1.55 s -- it saves at most one outer sweep
  // there, three on one GPU); 16384^2 was measured with two and stays there.
  int inner = (m >= 8 * n) ? 3 : (n >= 16384) ? 2 : 1;
  if (const char *env = std::getenv("SVDJ_INNER_SWEEPS")) {
    if (std::atoi(env) > 0) inner = std::atoi(env);
  }
  // Ordering the columns of a block pair by size (as on one GPU) is off: on the ring it needs the second inner
  // sweep to save outer sweeps (4096^2, 8 ranks: 14 instead of 17) and the two together cost what they save
  // (2 GPUs, 4096^2: 0.251 s against 0.249 s; 8192^2: 1.52 s against 1.46 s).  SVDJ_MG_SORT=1 turns it on.
  int sort_pairs = 0;
  if (const char *env = std::getenv("SVDJ_MG_SORT")) sort_pairs = std::atoi(env) != 0;
  std::vector<int> id_top, id_bot;
  arrangement(P, 0, id_top, id_bot);
  for (int sweep = 0; sweep < o.max_sweeps; ++sweep) {
    SVDJ_CUDA_TRY(cudaMemsetAsync(R->d_counters, 0, sizeof(SweepCounters), s));
    for (int rd = 0; rd < rounds; ++rd) {
      advance_roles(P, roles, next, tr);
      double *dst_top = nullptr, *dst_bot = nullptr;
      for (const Transfer &t : tr)
        if (t.src_rank == me) (t.src_buf == roles[me].top ? dst_top : dst_bot) = R->peer_buf[t.dst_rank][t.dst_buf];
      double *top = R->buf[roles[me].top], *bot = R->buf[roles[me].bottom];
      // with the ordering on, the block with the lower global index goes in as "left": the eigen-solve then sends
      // the larger columns of every pair towards the lower index, the same direction in every round on every rank
      // (without the ordering the roles stay as they are: 4096^2 on 8 ranks took 18 sweeps swapped, 17 unswapped)
      if (sort_pairs && id_top[me] > id_bot[me]) {
        std::swap(top, bot);
        std::swap(dst_top, dst_bot);
      }
      MG_TRY(ring_phase(ws, m, npad, w, R->A_of(top), R->A_of(bot), R->lda, R->V_of(top), R->V_of(bot), R->ldv,
                        rotate_v ? 1 : 0, rd == 0 ? SVDJ_PHASE_ALL : SVDJ_PHASE_CROSS, o.tol, inner,
                        reinterpret_cast<unsigned long long *>(R->d_counters), s, dst_top ? R->A_of(dst_top) : nullptr,
                        dst_bot ? R->A_of(dst_bot) : nullptr, dst_top ? R->V_of(dst_top) : nullptr,
                        dst_bot ? R->V_of(dst_bot) : nullptr, P > 1, &launches, &ps, sort_pairs));
      move_ids(P, id_top, id_bot);
      if (P > 1) {
        const u64 stamp = ++R->round_stamp;
        u64 *sig[2] = {nullptr, nullptr};
        const u64 *wt[2] = {nullptr, nullptr};
        int ns = 0, nw = 0;
        for (const Transfer &t : tr) {  // at most one block to, and one from, each of the two neighbours
          if (t.src_rank == me && ns < 2) sig[ns++] = R->peer_ctl[t.dst_rank] + (me < t.dst_rank ? 0 : 16);
          if (t.dst_rank == me && nw < 2) wt[nw++] = R->ctl + (t.src_rank < me ? 0 : 16);
        }
        mg_signal_kernel<<<1, 1, 0, s>>>(sig[0], sig[1], stamp);
        for (const Transfer &t : tr)
          if (t.src_rank == me) R->comm->publish(kChanRound, t.dst_rank, stamp, s);
        for (const Transfer &t : tr)
          if (t.dst_rank == me) R->comm->await(kChanRound, t.src_rank, stamp, s);
        mg_wait_kernel<<<1, 1, 0, s>>>(wt[0], wt[1], stamp, err, kTimeoutNs);
        launches += 2;
      }
      roles.swap(next);
      if (trace) std::fprintf(stderr, "[svdj mg] rank %d queued sweep %d round %d (stamp %llu)\n", me, sweep + 1, rd,
                              (unsigned long long)R->round_stamp);
    }
    // convergence: every rank posts its counters to every rank, all take the same decision
    const u64 sstamp = ++R->sweep_stamp;
    const int par = (int)(sstamp & 1);
    mg_post_counters_kernel<<<1, 32, 0, s>>>(R->d_counters, R->d_mail_ptrs[par], me, P, sstamp);
    for (int q = 0; q < P; ++q) R->comm->publish(kChanMail, q, sstamp, s);
    for (int q = 0; q < P; ++q) R->comm->await(kChanMail, q, sstamp, s);
    mg_collect_kernel<<<1, 1, 0, s>>>(R->ctl + R->mail_off(par), P, sstamp, R->d_out, err, kTimeoutNs);
    launches += 2;
    SVDJ_CUDA_TRY(cudaMemcpyAsync(R->h_out, R->d_out, 2 * sizeof(u64), cudaMemcpyDeviceToHost, s));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
    if (R->h_out[0] == ~0ull) {
      u64 code = 0;
      cudaMemcpyAsync(&code, err, sizeof(u64), cudaMemcpyDeviceToHost, s);
      cudaStreamSynchronize(s);
      cudaMemsetAsync(err, 0, sizeof(u64), s);
      set_error("svdj_mg: rank %d waited more than %llu s for its peers in sweep %d (a peer failed or never called; "
                "wait kind %llu, index %llu, stamp %llu; 1 = round flag, 2 = sweep mailbox, 3 = start)",
                me, (unsigned long long)(kTimeoutNs / 1000000000ull), sweep + 1, code >> 56, (code >> 48) & 0xffull,
                code & 0xffffffffffffull);
      return SVDJ_ERR_CUDA;
    }
    for (int i = 0; i < ps.used; ++i) {
      float pms = 0.f;
      if (cudaEventElapsedTime(&pms, ps.probes[i].before, ps.probes[i].after) == cudaSuccess) probe_ms += pms;
    }
    probe_launches += (unsigned long long)ps.used;
    ps.used = 0;
    st.sweeps = sweep + 1;
    st.applied += R->h_out[0];
    double off;
    std::memcpy(&off, &R->h_out[1], sizeof(double));
    st.last_offdiag = off;
    st.tested += (unsigned long long)npad * (npad - 1) / 2;
    if (trace) std::fprintf(stderr, "[svdj mg] rank %d sweep %d done\n", me, sweep + 1);
    if (std::getenv("SVDJ_TRACE") != nullptr && root)
      std::fprintf(stderr, "[svdj mg] sweep %2d: rotations %llu, max off-diagonal cosine %.3e\n", sweep + 1,
                   (unsigned long long)R->h_out[0], off);
    if (R->h_out[0] == 0 || off <= 2.0 * o.tol) {  // as in the single-GPU driver (svdj_api.cu)
      converged = true;
      break;
    }
  }
  SVDJ_CUDA_TRY(cudaEventRecord(ev1.e, s));
  st.converged = converged ? 1 : 0;

  // ---- sigma of the local columns, U = A / sigma, and the sigmas to the root ----
  double *top = R->buf[roles[me].top], *bot = R->buf[roles[me].bottom];
  MG_TRY(svdj_ring_finalize_device(ws, m, w, R->A_of(top), R->lda, R->d_sigma, want_u ? 1 : 0, s));
  MG_TRY(svdj_ring_finalize_device(ws, m, w, R->A_of(bot), R->lda, R->d_sigma + w, want_u ? 1 : 0, s));
  launches += want_u ? 4 : 2;
  double *root_mail = reinterpret_cast<double *>(R->peer_ctl[0] + R->sigma_off()) + (size_t)me * (2 * R->w_max + 2);
  mg_post_sigma_kernel<<<1, 256, 0, s>>>(R->d_sigma, (unsigned)(2 * w), root_mail, epoch);
  R->comm->publish(kChanSigma, 0, epoch, s);
  ++launches;
  SVDJ_CUDA_TRY(cudaGetLastError());

  if (root) {
    // wait for everybody's sigmas (the stamp word of each slot), sort, gather in sigma order
    const size_t slot = 2 * R->w_max + 2;
    double *mail = reinterpret_cast<double *>(R->ctl + R->sigma_off());
    for (int q = 0; q < P; ++q) {
      R->comm->await(kChanSigma, q, epoch, s);
      mg_wait_kernel<<<1, 1, 0, s>>>(reinterpret_cast<const u64 *>(mail + (size_t)q * slot + 2 * w), nullptr, epoch, err,
                                     kTimeoutNs, 4);
      ++launches;
    }
    std::vector<double> hs((size_t)P * slot);
    SVDJ_CUDA_TRY(cudaMemcpyAsync(hs.data(), mail, sizeof(double) * hs.size(), cudaMemcpyDeviceToHost, s));
    u64 herr = 0;
    SVDJ_CUDA_TRY(cudaMemcpyAsync(&herr, err, sizeof(u64), cudaMemcpyDeviceToHost, s));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
    if (herr != 0) {
      set_error("svdj_mg: the root timed out waiting for the ranks' singular values");
      return SVDJ_ERR_CUDA;
    }
    // which block sits where: the arrangement after all the moves made so far in this solve
    std::vector<int> atop, abot;
    arrangement(P, 0, atop, abot);  // whole sweeps only: 2P-1 moves bring every block back to where it started
    struct Col {
      double sigma;
      unsigned id;
      const double *a, *v;
    };
    std::vector<Col> cols;
    cols.reserve(n);
    for (int q = 0; q < P; ++q) {
      for (int half = 0; half < 2; ++half) {
        const int bid = half == 0 ? atop[q] : abot[q];
        const double *b = R->peer_buf[q][half == 0 ? roles[q].top : roles[q].bottom];
        for (size_t j = 0; j < w; ++j) {
          const size_t id = (size_t)bid * w + j;
          if (id >= n) continue;  // padding column
          cols.push_back({hs[(size_t)q * slot + (size_t)half * w + j], (unsigned)id, b + j * R->lda,
                          b + R->w_max * R->lda + j * R->ldv});
        }
      }
    }
    if (o.sort_desc) {
      std::stable_sort(cols.begin(), cols.end(), [](const Col &x, const Col &y) { return x.sigma > y.sigma; });
    } else {
      std::sort(cols.begin(), cols.end(), [](const Col &x, const Col &y) { return x.id < y.id; });
    }
    std::vector<const double *> src(2 * n);
    std::vector<double> sig(n);
    for (size_t k = 0; k < n; ++k) {
      src[k] = cols[k].a;
      src[n + k] = cols[k].v;
      sig[k] = cols[k].sigma;
    }
    SVDJ_CUDA_TRY(cudaMemcpyAsync(R->d_src, src.data(), sizeof(double *) * 2 * n, cudaMemcpyHostToDevice, s));
    SVDJ_CUDA_TRY(cudaMemcpyAsync(ds, sig.data(), sizeof(double) * std::min(m, n), cudaMemcpyHostToDevice, s));
    mg_gather_kernel<<<(unsigned)n, 256, 0, s>>>(R->d_src, dA, lda_in, (unsigned)m);
    ++launches;
    if (want_v) {
      mg_gather_kernel<<<(unsigned)n, 256, 0, s>>>(R->d_src + n, dV, ldv_in, (unsigned)n);
      ++launches;
    }
    for (int q = 0; q < P; ++q) {
      mg_signal_kernel<<<1, 1, 0, s>>>(R->peer_ctl[q] + 48, nullptr, epoch);
      R->comm->publish(kChanDone, q, epoch, s);
      ++launches;
    }
    SVDJ_CUDA_TRY(cudaGetLastError());
    SVDJ_CUDA_TRY(cudaStreamSynchronize(s));  // src / sig must outlive their copies
  }
  // nobody reuses its buffers before the root has read them
  R->comm->await(kChanDone, 0, epoch, s);
  mg_wait_kernel<<<1, 1, 0, s>>>(R->ctl + 48, nullptr, epoch, err, kTimeoutNs, 5);
  ++launches;
  u64 herr = 0;
  SVDJ_CUDA_TRY(cudaMemcpyAsync(&herr, err, sizeof(u64), cudaMemcpyDeviceToHost, s));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
  if (herr != 0) {
    set_error("svdj_mg: rank %d timed out waiting for a peer (stamp %llu)", me, herr);
    SVDJ_CUDA_TRY(cudaMemsetAsync(err, 0, sizeof(u64), s));
    return SVDJ_ERR_CUDA;
  }
  float ms = 0.f;
  SVDJ_CUDA_TRY(cudaEventElapsedTime(&ms, ev0.e, ev1.e));
  st.seconds_sweeps = ms * 1e-3;
  {  // the update kernel on THIS rank: mean probed launch and the block pairs those launches updated
    u64 pairs = 0;
    SVDJ_CUDA_TRY(cudaMemcpyAsync(&pairs, R->d_probe_pairs, sizeof(u64), cudaMemcpyDeviceToHost, s));
    SVDJ_CUDA_TRY(cudaStreamSynchronize(s));
    st.step_kernel_ms = probe_launches ? probe_ms / (double)probe_launches : 0.0;
    st.step_kernel_launches = probe_launches;
    st.probe_block_pairs = pairs;
  }
  st.launches += launches;
  st.variant = SVDJ_VARIANT_BLOCK;
  st.reserved = (w % 32 == 0) ? 32 : 16;
  st.seconds_total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  if (!converged) return SVDJ_NOT_CONVERGED;
  return SVDJ_OK;
}

}  // namespace svdj

using namespace svdj;

// ---------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------
static svdj_opts mg_opts(const svdj_opts *in, size_t m, int jobu, int jobv) {
  svdj_opts o;
  svdj_default_opts(&o);
  if (in) o = *in;
  if (o.max_sweeps <= 0) o.max_sweeps = 60;
  if (o.tol <= 0.0) o.tol = default_tol(m, jobu, jobv);
  return o;
}

extern "C" {

int svdj_mg_rank_create(const svdj_mg_comm *comm, size_t max_m, size_t max_n, void **handle) {
  clear_error();
  if (handle == nullptr) return -4;
  *handle = nullptr;
  if (comm == nullptr || comm->size < 1 || comm->rank < 0 || comm->rank >= comm->size ||
      (comm->size > 1 && comm->allgather == nullptr))
    return -1;
  std::unique_ptr<Comm> c;
  if (comm->size == 1) {  // a one-rank communicator needs no callback
    struct Self : Comm {
      int allgather(const void *send, void *recv, size_t bytes) override {
        std::memcpy(recv, send, bytes);
        return 0;
      }
    };
    c.reset(new Self());
  } else {
    CallbackComm *cc = new CallbackComm();
    cc->fn = comm->allgather;
    cc->ctx = comm->ctx;
    c.reset(cc);
  }
  c->rank = comm->rank;
  c->size = comm->size;
  MgRank *R = nullptr;
  const int rc = mg_create(std::move(c), max_m, max_n, &R);
  *handle = R;
  return rc;
}

int svdj_mg_plan(int P, int round, int *top_block, int *bottom_block, int *top_buf, int *bottom_buf, int *transfers,
                 int *ntransfers) {
  if (P < 1 || round < 0) return -1;
  std::vector<Role> roles = initial_roles(P), next;
  std::vector<Transfer> tr;
  for (int it = 0; it < round; ++it) {
    advance_roles(P, roles, next, tr);
    roles.swap(next);
  }
  std::vector<int> top, bottom;
  arrangement(P, round, top, bottom);
  advance_roles(P, roles, next, tr);
  for (int r = 0; r < P; ++r) {
    if (top_block) top_block[r] = top[r];
    if (bottom_block) bottom_block[r] = bottom[r];
    if (top_buf) top_buf[r] = roles[r].top;
    if (bottom_buf) bottom_buf[r] = roles[r].bottom;
  }
  if (ntransfers) *ntransfers = (int)tr.size();
  for (size_t i = 0; transfers != nullptr && i < tr.size(); ++i) {
    transfers[4 * i + 0] = tr[i].src_rank;
    transfers[4 * i + 1] = tr[i].src_buf;
    transfers[4 * i + 2] = tr[i].dst_rank;
    transfers[4 * i + 3] = tr[i].dst_buf;
  }
  return SVDJ_OK;
}

int svdj_mg_rank_destroy(void *handle) {
  mg_destroy(static_cast<MgRank *>(handle));
  return SVDJ_OK;
}

static int rank_solve_device(MgRank *R, int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds,
                             double *dV, size_t ldv, const svdj_opts *opts, svdj_stats &st, int pre_status) {
  std::memset(&st, 0, sizeof(st));
  if (m == 0 || n == 0) return SVDJ_OK;
  int cur = -1;
  SVDJ_CUDA_TRY(cudaGetDevice(&cur));
  if (cur != R->device) SVDJ_CUDA_TRY(cudaSetDevice(R->device));
  const svdj_opts o = mg_opts(opts, m, jobu, jobv);
  // QR front-end (opts.precondition = 1 or 2; "library decides" means none here): rank 0 factors on its GPU, the
  // ring iterates on the n x n triangular factor, rank 0 maps the result back.  Every rank works out the same
  // inner sizes from the same arguments.
  svdj_opts oq = o;
  if (oq.precondition < 0) oq.precondition = 0;
  const int steps = precondition_steps(oq, m, n);
  int rc;
  if (steps == 0) {
    rc = mg_solve_device(R, jobu, jobv, m, n, dA, lda, ds, dV, ldv, o, st, pre_status);
  } else {
    const size_t rows_l = (n + 1) & ~size_t(1);
    const bool want_u = jobu != SVDJ_NOVEC, want_v = jobv != SVDJ_NOVEC;
    const bool need_j = steps >= 2 ? want_v : want_u, need_w = steps >= 2 ? want_u : want_v;
    if (R->rank == 0 && pre_status == SVDJ_OK) {
      rc = solve_with_qr(jobu, jobv, m, n, dA, lda, ds, dV, ldv, steps, oq, st, R->stream,
                         [R](int pre, int ju, int jv, size_t mi, size_t ni, double *A, size_t la, double *s, double *V,
                             size_t lv, size_t, const svdj_opts &oi, svdj_stats &sti) -> int {
                           return mg_solve_device(R, ju, jv, mi, ni, A, la, s, V, lv, oi, sti, pre);
                         });
    } else {
      rc = mg_solve_device(R, need_w ? SVDJ_ALLVEC : SVDJ_NOVEC, need_j ? SVDJ_ALLVEC : SVDJ_NOVEC, rows_l, n, nullptr, 0,
                           nullptr, nullptr, 0, o, st, pre_status);
    }
  }
  if (cur != R->device && cur >= 0) cudaSetDevice(cur);
  return rc;
}

// argument errors are reported through the collective too (pre_status), so that in-process peers do not wait
static int root_arg_status(int jobu, int jobv, size_t m, size_t n, const double *A, size_t lda, const double *s,
                           const double *V, size_t ldv) {
  if (A == nullptr) return -6;
  if (lda < m) return -7;
  if (s == nullptr) return -8;
  if (jobv != SVDJ_NOVEC && (V == nullptr || ldv < n)) return -9;
  (void)jobu;
  return 0;
}

int svdj_mg_rank_dgesvd_device(void *handle, int jobu, int jobv, size_t m, size_t n, double *dA, size_t lda, double *ds,
                               double *dV, size_t ldv, const svdj_opts *opts, svdj_stats *stats) {
  clear_error();
  MgRank *R = static_cast<MgRank *>(handle);
  if (R == nullptr) return -1;
  if (jobu < SVDJ_ALLVEC || jobu > SVDJ_NOVEC) return -2;
  if (jobv < SVDJ_ALLVEC || jobv > SVDJ_NOVEC) return -3;
  svdj_stats local;
  svdj_stats &st = stats ? *stats : local;
  int pre = SVDJ_OK;
  if (R->rank == 0 && m * n > 0) {
    pre = root_arg_status(jobu, jobv, m, n, dA, lda, ds, dV, ldv);
    if (pre) set_error("svdj_mg_rank_dgesvd_device: argument %d is invalid", -pre);
  }
  return rank_solve_device(R, jobu, jobv, m, n, dA, lda, ds, dV, ldv, opts, st, pre);
}

int svdj_mg_rank_dgesvd(void *handle, int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s,
                        double *V, size_t ldv, const svdj_opts *opts, svdj_stats *stats) {
  clear_error();
  MgRank *R = static_cast<MgRank *>(handle);
  if (R == nullptr) return -1;
  if (jobu < SVDJ_ALLVEC || jobu > SVDJ_NOVEC) return -2;
  if (jobv < SVDJ_ALLVEC || jobv > SVDJ_NOVEC) return -3;
  svdj_stats local;
  svdj_stats &st = stats ? *stats : local;
  if (R->rank != 0) return rank_solve_device(R, jobu, jobv, m, n, nullptr, 0, nullptr, nullptr, 0, opts, st, SVDJ_OK);
  if (m == 0 || n == 0) {
    std::memset(&st, 0, sizeof(st));
    return SVDJ_OK;
  }
  const bool want_v = jobv != SVDJ_NOVEC;
  int cur = -1;
  SVDJ_CUDA_TRY(cudaGetDevice(&cur));
  if (cur != R->device) SVDJ_CUDA_TRY(cudaSetDevice(R->device));
  struct Restore {
    int cur, dev;
    ~Restore() {
      if (cur != dev && cur >= 0) cudaSetDevice(cur);
    }
  } restore{cur, R->device};
  const size_t ld_d = (m + 15) & ~size_t(15), ldv_d = (n + 15) & ~size_t(15);
  DevBuf dA, dV, dS;
  // a failure here still enters the collective (with its status) so that the other ranks return too
  int pre = root_arg_status(jobu, jobv, m, n, A, lda, s, V, ldv);
  if (pre) set_error("svdj_mg_rank_dgesvd: argument %d is invalid", -pre);
  if (pre == SVDJ_OK) pre = dA.alloc(sizeof(double) * ld_d * n);
  if (pre == SVDJ_OK) pre = dS.alloc(sizeof(double) * n);
  if (pre == SVDJ_OK && want_v) pre = dV.alloc(sizeof(double) * ldv_d * n);
  const auto t0 = std::chrono::steady_clock::now();
  if (pre == SVDJ_OK) {
    cudaError_t e = cudaMemcpy2DAsync(dA.p, ld_d * sizeof(double), A, lda * sizeof(double), m * sizeof(double), n,
                                      cudaMemcpyHostToDevice, R->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(R->stream);
    if (e != cudaSuccess) {
      set_error("svdj_mg_rank_dgesvd: host -> device copy failed: %s", cudaGetErrorString(e));
      cudaGetLastError();
      pre = SVDJ_ERR_CUDA;
    }
  }
  const double h2d = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  const int rc = rank_solve_device(R, jobu, jobv, m, n, dA.as<double>(), ld_d, dS.as<double>(), dV.as<double>(), ldv_d,
                                   opts, st, pre);
  if (rc != SVDJ_OK && rc != SVDJ_NOT_CONVERGED) return rc;
  const auto t1 = std::chrono::steady_clock::now();
  SVDJ_CUDA_TRY(cudaMemcpy2DAsync(A, lda * sizeof(double), dA.p, ld_d * sizeof(double), m * sizeof(double), n,
                                  cudaMemcpyDeviceToHost, R->stream));
  SVDJ_CUDA_TRY(cudaMemcpyAsync(s, dS.p, sizeof(double) * std::min(m, n), cudaMemcpyDeviceToHost, R->stream));
  if (want_v)
    SVDJ_CUDA_TRY(cudaMemcpy2DAsync(V, ldv * sizeof(double), dV.p, ldv_d * sizeof(double), n * sizeof(double), n,
                                    cudaMemcpyDeviceToHost, R->stream));
  SVDJ_CUDA_TRY(cudaStreamSynchronize(R->stream));
  st.seconds_h2d = h2d;
  st.seconds_d2h = std::chrono::duration<double>(std::chrono::steady_clock::now() - t1).count();
  st.seconds_total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return rc;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// single process, N devices: the ranks are threads; the handles are kept between calls
// ---------------------------------------------------------------------------------------------------------
namespace svdj {
struct LocalTeam {
  int ngpus = 0;
  std::vector<int> devices;
  size_t max_m = 0, max_n = 0;
  LocalHub hub;
  std::vector<MgRank *> ranks;
};
static std::mutex g_team_mu;
static std::unique_ptr<LocalTeam> g_team;

static void team_destroy(LocalTeam *t) {
  if (!t) return;
  for (MgRank *r : t->ranks) mg_destroy(r);
  t->ranks.clear();
  for (cudaEvent_t &e : t->hub.ev) {
    if (e) cudaEventDestroy(e);
    e = nullptr;
  }
  cudaGetLastError();
}

void mg_release_team() {
  std::lock_guard<std::mutex> lk(g_team_mu);
  if (g_team) team_destroy(g_team.get());
  g_team.reset();
}
}  // namespace svdj

extern "C" int svdj_dgesvd_mg(int jobu, int jobv, size_t m, size_t n, double *A, size_t lda, double *s, double *V,
                              size_t ldv, const svdj_opts *opts, svdj_stats *stats, int ngpus, const int *devices) {
  clear_error();
  if (ngpus < 1) return -12;
  int avail = 0;
  SVDJ_CUDA_TRY(cudaGetDeviceCount(&avail));
  if (devices == nullptr && ngpus > avail) {
    set_error("svdj_dgesvd_mg: %d GPUs asked for, %d visible", ngpus, avail);
    return -12;
  }
  // (an explicit device list may name a device more than once: several ranks then share that GPU -- how the
  // tests run the whole driver on a one-GPU box)
  for (int i = 0; devices != nullptr && i < ngpus; ++i) {
    if (devices[i] < 0 || devices[i] >= avail) {
      set_error("svdj_dgesvd_mg: devices[%d] = %d, %d visible", i, devices[i], avail);
      return -13;
    }
  }
  if (ngpus == 1) {
    svdj_opts o;
    svdj_default_opts(&o);
    if (opts) o = *opts;
    if (devices) o.device = devices[0];
    return svdj_dgesvd(jobu, jobv, m, n, A, lda, s, V, ldv, &o, stats);
  }
  if (m == 0 || n == 0) return SVDJ_OK;
  std::lock_guard<std::mutex> lk(g_team_mu);
  std::vector<int> devs((size_t)ngpus);
  for (int i = 0; i < ngpus; ++i) devs[i] = devices ? devices[i] : i;
  if (!g_team || g_team->ngpus != ngpus || g_team->devices != devs || g_team->max_m < m || g_team->max_n < n) {
    if (g_team) team_destroy(g_team.get());
    g_team.reset(new LocalTeam());
    g_team->ngpus = ngpus;
    g_team->devices = devs;
    g_team->max_m = m;
    g_team->max_n = n;
    g_team->hub.init_board(ngpus);
    g_team->ranks.assign((size_t)ngpus, nullptr);
    std::vector<int> rcs((size_t)ngpus, 0);
    std::vector<std::string> errs((size_t)ngpus);
    std::vector<std::thread> th;
    for (int r = 0; r < ngpus; ++r) {
      th.emplace_back([&, r] {
        if (cudaSetDevice(devs[r]) != cudaSuccess) {
          rcs[r] = SVDJ_ERR_CUDA;
          errs[r] = "cudaSetDevice failed";
          // still take part in the all-gathers so that the others do not wait for ever
        }
        std::unique_ptr<LocalComm> c(new LocalComm());
        c->rank = r;
        c->size = ngpus;
        c->hub = &g_team->hub;
        MgRank *R = nullptr;
        const int rc = mg_create(std::move(c), m, n, &R);
        g_team->ranks[r] = R;
        if (rc != SVDJ_OK) {
          rcs[r] = rc;
          errs[r] = svdj_last_error();
        }
      });
    }
    for (auto &t : th) t.join();
    for (int r = 0; r < ngpus; ++r) {
      if (rcs[r] != SVDJ_OK) {
        set_error("svdj_dgesvd_mg: rank %d set-up failed: %s", r, errs[r].c_str());
        team_destroy(g_team.get());
        g_team.reset();
        return rcs[r];
      }
    }
  }
  std::vector<int> rcs((size_t)ngpus, 0);
  std::vector<std::string> errs((size_t)ngpus);
  std::vector<svdj_stats> sts((size_t)ngpus);
  std::vector<std::thread> th;
  for (int r = 0; r < ngpus; ++r) {
    th.emplace_back([&, r] {
      cudaSetDevice(devs[r]);
      rcs[r] = svdj_mg_rank_dgesvd(g_team->ranks[r], jobu, jobv, m, n, r == 0 ? A : nullptr, lda, r == 0 ? s : nullptr,
                                   r == 0 ? V : nullptr, ldv, opts, &sts[r]);
      if (rcs[r] != SVDJ_OK) errs[r] = svdj_last_error();
    });
  }
  for (auto &t : th) t.join();
  if (stats) {
    *stats = sts[0];
    for (int r = 1; r < ngpus; ++r) stats->launches += sts[r].launches;
  }
  // report the cause, not its echo: a rank that timed out waiting is usually the victim of another rank's failure
  int bad = -1;
  for (int r = 0; r < ngpus; ++r) {
    if (rcs[r] == SVDJ_OK || rcs[r] == SVDJ_NOT_CONVERGED) continue;
    const bool echo = errs[r].find("waited more than") != std::string::npos ||
                      errs[r].find("another rank failed") != std::string::npos;
    if (bad < 0 || !echo) {
      const bool bad_is_echo = bad >= 0 && (errs[bad].find("waited more than") != std::string::npos ||
                                            errs[bad].find("another rank failed") != std::string::npos);
      if (bad < 0 || bad_is_echo) bad = r;
    }
  }
  if (bad >= 0) {
    std::string all;
    for (int r = 0; r < ngpus; ++r)
      if (rcs[r] != SVDJ_OK && rcs[r] != SVDJ_NOT_CONVERGED) all += " [rank " + std::to_string(r) + ": " + errs[r] + "]";
    set_error("svdj_dgesvd_mg: rank %d failed (status %d):%s", bad, rcs[bad], all.c_str());
    return rcs[bad];
  }
  for (int r = 1; r < ngpus; ++r)
    if (rcs[r] != rcs[0]) {
      set_error("svdj_dgesvd_mg: the ranks disagree on convergence (rank 0: %d, rank %d: %d)", rcs[0], r, rcs[r]);
      return SVDJ_ERR_CUDA;
    }
  return rcs[0];
}
