// comm.cu -- see comm.h.  CUDA VMM + POSIX-fd exchange over abstract Unix sockets + device flag barrier.
#include "comm.h"
#include "common.cuh"
#include <cuda.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <unistd.h>
#include <time.h>
#include <cerrno>
#include <cstring>
#include <map>

namespace mf {

// ---- driver entry points (resolved through the runtime so the library has no link-time dependency on libcuda) ----
namespace {
struct Drv {
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemExport)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImport)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemGetGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*GetErrorString)(CUresult, const char**);
  bool ok;
};
Drv g_drv = {};

template <class F> void drvSym(const char* name, F& f) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  CUDA_CHECK(cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q));
  mfAssert(p && q == cudaDriverEntryPointSuccess, "mantacuda: CUDA driver entry point " << name << " not available");
  f = reinterpret_cast<F>(p);
}
void drvInit() {
  if (g_drv.ok) return;
  drvSym("cuMemCreate", g_drv.MemCreate);
  drvSym("cuMemExportToShareableHandle", g_drv.MemExport);
  drvSym("cuMemImportFromShareableHandle", g_drv.MemImport);
  drvSym("cuMemAddressReserve", g_drv.MemAddressReserve);
  drvSym("cuMemMap", g_drv.MemMap);
  drvSym("cuMemSetAccess", g_drv.MemSetAccess);
  drvSym("cuMemUnmap", g_drv.MemUnmap);
  drvSym("cuMemRelease", g_drv.MemRelease);
  drvSym("cuMemAddressFree", g_drv.MemAddressFree);
  drvSym("cuMemGetAllocationGranularity", g_drv.MemGetGranularity);
  drvSym("cuGetErrorString", g_drv.GetErrorString);
  g_drv.ok = true;
}
#define CU_CHECK(x) do { CUresult _r = (x); if (_r != CUDA_SUCCESS) { const char* _m = nullptr; g_drv.GetErrorString(_r, &_m); \
  mfErr("CUDA driver error: " << (_m ? _m : "?") << " (" << (int)_r << ") in " #x); } } while (0)
}  // namespace

struct Piece { size_t off, size; CUmemGenericAllocationHandle h; };
struct AllocRec { size_t va; std::vector<Piece> pieces; bool plain; };

struct Comm {
  int rank, nranks, device;
  bool emulated;
  size_t gran;
  std::vector<int> sock;          // connected stream socket to each peer, -1 for self
  int listenFd;
  CommShared* shared;             // global array: control granule of rank r at shared + r*ctrlStride bytes
  size_t ctrlStride;
  unsigned long long epoch;
  int* errHost; int* errDev;      // mapped pinned flag set by a barrier that timed out
  long long barriers;
  std::map<void*, AllocRec> allocs;
};

int comm_rank(const Comm* c) { return c->rank; }
int comm_nranks(const Comm* c) { return c->nranks; }
int comm_device(const Comm* c) { return c->device; }
bool comm_emulated(const Comm* c) { return c->emulated; }
size_t comm_granularity(const Comm* c) { return c->gran; }
long long comm_barriers_issued(const Comm* c) { return c->barriers; }
int comm_error(Comm* c) { return c->errHost ? *(volatile int*)c->errHost : 0; }
CommShared* comm_shared(Comm* c, int r) { return reinterpret_cast<CommShared*>(reinterpret_cast<char*>(c->shared) + (size_t)r * c->ctrlStride); }

// ---- sockets -----------------------------------------------------------------------------------------------------
static void sockName(sockaddr_un& a, socklen_t& len, const char* session, int rank) {
  memset(&a, 0, sizeof(a));
  a.sun_family = AF_UNIX;
  char name[96];
  snprintf(name, sizeof(name), "mantacuda.%s.%d", session, rank);
  const size_t n = strlen(name);
  memcpy(a.sun_path + 1, name, n);                      // abstract namespace: leading NUL, no file system entry
  len = (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}
static void sendAll(int fd, const void* p, size_t n) {
  const char* b = (const char*)p;
  while (n) { ssize_t w = ::send(fd, b, n, MSG_NOSIGNAL); if (w <= 0) { if (errno == EINTR) continue; mfErr("mantacuda comm: send failed: " << strerror(errno)); } b += w; n -= (size_t)w; }
}
static void recvAll(int fd, void* p, size_t n) {
  char* b = (char*)p;
  while (n) { ssize_t r = ::recv(fd, b, n, 0); if (r <= 0) { if (r < 0 && errno == EINTR) continue; mfErr("mantacuda comm: peer closed the connection"); } b += r; n -= (size_t)r; }
}
static void sendFd(int sock, int fd) {
  msghdr msg; memset(&msg, 0, sizeof(msg));
  char data = 'F'; iovec io = { &data, 1 };
  msg.msg_iov = &io; msg.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))]; memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
  cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  cm->cmsg_level = SOL_SOCKET; cm->cmsg_type = SCM_RIGHTS; cm->cmsg_len = CMSG_LEN(sizeof(int));
  memcpy(CMSG_DATA(cm), &fd, sizeof(int));
  for (;;) { ssize_t w = ::sendmsg(sock, &msg, MSG_NOSIGNAL); if (w == 1) break; if (w < 0 && errno == EINTR) continue; mfErr("mantacuda comm: sendmsg(fd) failed: " << strerror(errno)); }
}
static int recvFd(int sock) {
  msghdr msg; memset(&msg, 0, sizeof(msg));
  char data = 0; iovec io = { &data, 1 };
  msg.msg_iov = &io; msg.msg_iovlen = 1;
  alignas(cmsghdr) char ctrl[CMSG_SPACE(sizeof(int))]; memset(ctrl, 0, sizeof(ctrl));
  msg.msg_control = ctrl; msg.msg_controllen = sizeof(ctrl);
  for (;;) { ssize_t r = ::recvmsg(sock, &msg, 0); if (r == 1) break; if (r < 0 && errno == EINTR) continue; mfErr("mantacuda comm: recvmsg(fd) failed"); }
  cmsghdr* cm = CMSG_FIRSTHDR(&msg);
  mfAssert(cm && cm->cmsg_level == SOL_SOCKET && cm->cmsg_type == SCM_RIGHTS, "mantacuda comm: no file descriptor in message");
  int fd = -1; memcpy(&fd, CMSG_DATA(cm), sizeof(int));
  return fd;
}

void comm_host_barrier(Comm* c) {
  if (c->emulated || c->nranks == 1) return;
  char t = 'B';
  for (int p = 0; p < c->nranks; p++) if (p != c->rank) sendAll(c->sock[p], &t, 1);
  for (int p = 0; p < c->nranks; p++) if (p != c->rank) recvAll(c->sock[p], &t, 1);
}

static void connectAll(Comm* c, const char* session) {
  const int P = c->nranks, me = c->rank;
  c->sock.assign(P, -1);
  c->listenFd = ::socket(AF_UNIX, SOCK_STREAM, 0);
  mfAssert(c->listenFd >= 0, "mantacuda comm: socket() failed");
  sockaddr_un a; socklen_t len;
  sockName(a, len, session, me);
  mfAssert(::bind(c->listenFd, (sockaddr*)&a, len) == 0, "mantacuda comm: bind failed (session name already in use?): " << strerror(errno));
  mfAssert(::listen(c->listenFd, P) == 0, "mantacuda comm: listen failed");
  // connect to lower ranks (retry until they listen), accept from higher ranks
  for (int p = 0; p < me; p++) {
    sockName(a, len, session, p);
    int fd = -1;
    for (int tries = 0; tries < 1200; tries++) {
      fd = ::socket(AF_UNIX, SOCK_STREAM, 0);
      if (::connect(fd, (sockaddr*)&a, len) == 0) break;
      ::close(fd); fd = -1;
      timespec ts = { 0, 50 * 1000 * 1000 }; nanosleep(&ts, nullptr);
    }
    mfAssert(fd >= 0, "mantacuda comm: rank " << me << " could not reach rank " << p << " (60 s)");
    sendAll(fd, &me, sizeof(int));
    c->sock[p] = fd;
  }
  for (int q = me + 1; q < P; q++) {
    int fd = ::accept(c->listenFd, nullptr, nullptr);
    mfAssert(fd >= 0, "mantacuda comm: accept failed");
    int who = -1; recvAll(fd, &who, sizeof(int));
    mfAssert(who > me && who < P && c->sock[who] < 0, "mantacuda comm: unexpected peer rank " << who);
    c->sock[who] = fd;
  }
}

// ---- global allocations ----------------------------------------------------------------------------------------
void* comm_alloc(Comm* c, size_t bytesPerComp, int ncomp, size_t compStride, const size_t* byteSplit) {
  mfAssert(bytesPerComp > 0 && ncomp >= 1, "comm_alloc: bad size");
  if (c->emulated) {
    const size_t total = ncomp > 1 ? compStride * (size_t)(ncomp - 1) + bytesPerComp : bytesPerComp;
    void* p = nullptr;
    CUDA_CHECK(cudaMalloc(&p, total));
    CUDA_CHECK(cudaMemset(p, 0, total));
    AllocRec r; r.va = total; r.plain = true;
    c->allocs[p] = r;
    return p;
  }
  const size_t g = c->gran;
  const size_t G = (bytesPerComp + g - 1) / g;             // granules per component
  if (ncomp == 1) compStride = G * g;
  mfAssert(compStride % g == 0 && compStride >= G * g, "comm_alloc: component stride must be a multiple of the allocation granularity");
  const size_t va = compStride * (size_t)ncomp;
  CUdeviceptr base = 0;
  CU_CHECK(g_drv.MemAddressReserve(&base, va, g, 0, 0));
  AllocRec rec; rec.va = va; rec.plain = false;
  CUmemAllocationProp prop; memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  const int P = c->nranks;
  // granule boundaries of the hosting split: rank r hosts granules [gb[r], gb[r+1]) of every component
  size_t gb[COMM_MAX_RANKS + 1];
  gb[0] = 0; gb[P] = G;
  for (int r = 1; r < P; r++) {
    gb[r] = byteSplit ? (byteSplit[r] >= bytesPerComp ? G : (byteSplit[r] + g / 2) / g) : G * (size_t)r / P;
    if (gb[r] > G) gb[r] = G;
    if (gb[r] < gb[r - 1]) gb[r] = gb[r - 1];
  }
  for (int q = 0; q < ncomp; q++) {
    for (int owner = 0; owner < P; owner++) {
      const size_t g0 = gb[owner], g1 = gb[owner + 1];
      if (g1 == g0) continue;
      Piece pc; pc.off = compStride * (size_t)q + g0 * g; pc.size = (g1 - g0) * g; pc.h = 0;
      if (owner == c->rank) {
        prop.location.id = c->device;
        CU_CHECK(g_drv.MemCreate(&pc.h, pc.size, &prop, 0));
        int fd = -1;
        CU_CHECK(g_drv.MemExport(&fd, pc.h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
        for (int p = 0; p < P; p++) if (p != c->rank) sendFd(c->sock[p], fd);
        ::close(fd);
      } else {
        int fd = recvFd(c->sock[owner]);
        CU_CHECK(g_drv.MemImport(&pc.h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
        ::close(fd);
      }
      CU_CHECK(g_drv.MemMap(base + pc.off, pc.size, 0, pc.h, 0));
      rec.pieces.push_back(pc);
    }
  }
  CUmemAccessDesc acc; memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc.location.id = c->device; acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  for (const Piece& pc : rec.pieces) CU_CHECK(g_drv.MemSetAccess(base + pc.off, pc.size, &acc, 1));
  // every rank clears what it hosts
  for (int q = 0; q < ncomp; q++) {
    const size_t g0 = gb[c->rank], g1 = gb[c->rank + 1];
    if (g1 > g0) CUDA_CHECK(cudaMemset((void*)(base + compStride * (size_t)q + g0 * g), 0, (g1 - g0) * g));
  }
  CUDA_CHECK(cudaDeviceSynchronize());
  comm_host_barrier(c);
  void* p = (void*)base;
  c->allocs[p] = rec;
  return p;
}

void comm_free(Comm* c, void* p) {
  if (!p) return;
  auto it = c->allocs.find(p);
  if (it == c->allocs.end()) return;
  cudaDeviceSynchronize();
  if (it->second.plain) { cudaFree(p); c->allocs.erase(it); return; }
  comm_host_barrier(c);                                     // nobody may still be reading the memory this rank hosts
  for (const Piece& pc : it->second.pieces) { g_drv.MemUnmap((CUdeviceptr)p + pc.off, pc.size); g_drv.MemRelease(pc.h); }
  g_drv.MemAddressFree((CUdeviceptr)p, it->second.va);
  c->allocs.erase(it);
}

// ---- device barrier -------------------------------------------------------------------------------------------
__device__ __forceinline__ void stReleaseSys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ldAcquireSys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
// thread t signals rank t (writes this rank's epoch into the flag array hosted on rank t) and then waits for rank t's signal
__global__ void k_comm_barrier(char* shared, size_t stride, int rank, int nranks, unsigned long long epoch, int* err) {
  MF_PDL_PROLOGUE
  const int t = threadIdx.x;
  if (t >= nranks || t == rank) return;
  // no separate __threadfence_system(): griddepcontrol.wait above made the preceding kernels' writes visible to this thread, and
  // the release store below is cumulative over them at system scope
  CommShared* theirs = reinterpret_cast<CommShared*>(shared + (size_t)t * stride);
  CommShared* mine = reinterpret_cast<CommShared*>(shared + (size_t)rank * stride);
  stReleaseSys(&theirs->flag[rank], epoch);
  const long long t0 = clock64();
  while (ldAcquireSys(&mine->flag[t]) < epoch) {
    if (clock64() - t0 > 20000000000LL) { *err = 1; break; }   // ~10 s at 2 GHz: a peer died; never hang the GPU
  }
}

void comm_barrier(Comm* c, cudaStream_t stream) {
  if (c->emulated || c->nranks == 1) return;
  c->epoch++;
  c->barriers++;
  launchPdl(stream, k_comm_barrier, dim3(1), dim3(32), reinterpret_cast<char*>(c->shared), c->ctrlStride, c->rank, c->nranks, c->epoch, c->errDev);
}

// ---- life cycle ---------------------------------------------------------------------------------------------------
Comm* comm_create_emulated(int nranks, int device) {
  mfAssert(nranks >= 1 && nranks <= COMM_MAX_RANKS, "mantacuda comm: between 1 and " << COMM_MAX_RANKS << " ranks");
  CUDA_CHECK(cudaSetDevice(device));
  Comm* c = new Comm();
  c->rank = 0; c->nranks = nranks; c->device = device; c->emulated = true; c->gran = 128; c->listenFd = -1;
  c->epoch = 0; c->errHost = nullptr; c->errDev = nullptr; c->barriers = 0;
  c->ctrlStride = (sizeof(CommShared) + 255) / 256 * 256;
  c->shared = (CommShared*)comm_alloc(c, c->ctrlStride * nranks);
  return c;
}

Comm* comm_create(int rank, int nranks, int device, const char* session) {
  mfAssert(nranks >= 1 && nranks <= COMM_MAX_RANKS, "mantacuda comm: between 1 and " << COMM_MAX_RANKS << " ranks");
  mfAssert(rank >= 0 && rank < nranks, "mantacuda comm: bad rank");
  mfAssert(session && *session && strlen(session) < 64, "mantacuda comm: session name missing or too long");
  int ndev = 0;
  mfAssert(cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0, "mantacuda: no CUDA device available (there is no CPU fallback)");
  mfAssert(device >= 0 && device < ndev, "mantacuda: invalid device index " << device);
  CUDA_CHECK(cudaSetDevice(device));
  CUDA_CHECK(cudaFree(0));
  drvInit();
  Comm* c = new Comm();
  c->rank = rank; c->nranks = nranks; c->device = device; c->emulated = false; c->listenFd = -1;
  c->epoch = 0; c->barriers = 0; c->errHost = nullptr; c->errDev = nullptr;
  CUmemAllocationProp prop; memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CU_CHECK(g_drv.MemGetGranularity(&c->gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM));
  if (nranks > 1) connectAll(c, session);
  else c->sock.assign(1, -1);
  CUDA_CHECK(cudaHostAlloc(&c->errHost, sizeof(int), cudaHostAllocMapped));
  *c->errHost = 0;
  CUDA_CHECK(cudaHostGetDevicePointer(&c->errDev, c->errHost, 0));
  // one control granule per rank, hosted on that rank
  c->ctrlStride = c->gran;
  mfAssert(sizeof(CommShared) <= c->gran, "mantacuda comm: control block larger than a granule");
  c->shared = (CommShared*)comm_alloc(c, c->ctrlStride * (size_t)nranks);
  return c;
}

void comm_destroy(Comm* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  while (!c->allocs.empty()) {
    void* p = c->allocs.begin()->first;
    try { comm_free(c, p); } catch (...) { c->allocs.erase(p); }
  }
  for (int fd : c->sock) if (fd >= 0) ::close(fd);
  if (c->listenFd >= 0) ::close(c->listenFd);
  if (c->errHost) cudaFreeHost(c->errHost);
  delete c;
}

// ---- self test: pattern check through peer mappings, barrier latency, NVLink read/write bandwidth ----------------
__global__ void k_comm_fill(float* p, long long n, float base) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += st) p[i] = base + (float)(i & 1023);
}
__global__ void k_comm_check(const float* p, long long n, float base, int* bad) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  int b = 0;
  for (; i < n; i += st) b += (p[i] != base + (float)(i & 1023));
  if (b) atomicAdd(bad, b);
}
__global__ void k_comm_copy(float4* dst, const float4* src, long long n4) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long st = (long long)gridDim.x * blockDim.x;
  for (; i < n4; i += st) dst[i] = src[i];
}

}  // namespace mf

using namespace mf;

extern "C" {

mf_comm* mf_comm_create(int rank, int nranks, int device, const char* session) {
  try { return reinterpret_cast<mf_comm*>(comm_create(rank, nranks, device, session)); }
  catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
// all ranks of the job inside this process on one device: launches run rank after rank on one stream (single-GPU testing of
// the partitioned code path)
mf_comm* mf_comm_create_emulated(int nranks, int device) {
  try { return reinterpret_cast<mf_comm*>(comm_create_emulated(nranks, device)); }
  catch (const std::exception& e) { set_last_error(e.what()); return nullptr; }
}
void mf_comm_destroy(mf_comm* c) { comm_destroy(reinterpret_cast<Comm*>(c)); }

// out[0] mismatching elements seen through peer mappings, out[1] device-barrier latency (us), out[2] peer read GB/s,
// out[3] peer write GB/s, out[4] allocation granularity (bytes), out[5] local copy GB/s (read+write bytes)
int mf_comm_selftest(mf_comm* cp, long long bytesPerRank, int iters, double* out) {
  MF_API_BEGIN
  mfAssert(cp && out && bytesPerRank >= 4096 && iters > 0, "mf_comm_selftest: bad argument");
  Comm* c = reinterpret_cast<Comm*>(cp);
  mfAssert(!c->emulated, "mf_comm_selftest: needs a multi-process communicator");
  CUDA_CHECK(cudaSetDevice(c->device));
  const int P = c->nranks, me = c->rank;
  const size_t g = c->gran;
  const size_t slab = ((size_t)bytesPerRank + g - 1) / g * g;      // so that rank r hosts exactly [r*slab, (r+1)*slab)
  float* a = (float*)comm_alloc(c, slab * P);
  float* scratch = nullptr; CUDA_CHECK(cudaMalloc(&scratch, slab));
  int* bad = nullptr; CUDA_CHECK(cudaMalloc(&bad, sizeof(int))); CUDA_CHECK(cudaMemset(bad, 0, sizeof(int)));
  cudaStream_t st; CUDA_CHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  const long long n = (long long)(slab / 4);
  k_comm_fill<<<1024, 256, 0, st>>>(a + (size_t)me * n, n, (float)(me * 4096));
  comm_barrier(c, st);
  for (int r = 0; r < P; r++) k_comm_check<<<1024, 256, 0, st>>>(a + (size_t)r * n, n, (float)(r * 4096), bad);
  comm_barrier(c, st);
  int hb = 0;
  CUDA_CHECK(cudaMemcpyAsync(&hb, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_CHECK(cudaStreamSynchronize(st));
  out[0] = hb;
  cudaEvent_t e0, e1; CUDA_CHECK(cudaEventCreate(&e0)); CUDA_CHECK(cudaEventCreate(&e1));
  float ms = 0.f;
  for (int w = 0; w < 10; w++) comm_barrier(c, st);
  CUDA_CHECK(cudaEventRecord(e0, st));
  for (int i = 0; i < iters; i++) comm_barrier(c, st);
  CUDA_CHECK(cudaEventRecord(e1, st));
  CUDA_CHECK(cudaEventSynchronize(e1)); CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
  out[1] = 1e3 * ms / iters;
  const int nb = (P > 1) ? (me + 1) % P : me;
  const int reps = 5;
  auto timeCopy = [&](float* dst, const float* src) {
    k_comm_copy<<<148 * 8, 256, 0, st>>>((float4*)dst, (const float4*)src, n / 4);
    CUDA_CHECK(cudaEventRecord(e0, st));
    for (int i = 0; i < reps; i++) k_comm_copy<<<148 * 8, 256, 0, st>>>((float4*)dst, (const float4*)src, n / 4);
    CUDA_CHECK(cudaEventRecord(e1, st));
    CUDA_CHECK(cudaEventSynchronize(e1)); CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    return (double)slab * reps / (ms * 1e-3) / 1e9;
  };
  comm_barrier(c, st);
  out[2] = timeCopy(scratch, a + (size_t)nb * n);              // read the neighbour's slab over NVLink
  comm_barrier(c, st);
  CUDA_CHECK(cudaStreamSynchronize(st)); comm_host_barrier(c);
  // write into the second half of the neighbour's slab while it writes into its own neighbour's: disjoint targets
  out[3] = timeCopy(a + (size_t)nb * n, scratch);
  comm_barrier(c, st);
  out[5] = 2.0 * timeCopy(scratch, a + (size_t)me * n);
  out[4] = (double)g;
  // launch-chain cost of tiny dependent kernels by the memory they touch: out[6] plain cudaMalloc, out[7] own slab of the
  // peer-mapped array, out[8] the neighbour's slab (us per launch)
  auto timeChain = [&](float* p) {
    for (int i = 0; i < 20; i++) k_comm_fill<<<8, 128, 0, st>>>(p, 1024, 1.f);
    CUDA_CHECK(cudaEventRecord(e0, st));
    for (int i = 0; i < 500; i++) k_comm_fill<<<8, 128, 0, st>>>(p, 1024, (float)i);
    CUDA_CHECK(cudaEventRecord(e1, st));
    CUDA_CHECK(cudaEventSynchronize(e1)); CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    return 1e3 * ms / 500;
  };
  out[6] = timeChain(scratch);
  out[7] = timeChain(a + (size_t)me * n);
  comm_barrier(c, st);
  out[8] = timeChain(a + (size_t)nb * n + n / 2);
  comm_barrier(c, st);
  CUDA_CHECK(cudaStreamSynchronize(st));
  mfAssert(!comm_error(c), "mf_comm_selftest: a device barrier timed out");
  comm_host_barrier(c);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st);
  cudaFree(scratch); cudaFree(bad);
  comm_free(c, a);
  MF_API_END
}

}  // extern "C"
