// comm.h -- multi-GPU plumbing of libmantacuda: one process per GPU, ONE global virtual address space.
//
// The reference is a shared-memory program: its KERNELs split the z range over TBB/OpenMP threads
// (source/preprocessor/codegen_kernel.cpp:187-190) and every thread reads its neighbours' planes directly.
// The B200 equivalent keeps that model: every grid is one contiguous virtual range that all ranks map at
// allocation time (CUDA VMM: cuMemCreate on the hosting GPU, POSIX-fd export over a Unix socket,
// cuMemMap + cuMemSetAccess on every peer), physically split into z-slabs over the GPUs.  Kernels are launched SPMD on the
// rank's own z range; neighbour planes are plain loads that travel over NVLink 5 / NVSwitch, so there is no separate
// halo-exchange step and no halo copies.  Launches that depend on other ranks' output are separated by a device-side
// flag barrier (k_comm_barrier: st.release.sys to every peer, ld.acquire.sys spin with a time-out), and grid-wide
// reductions push per-rank partials into every rank's slot array before that barrier (fixed rank order => all ranks
// compute bit-identical scalars).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace mf {

constexpr int COMM_MAX_RANKS = 16;
constexpr int COMM_RED_K = 8;          // doubles per reduction slot

struct CommShared {                     // lives at the start of each rank's control granule (hosted on that rank)
  unsigned long long flag[COMM_MAX_RANKS];            // flag[q]: last barrier epoch signalled by rank q
  double red[2][COMM_MAX_RANKS][COMM_RED_K];           // red[parity][q][k]: rank q's partials of the current reduction
};

struct Comm;

// collective over all ranks of the session (same argument values except rank/device)
Comm* comm_create(int rank, int nranks, int device, const char* session);
// all ranks live in this process on one device (testing on a single GPU): plain cudaMalloc, barriers are stream order
Comm* comm_create_emulated(int nranks, int device);
void  comm_destroy(Comm* c);

int   comm_rank(const Comm* c);
int   comm_device(const Comm* c);
int   comm_nranks(const Comm* c);
bool  comm_emulated(const Comm* c);
size_t comm_granularity(const Comm* c);

// Global array of ncomp components, component q at byte offset q*compStride (compStride a multiple of the granularity
// when ncomp > 1).  Rank r hosts the granules nearest to bytes [byteSplit[r], byteSplit[r+1]) of every component
// (byteSplit == nullptr: an even split), so that the hosting follows the z range the rank computes.  Collective.  Zero-filled.
void* comm_alloc(Comm* c, size_t bytesPerComp, int ncomp = 1, size_t compStride = 0, const size_t* byteSplit = nullptr);
void  comm_free(Comm* c, void* p);      // collective (host barrier first)

// control block of rank r in the global address space (every rank can read/write every block)
CommShared* comm_shared(Comm* c, int r);
// enqueue the device-side barrier on `stream` (no-op when emulated)
void  comm_barrier(Comm* c, cudaStream_t stream);
// host-side barrier over the sockets (no-op when emulated)
void  comm_host_barrier(Comm* c);
// non-zero once a device barrier has timed out
int   comm_error(Comm* c);
long long comm_barriers_issued(const Comm* c);

}  // namespace mf
