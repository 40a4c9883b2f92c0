// b200_mirror.h -- device mirrors of the reference's host grids for the `manta_b200` binary.
//
// The reference's sources stay untouched: instead of adding members to GridBase / FluidSolver, the mirror keeps a side table
// PbClass* -> { mf_grid*, host valid, device valid } and FluidSolver* -> mf_solver*.  Plugin bodies that were replaced by C-ABI
// calls (b200_plugins.cpp) ask for `dev(grid)` (uploads when the host copy is newer) and mark their outputs `wroteDevice`.
// Everything ELSE the reference does with a grid -- an un-ported plugin, a KERNEL, Grid::save, a Python method -- reaches the
// object through the generated Python wrapper, which this binary intercepts at link time (`-Wl,--wrap`, b200_mirror.cpp):
//   pbPreparePlugin(parent, name)   pclass.cpp:36     -> is `name` one of the ported plugins?
//   Pb::objFromPy(obj)              registry.cpp:633  -> every PbClass argument / self of the call that follows
//   ArgLocker::add(p)               pconvert.cpp      -> every locked argument
//   PbClass::~PbClass()             pclass.cpp        -> drop the mirror
// For an un-ported call every object it receives is brought up to date on the host first (download if the device copy is
// newer) and its device copy is invalidated (the callee may write it).  That is the host-mirror / dirty-bit scheme of
// source/kernel.h:65-83 + pconvert.h:31-36 without editing either.
#pragma once
#include "mantacuda.h"

namespace Manta { class PbClass; class GridBase; class FluidSolver; }

namespace B200 {

//! device handle of a grid, uploaded first when the host copy is the newer one
mf_grid* dev(const Manta::GridBase* g);
inline mf_grid* devOpt(const Manta::GridBase* g) { return g ? dev(g) : nullptr; }
//! the device copy was written by a ported plugin: the host copy is stale from now on
void wroteDevice(const Manta::GridBase* g);
//! device solver of a FluidSolver (created on first use; dt is refreshed on every call)
mf_solver* devSolver(Manta::FluidSolver* s);
//! status code of the C-ABI -> Manta::Error (general.h:75-77) -> Python RuntimeError
void check(int status);
//! counters for tests / the `b200Stats` plugin: uploads, downloads, ported calls, un-ported calls that touched a mirrored grid
struct Stats { long long uploads, downloads, portedCalls, hostCalls, uploadBytes, downloadBytes; };
Stats& stats();

}  // namespace B200
