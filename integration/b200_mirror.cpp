// b200_mirror.cpp -- see b200_mirror.h.  Compiled against the reference's (prep-generated) headers; linked with
//   -Wl,--wrap=<pbPreparePlugin> -Wl,--wrap=<pbFinalizePlugin> -Wl,--wrap=<pbSetError> -Wl,--wrap=<Pb::objFromPy> -Wl,--wrap=<PbClass::~PbClass (D2)>
// so that every call the reference's generated wrappers make to those functions lands in the __wrap_ functions below first.
#include "b200_mirror.h"

#include <cstring>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

#include "general.h"
#include "fluidsolver.h"
#include "grid.h"

using namespace Manta;

namespace B200 {

struct Mirror { mf_grid* h; bool hostValid, devValid; int type; Vec3i size; FluidSolver* parent; };

static std::unordered_map<const PbClass*, Mirror>& table() { static std::unordered_map<const PbClass*, Mirror> t; return t; }
static std::unordered_map<const FluidSolver*, mf_solver*>& solvers() { static std::unordered_map<const FluidSolver*, mf_solver*> t; return t; }
static std::vector<PbClass*> g_pending;     // objects seen by objFromPy since the last plugin boundary
static bool g_portedActive = false;         // inside a plugin whose body is a C-ABI call
static bool g_hostActive = false;           // inside any other plugin / method: host code may read and write its arguments
static Stats g_stats = { 0, 0, 0, 0, 0, 0 };
Stats& stats() { return g_stats; }

void check(int status) { if (status) errMsg(mf_last_error()); }

// the plugins of b200_plugins.cpp (names as pbPreparePlugin receives them)
static bool isPorted(const std::string& n) {
	static const std::set<std::string> names = {
		"advectSemiLagrange", "addGravity", "addGravityNoScale", "addBuoyancy", "setOpenBound", "resetOutflow", "setInflowBcs",
		"setWallBcs", "releaseMG", "computePressureRhs", "solvePressureSystem", "correctVelocity", "solvePressure", "b200Stats" };
	return names.count(n) != 0;
}

static void* hostData(const GridBase* g, size_t& bytes) {
	const size_t n = (size_t)g->getSizeX() * g->getSizeY() * g->getSizeZ();
	const int t = g->getType();
	if (t & GridBase::TypeInt)  { bytes = n * sizeof(int);  return &(*(Grid<int>*)g)[0]; }
	if (t & GridBase::TypeReal) { bytes = n * sizeof(Real); return &(*(Grid<Real>*)g)[0]; }
	if (t & GridBase::TypeVec3) { bytes = n * sizeof(Vec3); return &(*(Grid<Vec3>*)g)[0]; }
	errMsg("manta_b200: grid type " << t << " has no device mirror");
}

static int devType(int t) {      // GridBase::GridType bits == MF_TYPE_* bits (grid.h:33-41, mantacuda.h)
	static_assert(GridBase::TypeReal == MF_TYPE_REAL && GridBase::TypeInt == MF_TYPE_INT && GridBase::TypeVec3 == MF_TYPE_VEC3 &&
	              GridBase::TypeMAC == MF_TYPE_MAC && GridBase::TypeLevelset == MF_TYPE_LEVELSET && GridBase::TypeFlags == MF_TYPE_FLAGS, "grid type bits");
	return t & (MF_TYPE_REAL | MF_TYPE_INT | MF_TYPE_VEC3 | MF_TYPE_MAC | MF_TYPE_LEVELSET | MF_TYPE_FLAGS);
}

mf_solver* devSolver(FluidSolver* s) {
	assertMsg(s, "manta_b200: grid without a parent solver");
	mf_solver*& h = solvers()[s];
	if (!h) {
		static_assert(sizeof(Real) == sizeof(float), "manta_b200 binds the single-precision build");
		const Vec3i gs = s->getGridSize();
		h = mf_solver_create(gs.x, gs.y, gs.z, s->is3D() ? 3 : 2, 0);
		if (!h) errMsg(mf_last_error());
	}
	check(mf_solver_set_dt(h, s->getDt()));
	return h;
}

static Mirror& mirrorOf(const GridBase* g) {
	auto& t = table();
	auto it = t.find(g);
	if (it != t.end()) {
		Mirror& m = it->second;
		if (m.type == g->getType() && m.size == g->getSize() && m.parent == g->getParent()) return m;
		mf_grid_destroy(m.h);        // (an object of another kind now lives at this address)
		t.erase(it);
	}
	Mirror m;
	m.type = g->getType(); m.size = g->getSize(); m.parent = g->getParent();
	m.h = mf_grid_create(devSolver(g->getParent()), devType(m.type));
	if (!m.h) errMsg(mf_last_error());
	m.hostValid = true; m.devValid = false;
	return t.emplace(g, m).first->second;
}

mf_grid* dev(const GridBase* g) {
	Mirror& m = mirrorOf(g);
	devSolver(g->getParent());       // refreshes dt
	if (!m.devValid || g_hostActive) {     // inside an un-ported plugin the host copy may have changed since the last upload
		size_t bytes = 0; void* p = hostData(g, bytes);
		check(mf_grid_upload(m.h, p));
		g_stats.uploads++; g_stats.uploadBytes += (long long)bytes;
		m.devValid = true;
	}
	return m.h;
}

void wroteDevice(const GridBase* g) {
	Mirror& m = mirrorOf(g);
	m.devValid = true; m.hostValid = false;
	if (g_hostActive) {               // called from inside host code that goes on reading its grids: write back at once
		size_t bytes = 0; void* p = hostData(g, bytes);
		check(mf_grid_download(m.h, p));
		g_stats.downloads++; g_stats.downloadBytes += (long long)bytes;
		m.hostValid = true;
	}
}

// an object is about to be handed to host code: make the host copy current, and distrust the device copy afterwards
static void syncToHost(PbClass* p) {
	auto it = table().find(p);
	if (it == table().end()) return;
	Mirror& m = it->second;
	const GridBase* g = static_cast<const GridBase*>(p);
	if (!m.hostValid) {
		size_t bytes = 0; void* d = hostData(g, bytes);
		check(mf_grid_download(m.h, d));
		g_stats.downloads++; g_stats.downloadBytes += (long long)bytes;
		m.hostValid = true;
	}
	m.devValid = false;
}

static void dropMirror(const PbClass* p) {
	auto it = table().find(p);
	if (it != table().end()) { mf_grid_destroy(it->second.h); table().erase(it); }
	auto is = solvers().find(static_cast<const FluidSolver*>(p));     // (pointer identity only: no cast of a dead object is dereferenced)
	if (is != solvers().end()) {
		for (auto jt = table().begin(); jt != table().end();) { if (jt->second.parent == is->first) { mf_grid_destroy(jt->second.h); jt = table().erase(jt); } else ++jt; }
		mf_solver_destroy(is->second); solvers().erase(is);
	}
}

}  // namespace B200

// ---- link-time interposition (ld --wrap): symbol names are the Itanium manglings of the reference's functions ---------------
extern "C" {

typedef struct _object PyObject;

void __real__ZN5Manta15pbPreparePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(FluidSolver*, const std::string&, bool);
void __wrap__ZN5Manta15pbPreparePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(FluidSolver* parent, const std::string& name, bool doTime) {
	__real__ZN5Manta15pbPreparePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(parent, name, doTime);
	if (B200::isPorted(name)) {
		B200::g_portedActive = true; B200::g_hostActive = false;
		B200::g_stats.portedCalls++;
	} else {
		B200::g_hostActive = true; B200::g_portedActive = false;
		bool touched = false;
		for (PbClass* p : B200::g_pending) { if (B200::table().count(p)) touched = true; B200::syncToHost(p); }
		if (touched) B200::g_stats.hostCalls++;
	}
	B200::g_pending.clear();
}

void __real__ZN5Manta16pbFinalizePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(FluidSolver*, const std::string&, bool);
void __wrap__ZN5Manta16pbFinalizePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(FluidSolver* parent, const std::string& name, bool doTime) {
	B200::g_portedActive = false; B200::g_hostActive = false; B200::g_pending.clear();
	__real__ZN5Manta16pbFinalizePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb(parent, name, doTime);
}

void __real__ZN5Manta10pbSetErrorERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEES7_(const std::string&, const std::string&);
void __wrap__ZN5Manta10pbSetErrorERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEES7_(const std::string& fn, const std::string& ex) {
	B200::g_portedActive = false; B200::g_hostActive = false; B200::g_pending.clear();      // the wrapper left through its catch block
	__real__ZN5Manta10pbSetErrorERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEES7_(fn, ex);
}

PbClass* __real__ZN2Pb9objFromPyEP7_object(PyObject*);
PbClass* __wrap__ZN2Pb9objFromPyEP7_object(PyObject* obj) {
	PbClass* p = __real__ZN2Pb9objFromPyEP7_object(obj);
	if (p) {
		if (B200::g_hostActive) B200::syncToHost(p);
		else if (!B200::g_portedActive) B200::g_pending.push_back(p);
	}
	return p;
}

void __real__ZN5Manta7PbClassD2Ev(PbClass*);
void __wrap__ZN5Manta7PbClassD2Ev(PbClass* self) {
	B200::dropMirror(self);
	for (auto& q : B200::g_pending) if (q == self) q = nullptr;
	__real__ZN5Manta7PbClassD2Ev(self);
}

}  // extern "C"
