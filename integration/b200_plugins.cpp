// b200_plugins.cpp -- the plugin bodies of the Eulerian grid hot path, replaced by calls into the C-ABI of libmantacuda.so.
// This file is run through the reference's own preprocessor `prep` (it contains PYTHON() plugins) and linked IN FRONT OF the
// reference's objects into `manta_b200`; the reference's advection.cpp / extforces.cpp / pressure.cpp are compiled unmodified
// with their plugin functions renamed by -D (so that everything else in those files -- vorticityConfinement, the KERNELs, the
// helper plugins -- stays available on the host).  Signatures, defaults and checks are the reference's:
//   advectSemiLagrange  source/plugin/advection.cpp:443
//   addGravity :62, addGravityNoScale :68, addBuoyancy :84, setOpenBound :106, resetOutflow :134, setInflowBcs :171,
//   setWallBcs :327                                              source/plugin/extforces.cpp
//   releaseMG :252, computePressureRhs :277, solvePressureSystem :312, correctVelocity :457, solvePressure :482
//                                                                 source/plugin/pressure.cpp
#include "grid.h"
#include "particle.h"
#include "b200_mirror.h"

using namespace std;

namespace Manta {

using B200::dev; using B200::devOpt; using B200::wroteDevice; using B200::check;

//! Preconditioner for CG solver (pressure.cpp:24), kept so that scenes find the same enum values
enum Preconditioner { PcNone = 0, PcMIC = 1, PcMGDynamic = 2, PcMGStatic = 3 };

PYTHON() void advectSemiLagrange (const FlagGrid* flags, const MACGrid* vel, GridBase* grid,
                                  int order = 1, Real strength = 1.0, int orderSpace = 1, bool openBounds = false, int boundaryWidth = -1, int clampMode = 2, int orderTrace = 1)
{
	assertMsg(order==1 || order==2, "AdvectSemiLagrange: Only order 1 (regular SL) and 2 (MacCormack) supported");
	if((boundaryWidth!=-1)||(openBounds)) { debMsg("Warning: boundaryWidth and openBounds parameters in AdvectSemiLagrange plugin are deprecated (and have no more effect), please remove.", 0); }
	if (!(grid->getType() & (GridBase::TypeReal | GridBase::TypeMAC | GridBase::TypeVec3)))
		errMsg("AdvectSemiLagrange: Grid Type is not supported (only Real, Vec3, MAC, Levelset)");
	check(mf_advect_semi_lagrange(dev(flags), dev(vel), dev(grid), order, strength, orderSpace, clampMode, orderTrace));
	wroteDevice(grid);
}

PYTHON() void addGravity(const FlagGrid& flags, MACGrid& vel, Vec3 gravity, const Grid<Real>* exclude=NULL, bool scale=true) {
	check(mf_add_gravity(dev(&flags), dev(&vel), gravity.x, gravity.y, gravity.z, devOpt(exclude), scale ? 1 : 0));
	wroteDevice(&vel);
}
//! Deprecated: use addGravity(scale=false) instead
PYTHON() void addGravityNoScale(const FlagGrid& flags, MACGrid& vel, const Vec3& gravity, const Grid<Real>* exclude=NULL) {
	addGravity(flags, vel, gravity, exclude, false);
}

PYTHON() void addBuoyancy(const FlagGrid& flags, const Grid<Real>& density, MACGrid& vel, Vec3 gravity, Real coefficient=1., bool scale=true) {
	check(mf_add_buoyancy(dev(&flags), dev(&density), dev(&vel), gravity.x, gravity.y, gravity.z, coefficient, scale ? 1 : 0));
	wroteDevice(&vel);
}

PYTHON() void setOpenBound(FlagGrid& flags, int bWidth, string openBound = "", int type = FlagGrid::TypeOutflow | FlagGrid::TypeEmpty){
	if (openBound == "") return;
	check(mf_flags_set_open_bound(dev(&flags), bWidth, openBound.c_str(), type));
	wroteDevice(&flags);
}

PYTHON() void resetOutflow(FlagGrid& flags, Grid<Real>* phi = 0, BasicParticleSystem* parts = 0, Grid<Real>* real = 0, Grid<int>* index = 0, ParticleIndexSystem* indexSys = 0){
	if (parts) errMsg("resetOutflow: particle systems are not part of the device grid path");
	check(mf_reset_outflow(dev(&flags), devOpt(phi), devOpt(real)));
	wroteDevice(&flags);
	if (phi) wroteDevice(phi);
	if (real) wroteDevice(real);
}

PYTHON() void setInflowBcs(MACGrid& vel, string dir, Vec3 value) {
	check(mf_set_inflow_bcs(dev(&vel), dir.c_str(), value.x, value.y, value.z));
	wroteDevice(&vel);
}

PYTHON() void setWallBcs(const FlagGrid& flags, MACGrid& vel, const MACGrid* obvel = 0, const MACGrid* fractions = 0, const Grid<Real>* phiObs = 0, int boundaryWidth=0) {
	check(mf_set_wall_bcs(dev(&flags), dev(&vel), devOpt(obvel), devOpt(fractions), devOpt(phiObs), boundaryWidth));
	wroteDevice(&vel);
}

PYTHON() void releaseMG(FluidSolver* solver=nullptr) {
	// without an argument the reference drops every solver's hierarchy; the device library keeps one per solver handle
	if (solver) check(mf_release_mg(B200::devSolver(solver)));
}

static mf_pressure_params pressureParams(Real cgAccuracy, const Grid<Real>* phi, const Grid<Real>* perCellCorr, const MACGrid* fractions, const MACGrid* obvel,
	Real gfClamp, Real cgMaxIterFac, bool precondition, int preconditioner, bool enforceCompatibility, bool useL2Norm, bool zeroPressureFixing,
	const Grid<Real>* curv, Real surfTens, Grid<Real>* retRhs)
{
	mf_pressure_params p;
	mf_pressure_params_default(&p);
	p.cgAccuracy = cgAccuracy; p.phi = devOpt(phi); p.perCellCorr = devOpt(perCellCorr); p.fractions = devOpt(fractions); p.obvel = devOpt(obvel);
	p.gfClamp = gfClamp; p.cgMaxIterFac = cgMaxIterFac; p.precondition = precondition ? 1 : 0; p.preconditioner = preconditioner;
	p.enforceCompatibility = enforceCompatibility ? 1 : 0; p.useL2Norm = useL2Norm ? 1 : 0; p.zeroPressureFixing = zeroPressureFixing ? 1 : 0;
	p.curv = devOpt(curv); p.surfTens = surfTens; p.retRhs = devOpt(retRhs);
	return p;
}

PYTHON() void computePressureRhs(
	Grid<Real>& rhs, const MACGrid& vel,
	const Grid<Real>& pressure, const FlagGrid& flags, Real cgAccuracy = 1e-3,
	const Grid<Real>* phi = 0,
	const Grid<Real>* perCellCorr = 0,
	const MACGrid* fractions = 0,
	const MACGrid* obvel = 0,
	Real gfClamp = 1e-04,
	Real cgMaxIterFac = 1.5,
	bool precondition = true,
	int preconditioner = PcMIC,
	bool enforceCompatibility = false,
	bool useL2Norm = false,
	bool zeroPressureFixing = false,
	const Grid<Real> *curv = NULL,
	const Real surfTens = 0. )
{
	mf_pressure_params p = pressureParams(cgAccuracy, phi, perCellCorr, fractions, obvel, gfClamp, cgMaxIterFac, precondition, preconditioner,
		enforceCompatibility, useL2Norm, zeroPressureFixing, curv, surfTens, NULL);
	check(mf_compute_pressure_rhs(dev(&rhs), dev(&vel), dev(&pressure), dev(&flags), &p));
	wroteDevice(&rhs);
}

PYTHON() void solvePressureSystem(
	Grid<Real>& rhs, MACGrid& vel,
	Grid<Real>& pressure, const FlagGrid& flags, Real cgAccuracy = 1e-3,
	const Grid<Real>* phi = 0,
	const Grid<Real>* perCellCorr = 0,
	const MACGrid* fractions = 0,
	Real gfClamp = 1e-04,
	Real cgMaxIterFac = 1.5,
	bool precondition = true,
	int preconditioner = PcMIC,
	const bool enforceCompatibility = false,
	const bool useL2Norm = false,
	const bool zeroPressureFixing = false,
	const Grid<Real> *curv = NULL,
	const Real surfTens = 0.)
{
	mf_pressure_params p = pressureParams(cgAccuracy, phi, perCellCorr, fractions, NULL, gfClamp, cgMaxIterFac, precondition, preconditioner,
		enforceCompatibility, useL2Norm, zeroPressureFixing, curv, surfTens, NULL);
	mf_solve_info info;
	check(mf_solve_pressure_system(dev(&rhs), dev(&vel), dev(&pressure), dev(&flags), &p, &info));
	wroteDevice(&rhs); wroteDevice(&pressure);
	debMsg("FluidSolver::solvePressure done. Iterations:"<<info.iterations<<", residual:"<<info.resNorm, 2);
}

PYTHON() void correctVelocity(
	MACGrid& vel, Grid<Real>& pressure, const FlagGrid& flags, Real cgAccuracy = 1e-3,
	const Grid<Real>* phi = 0,
	const Grid<Real>* perCellCorr = 0,
	const MACGrid* fractions = 0,
	Real gfClamp = 1e-04,
	Real cgMaxIterFac = 1.5,
	bool precondition = true,
	int preconditioner = PcMIC,
	bool enforceCompatibility = false,
	bool useL2Norm = false,
	bool zeroPressureFixing = false,
	const Grid<Real> *curv = NULL,
	const Real surfTens = 0.)
{
	mf_pressure_params p = pressureParams(cgAccuracy, phi, perCellCorr, fractions, NULL, gfClamp, cgMaxIterFac, precondition, preconditioner,
		enforceCompatibility, useL2Norm, zeroPressureFixing, curv, surfTens, NULL);
	check(mf_correct_velocity(dev(&vel), dev(&pressure), dev(&flags), &p));
	wroteDevice(&vel);
}

PYTHON() void solvePressure(
	MACGrid& vel, Grid<Real>& pressure, const FlagGrid& flags, Real cgAccuracy = 1e-3,
	const Grid<Real>* phi = 0,
	const Grid<Real>* perCellCorr = 0,
	const MACGrid* fractions = 0,
	const MACGrid* obvel = 0,
	Real gfClamp = 1e-04,
	Real cgMaxIterFac = 1.5,
	bool precondition = true,
	int preconditioner = PcMIC,
	bool enforceCompatibility = false,
	bool useL2Norm = false,
	bool zeroPressureFixing = false,
	const Grid<Real> *curv = NULL,
	const Real surfTens = 0.,
	Grid<Real>* retRhs = NULL)
{
	mf_pressure_params p = pressureParams(cgAccuracy, phi, perCellCorr, fractions, obvel, gfClamp, cgMaxIterFac, precondition, preconditioner,
		enforceCompatibility, useL2Norm, zeroPressureFixing, curv, surfTens, retRhs);
	mf_solve_info info;
	check(mf_solve_pressure(dev(&vel), dev(&pressure), dev(&flags), &p, &info));
	wroteDevice(&vel); wroteDevice(&pressure);
	if (retRhs) wroteDevice(retRhs);
	debMsg("FluidSolver::solvePressure done. Iterations:"<<info.iterations<<", residual:"<<info.resNorm, 2);
}

//! transfer statistics of the device mirror (no counterpart in the reference): "uploads downloads portedCalls hostCalls uploadBytes downloadBytes"
PYTHON() std::string b200Stats() {
	const B200::Stats& s = B200::stats();
	std::ostringstream o;
	o << s.uploads << " " << s.downloads << " " << s.portedCalls << " " << s.hostCalls << " " << s.uploadBytes << " " << s.downloadBytes;
	return o.str();
}

} // namespace
