#!/bin/bash
# Builds integration/_build/manta_b200: the reference's own executable (its prep, pwrapper, Python API, every un-ported plugin,
# compiled from the UNMODIFIED sources under $MANTA_REF by oracle/build_ref.sh) with the plugin bodies of the Eulerian grid hot
# path replaced by C-ABI calls into mantaflow_b200/libmantacuda.so (b200_plugins.cpp) and the host-mirror / dirty-bit
# bookkeeping interposed at link time (b200_mirror.cpp, ld --wrap).  Scenes run as `manta_b200 scene.py`, exactly like
# `manta scene.py`.  Needs the reference checkout; on the GPU box the prebuilt binary travels with the snapshot.
#
#   1. prep generate  b200_plugins.cpp                       (the reference's preprocessor builds the Python wrappers)
#   2. g++            b200_plugins.o b200_mirror.o
#   3. g++            advection.cpp / extforces.cpp / pressure.cpp of the reference, unmodified, with their ported plugin
#                     functions renamed by -D<name>=ref_<name> (their other plugins and KERNELs stay available on the host)
#   4. link           ours FIRST (WrapperRegistry::addMethod keeps the first registration of a name, registry.cpp:265-273),
#                     then the oracle's objects, -lmantacuda, --wrap of the four hook points
set -e
REF=${MANTA_REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
ROOT=$(dirname "$HERE")
OB=$ROOT/oracle/_ref/build
OUT=$HERE/_build
if [ ! -d "$REF/source" ]; then echo "build_manta_b200: $REF not present, skipping (prebuilt integration/_build/manta_b200 is used)"; exit 0; fi
if [ ! -x "$OB/prep" ] || [ ! -d "$OB/obj_manta" ]; then bash "$ROOT/oracle/build_ref.sh"; fi
if [ ! -f "$ROOT/mantaflow_b200/libmantacuda.so" ]; then make -s -C "$ROOT/mantaflow_b200/csrc" -j"$(nproc)"; fi
mkdir -p "$OUT/pp" "$OUT/obj"
PYINC=$(python3 -c "import sysconfig;print(sysconfig.get_paths()['include'])")
PYLIB=$(python3 -c "import sysconfig;print('python'+sysconfig.get_config_var('LDVERSION'))")
PYLIBDIR=$(python3 -c "import sysconfig;print(sysconfig.get_config_var('LIBDIR'))")
PP=$OB/pp_manta/source
INC="-I$PP -I$PP/util -I$PP/fileio -I$REF/source/pwrapper -I$REF/source/util -I$REF/source/fileio -I$REF/dependencies/cnpy -I$PYINC -I$ROOT/include -I$HERE"
CF="-std=c++14 -O3 -fopenmp -w -DMANTAVERSION=\"0.13\" -DLINUX=1 -DMANTA_MT=1 -DOPENMP=1"

# 1 + 2: our plugin bodies through the reference's preprocessor, and the mirror
"$OB/prep" generate 0 OPENMP "$HERE/" b200_plugins.cpp "$OUT/pp/b200_plugins.cpp" > /dev/null
g++ $CF $INC -c "$OUT/pp/b200_plugins.cpp" -o "$OUT/obj/b200_plugins.o"
g++ $CF $INC -c "$HERE/b200_mirror.cpp" -o "$OUT/obj/b200_mirror.o"

# 3: the three reference files that hold the ported plugins, unmodified, ported names moved out of the way
ren() { local d=""; for n in "$@"; do d="$d -D$n=ref_$n -DPbRegister_$n=PbRegister_ref_$n"; done; echo "$d"; }
g++ $CF $INC $(ren advectSemiLagrange) -c "$PP/plugin/advection.cpp" -o "$OUT/obj/ref_advection.o" &
g++ $CF $INC $(ren addGravity addGravityNoScale addBuoyancy setOpenBound resetOutflow setInflowBcs setWallBcs) -c "$PP/plugin/extforces.cpp" -o "$OUT/obj/ref_extforces.o" &
g++ $CF $INC $(ren releaseMG computePressureRhs solvePressureSystem correctVelocity solvePressure) -c "$PP/plugin/pressure.cpp" -o "$OUT/obj/ref_pressure.o" &
wait

# 4: link
REFOBJS=$(ls "$OB"/obj_manta/*.o | grep -v -e 'plugin_advection.cpp.o' -e 'plugin_extforces.cpp.o' -e 'plugin_pressure.cpp.o')
WRAP=""
for s in _ZN5Manta15pbPreparePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb \
         _ZN5Manta16pbFinalizePluginEPNS_11FluidSolverERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEEb \
         _ZN5Manta10pbSetErrorERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEES7_ \
         _ZN2Pb9objFromPyEP7_object _ZN5Manta7PbClassD2Ev; do WRAP="$WRAP -Wl,--wrap=$s"; done
g++ -fopenmp -o "$OUT/manta_b200" "$OUT/obj/b200_plugins.o" "$OUT/obj/b200_mirror.o" "$OUT/obj/ref_advection.o" "$OUT/obj/ref_extforces.o" "$OUT/obj/ref_pressure.o" \
    $REFOBJS $WRAP -L"$ROOT/mantaflow_b200" -lmantacuda -Wl,-rpath,'$ORIGIN/../../mantaflow_b200' -L"$PYLIBDIR" -l"$PYLIB" -lz -ldl -lm
echo "build_manta_b200: built $OUT/manta_b200"
