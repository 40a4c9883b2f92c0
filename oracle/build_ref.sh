#!/bin/bash
# TEST INFRASTRUCTURE ONLY -- builds the UNMODIFIED reference (zoharl3/mantaflow) CPU solver
# from the sources where they lie under $MANTA_REF (default /root/reference) into
# oracle/_ref/manta (float, OpenMP) and, with DOUBLE=1, oracle/_ref/mantad (double).
#
# The reference's own CMakeLists.txt does not configure on Linux (hard-coded
# c:/prj/... include at CMakeLists.txt:275 and `attrib` commands at :627-629), so this
# script restates its pipeline (CMakeLists.txt:150-165 prep tool, :175-256 source lists,
# :597-657 generate/link/register, :803 -O3):
#   1. g++ the reference's source-to-source preprocessor `prep`
#   2. prep generate / link / register  -> build/pp*/  (generated code, never committed)
#   3. g++ -O3 -fopenmp everything, link against libpython3.12 + zlib
# Nothing is copied from the reference into the repository; all outputs live under
# oracle/_ref/ (git-ignored; build/ below it is also gpurun-ignored).
set -e
REF=${MANTA_REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
BUILD=$OUT/build
JOBS=${JOBS:-$(nproc)}
if [ ! -d "$REF/source" ]; then echo "build_ref: $REF not present, skipping (prebuilt oracle/_ref is used)"; exit 0; fi
PYINC=$(python3 -c "import sysconfig;print(sysconfig.get_paths()['include'])")
PYLIB=$(python3 -c "import sysconfig;print('python'+sysconfig.get_config_var('LDVERSION'))")
PYLIBDIR=$(python3 -c "import sysconfig;print(sysconfig.get_config_var('LIBDIR'))")
mkdir -p "$BUILD"
cd "$BUILD"

if [ ! -x prep ]; then
  P=$REF/source/preprocessor
  g++ -O2 -w $P/main.cpp $P/code.cpp $P/tokenize.cpp $P/parse.cpp $P/util.cpp $P/merge.cpp \
      $P/codegen_python.cpp $P/codegen_kernel.cpp -o prep
fi

PP_SOURCES="general.cpp fluidsolver.cpp conjugategrad.cpp multigrid.cpp grid.cpp grid4d.cpp levelset.cpp fastmarch.cpp shapes.cpp mesh.cpp particle.cpp movingobs.cpp fileio/ioutil.cpp fileio/iogrids.cpp fileio/iomeshes.cpp fileio/ioparticles.cpp fileio/iovdb.cpp fileio/mantaio.cpp noisefield.cpp kernel.cpp vortexsheet.cpp vortexpart.cpp turbulencepart.cpp timing.cpp edgecollapse.cpp plugin/advection.cpp plugin/extforces.cpp plugin/apic.cpp plugin/flip.cpp plugin/fire.cpp plugin/fluidguiding.cpp plugin/kepsilon.cpp plugin/initplugins.cpp plugin/meshplugins.cpp plugin/pressure.cpp plugin/ptsplugins.cpp plugin/secondaryparticles.cpp plugin/surfaceturbulence.cpp plugin/vortexplugins.cpp plugin/waveletturbulence.cpp plugin/waves.cpp plugin/implicitdensityprojection.cpp python/defines.py test.cpp"
PP_HEADERS="general.h commonkernels.h conjugategrad.h multigrid.h fastmarch.h fluidsolver.h grid.h grid4d.h mesh.h particle.h levelset.h shapes.h noisefield.h vortexsheet.h kernel.h timing.h movingobs.h fileio/mantaio.h edgecollapse.h vortexpart.h turbulencepart.h"
NOPP="$REF/source/pwrapper/pymain.cpp $REF/source/pwrapper/pclass.cpp $REF/source/pwrapper/pvec3.cpp $REF/source/pwrapper/pconvert.cpp $REF/source/pwrapper/registry.cpp $REF/source/util/vectorbase.cpp $REF/source/util/vector4d.cpp $REF/source/util/simpleimage.cpp $REF/dependencies/cnpy/cnpy.cpp"

build_variant() {   # $1 = binary name, $2 = extra defines, $3 = MT mode for prep, $4 = extra cflags
  local NAME=$1 XDEF=$2 MT=$3 XCF=$4
  local PP=pp_$NAME OBJ=obj_$NAME
  if [ -x "$OUT/$NAME" ] && [ "$FORCE" != "1" ]; then echo "build_ref: $OUT/$NAME exists"; return; fi
  mkdir -p $PP/source/plugin $PP/source/fileio $PP/source/python $PP/source/util $OBJ
  local REGS=""
  for f in $PP_SOURCES $PP_HEADERS; do
    ./prep generate 0 $MT $REF/source/ $f $PP/source/$f > /dev/null
    case $f in *.h|*.py) REGS="$REGS $PP/source/$f.reg";; esac
  done
  ./prep link $REGS > /dev/null
  echo '#define MANTA_GIT_VERSION "commit oracle"' > $PP/source/gitinfo.h
  local PPS=$(find $PP/source -name '*.cpp' ! -name 'registration.cpp' | sort)
  ./prep register $NOPP $PPS $PP/source/registration.cpp > /dev/null
  local INC="-I$PP/source -I$PP/source/util -I$PP/source/fileio -I$REF/source/pwrapper -I$REF/source/util -I$REF/source/fileio -I$REF/dependencies/cnpy -I$PYINC"
  local DEFS="-DMANTAVERSION=\\\"0.13\\\" -DLINUX=1 $XDEF"
  local OBJS=""
  : > Makefile.$NAME.objs
  for f in $NOPP $PPS $PP/source/registration.cpp; do
    local o=$OBJ/$(echo $f | tr '/' '_').o
    printf '%s: %s\n\tg++ -std=c++14 -O3 %s -w %s %s -c %s -o %s\n' "$o" "$f" "$XCF" "$DEFS" "$INC" "$f" "$o" >> Makefile.$NAME.objs
    OBJS="$OBJS $o"
  done
  printf '%s: %s\n\tg++ %s -o %s %s -L%s -l%s -lz -ldl -lm\n' "$NAME" "$OBJS" "$XCF" "$NAME" "$OBJS" "$PYLIBDIR" "$PYLIB" > Makefile.$NAME
  cat Makefile.$NAME.objs >> Makefile.$NAME
  make -s -j$JOBS -f Makefile.$NAME $NAME
  cp $NAME "$OUT/$NAME"
  echo "build_ref: built $OUT/$NAME"
}

build_variant manta "-DMANTA_MT=1 -DOPENMP=1" OPENMP "-fopenmp"
if [ "$DOUBLE" = "1" ]; then
  build_variant mantad "-DMANTA_MT=1 -DOPENMP=1 -DFLOATINGPOINT_PRECISION=2" OPENMP "-fopenmp"
fi
if [ "$SERIAL" = "1" ]; then
  build_variant manta1 "" NONE ""
fi
