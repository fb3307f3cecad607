#!/bin/bash
# Builds the REFERENCE's own scalar_rgb variant (SURVEY.md 8c recipe) from a scratch copy of
# /root/reference and installs the importable python tree into the git-ignored baseline/_ref/.
# Test infrastructure only: used to generate tests/golden/ref_*.npz (tests/golden/gen_ref_renders.py)
# and as the timed CPU arm (bench.py --impl reference, cpu_baseline.kind = "reference").
# Nothing of the reference is committed; all edits below are build glue outside the hot path.
#
#   tools/build_reference.sh [pinned|native]
#     pinned : plain x86-64, -ffp-contract=off (every fmadd = rounded mul + rounded add; the
#              arithmetic the oracle restates and the CUDA "strict" namespace reproduces)
#     native : -march=x86-64-v3 -ffp-contract=fast (what a user's build does, minus AVX-512 so that
#              it runs on any GPU-box host) -- the one that is timed
set -euo pipefail
FLAVOUR=${1:-pinned}
REF=${REF:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
WORK=${WORK:-/tmp/mi3ref_$FLAVOUR}
SRC=$WORK/src
BLD=$WORK/build
OUT=$ROOT/baseline/_ref/$FLAVOUR
JOBS=${JOBS:-$(nproc)}

if [ ! -d "$SRC" ]; then
  mkdir -p "$WORK"
  # docs/tutorials/notebooks are not needed by the build
  rsync -a --exclude docs --exclude tutorials --exclude notebooks --exclude .git "$REF/" "$SRC/" 2>/dev/null || \
    (mkdir -p "$SRC" && cd "$REF" && tar cf - --exclude=./docs --exclude=./tutorials --exclude=./notebooks . | (cd "$SRC" && tar xf -))
  chmod -R u+w "$SRC"
  cd "$SRC"
  # (2) ext/asmjit is an empty directory in this archive: drop it and force the non-JIT struct converter
  python3 - <<'EOF'
import re, io
def sub(path, pat, rep, count=1, flags=re.S):
    s = open(path).read()
    n = re.subn(pat, rep, s, count=count, flags=flags)
    assert n[1] >= 1, (path, pat)
    open(path, "w").write(n[0])
sub("ext/CMakeLists.txt", r'if \(\$\{CMAKE_SYSTEM_PROCESSOR\} MATCHES "x86_64\|AMD64"\)\n  # Build asmjit', 'if (FALSE)\n  # Build asmjit')
sub("CMakeLists.txt", r'set\(MI_DEPEND \$\{MI_DEPEND\} asmjit\)', '')
for f in ("src/core/CMakeLists.txt", "src/mitsuba/CMakeLists.txt", "src/python/CMakeLists.txt"):
    sub(f, r'target_link_libraries\((\S+) PRIVATE asmjit\)', '')
sub("include/mitsuba/core/struct.h", r'#if defined\(DRJIT_X86_64\) && defined\(DRJIT_X86_SSE42\)', '#if 0')
sub("include/mitsuba/core/jit.h", r"#if defined\(DRJIT_X86_64\)", "#if 0", count=0)
# (4) no IPO: lto-wrapper is not runnable here
sub("src/CMakeLists.txt", r'INTERPROCEDURAL_OPTIMIZATION ON', 'INTERPROCEDURAL_OPTIMIZATION OFF')
sub("src/core/CMakeLists.txt", r'INTERPROCEDURAL_OPTIMIZATION ON', 'INTERPROCEDURAL_OPTIMIZATION OFF')
sub("src/render/CMakeLists.txt", r'INTERPROCEDURAL_OPTIMIZATION ON', 'INTERPROCEDURAL_OPTIMIZATION OFF')
sub("ext/drjit/ext/drjit-core/ext/nanothread/CMakeLists.txt", r'INTERPROCEDURAL_OPTIMIZATION_RELEASE TRUE', 'INTERPROCEDURAL_OPTIMIZATION_RELEASE FALSE')
EOF
  if [ "$FLAVOUR" = pinned ]; then
    find . -path '*cmake-defaults/CMakeLists.txt' -exec sed -i 's/-ffp-contract=fast/-ffp-contract=off/' {} +
  fi
fi

mkdir -p "$BLD"
cd "$BLD"
EXR=$SRC/ext/openexr
# cmake wants the (missing) header to exist when it generates; the real one is written right after
[ -e "$EXR/OpenEXR/IlmImf/dwaLookups.h" ] || echo "// PLACEHOLDER" > "$EXR/OpenEXR/IlmImf/dwaLookups.h"
if [ "$FLAVOUR" = pinned ]; then NATIVE=""; else NATIVE="-march=x86-64-v3"; fi
cmake -GNinja -S "$SRC" -B "$BLD" -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DCMAKE_BUILD_TYPE=Release \
  -DMI_DEFAULT_VARIANTS=scalar_rgb -DMI_ENABLE_EMBREE=OFF -DMI_ENABLE_GUI=OFF \
  -DMI_NATIVE_FLAGS="$NATIVE" -DDRJIT_NATIVE_FLAGS="$NATIVE" -DNANOTHREAD_NATIVE_FLAGS="$NATIVE" \
  -DCMAKE_CXX_FLAGS="-include cstdint -include cmath" \
  -DPython_EXECUTABLE="$(which python3)" > "$WORK/cmake.log" 2>&1 || { tail -40 "$WORK/cmake.log"; exit 1; }
# (3) ext/openexr/OpenEXR/IlmImf/dwaLookups.h is a missing large blob in this archive: regenerate it with the
# generator the tree ships (dwaLookups.cpp), using the config headers cmake just wrote
if grep -q PLACEHOLDER "$EXR/OpenEXR/IlmImf/dwaLookups.h"; then
  CFG=$(dirname "$(find "$BLD" -name OpenEXRConfig.h | head -1)")
  ICFG=$(dirname "$(find "$BLD" -name IlmBaseConfig.h | head -1)")
  g++ -O1 -std=c++14 -I"$CFG" -I"$ICFG" -I$EXR/OpenEXR/IlmImf -I$EXR/IlmBase/Half -I$EXR/IlmBase/IlmThread \
      -I$EXR/IlmBase/Iex -I$EXR/IlmBase/Imath -o "$WORK/dwaLookups" \
      $EXR/OpenEXR/IlmImf/dwaLookups.cpp $EXR/IlmBase/Half/half.cpp \
      $EXR/IlmBase/IlmThread/IlmThread.cpp $EXR/IlmBase/IlmThread/IlmThreadPosix.cpp \
      $EXR/IlmBase/IlmThread/IlmThreadSemaphore.cpp $EXR/IlmBase/IlmThread/IlmThreadSemaphorePosix.cpp \
      $EXR/IlmBase/IlmThread/IlmThreadSemaphorePosixCompat.cpp $EXR/IlmBase/IlmThread/IlmThreadMutex.cpp \
      $EXR/IlmBase/IlmThread/IlmThreadMutexPosix.cpp $EXR/IlmBase/IlmThread/IlmThreadPool.cpp \
      $EXR/IlmBase/Iex/IexBaseExc.cpp $EXR/IlmBase/Iex/IexThrowErrnoExc.cpp -lpthread
  "$WORK/dwaLookups" > "$EXR/OpenEXR/IlmImf/dwaLookups.h" 2>/dev/null
  ls -la "$EXR/OpenEXR/IlmImf/dwaLookups.h"
fi
ninja -j "$JOBS" > "$WORK/ninja.log" 2>&1 || { tail -60 "$WORK/ninja.log"; exit 1; }

# install: the python tree (mitsuba + drjit packages), libraries and plugins
rm -rf "$OUT"; mkdir -p "$OUT"
cp -a "$BLD/python" "$OUT/python"
# shared libraries the extension modules need travel next to them (rpath is $ORIGIN-relative in the build tree)
for f in "$BLD"/*.so* "$BLD"/lib/*.so*; do [ -e "$f" ] && cp -a "$f" "$OUT/python/mitsuba/" 2>/dev/null || true; done
[ -d "$BLD/plugins" ] && cp -a "$BLD/plugins" "$OUT/python/mitsuba/" 2>/dev/null || true
[ -d "$BLD/data" ] && cp -a "$BLD/data" "$OUT/python/mitsuba/" 2>/dev/null || true
du -sh "$OUT"
echo "reference ($FLAVOUR) installed in $OUT; use PYTHONPATH=$OUT/python"
