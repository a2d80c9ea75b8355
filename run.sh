#!/bin/sh
# ./run.sh fluid  -- drop-in for the reference's run.sh (run.sh:1-7): removes the old binary and
# mygrids.vdb, compiles the host driver, runs it.  The link line swaps OpenVDB/TBB/Boost for
# libflipb200.so (the FLIP substep on the GPU) + the self-contained VDB writer + zlib.
HERE=$(cd "$(dirname "$0")" && pwd)
PKG="$HERE/fluid-simulation_b200"
rm $1 2> /dev/null
rm mygrids.vdb 2> /dev/null
if [ "$1" != "fluid" ]; then echo "only the FLIP driver is provided: ./run.sh fluid"; exit 1; fi
[ -f "$PKG/libflipb200.so" ] || python "$PKG/build.py" > /dev/null || exit 1
/usr/bin/g++ -std=c++11 -pthread -O3 -DNDEBUG -I "$HERE/include" -I "$PKG/host" -o $1 \
    "$PKG/host/fluid_b200.cc" "$PKG/host/vdb_writer.cc" \
    -Wl,-rpath,"$PKG" -L"$PKG" -lflipb200 -lz -ldl -lm || exit 1
echo "Compiled"
./$1
