#!/bin/bash
# Builds the C++ drop-in (CudaRaytracer) and the headless CLI against the reference's OWN headers and sources
# (SceneLoader.cpp, ObjectData.cpp, PPMExporter.cpp compiled where they lie -- nothing is copied into the repo)
# and stages the four scene files into the git-ignored baseline/_ref/ so that they travel to the GPU box.
# Needs /root/reference (build container only); on the GPU box the prebuilt build/rt_cli is used.
set -euo pipefail
ROOT=$(cd "$(dirname "$0")/.." && pwd)
REF=${REF_DIR:-/root/reference}
if [ ! -f "$REF/IRaytracer.hpp" ]; then echo "build_host: $REF not present, keeping prebuilt build/ if any"; exit 0; fi
PKG="$ROOT/opencl-raytracer_b200"
mkdir -p "$ROOT/build" "$ROOT/baseline/_ref"
CXXFLAGS="-std=c++17 -O2 -ffp-contract=off -include cmath -I$PKG/compat -I$REF -I$ROOT/include -I$PKG/host"
g++ $CXXFLAGS -o "$ROOT/build/rt_cli" "$PKG/host/rt_cli.cpp" "$PKG/host/CudaRaytracer.cpp" \
    "$REF/SceneLoader.cpp" "$REF/ObjectData.cpp" "$REF/PPMExporter.cpp" \
    -L"$PKG" -lrt_b200 -Wl,-rpath,'$ORIGIN/../opencl-raytracer_b200'
cp -f "$REF"/simpleSphere.txt "$REF"/multipleSpheres.txt "$REF"/simpleScene.txt "$REF"/roundedCube.txt "$ROOT/baseline/_ref/"
echo "build_host: built build/rt_cli, staged scene files in baseline/_ref/"
