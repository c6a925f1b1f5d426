#!/bin/bash
# Regenerates tests/golden/scenes/*.scene.bin from the reference's scene files with the
# reference's own loader.  Needs /root/reference (build container only).
set -euo pipefail
ROOT=$(cd "$(dirname "$0")/.." && pwd)
REF=${REF_DIR:-/root/reference}
T=$(mktemp -d /tmp/rtfix.XXXXXX)
g++ -std=c++17 -O2 -ffp-contract=off -include cmath -I"$ROOT/opencl-raytracer_b200/compat" -I"$REF" -I"$ROOT/include" \
    "$ROOT/tools/dump_scene.cpp" "$REF/SceneLoader.cpp" "$REF/ObjectData.cpp" -o "$T/dump_scene"
mkdir -p "$ROOT/tests/golden/scenes"
for s in simpleSphere multipleSpheres simpleScene roundedCube; do
    "$T/dump_scene" "$REF/$s.txt" "$ROOT/tests/golden/scenes/$s.scene.bin"
done
# cfg4 derived variant: simpleScene plus two more lights, to exercise
# last-light-wins (SAR:238) and the stale specular term (SAR:231-232).
cp "$REF/simpleScene.txt" "$T/simpleScene3L.txt"
printf '\ntranslate -8 6 4\n  light globalLight\ntranslate 0 -12 6\n  light globalLight\n' >> "$T/simpleScene3L.txt"
"$T/dump_scene" "$T/simpleScene3L.txt" "$ROOT/tests/golden/scenes/simpleScene3L.scene.bin"
rm -rf "$T"
