#!/usr/bin/env bash
# Builds the reference's own CPU path tracer, headless, from the sources where they lie under
# /root/reference (never copied into this repo) into oracle/_ref/ (git-ignored, travels to the GPU box).
#   oracle/_ref/ref_bh    unmodified reference (black hole always on, src/bvh.cpp:103-113)
#   oracle/_ref/ref_flat  same sources, BVHAccel::intersect(ray,i,node) forwards to the reference's own
#                         flat traversal intersect_micro (src/bvh.cpp:115-138); the one-function edit is made
#                         by sed into oracle/_ref/gen/ at build time.
#   oracle/_ref/ref_probe[_flat]  our probe TU (oracle/ref_probe.cpp) linked against the reference objects:
#                         dumps flattened scenes, KAT vectors and HDR renders for tests/golden/.
#   oracle/_ref/exr_probe the reference's `-e map.exr` reader alone (tinyexr as vendored by the reference), for the OpenEXR fixtures
#   oracle/_ref/dae/      the bundled scenes the configs use (data, needed by the CPU baseline on the GPU box)
# Does not run the reference's CMake (it requires OpenGL/Freetype, absent here).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${RRT_REFERENCE_ROOT:-/root/reference}/pathtracer"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then echo "[build_ref] $REF not present; keeping prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT/obj" "$OUT/obj_cgl" "$OUT/gen" "$OUT/dae"
CXX="${RRT_CXX:-/usr/bin/g++}"   # the image's $CXX (/opt/gcc) has no libgomp.spec
FLAGS="-std=gnu++11 -O3 -fopenmp -w -include cstdint -DGLFW_INCLUDE_NONE -I$HERE/ref_stubs -I$REF/CGL/include -I$REF/CGL/include/CGL -I$REF/src"
SRCS=$(ls $REF/src/collada/*.cpp $REF/src/dynamic_scene/*.cpp $REF/src/static_scene/*.cpp \
  $REF/src/{halfEdgeMesh,meshEdit,bsdf,camera,sampler,bbox,bvh,pathtracer,part1_code,application,main}.cpp \
  $REF/src/misc/sphere_drawing.cpp)
CGLS=$(ls $REF/CGL/src/{vector2D,vector3D,vector4D,matrix3x3,matrix4x4,quaternion,complex,color,spectrum,tinyxml2,lodepng,path}.cpp)
compile() { # src outdir
  local o="$2/$(echo "$1" | sed "s#$REF/##; s#/#_#g; s#\.cpp\$#.o#")"
  if [ ! -f "$o" ] || [ "$1" -nt "$o" ]; then $CXX $FLAGS -c "$1" -o "$o"; fi
}
export -f compile; export CXX FLAGS REF
printf '%s\n' $SRCS | xargs -P"$(nproc)" -I{} bash -c 'compile {} '"$OUT/obj"
printf '%s\n' $CGLS | xargs -P"$(nproc)" -I{} bash -c 'compile {} '"$OUT/obj_cgl"
# flat variant of the single function
python3 - "$REF/src/bvh.cpp" "$OUT/gen/bvh_flat.cpp" <<'PY'
import re,sys
s=open(sys.argv[1]).read()
pat=re.compile(r'(bool BVHAccel::intersect\(const Ray& ray, Intersection\* i, BVHNode \*node\) const \{)(.*?)(\n\}\n)',re.S)
assert pat.search(s)
s=pat.sub(lambda m:m.group(1)+"\n  return intersect_micro(ray, i, node);"+m.group(3),s,count=1)
open(sys.argv[2],'w').write(s)
PY
$CXX $FLAGS -I$REF/src -c "$OUT/gen/bvh_flat.cpp" -o "$OUT/gen/bvh_flat.o"
$CXX $FLAGS -c "$HERE/ref_link_stubs.cpp" -o "$OUT/gen/link_stubs_cgl.o"
gcc -O1 -w -c "$HERE/ref_link_stubs_gl.c" -o "$OUT/gen/link_stubs_gl.o"
STUBS="$OUT/gen/link_stubs_cgl.o $OUT/gen/link_stubs_gl.o"
$CXX $FLAGS -c "$HERE/ref_probe.cpp" -o "$OUT/gen/ref_probe.o"
ALL=$(ls $OUT/obj/*.o $OUT/obj_cgl/*.o)
NOBVH=$(echo "$ALL" | grep -v 'src_bvh.o')
NOMAIN=$(echo "$ALL" | grep -v 'src_main.o')
NOMAINBVH=$(echo "$NOMAIN" | grep -v 'src_bvh.o')
$CXX -O3 -fopenmp -o "$OUT/ref_bh"   $ALL   $STUBS -lpthread
$CXX -O3 -fopenmp -o "$OUT/ref_flat" $NOBVH "$OUT/gen/bvh_flat.o" $STUBS -lpthread
$CXX -O3 -fopenmp -o "$OUT/ref_probe"      $NOMAIN    "$OUT/gen/ref_probe.o" $STUBS -lpthread
$CXX -O3 -fopenmp -o "$OUT/ref_probe_flat" $NOMAINBVH "$OUT/gen/bvh_flat.o" "$OUT/gen/ref_probe.o" $STUBS -lpthread
# the reference's OpenEXR reader on its own (oracle/exr_probe.cpp): pins the PIZ fixtures of tests/golden/
$CXX -std=gnu++11 -O2 -w -I$REF/CGL/include -o "$OUT/exr_probe" "$HERE/exr_probe.cpp"
for f in CBspheres_lambertian CBspheres CBspheres_microfacet_al_ag CBbunny CBbunny_microfacet_cu CBgems CBcoil CBempty; do
  cp -f "$REF/dae/sky/$f.dae" "$OUT/dae/" 2>/dev/null || true
done
cp -f "$REF/dae/keenan/banana.dae" "$OUT/dae/" 2>/dev/null || true   # <ambient> + <directional> lights
echo "[build_ref] built $(ls $OUT | tr '\n' ' ')"
