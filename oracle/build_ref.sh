#!/bin/sh
# Compiles the reference's OWN native SPRITE A-step code (Cython + C++,
# igm/cython_compiled/{sprite.pyx,cpp_sprite_assignment.cpp,.h}) from the sources
# where they lie under /root/reference into oracle/_ref/ (git-ignored; travels to
# the GPU box with the snapshot).  Nothing is copied into the repo: the generated
# .cpp and the objects live in a temp dir.  No-op when the reference tree is absent.
set -e
REF=${IGM_REFERENCE_ROOT:-/root/reference}/igm/cython_compiled
HERE=$(cd "$(dirname "$0")" && pwd)
[ -f "$REF/sprite.pyx" ] || { echo "reference tree not present: skipping oracle/_ref"; exit 0; }
OUT="$HERE/_ref"
mkdir -p "$OUT"
TMP=$(mktemp -d)
PY=${PYTHON:-python}
EXT=$($PY -c "import sysconfig; print(sysconfig.get_config_var('EXT_SUFFIX'))")
[ -f "$OUT/sprite$EXT" ] && [ "$OUT/sprite$EXT" -nt "$REF/sprite.pyx" ] && exit 0
INC_PY=$($PY -c "import sysconfig; print(sysconfig.get_paths()['include'])")
INC_NP=$($PY -c "import numpy; print(numpy.get_include())")
$PY -m cython --cplus -3 -I "$REF" "$REF/sprite.pyx" -o "$TMP/sprite.cpp"
g++ -O2 -fPIC -shared -std=c++11 -w -DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION \
    -I"$INC_PY" -I"$INC_NP" -I"$REF" "$TMP/sprite.cpp" "$REF/cpp_sprite_assignment.cpp" \
    -o "$OUT/sprite$EXT"
rm -rf "$TMP"
echo "built $OUT/sprite$EXT"
