#!/bin/sh
# Compiles the reference's OWN native SPRITE A-step code (Cython + C++,
# igm/cython_compiled/{sprite.pyx,cpp_sprite_assignment.cpp,.h}) from the sources
# where they lie under /root/reference into oracle/_ref/ (git-ignored; travels to
# the GPU box with the snapshot), and copies the demo's data files next to it.  No source is copied into the repo: the generated
# .cpp and the objects live in a temp dir.  No-op when the reference tree is absent.
set -e
REF=${IGM_REFERENCE_ROOT:-/root/reference}/igm/cython_compiled
HERE=$(cd "$(dirname "$0")" && pwd)
[ -f "$REF/sprite.pyx" ] || { echo "reference tree not present: skipping oracle/_ref"; exit 0; }
OUT="$HERE/_ref"
mkdir -p "$OUT"
# The reference's demo inputs and its recorded demo output (data, not sources): tests/test_gpu_demo.py runs the demo on
# the GPU box against them, tools/run_demo.py takes them as arguments.
DEMO=${IGM_REFERENCE_ROOT:-/root/reference}/demo
if [ -f "$DEMO/WTC11_HiC_2Mb.hcs" ]; then
    mkdir -p "$OUT/demo"
    for f in "$DEMO/WTC11_HiC_2Mb.hcs" "$DEMO/demo_sample_outputs/igm-model.hss.randomInit" "$DEMO/demo_sample_outputs/igm-model.hss.T"; do
        [ -f "$OUT/demo/$(basename "$f")" ] || cp "$f" "$OUT/demo/"
    done
fi
# The pure-Python data model of the reference (igm/model, igm/restraints, igm/utils/log.py) for the test that drives the
# b200 kernel through the REAL Model / Particle / Force / Restraint classes on the GPU box (tests/test_gpu_real_classes.py).
# Staged copies, git-ignored like everything under oracle/_ref; the package __init__ files written here are empty so that
# nothing else of igm is imported.
SRC=${IGM_REFERENCE_ROOT:-/root/reference}/igm
if [ -d "$SRC/model" ] && [ ! -f "$OUT/igm_py/igm/model/model.py" ]; then
    mkdir -p "$OUT/igm_py/igm/model" "$OUT/igm_py/igm/restraints" "$OUT/igm_py/igm/utils"
    cp "$SRC/model/__init__.py" "$SRC/model/model.py" "$SRC/model/particle.py" "$SRC/model/forces.py" "$OUT/igm_py/igm/model/"
    cp "$SRC"/restraints/*.py "$OUT/igm_py/igm/restraints/"
    cp "$SRC/utils/log.py" "$OUT/igm_py/igm/utils/"
    : > "$OUT/igm_py/igm/__init__.py"
    : > "$OUT/igm_py/igm/utils/__init__.py"
fi
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
