#!/bin/bash
# Install the UNMODIFIED reference (Chen001117/mappo, package `onpolicy`) into the git-ignored baseline/_ref/ so that
# it travels to the GPU box with the repo snapshot (like the built .so) and `bench.py --impl reference` /
# tests/golden/make_golden.py popart can run the reference's own code there.  Run in the build container (where
# /root/reference exists); __graft_entry__.build() calls it.  Nothing is edited: the files are byte-for-byte the
# reference's.
#
# 1. the contract's recipe: pip install --target from a copy (the source tree is read-only and setuptools writes
#    build/ and *.egg-info into it).  `find_packages()` in the reference's setup.py only picks directories that have
#    an __init__.py, so that wheel LACKS onpolicy/runner/**, onpolicy/algorithms/utils/** and
#    onpolicy/algorithms/r_mappo/algorithm/** (23 of 63 modules) -- it cannot run MAPPO.
# 2. therefore the missing directories are completed with a plain copy of /root/reference/onpolicy (namespace
#    packages import fine on python 3).
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
SRC="${1:-/root/reference}"
DST="$ROOT/baseline/_ref"
[ -d "$SRC/onpolicy" ] || { echo "install_ref: $SRC/onpolicy not found (only the build container has it)"; exit 0; }
rm -rf "$DST" /tmp/_mappo_refcopy
mkdir -p "$DST"
cp -r "$SRC" /tmp/_mappo_refcopy
python -m pip install -q --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
    --target "$DST" /tmp/_mappo_refcopy >/dev/null 2>&1 || echo "install_ref: pip step failed, using the plain copy only"
rm -rf /tmp/_mappo_refcopy
cp -r "$SRC/onpolicy" "$DST/"
find "$DST" -name "__pycache__" -type d -prune -exec rm -rf {} +
n=$(find "$DST/onpolicy" -name "*.py" | wc -l)
echo "install_ref: $n modules under $DST/onpolicy"
