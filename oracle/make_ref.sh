#!/bin/sh
# oracle/make_ref.sh -- stage the UNMODIFIED reference for the checks that need it off this container.
#
# The reference is pure Python (no build step, no setup.py), so "building" it is copying its modules, byte for byte,
# from where they lie under /root/reference into oracle/_ref/ (git-ignored: reference sources never enter the
# history; NOT gpurun-ignored: the directory travels to the GPU box like a built .so).  Used by
#   * bench.py --impl reference / cpu_baseline.kind = "reference": the reference's own dm_nerf / manipulator on the
#     box's host cores,
#   * tests/test_install.py: the reference's drivers binding to the B200 path after dm_nerf_b200.install().
# Nothing under dm_nerf_b200/ reads this directory.  When /root/reference is absent (the GPU box) the script leaves an
# existing oracle/_ref alone.
set -e
REF="${DMNERF_REFERENCE:-/root/reference}"
DST="$(cd "$(dirname "$0")" && pwd)/_ref"
if [ ! -d "$REF/networks" ]; then
    echo "make_ref: $REF not present; keeping $DST as is" >&2
    exit 0
fi
rm -rf "$DST"
mkdir -p "$DST/networks" "$DST/tools" "$DST/datasets" "$DST/data"
cp "$REF"/networks/*.py "$DST/networks/"
cp "$REF"/tools/*.py "$DST/tools/"
cp "$REF"/datasets/*.py "$DST/datasets/"
cp "$REF"/config.py "$REF"/train_dmsr.py "$REF"/test_dmsr.py "$REF"/train_replica.py "$REF"/test_replica.py \
   "$REF"/train_scannet.py "$REF"/test_scannet.py "$DST/"
[ -f "$REF/data/color_dict.json" ] && cp "$REF/data/color_dict.json" "$DST/data/"
( cd "$REF" && find networks tools datasets config.py train_*.py test_*.py -name '*.py' | sort | xargs sha256sum ) > "$DST/SHA256SUMS"
echo "make_ref: staged $(wc -l < "$DST/SHA256SUMS") reference files into $DST"
