#!/bin/bash
# Stages the UNMODIFIED reference files of the hot path under baseline/_ref/ (git-ignored, but shipped to
# the GPU box with the gpurun snapshot, where /root/reference does not exist).  Used by
#   * bench.py --impl reference  (times baseline/_ref/utils.sgc_precompute on the host cores), and
#   * tests/test_gpu_reference_scripts.py (runs baseline/_ref/citation.py, reddit.py, tuning.py against
#     this repo's drop-in utils.py / normalization.py).
# Nothing is edited: files are byte-for-byte copies (sha256 manifest written next to them).
set -eu
SRC="${1:-/root/reference}"
HERE="$(cd "$(dirname "$0")/.." && pwd)"
DST="$HERE/baseline/_ref"
if [ ! -d "$SRC" ]; then
    echo "stage_reference: $SRC not present (GPU box?) -- keeping whatever is staged in $DST" >&2
    exit 0
fi
mkdir -p "$DST/data" "$DST/SGC-tuning"
for f in normalization.py utils.py models.py metrics.py args.py citation.py reddit.py tuning.py; do
    cp -f "$SRC/$f" "$DST/$f"
done
cp -f "$SRC"/SGC-tuning/*.txt "$DST/SGC-tuning/"
for d in cora citeseer pubmed; do
    cp -f "$SRC"/data/ind.$d.* "$DST/data/" 2>/dev/null || true
done
chmod -R u+w "$DST"
( cd "$DST" && find . -type f ! -name MANIFEST.sha256 ! -name '*.pyc' ! -path './__pycache__/*' | sort | xargs sha256sum > MANIFEST.sha256 )
echo "staged $(wc -l < "$DST/MANIFEST.sha256") files into $DST ($(du -sh "$DST" | cut -f1))"
