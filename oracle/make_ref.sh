#!/bin/sh
# oracle/_ref: the reference's own implementation of the path, UNMODIFIED, for the CPU arm of bench.py
# (`--impl reference`, cpu_baseline.kind = "reference") and for cross-checking the numpy restatement in oracle/.
# The reference is three pure-Python files; this recipe copies them as they are from where they lie under
# /root/reference (nothing is built, nothing is edited).  oracle/_ref/ is git-ignored -- reference sources never enter
# the repository's history -- but not gpurun-ignored, so the copy travels to the GPU box with the snapshot.
# Run by __graft_entry__.build() when /root/reference is present; a no-op elsewhere.
set -e
SRC="${1:-/root/reference/src/featgen}"
DST="$(cd "$(dirname "$0")" && pwd)/_ref"
if [ ! -d "$SRC" ]; then
  echo "make_ref.sh: $SRC not present, nothing to do"
  exit 0
fi
mkdir -p "$DST"
for f in features.py computeFDLPSpectrogram.py computeMelSpectrum.py; do
  cp "$SRC/$f" "$DST/$f"
done
(cd "$DST" && sha1sum features.py computeFDLPSpectrogram.py computeMelSpectrum.py > SHA1SUMS)
echo "make_ref.sh: copied 3 files to $DST"
