#!/usr/bin/env bash
# Stage the LIVE reference's hot-path modules for timing on the GPU box (SURVEY §8d "as-shipped" CPU baseline).
#
# /root/reference does not exist on the GPU box, and its hot path is two pure-Python files with no package metadata
# (nothing to pip-install).  This copies them, unmodified, into the git-ignored oracle/_ref/ - which is NOT
# gpurun-ignored, so it travels with the snapshot like a built .so.  bench.py then times the real thing
# (`cpu_baseline.kind: "reference"`, `--impl reference`); without the staged copy it falls back to the C port
# (`kind: "port"`).  Nothing under oracle/_ref/ is ever committed, imported by the product, or used as a checker.
set -euo pipefail
REF="${SYN3R_REFERENCE:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
DST="$HERE/oracle/_ref/solver_utils"
if [ ! -f "$REF/solver_utils/forward_warp.py" ]; then
    echo "stage_reference: $REF/solver_utils/forward_warp.py not found (nothing staged)" >&2
    exit 0
fi
mkdir -p "$DST"
cp "$REF/solver_utils/forward_warp.py" "$REF/solver_utils/consistency.py" "$DST/"
: > "$DST/__init__.py"
( cd "$DST" && sha256sum forward_warp.py consistency.py ) > "$HERE/oracle/_ref/SHA256SUMS"
echo "stage_reference: staged $(ls "$DST" | tr '\n' ' ')-> oracle/_ref/solver_utils"
