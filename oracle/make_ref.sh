#!/bin/bash
# TEST / BASELINE INFRASTRUCTURE -- never imported by the product.
# Makes the UNMODIFIED reference importable on the GPU box: copies the reference's Python package (src/**/*.py, 68 files,
# ~11k lines, pure Python) from /root/reference into oracle/_ref/ (git-ignored, shipped to the box by gpurun like a built
# .so).  /root/reference does not exist on the GPU box, and the reference has no setup.py / pyproject to pip-install.
# Used only by bench.py's CPU legs (cpu_baseline.kind == "reference", `--impl reference --workload assessor`) through
# oracle/ref_loader.py (PSOB_REFERENCE_ROOT=oracle/_ref).  Nothing is edited: `diff -r` against /root/reference/src is empty.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="${1:-/root/reference}"
if [ ! -d "${SRC}/src/pso" ]; then
  echo "make_ref: ${SRC}/src/pso not found (GPU box?): keeping whatever oracle/_ref already holds" >&2
  exit 0
fi
rm -rf "${HERE}/_ref"
mkdir -p "${HERE}/_ref"
(cd "${SRC}" && find src -name '*.py' -print0 | tar --null -T - -cf -) | tar -xf - -C "${HERE}/_ref"
echo "make_ref: $(find "${HERE}/_ref" -name '*.py' | wc -l) files -> ${HERE}/_ref"
