#!/usr/bin/env bash
# Builds librx_metis.so — METIS 5 (serial multilevel k-way, the algorithm family behind the reference's ParMETIS_V3_PartKway call,
# Common/src/geometry_structure.cpp:11524) — from the copy vendored in the reference tree (externals/metis: GKlib + libmetis, plain C),
# compiled where the sources lie; nothing of it is copied into this repository.  The library is optional: rx_partition_kway uses it
# when it is present next to librx_cuda.so (or named by RX_METIS_LIB) and falls back to the library's own partitioner otherwise.
# The built .so is git-ignored and travels to the GPU box with the snapshot, like the other built artefacts.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${SU2_REFERENCE:-/root/reference}"
M="$REF/externals/metis"
OUT="$HERE/../../librx_metis.so"
if [ ! -d "$M/libmetis" ]; then echo "build_metis.sh: $M not present — nothing to do (a prebuilt librx_metis.so is used if it exists)"; exit 0; fi
if [ -e "$OUT" ]; then echo "librx_metis.so is up to date"; exit 0; fi
T="$(mktemp -d)"
trap 'rm -rf "$T"' EXIT
J="${JOBS:-$(nproc)}"
( ls "$M"/GKlib/*.c | sed 's|$| gk|'; ls "$M"/libmetis/*.c | sed 's|$| lm|' ) | \
  xargs -P "$J" -L 1 bash -c 'gcc -O2 -w -fPIC -DNDEBUG -DNDEBUG2 -I'"$M"'/include -I'"$M"'/GKlib -I'"$M"'/libmetis -c "$0" -o '"$T"'/"$1"_$(basename "$0" .c).o'
gcc -shared -o "$OUT" "$T"/*.o -lm
echo "built $OUT"
