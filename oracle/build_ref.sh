#!/usr/bin/env bash
# build_ref.sh — compile the REFERENCE'S OWN LINES for the hot path into
# oracle/_ref/libsl_ref.so, to pin the restatement in sl_oracle.c.
#
# TEST INFRASTRUCTURE ONLY.  The reference's translation units cannot be compiled
# as they stand (no gRPC/protobuf toolchain here, and master.cc / worker.cc have
# compile errors of their own on exactly these lines — SURVEY.md Appendix A).  So
# this recipe lifts the hot-path statements, by line number, from the sources
# where they lie under /root/reference, applies only the listed repairs with sed,
# and compiles them inside the thin wrappers of ref_shim.cc against a 20-line
# stand-in for the generated `Update` message (delta_size / delta / add_delta).
# Nothing from the reference is kept: the extracted fragments live in a temporary
# directory for the duration of the compile only; the built library alone goes to
# oracle/_ref/, which is git-ignored (but travels to the GPU box).
#
#   fragment            reference lines                     repair
#   file0_consts.inc    src/file_server.cc:40,43,46         none
#   file0_body.inc      src/file_server.cc:152-156          none (verbatim)
#   state.inc           src/master.cc:58-60                 none
#   merge_body.inc      src/master.cc:99-110                R2  updates-> => update->
#   apply_body.inc      src/worker.cc:155-162,164           R3  update_back-> => update_back.
#                                                           R5  line 163 dropped
#   delta_body.inc      src/worker.cc:205-208               R3  update-> => update.
#   oldcopy_body.inc    src/worker.cc:215                   none
#   tick_body.inc       src/worker.cc:225-227               none
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SL_REFERENCE_DIR:-/root/reference}/src"
OUT="$HERE/_ref"
if [ ! -f "$REF/master.cc" ]; then
  echo "build_ref.sh: $REF not present; keeping any prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
rm -f "$OUT"/*.inc
FRAG="$(mktemp -d)"
trap 'rm -rf "$FRAG"' EXIT
sed -n '40p;43p;46p' "$REF/file_server.cc" > "$FRAG/file0_consts.inc"
sed -n '152,156p' "$REF/file_server.cc" > "$FRAG/file0_body.inc"
sed -n '58,60p' "$REF/master.cc" > "$FRAG/state.inc"
sed -n '99,110p' "$REF/master.cc" | sed 's/updates->/update->/' > "$FRAG/merge_body.inc"
sed -n '155,162p;164p' "$REF/worker.cc" | sed 's/update_back->/update_back./g' > "$FRAG/apply_body.inc"
sed -n '205,208p' "$REF/worker.cc" | sed 's/update->/update./' > "$FRAG/delta_body.inc"
sed -n '215p' "$REF/worker.cc" > "$FRAG/oldcopy_body.inc"
sed -n '225,227p' "$REF/worker.cc" > "$FRAG/tick_body.inc"
# -O0, no -march: the reference's own flags are `-std=c++11` and nothing else
# (src/Makefile:8); file 0 is built with -O2 only to keep the 1e8-iteration loop short.
g++ -std=c++11 -O2 -fPIC -shared -ffp-contract=off -w -I"$FRAG" "$HERE/ref_shim.cc" -o "$OUT/libsl_ref.so"
echo "built $OUT/libsl_ref.so"
