#!/bin/bash
# Build the comm_norender / nocomm_norender drop-in executables against the in-tree liblearncomms.so.
set -e
HERE="$(cd "$(dirname "$0")/.." && pwd)"
OUT="$HERE/learn_comms_b200/bin"
mkdir -p "$OUT"
for flavour in comm nocomm; do
  DEF=""; [ "$flavour" = nocomm ] && DEF="-DNO_COMM"
  g++ -std=c++17 -O2 $DEF "$HERE/learn_comms_b200/cli/comm_norender.cpp" -o "$OUT/${flavour}_norender" \
      -L"$HERE/learn_comms_b200" -l:liblearncomms.so -Wl,-rpath,'$ORIGIN/..' -Wl,-rpath,/usr/local/cuda/lib64
  # pso_opt.py's validation leg (pso_opt.py:271-285) launches the *_render flavours with --record: there is no graphics
  # stack on this box, so they are the same headless executable (--record is accepted, no video is written)
  cp "$OUT/${flavour}_norender" "$OUT/${flavour}_render"
done
echo "$OUT"
