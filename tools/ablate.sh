#!/bin/bash
# Ablation builds of the fragment-layout epilogue (csrc/build.py GANCE_VARIANT) timed layer by layer: which part of the
# halo kernels' epilogue costs what. Build here (CPU), run under gpurun:  bash tools/ablate.sh build | run
variants="nostore:GANCE_ABL_NOSTORE noloads:GANCE_ABL_NOLOADS noepi:GANCE_ABL_NOEPI noframes:GANCE_ABL_NOFRAMES noloadsframes:GANCE_ABL_NOLOADS,GANCE_ABL_NOFRAMES"
if [ "$1" == "build" ]; then
  for v in $variants; do GANCE_VARIANT=$v python gance_b200/csrc/build.py | tail -1; done
else
  mkdir -p gpurun_out
  for lib in "" $(for v in $variants; do echo _${v%%:*}; done); do
    echo "== libgance_b200$lib.so"
    python tools/profile_layers.py --mode fp16 --perturbed --noise random --pairing 0 --library gance_b200/csrc/libgance_b200$lib.so 2>/dev/null | grep -E "512/conv1|1024/conv|total"
  done
fi
