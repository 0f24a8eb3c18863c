#!/bin/bash
# dev: A/B runs of library variants built into build/variants/ (each replaces lib/libvcrt.so on the scratch copy of the GPU box).
# usage: tools/variant_run.sh "W H spp trav chunks scene reps shape" variant...   ("base" = the committed build)
args="$1"; shift
cp vulkancomputeraytracing_b200/lib/libvcrt.so /tmp/libvcrt_base.so
for v in "$@"; do
  if [ "$v" = base ]; then cp /tmp/libvcrt_base.so vulkancomputeraytracing_b200/lib/libvcrt.so; else cp build/variants/libvcrt_$v.so vulkancomputeraytracing_b200/lib/libvcrt.so; fi
  echo "== $v: $(python tools/prof_run.py $args 2>&1 | grep -v 'fp32 peak' | tail -2 | tr '\n' '|' | cut -c1-400)"
done
cp /tmp/libvcrt_base.so vulkancomputeraytracing_b200/lib/libvcrt.so
