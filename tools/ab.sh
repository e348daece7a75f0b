#!/bin/bash
# A/B timing of librt variants on the GPU box:  tools/ab.sh "v1 v2 ..." [profile_frame args]
V=$1; shift
for v in $V; do
  echo "== $v"
  RT_LIBRT_PATH=$PWD/raytracer_b200/csrc/variants/librt_$v.so python tools/profile_frame.py "$@" | tail -2
done
