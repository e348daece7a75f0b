#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for v in res4 res5 res6; do
  echo "== $v"
  RT_LIBRT_PATH=$PWD/raytracer_b200/csrc/variants/librt_$v.so python tools/profile_frame.py --scene conference --restir --frames 32 --depth 3 --repeat 3 | tail -n 2
done
