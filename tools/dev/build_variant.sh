#!/bin/bash
# Developer tool: build an alternative libhmj_b200.so beside the product one, for A/B runs with HMJ_B200_LIB.
# usage: tools/dev/build_variant.sh <name> "<extra nvcc flags>"   ->  humanoid-mujoco_b200/lib_<name>/libhmj_b200.so
set -e
cd "$(dirname "$0")/../../humanoid-mujoco_b200/csrc"
make OUT=../lib_$1/libhmj_b200.so OBJ=../lib_$1/obj XFLAGS="$2" 2>&1 | grep -i "error\|warning" || true
ls -la ../lib_$1/libhmj_b200.so
