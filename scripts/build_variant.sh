#!/bin/bash
# build_variant.sh NAME [make variables...]  ->  variants/NAME/libghost_b200.so  (git-ignored scratch for A/B runs:
# GHOST_B200_LIB=variants/NAME/libghost_b200.so python bench.py ...)
set -e
root=$(cd "$(dirname "$0")/.." && pwd)
name=$1; shift
mkdir -p "$root/variants/$name"
make -s -C "$root/ghost_b200/csrc" OUT="$root/variants/$name/libghost_b200.so" OBJ="$root/variants/$name/build" "$@" > "$root/variants/$name/make.log" 2>&1 || { tail -20 "$root/variants/$name/make.log"; exit 1; }
grep -A3 "search_kernelILi32ELb1ELb1Et" "$root/variants/$name/build/launch_t32.ptxas.log" | grep "Used\|spill" | tr '\n' ' '; echo " <- $name"
