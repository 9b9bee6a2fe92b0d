#!/bin/bash
# Developer tool: build a flavour of libita_b200.so into build_tools/<name>/ (git-ignored).   usage: build_variant.sh <name> [make args, e.g. FAST=1 TRACE=1 EXTRA=-DFOO]
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p $root/build_tools/$name/csrc $root/build_tools/include
cp $root/include/*.h $root/build_tools/include/
cp $root/drone-oa-iree-vit-accelerator_b200/csrc/*.cu $root/drone-oa-iree-vit-accelerator_b200/csrc/*.cuh $root/drone-oa-iree-vit-accelerator_b200/csrc/*.c $root/drone-oa-iree-vit-accelerator_b200/csrc/Makefile $root/build_tools/$name/csrc/
make -C $root/build_tools/$name/csrc -j8 ../libita_b200.so "$@" 2>&1 | grep -E "error|Error" || true
ls -la $root/build_tools/$name/libita_b200.so
