#!/bin/bash
# usage: scripts/mkvariants.sh name1:"flags" name2:"flags" ...   -> build/variants/<name>.so (git-ignored; travels to the GPU box)
cd "$(dirname "$0")/.." || exit 1
mkdir -p build/variants; rm -f build/variants/*.so
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  UHDR_NVCC_EXTRA="$flags" python -m libultrahdr_dev_b200.build --force -v 2>&1 | grep -A3 "apply_q_kernelILi0ELb0ELb1ELb0ELb0EE" | grep -E "Used|spill" | tr '\n' ' '; echo " <- $name ($flags)"
  cp libultrahdr_dev_b200/libuhdr_b200.so build/variants/$name.so
done
python -m libultrahdr_dev_b200.build --force > /dev/null
