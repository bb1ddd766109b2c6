#!/bin/bash
# usage: scratch/mkvariants.sh name1:"flags" name2:"flags" ...   -> scratch/variants/<name>.so
cd /root/repo; mkdir -p scratch/variants; rm -f scratch/variants/*.so
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  UHDR_NVCC_EXTRA="$flags" python -m libultrahdr_dev_b200.build --force -v 2>&1 | grep -A3 "apply_q_kernelILi0ELb0ELb1ELb0ELb0EE" | grep -E "Used|spill" | tr '\n' ' '; echo " <- $name"
  cp libultrahdr_dev_b200/libuhdr_b200.so scratch/variants/$name.so
done
python -m libultrahdr_dev_b200.build --force > /dev/null
