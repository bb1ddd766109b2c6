#!/bin/bash
# usage: scratch/sassq.sh  -> builds, dumps apply_q_kernel<false> SASS to /tmp/q0.sass, prints stats
cd /root/repo
python -m libultrahdr_dev_b200.build --force -v 2>&1 | grep -E "apply_q_kernelILb0ELb1" -A3 | grep -E "Used|spill|error"
cuobjdump -sass libultrahdr_dev_b200/libuhdr_b200.so > /tmp/all.sass
awk '/Function : .*apply_q_kernelILb0ELb1/{f=1} f&&/Function : /&&!/apply_q_kernelILb0ELb1/{f=0} f' /tmp/all.sass | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/\s+\/\* 0x[0-9a-f]+ \*\/$//' > /tmp/q0.sass
wc -l /tmp/q0.sass
grep -nE "BRA|STG|BSSY|BSYNC|EXIT" /tmp/q0.sass | cut -c1-90
