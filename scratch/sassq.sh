#!/bin/bash
# usage: scratch/sassq.sh [mangled-template-args, default Li0ELb0ELb1ELb0ELb0E]  -> dumps that apply_q_kernel's SASS to /tmp/q0.sass
cd /root/repo
K="apply_q_kernelI${1:-Li0ELb0ELb1ELb0ELb0EE}"
cuobjdump -sass libultrahdr_dev_b200/libuhdr_b200.so > /tmp/all.sass
awk -v k="$K" '/Function : /{f = index($0, k) > 0} f' /tmp/all.sass | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/\s+\/\* 0x[0-9a-f]+ \*\/$//' > /tmp/q0.sass
wc -l /tmp/q0.sass
grep -nE "BRA|STG|BSSY|BSYNC|EXIT" /tmp/q0.sass | cut -c1-90
