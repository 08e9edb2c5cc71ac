#!/bin/bash
# per-phase instruction/sample shares of export_kernel<FAST> from an ncu report:  tools/ncu_phase.sh report.ncu-rep
set -e
REP=$(realpath $1); W=/tmp/prof_$$; mkdir -p $W; cd $W
ncu -i $REP --page source --csv > sass_all.csv 2>/dev/null
L=$(grep -n "Kernel Name.*export_kernel<(bool)1>" sass_all.csv | tail -1 | cut -d: -f1)
sed -n "$((L+1)),\$p" sass_all.csv > sass_fast.csv
cuobjdump -xelf all /root/repo/madrona-bots_b200/lib/libmbots.so >/dev/null
nvdisasm -g -c sim_kernels*.cubin > sk.asm 2>/dev/null
python /root/repo/tools/ncu_src_top.py sass_fast.csv sk.asm export_kernelILb1 ${2:-400} > all_lines.txt
head -1 all_lines.txt
echo $W
