#!/bin/bash
# Blackwell proof: which tensor-core / tensor-memory / bulk-copy SASS the shipped library
# holds, and the register / shared-memory table of every kernel. Runs here (no GPU needed); output is committed
# as profiles/<TAG>_sass_summary.txt.
TAG=${1:-r2}
LIB=madrona-bots_b200/lib/libmbots.so
OUT=profiles/${TAG}_sass_summary.txt
{
  echo "# $LIB  ($(date -u +%Y-%m-%dT%H:%MZ), $(/usr/local/cuda/bin/nvcc --version | tail -2 | head -1))"
  echo "# cuobjdump -sass | mnemonic histogram of the Blackwell-specific instructions"
  /usr/local/cuda/bin/cuobjdump -sass $LIB | grep -oE '\b(UTCHMMA|UTCQMMA|UTCBAR|UTCCP|LDTM|STTM|UBLKCP|UBLKPF|UTMALDG|UTMASTG|HMMA\.[0-9]+|SYNCS[.A-Z0-9_]*|ATOMS[.A-Z0-9_]*|REDUX[.A-Z0-9_]*|MUFU\.[A-Z0-9]+|ACQBULK|CCTL[.A-Z]*|ELECT|LDGSTS[.A-Z0-9_]*|FENCE[.A-Z0-9_]*)' \
      | sort | uniq -c | sort -rn
  echo
  echo "# per kernel: arch, registers, shared memory (cuobjdump -res-usage)"
  /usr/local/cuda/bin/cuobjdump -res-usage $LIB | grep -E 'Function|REG' | paste - - | sed -E 's/Fatbin elf code://' \
      | c++filt | awk '{$1=$1};1'
  echo
  echo "# per kernel: count of Blackwell-specific mnemonics"
  /usr/local/cuda/bin/cuobjdump -sass $LIB | awk '/Function :/ {fn=$3} /UTCHMMA|LDTM|STTM|UBLKCP|HMMA\./ {split($0,a," "); for (i in a) if (a[i] ~ /^(UTCHMMA|LDTM|STTM|UBLKCP|HMMA)/) {sub(/;$/,"",a[i]); split(a[i],b,"."); c[fn" "b[1]]++}} END {for (k in c) print c[k], k}' | c++filt | sort -k2
} > $OUT
wc -l $OUT; head -30 $OUT
