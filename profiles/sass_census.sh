#!/bin/bash
# Per-kernel census of the Blackwell-specific SASS opcodes in the shipped library (run here, no GPU needed):
#   UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA tensor load, UBLKCP = bulk copy, UTCBAR = tcgen05.commit,
#   HMMA = legacy mma.sync path (must be 0)
# usage: bash profiles/sass_census.sh > profiles/r02_sass_census.txt
lib=${1:-flow-q-learning_b200/lib/libfqlpop.so}
echo "cuobjdump -sass $lib  ($(date -u +%F))"
printf "%8s %6s %6s %8s %7s %7s %6s %7s  %s\n" UTCHMMA LDTM STTM UTMALDG UBLKCP UTCBAR HMMA instr kernel
cuobjdump -sass "$lib" | awk '
BEGIN { ops[1]="UTCHMMA"; ops[2]="LDTM"; ops[3]="STTM"; ops[4]="UTMALDG"; ops[5]="UBLKCP"; ops[6]="UTCBAR"; ops[7]=" HMMA" }
/Function :/ { if (name != "") out(); name=$3; n=0; split("", c) }
/^ +\/\*[0-9a-f]+\*\// { n++; for (k in ops) if (index($0, ops[k])) c[ops[k]]++ }
function out() { printf "%8d %6d %6d %8d %7d %7d %6d %7d  %s\n", c["UTCHMMA"], c["LDTM"], c["STTM"], c["UTMALDG"], c["UBLKCP"], c["UTCBAR"], c[" HMMA"], n, name }
END { if (name != "") out() }' | c++filt | sed -e 's/fqlpop::(anonymous namespace):://' -e 's/(fqlpop::[A-Za-z]*)$//' | sort -k1,1nr -k8,8nr | cut -c1-170
