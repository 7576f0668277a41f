#!/bin/bash
# SASS evidence of the Blackwell-specific instructions in the built library (no GPU needed):
# tools/sass_extract.sh > profiles/<name>.txt
SO=${1:-mrs_optic_flow_b200/libmof_b200.so}
cuobjdump -sass $SO > /tmp/mof_all.sass
echo "cuobjdump -sass $SO: mnemonic counts over the whole library"
for m in LDTM STTM FADD2 FMUL2 FFMA2 UTMALDG UBLKCP LDGSTS VABSDIFF4 IDP.4A "LDG.E.128" "STS.128" "LDS.64" MUFU.RSQ PRMT SYNCS "BAR.SYNC" "BAR.ARV"; do
  printf "%-12s %8d\n" "$m" "$(grep -c -- "$m" /tmp/mof_all.sass)"
done
echo
echo "per kernel (function : instructions, LDTM, STTM, packed FP32x2, LDGSTS, UTMALDG)"
awk '/Function :/{name=$3} /^[ \t]+\/\*[0-9a-f]+\*\/[ \t]+[A-Z@]/{n[name]++; if($0~/LDTM/)a[name]++; if($0~/STTM/)b[name]++; if($0~/F(ADD|MUL|FMA)2/)c[name]++; if($0~/LDGSTS/)d[name]++; if($0~/UTMALDG/)e[name]++} END{for(k in n) printf "%s : %d %d %d %d %d %d\n", k, n[k], a[k], b[k], c[k], d[k], e[k]}' /tmp/mof_all.sass | sort | cut -c1-140
echo
echo "first tensor-memory / TMA / packed sites"
grep -m3 "LDTM" /tmp/mof_all.sass | cut -c1-110
grep -m3 "STTM" /tmp/mof_all.sass | cut -c1-110
grep -m3 "UTMALDG" /tmp/mof_all.sass | cut -c1-110
grep -m3 "FFMA2" /tmp/mof_all.sass | cut -c1-110
grep -m2 "LDGSTS" /tmp/mof_all.sass | cut -c1-110
