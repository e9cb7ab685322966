#!/bin/bash
# opcode evidence from the built library (no GPU needed): per kernel family, the counts of the SASS mnemonics that prove the sm_100a features
L=ggml-hexagon_b200/lib/libb200ops.so
echo "# cuobjdump -sass $L ($(git rev-parse --short HEAD)): kernels and the tcgen05 / TMA / bulk-copy / cluster mnemonics they contain"
cuobjdump -sass $L 2>/dev/null | awk '
/Function :/ { fn=$3; next }
{ for (i = 1; i <= NF; i++) { op=$i; sub(/;$/, "", op);
    if (op ~ /^(UTCHMMA|UTCIMMA|UTCQMMA|UTMALDG|UTMASTG|UBLKCP|UTCBAR|LDTM|STTM|UTCATOMSWS|SYNCS|LDGSTS|USETMAXREG|UCGABAR_ARV|UCGABAR_WAIT|ACQBULK|FENCE|ELECT|HMMA|IMMA)/) { split(op, p, "."); key=p[1]; if (op ~ /2CTA/) key=key".2CTA"; if (op ~ /MULTICAST/) key=key".MULTICAST"; c[fn" "key]++ } } }
END { for (k in c) print k, c[k] }' | sort | c++filt 2>/dev/null | awk '{ n=$NF; $NF=""; k=$(NF-1); $(NF-1)=""; name=$0; gsub(/ +$/, "", name); if (length(name) > 110) name=substr(name, 1, 110) "..."; printf "%-114s %-22s %5d\n", name, k, n }'
