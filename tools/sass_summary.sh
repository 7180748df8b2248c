#!/bin/bash
# Per-kernel counts of the SASS mnemonics that prove the Blackwell paths (B200_PROFILING.md): UTCHMMA (tcgen05.mma),
# LDTM (tcgen05.ld), UTMALDG (tensor-map TMA), UBLKCP (bulk TMA), DMMA (fp64 tensor core), SHFL, from the shipped library.
LIB="${1:-iddiff_b200/lib/libiddiff_b200.so}"
cuobjdump -sass "$LIB" | awk '
/Function :/ { fn=$3; order[++n]=fn }
/UTCHMMA/ {c[fn,"UTCHMMA"]++} /LDTM/ {c[fn,"LDTM"]++} /UTMALDG/ {c[fn,"UTMALDG"]++} /UBLKCP/ {c[fn,"UBLKCP"]++}
/DMMA/ {c[fn,"DMMA"]++} /SHFL/ {c[fn,"SHFL"]++} /MUFU/ {c[fn,"MUFU"]++}
END { printf "%-8s %-6s %-8s %-7s %-6s %-6s %-6s %s\n","UTCHMMA","LDTM","UTMALDG","UBLKCP","DMMA","SHFL","MUFU","kernel";
  for(i=1;i<=n;i++){fn=order[i]; printf "%-8d %-6d %-8d %-7d %-6d %-6d %-6d %s\n", c[fn,"UTCHMMA"],c[fn,"LDTM"],c[fn,"UTMALDG"],c[fn,"UBLKCP"],c[fn,"DMMA"],c[fn,"SHFL"],c[fn,"MUFU"],fn} }' | (read h; echo "$h"; sort -k8 | while read a b c d e f g k; do printf "%-8s %-6s %-8s %-7s %-6s %-6s %-6s %s\n" $a $b $c $d $e $f $g "$(echo $k | c++filt)"; done)
