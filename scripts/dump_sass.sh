#!/bin/bash
# SASS evidence for the Blackwell-native kernels: mnemonic counts per object file into profiles/sass/.
cd "$(dirname "$0")/.."
mkdir -p profiles/sass
for f in gemm_tc comm stencil pagerank mlp_elem mlp_runtime; do
  cuobjdump -sass build/obj/$f.o > /tmp/$f.sass 2>/dev/null
  {
    echo "# $f.cu — SASS mnemonic counts (cuobjdump -sass build/obj/$f.o, sm_100a)"
    grep -oE "^\s+/\*[0-9a-f]+\*/\s+[A-Z0-9_.]+" /tmp/$f.sass | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -40
    echo; echo "# tensor-core / TMA / multimem / peer instructions"
    grep -oE "UTC[A-Z]*MMA[.A-Z0-9_]*|UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|LDTM[.A-Z0-9_x]*|UTCBAR[.A-Z0-9_]*|SYNCS[.A-Z0-9_]*|LDGMC[.A-Za-z0-9_]*|REDG[.A-Z0-9_]*|LDG.E[.A-Z0-9_]*SYS[.A-Z0-9_]*|LD.E[.A-Z0-9_]*SYS[.A-Z0-9_]*" /tmp/$f.sass | sort | uniq -c | sort -rn
  } > profiles/sass/$f.txt
done
ls -la profiles/sass
