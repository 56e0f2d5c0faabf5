#!/bin/bash
# SASS evidence for every translation unit of the device library: per-file mnemonic counts and
# the tensor-core / TMA / multimem / peer instructions, plus FULL listings of the hot kernels.
#   bash scripts/dump_sass.sh        (after python -m cme213_sp18_b200.build)
cd "$(dirname "$0")/.."
mkdir -p profiles/sass
for src in csrc/*.cu; do
  f=$(basename $src .cu)
  [ -f build/obj/$f.o ] || continue
  cuobjdump -sass build/obj/$f.o > /tmp/$f.sass 2>/dev/null
  {
    echo "# $f.cu -- SASS mnemonic counts (cuobjdump -sass build/obj/$f.o, sm_100a), $(date -u +%Y-%m-%d)"
    echo "# kernels:"; grep -oE "Function : [A-Za-z0-9_]+" /tmp/$f.sass | sed 's/Function : /#   /' | c++filt | cut -c1-160
    grep -oE "^\s+/\*[0-9a-f]+\*/\s+[A-Z0-9_.]+" /tmp/$f.sass | awk '{print $2}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -40
    echo; echo "# tensor-core / TMA / multimem / peer / fence instructions"
    grep -oE "UTC[A-Z]*MMA[.A-Z0-9_]*|UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|LDTM[.A-Z0-9_x]*|UTCBAR[.A-Z0-9_]*|SYNCS[.A-Z0-9_]*|LDGMC[.A-Za-z0-9_]*|STGMC[.A-Za-z0-9_]*|REDG[.A-Z0-9_]*|RED\.E[.A-Z0-9_]*|LDG?\.E[.A-Z0-9_]*SYS[.A-Z0-9_]*|MEMBAR[.A-Z0-9_]*|DFMA|DMMA[.A-Z0-9_]*" /tmp/$f.sass | sort | uniq -c | sort -rn
  } > profiles/sass/$f.txt
done
# full listings of the hot kernels (one file per kernel)
dump_kernel() { # object, kernel-name substring, output
  cuobjdump -sass -fun "$(cuobjdump -elf build/obj/$1.o 2>/dev/null | grep -oE "_Z[A-Za-z0-9_]*$2[A-Za-z0-9_]*" | sort -u | head -1)" build/obj/$1.o > profiles/sass/$3 2>/dev/null
}
dump_kernel fwd_fused fwd_rider_kernel kernel_fwd_rider.sass
dump_kernel gemm_tc gemm_mnmajor2_kernel kernel_dw1_update.sass
dump_kernel gemm_tc bwd_tail_kernel kernel_bwd_tail.sass
dump_kernel comm update_stream_kernel kernel_update_stream.sass
ls -la profiles/sass | head -30
