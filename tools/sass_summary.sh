#!/bin/bash
# counts of the Blackwell-native SASS mnemonics per object file and kernel (cuobjdump -sass on the built objects):
# UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG / UTMAREDG = TMA load / store / reduce,
# HMMA = warp-level mma.sync (kept only where a 16-row problem cannot fill a tcgen05 tile, and for the 3xTF32 circulant
# passes of freq_split3d_mma_kernel in signal.o)
for o in hivae_b200/build/gemm_tc.o hivae_b200/build/attn_tc2.o hivae_b200/build/attn_tc.o hivae_b200/build/attn_bwd_tc.o hivae_b200/build/attn_bwd.o hivae_b200/build/attn_small.o hivae_b200/build/attn_small_bwd.o hivae_b200/build/signal.o; do
  echo "== $o   (BRA.U.ANY = ELECT / R2UR.BROADCAST issue loops: $(cuobjdump -sass $o | grep -c BRA.U.ANY))"
  cuobjdump -sass $o | awk '
    /Function :/ { fn=$3; sub(/^_ZN[0-9]+_GLOBAL__N__[0-9a-f_]+cu_[0-9a-f]+/, "", fn); next }
    { for (i=1;i<=NF;i++) if ($i ~ /^(UTCHMMA|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|HMMA|UTCBAR|SYNCS|MUFU\.EX2)/) { split($i,a,"."); k=a[1]; if (k=="MUFU") k="MUFU.EX2"; c[fn" "k]++ } }
    END { for (x in c) print c[x], x }' | sort -k2,2 -k3,3 | awk '{printf "  %-90s %-10s %d\n", substr($2,1,90), $3, $1}'
done
