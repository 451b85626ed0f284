#!/bin/bash
# Per-kernel counts of the Blackwell-only SASS instructions in the built library (evidence that the GEMM path is
# tcgen05 / TMEM / TMA and which kernels use what).  Usage: scripts/sass_grep.sh > profiles/r02_sass_grep.md
LIB=${1:-deltaconv_b200/libdeltaconv_b200.so}
echo "SASS instruction counts per kernel, \`cuobjdump --dump-sass $LIB\` (sm_100a), $(date -u +%Y-%m-%d)."
echo
echo "UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), UTMALDG / UTMASTG = TMA tensor load / store, LDTM / STTM = tcgen05.ld / st (tensor memory),"
echo "UTCBAR = tcgen05.commit (mbarrier arrive, .MULTICAST = to both CTAs of a pair), SYNCS = mbarrier ops, MUFU = special-function unit, REDUX/CREDUX = warp reductions."
echo
echo "| kernel | instructions | UTCHMMA | UTCHMMA.2CTA | UTMALDG | UTMASTG | LDTM | STTM | UTCBAR | SYNCS | MUFU | ATOM/RED |"
echo "|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|"
cuobjdump --dump-sass "$LIB" 2>/dev/null | awk '
  /Function : /{ if (name != "") emit(); name=$3; n=0; mma=0; mma2=0; ldg=0; stg=0; ldtm=0; sttm=0; bar=0; syncs=0; mufu=0; atom=0; next }
  /^ *\/\*[0-9a-f]+\*\/ /{ n++;
     if ($0 ~ /UTCHMMA\.2CTA|UTCQMMA\.2CTA|UTCIMMA\.2CTA/) mma2++; else if ($0 ~ /UTCHMMA|UTCQMMA|UTCIMMA/) mma++;
     if ($0 ~ /UTMALDG/) ldg++; if ($0 ~ /UTMASTG/) stg++; if ($0 ~ /LDTM/) ldtm++; if ($0 ~ /STTM/) sttm++;
     if ($0 ~ /UTCBAR/) bar++; if ($0 ~ /SYNCS/) syncs++; if ($0 ~ /MUFU/) mufu++; if ($0 ~ / ATOM|ATOMG|ATOMS| RED\.|REDG/) atom++ }
  function emit() { printf "%s\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\n", name, n, mma, mma2, ldg, stg, ldtm, sttm, bar, syncs, mufu, atom }
  END{ if (name != "") emit() }' | while IFS=$'\t' read -r name rest; do
    pretty=$(echo "$name" | c++filt | sed 's/(anonymous namespace):://g; s/(.*//' | cut -c1-90)
    echo "| \`$pretty\` | $(echo "$rest" | sed 's/\t/ | /g') |"
  done | sort -t'|' -k3 -n -r
