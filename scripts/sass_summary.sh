#!/bin/bash
# Per-kernel counts of the Blackwell-native SASS mnemonics in the built library (evidence that the contraction kernels are
# tcgen05 / TMEM / TMA code):  bash scripts/sass_summary.sh > profiles/sass_summary.txt
LIB=${1:-ace-step-1.5-ggml_b200/acestep_b200/lib/libacestep_b200.so}
echo "# $(basename $LIB): cuobjdump -sass, per kernel: UTCHMMA (tcgen05.mma; .2CTA = cta_group::2), LDTM/STTM (tcgen05.ld/st), UTMALDG/UTMAREDG (TMA tensor load / reduce), UBLKCP (bulk copy), HMMA (mma.sync)"
cuobjdump -sass "$LIB" | awk '
  /Function :/ { if (name != "") emit(); name=$3; delete c; next }
  { for (k in pat) if ($0 ~ pat[k]) c[k]++ }
  function emit(   s, k, any) { s=""; any=0; for (i=1;i<=n;i++) { k=ord[i]; if (c[k]>0) { s=s sprintf(" %s=%d", k, c[k]); any=1 } } if (any) print s " | " name }
  BEGIN { n=split("UTCHMMA UTCHMMA.2CTA LDTM STTM UTMALDG UTMAREDG UBLKCP HMMA", ord, " ");
          pat["UTCHMMA"]="UTCHMMA"; pat["UTCHMMA.2CTA"]="UTCHMMA\\.2CTA"; pat["LDTM"]="LDTM"; pat["STTM"]="STTM"; pat["UTMALDG"]="UTMALDG"; pat["UTMAREDG"]="UTMAREDG"; pat["UBLKCP"]="UBLKCP"; pat["HMMA"]="[^C]HMMA\\." }
  END { if (name != "") emit() }' | c++filt | sed -E 's/\(anonymous namespace\):://g; s/\(CUtensorMap.*//; s/\((const )?[a-z_A-Z:]+ ?[*&]?,.*//' | sort -t'|' -k2
echo "# totals"
for m in UTCHMMA 'UTCHMMA\.2CTA' LDTM STTM UTMALDG UTMAREDG UBLKCP; do echo "$m $(cuobjdump -sass "$LIB" | grep -c "$m")"; done
