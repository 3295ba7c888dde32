#!/bin/bash
# Regenerate profiles/sass_evidence.txt from the built library (no GPU needed).
# Shows the SASS mnemonics that prove the Blackwell paths are what runs:
#   UTCHMMA / UTCBAR / LDTM / UTMALDG  = tcgen05.mma / commit / tcgen05.ld / TMA
#   UBLKCP / LDSM / HMMA               = cp.async.bulk / ldmatrix / mma.sync (linlink forward product)
#   LDGMC / STG ... (multimem)         = NVLink-switch reduction / multicast store
#   FFMA2 / FMUL2 / FADD2, MUFU        = packed fp32 pipe + SFU of the erfc loop
#   *.STRONG.SYS                       = system-scope peer loads/stores (NVLink)
set -e
cd "$(dirname "$0")/.."
LIB=multigrad_b200/lib/libmgb200.so
OUT=profiles/sass_evidence.txt
TMP=$(mktemp)
cuobjdump -sass "$LIB" > "$TMP"
{
  echo "# SASS evidence (tools/sass_evidence.sh: cuobjdump -sass $LIB), sm_100a"
  echo "## instruction counts over the whole library"
  grep -oE "UCGABAR_ARV|UCGABAR_WAIT|HMMA\.16816\.F32\.BF16|LDSM\.16\.M88\.4|UBLKCP[.A-Z0-9]*|UTCHMMA|UTCBAR|UTCATOMSWS|LDTM|UTMALDG|LDGMC[.A-Z0-9x]*|MUFU\.(EX2|RCP)|FFMA2|FMUL2|FADD2|SYNCS\.[A-Z]+\.|ACQBULK|CCTL[.A-Z]*|REDG[.A-Z0-9]*|ST[GS]?\.E\.[A-Z0-9.]*MMIO[.A-Z]*|RED\.E\.[.A-Z0-9]*|MULTIMEM[.A-Z0-9]*" "$TMP" | sort | uniq -c
  for k in "linlink_step_kernelILi64" "nvls_allreduce_f32_kernelILb1" "oneshot_allreduce_f32_kernel" "group_step_kernel"; do
    echo
    echo "## $k: tensor-core / TMA / TMEM / multimem instructions and system-scope (peer) accesses"
    awk -v k="$k" '/Function :/ {on = index($0, k) > 0} on' "$TMP" |
      grep -E "UTCHMMA|UTCBAR|LDTM|UTMALDG|UBLKCP|LDSM|HMMA|LDGMC|STRONG\.SYS|MULTIMEM|REDG|\.MMIO" | head -48
  done
  echo
  echo "## erfbin_step_kernel<11,false>: thread-block cluster / DSMEM reduction (barrier.cluster = UCGABAR_*,"
  echo "## %cluster_ctarank = SR_CgaCtaId, mapa = PRMT 0x654 on the shared address, st.shared::cluster = ST.E.64),"
  echo "## persistent-fit hand-over (LDG/STG.STRONG.GPU on the step counter), tile claims (ATOMG), PDL (ACQBULK/PREEXIT)"
  awk '/Function :/ {on = index($0, "erfbin_step_kernelILi11ELb0") > 0} on' "$TMP" |
    grep -E "UCGABAR|SR_Cga|PRMT.*0x654|ST\.E\.64 |STRONG\.GPU|ATOMG|ACQBULK|PREEXIT|CCTL" | head -40
  echo
  echo "## erfbin_step_kernel<11,false>: packed fp32 main loop (excerpt) and the system-scope LL accesses"
  awk '/Function :/ {on = index($0, "erfbin_step_kernelILi11ELb0") > 0} on' "$TMP" |
    grep -E "FFMA2|FMUL2|FADD2|MUFU|STRONG\.SYS|LDG.*CONSTANT" | head -60
} > "$OUT"
rm -f "$TMP"
wc -l "$OUT"
