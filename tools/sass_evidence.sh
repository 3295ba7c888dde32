#!/bin/bash
# Regenerate profiles/sass_evidence.txt from the built library (no GPU needed).
# Shows the SASS mnemonics that prove the Blackwell paths are what runs:
#   UTCHMMA / UTCBAR / LDTM / UTMALDG  = tcgen05.mma / commit / tcgen05.ld / TMA
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
  grep -oE "UTCHMMA|UTCBAR|UTCATOMSWS|LDTM|UTMALDG|LDGMC[.A-Z0-9x]*|MUFU\.(EX2|RCP)|FFMA2|FMUL2|FADD2|SYNCS\.[A-Z]+\.|ACQBULK|CCTL[.A-Z]*|REDG[.A-Z0-9]*|ST[GS]?\.E\.[A-Z0-9.]*MMIO[.A-Z]*|RED\.E\.[.A-Z0-9]*|MULTIMEM[.A-Z0-9]*" "$TMP" | sort | uniq -c
  for k in "linlink_step_kernelILi64" "nvls_allreduce_f32_kernelILb1" "oneshot_allreduce_f32_kernel" "group_step_kernel"; do
    echo
    echo "## $k: tensor-core / TMA / TMEM / multimem instructions and system-scope (peer) accesses"
    awk -v k="$k" '/Function :/ {on = index($0, k) > 0} on' "$TMP" |
      grep -E "UTCHMMA|UTCBAR|LDTM|UTMALDG|LDGMC|STRONG\.SYS|MULTIMEM|REDG|\.MMIO" | head -40
  done
  echo
  echo "## erfbin_step_kernel<4,false>: packed fp32 main loop (excerpt), PDL, and the system-scope LL accesses"
  awk '/Function :/ {on = index($0, "erfbin_step_kernelILi4ELb0") > 0} on' "$TMP" |
    grep -E "FFMA2|FMUL2|FADD2|MUFU|STRONG\.SYS|ACQBULK|DEPBAR|PREEXIT|LDG.*CONSTANT" | head -60
} > "$OUT"
rm -f "$TMP"
wc -l "$OUT"
