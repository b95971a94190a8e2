#!/bin/bash
# SASS evidence for the two gather variants: instruction mix of row_move_kernel<0,0,0> (LDG/STG.128
# gather, the default) and row_gather_bulk_kernel (TMA bulk copies: UBLKCP + mbarrier), from the built library.
LIB=reth_b200/_lib/libreth_replay.so
OUT=profiles/r02_sass_gather.md
cuobjdump -sass $LIB > /tmp/reth_sass.txt
fn() { awk -v pat="$1" '/Function :/{p = index($0, pat) > 0} p' /tmp/reth_sass.txt; }
mix() { grep -oE "^\s+/\*[0-9a-f]{4}\*/\s+(@!?U?P[0-9T]+ )?[A-Z0-9_.]+" | sed -E 's/.*\*\/\s+(@!?U?P[0-9T]+ )?//' | sed -E 's/\.(E|U8|U16|S8|S16|CONSTANT|STRONG|GPU|SYS|SM|LTC[0-9A-Z]+|EF|EL|EN|64|128|32|LU|NA)\b//g' | sort | uniq -c | sort -rn | head -${2:-22}; }
{
echo "# SASS of the gather kernels (sm_100a, \`cuobjdump -sass $LIB\`)"
echo
for K in "row_move_kernelILb0ELb0ELb0EE" "row_gather_bulk_kernel" "per_sample_kernel" "per_update_kernelILi5ELi2ELb0EE"; do
  echo "## $K"; echo; echo '```'
  echo "instructions: $(fn $K | grep -cE '^\s+/\*[0-9a-f]{4}\*/')"
  echo "-- opcode mix (top) --"; fn $K | mix
  echo "-- memory / async instructions as emitted --"
  fn $K | grep -oE "(LDG|STG|LD|ST|UBLKCP|SYNCS|ATOMG|RED|LDS|STS|UTMA|CCTL|MEMBAR|ERRBAR|UCGABAR|BAR)[A-Z0-9_.]*" | sort | uniq -c | sort -rn | head -20
  echo '```'; echo
done
echo "## first LDG/STG.128 group of row_move_kernel<false,false,false> (4 independent 16-byte loads, then 4 stores)"; echo; echo '```'
fn "row_move_kernelILb0ELb0ELb0EE" | grep -E "LDG|STG" | sed -E 's/\s+\/\* 0x[0-9a-f]+ \*\/$//' | head -12
echo '```'; echo
echo "## bulk-copy issue of row_gather_bulk_kernel (cp.async.bulk global->shared on an mbarrier, shared->global in bulk groups)"; echo; echo '```'
fn "row_gather_bulk_kernel" | grep -E "UBLKCP|SYNCS|UTMA" | sed -E 's/\s+\/\* 0x[0-9a-f]+ \*\/$//' | head -16
echo '```'
} > $OUT
wc -l $OUT
