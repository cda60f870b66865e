#!/bin/bash
# Opcode digest of every kernel in the shipped library (no GPU needed): registers / shared memory from the ELF, static
# instruction count and the Blackwell-specific opcodes (TMA tensor / bulk copies, mbarrier, cp.async, packed f32, warp votes).
#   tools/sass_digest.sh > profiles/rNN_sass_digest.txt
set -e
cd "$(dirname "$0")/.."
SO=ravengine_b200/librve_b200.so
echo "# $(sha256sum $SO | cut -c1-16)  $SO  ($(stat -c %s $SO) bytes), $(nvcc --version | tail -2 | head -1)"
echo "# kernel | regs | static instr | UTMALDG (TMA tensor) | UBLKCP (TMA bulk) | SYNCS (mbarrier) | LDGSTS (cp.async) | FFMA2/FMUL2/FADD2 | VOTE/MATCH | ATOM/RED | BAR"
cuobjdump -res-usage $SO 2>/dev/null | awk '/Function /{name=$2; sub(/:$/,"",name)} /REG:/{for(i=1;i<=NF;i++) if($i ~ /^REG:/){split($i,a,":"); regs[name]=a[2]}} END{for(k in regs) print k, regs[k]}' > /tmp/_regs.txt
cuobjdump -sass $SO | awk '
  /Function : /{name=$3}
  /^[[:space:]]+\/\*[0-9a-f]{4}\*\//{
    line=$0; sub(/^[[:space:]]+\/\*[0-9a-f]{4}\*\/[[:space:]]+/,"",line); op=line; if (op ~ /^@/) {sub(/^@[!A-Z0-9]+[[:space:]]+/,"",op)}
    sub(/[[:space:];].*/,"",op); n[name]++
    if (op ~ /^UTMALDG/) t[name]++; if (op ~ /^UBLKCP/) b[name]++; if (op ~ /^SYNCS/) s[name]++; if (op ~ /^LDGSTS/) l[name]++;
    if (op ~ /^(FFMA2|FMUL2|FADD2)/) f[name]++; if (op ~ /^(VOTE|MATCH)/) v[name]++; if (op ~ /^(ATOM|RED|ATOMG|ATOMS)/) a[name]++; if (op ~ /^BAR/) r[name]++
  }
  END{for(k in n) printf "%s %d %d %d %d %d %d %d %d %d\n", k, n[k], t[k], b[k], s[k], l[k], f[k], v[k], a[k], r[k]}' | sort > /tmp/_ops.txt
while read k n t b s l f v a r; do
  regs=$(grep -F "$k " /tmp/_regs.txt | head -1 | awk '{print $2}')
  short=$(echo "$k" | c++filt | sed -E 's/\(.*//; s/.*:://; s/^void //')
  printf "%-34s | %3s | %5d | %3d | %3d | %3d | %3d | %3d | %3d | %3d | %3d\n" "$short" "${regs:-?}" $n $t $b $s $l $f $v $a $r
done < /tmp/_ops.txt
