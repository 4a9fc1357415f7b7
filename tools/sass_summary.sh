#!/bin/bash
# profiles/sass_summary.txt: per kernel of liborlg.so - SASS instructions, registers, TMA bulk copies (UBLKCP),
# mbarrier waits (SYNCS), local-memory instructions.   tools/sass_summary.sh > profiles/sass_summary.txt
cd "$(dirname "$0")/.."
lib=optical_rl_gym_qot_aware_b200/liborlg.so
echo "# $(basename $lib): cuobjdump -sass / -res-usage (sm_100a), $(date -u +%Y-%m-%d)"
cuobjdump -res-usage $lib 2>/dev/null | awk '/Function/{sub(/:$/,"",$2); name=$2} /REG:/{for(i=1;i<=NF;i++){if($i ~ /^REG:/)r=$i; if($i ~ /^SHARED:/)s=$i; if($i ~ /^STACK:/)st=$i}; print name, r, s, st}' | sort > /tmp/sass_res.txt
cuobjdump -sass $lib 2>/dev/null | awk '
  /Function :/ {name=$3}
  /^ +\/\*[0-9a-f]+\*\/ / {n[name]++; if ($0 ~ /UBLKCP/) u[name]++; if ($0 ~ /SYNCS/) y[name]++; if ($0 ~ / LDL| STL/) l[name]++; if ($0 ~ /BSYNC/) b[name]++}
  END {for (k in n) printf "%s insts=%d UBLKCP=%d SYNCS=%d LDL/STL=%d BSYNC=%d\n", k, n[k], u[k], y[k], l[k], b[k]}' | sort > /tmp/sass_cnt.txt
join /tmp/sass_cnt.txt /tmp/sass_res.txt | while read -r name rest; do echo "$(echo $name | c++filt | cut -c1-110) | $rest"; done
