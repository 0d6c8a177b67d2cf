#!/bin/bash
# Per-kernel SASS evidence: counts of the tensor-core / TMEM / TMA / peer-memory instructions in libnxd_b200.so.
# Usage: tools/sass_evidence.sh > profiles/sass_evidence.txt    (no GPU needed)
LIB=${1:-neuronx_distributed_llama3_2_b200/ops/libnxd_b200.so}
echo "# cuobjdump -sass $LIB : instruction counts per kernel (sm_100a)"
echo "# UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st (TMEM), UTMALDG/UTMASTG/UTMAREDG = TMA tensor load/store/reduce,"
echo "# UBLKCP = cp.async.bulk (1-D), SYNCS = mbarrier ops, HMMA = mma.sync (portable fallback kernels), MEMBAR.*SYS / *.SYS = system-scope (peer) traffic"
cuobjdump -sass "$LIB" | awk '
/Function :/ { if (name != "") flush(); name = $3; delete c; total = 0; next }
/^[ \t]+\/\*[0-9a-f]+\*\// {
  total++
  if ($0 ~ /UTC[A-Z]*MMA/) c["UTCMMA"]++
  if ($0 ~ /LDTM/) c["LDTM"]++
  if ($0 ~ /STTM/) c["STTM"]++
  if ($0 ~ /UTMALDG/) c["UTMALDG"]++
  if ($0 ~ /UTMASTG/) c["UTMASTG"]++
  if ($0 ~ /UTMAREDG/) c["UTMAREDG"]++
  if ($0 ~ /UBLKCP/) c["UBLKCP"]++
  if ($0 ~ /UBLKRED/) c["UBLKRED"]++
  if ($0 ~ /SYNCS/) c["SYNCS"]++
  if ($0 ~ /[ \t]HMMA/) c["HMMA"]++
  if ($0 ~ /\.SYS/) c["SYS"]++
  if ($0 ~ /UCGABAR|CGA/) c["CLUSTER"]++
}
function flush() {
  line = sprintf("%-110s insts=%-6d", substr(name, 1, 110), total)
  n = split("UTCMMA LDTM STTM UTMALDG UTMASTG UTMAREDG UBLKCP UBLKRED SYNCS HMMA SYS CLUSTER", keys, " ")
  for (i = 1; i <= n; i++) if (c[keys[i]] > 0) line = line sprintf(" %s=%d", keys[i], c[keys[i]])
  print line
}
END { if (name != "") flush() }' | sed 's/_ZN[0-9]*_GLOBAL__N__[0-9a-f]*_[0-9]*_//' | c++filt 2>/dev/null | sort
