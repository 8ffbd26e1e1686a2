#!/bin/bash
# Per-kernel SASS evidence (CPU only): which Blackwell-native instructions each kernel of
# libb200ps.so contains.  UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG = TMA tensor load,
# UBLKCP = cp.async.bulk, LDGSTS = cp.async, SYNCS = mbarrier, RED/ATOM = reductions/atomics, MEMBAR = fences,
# STG/LDG.E.64.STRONG.SYS = the 8-byte {value, epoch} packets of the low-latency push, LDGMC = multimem.ld_reduce (NVLS).
lib=${1:-b200ps/_lib/libb200ps.so}
cuobjdump -sass "$lib" | awk '
/Function :/ { fn=$3; next }
fn != "" {
  if (match($0, /UTC[A-Z]*MMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UBLKCP|SYNCS\.[A-Z.0-9]*|HMMA|REDG?\.E[A-Z.0-9]*|RED\.E[A-Z.0-9]*|ATOMG?\.E[A-Z.0-9]*|MEMBAR\.[A-Z.]*|ERRBAR|CCTL\.IVALL|ACQBULK|UCGABAR[A-Z_]*|LDG\.E[A-Z.0-9]*SYS[A-Z.0-9]*|STG\.E[A-Z.0-9]*SYS[A-Z.0-9]*/)) {
    m=substr($0, RSTART, RLENGTH); c[fn" "m]++
  }
}
END { for (k in c) print k, c[k] }' | sort | c++filt | awk '{n=$NF; $NF=""; print}' >/dev/null
cuobjdump -sass "$lib" | awk '
/Function :/ { fn=$3; next }
fn != "" {
  if (match($0, /UTC[A-Z]*MMA|UTCBAR|LDTM|UTMALDG|LDGMC[A-Z.0-9x]*|UBLKCP[A-Z.]*|LDGSTS[A-Z.0-9]*|STG\.E\.64\.STRONG\.SYS|LDG\.E\.64\.STRONG\.SYS|SYNCS\.ARRIVE\.TRANS64|SYNCS\.PHASECHK[A-Z.0-9]*|HMMA|RED\.E\.ADD[A-Z.0-9]*|REDG\.E[A-Z.0-9]*|ATOMG\.E[A-Z.0-9]*|ATOM\.E[A-Z.0-9]*|MEMBAR\.[A-Z.]*/)) {
    m=substr($0, RSTART, RLENGTH); c[fn"\t"m]++
  }
}
END { for (k in c) print k"\t"c[k] }' | sort | while IFS=$'\t' read fn m n; do printf "%-110s %-34s %s\n" "$(echo $fn | c++filt | cut -c1-110)" "$m" "$n"; done
