#!/bin/bash
# usage: tools/dump_sass.sh   — SASS listings + mnemonic histograms of the hot kernels into profiles/sass/
# (cuobjdump -sass on the in-tree library; run after python -m concrete_b200.build)
cd "$(dirname "$0")/.."
LIB=concrete_b200/libconcrete_b200.so
OUT=profiles/sass
mkdir -p $OUT
dump() { # <output name> <grep pattern on the mangled kernel name>
  local sym
  sym=$(cuobjdump -elf $LIB 2>/dev/null | grep -o "\.text\.[A-Za-z0-9_]*" | sort -u | sed 's/^\.text\.//' | grep -E "$2" | head -1)
  [ -z "$sym" ] && { echo "no kernel matches $2"; return; }
  cuobjdump -sass -fun "$sym" $LIB 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,6}\*/" | sed -E 's#^\s+/\*([0-9a-f]{4,6})\*/\s+#\1 #; s#\s*/\* 0x[0-9a-f]+ \*/##' | gzip -9 > $OUT/$1.sass.gz
  { echo "# $1: $(echo $sym | c++filt)"; echo "# instructions: $(zcat $OUT/$1.sass.gz | wc -l)"; echo "# mnemonic histogram (top 40):"
    zcat $OUT/$1.sass.gz | awk '{m=$2; if (m ~ /^@/) m=$3; sub(/;$/,"",m); print m}' | sort | uniq -c | sort -rn | head -40
    echo "# Blackwell-specific: $(zcat $OUT/$1.sass.gz | grep -cE 'LDTM|STTM') LDTM/STTM (tcgen05.ld/st), $(zcat $OUT/$1.sass.gz | grep -c UBLKCP) UBLKCP (cp.async.bulk), $(zcat $OUT/$1.sass.gz | grep -c SYNCS) SYNCS (mbarrier), $(zcat $OUT/$1.sass.gz | grep -cE 'UTCIMMA|UTCHMMA|UTCQMMA') UTC*MMA (tcgen05.mma), $(zcat $OUT/$1.sass.gz | grep -c UTMALDG) UTMALDG (TMA tensor), $(zcat $OUT/$1.sass.gz | grep -cE 'LDL|STL') LDL/STL (local memory), $(zcat $OUT/$1.sass.gz | grep -c 'IMMA') IMMA, $(zcat $OUT/$1.sass.gz | grep -cE 'DFMA|DMUL|DADD') FP64, $(zcat $OUT/$1.sass.gz | grep -cE 'UCGABAR|CGA') cluster-barrier"
  } > $OUT/$1.summary.txt
  cat $OUT/$1.summary.txt | head -3; tail -1 $OUT/$1.summary.txt
}
dump pbs_kernel_w32v3 'pbs_kernel_w32v3'
dump ks_umma_kernel 'ks_umma_kernel'
dump pbs_kernel_v4_N8192 'pbs_kernel_v4INS_3CfgILi1ELi13ELi4ELi1ELb0.*ELb1ELi1EEE'
dump pbs_kernel_v4_N4096 'pbs_kernel_v4INS_3CfgILi1ELi12ELi4ELi2ELb0.*ELb1ELi1EEE'
dump pbs_kernel_v4_cluster_N16384 'pbs_kernel_v4INS_3CfgILi1ELi14ELi5ELi1ELb1ELi4ELi3ELb1ELb0EEELb1ELi2EEE'
dump pbs_kernel_v4_k2_N1024 'pbs_kernel_v4INS_3CfgILi2ELi10ELi4ELi5ELb0ELi8ELi2ELb0ELb0EEELb1ELi1EEE'
dump pbs_kernel_grp_k2_N1024 'pbs_kernel_grpILi2ELi10ELb1ELi1'
dump mask_fill_kernel 'mask_fill_kernel'
