#!/bin/bash
# usage (here): tools/hotspots.sh <ncu-rep> <mangled-kernel-substring> [top]   -- needs the same libprt_b200.so build
set -e
REP=$1; KSUB=$2; TOP=${3:-40}
T=$(mktemp -d)
ncu -i $REP --page source --csv > $T/src.csv 2>/dev/null
(cd $T && cuobjdump -xelf all /root/repo/prosper_b200/libprt_b200.so >/dev/null 2>&1 && nvdisasm -g -c prt_api.sm_100a.cubin > all.sass 2>/dev/null)
awk -v k="$KSUB" '/^\/\/-+ \.text\./{p=index($0,k)>0} p' $T/all.sass > $T/k.sass
python /root/repo/tools/sass_hotspots.py $T/src.csv $T/k.sass $TOP $4
rm -rf $T
