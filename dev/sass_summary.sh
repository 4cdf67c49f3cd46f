#!/bin/bash
# profiles/sass_summary.txt: per object file of the product library, how many tcgen05 / TMA / TMEM SASS
# instructions the sm_100a build contains (cuobjdump -sass | grep -c); proves the tensor-core kernels are
# genuine tcgen05 (UTCHMMA = tcgen05.mma kind::f16, UTMALDG = cp.async.bulk.tensor, LDTM/STTM = tcgen05.ld/st).
cd "$(dirname "$0")/../cp2_b200/csrc" || exit 1
out=../../profiles/sass_summary.txt
{
  echo "# cuobjdump -sass <object> | grep -c <mnemonic>   (nvcc $(nvcc --version | grep -o 'release [0-9.]*'), -gencode arch=compute_100a,code=sm_100a)"
  printf "%-22s %8s %8s %8s %8s %8s %10s %8s\n" object UTCHMMA UTMALDG LDTM STTM UTCBAR SYNCS.ARR MUFU.EX2
  for o in dense_umma.o dense_small.o dense_wsmall.o dense_wstream.o densecl_match.o global_umma.o weighted.o metrics.o prep.o composite.o ema.o enqueue.o combine.o ids.o mirror.o densecl.o; do
    [ -f $o ] || continue
    s=$(cuobjdump -sass $o)
    c() { echo "$s" | grep -c "$1"; }
    printf "%-22s %8d %8d %8d %8d %8d %10d %8d\n" $o $(c UTCHMMA) $(c UTMALDG) $(c LDTM) $(c STTM) $(c UTCBAR) $(c "SYNCS.ARRIVE") $(c "MUFU.EX2")
  done
  echo
  echo "# per kernel (entry symbol) of the tensor-core objects"
  for o in dense_umma.o dense_small.o dense_wsmall.o dense_wstream.o densecl_match.o global_umma.o; do
    [ -f $o ] || continue
    cuobjdump -sass $o | awk -v obj=$o '
      /Function :/ { name=$3 }
      /UTCHMMA/ { m[name]++ } /UTMALDG/ { t[name]++ } /LDTM/ { l[name]++ } /STTM/ { s[name]++ }
      END { for (n in m) printf "%-14s UTCHMMA %4d  UTMALDG %4d  LDTM %4d  STTM %4d  %s\n", obj, m[n], t[n], l[n], s[n], substr(n,1,110) }'
  done
} > $out
cat $out
