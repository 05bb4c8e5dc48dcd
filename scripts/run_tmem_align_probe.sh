# usage: bash scripts/run_tmem_align_probe.sh [full]   (on a B200; one process per case)
P=tests/_build/tmem_align_probe
if [ "$1" = full ]; then
for op in st ld; do for N in 1 2 4 8 16 32; do for c in 1 2 4 8 16; do
  if [ "$op" = st ] && [ $N = 32 ]; then continue; fi
  timeout 20 $P $op $N $c; done; done; done
for c in 16 8 4 2 1; do timeout 20 $P mma 16 $c 0; done
fi
# accumulator (D) column alignment; A operand at column 8 (TMEM) -- D columns must not overlap A
for N in 16 32 48 64; do for d in 16 32 48 64 80 400 448; do timeout 20 $P mma $N 8 $d; done; done
