# ncu --set full of the streaming rating kernels on ladders whose rows do not fit L2:
# config 4's ladder (125M 3v3 / 12.5M players) with K=1 and K=2 pairs per player, and config 3.
set -x
i=0
for a in "125e6 12.5e6 6 x 0 ladder" "125e6 12.5e6 6 x 0 pair" "100e6 10e6 10 zipf 0.02 ladder"; do
  i=$((i+1))
  python scripts/prof_big.py $a > gpurun_out/prof_big_plain$i.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_rate_stream' -c 1 -f -o gpurun_out/r2_big$i python scripts/prof_big.py $a > gpurun_out/prof_big_ncu$i.log 2>&1
  tail -2 gpurun_out/prof_big_ncu$i.log | cut -c1-200
done
