# Builds of libtsrate.so with other compile-time tuning macros, for A/B runs on the GPU box:
#   scripts/build_variants.sh name "-DTS_SORT_LOOKBACK=4 -DTS_SORT_MIN_BLOCKS=4" ...
# -> variants/libtsrate_<name>.so (select with TSRATE_LIB=...).
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  (cd analyzer_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --extended-lambda -shared -Xcompiler -fPIC $flags \
     -o ../../variants/libtsrate_$name.so api.cu gen.cu sched.cu rate.cu stream.cu stream_pair.cu shard.cu static_sched.cu) &
done
wait
ls -la variants/
