#!/bin/sh
# Tuning helper: build a kernel variant of the library next to the product one.
#   ./build_variant.sh NAME -DVIT_MINBLK=12 -DDEMOD_MINBLK=6 ...   ->  ../variants/libw80211rx_NAME.so
# Select it with bench.py --resident-only --lib <path>; never used by tests or the default bench.
set -e
cd "$(dirname "$0")"
name=$1; shift
mkdir -p ../variants /tmp/w80211_$name
for f in hdr demod viterbi post iqconv detect capi; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -Xcudafe --diag_suppress=177 -Xcudafe --diag_suppress=20091 "$@" -c $f.cu -o /tmp/w80211_$name/$f.o &
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC -x cu -c tables.cc -o /tmp/w80211_$name/tables.o &
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libw80211rx_$name.so /tmp/w80211_$name/*.o -lcudart -ldl
