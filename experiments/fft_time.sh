#!/bin/bash
# config-2 timings of the general FFT kernels for several builds in one gpurun call: experiments/fft_time.sh lib1.so lib2.so ...
cd "$(dirname "$0")/.."
for lib in "$@"; do
  if [ "$lib" = "-" ]; then unset KK_LIB; else export KK_LIB=$PWD/$lib; fi
  timeout 300 python experiments/fft_ab.py gpurun_out/fft_time_tmp.json > /dev/null 2> gpurun_out/fft_time.err || tail -3 gpurun_out/fft_time.err
  python - "$lib" <<'PY'
import json, sys
d = json.load(open("gpurun_out/fft_time_tmp.json"))["B256_n720000"]
print(sys.argv[1], {k: round(v, 4) for k, v in d.items() if k.endswith("_ms")})
PY
done
