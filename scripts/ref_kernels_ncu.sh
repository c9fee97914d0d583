#!/bin/bash
# Device time of the reference's own three kernels (kernel.cu:192-195) on one B200: ncu on the
# unmodified, shim-built binary. Run under gpurun; parse with scripts/ref_kernels_ncu_parse.py.
set -e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for WH in "7680 4320" "1920 1080"; do
  set -- $WH
  python - "$1" "$2" <<'PY'
import sys
sys.path.insert(0, "baseline")
from synth_np import synth_image, write_ppm_from_bgr
W, H = int(sys.argv[1]), int(sys.argv[2])
write_ppm_from_bgr(f"/tmp/ref_in_{W}x{H}.ppm", synth_image(W, H, 3, seed=1))
PY
  export B2I_REF_INPUT=/tmp/ref_in_$1x$2.ppm B2I_REF_OUTDIR=/tmp
  baseline/_ref/ref_kernel_cu > gpurun_out/ref_plain_$1x$2.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
      --log-file gpurun_out/ref_ncu_$1x$2.csv baseline/_ref/ref_kernel_cu > gpurun_out/ref_ncu_$1x$2.log 2>&1
done
