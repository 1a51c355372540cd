run() { VO_NVCC_EXTRA="$1" python -m acs_visual_odometry_b200.build --force 2>&1 | grep -A3 "k_stencilILb0ELb0ELb1" | grep -E "registers" 
  echo "== $2 $1"; python tools/kernel_times.py 256 ${3:-64} 32 | python -c "import json,sys; d=json.load(sys.stdin); print(d['us_per_launch']['stencil'], d['frames_per_s'])"; }
run "" row116
run "-DVO_NO_SKIP" row116
run "-DVO_CTAS=5" row116
run "-DVO_NO_SKIP" row116 71
run "-DVO_NO_SKIP" row116 107
cp tools/stencil_pair.cu.txt acs_visual_odometry_b200/csrc/stencil.cu
run "-DVO_CTAS=5" pair116
