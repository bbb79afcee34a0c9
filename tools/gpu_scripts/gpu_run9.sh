mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
for cfg in "16 8 8 64" "16 8 8 0" "24 8 8 64" "24 8 16 64" "20 16 16 64" "24 16 32 64" "12 8 8 64" "28 8 8 64" "24 4 8 64"; do
  set -- $cfg
  echo "FAST $1 GROUP $2 XGROUP $3 XSLEEP $4"
  GMG_FLOW_FAST=$1 GMG_FLOW_GROUP=$2 GMG_FLOW_XGROUP=$3 GMG_FLOW_XSLEEP=$4 timeout 300 python tools/prof_build.py C3 0.25 exact 0 2 2>&1 | tail -1
  GMG_FLOW_FAST=$1 GMG_FLOW_GROUP=$2 GMG_FLOW_XGROUP=$3 GMG_FLOW_XSLEEP=$4 timeout 300 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
done
timeout 600 python tools/prof_build.py C3 1.0 parallel 0 2 2>&1 | tail -1
