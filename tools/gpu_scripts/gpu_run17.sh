mkdir -p gpurun_out
GMG_FLOW_CLAIM=1 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
for cfg in "16 8 8 64 0" "16 8 8 64 1" "16 8 16 64 1" "16 8 32 0 1" "20 8 8 64 1" "12 8 8 64 1" "16 16 8 64 1" "16 8 8 0 1" "24 8 16 32 1"; do
  set -- $cfg
  echo "FAST $1 GROUP $2 XGROUP $3 XSLEEP $4 CLAIM $5"
  GMG_FLOW_FAST=$1 GMG_FLOW_GROUP=$2 GMG_FLOW_XGROUP=$3 GMG_FLOW_XSLEEP=$4 GMG_FLOW_CLAIM=$5 timeout 300 python tools/prof_build.py C3 1.0 exact 0 2 2>&1 | tail -1
done
GMG_FLOW_CLAIM=1 timeout 300 python tools/prof_build.py C4 0.1 exact 0 2 2>&1 | tail -1
timeout 300 python tools/prof_build.py C4 0.1 exact 0 2 2>&1 | tail -1
