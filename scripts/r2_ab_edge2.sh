set -x
bash scripts/r2_ab_edge.sh
MACBF_B200_LIB=$PWD/macbf-gnn_b200/libmacbf_b200_trace.so python macbf-gnn_b200/tools/ce2_trace.py > gpurun_out/ce2_trace.txt 2>&1; tail -3 gpurun_out/ce2_trace.txt
