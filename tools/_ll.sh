cd $GRAFT_REPO_ROOT
M=$1; B=$2
timeout 300 python tools/eager_forward.py $M $B > gpurun_out/ll_plain_$M.log 2>&1; echo "plain rc=$?"
if grep -q done gpurun_out/ll_plain_$M.log; then
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02f_eager_${M}_b$B.csv python tools/eager_forward.py $M $B > gpurun_out/ll_ncu_$M.log 2>&1; echo "ncu rc=$?"
fi
