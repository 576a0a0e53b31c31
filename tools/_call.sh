cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_conv_gpu.py -k shuffle_tail -q > gpurun_out/a_unit.log 2>&1; echo "unit rc=$?" >> gpurun_out/a_unit.log
grep -E "passed|failed|rc=" gpurun_out/a_unit.log | head
rm -f gpurun_out/a_ab.log
for r in 1 2; do
  for f in 0 1; do
    B200_FUSE_SHUFFLE=$f timeout 300 python tools/model_throughput.py shufflenet_v2_x1_0 512 50 2>&1 | tail -1 | sed "s/^/fuse=$f /" >> gpurun_out/a_ab.log
  done
done
cat gpurun_out/a_ab.log
timeout 600 python bench.py --model shufflenet_v2_x1_0 --steps 1 --warmup 3 --sustain-s 0 --no-serving --no-cpu-baseline --parity-rows 32 > gpurun_out/a_bench1.json 2> gpurun_out/a_bench1.err; echo "bench rc=$?"
if [ -s gpurun_out/a_bench1.json ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/a_launches_shufflenet.csv python bench.py --model shufflenet_v2_x1_0 --steps 1 --warmup 3 --sustain-s 0 --no-serving --no-cpu-baseline --parity-rows 1 > gpurun_out/a_ncu.log 2>&1
fi
