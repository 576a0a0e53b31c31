cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -3 gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/f_smoke.log 2>&1; tail -1 gpurun_out/f_smoke.log
timeout 600 python bench.py > gpurun_out/r02f_bench_resnet50.json 2> gpurun_out/r02f_bench_resnet50.err; echo "resnet rc=$?"
timeout 600 python bench.py --model shufflenet_v2_x1_0 > gpurun_out/r02f_bench_shufflenet_v2_x1_0.json 2> gpurun_out/r02f_bench_shufflenet_v2_x1_0.err; echo "shuffle rc=$?"
timeout 600 python bench.py --model yolov8n --no-serving > gpurun_out/r02f_bench_yolov8n.json 2> gpurun_out/r02f_bench_yolov8n.err; echo "yolo rc=$?"
timeout 600 python bench.py --model efficientnet_v2_s --no-serving > gpurun_out/r02f_bench_efficientnet_v2_s.json 2> gpurun_out/r02f_bench_efficientnet_v2_s.err; echo "effnet rc=$?"
