set -x
python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > gpurun_out/r02_tests4.log
python bench.py --model yolov8n --steps 20 --warmup 5 > gpurun_out/r02c_bench_yolov8n.json 2> gpurun_out/r02c_bench_yolov8n.err
python bench.py --steps 20 --warmup 5 --inflight 3 --no-serving --no-cpu-baseline --sustain-s 0 > gpurun_out/r02c_bench_resnet_inflight3.json 2> gpurun_out/r02c_bench_resnet_inflight3.err
python tools/serve_sim.py --models resnet --gpus 1 --rate 85000 --seconds 4 --clients 3 --out gpurun_out/r02c_serve.jsonl > /dev/null 2> gpurun_out/r02c_serve_a.err
python tools/serve_sim.py --models resnet --gpus 1 --rate 85000 --seconds 4 --clients 3 --pipeline-depth 3 --out gpurun_out/r02c_serve.jsonl > /dev/null 2> gpurun_out/r02c_serve_b.err
python tools/serve_sim.py --models resnet --gpus 1 --rate 60000 --seconds 4 --clients 2 --out gpurun_out/r02c_serve.jsonl > /dev/null 2> gpurun_out/r02c_serve_c.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 1 --rates 20000,8000,4000 --seconds 4 --clients 1 --out gpurun_out/r02c_serve.jsonl > /dev/null 2> gpurun_out/r02c_serve_d.err
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt; df -h /dev/shm >> gpurun_out/nproc.txt
