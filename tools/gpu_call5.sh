set -x
python -m pytest tests/test_yolo_gpu.py -m gpu -q -x 2>&1 | tail -5 > gpurun_out/r02_tests5.log
python bench.py --model yolov8n --steps 20 --warmup 5 > gpurun_out/r02d_bench_yolov8n.json 2> gpurun_out/r02d_bench_yolov8n.err
rm -f gpurun_out/r02d_serve.jsonl
python tools/serve_sim.py --models resnet --gpus 1 --rate 64000 --seconds 5 --clients 2 --out gpurun_out/r02d_serve.jsonl > /dev/null 2> gpurun_out/r02d_serve_a.err
python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 5 --clients 3 --out gpurun_out/r02d_serve.jsonl > /dev/null 2> gpurun_out/r02d_serve_b.err
python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 5 --clients 3 --pipeline-depth 3 --out gpurun_out/r02d_serve.jsonl > /dev/null 2> gpurun_out/r02d_serve_c.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 1 --rates 20000,6000,3000 --seconds 5 --clients 1 --out gpurun_out/r02d_serve.jsonl > /dev/null 2> gpurun_out/r02d_serve_d.err
python tools/serve_sim.py --models resnet --gpus 1 --slope 3000 --seconds 8 --clients 2 --out gpurun_out/r02d_serve.jsonl > /dev/null 2> gpurun_out/r02d_serve_e.err
