set -x
rm -f gpurun_out/r02l_serve2.jsonl
python tools/serve_sim.py --models resnet --gpus 2 --rate 400000 --seconds 4 --clients 4 --out gpurun_out/r02l_serve2.jsonl > /dev/null 2> gpurun_out/r02l_serve_a.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 2 --rates 60000,14000,7000 --seconds 4 --clients 2 --out gpurun_out/r02l_serve2.jsonl > /dev/null 2> gpurun_out/r02l_serve_b.err
python -m pytest tests/test_worker_gpu.py tests/test_collate_resize_gpu.py -m gpu -q -x 2>&1 | tail -3 > gpurun_out/r02l_tests.log
