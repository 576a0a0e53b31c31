set -x
nproc > gpurun_out/nproc8.txt; free -g >> gpurun_out/nproc8.txt; nvidia-smi -L >> gpurun_out/nproc8.txt
rm -f gpurun_out/r02e_serve8.jsonl
python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 5 --clients 3 --out gpurun_out/r02e_serve8.jsonl > /dev/null 2> gpurun_out/r02e_serve_a.err
python tools/serve_sim.py --models resnet --gpus 8 --rate 680000 --seconds 5 --clients 24 --out gpurun_out/r02e_serve8.jsonl > /dev/null 2> gpurun_out/r02e_serve_b.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 240000,60000,30000 --seconds 6 --clients 8 --out gpurun_out/r02e_serve8.jsonl > /dev/null 2> gpurun_out/r02e_serve_c.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 240000,60000,30000 --seconds 8 --clients 8 --dynamic --out gpurun_out/r02e_serve8.jsonl > /dev/null 2> gpurun_out/r02e_serve_d.err
