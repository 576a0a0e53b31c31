set -x
rm -f gpurun_out/r02k_serve8.jsonl
python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 4 --clients 3 --out gpurun_out/r02k_serve8.jsonl > /dev/null 2> gpurun_out/r02k_serve_a.err
python tools/serve_sim.py --models resnet --gpus 8 --rate 680000 --seconds 4 --clients 20 --out gpurun_out/r02k_serve8.jsonl > /dev/null 2> gpurun_out/r02k_serve_b.err
python tools/serve_sim.py --models resnet --gpus 8 --rate 470000 --seconds 4 --clients 16 --out gpurun_out/r02k_serve8.jsonl > /dev/null 2> gpurun_out/r02k_serve_b2.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 240000,60000,30000 --seconds 5 --clients 6 --out gpurun_out/r02k_serve8.jsonl > /dev/null 2> gpurun_out/r02k_serve_c.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 240000,60000,30000 --seconds 7 --clients 6 --dynamic --out gpurun_out/r02k_serve8.jsonl > /dev/null 2> gpurun_out/r02k_serve_d.err
