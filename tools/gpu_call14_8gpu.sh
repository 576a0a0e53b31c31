set -x
rm -f gpurun_out/r02m_serve8.jsonl
python tools/serve_sim.py --models resnet --gpus 8 --rate 680000 --seconds 4 --clients 22 --out gpurun_out/r02m_serve8.jsonl > /dev/null 2> gpurun_out/r02m_serve_a.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 8 --rates 240000,56000,28000 --seconds 5 --clients 6 --out gpurun_out/r02m_serve8.jsonl > /dev/null 2> gpurun_out/r02m_serve_b.err
