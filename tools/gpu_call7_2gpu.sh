set -x
nproc > gpurun_out/nproc2.txt
rm -f gpurun_out/r02f_serve2.jsonl
python tools/serve_sim.py --models resnet --gpus 2 --rate 170000 --seconds 5 --clients 6 --out gpurun_out/r02f_serve2.jsonl > /dev/null 2> gpurun_out/r02f_serve_a.err
python tools/serve_sim.py --models resnet --gpus 2 --rate 130000 --seconds 5 --clients 4 --out gpurun_out/r02f_serve2.jsonl > /dev/null 2> gpurun_out/r02f_serve_b.err
python tools/serve_sim.py --models resnet,efficientnet,yolov8 --gpus 2 --rates 60000,14000,7000 --seconds 6 --clients 2 --dynamic --out gpurun_out/r02f_serve2.jsonl > /dev/null 2> gpurun_out/r02f_serve_c.err
