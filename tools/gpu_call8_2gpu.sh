set -x
rm -f gpurun_out/r02g_serve2.jsonl
python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 4 --clients 3 --out gpurun_out/r02g_serve2.jsonl > /dev/null 2> gpurun_out/r02g_serve_a.err
python tools/serve_sim.py --models resnet --gpus 2 --rate 170000 --seconds 4 --clients 6 --out gpurun_out/r02g_serve2.jsonl > /dev/null 2> gpurun_out/r02g_serve_b.err
