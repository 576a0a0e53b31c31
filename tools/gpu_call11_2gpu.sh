set -x
rm -f gpurun_out/r02j_serve2.jsonl
python tools/serve_sim.py --models resnet --gpus 2 --rate 170000 --seconds 4 --clients 6 --out gpurun_out/r02j_serve2.jsonl > /dev/null 2> gpurun_out/r02j_serve_a.err
python tools/serve_sim.py --models resnet --gpus 2 --rate 170000 --seconds 4 --clients 6 --no-shards --out gpurun_out/r02j_serve2.jsonl > /dev/null 2> gpurun_out/r02j_serve_b.err
