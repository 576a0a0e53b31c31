set -x
rm -f gpurun_out/r02i_*.jsonl
# H2: two independent single-GPU deployments side by side
CUDA_VISIBLE_DEVICES=0 python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 4 --clients 3 --metrics-file gpurun_out/m0.json --out gpurun_out/r02i_ind0.jsonl > /dev/null 2> gpurun_out/r02i_ind0.err &
CUDA_VISIBLE_DEVICES=1 python tools/serve_sim.py --models resnet --gpus 1 --rate 90000 --seconds 4 --clients 3 --metrics-file gpurun_out/m1.json --out gpurun_out/r02i_ind1.jsonl > /dev/null 2> gpurun_out/r02i_ind1.err &
wait
# H1: one deployment, 2 GPUs, only 3 clients
python tools/serve_sim.py --models resnet --gpus 2 --rate 120000 --seconds 4 --clients 3 --out gpurun_out/r02i_c3.jsonl > /dev/null 2> gpurun_out/r02i_c3.err
