set -x
python tools/model_throughput.py shufflenet_v2_x1_0 512 6 > gpurun_out/plain_shuffle.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stem_pool -s 2 -c 1 -f -o gpurun_out/r02_stempool_3x3 python tools/model_throughput.py shufflenet_v2_x1_0 512 6 > gpurun_out/ncu_stempool_3x3.log 2>&1
python tools/model_throughput.py resnet50 256 6 > gpurun_out/plain_resnet.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stem_pool -s 2 -c 1 -f -o gpurun_out/r02_stempool_7x7 python tools/model_throughput.py resnet50 256 6 > gpurun_out/ncu_stempool_7x7.log 2>&1
cat gpurun_out/plain_shuffle.log gpurun_out/plain_resnet.log
