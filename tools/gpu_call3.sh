set -x
for m in shufflenet_v2_x1_0 efficientnet_v2_s yolov8n; do
  python bench.py --model $m --steps 20 --warmup 5 > gpurun_out/r02_bench_$m.json 2> gpurun_out/r02_bench_$m.err
done
python bench.py --model efficientnet_v2_s --hw 384 --steps 10 --warmup 3 --no-serving > gpurun_out/r02_bench_efficientnet_v2_s_384.json 2> gpurun_out/r02_bench_efficientnet_v2_s_384.err
( time python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/r02_ref_resnet50.json 2> gpurun_out/r02_ref_resnet50.err
mkdir -p gpurun_out/profiles_csv
python tools/profile_batches.py resnet50 gpurun_out/profiles_csv/resnet50_b200_summary.csv 512 > gpurun_out/prof_resnet.log 2>&1
python tools/profile_batches.py shufflenet_v2_x1_0 gpurun_out/profiles_csv/shufflenet_b200_summary.csv 512 > gpurun_out/prof_shuffle.log 2>&1
python tools/profile_batches.py efficientnet_v2_s gpurun_out/profiles_csv/efficientnetv2_b200_summary.csv 512 > gpurun_out/prof_eff.log 2>&1
python tools/profile_batches.py yolov8n gpurun_out/profiles_csv/yolov8n_b200_summary.csv 128 > gpurun_out/prof_yolo.log 2>&1
for spec in "resnet50 256" "shufflenet_v2_x1_0 512" "efficientnet_v2_s 128" "yolov8n 32"; do
  set -- $spec
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02_launches_$1_b$2.csv python bench.py --model $1 --steps 1 --warmup 3 --sustain-s 0 --no-serving --no-cpu-baseline --parity-rows 1 > gpurun_out/ncu_$1.log 2>&1
done
