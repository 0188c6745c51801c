set -x
python -m pytest tests/test_gpu_train_driver.py -m gpu -q 2>&1 | tail -15 > gpurun_out/t3.log
python bench.py --steps 5 --warmup 3 --mode train --no-cpu-baseline > gpurun_out/bench_train1.json 2> gpurun_out/bench_train1.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:cbf_forward -s 2 -c 2 -o gpurun_out/prof_cbf_r1 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
cat gpurun_out/t3.log; cat gpurun_out/bench_train1.json; tail -3 gpurun_out/bench_train1.err
