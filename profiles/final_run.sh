#!/bin/bash
# round-1 final single-GPU evidence: stress tests, bench line, launch list of one timed step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k stress 2>&1 | grep -E "stress Q|passed|failed|Error" > gpurun_out/final_stress.log
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err || exit 1
tail -c 600 gpurun_out/final_bench.json
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 1500 -c 900 --csv \
  --log-file gpurun_out/final_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/final_ncu.log 2>&1
python profiles/legs.py > gpurun_out/final_legs.json 2> gpurun_out/final_legs.err
cat gpurun_out/final_stress.log
