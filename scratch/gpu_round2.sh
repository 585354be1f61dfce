#!/bin/bash
# Round-2 evidence run on the GPU box: tests, bench lines, ncu launch list, ncu full captures (each after its command passed without ncu).
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/r02_gputests.log
python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"
python bench.py --workload cfg5 > gpurun_out/r02_bench_cfg5.json 2> gpurun_out/r02_bench_cfg5.err; echo "bench cfg5 rc=$?"
python bench.py --workload cfg4 > gpurun_out/r02_bench_cfg4.json 2> gpurun_out/r02_bench_cfg4.err; echo "bench cfg4 rc=$?"
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err; echo "ref rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02_ncu_launches.log 2>&1; echo "ncu list rc=$?"
python scratch/prof_case.py 64 8192 32 2 2 && \
ncu --set full --clock-control none --import-source on -k regex:"factor_kernel|solve_quad32" -s 2 -c 2 -f -o gpurun_out/r02_prof_cfg3 python scratch/prof_case.py 64 8192 32 2 2 > gpurun_out/r02_ncu_cfg3.log 2>&1; echo "ncu cfg3 rc=$?"
python scratch/prof_case.py 4096 512 8 2 2 && \
ncu --set full --clock-control none --import-source on -k regex:"factor_kernel|solve_chain_pf" -s 2 -c 2 -f -o gpurun_out/r02_prof_cfg5 python scratch/prof_case.py 4096 512 8 2 2 > gpurun_out/r02_ncu_cfg5.log 2>&1; echo "ncu cfg5 rc=$?"
ls -la gpurun_out
