#!/bin/bash
# round-1 session-c GPU pass: parity tests, both bench arms, ncu launch list + one full capture
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1c_pytest.log
python bench.py --impl reference --steps 100 --warmup 3 > gpurun_out/r1c_bench_ref.json 2> gpurun_out/r1c_bench_ref.err
python bench.py > gpurun_out/r1c_bench_full.json 2> gpurun_out/r1c_bench_full.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1c_smoke.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1c_launches.csv python bench.py --steps 20 --warmup 3 > gpurun_out/r1c_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rmpc_closed_loop -s 5 -c 1 -o gpurun_out/r1c_full -f python scratch/prof_chunk.py 1 > gpurun_out/r1c_ncu_full.log 2>&1
ncu -i gpurun_out/r1c_full.ncu-rep --page raw --csv > gpurun_out/r1c_full_raw.csv 2>/dev/null
tail -3 gpurun_out/r1c_pytest.log; cat gpurun_out/r1c_bench_full.json; cat gpurun_out/r1c_bench_ref.json; cat gpurun_out/r1c_smoke.log | tail -2
