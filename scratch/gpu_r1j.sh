#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1j_pytest.log
tail -5 gpurun_out/r1j_pytest.log
python scratch/ur5_sim.py > gpurun_out/r1j_ur5_sim.json 2> gpurun_out/r1j_ur5_sim.err; cut -c1-600 gpurun_out/r1j_ur5_sim.json; tail -3 gpurun_out/r1j_ur5_sim.err
python bench.py --no-cpu-baseline > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err; cut -c1-300 gpurun_out/r1j_bench.json
