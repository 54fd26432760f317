#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r1k_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1k_pytest.log
tail -12 gpurun_out/r1k_pytest.log
python bench.py --no-cpu-baseline > gpurun_out/r1k_bench.json 2> gpurun_out/r1k_bench.err; cut -c1-300 gpurun_out/r1k_bench.json
