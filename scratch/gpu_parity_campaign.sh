#!/bin/bash
mkdir -p gpurun_out
P=profiles/tools/parity_campaign.py
O=gpurun_out/r1j_parity_campaign.jsonl
: > $O
timeout 600 python $P --batch 4096 --steps 60 --horizon 100 --solver SQP_RTI >> $O 2>gpurun_out/pc.err
timeout 600 python $P --batch 1024 --steps 40 --horizon 20 --solver SQP --seed 43 >> $O 2>>gpurun_out/pc.err
timeout 600 python $P --batch 1024 --steps 60 --horizon 30 --solver SQP_RTI --robot ur5 --seed 45 >> $O 2>>gpurun_out/pc.err
timeout 600 python $P --batch 512 --steps 30 --horizon 300 --solver SQP_RTI --seed 46 >> $O 2>>gpurun_out/pc.err
cut -c1-700 $O; tail -3 gpurun_out/pc.err
