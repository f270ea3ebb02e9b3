#!/bin/bash
# usage: run_ncu_list.sh <tag> [env assignments...]   -> gpurun_out/launches_<tag>.csv
tag=$1; shift
env "$@" python experiments/bf16_infer_once.py 2 > gpurun_out/plain_$tag.log 2>&1 && env "$@" ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_$tag.csv python experiments/bf16_infer_once.py 2 > gpurun_out/ncu_$tag.log 2>&1
