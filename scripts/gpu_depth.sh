#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/c24_tests.log 2>&1; tail -2 $O/c24_tests.log
python bench_configs.py env_step_api > $O/c24_api.jsonl 2>&1; cut -c1-260 $O/c24_api.jsonl
SLB_NO_PYFAST=1 python bench_configs.py env_step_api >> $O/c24_api.jsonl 2>&1; tail -1 $O/c24_api.jsonl | cut -c1-260
