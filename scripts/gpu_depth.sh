#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu > $O/c27_tests.log 2>&1; tail -2 $O/c27_tests.log
python bench_configs.py env_step_api > $O/c27_api.jsonl 2>&1; cut -c1-200 $O/c27_api.jsonl
