#!/bin/bash
O=gpurun_out
python -m pytest tests -x -q -m gpu -k "back_to_back" > $O/c28_tests.log 2>&1; tail -3 $O/c28_tests.log
