#!/bin/bash
out=gpurun_out; tag=r02q
timeout 600 python -m pytest tests/test_gpu_train_api.py -m gpu -q -s -k pipelined > $out/${tag}_pytest_pipe.log 2>&1; grep -n "^sync\|^pipe\|passed\|failed\|Error" $out/${tag}_pytest_pipe.log | cut -c1-300 | head -40
