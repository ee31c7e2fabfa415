#!/bin/bash
# r03i (2 GPUs): the whole GPU suite once more after the test fix
out=gpurun_out; tag=r03i
timeout 1500 python -X faulthandler -m pytest tests -x -q -m gpu > $out/${tag}_pytest.log 2>&1; tail -6 $out/${tag}_pytest.log | cut -c1-400
timeout 300 python -m pytest tests/test_gpu_train_api.py -x -q -m gpu -k pipelined > $out/${tag}_pytest_pipe2.log 2>&1; tail -2 $out/${tag}_pytest_pipe2.log
timeout 300 python -m pytest tests/test_gpu_train_api.py -x -q -m gpu -k pipelined > $out/${tag}_pytest_pipe3.log 2>&1; tail -2 $out/${tag}_pytest_pipe3.log
