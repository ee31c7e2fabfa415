#!/bin/bash
# r03m (1 GPU, half a minute): Trainer.valid on the device
out=gpurun_out; tag=r03m
timeout 100 python -m pytest tests/test_gpu_train_api.py -x -q -m gpu -k "valid or epoch" > $out/${tag}_pytest.log 2>&1; tail -15 $out/${tag}_pytest.log | cut -c1-400
