#!/bin/bash
# r02m (2 GPUs): whole GPU suite (the no-alloc test runs last), then that test once more alone
out=gpurun_out; tag=r02m
timeout 1200 python -X faulthandler -m pytest tests -m gpu -q > $out/${tag}_pytest.log 2>&1; tail -15 $out/${tag}_pytest.log | cut -c1-500
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $out/${tag}_noalloc_alone.log 2>&1; tail -3 $out/${tag}_noalloc_alone.log | cut -c1-400
