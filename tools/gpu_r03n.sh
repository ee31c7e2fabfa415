#!/bin/bash
# r03n (1 GPU): the whole GPU suite once more after the host-side changes (reader pool, segment reads, Trainer.valid)
out=gpurun_out; tag=r03n
timeout 100 python -m pytest tests -x -q -m gpu > $out/${tag}_pytest.log 2>&1; tail -4 $out/${tag}_pytest.log | cut -c1-300
