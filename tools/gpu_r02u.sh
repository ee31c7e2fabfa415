#!/bin/bash
# r02u (1 GPU): in-step timeline of the training step (events in front of every compute-stream launch)
out=gpurun_out; tag=r02u
XV_TRAIN_TIMELINE=100 python bench.py --workload train --steps 150 --warmup 30 --no-cpu-baseline > $out/${tag}_a.json 2> $out/${tag}_timeline_default.txt; grep timeline $out/${tag}_timeline_default.txt | cut -c1-3000
XV_TRAIN_WGRAD_STREAM=0 XV_TRAIN_TIMELINE=100 python bench.py --workload train --steps 150 --warmup 30 --no-cpu-baseline > $out/${tag}_b.json 2> $out/${tag}_timeline_one_stream.txt; grep timeline $out/${tag}_timeline_one_stream.txt | cut -c1-3000
XV_TRAIN_BULK=0 XV_TRAIN_WGRAD_STREAM=0 XV_TRAIN_TIMELINE=100 python bench.py --workload train --steps 150 --warmup 30 --no-cpu-baseline > $out/${tag}_c.json 2> $out/${tag}_timeline_one_stream_nobulk.txt; grep "timeline. total" $out/${tag}_timeline_one_stream_nobulk.txt | cut -c1-3000
