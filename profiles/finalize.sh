#!/bin/bash
# profiles/finalize.sh <batch tag, e.g. b23> — copies one confirmation batch (profiles/r02/batch23.sh style) out of the
# scratch gpurun_out/ into the tracked profiles/r02_final_* files and regenerates the SASS histogram.
set -e
T=$1; O=gpurun_out
tail -1 $O/${T}_bench.json > profiles/r02_final_bench.json
tail -1 $O/${T}_bench_ref.json > profiles/r02_final_bench_reference.json
cp $O/${T}_launches.csv profiles/r02_final_launches.csv
cp $O/${T}_summary.txt profiles/r02_final_gputests.txt
python profiles/summarize.py r02_final $O/${T}_launches.csv $O/${T}_gather_full.ncu-rep $O/${T}_bench.json > /dev/null
python profiles/sass_histogram.py > /dev/null
ls -la profiles/r02_final_* profiles/r02_sass_histogram.txt
