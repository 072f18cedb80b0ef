#!/bin/bash
# Bench line + ncu launch list + one ncu --set full capture of the refinement kernel.  Usage: tools/profile_round.sh <tag>
tag=${1:-prof}
o=gpurun_out
small="--steps 2 --warmup 1 --no-e2e --no-cpu --no-adapter-e2e"
timeout 900 python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err; echo "bench rc=$?"; cat $o/${tag}_bench.json
timeout 300 python bench.py $small > $o/${tag}_plain.log 2>&1 || { echo "plain run failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $o/${tag}_launches.csv \
    python bench.py $small > $o/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:refine_tasks_kernel -c 1 -f -o $o/${tag}_refine_tasks_full \
    python bench.py $small > $o/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
grep -c fmpnp $o/${tag}_launches.csv
