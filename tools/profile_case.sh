#!/bin/bash
# One ncu --set full capture of the refinement kernel for a bench.py case.  Usage: tools/profile_case.sh <tag> <bench args...>
tag=$1; shift
o=gpurun_out
small="--steps 1 --warmup 1 --no-e2e --no-cpu --no-adapter-e2e"
timeout 300 python bench.py $small "$@" > $o/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 $o/${tag}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:refine_tasks_kernel -c 1 -f -o $o/${tag} \
    python bench.py $small "$@" > $o/${tag}_ncu.log 2>&1; echo "ncu rc=$?"
