#!/bin/bash
# Round 2, first GPU pass: tests, step timing, bench (both arms), launch list.
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r2_tests.txt 2>&1; echo "pytest exit $?" >> $O/r2_tests.txt
tail -5 $O/r2_tests.txt
timeout 300 python profiles/stepbench.py > $O/r2_stepbench.txt 2>&1; cat $O/r2_stepbench.txt
timeout 600 python bench.py > $O/r2_bench.json 2> $O/r2_bench.err; tail -c 3000 $O/r2_bench.json; tail -3 $O/r2_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference.json 2>> $O/r2_bench.err
