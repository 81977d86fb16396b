#!/bin/bash
# smoke() and the default bench line, as the driver runs them
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== smoke"; timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
echo "== bench default"; ( time timeout 900 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo rc=$?; cat gpurun_out/bench_default.json; tail -4 gpurun_out/bench_default.err
