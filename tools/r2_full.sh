#!/bin/bash
# round-2 check on one GPU box: the whole GPU suite (all failures listed), smoke(), the default bench line and a short reference arm
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -p no:cacheprovider > gpurun_out/noalloc_alone.log 2>&1; tail -5 gpurun_out/noalloc_alone.log
timeout 2400 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/full_tests.log 2>&1; tail -30 gpurun_out/full_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
if [ "$1" != "nobench" ]; then
timeout 900 python bench.py --kernels-out gpurun_out/kernels.json > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 1500 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 400 gpurun_out/bench_ref.json
python tools/show_bench.py gpurun_out/bench.json
fi
