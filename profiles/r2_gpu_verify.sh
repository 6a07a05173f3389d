#!/bin/bash
# full GPU suite on the product build and on the bounds-checked build, smoke, default bench line
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/verify_pytest.txt; cat gpurun_out/verify_pytest.txt
EAGS_RASTER_LIB=$PWD/variants/checks.so timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4 > gpurun_out/verify_pytest_checks.txt; cat gpurun_out/verify_pytest_checks.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/verify_smoke.txt 2>&1; tail -1 gpurun_out/verify_smoke.txt
timeout 900 python bench.py > gpurun_out/verify_bench.json 2> gpurun_out/verify_bench.err; tail -c 300 gpurun_out/verify_bench.json
