#!/bin/bash
# the no-alloc entry point alone in a fresh process, then the whole GPU suite (it runs last there too)
TAG=${1:-r02d}
python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/${TAG}_noalloc_alone.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_noalloc_alone.log
tail -5 gpurun_out/${TAG}_noalloc_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests_all.log 2>&1; echo "rc=$?" >> gpurun_out/${TAG}_tests_all.log
tail -6 gpurun_out/${TAG}_tests_all.log
