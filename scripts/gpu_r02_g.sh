#!/bin/bash
# round-2 call G: the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last there)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r02g_pytest_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r02g_pytest_noalloc_alone.log | cut -c1-250
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02g_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02g_pytest.log | cut -c1-250
