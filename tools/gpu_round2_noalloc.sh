#!/bin/bash
# sb200_prune_rows_static: its test alone in a fresh process, then the whole GPU suite (it is collected last).
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > $O/r02_noalloc_alone.log 2>&1; echo "pytest_rc=$?" >> $O/r02_noalloc_alone.log; tail -4 $O/r02_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_noalloc_suite.log 2>&1; echo "pytest_rc=$?" >> $O/r02_noalloc_suite.log; tail -4 $O/r02_noalloc_suite.log
