#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2z_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r2z_tests.log | head -30 | cut -c1-300
python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
python bench.py --impl reference --steps 3 --warmup 1 2>/dev/null | grep "^{" | cut -c1-400
