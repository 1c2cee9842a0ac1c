#!/bin/bash
# the static-scratch entry point: alone in a fresh process, then the whole GPU suite (it is collected last)
set -u
mkdir -p gpurun_out
echo "== alone"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -4
echo "== whole suite"; timeout 2400 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout 600 python bench.py --steps 3 --warmup 3 --cards 64 --e2e-cards 0 --no-cpu-baseline --hough-batch 0 --default-op-cards 32 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); c=d['configs']['c5_huge_scan']; print('c5', c['seconds'], c['oracle_equal'], 'value', d['value'])"
