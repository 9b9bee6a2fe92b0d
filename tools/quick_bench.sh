#!/bin/bash
# Developer tool: kernel-only frames/s of the bench workload for one library build.   usage: quick_bench.sh [lib.so] [ENV=VAL ...]
lib=$1; shift
env ITA_B200_LIB=$lib "$@" python bench.py --no-cpu-baseline --steps 50 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lib $*', round(d['value']/1e6,2), 'Mframes/s', d['roofline']['kernel_ms'], 'ms parity', d.get('parity_spot_check'))"
