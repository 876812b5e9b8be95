#!/bin/bash
# Times scenes with several library variants (tools/build_variant.sh): tools/variant_probe.sh "VARIANT_DIRS" CASE...   ("-" = the default build)
VARS=$1; shift
for v in $VARS; do
  d=$v; [ "$v" = "-" ] && d=""
  echo "variant=$v"
  RTC_LIB_DIR=$d python tools/perf_probe.py "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print(' ', d['scene'], d['res'], 'ms', d['ms_kernels'], d['serial_ms'])"
done
