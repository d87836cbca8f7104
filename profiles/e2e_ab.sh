#!/bin/bash
# A/B of the packed form downloads (LITAP_FORM_PACK) on one box: pinned and pageable end-to-end times of litap_flow_form, 8192^2
cd "$(dirname "$0")/.."
for rep in 1 2 3; do
  for pack in 1 0; do
    LITAP_FORM_PACK=$pack python bench.py --steps 5 --no-cpu --no-weti --no-files 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('pack=$pack', 'pinned %.1f ms' % d['e2e']['ms_per_step'], 'pageable %.1f ms' % d['e2e']['pageable']['ms_per_step'])"
  done
done
