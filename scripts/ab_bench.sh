#!/bin/bash
# A/B kernel timing on one box: the baseline worktree (_base, built from an earlier commit) against the working tree,
# same bench.py workload.  usage: scripts/ab_bench.sh [bench args]
for d in _base _v1 _v2 _v3 .; do
  if [ -d "$d/enarf-gan_b200" ]; then
    (cd $d && python bench.py --no-cpu-baseline --no-gpu-reference "$@" 2>/dev/null | python -c "
import json,sys; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=j['roofline']['kernel_ms_all']
print('$d', 'step %.3f e2e %.3f' % (j['ms_per_step'], j['e2e']['ms_per_step']), ' '.join('%s %.3f' % (n, v) for n, v in k.items()))")
  fi
done
