#!/bin/bash
# quick bench: prints tok/s, roofline fraction and the per-kernel step profile for the given env settings
export B2_MODEL_DIR=/tmp/b2gguf-models
python bench.py --steps ${STEPS:-64} --warmup 8 --no-cpu --no-batch --no-tp ${QB_ARGS:-} 2>&1 | python -c "
import sys,json
for l in sys.stdin.read().strip().splitlines():
    try: d=json.loads(l)
    except Exception: print(l); continue
    print('tok/s %.1f e2e %.1f frac %.3f prof %s' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['profile_step_ms']))
"
