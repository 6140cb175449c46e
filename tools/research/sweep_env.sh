#!/bin/bash
# usage: sweep_env.sh VAR v1 v2 ...   -> bench line (value only) per setting
var=$1; shift
for v in "$@"; do
  echo -n "$var=$v  "
  env $var=$v python bench.py --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), d['ms_per_step'])"
done
