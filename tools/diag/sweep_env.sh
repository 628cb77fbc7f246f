#!/bin/bash
# usage: sweep_env.sh VAR v1 v2 ...   -> stage times + value of a short bench run per setting (tuning helper)
var=$1; shift
for v in "$@"; do
  env $var=$v python bench.py --steps 3 --warmup 3 --no-extras --no-allpairs --skip-cpu > /tmp/sw.json 2> /tmp/sw.err || { tail -5 /tmp/sw.err; continue; }
  python - "$var" "$v" <<'PY'
import json,sys
d=json.loads(open('/tmp/sw.json').read().strip().splitlines()[-1])
print(sys.argv[1], sys.argv[2], "value %.0f e2e %.0f" % (d["value"], d["e2e"]["value"]), d["roofline"]["stage_ms_per_image_batch"], d["roofline"]["stereo_match_ms"])
PY
done
