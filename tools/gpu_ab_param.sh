#!/bin/bash
# A/B of kmc_set_param settings in the timed loop of bench.py, alternating.  $1 = "ID=V ID=V ..." variants separated by
# commas (an empty variant = defaults); further arguments go to bench.py.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
IFS=',' read -ra VARS <<< "$1"; shift
for rep in 1 2; do for v in "${VARS[@]}"; do
  extra=""; for kv in $v; do extra="$extra --param $kv"; done
  tag=$(echo "$v" | tr ' =' '__'); [ -z "$tag" ] && tag=default
  timeout 900 python bench.py "$@" --no-e2e --no-cpu-baseline $extra > gpurun_out/abp_$tag.json 2> gpurun_out/abp_$tag.err
  python - gpurun_out/abp_$tag.json "$tag" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); c=d["config"]
    print(sys.argv[2], round(d["ms_per_step"],3), d["roofline"]["phase_ms"], "refined", c["refined_rows_last_step"], "parity", c["checks"].get("parity",{}).get("ok"), d["clocks"]["sm_mhz"])
except Exception as e:
    print(sys.argv[2], "ERR", e); print(open(sys.argv[1].replace(".json",".err")).read()[-800:])
PY
done; done
