#!/bin/bash
# A/B of library variants on the GPU box: tools/ab_bench.sh  (writes gpurun_out/ab_*.json; one bench line per variant)
cd "$(dirname "$0")/.."
PK=$(ls -d vistaf-*_b200)
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 400 python bench.py --no-cpu --no-maps --steps 5 --warmup 3 > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ab_$name.json").read().strip().splitlines()[-1])
    st = {s["stage"]: round(s["ms_per_frame"], 4) for s in d.get("stages", [])}
    print(round(d["value"], 1), d["e2e"]["value"] and round(d["e2e"]["value"], 1), d["check"], {k: st[k] for k in st if "irls" in k or "median" in k or "percentil" in k or "quality" in k})
except Exception as e:
    print("ERR", e)
PY
)"; }
cp $PK/lib/libvistaf.so /tmp/libvistaf_main.so
for v in "$@"; do
  case $v in
    alt:*) cp $PK/lib/libvistaf_${v#alt:}.so $PK/lib/libvistaf.so; run ${v#alt:}; cp /tmp/libvistaf_main.so $PK/lib/libvistaf.so ;;
    main) run main ;;
    main2) run main2 ;;
    env:*) run main_${v#env:} ${v#env:} ;;
  esac
done
