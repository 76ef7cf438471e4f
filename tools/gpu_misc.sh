#!/bin/bash
# the other BASELINE configs and the sustained run (1 GPU); JSON lines kept under gpurun_out/
mkdir -p gpurun_out
run() { name=$1; shift; timeout 600 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; tail -c 400 gpurun_out/$name.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/$name.json"))
    print("$name", "value %.4g" % d["value"], "e2e %.4g" % d["e2e"]["value"], "lag_ok", d["e2e"].get("lag_known_answer_ok"), {k: "%.4g" % v["value"] for k,v in d.get("sustained",{}).items()}, d.get("clocks"))
except Exception as e:
    print("$name failed", e)
PY
}
run r2_sustained --sustained 1000 --no-cpu
run r2_config1_four --config four --no-cpu
run r2_config3_big65536 --config big65536 --steps 16 --no-cpu
run r2_config3_big65536_track --config big65536 --steps 16 --no-cpu --mode track
run r2_config5_stream_mixed --config stream --mode mixed --no-cpu --steps 50
run r2_bench_1gpu
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; head -c 1500 gpurun_out/r2_bench_reference.json
