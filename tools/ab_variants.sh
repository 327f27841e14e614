# A/B harness for kernel variants built into quest_b200/lib/variants/lib<X>.so: same box, same bench command.
# usage (on the GPU box): bash tools/ab_variants.sh A C H ...   (a variant suffixed with + also runs the fusion parity tests)
cp quest_b200/lib/libQuEST.so /tmp/keep.so
for spec in "$@"; do
  v=${spec%+}
  cp quest_b200/lib/variants/lib$v.so quest_b200/lib/libQuEST.so
  if [ "$spec" != "$v" ]; then
    python -m pytest tests/test_gpu_parity.py -x -q -k "fus or fuzz or bit_exact or golden" 2>&1 | tail -2
  fi
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || echo "variant $v failed"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/ab_$v.json"))
    r=d["roofline"]
    print("$v", "ms/step", round(d["ms_per_step"],2), "fused ms", round(r["ms_per_launch"],3), "launches", r["launches_timed"], "fma", d.get("contracted_arithmetic",{}).get("ms_per_step"), "clk", d["clocks"]["sm_mhz"])
except Exception as e: print("$v", "ERR", e)
PY
done
cp /tmp/keep.so quest_b200/lib/libQuEST.so
