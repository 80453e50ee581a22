# scratch: the other BASELINE configs on one GPU (outputs -> gpurun_out/<tag>_*.json)
TAG=${1:-r01i}
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_cfg3.json 2> gpurun_out/${TAG}_cfg3.err
python bench.py --forward-only --points 6000000 --width 3840 --height 2160 --steps 20 --warmup 5 > gpurun_out/${TAG}_cfg5_sh3.json 2> gpurun_out/${TAG}_cfg5_sh3.err
python bench.py --forward-only --points 6000000 --width 3840 --height 2160 --sh-degree 0 --steps 20 --warmup 5 > gpurun_out/${TAG}_cfg5_sh0.json 2> gpurun_out/${TAG}_cfg5_sh0.err
python bench.py --points 6000000 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_cfg4_1gpu.json 2> gpurun_out/${TAG}_cfg4_1gpu.err
python - <<PY
import json
for n in ("cfg3","cfg5_sh3","cfg5_sh0","cfg4_1gpu"):
    try:
        d=json.loads(open("gpurun_out/${TAG}_%s.json"%n).read().strip().splitlines()[-1])
        st=d.get("stages_ms") or {k:v["ms"] for k,v in d["stages"].items()}
        print(n, d["metric"], round(d["value"],1), d["unit"], "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), {k:round(v,3) for k,v in st.items()}, d["config"]["counts"].get("R"))
    except Exception as e:
        print(n, "FAIL", e, open("gpurun_out/${TAG}_%s.err"%n).read()[-1500:])
PY
