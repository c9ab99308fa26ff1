# usage: bash scripts/gpu_check.sh [suffixes...]   -- GPU parity tests, then bench.py on each build variant
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -15
for v in "" "$@"; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --lib minicfd_b200/lib/libminicfd_b200$v.so > gpurun_out/bench_v$v.log 2>&1; python - <<PY
import json
l=[x for x in open("gpurun_out/bench_v$v.log") if x.startswith("{")]
d=json.loads(l[-1]) if l else None
print("variant '$v'", d and dict(value=d["value"], ms_step=d["ms_per_step"], ms_face=d["roofline"]["ms_face_per_stage"], ms_cell=d["roofline"]["ms_cell_per_stage"], e2e=d["e2e"]["value"]))
PY
done
